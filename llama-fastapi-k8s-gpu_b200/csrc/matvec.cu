// Decode mat-vec: host side (planning, launches). The kernel and its design notes are in matvec_body.cuh; this file
// instantiates it for the block formats of llama.cpp's common mixes, matvec_extra.cu for the rest.
#include "matvec_body.cuh"
#include <cuda.h>

namespace b2 {

cudaError_t mv_extra_init(int smem_max);                                           // matvec_extra.cu
cudaError_t mv_extra_launch(cudaLaunchConfig_t* cfg, const MvParams& p, bool act2);  // matvec_extra.cu

// ---------------------------------------------------------------- host side
static int g_stage_cap = 0;      // B2_MV_STAGES (tuning): read once in mv_init
static int g_n_prod = NPROD;     // B2_MV_PROD (tuning)
static int g_no_tma2d = 0;       // B2_MV_NO_TMA2D=1 (tuning): always four 1-D copies per tile
static int g_tma_modes = 15;     // B2_MV_TMA_MODES (diagnosis): bit m set = launches of MvMode m may use the tensor copies

typedef CUresult (*MvEncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static MvEncodeTiledFn g_mv_encode = nullptr;

cudaError_t mv_init() {   // per device: allow > 48 KB dynamic shared memory; read the tuning knobs once
    static bool env_read = false;
    if (!env_read) {
        if (const char* s = getenv("B2_MV_STAGES")) g_stage_cap = atoi(s);
        if (const char* s = getenv("B2_MV_PROD")) { const int v = atoi(s); if (v >= 1 && v <= NPROD) g_n_prod = v; }
        if (const char* s = getenv("B2_MV_NO_TMA2D")) g_no_tma2d = atoi(s);
        if (const char* s = getenv("B2_MV_TMA_MODES")) g_tma_modes = atoi(s);
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess) g_mv_encode = (MvEncodeTiledFn)fn;
        cudaGetLastError();
        env_read = true;
    }
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(matvec_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, MV_SMEM_MAX)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(matvec_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, MV_SMEM_MAX)) != cudaSuccess) return e;
    return mv_extra_init(MV_SMEM_MAX);
}

int mv_plan(MvParams& p, int n_ctas) {
    if (p.K % 256 || p.K <= 0) return -1;
    if (p.K > 32768) return -2;
    int max_sb = 0;
    for (int s = 0; s < p.nseg; s++) {
        MvSeg& sg = p.seg[s];
        const int be = block_elems(sg.type), bb = block_bytes(sg.type);
        if (!bb || be == 1) return -3;
        sg.row_bytes = p.K / be * bb;
        if (sg.n_rows % 2) return -4;
        sg.slice_bytes = slice_bytes(sg.type);
        int sb = sg.slice_bytes;
        if (sb > sg.row_bytes) sb = sg.row_bytes;
        max_sb = sb > max_sb ? sb : max_sb;
    }
    if (p.mode == MV_GATEUP) {
        if (p.nseg != 2 || p.seg[0].n_rows != p.seg[1].n_rows) return -6;
        if (p.seg[0].type != p.seg[1].type) return -8;     // a mixed pair is run unfused by the caller
    }
    // One 2-D tensor copy per tile where every segment allows it: rows and slices on 16-byte boundaries, a slice within
    // the 256-element box limit (8-byte elements: 2048 B; Q8_0 slices are 2176 B), rows at least a slice long.
    static_assert(sizeof(CUtensorMap) == sizeof(p.seg[0].tmap), "tensor map storage");
    p.use_tma = (g_mv_encode && !g_no_tma2d && ((g_tma_modes >> p.mode) & 1)) ? 1 : 0;
    for (int s = 0; s < p.nseg && p.use_tma; s++) {
        const MvSeg& sg = p.seg[s];
        if (sg.row_bytes % 16 || ((uintptr_t)sg.w & 15) || sg.slice_bytes % 16 || sg.row_bytes < sg.slice_bytes) p.use_tma = 0;
        else if (sg.slice_bytes > 2048) {       // wider than one box of 256 8-byte elements: two half-slices (every segment then)
            if (sg.slice_bytes % 32 || sg.slice_bytes > 4096 || sg.row_bytes % (sg.slice_bytes / 2) || (s > 0 && p.use_tma != 2)) p.use_tma = 0;
            else p.use_tma = 2;
        } else if (p.use_tma == 2) p.use_tma = 0;
    }
    for (int s = 0; s < p.nseg && p.use_tma; s++) {
        MvSeg& sg = p.seg[s];
        const cuuint32_t estr[3] = {1, 1, 1};
        CUresult r;
        if (p.use_tma == 1) {
            const cuuint64_t dims[2] = {(cuuint64_t)sg.row_bytes / 8, (cuuint64_t)sg.n_rows};
            const cuuint64_t strides[1] = {(cuuint64_t)sg.row_bytes};
            const cuuint32_t boxd[2] = {(cuuint32_t)sg.slice_bytes / 8, (cuuint32_t)R};
            r = g_mv_encode(reinterpret_cast<CUtensorMap*>(sg.tmap), CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, (void*)sg.w, dims, strides, boxd, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        } else {
            const int half = sg.slice_bytes / 2;
            const cuuint64_t dims[3] = {(cuuint64_t)half / 8, (cuuint64_t)(sg.row_bytes / half), (cuuint64_t)sg.n_rows};
            const cuuint64_t strides[2] = {(cuuint64_t)half, (cuuint64_t)sg.row_bytes};
            const cuuint32_t boxd[3] = {(cuuint32_t)half / 8, 2, (cuuint32_t)R};
            r = g_mv_encode(reinterpret_cast<CUtensorMap*>(sg.tmap), CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, (void*)sg.w, dims, strides, boxd, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        }
        if (r != CUDA_SUCCESS) p.use_tma = 0;
    }
    p.n_slices = (p.K + B2_SLICE - 1) / B2_SLICE;
    p.slot_bytes = (max_sb + 16 + 15) / 16 * 16;           // + slack for the 16-byte alignment of each copy
    p.stage_bytes = (R * p.slot_bytes + 127) / 128 * 128;
    // shared memory: [header][plane a][plane b][K/8 floats of packed scales][partial-sum slots][ring]
    p.act_plane = p.n_slices * B2_SLICE;
    const int act_block = (2 * p.act_plane + p.K / 2 + 127) / 128 * 128;
    const int tpg = p.n_slices * (p.mode == MV_GATEUP ? 2 : 1);
    p.act_off = MV_SHARED_HDR;
    p.part_off = MV_SHARED_HDR + act_block;
    p.ring_off = p.part_off + (tpg > 1 ? MV_GSLOTS * tpg * 16 : 0);
    p.ring_off = (p.ring_off + 127) / 128 * 128;
    int S = (MV_SMEM_MAX - p.ring_off) / p.stage_bytes;
    if (S > MV_MAX_STAGES) S = MV_MAX_STAGES;
    if (g_stage_cap >= TPB && g_stage_cap < S) S = g_stage_cap;
    if (S < TPB) return -5;                                // a producer warp fills TPB stages per round
    p.n_prod = g_n_prod;
    while (p.n_prod > 1 && S < TPB * p.n_prod) p.n_prod--;
    S -= S % p.n_prod;                                     // every stage belongs to one producer (entry % n_prod)
    p.n_stages = S;
    // CTA split proportional to bytes (gate/up: every CTA handles both matrices)
    if (p.mode == MV_GATEUP) {
        p.seg[0].cta0 = p.seg[1].cta0 = 0; p.seg[0].ncta = p.seg[1].ncta = n_ctas;
    } else {
        if (n_ctas < p.nseg) return -7;
        double tot = 0;
        for (int s = 0; s < p.nseg; s++) tot += (double)p.seg[s].n_rows * p.seg[s].row_bytes;
        int left = n_ctas, c0 = 0;
        for (int s = 0; s < p.nseg; s++) {
            int nc;
            if (s == p.nseg - 1) nc = left;
            else {
                nc = (int)((double)p.seg[s].n_rows * p.seg[s].row_bytes / tot * n_ctas + 0.5);
                const int remaining_segs = p.nseg - 1 - s;
                if (nc < 1) nc = 1;
                if (nc > left - remaining_segs) nc = left - remaining_segs;
            }
            p.seg[s].cta0 = c0; p.seg[s].ncta = nc;
            c0 += nc; left -= nc;
        }
    }
    return 0;
}

int mv_ctas_per_sm() { return 1; }
size_t mv_smem_bytes(const MvParams& p) { return (size_t)p.ring_off + (size_t)p.n_stages * p.stage_bytes; }

cudaError_t mv_launch(const MvParams& p, int n_ctas, cudaStream_t stream, bool act2, bool pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(n_ctas); cfg.blockDim = dim3(MV_THREADS);
    cfg.dynamicSmemBytes = mv_smem_bytes(p); cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    bool extra = false;
    for (int i = 0; i < p.nseg; i++) {
        const int t = p.seg[i].type;
        if (!(t == T_Q4_K || t == T_Q6_K || t == T_Q8_0 || t == T_Q4_0 || t == T_Q5_K)) extra = true;
    }
    if (extra) return mv_extra_launch(&cfg, p, act2);
    return act2 ? cudaLaunchKernelEx(&cfg, matvec_kernel<true, false>, p) : cudaLaunchKernelEx(&cfg, matvec_kernel<false, false>, p);
}

}  // namespace b2
