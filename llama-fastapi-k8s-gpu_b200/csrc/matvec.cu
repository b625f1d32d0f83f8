// Decode mat-vec with in-kernel dequantisation: y = W x for GGUF-quantised W (Q4_0, Q8_0, Q4_K,
// Q5_K, Q6_K), weights streamed unchanged from HBM. Replaces ggml-cuda's quantize_q8_1 +
// mul_mat_vec_q + the separate rms_norm / rope / cpy / silu / add launches that the reference
// runs under llm.create_chat_completion (/root/reference/api.py:55-63; SURVEY.md 2.3 K2-K7,K11).
//
// Structure (one CTA per SM, 9 warps):
//   * warp 8: TMA producer. A tile = one K-slice (2048 elements) of a group of R=4 consecutive
//     rows = 4 row-slices of 1-2 KB, brought into one ring stage with 1-D bulk copies
//     (cp.async.bulk + mbarrier complete_tx); 4 lanes serve a tile, 8 tiles are issued per round.
//     The producer never touches activations, so with programmatic dependent launch it starts
//     prefetching weights while the previous kernel is still running.
//   * warps 0-7: autonomous consumers. Tile q belongs to warp q % 8; a warp walks its row groups,
//     and for each group the K-slices in order, keeping 4 row accumulators in registers. Per
//     slice a lane loads its 64 quantised activations (two int8 planes) from shared memory once
//     and reuses them for the 4 rows; weights stay packed and go through dp4a. No cross-warp
//     reduction and no block-wide barrier in the main loop: a warp only waits on its own tile's
//     mbarrier, reduces its 4 rows with a transposed butterfly (6 shuffles) and writes them out.
//   * prologue (all consumers, after griddepcontrol.wait): x -> optional fused RMSNorm -> two
//     int8 planes + per-32 scale + per-16 sums in shared memory.
//   * epilogue (lanes 0 and 16, a row pair each): plain store, residual add, RoPE + paged-KV
//     store for the fused QKV projection, or SiLU(gate) * up for the fused FFN projection.
//
// Roofline: HBM. Algorithmic bytes per launch = sum of the weight bytes of its segments.
#include "matvec_phase.cuh"
#include <stdlib.h>

namespace b2 {

template <int NCW, bool ACT2>
__global__ void __launch_bounds__((NCW + (NCW <= 8 ? 1 : MV_MAX_PROD)) * 32, NCW <= 8 ? 2 : 1) matvec_kernel(const MvParams p) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    MvShared<NCW>& sh = *reinterpret_cast<MvShared<NCW>*>(smem_raw);
    uint8_t* ring = smem_raw + p.ring_off;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    RingGeom geo; geo.slot_bytes = p.slot_bytes; geo.stage_bytes = p.stage_bytes; geo.n_stages = p.n_stages;
    const CtaWork work = cta_work(p, blockIdx.x, gridDim.x);

    if (threadIdx.x == 0) mv_ring_init<NCW>(sh, geo.n_stages);
    __syncthreads();
    griddep_launch();   // let the next kernel start prefetching its own weights

    if (warp >= NCW) {  // producer warps never touch activations: they run ahead of the dependency wait
        mv_produce<NCW>(p, work, sh, ring, geo, 0u, warp - NCW, p.n_prod, lane);
        return;
    }
    griddep_wait();     // activations written by the previous kernel are visible from here on
    mv_consume_phase<NCW, ACT2>(p, work, sh, smem_raw + MV_SHARED_HDR, ring, geo, 0u, warp, lane);
}

// ---------------------------------------------------------------- host side
static int g_ncw = 0;   // consumer warps per CTA: 8 (two CTAs per SM) or 16 / 20 (one CTA per SM)

static int mv_ncw() {
    if (!g_ncw) {
        g_ncw = 16;
        if (const char* s = getenv("B2_MV_WARPS")) { int v = atoi(s); if (v == 8 || v == 16 || v == 20) g_ncw = v; }
    }
    return g_ncw;
}
int mv_ctas_per_sm_default() { return mv_ncw() <= 8 ? 2 : 1; }
static int mv_smem_max() { return mv_ncw() <= 8 ? 112 * 1024 : 224 * 1024; }

template <int NCW>
static cudaError_t set_attr(int bytes) {
    cudaError_t e = cudaFuncSetAttribute(matvec_kernel<NCW, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(matvec_kernel<NCW, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

cudaError_t mv_init() {   // per device: allow > 48 KB dynamic shared memory
    cudaError_t e;
    if ((e = set_attr<8>(112 * 1024)) != cudaSuccess) return e;
    if ((e = set_attr<16>(224 * 1024)) != cudaSuccess) return e;
    return set_attr<20>(224 * 1024);
}

int mv_plan(MvParams& p, int n_ctas) {
    if (p.K % 256 || p.K <= 0) return -1;
    if (p.K > 32768) return -2;
    int max_sb = 0;
    for (int s = 0; s < p.nseg; s++) {
        int be = block_elems(p.seg[s].type), bb = block_bytes(p.seg[s].type);
        if (!bb || be == 1) return -3;
        p.seg[s].row_bytes = p.K / be * bb;
        if (p.seg[s].n_rows % 2) return -4;
        int sb = slice_bytes(p.seg[s].type);
        if (sb > p.seg[s].row_bytes) sb = p.seg[s].row_bytes;
        max_sb = sb > max_sb ? sb : max_sb;
    }
    p.n_slices = (p.K + B2_SLICE - 1) / B2_SLICE;
    p.slot_bytes = (max_sb + 16 + 15) / 16 * 16;           // + slack for the 16-byte alignment of each copy
    p.stage_bytes = (R * p.slot_bytes + 127) / 128 * 128;
    // short rows: a group of R consecutive full rows is one contiguous run -> one bulk copy per tile
    int max_rb = 0;
    for (int s = 0; s < p.nseg; s++) max_rb = p.seg[s].row_bytes > max_rb ? p.seg[s].row_bytes : max_rb;
    p.whole_rows = 0;   // measured slower than per-slice tiles on B200 (fewer tiles in flight); kept as a knob
    if (const char* s = getenv("B2_MV_WHOLE_ROWS")) p.whole_rows = atoi(s) && (R * max_rb + 32 <= 48 * 1024);
    if (p.whole_rows) p.stage_bytes = (R * max_rb + 32 + 127) / 128 * 128;
    p.act_layout = act_layout_of(p.seg[0].type);
    p.act_dual = 0;
    for (int s = 1; s < p.nseg; s++)
        if (p.mode == MV_GATEUP && act_layout_of(p.seg[s].type) != p.act_layout) p.act_dual = 1;
    int act_bytes = (p.act_dual ? 4 : 2) * p.n_slices * B2_SLICE + (p.K / 32) * 4 + (p.K / 16) * 4;
    p.ring_off = (MV_SHARED_HDR + act_bytes + 127) / 128 * 128;
    int S = (mv_smem_max() - p.ring_off) / p.stage_bytes;
    if (S > MV_MAX_STAGES) S = MV_MAX_STAGES;
    if (const char* cap = getenv("B2_MV_STAGES")) { int c = atoi(cap); if (c >= TPB && c < S) S = c; }   // tuning knob
    if (S < TPB) return -5;                                // the producer fills TPB stages per round
    p.n_stages = S;
    p.skip_compute = getenv("B2_MV_SKIP_COMPUTE") ? 1 : 0;   // experiment: measure the TMA ring alone
    p.n_prod = mv_ncw() <= 8 ? 1 : 2;
    if (const char* s = getenv("B2_MV_PROD")) { int v = atoi(s); if (v >= 1 && v <= MV_MAX_PROD && mv_ncw() > 8) p.n_prod = v; }
    if (S < TPB * p.n_prod) p.n_prod = 1;                  // each producer warp fills TPB stages per round
    S -= S % p.n_prod;                                     // every stage belongs to one producer (entry % n_prod)
    p.n_stages = S;
    // CTA split proportional to bytes (gate/up: every CTA handles both matrices)
    if (p.mode == MV_GATEUP) {
        if (p.nseg != 2 || p.seg[0].n_rows != p.seg[1].n_rows) return -6;
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
                int remaining_segs = p.nseg - 1 - s;
                if (nc < 1) nc = 1;
                if (nc > left - remaining_segs) nc = left - remaining_segs;
            }
            p.seg[s].cta0 = c0; p.seg[s].ncta = nc;
            c0 += nc; left -= nc;
        }
    }
    return 0;
}

size_t mv_smem_bytes(const MvParams& p) { return (size_t)p.ring_off + (size_t)p.n_stages * p.stage_bytes; }

template <int NCW>
static cudaError_t launch_t(const MvParams& p, int n_ctas, cudaStream_t stream, bool act2, bool pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(n_ctas); cfg.blockDim = dim3((NCW + p.n_prod) * 32);
    cfg.dynamicSmemBytes = mv_smem_bytes(p); cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    if (act2) return cudaLaunchKernelEx(&cfg, matvec_kernel<NCW, true>, p);
    return cudaLaunchKernelEx(&cfg, matvec_kernel<NCW, false>, p);
}

cudaError_t mv_launch(const MvParams& p, int n_ctas, cudaStream_t stream, bool act2, bool pdl) {
    switch (mv_ncw()) {
        case 8: return launch_t<8>(p, n_ctas, stream, act2, pdl);
        case 20: return launch_t<20>(p, n_ctas, stream, act2, pdl);
        default: return launch_t<16>(p, n_ctas, stream, act2, pdl);
    }
}

}  // namespace b2
