// Decode mat-vec with in-kernel dequantisation: y = W x for GGUF-quantised W (Q4_0, Q8_0, Q4_K,
// Q5_K, Q6_K), weights streamed unchanged from HBM. Replaces ggml-cuda's quantize_q8_1 +
// mul_mat_vec_q + the separate rms_norm / rope / cpy / silu / add launches that the reference
// runs under llm.create_chat_completion (/root/reference/api.py:55-63; SURVEY.md 2.3 K2-K7,K11).
//
// Structure (one CTA per SM, 9 warps):
//   * warp 8, one lane: TMA producer. Streams this CTA's contiguous row range through a ring of
//     shared-memory stages with 1-D bulk copies (cp.async.bulk + mbarrier complete_tx). It never
//     touches activations, so with programmatic dependent launch it starts prefetching weights
//     while the previous kernel is still running.
//   * warps 0-7: consumers. Each lane owns one 64-element chunk of the input vector for the whole
//     kernel: it loads its 64 floats once (after griddepcontrol.wait), optionally applies the fused
//     RMSNorm, quantises them to two int8 planes held in registers, and then only streams weight
//     chunks out of shared memory through dp4a. W_k = ceil(K/2048) warps cover one row; the 8
//     warps form G = 8 / W_k groups that work on different row pairs of a stage.
//   * epilogue (one lane per row pair): plain store, residual add, RoPE + paged-KV store for the
//     fused QKV projection, or SiLU(gate) * up for the fused FFN projection.
//
// Roofline: HBM. Algorithmic bytes per launch = sum of the weight bytes of its segments.
#include "kernels.h"
#include "ptx.cuh"
#include "quant_blocks.cuh"
#include <cuda_fp16.h>

namespace b2 {

constexpr int NCW = 8;                 // consumer warps
constexpr int NTHREADS = (NCW + 1) * 32;

struct __align__(16) MvShared {
    uint64_t full[MV_MAX_STAGES];
    uint64_t empty[MV_MAX_STAGES];
    float red[NCW];                    // sum-of-squares partials
    float part[2][NCW][MV_MAX_PAIRS][2];   // cross-warp partial dot products, double buffered by tile
    float gate[NCW][MV_MAX_PAIRS][2];  // gate results waiting for the matching up tile
};

struct TileInfo {
    const uint8_t* src;    // 16-byte aligned start of the bulk copy
    uint32_t bytes;        // multiple of 16
    uint32_t off;          // offset of the first row inside the stage
    int seg;               // weight matrix index
    int row0;              // first row (within the segment)
    int nrows;
};

// CTA-uniform description of the work: which segment(s), which rows.
struct CtaWork {
    int mode_gateup;
    int seg;        // segment for non-gateup modes
    int r0, r1;     // row range [r0, r1)
    int R;          // rows per tile
    int ntiles;
};

__device__ __forceinline__ CtaWork cta_work(const MvParams& p) {
    CtaWork w;
    w.mode_gateup = (p.mode == MV_GATEUP);
    int cta = blockIdx.x, s = 0, c0 = 0, nc = gridDim.x;
    if (!w.mode_gateup) {
        for (s = 0; s < p.nseg - 1; s++) {
            if (cta < p.seg[s].cta0 + p.seg[s].ncta) break;
        }
        c0 = p.seg[s].cta0; nc = p.seg[s].ncta;
    }
    w.seg = s;
    int units = p.seg[s].n_rows >> 1;                    // row pairs
    int i = cta - c0;
    long u0 = (long)units * i / nc, u1 = (long)units * (i + 1) / nc;
    w.r0 = (int)u0 * 2; w.r1 = (int)u1 * 2;
    w.R = p.rows_per_tile;
    int nt = (w.r1 - w.r0 + w.R - 1) / w.R;
    w.ntiles = w.mode_gateup ? 2 * nt : nt;
    return w;
}

__device__ __forceinline__ TileInfo tile_info(const MvParams& p, const CtaWork& w, int t) {
    TileInfo ti;
    int tt = w.mode_gateup ? (t >> 1) : t;
    ti.seg = w.mode_gateup ? (t & 1) : w.seg;
    ti.row0 = w.r0 + tt * w.R;
    ti.nrows = min(w.R, w.r1 - ti.row0);
    const MvSeg& sg = p.seg[ti.seg];
    uintptr_t start = (uintptr_t)sg.w + (uintptr_t)ti.row0 * sg.row_bytes;
    uintptr_t end = start + (uintptr_t)ti.nrows * sg.row_bytes;
    uintptr_t a0 = start & ~(uintptr_t)15, a1 = (end + 15) & ~(uintptr_t)15;
    ti.src = (const uint8_t*)a0;
    ti.bytes = (uint32_t)(a1 - a0);
    ti.off = (uint32_t)(start - a0);
    return ti;
}

// one lane finishes a row pair
__device__ __forceinline__ void epilogue_pair(const MvParams& p, int seg, int row, float v0, float v1) {
    const MvSeg& sg = p.seg[seg];
    if (p.mode == MV_STORE) {
        float* y = p.y + sg.out_off;
        y[row] = v0; y[row + 1] = v1;
    } else if (p.mode == MV_RESID) {
        float* y = p.y + sg.out_off;
        const float* r = p.resid + sg.out_off;
        float a = r[row], b = r[row + 1];
        y[row] = a + v0; y[row + 1] = b + v1;
    } else if (p.mode == MV_QKV) {
        int hd = p.head_dim;
        int pos = *p.pos_ptr;
        if (seg < 2) {                      // Q or K: rotate the adjacent pair (ggml "norm" RoPE)
            int i = (row % hd) >> 1;
            float ang = (float)pos * p.rope_freq[i];
            float s, c;
            sincosf(ang, &s, &c);
            float a = v0 * c - v1 * s, b = v0 * s + v1 * c;
            v0 = a; v1 = b;
        }
        if (seg == 0) {
            p.q_out[row] = v0; p.q_out[row + 1] = v1;
        } else {
            int bt = p.block_tokens;
            int page = p.block_table[pos / bt], slot = pos % bt;
            int kvh = row / hd, d = row % hd;
            size_t idx = (((size_t)page * p.n_kv + kvh) * bt + slot) * hd + d;
            __half* dst = (seg == 1 ? p.k_cache : p.v_cache) + idx;
            *reinterpret_cast<__half2*>(dst) = __floats2half2_rn(v0, v1);
        }
    }
}

template <bool ACT2>
__global__ void __launch_bounds__(NTHREADS, 2) matvec_kernel(const MvParams p) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    MvShared& sh = *reinterpret_cast<MvShared*>(smem_raw);
    uint8_t* ring = smem_raw + ((sizeof(MvShared) + 127) & ~127);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int S = p.n_stages;
    const int Wk = p.warps_per_row, G = NCW / Wk;
    const CtaWork work = cta_work(p);

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; s++) { mbar_init(&sh.full[s], 1); mbar_init(&sh.empty[s], G * Wk); }
        fence_mbar_init();
    }
    __syncthreads();
    griddep_launch();   // let the next kernel start prefetching its own weights

    if (warp == NCW) {
        // ------------------------------------------------------------ producer
        if (lane == 0) {
            for (int t = 0; t < work.ntiles; t++) {
                int s = t % S; uint32_t ph = (t / S) & 1;
                mbar_wait(&sh.empty[s], ph ^ 1);
                TileInfo ti = tile_info(p, work, t);
                mbar_arrive_expect_tx(&sh.full[s], ti.bytes);
                tma_load_1d(ring + (size_t)s * p.stage_bytes, ti.src, ti.bytes, &sh.full[s]);
            }
        }
        return;
    }

    // ---------------------------------------------------------------- consumers
    const int g = warp / Wk, wk = warp % Wk;
    const bool warp_active = g < G;
    const int u = wk * 32 + lane;                       // this lane's 64-element chunk
    const bool lane_active = warp_active && (u * 64 < p.K);

    ActChunk act;
    float ssq = 0.f;
    {
        const bool has_norm = p.norm_w != nullptr;
        griddep_wait();  // activations written by the previous kernel are visible from here on
        if (lane_active) {
            const float4* xp4 = reinterpret_cast<const float4*>(p.x + 64 * u);
            const float4* np4 = reinterpret_cast<const float4*>(p.norm_w + 64 * u);
#pragma unroll
            for (int h = 0; h < 2; h++) {
                float t[32];
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    float4 v = xp4[8 * h + i];
                    ssq += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
                    if (has_norm) { float4 w4 = __ldg(np4 + 8 * h + i); v.x *= w4.x; v.y *= w4.y; v.z *= w4.z; v.w *= w4.w; }
                    t[4 * i] = v.x; t[4 * i + 1] = v.y; t[4 * i + 2] = v.z; t[4 * i + 3] = v.w;
                }
                float lo, hi;
                float amax = quantize32(t, act.a + 8 * h, act.b + 8 * h, lo, hi);
                act.dx[h] = amax / 127.0f;
                act.s16[2 * h] = lo; act.s16[2 * h + 1] = hi;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 16; i++) { act.a[i] = 0; act.b[i] = 0; }
            act.dx[0] = act.dx[1] = 0.f;
            act.s16[0] = act.s16[1] = act.s16[2] = act.s16[3] = 0.f;
        }
        if (has_norm) {
            // rsqrt(mean(x^2) + eps) scales dx and the group sums; the int8 planes do not change
            ssq = warp_sum(ssq);
            if (lane == 0) sh.red[warp] = (g == 0) ? ssq : 0.f;
            named_bar_sync(1, NCW * 32);
            float tot = 0.f;
            for (int i = 0; i < NCW; i++) tot += sh.red[i];
            float rs = rsqrtf(tot / (float)p.K + p.eps);
            act.dx[0] *= rs; act.dx[1] *= rs;
#pragma unroll
            for (int i = 0; i < 4; i++) act.s16[i] *= rs;
        }
    }
    if (!warp_active) return;

    const int bar_id = 2 + g, bar_n = Wk * 32;
    for (int t = 0; t < work.ntiles; t++) {
        int s = t % S; uint32_t ph = (t / S) & 1;
        TileInfo ti = tile_info(p, work, t);
        const MvSeg& sg = p.seg[ti.seg];
        mbar_wait(&sh.full[s], ph);
        const uint8_t* base = ring + (size_t)s * p.stage_bytes + ti.off;
        int npairs = ti.nrows >> 1;
        int pi = 0;
        for (int pp = g; pp < npairs; pp += G, pi++) {
            const uint8_t* r0p = base + (size_t)(2 * pp) * sg.row_bytes;
            float a0 = 0.f, a1 = 0.f;
            if (lane_active) {
                a0 = chunk_dot<ACT2>(sg.type, r0p, u, act);
                a1 = chunk_dot<ACT2>(sg.type, r0p + sg.row_bytes, u, act);
            }
            // 2-row transposed butterfly: lanes 0-15 end with row 0's sum, lanes 16-31 with row 1's
            float send = (lane & 16) ? a0 : a1, keep = (lane & 16) ? a1 : a0;
            keep += __shfl_xor_sync(0xffffffffu, send, 16);
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) keep += __shfl_xor_sync(0xffffffffu, keep, o);
            if ((lane & 15) == 0) sh.part[t & 1][warp][pi][lane >> 4] = keep;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&sh.empty[s]);       // stage can be refilled
        if (Wk > 1) named_bar_sync(bar_id, bar_n); else __syncwarp();
        if (wk == 0 && lane < pi) {
            // lane i finishes pair i of this group
            int pp = g + lane * G;
            float v0 = 0.f, v1 = 0.f;
            for (int w2 = 0; w2 < Wk; w2++) {
                v0 += sh.part[t & 1][g * Wk + w2][lane][0];
                v1 += sh.part[t & 1][g * Wk + w2][lane][1];
            }
            int row = ti.row0 + 2 * pp;
            if (p.mode == MV_GATEUP) {
                if (ti.seg == 0) { sh.gate[g][lane][0] = v0; sh.gate[g][lane][1] = v1; }
                else {
                    float g0 = sh.gate[g][lane][0], g1 = sh.gate[g][lane][1];
                    float* y = p.y;
                    y[row] = g0 / (1.0f + __expf(-g0)) * v0;
                    y[row + 1] = g1 / (1.0f + __expf(-g1)) * v1;
                }
            } else {
                epilogue_pair(p, ti.seg, row, v0, v1);
            }
        }
    }
}

// ---------------------------------------------------------------- host side
cudaError_t mv_init() {   // per device: allow > 48 KB dynamic shared memory
    cudaError_t e = cudaFuncSetAttribute(matvec_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(matvec_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024);
}

int mv_plan(MvParams& p, int n_ctas) {
    // decide tile shape, ring depth and the CTA split between segments
    if (p.K % 256 || p.K <= 0) return -1;
    int chunks = p.K / 64;
    int Wk = (chunks + 31) / 32;
    if (Wk > NCW) return -2;          // K > 16384 needs 2 chunks per lane (not built yet)
    p.warps_per_row = Wk;
    int G = NCW / Wk;
    int max_rb = 0;
    for (int s = 0; s < p.nseg; s++) {
        int be = block_elems(p.seg[s].type), bb = block_bytes(p.seg[s].type);
        if (!bb || be == 1) return -3;
        p.seg[s].row_bytes = p.K / be * bb;
        if (p.seg[s].n_rows % 2) return -4;
        max_rb = p.seg[s].row_bytes > max_rb ? p.seg[s].row_bytes : max_rb;
    }
    const int ring_budget = 104 * 1024;
    int R = 2 * G;                                           // one pair per group
    while (R + 2 * G <= 2 * G * MV_MAX_PAIRS && (long)(R + 2 * G) * max_rb + 32 <= 28 * 1024) R += 2 * G;
    while (R > 2 && (long)R * max_rb + 32 > ring_budget / 2) R -= 2;
    p.rows_per_tile = R;
    p.stage_bytes = (int)((((long)R * max_rb + 32) + 127) / 128 * 128);
    int S = ring_budget / p.stage_bytes;
    if (S > MV_MAX_STAGES) S = MV_MAX_STAGES;
    if (S < 2) return -5;
    p.n_stages = S;
    // CTA split proportional to bytes (gate/up: every CTA handles both matrices)
    if (p.mode == MV_GATEUP) {
        if (p.nseg != 2 || p.seg[0].n_rows != p.seg[1].n_rows) return -6;
        p.seg[0].cta0 = p.seg[1].cta0 = 0; p.seg[0].ncta = p.seg[1].ncta = n_ctas;
    } else {
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

size_t mv_smem_bytes(const MvParams& p) {
    return ((sizeof(MvShared) + 127) & ~127) + (size_t)p.n_stages * p.stage_bytes;
}

cudaError_t mv_launch(const MvParams& p, int n_ctas, cudaStream_t stream, bool act2, bool pdl) {
    size_t smem = mv_smem_bytes(p);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(n_ctas); cfg.blockDim = dim3(NTHREADS);
    cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    if (act2) return cudaLaunchKernelEx(&cfg, matvec_kernel<true>, p);
    return cudaLaunchKernelEx(&cfg, matvec_kernel<false>, p);
}

}  // namespace b2
