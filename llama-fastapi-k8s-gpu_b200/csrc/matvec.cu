// Decode mat-vec with in-kernel dequantisation: y = W x for GGUF-quantised W (Q4_0, Q8_0, Q4_K,
// Q5_K, Q6_K), weights streamed unchanged from HBM. Replaces ggml-cuda's quantize_q8_1 +
// mul_mat_vec_q + the separate rms_norm / rope / cpy / silu / add launches that the reference
// runs under llm.create_chat_completion (/root/reference/api.py:55-63; SURVEY.md 2.3 K2-K7,K11).
//
// Structure: 148 CTAs of 8 warps and at most 112 KB of shared memory, so that TWO CTAs fit an SM:
// with programmatic dependent launch the CTAs of the next kernel in the stream become resident
// beside the running ones and fill their rings with weights while this kernel still computes --
// HBM keeps streaming across the kernel boundary.
//   * Tiles. A CTA's rows are cut into groups of R=2 consecutive rows (one RoPE pair); a group is
//     one tile when the two rows fit a ring stage (they are contiguous in memory: ONE bulk copy of
//     4-13 KB), else `parts` tiles of a few K-slices each (one copy per row). Measured on B200: a
//     warp issues a cp.async.bulk every ~110 cycles, so copies of 1-2 KB cap a CTA at ~3 TB/s chip
//     rate no matter how deep the ring is; copies of >= 4 KB do not.
//   * Ring. Tile q of the CTA lives in stage q % S. There is no producer warp and there are no
//     "empty" barriers: the warp that has just read stage s refills it itself with tile q + S
//     (self-refill; eight issuers in parallel). An mbarrier only exposes the parity of its phase and
//     consecutive uses of a stage belong to different warps, so a warp that looked at a stage before
//     the previous use had even landed would read "complete" for the wrong phase: the refilling warp
//     therefore publishes `issued[s] = use + 1` (release) after arming the barrier, and a reader
//     waits for that value (acquire; by then the previous use has been consumed, hence landed)
//     before it waits on the parity.
//   * Warps. Warp w owns the row groups g = w, w+8, ... for ALL of K: sums stay in registers over
//     the whole row, one lane reduction (5 shuffles per row pair) per group, no shared-memory
//     atomics. Per slice a lane loads its 64 quantised activations (two int8 planes) once and
//     reuses them for both rows; weights stay packed and go through dp4a.
//   * The first S tiles are requested BEFORE griddepcontrol.wait (weights do not depend on the
//     previous kernel).
//   * prologue (after the wait): x -> optional fused RMSNorm weight -> two int8 planes + packed
//     per-block scales in shared memory; the loads of x are issued in one batch (one L2 round
//     trip), the norm weights were prefetched to L1 before the wait.
//   * epilogue (lane 0 of the owning warp, a row pair): plain store, residual add, RoPE +
//     paged-KV store for the fused QKV projection, or SiLU(gate) * up for the fused FFN projection.
//
// Roofline: HBM. Algorithmic bytes per launch = sum of the weight bytes of its segments.
#include "kernels.h"
#include "ptx.cuh"
#include "quant_blocks.cuh"
#include <cuda_fp16.h>
#include <stdlib.h>

namespace b2 {

#ifndef B2_ROW_UNROLL
#define B2_ROW_UNROLL 2
#endif
constexpr int ROW_UNROLL = B2_ROW_UNROLL;   // rows of a tile interleaved in the instruction stream
constexpr int NCW = MV_NCW;
constexpr int R = MV_ROWS;
static_assert(R == 2, "a row group is one RoPE pair");
constexpr int MV_THREADS = NCW * 32;
constexpr int MV_SMEM_CO = 112 * 1024;      // two CTAs per SM
constexpr int MV_SMEM_SOLO = 224 * 1024;    // shapes whose activations leave too little ring: one CTA per SM
constexpr int MV_TILE_MAX = 14 * 1024;      // largest tile (both rows)

struct __align__(16) MvShared {
    uint64_t full[MV_MAX_STAGES];
    unsigned int issued[MV_MAX_STAGES];     // use number + 1 last armed on each stage
    float red[NCW];
};
static_assert(sizeof(MvShared) <= MV_SHARED_HDR, "header does not fit");

struct CtaWork {
    int seg;            // segment (non gate/up modes)
    int r0, r1;         // row range [r0, r1), both even
    int ng;             // row groups of R rows
    int parts0;         // tiles per group of the first (or only) matrix
    int tpg;            // tiles per group (all matrices)
};

__device__ __forceinline__ CtaWork cta_work(const MvParams& p, int cta, int nctas) {
    CtaWork w;
    const bool gu = (p.mode == MV_GATEUP);
    int s = 0, c0 = 0, nc = nctas;
    if (!gu) {
        for (s = 0; s < p.nseg - 1; s++)
            if (cta < p.seg[s].cta0 + p.seg[s].ncta) break;
        c0 = p.seg[s].cta0; nc = p.seg[s].ncta;
    }
    w.seg = s;
    const int units = p.seg[s].n_rows >> 1;                    // row pairs
    const int i = cta - c0;
    const long u0 = (long)units * i / nc, u1 = (long)units * (i + 1) / nc;
    w.r0 = (int)u0 * 2; w.r1 = (int)u1 * 2;
    w.ng = (w.r1 - w.r0) / R;
    w.parts0 = p.seg[s].parts;
    w.tpg = gu ? p.seg[0].parts + p.seg[1].parts : p.seg[s].parts;
    return w;
}

// lane 0 of the owning warp finishes a row pair
__device__ __forceinline__ void epilogue_pair(const MvParams& p, int seg, int row, float v0, float v1, float2 res) {
    const MvSeg& sg = p.seg[seg];
    if (p.mode == MV_STORE) {
        *reinterpret_cast<float2*>(p.y + sg.out_off + row) = make_float2(v0, v1);
    } else if (p.mode == MV_RESID) {
        *reinterpret_cast<float2*>(p.y + sg.out_off + row) = make_float2(res.x + v0, res.y + v1);
    } else if (p.mode == MV_QKV) {
        const int hd = p.head_dim;
        const int pos = *p.pos_ptr;
        if (seg < 2) {                      // Q or K: rotate the adjacent pair (ggml "norm" RoPE)
            const int i = (row % hd) >> 1;
            const float ang = (float)pos * p.rope_freq[i];
            float s, c;
            sincosf(ang, &s, &c);
            const float a = v0 * c - v1 * s, b = v0 * s + v1 * c;
            v0 = a; v1 = b;
        }
        if (seg == 0) {
            *reinterpret_cast<float2*>(p.q_out + row) = make_float2(v0, v1);
        } else {
            const int bt = p.block_tokens;
            const int page = p.block_table[pos / bt], slot = pos % bt;
            const int kvh = row / hd, d = row % hd;
            const size_t idx = (((size_t)page * p.n_kv + kvh) * bt + slot) * hd + d;
            __half* dst = (seg == 1 ? p.k_cache : p.v_cache) + idx;
            *reinterpret_cast<__half2*>(dst) = __floats2half2_rn(v0, v1);
        }
    }
}

// One K-slice of a tile: both rows against the lane's 64 activations of that slice. row[r] points at the first
// byte of row r's slice in the ring stage (Q6_K rows may sit off a 4-byte boundary: row_dot copes).
template <int T, bool ACT2>
__device__ __forceinline__ void slice_dots(const uint8_t* row0, const uint8_t* row1, int slice, int lane, const ActView& av,
                                           float (&acc)[R]) {
    LaneAct la;
    load_act<T, ACT2>(la, av, slice * B2_SLICE, lane);
#pragma unroll ROW_UNROLL
    for (int r = 0; r < R; r++) {
        const float d = row_dot<T, ACT2>(r ? row1 : row0, lane, la);
#pragma unroll
        for (int k = 0; k < R; k++) if (k == r) acc[k] += d;
    }
}

// cold path: second matrix of a gate/up pair whose quant type differs from the first
struct AccR { float v[R]; };
template <bool ACT2>
__device__ __noinline__ AccR slice_dots_any(int type, const uint8_t* row0, const uint8_t* row1, int slice, int lane, ActView av, AccR in) {
    float acc[R];
#pragma unroll
    for (int r = 0; r < R; r++) acc[r] = in.v[r];
    switch (type) {
        case T_Q4_K: slice_dots<T_Q4_K, ACT2>(row0, row1, slice, lane, av, acc); break;
        case T_Q6_K: slice_dots<T_Q6_K, ACT2>(row0, row1, slice, lane, av, acc); break;
        case T_Q8_0: slice_dots<T_Q8_0, ACT2>(row0, row1, slice, lane, av, acc); break;
        case T_Q4_0: slice_dots<T_Q4_0, ACT2>(row0, row1, slice, lane, av, acc); break;
        case T_Q5_K: slice_dots<T_Q5_K, ACT2>(row0, row1, slice, lane, av, acc); break;
    }
    AccR out;
#pragma unroll
    for (int r = 0; r < R; r++) out.v[r] = acc[r];
    return out;
}

// Where tile t of group g comes from (global) and how it is laid out in its stage.
struct TileGeo {
    int seg, sl0, nsl;          // matrix, first K-slice, number of K-slices
    uint32_t off[R];            // byte offset of row r's first slice inside the stage
};
__device__ __forceinline__ TileGeo tile_geo(const MvParams& p, const CtaWork& w, int g, int t) {
    TileGeo tg;
    const bool second = (p.mode == MV_GATEUP) && t >= w.parts0;
    tg.seg = (p.mode == MV_GATEUP) ? (second ? 1 : 0) : w.seg;
    const int part = second ? t - w.parts0 : t;
    const MvSeg& sg = p.seg[tg.seg];
    tg.sl0 = part * sg.chunk;
    tg.nsl = min(sg.chunk, p.n_slices - tg.sl0);
    const uint32_t row_lo = (uint32_t)(uintptr_t)sg.w + (uint32_t)(w.r0 + g * R) * (uint32_t)sg.row_bytes;   // low address bits of row 0
    if (sg.parts == 1) {        // both rows in one copy that started at the 16-byte boundary below row 0
        tg.off[0] = row_lo & 15u; tg.off[1] = tg.off[0] + (uint32_t)sg.row_bytes;
    } else {                    // one copy per row into its own slot, each from the 16-byte boundary below its start
        const uint32_t s_lo = row_lo + (uint32_t)tg.sl0 * (uint32_t)slice_bytes(sg.type);
        tg.off[0] = s_lo & 15u; tg.off[1] = (uint32_t)p.slot_bytes + ((s_lo + (uint32_t)sg.row_bytes) & 15u);
    }
    return tg;
}

// Request tile t of group g into `stage` (lanes 0..R-1 of the calling warp take part; call with the warp converged).
__device__ __forceinline__ void st_release_shared(unsigned int* p, unsigned int v) {
    asm volatile("st.release.cta.shared::cta.u32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_shared(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.cta.shared::cta.u32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
    return v;
}
__device__ __forceinline__ void tile_issue(const MvParams& p, const CtaWork& w, int g, int t, uint8_t* stage, uint64_t* bar,
                                           unsigned int* issued, unsigned int use, int lane) {
    const bool second = (p.mode == MV_GATEUP) && t >= w.parts0;
    const int seg = (p.mode == MV_GATEUP) ? (second ? 1 : 0) : w.seg;
    const int part = second ? t - w.parts0 : t;
    const MvSeg& sg = p.seg[seg];
    const int row0 = w.r0 + g * R;
    if (sg.parts == 1) {
        if (lane == 0) {
            const uintptr_t start = (uintptr_t)sg.w + (size_t)row0 * sg.row_bytes;
            const uintptr_t a0 = start & ~(uintptr_t)15, a1 = (start + (size_t)R * sg.row_bytes + 15) & ~(uintptr_t)15;
            mbar_arrive_expect_tx(bar, (uint32_t)(a1 - a0));
            st_release_shared(issued, use + 1);
            tma_load_1d(stage, (const void*)a0, (uint32_t)(a1 - a0), bar);
        }
    } else {
        const int sb = slice_bytes(sg.type), sl0 = part * sg.chunk;
        const int len = min(sg.chunk * sb, sg.row_bytes - sl0 * sb);            // the last part may be shorter
        const uintptr_t start = (uintptr_t)sg.w + (size_t)(row0 + (lane & (R - 1))) * sg.row_bytes + (size_t)sl0 * sb;
        const uintptr_t a0 = start & ~(uintptr_t)15, a1 = (start + len + 15) & ~(uintptr_t)15;
        uint32_t bytes = (uint32_t)(a1 - a0), tot = bytes;
#pragma unroll
        for (int o = 1; o < R; o <<= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
        if (lane == 0) { mbar_arrive_expect_tx(bar, tot); st_release_shared(issued, use + 1); }
        __syncwarp();
        if (lane < R) tma_load_1d(stage + lane * p.slot_bytes, (const void*)a0, bytes, bar);
    }
}

// Consumer prologue: x (float32, global) -> optional RMSNorm weight -> two int8 planes + packed per-block
// scales in shared memory, in the layout(s) `layouts` asks for (bit 0: layout 0, bit 1: layout 1). Four
// threads share a 32-element block (8 values each; max and sums combined with two shuffles). The RMS factor
// is a scalar, so it is NOT applied here: the per-warp sums of squares go to sh.red[] and mv_rms_scale()
// turns them into the factor the epilogue multiplies the finished row sums with.
template <bool ACT2>
__device__ __forceinline__ void mv_quantize_input(const MvParams& p, MvShared& sh, uint8_t* smem, int layouts, int warp, int lane) {
    const bool has_norm = p.norm_w != nullptr;
    const int t = warp * 32 + lane, sub = t & 3, quad = t >> 2;
    const int nblk = p.K / 32;                         // a multiple of 8: a warp's 8 quads are in or out together
    const float MAGIC = 12582912.0f;                   // 1.5 * 2^23: float add rounds to nearest integer
    const float fs = ACT2 ? 1.0f / B2_ACT_RES : 1.0f;
    uint8_t* const lay0 = smem + p.act_off[0];
    uint8_t* const lay1 = smem + p.act_off[1];
    constexpr int U = 8;                               // blocks per thread whose loads are in flight together
    constexpr int QUADS = NCW * 8;
    float ssq = 0.f;
    for (int b0 = quad; b0 < nblk; b0 += QUADS * U) {
        float4 xv[U][2];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int blk = b0 + QUADS * u;
            if (blk < nblk) {
                const float4* xp4 = reinterpret_cast<const float4*>(p.x + 32 * blk + 8 * sub);
                xv[u][0] = __ldcg(xp4); xv[u][1] = __ldcg(xp4 + 1);       // written by other SMs (previous kernel)
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int blk = b0 + QUADS * u;
            if (blk >= nblk) continue;
            float v[8] = {xv[u][0].x, xv[u][0].y, xv[u][0].z, xv[u][0].w, xv[u][1].x, xv[u][1].y, xv[u][1].z, xv[u][1].w};
#pragma unroll
            for (int i = 0; i < 8; i++) ssq += v[i] * v[i];
            if (has_norm) {
                const float4* np4 = reinterpret_cast<const float4*>(p.norm_w + 32 * blk + 8 * sub);
                const float4 w0 = __ldg(np4), w1 = __ldg(np4 + 1);
                v[0] *= w0.x; v[1] *= w0.y; v[2] *= w0.z; v[3] *= w0.w; v[4] *= w1.x; v[5] *= w1.y; v[6] *= w1.z; v[7] *= w1.w;
            }
            float amax = 0.f, s8 = 0.f;
#pragma unroll
            for (int i = 0; i < 8; i++) { amax = fmaxf(amax, fabsf(v[i])); s8 += v[i]; }
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
            const float s16 = s8 + __shfl_xor_sync(0xffffffffu, s8, 1);      // sum of this 16-element half
            const float s32 = s16 + __shfl_xor_sync(0xffffffffu, s16, 2);    // sum of the block
            const float id = amax > 0.f ? 127.0f / amax : 0.f, dxl = amax / 127.0f, id2 = id * B2_ACT_RES;
            uint32_t aw[2] = {0u, 0u}, bw[2] = {0u, 0u};
#pragma unroll
            for (int i = 0; i < 8; i++) {                 // same arithmetic as quantize32 (quant_blocks.cuh)
                const float ra = v[i] * id + MAGIC;
                const float af = ra - MAGIC;
                const float r = v[i] - af * dxl;
                const float rb = r * id2 + MAGIC;
                aw[i >> 2] |= (__float_as_uint(ra) & 0xffu) << (8 * (i & 3));
                bw[i >> 2] |= (__float_as_uint(rb) & 0xffu) << (8 * (i & 3));
            }
            // 16-byte unit (2 per block) permuted inside the slice for conflict-free lane reads; 8 bytes per thread
            const int un = 2 * blk + (sub >> 1), sbase = (un >> 7) * B2_SLICE, half = 8 * (sub & 1);
            if (layouts & 1) {
                const int pos = sbase + 16 * act_unit_pos(0, un & 127) + half;
                *reinterpret_cast<uint2*>(lay0 + pos) = make_uint2(aw[0], aw[1]);
                *reinterpret_cast<uint2*>(lay0 + p.act_plane + pos) = make_uint2(bw[0], bw[1]);
                float* f = reinterpret_cast<float*>(lay0 + 2 * p.act_plane);
                if (sub == 0) { f[4 * (blk >> 1) + (blk & 1)] = dxl * fs; f[4 * (blk >> 1) + 2 + (blk & 1)] = s32; }
            }
            if (layouts & 2) {
                const int pos = sbase + 16 * act_unit_pos(1, un & 127) + half;
                *reinterpret_cast<uint2*>(lay1 + pos) = make_uint2(aw[0], aw[1]);
                *reinterpret_cast<uint2*>(lay1 + p.act_plane + pos) = make_uint2(bw[0], bw[1]);
                float* f = reinterpret_cast<float*>(lay1 + 2 * p.act_plane);
                float* g = f + p.K / 16;
                if (sub == 0) f[blk] = dxl * fs;
                if ((sub & 1) == 0) g[act_s16p_index(un)] = 32.0f * s16;
            }
        }
    }
    if (has_norm) {
        ssq = warp_sum(ssq);
        if (lane == 0) sh.red[warp] = ssq;
    }
    named_bar_sync(1, NCW * 32);
}

// rsqrt(mean(x^2) + eps) of the input (1 when there is no fused RMSNorm); call after the prologue
__device__ __forceinline__ float mv_rms_scale(const MvParams& p, const MvShared& sh) {
    if (p.norm_w == nullptr) return 1.0f;
    float tot = 0.f;
#pragma unroll
    for (int i = 0; i < NCW; i++) tot += sh.red[i];
    return rsqrtf(tot / (float)p.K + p.eps);
}

__device__ __forceinline__ ActView act_view(const MvParams& p, uint8_t* smem, int layout) {
    uint8_t* base = smem + p.act_off[layout];
    ActView v;
    v.a = base; v.b = base + p.act_plane;
    v.f = reinterpret_cast<const float*>(base + 2 * p.act_plane);
    v.g = v.f + p.K / 16;
    return v;
}

// sum over the lanes of a row pair: lanes 0-15 end with row 0, lanes 16-31 with row 1; returns both on lane 0
__device__ __forceinline__ void pair_reduce(const float (&acc)[R], int lane, float& v0, float& v1) {
    const bool hi = lane & 16;
    float keep = hi ? acc[1] : acc[0];
    const float send = hi ? acc[0] : acc[1];
    keep += __shfl_xor_sync(0xffffffffu, send, 16);
    keep += __shfl_xor_sync(0xffffffffu, keep, 8);
    keep += __shfl_xor_sync(0xffffffffu, keep, 4);
    keep += __shfl_xor_sync(0xffffffffu, keep, 2);
    keep += __shfl_xor_sync(0xffffffffu, keep, 1);
    v0 = keep;
    v1 = __shfl_sync(0xffffffffu, keep, 16);
}

// One warp: its row groups, every tile of a group in ring order, sums in registers; after reading a stage the
// warp refills it with the tile S places further on in the CTA's tile order.
template <int T, bool ACT2>
__device__ __forceinline__ void consume(const MvParams& p, const CtaWork& work, MvShared& sh, uint8_t* ring,
                                        const ActView& av0, const ActView& av1, float rs, int warp, int lane) {
    const bool gateup = (p.mode == MV_GATEUP);
    const int S = p.n_stages, tpg = work.tpg, ng = work.ng;
    if (warp >= ng) return;
    // ring position of this warp's current tile q = g * tpg + t, and the tile q + S that will replace it
    int s, u, g2, t2;
    { const int q = warp * tpg; s = q % S; u = q / S; g2 = warp + S / tpg; t2 = S % tpg; }
    const int jump = (NCW - 1) * tpg + 1, jump_s = jump % S, jump_u = jump / S;     // last tile of a group -> first of the next
    for (int g = warp; g < ng; g += NCW) {
        const int row0 = work.r0 + g * R;
        float2 res = make_float2(0.f, 0.f);
        if (p.mode == MV_RESID && lane == 0)           // early: the round trip hides behind the group's tiles
            res = __ldcg(reinterpret_cast<const float2*>(p.resid + p.seg[work.seg].out_off + row0));
        float acc[R], gacc[R];
#pragma unroll
        for (int r = 0; r < R; r++) { acc[r] = 0.f; gacc[r] = 0.f; }
        int seg = work.seg;
        for (int t = 0; t < tpg; t++) {
            if (gateup && t == work.parts0) {          // gate done: keep its sums, go on with up
#pragma unroll
                for (int r = 0; r < R; r++) { gacc[r] = acc[r]; acc[r] = 0.f; }
            }
            const TileGeo tg = tile_geo(p, work, g, t);
            seg = tg.seg;
            const int type = p.seg[seg].type;
            const int sb = slice_bytes(type);
            uint8_t* stage = ring + (size_t)s * p.stage_bytes;
            {   // this use of the stage has been armed (see the ring note at the top), then: its bytes have landed
                uint32_t spins = 0;
                while (ld_acquire_shared(&sh.issued[s]) != (unsigned)u + 1u) {
                    __nanosleep(64);
                    if (++spins > (1u << 22)) {
                        volatile unsigned int* e = reinterpret_cast<volatile unsigned int*>(p.err);
                        if (e && e[0] == 0) { e[1] = blockIdx.x; e[2] = threadIdx.x; e[3] = (unsigned)g; e[4] = (unsigned)u; e[0] = 0x10003u; }
                        break;
                    }
                }
            }
            mbar_wait(&sh.full[s], u & 1, p.err, 0x10002u, (unsigned)g);
            const uint8_t* r0p = stage + tg.off[0];
            const uint8_t* r1p = stage + tg.off[1];
            for (int i = 0; i < tg.nsl; i++, r0p += sb, r1p += sb) {
                const int sl = tg.sl0 + i;
#ifndef B2_MV_SKIP_COMPUTE      // experiment builds only (scripts/mv_variants.sh): stream the tiles, skip the dot products
                if (sl * B2_SLICE + lane * 64 < p.K)
#else
                if (false)
#endif
                {
                    if (type == T) slice_dots<T, ACT2>(r0p, r1p, sl, lane, av0, acc);
                    else {
                        AccR in;
#pragma unroll
                        for (int r = 0; r < R; r++) in.v[r] = acc[r];
                        const AccR out = slice_dots_any<ACT2>(type, r0p, r1p, sl, lane,
                                                              act_layout_of(type) == act_layout_of(T) ? av0 : av1, in);
#pragma unroll
                        for (int r = 0; r < R; r++) acc[r] = out.v[r];
                    }
                }
            }
            fence_proxy_async();    // generic-proxy reads of the stage, ordered before the TMA (async proxy) refill
            __syncwarp();
            if (g2 < ng) tile_issue(p, work, g2, t2, stage, &sh.full[s], &sh.issued[s], (unsigned)u + 1u, lane);
            // next tile of this warp (and the tile S places after it)
            if (t + 1 < tpg) { if (++s == S) { s = 0; u++; } }
            else { s += jump_s; u += jump_u; if (s >= S) { s -= S; u++; } g2 += NCW - 1; }
            if (++t2 == tpg) { t2 = 0; g2++; }
        }
        float v0, v1, g0 = 0.f, g1 = 0.f;
        pair_reduce(acc, lane, v0, v1);
        if (gateup) pair_reduce(gacc, lane, g0, g1);
        if (lane == 0) {
            v0 *= rs; v1 *= rs;                            // fused RMSNorm factor (a scalar: applied to the sums)
            if (gateup) {
                g0 *= rs; g1 *= rs;
                const float h0 = g0 / (1.0f + __expf(-g0)) * v0;
                const float h1 = g1 / (1.0f + __expf(-g1)) * v1;
                *reinterpret_cast<float2*>(p.y + row0) = make_float2(h0, h1);
            } else {
                epilogue_pair(p, seg, row0, v0, v1, res);
            }
        }
    }
}

template <bool ACT2>
__global__ void __launch_bounds__(MV_THREADS, 2) matvec_kernel(const MvParams p) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    MvShared& sh = *reinterpret_cast<MvShared*>(smem_raw);
    uint8_t* ring = smem_raw + p.ring_off;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const CtaWork work = cta_work(p, blockIdx.x, gridDim.x);

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.n_stages; s++) { mbar_init(&sh.full[s], 1); sh.issued[s] = 0; }
        fence_mbar_init();
    }
    __syncthreads();
    if (!p.trigger_late) griddep_launch();   // the next kernel's CTAs may move in beside us and prefetch their weights
    {   // first pass over the ring: weights do not depend on the previous kernel, so this runs ahead of the wait
        const int n_first = min(p.n_stages, work.ng * work.tpg);
        int g = 0, t = warp;
        while (t >= work.tpg && work.tpg > 0) { t -= work.tpg; g++; }
        for (int q = warp; q < n_first; q += NCW) {
            tile_issue(p, work, g, t, ring + (size_t)q * p.stage_bytes, &sh.full[q], &sh.issued[q], 0u, lane);
            t += NCW;
            while (t >= work.tpg) { t -= work.tpg; g++; }
        }
    }
    if (p.norm_w) {     // weights, not activations: warm L1 while the previous kernel finishes
        for (int i = threadIdx.x * 32; i < p.K; i += MV_THREADS * 32)
            asm volatile("prefetch.global.L1 [%0];" ::"l"(p.norm_w + i));
    }
    griddep_wait();     // activations written by the previous kernel are visible from here on
    if (p.trigger_late) griddep_launch();
    // plane layout follows the quant type this CTA reads (gate/up: seg 0's, plus the other if the pair is mixed)
    const int type0 = (p.mode == MV_GATEUP) ? p.seg[0].type : p.seg[work.seg].type;
    const int lay = act_layout_of(type0);
    const int layouts = (p.act_dual ? 3 : (1 << lay));
    mv_quantize_input<ACT2>(p, sh, smem_raw, layouts, warp, lane);
    const ActView av0 = act_view(p, smem_raw, lay), av1 = act_view(p, smem_raw, p.act_dual ? 1 - lay : lay);
    const float rs = mv_rms_scale(p, sh);
    switch (type0) {
        case T_Q4_K: consume<T_Q4_K, ACT2>(p, work, sh, ring, av0, av1, rs, warp, lane); break;
        case T_Q6_K: consume<T_Q6_K, ACT2>(p, work, sh, ring, av0, av1, rs, warp, lane); break;
        case T_Q8_0: consume<T_Q8_0, ACT2>(p, work, sh, ring, av0, av1, rs, warp, lane); break;
        case T_Q4_0: consume<T_Q4_0, ACT2>(p, work, sh, ring, av0, av1, rs, warp, lane); break;
        case T_Q5_K: consume<T_Q5_K, ACT2>(p, work, sh, ring, av0, av1, rs, warp, lane); break;
    }
}

// ---------------------------------------------------------------- host side
static int g_stage_cap = 0;      // B2_MV_STAGES (tuning): read once in mv_init
static int g_solo = 0;           // B2_MV_SOLO=1: always one CTA per SM (224 KB of shared memory)
static int g_tile_max = MV_TILE_MAX;   // B2_MV_TILE (tuning): largest tile in bytes

cudaError_t mv_init() {   // per device: allow > 48 KB dynamic shared memory; read the tuning knobs once
    static bool env_read = false;
    if (!env_read) {
        if (const char* s = getenv("B2_MV_STAGES")) g_stage_cap = atoi(s);
        if (const char* s = getenv("B2_MV_SOLO")) g_solo = atoi(s);
        if (const char* s = getenv("B2_MV_TILE")) { int v = atoi(s); if (v >= 2048 && v <= 65536) g_tile_max = v; }
        env_read = true;
    }
    cudaError_t e = cudaFuncSetAttribute(matvec_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, MV_SMEM_SOLO);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(matvec_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, MV_SMEM_SOLO);
}

int mv_plan(MvParams& p, int n_ctas) {
    if (p.K % 256 || p.K <= 0) return -1;
    if (p.K > 65536) return -2;
    p.n_slices = (p.K + B2_SLICE - 1) / B2_SLICE;
    for (int s = 0; s < p.nseg; s++) {
        MvSeg& sg = p.seg[s];
        const int be = block_elems(sg.type), bb = block_bytes(sg.type);
        if (!bb || be == 1) return -3;
        sg.row_bytes = p.K / be * bb;
        if (sg.n_rows % 2) return -4;
    }
    p.act_dual = 0;
    for (int s = 1; s < p.nseg; s++)
        if (p.mode == MV_GATEUP && act_layout_of(p.seg[s].type) != act_layout_of(p.seg[0].type)) p.act_dual = 1;
    // one activation block per layout in use: [plane a][plane b][K/8 floats of packed scales]
    p.act_plane = p.n_slices * B2_SLICE;
    const int act_block = (2 * p.act_plane + p.K / 2 + 127) / 128 * 128;
    p.act_off[0] = MV_SHARED_HDR;
    p.act_off[1] = p.act_dual ? MV_SHARED_HDR + act_block : MV_SHARED_HDR;
    p.ring_off = MV_SHARED_HDR + (p.act_dual ? 2 : 1) * act_block;
    // Tile geometry for a given cap on the tile size. A row pair is one contiguous run of 2 * row_bytes: one tile
    // (ONE copy) if that fits, else the rows are cut into `parts` runs of `chunk` K-slices, one copy per row and run.
    auto geometry = [&](int tile_cap) {
        int max_tile = 0, max_slot = 0;
        for (int s = 0; s < p.nseg; s++) {
            MvSeg& sg = p.seg[s];
            if (R * sg.row_bytes + 32 <= tile_cap) {
                sg.parts = 1; sg.chunk = p.n_slices;
                const int tile = (R * sg.row_bytes + 32 + 15) / 16 * 16;      // + slack: the copy starts / ends on 16-byte boundaries
                max_tile = tile > max_tile ? tile : max_tile;
            } else {
                const int sb = slice_bytes(sg.type);
                int chunk = (tile_cap / R - 32) / sb;
                if (chunk < 1) chunk = 1;
                sg.parts = (p.n_slices + chunk - 1) / chunk;
                sg.chunk = (p.n_slices + sg.parts - 1) / sg.parts;             // even parts
                sg.parts = (p.n_slices + sg.chunk - 1) / sg.chunk;
                const int slot = (sg.chunk * sb + 32 + 15) / 16 * 16;
                max_slot = slot > max_slot ? slot : max_slot;
                max_tile = R * slot > max_tile ? R * slot : max_tile;
            }
        }
        p.slot_bytes = max_slot;
        p.stage_bytes = (max_tile + 127) / 128 * 128;
    };
    // Prefer two CTAs per SM (112 KB each) with at least one stage per warp; shrink the tiles (down to ~4 KB copies)
    // before giving that up; shapes that still leave < 6 stages take the SM alone (224 KB).
    int S = 0;
    if (!g_solo) {
        for (int cap = g_tile_max; cap >= 8192 && S < NCW; cap -= 2048) {
            geometry(cap);
            S = (MV_SMEM_CO - p.ring_off) / p.stage_bytes;
        }
    }
    if (S < 6) {
        geometry(g_tile_max);
        S = (MV_SMEM_SOLO - p.ring_off) / p.stage_bytes;
    }
    if (S > MV_MAX_STAGES) S = MV_MAX_STAGES;
    if (g_stage_cap >= 2 && g_stage_cap < S) S = g_stage_cap;
    if (S < 2) return -5;
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

size_t mv_smem_bytes(const MvParams& p) { return (size_t)p.ring_off + (size_t)p.n_stages * p.stage_bytes; }

cudaError_t mv_launch(const MvParams& p, int n_ctas, cudaStream_t stream, bool act2, bool pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(n_ctas); cfg.blockDim = dim3(MV_THREADS);
    cfg.dynamicSmemBytes = mv_smem_bytes(p); cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    if (act2) return cudaLaunchKernelEx(&cfg, matvec_kernel<true>, p);
    return cudaLaunchKernelEx(&cfg, matvec_kernel<false>, p);
}

}  // namespace b2
