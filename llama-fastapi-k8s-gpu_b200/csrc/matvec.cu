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
#include "kernels.h"
#include "ptx.cuh"
#include "quant_blocks.cuh"
#include <cuda_fp16.h>
#include <stdlib.h>

namespace b2 {

#ifndef B2_ROW_UNROLL
#define B2_ROW_UNROLL 1
#endif
constexpr int ROW_UNROLL = B2_ROW_UNROLL;  // rows of a group interleaved in the instruction stream
constexpr int R = MV_ROWS;             // rows per group
constexpr int TPB = 32 / R;            // tiles the producer warp issues per round

template <int NCW>
struct __align__(16) MvShared {
    uint64_t full[MV_MAX_STAGES];
    uint64_t empty[MV_MAX_STAGES];
    float red[NCW];
};

struct CtaWork {
    int seg;            // segment (non gate/up modes)
    int r0, r1;         // row range [r0, r1)
    int ng;             // row groups
    int tpg;            // tiles per group
    int rounds, rem;    // full rounds of NCW groups, groups in the last partial round
    int total;          // tiles for the whole CTA
};

template <int NCW>
__device__ __forceinline__ CtaWork cta_work(const MvParams& p) {
    CtaWork w;
    const bool gu = (p.mode == MV_GATEUP);
    int cta = blockIdx.x, s = 0, c0 = 0, nc = gridDim.x;
    if (!gu) {
        for (s = 0; s < p.nseg - 1; s++)
            if (cta < p.seg[s].cta0 + p.seg[s].ncta) break;
        c0 = p.seg[s].cta0; nc = p.seg[s].ncta;
    }
    w.seg = s;
    int units = p.seg[s].n_rows >> 1;                    // row pairs
    int i = cta - c0;
    long u0 = (long)units * i / nc, u1 = (long)units * (i + 1) / nc;
    w.r0 = (int)u0 * 2; w.r1 = (int)u1 * 2;
    w.ng = (w.r1 - w.r0 + R - 1) / R;
    w.tpg = (p.whole_rows ? 1 : p.n_slices) * (gu ? 2 : 1);
    w.rounds = w.ng / NCW; w.rem = w.ng % NCW;
    w.total = w.ng * w.tpg;
    return w;
}

struct TileId { int warp, gi, seg, slice; };

// q-th tile in issue order -> owner warp, row group, matrix, K-slice
template <int NCW>
__device__ __forceinline__ TileId tile_of(const MvParams& p, const CtaWork& w, int q) {
    TileId t;
    int full_tiles = w.rounds * w.tpg * NCW;
    int lt;
    if (q < full_tiles) { t.warp = q % NCW; lt = q / NCW; }
    else { int qq = q - full_tiles; t.warp = qq % w.rem; lt = w.rounds * w.tpg + qq / w.rem; }
    t.gi = t.warp + NCW * (lt / w.tpg);
    int within = lt % w.tpg;
    const int spt = p.whole_rows ? 1 : p.n_slices;      // ring tiles per (group, matrix)
    t.seg = (p.mode == MV_GATEUP) ? within / spt : w.seg;
    t.slice = within % spt;
    return t;
}
// issue-order index of warp `warp`'s lt-th tile
template <int NCW>
__device__ __forceinline__ int q_of(const CtaWork& w, int warp, int lt) {
    int full_lt = w.rounds * w.tpg;
    if (lt < full_lt) return lt * NCW + warp;
    return full_lt * NCW + (lt - full_lt) * w.rem + warp;
}

// one lane finishes a row pair
__device__ __forceinline__ void epilogue_pair(const MvParams& p, int seg, int row, float v0, float v1) {
    const MvSeg& sg = p.seg[seg];
    if (p.mode == MV_STORE) {
        float* y = p.y + sg.out_off;
        *reinterpret_cast<float2*>(y + row) = make_float2(v0, v1);
    } else if (p.mode == MV_RESID) {
        float* y = p.y + sg.out_off;
        float2 r = *reinterpret_cast<const float2*>(p.resid + sg.out_off + row);
        *reinterpret_cast<float2*>(y + row) = make_float2(r.x + v0, r.y + v1);
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
            *reinterpret_cast<float2*>(p.q_out + row) = make_float2(v0, v1);
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

// accumulate one tile (R row-slices in a ring stage) into acc[R].
// row_lo = low bits of the global address of row0's slice: each bulk copy started at the 16-byte
// boundary below its row-slice, so the data sits (address & 15) bytes into its slot. Only Q6_K
// rows can be off a 16-byte boundary (K % 2048 != 0); every other type has row_bytes % 16 == 0.
template <int T, bool ACT2>
__device__ __forceinline__ void tile_dots(const uint8_t* stage, int slot_bytes, uint32_t row_lo, uint32_t row_bytes,
                                          int nrows, int slice, int lane, const ActView& av, float (&acc)[R],
                                          bool contig = false) {
    LaneAct la;
    load_act<T, ACT2>(la, av, slice * B2_SLICE, lane);
    // Rows are NOT fully unrolled on purpose: the unrolled body (4 x ~100 instructions per quant type)
    // overflowed the instruction cache (31% of stall samples were "no instruction"); latency is
    // hidden by the 16-20 resident warps instead of by interleaving rows inside one warp.
    float v[R];
#pragma unroll
    for (int r = 0; r < R; r++) v[r] = 0.f;
#pragma unroll ROW_UNROLL
    for (int r = 0; r < nrows; r++) {
        uint32_t off = (T == T_Q6_K && !contig) ? ((row_lo + r * row_bytes) & 15u) : 0u;
        float d = row_dot<T, ACT2>(stage + r * slot_bytes + off, lane, la);
#pragma unroll
        for (int k = 0; k < R; k++) if (k == r) v[k] = d;
    }
#pragma unroll
    for (int r = 0; r < R; r++) acc[r] += v[r];
}

// cold path: second matrix of a gate/up pair whose quant family differs from the first
struct Acc4 { float v[R]; };
template <bool ACT2>
__device__ __noinline__ Acc4 tile_dots_any(int type, const uint8_t* stage, int slot_bytes, uint32_t row_lo,
                                           uint32_t row_bytes, int nrows, int slice, int lane, ActView av, Acc4 in,
                                           bool contig = false) {
    float acc[R];
#pragma unroll
    for (int r = 0; r < R; r++) acc[r] = in.v[r];
    switch (type) {
        case T_Q4_K: tile_dots<T_Q4_K, ACT2>(stage, slot_bytes, row_lo, row_bytes, nrows, slice, lane, av, acc, contig); break;
        case T_Q6_K: tile_dots<T_Q6_K, ACT2>(stage, slot_bytes, row_lo, row_bytes, nrows, slice, lane, av, acc, contig); break;
        case T_Q8_0: tile_dots<T_Q8_0, ACT2>(stage, slot_bytes, row_lo, row_bytes, nrows, slice, lane, av, acc, contig); break;
        case T_Q4_0: tile_dots<T_Q4_0, ACT2>(stage, slot_bytes, row_lo, row_bytes, nrows, slice, lane, av, acc, contig); break;
        case T_Q5_K: tile_dots<T_Q5_K, ACT2>(stage, slot_bytes, row_lo, row_bytes, nrows, slice, lane, av, acc, contig); break;
    }
    Acc4 out;
#pragma unroll
    for (int r = 0; r < R; r++) out.v[r] = acc[r];
    return out;
}

// One consumer warp's whole job for a CTA whose (first) matrix has quant type T.
template <int T, bool ACT2, int NCW>
__device__ __forceinline__ void consume(const MvParams& p, const CtaWork& work, MvShared<NCW>& sh, const uint8_t* ring,
                                        const ActView& av0, const ActView& av1, int warp, int lane) {
    const int S = p.n_stages;
    const int my_groups = work.rounds + (warp < work.rem ? 1 : 0);
    const bool gateup = (p.mode == MV_GATEUP);
    const int nmat = gateup ? 2 : 1;
    const int sb = p.slot_bytes;
    int lt = 0;
    for (int g = 0; g < my_groups; g++) {
        const int gi = warp + NCW * g;
        const int row0 = work.r0 + gi * R, nrows = min(R, work.r1 - row0);
        float gate0 = 0.f, gate1 = 0.f;
        for (int m = 0; m < nmat; m++) {
            const int seg = gateup ? m : work.seg;
            const MvSeg& sg = p.seg[seg];
            const int type = sg.type;
            const uint32_t row_bytes = (uint32_t)sg.row_bytes;
            const uint32_t row_lo0 = (uint32_t)(uintptr_t)sg.w + (uint32_t)row0 * row_bytes;   // low address bits suffice
            const uint32_t sbytes = (uint32_t)slice_bytes(type);
            float acc[R];
#pragma unroll
            for (int r = 0; r < R; r++) acc[r] = 0.f;
            if (p.whole_rows) {
                // one ring tile = the R full rows of this group, fetched with a single bulk copy
                const int q = q_of<NCW>(work, warp, lt++);
                const int st = q % S;
                mbar_wait(&sh.full[st], (q / S) & 1, p.err, 0x10002u);
                const uint8_t* stage = ring + (size_t)st * p.stage_bytes + (row_lo0 & 15u);
                for (int sl = 0; sl < p.n_slices; sl++) {
                    if (sl * B2_SLICE + lane * 64 < p.K && !p.skip_compute) {
                        if (type == T) tile_dots<T, ACT2>(stage + sl * sbytes, (int)row_bytes, 0u, row_bytes, nrows, sl, lane, av0, acc, true);
                        else {
                            Acc4 in;
#pragma unroll
                            for (int r = 0; r < R; r++) in.v[r] = acc[r];
                            Acc4 out = tile_dots_any<ACT2>(type, stage + sl * sbytes, (int)row_bytes, 0u, row_bytes, nrows, sl, lane,
                                                           act_layout_of(type) == act_layout_of(T) ? av0 : av1, in, true);
#pragma unroll
                            for (int r = 0; r < R; r++) acc[r] = out.v[r];
                        }
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&sh.empty[st]);
            } else
            for (int sl = 0; sl < p.n_slices; sl++, lt++) {
                const int q = q_of<NCW>(work, warp, lt);
                const int st = q % S;
                mbar_wait(&sh.full[st], (q / S) & 1, p.err, 0x10002u);
                const uint8_t* stage = ring + (size_t)st * p.stage_bytes;
                if (sl * B2_SLICE + lane * 64 < p.K && !p.skip_compute) {
                    const uint32_t row_lo = row_lo0 + (uint32_t)sl * sbytes;
                    if (type == T) tile_dots<T, ACT2>(stage, sb, row_lo, row_bytes, nrows, sl, lane, av0, acc);
                    else {
                        Acc4 in;
#pragma unroll
                        for (int r = 0; r < R; r++) in.v[r] = acc[r];
                        Acc4 out = tile_dots_any<ACT2>(type, stage, sb, row_lo, row_bytes, nrows, sl, lane,
                                                       act_layout_of(type) == act_layout_of(T) ? av0 : av1, in);
#pragma unroll
                        for (int r = 0; r < R; r++) acc[r] = out.v[r];
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&sh.empty[st]);      // stage can be refilled
            }
            // transposed butterfly over the 4 rows: lane 0 / 8 / 16 / 24 end with row 0 / 1 / 2 / 3
            const bool hi16 = lane & 16, hi8 = lane & 8;
            float s0 = hi16 ? acc[0] : acc[2], s1 = hi16 ? acc[1] : acc[3];
            float k0 = hi16 ? acc[2] : acc[0], k1 = hi16 ? acc[3] : acc[1];
            k0 += __shfl_xor_sync(0xffffffffu, s0, 16);
            k1 += __shfl_xor_sync(0xffffffffu, s1, 16);
            float snd = hi8 ? k0 : k1, kp = hi8 ? k1 : k0;
            kp += __shfl_xor_sync(0xffffffffu, snd, 8);
            kp += __shfl_xor_sync(0xffffffffu, kp, 4);
            kp += __shfl_xor_sync(0xffffffffu, kp, 2);
            kp += __shfl_xor_sync(0xffffffffu, kp, 1);
            float odd = __shfl_down_sync(0xffffffffu, kp, 8);   // lane 0 <- row 1, lane 16 <- row 3
            if ((lane & 15) == 0) {
                int pr = lane >> 4;                             // pair 0 or 1 of the group
                if (2 * pr < nrows) {
                    int row = row0 + 2 * pr;
                    if (gateup) {
                        if (m == 0) { gate0 = kp; gate1 = odd; }
                        else {
                            float h0 = gate0 / (1.0f + __expf(-gate0)) * kp;
                            float h1 = gate1 / (1.0f + __expf(-gate1)) * odd;
                            *reinterpret_cast<float2*>(p.y + row) = make_float2(h0, h1);
                        }
                    } else {
                        epilogue_pair(p, seg, row, kp, odd);
                    }
                }
            }
        }
    }
}

template <int NCW, bool ACT2>
__global__ void __launch_bounds__((NCW + (NCW <= 8 ? 1 : MV_MAX_PROD)) * 32, NCW <= 8 ? 2 : 1) matvec_kernel(const MvParams p) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    MvShared<NCW>& sh = *reinterpret_cast<MvShared<NCW>*>(smem_raw);
    const int Kp = p.n_slices * B2_SLICE;                 // planes padded to whole slices
    uint8_t* act_a = smem_raw + MV_SHARED_HDR;
    uint8_t* act_b = act_a + Kp;
    uint8_t* act_a2 = p.act_dual ? act_b + Kp : act_a;    // second layout (gate/up of different quant families)
    uint8_t* act_b2 = p.act_dual ? act_a2 + Kp : act_b;
    float* act_dx = reinterpret_cast<float*>((p.act_dual ? act_b2 : act_b) + Kp);
    float* act_s16 = act_dx + p.K / 32;
    uint8_t* ring = smem_raw + p.ring_off;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int S = p.n_stages;
    const CtaWork work = cta_work<NCW>(p);
    // plane layout follows the quant family this CTA reads (gate/up: seg 0's, plus a second copy if they differ)
    const int type0 = (p.mode == MV_GATEUP) ? p.seg[0].type : p.seg[work.seg].type;
    const int cta_layout = act_layout_of(type0);

    if (threadIdx.x == 0) {
        for (int s = 0; s < S; s++) { mbar_init(&sh.full[s], 1); mbar_init(&sh.empty[s], 1); }
        fence_mbar_init();
    }
    __syncthreads();
    griddep_launch();   // let the next kernel start prefetching its own weights

    if (warp >= NCW) {
        // ------------------------------------------------------------ producer warps (rounds interleaved)
        const int lanes_per_tile = p.whole_rows ? 1 : R;
        const int tsub = p.whole_rows ? lane : lane / R, rsub = p.whole_rows ? 0 : lane % R;
        for (int qb = (warp - NCW) * TPB; qb < work.total; qb += TPB * p.n_prod) {
            int q = qb + tsub;
            bool have = q < work.total && tsub < TPB;
            const uint8_t* src = nullptr; uint32_t bytes = 0; uint8_t* dst = nullptr; int st = 0;
            if (have) {
                TileId t = tile_of<NCW>(p, work, q);
                const MvSeg& sg = p.seg[t.seg];
                int row0 = work.r0 + t.gi * R, nrows = min(R, work.r1 - row0);
                st = q % S;
                if (p.whole_rows) {
                    uintptr_t start = (uintptr_t)sg.w + (size_t)row0 * sg.row_bytes;
                    uintptr_t a0 = start & ~(uintptr_t)15, a1 = (start + (size_t)nrows * sg.row_bytes + 15) & ~(uintptr_t)15;
                    src = (const uint8_t*)a0; bytes = (uint32_t)(a1 - a0);
                    dst = ring + (size_t)st * p.stage_bytes;
                } else if (rsub < nrows) {
                    int sb = slice_bytes(sg.type);
                    int len = min(sb, sg.row_bytes - t.slice * sb);        // last slice may be shorter
                    uintptr_t start = (uintptr_t)sg.w + (size_t)(row0 + rsub) * sg.row_bytes + (size_t)t.slice * sb;
                    uintptr_t a0 = start & ~(uintptr_t)15, a1 = (start + len + 15) & ~(uintptr_t)15;
                    src = (const uint8_t*)a0; bytes = (uint32_t)(a1 - a0);
                    dst = ring + (size_t)st * p.stage_bytes + rsub * p.slot_bytes;
                }
                if (rsub == 0) mbar_wait(&sh.empty[st], ((q / S) & 1) ^ 1, p.err, 0x10001u);
            }
            __syncwarp();
            // total bytes of the tile = sum over its lanes
            uint32_t tot = bytes;
            if (lanes_per_tile > 1) {
#pragma unroll
                for (int o = 1; o < R; o <<= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
            }
            if (have && rsub == 0) mbar_arrive_expect_tx(&sh.full[st], tot);
            __syncwarp();
            if (bytes) tma_load_1d(dst, src, bytes, &sh.full[st]);
        }
        return;
    }

    // ---------------------------------------------------------------- consumers: quantise x
    griddep_wait();     // activations written by the previous kernel are visible from here on
    {
        const bool has_norm = p.norm_w != nullptr;
        const int t = threadIdx.x;
        float ssq = 0.f;
        for (int blk = t; blk < p.K / 32; blk += NCW * 32) {
            const float4* xp4 = reinterpret_cast<const float4*>(p.x + 32 * blk);
            const float4* np4 = reinterpret_cast<const float4*>(p.norm_w + 32 * blk);
            float v[32];
#pragma unroll
            for (int i = 0; i < 8; i++) {
                float4 f = xp4[i];
                ssq += f.x * f.x + f.y * f.y + f.z * f.z + f.w * f.w;
                if (has_norm) { float4 w4 = __ldg(np4 + i); f.x *= w4.x; f.y *= w4.y; f.z *= w4.z; f.w *= w4.w; }
                v[4 * i] = f.x; v[4 * i + 1] = f.y; v[4 * i + 2] = f.z; v[4 * i + 3] = f.w;
            }
            int qa[8], qb[8]; float lo, hi;
            float amax = quantize32(v, qa, qb, lo, hi);
            // two 16-byte units per block, permuted inside the slice for conflict-free lane reads
            const int u0 = 2 * blk, sbase = (u0 >> 7) * B2_SLICE;
#pragma unroll
            for (int h2 = 0; h2 < 2; h2++) {
                const int4 wa = make_int4(qa[4 * h2], qa[4 * h2 + 1], qa[4 * h2 + 2], qa[4 * h2 + 3]);
                const int4 wb = make_int4(qb[4 * h2], qb[4 * h2 + 1], qb[4 * h2 + 2], qb[4 * h2 + 3]);
                const int pos = sbase + 16 * act_unit_pos(cta_layout, (u0 + h2) & 127);
                *reinterpret_cast<int4*>(act_a + pos) = wa;
                *reinterpret_cast<int4*>(act_b + pos) = wb;
                if (p.act_dual) {
                    const int pos2 = sbase + 16 * act_unit_pos(1 - cta_layout, (u0 + h2) & 127);
                    *reinterpret_cast<int4*>(act_a2 + pos2) = wa;
                    *reinterpret_cast<int4*>(act_b2 + pos2) = wb;
                }
            }
            act_dx[blk] = amax / 127.0f;
            act_s16[2 * blk] = lo; act_s16[2 * blk + 1] = hi;
        }
        if (has_norm) {
            // rsqrt(mean(x^2) + eps) scales dx and the group sums; the int8 planes do not change
            ssq = warp_sum(ssq);
            if (lane == 0) sh.red[warp] = ssq;
            named_bar_sync(1, NCW * 32);
            float tot = 0.f;
#pragma unroll
            for (int i = 0; i < NCW; i++) tot += sh.red[i];
            float rs = rsqrtf(tot / (float)p.K + p.eps);
            for (int blk = t; blk < p.K / 32; blk += NCW * 32) {
                act_dx[blk] *= rs; act_s16[2 * blk] *= rs; act_s16[2 * blk + 1] *= rs;
            }
        }
        named_bar_sync(1, NCW * 32);
    }
    ActView av0, av1;
    av0.a = act_a; av0.b = act_b; av0.dx = act_dx; av0.s16 = act_s16;
    av1.a = act_a2; av1.b = act_b2; av1.dx = act_dx; av1.s16 = act_s16;

    // ---------------------------------------------------------------- consumers: stream tiles
    switch (type0) {
        case T_Q4_K: consume<T_Q4_K, ACT2, NCW>(p, work, sh, ring, av0, av1, warp, lane); break;
        case T_Q6_K: consume<T_Q6_K, ACT2, NCW>(p, work, sh, ring, av0, av1, warp, lane); break;
        case T_Q8_0: consume<T_Q8_0, ACT2, NCW>(p, work, sh, ring, av0, av1, warp, lane); break;
        case T_Q4_0: consume<T_Q4_0, ACT2, NCW>(p, work, sh, ring, av0, av1, warp, lane); break;
        case T_Q5_K: consume<T_Q5_K, ACT2, NCW>(p, work, sh, ring, av0, av1, warp, lane); break;
    }
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
