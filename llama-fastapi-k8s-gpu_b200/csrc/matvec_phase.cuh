// Device-side building blocks of the streaming mat-vec (see matvec.cu for the design notes):
// a producer routine that TMA-copies one phase's tiles into the shared-memory ring and a consumer
// routine that quantises the input vector and eats the tiles. matvec.cu wraps them in a one-phase
// kernel; mega.cu chains many phases inside one persistent kernel (ring indices run on across phases).
#pragma once
#include "kernels.h"
#include "ptx.cuh"
#include "quant_blocks.cuh"
#include <cuda_fp16.h>

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

struct RingGeom { int slot_bytes, stage_bytes, n_stages; };

struct CtaWork {
    int seg;            // segment (non gate/up modes)
    int r0, r1;         // row range [r0, r1)
    int ng;             // row groups
    int tpg;            // tiles per group
    int rounds, rem;    // full rounds of NCW groups, groups in the last partial round
    int total;          // tiles for the whole CTA
};

template <int NCW>
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
        float2 r = __ldcg(reinterpret_cast<const float2*>(p.resid + sg.out_off + row));   // written by other SMs
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
                                        const RingGeom& geo, uint32_t qbase, const ActView& av0, const ActView& av1,
                                        int warp, int lane) {
    const uint32_t S = (uint32_t)geo.n_stages;
    const int my_groups = work.rounds + (warp < work.rem ? 1 : 0);
    const bool gateup = (p.mode == MV_GATEUP);
    const int nmat = gateup ? 2 : 1;
    const int sb = geo.slot_bytes;
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
                const uint32_t q = qbase + (uint32_t)q_of<NCW>(work, warp, lt++);
                const uint32_t st = q % S;
                mbar_wait(&sh.full[st], (q / S) & 1, p.err, 0x10002u, q);
                const uint8_t* stage = ring + (size_t)st * geo.stage_bytes + (row_lo0 & 15u);
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
                const uint32_t q = qbase + (uint32_t)q_of<NCW>(work, warp, lt);
                const uint32_t st = q % S;
                mbar_wait(&sh.full[st], (q / S) & 1, p.err, 0x10002u, q);
                const uint8_t* stage = ring + (size_t)st * geo.stage_bytes;
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

// Producer side of one mat-vec phase: issue this CTA's tiles, TPB per round, tiles interleaved
// between the n_prod producer warps. Tile q of the phase is ring entry qbase + q.
template <int NCW>
__device__ __forceinline__ void mv_produce(const MvParams& p, const CtaWork& work, MvShared<NCW>& sh, uint8_t* ring,
                                           const RingGeom& geo, uint32_t qbase, int prod_idx, int n_prod, int lane) {
    const uint32_t S = (uint32_t)geo.n_stages;
        const int lanes_per_tile = p.whole_rows ? 1 : R;
        const int tsub = p.whole_rows ? lane : lane / R, rsub = p.whole_rows ? 0 : lane % R;
        // Ring entry g = qbase + q is always handled by producer g % n_prod. With n_stages a multiple of
        // n_prod every stage then has exactly one producer, which sees all of its barrier phases in
        // program order (two producers sharing a stage could mistake an old phase of the same parity
        // for the one they wait for).
        const int first = (int)(((uint32_t)prod_idx + (uint32_t)n_prod - qbase % (uint32_t)n_prod) % (uint32_t)n_prod);
        for (int qb = first; qb < work.total; qb += TPB * n_prod) {
            int q = qb + tsub * n_prod;
            bool have = q < work.total && tsub < TPB;
            const uint8_t* src = nullptr; uint32_t bytes = 0, gq = 0; uint8_t* dst = nullptr; int st = 0;
            if (have) {
                TileId t = tile_of<NCW>(p, work, q);
                const MvSeg& sg = p.seg[t.seg];
                int row0 = work.r0 + t.gi * R, nrows = min(R, work.r1 - row0);
                gq = qbase + (uint32_t)q;
                st = (int)(gq % S);
                if (p.whole_rows) {
                    uintptr_t start = (uintptr_t)sg.w + (size_t)row0 * sg.row_bytes;
                    uintptr_t a0 = start & ~(uintptr_t)15, a1 = (start + (size_t)nrows * sg.row_bytes + 15) & ~(uintptr_t)15;
                    src = (const uint8_t*)a0; bytes = (uint32_t)(a1 - a0);
                    dst = ring + (size_t)st * geo.stage_bytes;
                } else if (rsub < nrows) {
                    int sb = slice_bytes(sg.type);
                    int len = min(sb, sg.row_bytes - t.slice * sb);        // last slice may be shorter
                    uintptr_t start = (uintptr_t)sg.w + (size_t)(row0 + rsub) * sg.row_bytes + (size_t)t.slice * sb;
                    uintptr_t a0 = start & ~(uintptr_t)15, a1 = (start + len + 15) & ~(uintptr_t)15;
                    src = (const uint8_t*)a0; bytes = (uint32_t)(a1 - a0);
                    dst = ring + (size_t)st * geo.stage_bytes + rsub * geo.slot_bytes;
                }
                if (rsub == 0) mbar_wait(&sh.empty[st], ((gq / S) & 1) ^ 1, p.err, 0x10001u, gq);
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
}

// Consumer prologue of one phase: x (float32, global) -> optional RMSNorm weight -> two int8 planes +
// per-32 scale + per-16 sums in shared memory. All NCW consumer warps take part (named barrier 1).
template <int NCW>
__device__ __forceinline__ void mv_quantize_input(const MvParams& p, MvShared<NCW>& sh, int cta_layout, uint8_t* act_a,
                                                  uint8_t* act_b, uint8_t* act_a2, uint8_t* act_b2, float* act_dx,
                                                  float* act_s16, int warp, int lane) {
    {
        const bool has_norm = p.norm_w != nullptr;
        const int t = warp * 32 + lane;
        float ssq = 0.f;
        for (int blk = t; blk < p.K / 32; blk += NCW * 32) {
            const float4* xp4 = reinterpret_cast<const float4*>(p.x + 32 * blk);
            const float4* np4 = reinterpret_cast<const float4*>(p.norm_w + 32 * blk);
            float v[32];
#pragma unroll
            for (int i = 0; i < 8; i++) {
                float4 f = __ldcg(xp4 + i);      // written by other SMs (previous kernel or previous phase)
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
}

// Carve the activation planes out of shared memory and run one whole consumer phase.
template <int NCW, bool ACT2>
__device__ __forceinline__ void mv_consume_phase(const MvParams& p, const CtaWork& work, MvShared<NCW>& sh, uint8_t* act_base,
                                                 const uint8_t* ring, const RingGeom& geo, uint32_t qbase, int warp, int lane) {
    const int Kp = p.n_slices * B2_SLICE;                 // planes padded to whole slices
    uint8_t* act_a = act_base;
    uint8_t* act_b = act_a + Kp;
    uint8_t* act_a2 = p.act_dual ? act_b + Kp : act_a;    // second layout (gate/up of different quant families)
    uint8_t* act_b2 = p.act_dual ? act_a2 + Kp : act_b;
    float* act_dx = reinterpret_cast<float*>((p.act_dual ? act_b2 : act_b) + Kp);
    float* act_s16 = act_dx + p.K / 32;
    // plane layout follows the quant family this CTA reads (gate/up: seg 0's, plus a second copy if they differ)
    const int type0 = (p.mode == MV_GATEUP) ? p.seg[0].type : p.seg[work.seg].type;
    const int cta_layout = act_layout_of(type0);
    mv_quantize_input<NCW>(p, sh, cta_layout, act_a, act_b, act_a2, act_b2, act_dx, act_s16, warp, lane);
    ActView av0, av1;
    av0.a = act_a; av0.b = act_b; av0.dx = act_dx; av0.s16 = act_s16;
    av1.a = act_a2; av1.b = act_b2; av1.dx = act_dx; av1.s16 = act_s16;
    switch (type0) {
        case T_Q4_K: consume<T_Q4_K, ACT2, NCW>(p, work, sh, ring, geo, qbase, av0, av1, warp, lane); break;
        case T_Q6_K: consume<T_Q6_K, ACT2, NCW>(p, work, sh, ring, geo, qbase, av0, av1, warp, lane); break;
        case T_Q8_0: consume<T_Q8_0, ACT2, NCW>(p, work, sh, ring, geo, qbase, av0, av1, warp, lane); break;
        case T_Q4_0: consume<T_Q4_0, ACT2, NCW>(p, work, sh, ring, geo, qbase, av0, av1, warp, lane); break;
        case T_Q5_K: consume<T_Q5_K, ACT2, NCW>(p, work, sh, ring, geo, qbase, av0, av1, warp, lane); break;
    }
}

}  // namespace b2
