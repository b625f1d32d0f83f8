// Device-side building blocks of the streaming mat-vec (see matvec.cu for the design notes):
// a producer routine that TMA-copies one phase's tiles into the shared-memory ring and a consumer
// routine that quantises the input vector and eats the tiles. matvec.cu wraps them in a one-phase
// kernel; mega.cu chains many phases inside one persistent kernel (ring entries run on across phases).
//
// Ring protocol. Ring entry g (a running number across phases) lives in stage g % S and is use
// number u = g / S of that stage. It is issued by producer warp g % n_prod (S is a multiple of n_prod,
// so a stage always has the same producer, which therefore sees that stage's `empty` phases in
// order) and consumed by consumer warp g % NCW. Different uses of a stage are consumed by different
// warps, and an mbarrier only exposes the parity of its phase, so a consumer that looked at `full`
// before the previous use had even landed would read "complete" for the wrong phase. The producer
// therefore publishes `issued[stage] = u + 1` after arming the barrier; a consumer waits for that
// value first (by then use u-1 has been consumed, hence landed) and only then on the parity.
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
constexpr int TPB = 32 / R;            // tiles a producer warp issues per round
constexpr int MV_GSLOTS = 64;          // row groups whose partial sums can be in flight at once

template <int NCW>
struct __align__(16) MvShared {
    uint64_t full[MV_MAX_STAGES];
    uint64_t empty[MV_MAX_STAGES];
    unsigned int issued[MV_MAX_STAGES];         // use number + 1 last armed on each stage
    float gacc[MV_GSLOTS][2][R];                // per group: row sums (second matrix of gate/up in [1])
    unsigned int gcnt[MV_GSLOTS];               // tiles of the group finished so far
    float red[NCW];
};

struct RingGeom { int slot_bytes, stage_bytes, n_stages; };

struct CtaWork {
    int seg;            // segment (non gate/up modes)
    int r0, r1;         // row range [r0, r1)
    int ng;             // row groups
    int tpg;            // tiles per group (slices x matrices)
    int total;          // tiles for the whole CTA
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
    int units = p.seg[s].n_rows >> 1;                    // row pairs
    int i = cta - c0;
    long u0 = (long)units * i / nc, u1 = (long)units * (i + 1) / nc;
    w.r0 = (int)u0 * 2; w.r1 = (int)u1 * 2;
    w.ng = (w.r1 - w.r0 + R - 1) / R;
    w.tpg = p.n_slices * (gu ? 2 : 1);
    w.total = w.ng * w.tpg;
    return w;
}

// one-time set-up of the ring (thread 0 of the CTA, before the first phase)
template <int NCW>
__device__ __forceinline__ void mv_ring_init(MvShared<NCW>& sh, int n_stages) {
    for (int s = 0; s < n_stages; s++) { mbar_init(&sh.full[s], 1); mbar_init(&sh.empty[s], 1); sh.issued[s] = 0; }
    for (int i = 0; i < MV_GSLOTS; i++) {
        sh.gcnt[i] = 0;
        for (int m = 0; m < 2; m++) for (int r = 0; r < R; r++) sh.gacc[i][m][r] = 0.f;
    }
    fence_mbar_init();
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

// Producer side of one mat-vec phase. Tile q of the phase is ring entry qbase + q; tiles are laid
// out group-major: q -> (group q / tpg, matrix, K-slice). A round is TPB tiles of this producer
// (entries n_prod apart), R lanes per tile (one row-slice copy each).
template <int NCW>
__device__ __forceinline__ void mv_produce(const MvParams& p, const CtaWork& work, MvShared<NCW>& sh, uint8_t* ring,
                                           const RingGeom& geo, uint32_t qbase, int prod_idx, int n_prod, int lane) {
    const uint32_t S = (uint32_t)geo.n_stages;
    const int tsub = lane / R, rsub = lane % R;
    const int first = (int)(((uint32_t)prod_idx + (uint32_t)n_prod - qbase % (uint32_t)n_prod) % (uint32_t)n_prod);
    for (int qb = first; qb < work.total; qb += TPB * n_prod) {
        const int q = qb + tsub * n_prod;
        const bool have = q < work.total;
        const uint8_t* src = nullptr; uint32_t bytes = 0, gq = 0; uint8_t* dst = nullptr; int st = 0;
        if (have) {
            const int gi = q / work.tpg, within = q % work.tpg;
            const int seg = (p.mode == MV_GATEUP) ? within / p.n_slices : work.seg;
            const int slice = within % p.n_slices;
            const MvSeg& sg = p.seg[seg];
            const int row0 = work.r0 + gi * R, nrows = min(R, work.r1 - row0);
            gq = qbase + (uint32_t)q;
            st = (int)(gq % S);
            if (rsub < nrows) {
                const int sb = slice_bytes(sg.type);
                const int len = min(sb, sg.row_bytes - slice * sb);        // last slice may be shorter
                const uintptr_t start = (uintptr_t)sg.w + (size_t)(row0 + rsub) * sg.row_bytes + (size_t)slice * sb;
                const uintptr_t a0 = start & ~(uintptr_t)15, a1 = (start + len + 15) & ~(uintptr_t)15;
                src = (const uint8_t*)a0; bytes = (uint32_t)(a1 - a0);
                dst = ring + (size_t)st * geo.stage_bytes + rsub * geo.slot_bytes;
            }
            if (rsub == 0) mbar_wait(&sh.empty[st], ((gq / S) & 1) ^ 1, p.err, 0x10001u, gq);
        }
        __syncwarp();
        uint32_t tot = bytes;                       // bytes of the tile = sum over its R lanes
#pragma unroll
        for (int o = 1; o < R; o <<= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
        if (have && rsub == 0) {
            mbar_arrive_expect_tx(&sh.full[st], tot);
            *reinterpret_cast<volatile unsigned int*>(&sh.issued[st]) = gq / S + 1;
        }
        __syncwarp();
        if (bytes) tma_load_1d(dst, src, bytes, &sh.full[st]);
    }
}

// Consumer prologue of one phase: x (float32, global) -> optional RMSNorm weight -> two int8 planes +
// per-32 scale + per-16 sums in shared memory. Four threads share a 32-element block (8 values each,
// max and sums combined with two shuffles). The RMS factor is a scalar, so it is NOT applied here: the
// per-warp sums of squares go to sh.red[] and mv_rms_scale() turns them into the factor the epilogue
// multiplies the finished row sums with. All NCW consumer warps take part (named barrier 1).
template <int NCW>
__device__ __forceinline__ void mv_quantize_input(const MvParams& p, MvShared<NCW>& sh, int cta_layout, uint8_t* act_a,
                                                  uint8_t* act_b, uint8_t* act_a2, uint8_t* act_b2, float* act_dx,
                                                  float* act_s16, int warp, int lane) {
    const bool has_norm = p.norm_w != nullptr;
    const int t = warp * 32 + lane, sub = t & 3;
    const float MAGIC = 12582912.0f;                  // 1.5 * 2^23: float add rounds to nearest integer
    float ssq = 0.f;
    for (int blk = t >> 2; blk < p.K / 32; blk += NCW * 8) {
        const float4* xp4 = reinterpret_cast<const float4*>(p.x + 32 * blk + 8 * sub);
        const float4* np4 = reinterpret_cast<const float4*>(p.norm_w + 32 * blk + 8 * sub);
        float v[8];
#pragma unroll
        for (int i = 0; i < 2; i++) {
            float4 f = __ldcg(xp4 + i);               // written by other SMs (previous kernel or previous phase)
            ssq += f.x * f.x + f.y * f.y + f.z * f.z + f.w * f.w;
            if (has_norm) { float4 w4 = __ldg(np4 + i); f.x *= w4.x; f.y *= w4.y; f.z *= w4.z; f.w *= w4.w; }
            v[4 * i] = f.x; v[4 * i + 1] = f.y; v[4 * i + 2] = f.z; v[4 * i + 3] = f.w;
        }
        float amax = 0.f, s8 = 0.f;
#pragma unroll
        for (int i = 0; i < 8; i++) { amax = fmaxf(amax, fabsf(v[i])); s8 += v[i]; }
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
        const float s16 = s8 + __shfl_xor_sync(0xffffffffu, s8, 1);      // sum of this 16-element half
        const float id = amax > 0.f ? 127.0f / amax : 0.f, dxl = amax / 127.0f, id2 = id * B2_ACT_RES;
        uint32_t aw[2] = {0u, 0u}, bw[2] = {0u, 0u};
#pragma unroll
        for (int i = 0; i < 8; i++) {                 // same arithmetic as quantize32 (quant_blocks.cuh)
            float ra = v[i] * id + MAGIC;
            float af = ra - MAGIC;
            float r = v[i] - af * dxl;
            float rb = r * id2 + MAGIC;
            aw[i >> 2] |= (__float_as_uint(ra) & 0xffu) << (8 * (i & 3));
            bw[i >> 2] |= (__float_as_uint(rb) & 0xffu) << (8 * (i & 3));
        }
        // 16-byte unit (2 per block) permuted inside the slice for conflict-free lane reads; 8 bytes per thread
        const int u = 2 * blk + (sub >> 1), sbase = (u >> 7) * B2_SLICE, half = 8 * (sub & 1);
        const int pos = sbase + 16 * act_unit_pos(cta_layout, u & 127) + half;
        *reinterpret_cast<uint2*>(act_a + pos) = make_uint2(aw[0], aw[1]);
        *reinterpret_cast<uint2*>(act_b + pos) = make_uint2(bw[0], bw[1]);
        if (p.act_dual) {
            const int pos2 = sbase + 16 * act_unit_pos(1 - cta_layout, u & 127) + half;
            *reinterpret_cast<uint2*>(act_a2 + pos2) = make_uint2(aw[0], aw[1]);
            *reinterpret_cast<uint2*>(act_b2 + pos2) = make_uint2(bw[0], bw[1]);
        }
        if (sub == 0) act_dx[blk] = dxl;
        if ((sub & 1) == 0) act_s16[2 * blk + (sub >> 1)] = s16;
    }
    if (has_norm) {
        ssq = warp_sum(ssq);
        if (lane == 0) sh.red[warp] = ssq;
    }
    named_bar_sync(1, NCW * 32);
}

// rsqrt(mean(x^2) + eps) of the phase's input (1 when the phase has no fused RMSNorm); call after the prologue
template <int NCW>
__device__ __forceinline__ float mv_rms_scale(const MvParams& p, const MvShared<NCW>& sh) {
    if (p.norm_w == nullptr) return 1.0f;
    float tot = 0.f;
#pragma unroll
    for (int i = 0; i < NCW; i++) tot += sh.red[i];
    return rsqrtf(tot / (float)p.K + p.eps);
}

// One consumer warp's share of a phase: every NCW-th ring entry. A tile is reduced over the lanes
// straight away; groups that span several tiles (K > 2048, or gate/up) collect their row sums in
// shared memory with atomics and the warp that brings in the last tile runs the epilogue.
template <int T, bool ACT2, int NCW>
__device__ __forceinline__ void consume(const MvParams& p, const CtaWork& work, MvShared<NCW>& sh, const uint8_t* ring,
                                        const RingGeom& geo, uint32_t qbase, const ActView& av0, const ActView& av1,
                                        float rs, int warp, int lane) {
    const uint32_t S = (uint32_t)geo.n_stages;
    const bool gateup = (p.mode == MV_GATEUP);
    const int sb = geo.slot_bytes;
    const int tpg = work.tpg;
    const int first = (int)(((uint32_t)warp + (uint32_t)NCW - qbase % (uint32_t)NCW) % (uint32_t)NCW);
    // tile bookkeeping advances by NCW per iteration: carried incrementally (integer division by a run-time
    // value is ~20 instructions, and there were six of them per tile)
    if (first >= work.total || tpg <= 0) return;     // no tile for this warp (tpg is 0 in CTAs without rows)
    int gi = first / tpg, within = first % tpg;
    uint32_t st = (qbase + (uint32_t)first) % S, use = (qbase + (uint32_t)first) / S;
    const int n_sl = p.n_slices;
    for (int q = first; q < work.total; q += NCW) {
        const int m = (gateup && within >= n_sl) ? 1 : 0;
        const int seg = gateup ? m : work.seg;
        const int sl = within - m * n_sl;
        const MvSeg& sg = p.seg[seg];
        const int type = sg.type;
        const int row0 = work.r0 + gi * R, nrows = min(R, work.r1 - row0);
        const uint32_t row_bytes = (uint32_t)sg.row_bytes;
        const uint32_t row_lo = (uint32_t)(uintptr_t)sg.w + (uint32_t)row0 * row_bytes + (uint32_t)sl * (uint32_t)slice_bytes(type);
        const uint32_t g = qbase + (uint32_t)q;
        const int gi_cur = gi;
        // next iteration's values
        within += NCW; while (within >= tpg) { within -= tpg; gi++; }
        const uint32_t st_cur = st, use_cur = use;
        st += NCW; while (st >= S) { st -= S; use++; }
        {   // the producer has armed this use of the stage (see the protocol note at the top)
            uint32_t spins = 0;
            while (*reinterpret_cast<volatile unsigned int*>(&sh.issued[st_cur]) != use_cur + 1) {
                if (++spins > (1u << 26)) {
                    volatile unsigned int* e = reinterpret_cast<volatile unsigned int*>(p.err);
                    if (e && e[0] == 0) { e[1] = blockIdx.x; e[2] = threadIdx.x; e[3] = g; e[4] = use_cur; e[0] = 0x10003u; }
                    break;
                }
            }
        }
        mbar_wait(&sh.full[st_cur], use_cur & 1, p.err, 0x10002u, g);
        const uint8_t* stage = ring + (size_t)st_cur * geo.stage_bytes;
        float acc[R];
#pragma unroll
        for (int r = 0; r < R; r++) acc[r] = 0.f;
        if (sl * B2_SLICE + lane * 64 < p.K && !p.skip_compute) {
            if (type == T) tile_dots<T, ACT2>(stage, sb, row_lo, row_bytes, nrows, sl, lane, av0, acc);
            else {
                Acc4 in;
#pragma unroll
                for (int r = 0; r < R; r++) in.v[r] = 0.f;
                Acc4 out = tile_dots_any<ACT2>(type, stage, sb, row_lo, row_bytes, nrows, sl, lane,
                                               act_layout_of(type) == act_layout_of(T) ? av0 : av1, in);
#pragma unroll
                for (int r = 0; r < R; r++) acc[r] = out.v[r];
            }
        }
        fence_proxy_async();    // generic-proxy reads of the stage, ordered before the TMA (async proxy) refill
        __syncwarp();
        if (lane == 0) mbar_arrive(&sh.empty[st_cur]);      // stage can be refilled
        // transposed butterfly over the 4 rows: lanes 0-7 / 8-15 / 16-23 / 24-31 end with row 0 / 1 / 2 / 3
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
        float v0, v1, g0 = 0.f, g1 = 0.f;                   // lanes 0 / 16 finish pair 0 / 1 of the group
        bool finish = true;
        if (tpg > 1) {
            const int slot = gi_cur % MV_GSLOTS;
            if ((lane & 7) == 0) atomicAdd(&sh.gacc[slot][m][lane >> 3], kp);
            __syncwarp();
            unsigned int old = 0;
            if (lane == 0) { __threadfence_block(); old = atomicAdd(&sh.gcnt[slot], 1u); }
            old = __shfl_sync(0xffffffffu, old, 0);
            finish = (old == (unsigned)tpg - 1);            // this was the group's last tile
            if (finish) {
                __threadfence_block();
                const int pr = lane >> 4;
                volatile float* ga = &sh.gacc[slot][0][0];
                v0 = ga[2 * pr]; v1 = ga[2 * pr + 1];
                if (gateup) { g0 = v0; g1 = v1; v0 = ga[R + 2 * pr]; v1 = ga[R + 2 * pr + 1]; }
                __syncwarp();
                if (lane < 2 * R) ga[lane] = 0.f;           // both matrices' sums: ready for group gi + MV_GSLOTS
                if (lane == 0) *reinterpret_cast<volatile unsigned int*>(&sh.gcnt[slot]) = 0u;
            }
        } else {
            v0 = kp;
            v1 = __shfl_down_sync(0xffffffffu, kp, 8);      // lane 0 <- row 1, lane 16 <- row 3
        }
        if (finish && (lane & 15) == 0) {
            const int pr = lane >> 4;
            if (2 * pr < nrows) {
                const int row = row0 + 2 * pr;
                v0 *= rs; v1 *= rs; g0 *= rs; g1 *= rs;         // fused RMSNorm factor (a scalar: applied to the sums)
                if (gateup) {
                    float h0 = g0 / (1.0f + __expf(-g0)) * v0;
                    float h1 = g1 / (1.0f + __expf(-g1)) * v1;
                    *reinterpret_cast<float2*>(p.y + row) = make_float2(h0, h1);
                } else {
                    epilogue_pair(p, seg, row, v0, v1);
                }
            }
        }
    }
}

// Carve the activation planes out of shared memory and run one whole consumer phase.
template <int NCW, bool ACT2>
__device__ __forceinline__ void mv_consume_phase(const MvParams& p, const CtaWork& work, MvShared<NCW>& sh, uint8_t* act_base,
                                                 const uint8_t* ring, const RingGeom& geo, uint32_t qbase, int warp, int lane,
                                                 unsigned long long* stamp = nullptr) {
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
    if (stamp) { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); *stamp = t; }   // prologue done
    ActView av0, av1;
    av0.a = act_a; av0.b = act_b; av0.dx = act_dx; av0.s16 = act_s16;
    av1.a = act_a2; av1.b = act_b2; av1.dx = act_dx; av1.s16 = act_s16;
    const float rs = mv_rms_scale<NCW>(p, sh);
    switch (type0) {
        case T_Q4_K: consume<T_Q4_K, ACT2, NCW>(p, work, sh, ring, geo, qbase, av0, av1, rs, warp, lane); break;
        case T_Q6_K: consume<T_Q6_K, ACT2, NCW>(p, work, sh, ring, geo, qbase, av0, av1, rs, warp, lane); break;
        case T_Q8_0: consume<T_Q8_0, ACT2, NCW>(p, work, sh, ring, geo, qbase, av0, av1, rs, warp, lane); break;
        case T_Q4_0: consume<T_Q4_0, ACT2, NCW>(p, work, sh, ring, geo, qbase, av0, av1, rs, warp, lane); break;
        case T_Q5_K: consume<T_Q5_K, ACT2, NCW>(p, work, sh, ring, geo, qbase, av0, av1, rs, warp, lane); break;
    }
}

}  // namespace b2
