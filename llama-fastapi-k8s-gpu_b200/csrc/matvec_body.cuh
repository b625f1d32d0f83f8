// Decode mat-vec with in-kernel dequantisation: y = W x for GGUF-quantised W (Q4_0, Q8_0, Q4_K,
// Q5_K, Q6_K), weights streamed unchanged from HBM. Replaces ggml-cuda's quantize_q8_1 +
// mul_mat_vec_q + the separate rms_norm / rope / cpy / silu / add launches that the reference
// runs under llm.create_chat_completion (/root/reference/api.py:55-63; SURVEY.md 2.3 K2-K7,K11).
//
// Structure: one CTA per SM, 19 consumer warps + 1 TMA producer warp (20 warps x 96 registers fill the register file),
// the SM's shared memory as a ring of tiles.
//   * Tiles. A tile = one K-slice (2048 elements) of a group of R = 4 consecutive rows, brought into one ring stage by ONE
//     tensor copy (cp.async.bulk.tensor, box = slice bytes x 4 rows of a tensor map over the raw rows as 8-byte elements;
//     Q8_0, whose slice is wider than a box, as a 3-D copy of two half-slices per row) + mbarrier complete_tx; rows land
//     densely. Launches whose rows or slices are off a 16-byte boundary (Q6_K / Q2_K / Q3_K with K % 2048 != 0, tiny K)
//     use four 1-D bulk copies per tile instead, into slots with room for the misalignment.
//     The producer never touches activations, so with programmatic dependent launch it starts on the weights while the
//     previous kernel is still running, and it keeps the ring full independently of how the consumers are doing. Its issue
//     rate matters: with 1-D copies (4 per tile) one producer warp could not keep up (2.02 ms/token against 1.87 with two),
//     with tensor copies one is enough and the freed warp computes (measured per token: 16+2 warps 2.01 ms, 17+3 2.00,
//     16+4 2.05, 18+2 1.87, 19+1 with tensor copies 1.82).
//   * Consumers. Tile q belongs to consumer warp q % 19, whatever row group and K-slice it is: the warps of a
//     CTA stay within one tile of each other to the end (the kernel is one link of a chain of ~160 dependent
//     launches per token, so the last warp of the last CTA is what the next kernel waits for). Per tile a lane
//     loads its 64 quantised activations (two int8 planes + packed scales) from shared memory and reuses them
//     for the 4 rows; weights stay packed and go through dp4a. A warp only ever waits on its own tile's
//     mbarrier: no block-wide barrier in the main loop.
//     (A round-2 rewrite with warp teams owning K-slices, register-resident activations and last-reader refill
//     was faster in a back-to-back micro-benchmark and slower inside the decode step -- 2.30 against 2.10 ms per
//     token on the same box -- because refills depended on the slowest reader and teams finished unevenly.)
//   * Row sums. A tile's 4 row sums are reduced over the lanes (transposed butterfly). Groups that span several
//     tiles (K > 2048, gate/up) park them in a shared-memory slot; the warp that brings the last tile adds the
//     partial sums in tile order -- the result does not depend on timing -- and runs the epilogue.
//   * An mbarrier only exposes the parity of its phase and consecutive uses of a stage belong to different
//     warps, so a consumer that looked at `full` before the previous use had even landed would read "complete"
//     for the wrong phase: the producer publishes `issued[stage] = use + 1` after arming the barrier and a
//     consumer waits for that value (by then the previous use has been consumed, hence landed) first.
//   * prologue (consumers, after griddepcontrol.wait): x -> optional fused RMSNorm weight -> two int8 planes +
//     packed per-block scales in shared memory (loads of x in one batch: one L2 round trip).
//   * epilogue (lane 0 of the finishing warp): plain store, residual add, RoPE (cos/sin table built once per
//     CTA) + paged-KV store for the fused QKV projection, or SiLU(gate) * up for the fused FFN projection.
//
// Roofline: HBM. Algorithmic bytes per launch = sum of the weight bytes of its segments.
#pragma once
#include "kernels.h"
#include "ptx.cuh"
#include "quant_blocks.cuh"
#include <cuda_fp16.h>
#include <stdlib.h>

namespace b2 {

#ifndef B2_ROW_UNROLL
#define B2_ROW_UNROLL 1
#endif
constexpr int ROW_UNROLL = B2_ROW_UNROLL;   // rows of a tile interleaved in the instruction stream (more overflowed the instruction cache)
constexpr int NCW = MV_NCW;                 // consumer warps
#ifndef B2_MV_NPROD
#define B2_MV_NPROD 1
#endif
constexpr int NPROD = B2_MV_NPROD;          // producer warps
constexpr int R = MV_ROWS;                  // rows per group
constexpr int TPB = 32 / R;                 // tiles a producer warp issues per round
static_assert(R == 4, "the lane reduction and the epilogue are written for groups of two RoPE pairs");
constexpr int MV_THREADS = (NCW + NPROD) * 32;
constexpr int MV_SMEM_MAX = 224 * 1024;
constexpr int MV_GSLOTS = 32;               // row groups whose partial sums can be in flight at once

struct __align__(16) MvShared {
    uint64_t full[MV_MAX_STAGES];           // a tile's bytes have landed (expect_tx)
    uint64_t empty[MV_MAX_STAGES];          // the tile has been read (one arrival: its consumer warp)
    unsigned int issued[MV_MAX_STAGES];     // use number + 1 last armed on each stage
    unsigned int gcnt[MV_GSLOTS];           // tiles of the slot's group handed in so far
    unsigned int gen[MV_GSLOTS];            // slot i serves group g with g % MV_GSLOTS == i and g / MV_GSLOTS == gen[i]
    float red[NCW];
    float2 rope[64];                        // QKV launches: (cos, sin) of pos * freq[i], i < head_dim / 2
};
static_assert(sizeof(MvShared) <= MV_SHARED_HDR, "header does not fit");

struct CtaWork {
    int seg;            // segment (non gate/up modes)
    int r0, r1;         // row range [r0, r1), both even
    int ng;             // row groups (the last one may hold a single pair)
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
    const int units = p.seg[s].n_rows >> 1;                    // row pairs
    const int i = cta - c0;
    const long u0 = (long)units * i / nc, u1 = (long)units * (i + 1) / nc;
    w.r0 = (int)u0 * 2; w.r1 = (int)u1 * 2;
    w.ng = (w.r1 - w.r0 + R - 1) / R;
    w.tpg = p.n_slices * (gu ? 2 : 1);
    w.total = w.ng * w.tpg;
    return w;
}

// lane 0 of the finishing warp completes a row pair
// (pos, page: position of the token and its KV page, fetched once per thread at kernel start; res: the residual pair,
// fetched when the warp started on the row pair -- no global round trip is left on this path)
__device__ __forceinline__ void epilogue_pair(const MvParams& p, const float2* rope, int seg, int row, float v0, float v1, int pos, int page, float2 res) {
    const MvSeg& sg = p.seg[seg];
    if (p.mode == MV_STORE) {
        *reinterpret_cast<float2*>(p.y + sg.out_off + row) = make_float2(v0, v1);
    } else if (p.mode == MV_RESID) {
        *reinterpret_cast<float2*>(p.y + sg.out_off + row) = make_float2(res.x + v0, res.y + v1);
    } else if (p.mode == MV_QKV) {
        const int hd = p.head_dim;
        if (seg < 2) {                      // Q or K: rotate the adjacent pair (ggml "norm" RoPE)
            const float2 cs = rope[(row % hd) >> 1];
            const float c = cs.x, s = cs.y;
            const float a = v0 * c - v1 * s, b = v0 * s + v1 * c;
            v0 = a; v1 = b;
        }
        if (seg == 0) {
            *reinterpret_cast<float2*>(p.q_out + row) = make_float2(v0, v1);
        } else {
            const int bt = p.block_tokens;
            const int slot = pos % bt;
            const int kvh = row / hd, d = row % hd;
            const size_t idx = (((size_t)page * p.n_kv + kvh) * bt + slot) * hd + d;
            __half* dst = (seg == 1 ? p.k_cache : p.v_cache) + idx;
            *reinterpret_cast<__half2*>(dst) = __floats2half2_rn(v0, v1);
        }
    }
}

// ---------------------------------------------------------------- small PTX helpers of the ring protocol
__device__ __forceinline__ void st_release_shared(unsigned int* p, unsigned int v) {
    asm volatile("st.release.cta.shared::cta.u32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_volatile_shared(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.volatile.shared::cta.u32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_shared(unsigned int* p, unsigned int v) {
    asm volatile("st.volatile.shared::cta.u32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}

// Producer warp `prod_idx` of n_prod: ring entries prod_idx, prod_idx + n_prod, ... (n_stages is a multiple of n_prod, so a
// stage always has the same producer, which therefore sees that stage's `empty` phases in order). Tiles are laid out
// group-major: q -> (group q / tpg, matrix, K-slice). A round is TPB tiles, R lanes per tile (one row-slice copy each).
__device__ __forceinline__ void mv_produce(const MvParams& p, const CtaWork& work, MvShared& sh, uint8_t* ring, int prod_idx, int lane) {
    const uint32_t S = (uint32_t)p.n_stages;
    const int n_prod = p.n_prod, tpg = work.tpg, n_sl = p.n_slices;
    if (tpg <= 0) return;
    const int tsub = lane / R, rsub = lane % R;
    const bool gateup = (p.mode == MV_GATEUP);
    // loop invariants out of the parameter block (a gate/up pair has one type, one K: mv_plan); the producers' issue rate
    // is what keeps the ring full, so the per-round bookkeeping is carried incrementally (no division in the loop)
    const MvSeg& sg0 = p.seg[gateup ? 0 : work.seg];
    const int sb = sg0.slice_bytes, row_bytes = sg0.row_bytes;
    const uintptr_t w0 = (uintptr_t)sg0.w, w1 = gateup ? (uintptr_t)p.seg[1].w : w0;
    const uint32_t stage_bytes = (uint32_t)p.stage_bytes, slot_off = (uint32_t)(rsub * p.slot_bytes);
    const int step = TPB * n_prod;
    const int step_g = step / tpg, step_w = step % tpg;
    const uint32_t step_u = (uint32_t)step / S, step_s = (uint32_t)step % S;
    int q = prod_idx + tsub * n_prod;
    int gi = q / tpg, within = q % tpg;
    uint32_t st = (uint32_t)q % S, use = (uint32_t)q / S;
    if (p.use_tma) {
        // one 2-D tensor copy per tile: box = (slice bytes, R rows) at (slice, row0); rows past the matrix and bytes past
        // the row are filled with zeros and still count towards the transaction bytes
        const void* tm0 = p.seg[gateup ? 0 : work.seg].tmap;
        const void* tm1 = gateup ? (const void*)p.seg[1].tmap : tm0;
        for (int qb = prod_idx; qb < work.total; qb += step) {
            if (q < work.total && rsub == 0) {
                const int m = (gateup && within >= n_sl) ? 1 : 0;
                const int slice = within - m * n_sl;
                mbar_wait(&sh.empty[st], (use & 1u) ^ 1u, p.err, 0x10001u, (unsigned)q);
                mbar_arrive_expect_tx(&sh.full[st], (uint32_t)(R * sb));
                st_volatile_shared(&sh.issued[st], use + 1u);
                if (p.use_tma == 2) tma_load_3d(ring + st * stage_bytes, m ? tm1 : tm0, 0, 2 * slice, work.r0 + gi * R, &sh.full[st]);
                else tma_load_2d(ring + st * stage_bytes, m ? tm1 : tm0, slice * (sb >> 3), work.r0 + gi * R, &sh.full[st]);
            }
            __syncwarp();       // rounds stay in order: a lane that ran ahead could arm a later use of a stage before an earlier one
            q += step;
            gi += step_g; within += step_w; if (within >= tpg) { within -= tpg; gi++; }
            use += step_u; st += step_s; if (st >= S) { st -= S; use++; }
        }
        return;
    }
    for (int qb = prod_idx; qb < work.total; qb += step) {
        const bool have = q < work.total;
        const uint8_t* src = nullptr; uint32_t bytes = 0; uint8_t* dst = nullptr;
        if (have) {
            const int m = (gateup && within >= n_sl) ? 1 : 0;
            const int slice = within - m * n_sl;
            const int row0 = work.r0 + gi * R, nrows = min(R, work.r1 - row0);
            if (rsub < nrows) {
                const int len = min(sb, row_bytes - slice * sb);           // the last slice may be shorter
                const uintptr_t start = (m ? w1 : w0) + (size_t)(row0 + rsub) * row_bytes + (size_t)(slice * sb);
                const uintptr_t a0 = start & ~(uintptr_t)15, a1 = (start + len + 15) & ~(uintptr_t)15;
                src = (const uint8_t*)a0; bytes = (uint32_t)(a1 - a0);
                dst = ring + st * stage_bytes + slot_off;
            }
            if (rsub == 0) mbar_wait(&sh.empty[st], (use & 1u) ^ 1u, p.err, 0x10001u, (unsigned)q);
        }
        __syncwarp();
        uint32_t tot = bytes;                       // bytes of the tile = sum over its R lanes
#pragma unroll
        for (int o = 1; o < R; o <<= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
        if (have && rsub == 0) {
            mbar_arrive_expect_tx(&sh.full[st], tot);
            st_volatile_shared(&sh.issued[st], use + 1u);
        }
        __syncwarp();
        if (bytes) tma_load_1d(dst, src, bytes, &sh.full[st]);
        q += step;
        gi += step_g; within += step_w; if (within >= tpg) { within -= tpg; gi++; }
        use += step_u; st += step_s; if (st >= S) { st -= S; use++; }
    }
}

// Consumer prologue: x (float32, global) -> optional RMSNorm weight -> two int8 planes + packed per-block
// scales in shared memory, in the layout(s) `layouts` asks for (bit 0: layout 0, bit 1: layout 1; SUMS16: + 16-element sums). Four
// threads share a 32-element block (8 values each; max and sums combined with two shuffles). The RMS factor
// is a scalar, so it is NOT applied here: the per-warp sums of squares go to sh.red[] and mv_rms_scale()
// turns them into the factor the epilogue multiplies the finished row sums with.
template <bool ACT2, bool SUMS16>
__device__ __forceinline__ void mv_quantize_input(const MvParams& p, MvShared& sh, uint8_t* smem, int layouts, int warp, int lane) {
    const bool has_norm = p.norm_w != nullptr;
    const int t = warp * 32 + lane, sub = t & 3, quad = t >> 2;
    const int nblk = p.K / 32;                         // a multiple of 8: a warp's 8 quads are in or out together
    const float MAGIC = 12582912.0f;                   // 1.5 * 2^23: float add rounds to nearest integer
    const float fs = ACT2 ? 1.0f / B2_ACT_RES : 1.0f;
    uint8_t* const lay0 = smem + p.act_off;
    uint8_t* const lay1 = smem + p.act_off;
    constexpr int U = 4;                               // blocks per thread whose loads are in flight together (18 warps: K = 18432 per pass)
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
                if (SUMS16 && (sub & 1) == 0) (f + p.K / 16)[un] = s16;        // Q2_K / Q3_K: sums of the 16-element runs, plain order
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

__device__ __forceinline__ ActView act_view(const MvParams& p, uint8_t* smem) {
    uint8_t* base = smem + p.act_off;
    ActView v;
    v.a = base; v.b = base + p.act_plane;
    v.f = reinterpret_cast<const float*>(base + 2 * p.act_plane);
    v.g = v.f + p.K / 16;
    return v;
}

// Sum the 4 row accumulators over the lanes (transposed butterfly, 6 shuffles + 3 broadcasts): every lane gets all sums.
__device__ __forceinline__ void rows_reduce(const float (&acc)[R], int lane, float (&out)[R]) {
    const bool hi16 = lane & 16, hi8 = lane & 8;
    const float s0 = hi16 ? acc[0] : acc[2], s1 = hi16 ? acc[1] : acc[3];
    float k0 = hi16 ? acc[2] : acc[0], k1 = hi16 ? acc[3] : acc[1];
    k0 += __shfl_xor_sync(0xffffffffu, s0, 16);
    k1 += __shfl_xor_sync(0xffffffffu, s1, 16);
    const float snd = hi8 ? k0 : k1;
    float kp = hi8 ? k1 : k0;
    kp += __shfl_xor_sync(0xffffffffu, snd, 8);
    kp += __shfl_xor_sync(0xffffffffu, kp, 4);
    kp += __shfl_xor_sync(0xffffffffu, kp, 2);
    kp += __shfl_xor_sync(0xffffffffu, kp, 1);      // lanes 0-7 / 8-15 / 16-23 / 24-31 hold row 0 / 1 / 2 / 3
    out[0] = kp;
    out[1] = __shfl_sync(0xffffffffu, kp, 8);
    out[2] = __shfl_sync(0xffffffffu, kp, 16);
    out[3] = __shfl_sync(0xffffffffu, kp, 24);
}

// One consumer warp's share of the launch: every NCW-th ring entry.
template <int T, bool ACT2>
__device__ __forceinline__ void consume(const MvParams& p, const CtaWork& work, MvShared& sh, const uint8_t* ring, float4* part,
                                        const ActView& av, float rs, int pos, int page, int warp, int lane) {
    const uint32_t S = (uint32_t)p.n_stages;
    const bool gateup = (p.mode == MV_GATEUP);
    const int tpg = work.tpg, n_sl = p.n_slices;
    if (warp >= work.total || tpg <= 0) return;      // no tile for this warp (tpg is 0 in CTAs without rows)
    // loop invariants out of the parameter block (a gate/up pair has one type, one K: mv_plan)
    const int seg0 = gateup ? 0 : work.seg;
    const uint32_t row_bytes = (uint32_t)p.seg[seg0].row_bytes;
    const uint32_t wlo0 = (uint32_t)(uintptr_t)p.seg[seg0].w, wlo1 = gateup ? (uint32_t)(uintptr_t)p.seg[1].w : wlo0;
    const uint32_t sb = (uint32_t)slice_bytes(T);
    const bool tma = p.use_tma != 0;                       // rows of a tile land densely at the 16-byte-aligned slice start
    const uint32_t stage_bytes = (uint32_t)p.stage_bytes, slot_bytes = tma ? sb : (uint32_t)p.slot_bytes;
    const int resid_off = p.seg[work.seg].out_off;
    // tile bookkeeping advances by NCW per iteration: carried incrementally (no integer division in the loop)
    int gi = warp / tpg, within = warp % tpg;
    uint32_t st = (uint32_t)warp % S, use = (uint32_t)warp / S;
    for (int q = warp; q < work.total; q += NCW) {
        const int m = (gateup && within >= n_sl) ? 1 : 0;
        const int sl = within - m * n_sl;
        const int row0 = work.r0 + gi * R, nrows = min(R, work.r1 - row0);
        const int gi_cur = gi, within_cur = within;
        within += NCW; while (within >= tpg) { within -= tpg; gi++; }
        const uint32_t st_cur = st, use_cur = use;
        st += NCW; while (st >= S) { st -= S; use++; }
        {   // the producer has armed this use of the stage (see the note at the top), then: its bytes have landed
            uint32_t spins = 0;
            while (ld_volatile_shared(&sh.issued[st_cur]) != use_cur + 1u) {
                if (++spins > (1u << 26)) {
                    volatile unsigned int* e = reinterpret_cast<volatile unsigned int*>(p.err);
                    if (e && e[0] == 0) { e[1] = blockIdx.x; e[2] = threadIdx.x; e[3] = (unsigned)q; e[4] = use_cur; e[0] = 0x10003u; }
                    break;
                }
            }
        }
        mbar_wait(&sh.full[st_cur], use_cur & 1u, p.err, 0x10002u, (unsigned)q);
        // residual of the group: requested now by every tile's lane 0 (16 bytes from L2), used by whichever warp ends up
        // finishing the group -- the round trip hides behind the dot products instead of sitting on the kernel's tail
        float2 res0 = make_float2(0.f, 0.f), res1 = res0;
        if (p.mode == MV_RESID && lane == 0) {
            const float* rp = p.resid + resid_off + row0;
            res0 = __ldcg(reinterpret_cast<const float2*>(rp));
            if (nrows == R) res1 = __ldcg(reinterpret_cast<const float2*>(rp + 2));
        }
        const uint8_t* stage = ring + st_cur * stage_bytes;
        float acc[R];
#pragma unroll
        for (int r = 0; r < R; r++) acc[r] = 0.f;
        if (sl * B2_SLICE + lane * 64 < p.K) {                     // lanes past K in a partial last slice idle
            LaneAct la;
            load_act<T, ACT2>(la, av, sl * B2_SLICE, lane);
            // each copy started at the 16-byte boundary below its row-slice (only Q6_K / Q2_K / Q3_K rows can sit off one)
            const uint32_t row_lo = (m ? wlo1 : wlo0) + (uint32_t)row0 * row_bytes + (uint32_t)sl * sb;
            // always R rows, results through a shift register (a rolled loop cannot index acc[] by r without a select per
            // element): a group with a single pair computes two rows of stale bytes that nobody looks at
#pragma unroll ROW_UNROLL
            for (int r = 0; r < R; r++) {
                const uint32_t off = ((T == T_Q6_K || T == T_Q2_K || T == T_Q3_K) && !tma) ? ((row_lo + r * row_bytes) & 15u) : 0u;
                const float d = row_dot<T, ACT2>(stage + r * slot_bytes + off, lane, la);
                acc[0] = acc[1]; acc[1] = acc[2]; acc[2] = acc[3]; acc[3] = d;
            }
        }
        fence_proxy_async();    // generic-proxy reads of the stage, ordered before the TMA (async proxy) refill
        __syncwarp();
        if (lane == 0) mbar_arrive(&sh.empty[st_cur]);             // the stage can be refilled
        float v[R], gv[R] = {0.f, 0.f, 0.f, 0.f};
        rows_reduce(acc, lane, v);
        if (lane != 0) continue;
        if (tpg > 1) {          // hand the tile's sums in; whoever brings the group's last tile adds them up in tile order
            const int slot = gi_cur % MV_GSLOTS;
            const unsigned int gen = (unsigned)(gi_cur / MV_GSLOTS);
            while (ld_volatile_shared(&sh.gen[slot]) != gen) __nanosleep(20);     // the group MV_GSLOTS back is done with the slot
            float4* ps = part + slot * tpg;
            ps[within_cur] = make_float4(v[0], v[1], v[2], v[3]);
            __threadfence_block();
            const unsigned int old = atomicAdd(&sh.gcnt[slot], 1u);
            if (old != (unsigned)tpg - 1u) continue;
            __threadfence_block();
#pragma unroll
            for (int r = 0; r < R; r++) v[r] = 0.f;
            if (gateup) {
                for (int i = 0; i < n_sl; i++) { const float4 c = ps[i]; gv[0] += c.x; gv[1] += c.y; gv[2] += c.z; gv[3] += c.w; }
                for (int i = n_sl; i < tpg; i++) { const float4 c = ps[i]; v[0] += c.x; v[1] += c.y; v[2] += c.z; v[3] += c.w; }
            } else {
                for (int i = 0; i < tpg; i++) { const float4 c = ps[i]; v[0] += c.x; v[1] += c.y; v[2] += c.z; v[3] += c.w; }
            }
            st_volatile_shared(&sh.gcnt[slot], 0u);
            st_release_shared(&sh.gen[slot], gen + 1u);
        }
#pragma unroll
        for (int r = 0; r < R; r++) { v[r] *= rs; gv[r] *= rs; }   // fused RMSNorm factor (a scalar: applied to the sums)
        if (gateup) {
            float h[R];
#pragma unroll
            for (int r = 0; r < R; r++) h[r] = gv[r] / (1.0f + __expf(-gv[r])) * v[r];
            *reinterpret_cast<float2*>(p.y + row0) = make_float2(h[0], h[1]);
            if (nrows == R) *reinterpret_cast<float2*>(p.y + row0 + 2) = make_float2(h[2], h[3]);
        } else {
            epilogue_pair(p, sh.rope, work.seg, row0, v[0], v[1], pos, page, res0);
            if (nrows == R) epilogue_pair(p, sh.rope, work.seg, row0 + 2, v[2], v[3], pos, page, res1);
        }
    }
}

// EXTRA = the instantiation for the block formats outside llama.cpp's common mixes (Q4_1, Q5_0, Q5_1, Q2_K, Q3_K): kept out of
// the main kernel, whose code layout they disturbed (+2 % per token on Q4_K_M with all ten formats in one kernel)
template <bool ACT2, bool EXTRA>
__global__ void __launch_bounds__(MV_THREADS, 1) matvec_kernel(const __grid_constant__ MvParams p) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    MvShared& sh = *reinterpret_cast<MvShared*>(smem_raw);
    uint8_t* ring = smem_raw + p.ring_off;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const CtaWork work = cta_work(p, blockIdx.x, gridDim.x);
    if (threadIdx.x == 0) tl_stamp(p.tl, 0);

    // barrier set-up spread over the threads (one thread doing ~100 mbarrier.init in a row cost over a microsecond per launch)
    if ((int)threadIdx.x < p.n_stages) {
        mbar_init(&sh.full[threadIdx.x], 1); mbar_init(&sh.empty[threadIdx.x], 1); sh.issued[threadIdx.x] = 0;
    } else if (threadIdx.x >= 64 && threadIdx.x < 64 + MV_GSLOTS) {
        sh.gcnt[threadIdx.x - 64] = 0; sh.gen[threadIdx.x - 64] = 0;
    }
    fence_mbar_init();
    __syncthreads();
    if (!p.trigger_late) griddep_launch();   // the next kernel may start (it will prefetch its own weights)

    if (warp >= NCW) {  // producer warps never touch activations: they run ahead of the dependency wait
        if (warp - NCW < p.n_prod) mv_produce(p, work, sh, ring, warp - NCW, lane);
        return;         // (with trigger_late they never trigger: the CTA's trigger is the consumers', after their wait)
    }
    if (p.norm_w) {     // weights, not activations: warm L1 while the previous kernel finishes
        for (int i = threadIdx.x * 32; i < p.K; i += NCW * 32 * 32)
            asm volatile("prefetch.global.L1 [%0];" ::"l"(p.norm_w + i));
    }
    griddep_wait();     // activations written by the previous kernel are visible from here on
    if (threadIdx.x == 0) tl_stamp(p.tl, 1);
    if (p.trigger_late) griddep_launch();
    int pos = 0, page = 0;      // QKV epilogue inputs: fetched once per thread, here
    if (p.mode == MV_QKV) {
        pos = *p.pos_ptr; page = p.block_table[pos / p.block_tokens];
        if ((int)threadIdx.x < p.head_dim / 2) {           // head_dim <= 128; published by the barrier that ends the prologue
            float sn, cs;
            sincosf((float)pos * __ldg(p.rope_freq + threadIdx.x), &sn, &cs);
            sh.rope[threadIdx.x] = make_float2(cs, sn);
        }
    }
    // plane layout follows the quant type this CTA reads (a gate/up pair has one type: mv_plan)
    const int type0 = (p.mode == MV_GATEUP) ? p.seg[0].type : p.seg[work.seg].type;
    mv_quantize_input<ACT2, EXTRA>(p, sh, smem_raw, 1 << act_layout_of(type0), warp, lane);
    const ActView av = act_view(p, smem_raw);
    const float rs = mv_rms_scale(p, sh);
    if (threadIdx.x == 0) tl_stamp(p.tl, 2);
    float4* part = reinterpret_cast<float4*>(smem_raw + p.part_off);
#define B2_MV_CASE(TY) case TY: consume<TY, ACT2>(p, work, sh, ring, part, av, rs, pos, page, warp, lane); break;
    if (!EXTRA) {
        switch (type0) { B2_MV_CASE(T_Q4_K) B2_MV_CASE(T_Q6_K) B2_MV_CASE(T_Q8_0) B2_MV_CASE(T_Q4_0) B2_MV_CASE(T_Q5_K) }
    } else {        // a launch with any of the other formats among its segments runs this instantiation for all of them
        switch (type0) {
            B2_MV_CASE(T_Q4_K) B2_MV_CASE(T_Q6_K) B2_MV_CASE(T_Q8_0) B2_MV_CASE(T_Q4_0) B2_MV_CASE(T_Q5_K)
            B2_MV_CASE(T_Q4_1) B2_MV_CASE(T_Q5_0) B2_MV_CASE(T_Q5_1) B2_MV_CASE(T_Q2_K) B2_MV_CASE(T_Q3_K)
        }
    }
#undef B2_MV_CASE
    if (p.tl) { named_bar_sync(1, NCW * 32); if (threadIdx.x == 0) tl_stamp(p.tl, 3); }
}

}  // namespace b2
