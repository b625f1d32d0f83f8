// Decode mat-vec with in-kernel dequantisation: y = W x for GGUF-quantised W (Q4_0, Q8_0, Q4_K,
// Q5_K, Q6_K), weights streamed unchanged from HBM. Replaces ggml-cuda's quantize_q8_1 +
// mul_mat_vec_q + the separate rms_norm / rope / cpy / silu / add launches that the reference
// runs under llm.create_chat_completion (/root/reference/api.py:55-63; SURVEY.md 2.3 K2-K7,K11).
//
// Structure: one CTA of 16 warps per SM, all of the SM's shared memory as a ring of weight tiles.
//   * Tiles. A tile is a row pair (one RoPE pair; a single row when two very long rows would not leave four
//     ring stages). Rows are contiguous in memory, so a tile is ONE bulk copy of 4-30 KB (cp.async.bulk +
//     mbarrier complete_tx). Measured on B200: a warp issues a cp.async.bulk only every ~110 cycles, so
//     copies of 1-2 KB cap a CTA at ~3 TB/s chip rate however deep the ring is; copies of >= 4 KB do not.
//   * Teams. K is cut into slices of 2048 elements; team t = the warps with warp % n_slices == t owns slice t
//     for the whole kernel and keeps that slice's 64 quantised activations per lane (two int8 planes +
//     scales, 40 registers) in REGISTERS: activations are read from shared memory once per kernel, not once
//     per tile (re-reading them per row pair took most of the shared-memory bandwidth and of the issue slots).
//     The row pairs of the CTA are dealt round-robin to the members of each team, so every tile is read by
//     exactly one warp of every team.
//   * Ring. Tile q of the CTA lives in stage q % S. There is no producer warp: the LAST of a tile's readers
//     (mbarrier `empty[s]` with one arrival per team; the arrival that finds one pending is the last) refills
//     the stage with tile q + S -- sixteen potential issuers. An mbarrier only exposes the parity of its
//     phase and consecutive uses of a stage belong to different warps, so a warp that looked at a stage
//     before the previous use had even landed would read "complete" for the wrong phase: the refilling warp
//     publishes `issued[s] = use + 1` after arming the barrier, and a reader waits for that value (by then
//     the previous use has been consumed, hence landed) before it waits on the parity.
//   * Row sums. Each warp reduces its slice of the row pair over its lanes (5 shuffles) and drops the partial
//     sums in a shared-memory slot; the warp that brings the last partial (mbarrier again) adds them in team
//     order -- the result does not depend on timing -- and runs the epilogue.
//   * The first S tiles are requested BEFORE griddepcontrol.wait (weights do not depend on the previous kernel).
//   * prologue (after the wait): x -> optional fused RMSNorm weight -> two int8 planes + packed per-block
//     scales in shared memory (loads of x in one batch: one L2 round trip; norm weights prefetched to L1
//     before the wait), then every warp loads its slice into registers.
//   * epilogue (lane 0 of the finishing warp): plain store, residual add, RoPE + paged-KV store for the fused
//     QKV projection, or SiLU(gate) * up for the fused FFN projection.
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
#ifndef B2_MV_CTAS_PER_SM
#define B2_MV_CTAS_PER_SM 1
#endif
constexpr int MV_CTAS_PER_SM = B2_MV_CTAS_PER_SM;      // experiment knob: 2 = two 8-warp CTAs per SM (build with -DB2_MV_NCW=8)
constexpr int MV_SMEM_MAX = (MV_CTAS_PER_SM == 1 ? 224 : 112) * 1024;
constexpr int MV_GSLOTS = 16;               // row pairs whose partial sums can be in flight at once

struct __align__(16) MvShared {
    uint64_t full[MV_MAX_STAGES];           // a tile's bytes have landed (expect_tx)
    uint64_t empty[MV_MAX_STAGES];          // one arrival per team: the tile has been read
    uint64_t gbar[MV_GSLOTS];               // one arrival per team: the row pair's partial sums are in
    unsigned int issued[MV_MAX_STAGES];     // use number + 1 last armed on each stage
    unsigned int gen[MV_GSLOTS];            // slot i serves row pair g with g % MV_GSLOTS == i and g / MV_GSLOTS == gen[i]
    float red[NCW];
    uint64_t probe;                         // scratch barrier: what does pending_count report for the last arrival?
    unsigned int last_cnt;                  // ... this value (measured once per CTA instead of trusting a reading of the PTX manual)
    float4 part[MV_GSLOTS][NCW];            // per team: {row0, row1, gate row0, gate row1}
};
static_assert(sizeof(MvShared) <= MV_SHARED_HDR, "header does not fit");

struct CtaWork {
    int seg;            // segment (non gate/up modes)
    int r0, r1;         // row range [r0, r1), both even
    int ng;             // row groups of p.nrows rows
    int tpg;            // tiles per group: 1, or 2 for gate/up (gate tile, up tile)
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
    w.ng = (w.r1 - w.r0) / p.nrows;
    w.tpg = gu ? 2 : 1;
    return w;
}

// lane 0 of the finishing warp completes a row pair
// (pos, page: position of the token and its KV page, fetched once per thread at kernel start; res: the residual pair,
// fetched when the warp started on the row pair -- no global round trip is left on this path)
__device__ __forceinline__ void epilogue_pair(const MvParams& p, int seg, int row, float v0, float v1, int pos, int page, float2 res) {
    const MvSeg& sg = p.seg[seg];
    if (p.mode == MV_STORE) {
        *reinterpret_cast<float2*>(p.y + sg.out_off + row) = make_float2(v0, v1);
    } else if (p.mode == MV_RESID) {
        *reinterpret_cast<float2*>(p.y + sg.out_off + row) = make_float2(res.x + v0, res.y + v1);
    } else if (p.mode == MV_QKV) {
        const int hd = p.head_dim;
        if (seg < 2) {                      // Q or K: rotate the adjacent pair (ggml "norm" RoPE)
            const int i = (row % hd) >> 1;
            const float ang = (float)pos * __ldg(p.rope_freq + i);
            float s, c;
            sincosf(ang, &s, &c);
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
// arrive on an mbarrier and report the pending-arrival count of the state the arrive returns
__device__ __forceinline__ uint32_t mbar_arrive_pending(uint64_t* bar, uint64_t& state) {
    uint32_t cnt;
    asm volatile("{\n\tmbarrier.arrive.shared::cta.b64 %0, [%2];\n\tmbarrier.pending_count.b64 %1, %0;\n\t}"
                 : "=l"(state), "=r"(cnt) : "r"(smem_u32(bar)) : "memory");
    return cnt;
}
__device__ __forceinline__ bool mbar_test_wait_state(uint64_t* bar, uint64_t state) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "l"(state) : "memory");
    return ok != 0;
}

// Request tile t (0, or 1 = the second matrix of gate/up) of group g into `stage`: ONE bulk copy (lane 0 of the calling warp).
__device__ __forceinline__ void tile_issue(const MvParams& p, const CtaWork& w, int g, int t, uint8_t* stage, uint64_t* bar,
                                           unsigned int* issued, unsigned int use, int lane) {
    if (lane != 0) return;
    const MvSeg& sg = p.seg[(p.mode == MV_GATEUP) ? t : w.seg];
    const uintptr_t start = (uintptr_t)sg.w + (size_t)(w.r0 + g * p.nrows) * sg.row_bytes;
    const uintptr_t a0 = start & ~(uintptr_t)15, a1 = (start + (size_t)p.nrows * sg.row_bytes + 15) & ~(uintptr_t)15;
    mbar_arrive_expect_tx(bar, (uint32_t)(a1 - a0));
    st_volatile_shared(issued, use + 1);
    tma_load_1d(stage, (const void*)a0, (uint32_t)(a1 - a0), bar);
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
    uint8_t* const lay0 = smem + p.act_off;
    uint8_t* const lay1 = smem + p.act_off;
    constexpr int U = 4;                               // blocks per thread whose loads are in flight together (16 warps: K = 16384 per pass)
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

__device__ __forceinline__ ActView act_view(const MvParams& p, uint8_t* smem) {
    uint8_t* base = smem + p.act_off;
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

// One warp = (team, member): K-slice `team` of the row groups g = member, member + team size, ...
template <int T, bool ACT2>
__device__ __forceinline__ void consume(const MvParams& p, const CtaWork& work, MvShared& sh, uint8_t* ring,
                                        const ActView& av, float rs, int pos, int page, int warp, int lane) {
    const bool gateup = (p.mode == MV_GATEUP);
    const int S = p.n_stages, tpg = work.tpg, ng = work.ng, n_teams = p.n_slices, nrows = p.nrows;
    const int team = warp % n_teams, member = warp / n_teams;
    const int team_size = (NCW - 1 - team) / n_teams + 1;
    const unsigned int n_tiles = (unsigned)(ng * tpg);
    const unsigned int last_cnt = sh.last_cnt;
    const bool active = team * B2_SLICE + lane * 64 < p.K;              // lanes past K in a partial last slice idle
    LaneAct la;
    if (active) load_act<T, ACT2>(la, av, team * B2_SLICE, lane);       // this warp's activations, for the whole kernel
    const int sb = slice_bytes(T);
    for (int g = member; g < ng; g += team_size) {
        const int row0 = work.r0 + g * nrows;
        float2 res = make_float2(0.f, 0.f);
        if (p.mode == MV_RESID && lane == 0) {         // early: the round trip hides behind the tile (any team may end up finishing the group)
            const float* rp = p.resid + p.seg[work.seg].out_off + row0;
            if (nrows == 2) res = __ldcg(reinterpret_cast<const float2*>(rp)); else res.x = __ldcg(rp);
        }
        float acc[R], gacc[R];
#pragma unroll
        for (int r = 0; r < R; r++) { acc[r] = 0.f; gacc[r] = 0.f; }
        for (int t = 0; t < tpg; t++) {
            const MvSeg& sg = p.seg[gateup ? t : work.seg];
            const unsigned int q = (unsigned)(g * tpg + t);
            const unsigned int u = __umulhi(q, p.inv_stages);
            const int s = (int)(q - u * (unsigned)S);
            uint8_t* stage = ring + (size_t)s * p.stage_bytes;
            {   // this use of the stage has been armed (see the ring note at the top), then: its bytes have landed
                uint32_t spins = 0;
                while (ld_volatile_shared(&sh.issued[s]) != u + 1u) {
                    __nanosleep(20);
                    if (++spins > (1u << 23)) {
                        volatile unsigned int* e = reinterpret_cast<volatile unsigned int*>(p.err);
                        if (e && e[0] == 0) { e[1] = blockIdx.x; e[2] = threadIdx.x; e[3] = (unsigned)g; e[4] = u; e[0] = 0x10003u; }
                        break;
                    }
                }
            }
            mbar_wait(&sh.full[s], u & 1u, p.err, 0x10002u, (unsigned)g);
            // the copy started at the 16-byte boundary below row 0 (only Q6_K rows can sit off one: row_dot copes)
            const uint32_t off0 = ((uint32_t)(uintptr_t)sg.w + (uint32_t)row0 * (uint32_t)sg.row_bytes) & 15u;
            const uint8_t* r0p = stage + off0 + team * sb;
#ifndef B2_MV_SKIP_COMPUTE      // experiment builds only (scripts/mv_variants.sh): stream the tiles, skip the dot products
            if (active) {
                if (nrows == 2) {
                    const uint8_t* r1p = r0p + sg.row_bytes;
#pragma unroll ROW_UNROLL
                    for (int r = 0; r < R; r++) {
                        const float d = row_dot<T, ACT2>(r ? r1p : r0p, lane, la);
#pragma unroll
                        for (int k = 0; k < R; k++) if (k == r) acc[k] += d;
                    }
                } else {
                    acc[0] += row_dot<T, ACT2>(r0p, lane, la);
                }
            }
#endif
            if (gateup && t == 0) {                    // gate done: keep its sums, go on with up
#pragma unroll
                for (int r = 0; r < R; r++) { gacc[r] = acc[r]; acc[r] = 0.f; }
            }
            // done with the stage: the last of its readers (one warp per team) refills it
            fence_proxy_async();    // generic-proxy reads of the stage, ordered before the TMA (async proxy) refill
            __syncwarp();
            bool last = false;
            if (lane == 0) { uint64_t st; last = mbar_arrive_pending(&sh.empty[s], st) == last_cnt; }
            last = __shfl_sync(0xffffffffu, last, 0);
            if (last && q + (unsigned)S < n_tiles) {
                const unsigned int q2 = q + (unsigned)S;
                const int g2 = (int)(tpg == 2 ? q2 >> 1 : q2), t2 = (int)(tpg == 2 ? q2 & 1u : 0u);
                tile_issue(p, work, g2, t2, stage, &sh.full[s], &sh.issued[s], u + 1u, lane);
            }
        }
        float v0, v1, g0 = 0.f, g1 = 0.f;
        if (nrows == 2) {
            pair_reduce(acc, lane, v0, v1);
            if (gateup) pair_reduce(gacc, lane, g0, g1);
        } else {
            v0 = warp_sum(acc[0]); v1 = 0.f;
        }
        if (lane == 0) {
            bool finish = true;
            if (n_teams > 1) {     // hand this team's partial sums in; the team that completes the group adds them in team order
                const int slot = g % MV_GSLOTS;
                const unsigned int gen = (unsigned)(g / MV_GSLOTS);
                while (ld_volatile_shared(&sh.gen[slot]) != gen) __nanosleep(20);     // the group MV_GSLOTS back is done with the slot
                sh.part[slot][team] = make_float4(v0, v1, g0, g1);
                uint64_t st;
                finish = mbar_arrive_pending(&sh.gbar[slot], st) == last_cnt;         // release: our partial sums are visible ...
                if (finish) {
                    while (!mbar_test_wait_state(&sh.gbar[slot], st)) {}              // ... acquire: and so are everybody else's
                    v0 = v1 = g0 = g1 = 0.f;
                    for (int i = 0; i < n_teams; i++) {
                        const float4 c = sh.part[slot][i];
                        v0 += c.x; v1 += c.y; g0 += c.z; g1 += c.w;
                    }
                    st_release_shared(&sh.gen[slot], gen + 1u);
                }
            }
            if (finish) {
                v0 *= rs; v1 *= rs;                        // fused RMSNorm factor (a scalar: applied to the sums)
                if (gateup) {
                    g0 *= rs; g1 *= rs;
                    const float h0 = g0 / (1.0f + __expf(-g0)) * v0;
                    const float h1 = g1 / (1.0f + __expf(-g1)) * v1;
                    *reinterpret_cast<float2*>(p.y + row0) = make_float2(h0, h1);
                } else if (nrows == 2) {
                    epilogue_pair(p, work.seg, row0, v0, v1, pos, page, res);
                } else {                                   // single rows: plain store / residual add only (mv_plan)
                    float* y = p.y + p.seg[work.seg].out_off + row0;
                    *y = res.x + v0;
                }
            }
        }
    }
}

template <bool ACT2>
__global__ void __launch_bounds__(MV_THREADS, MV_CTAS_PER_SM) matvec_kernel(const MvParams p) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    MvShared& sh = *reinterpret_cast<MvShared*>(smem_raw);
    uint8_t* ring = smem_raw + p.ring_off;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const CtaWork work = cta_work(p, blockIdx.x, gridDim.x);
    if (threadIdx.x == 0) tl_stamp(p.tl, 0);

    // barrier set-up spread over the threads (one thread doing ~110 mbarrier.init in a row cost over a microsecond per launch)
    if ((int)threadIdx.x < p.n_stages) {
        mbar_init(&sh.full[threadIdx.x], 1); mbar_init(&sh.empty[threadIdx.x], p.n_slices); sh.issued[threadIdx.x] = 0;
    } else if (threadIdx.x >= 64 && threadIdx.x < 64 + MV_GSLOTS) {
        mbar_init(&sh.gbar[threadIdx.x - 64], p.n_slices); sh.gen[threadIdx.x - 64] = 0;
    } else if (threadIdx.x == 96) {
        mbar_init(&sh.probe, 1);
        uint64_t st;
        sh.last_cnt = mbar_arrive_pending(&sh.probe, st);    // the only (= last) arrival of a one-arrival barrier
    }
    fence_mbar_init();
    __syncthreads();
    if (!p.trigger_late) griddep_launch();   // the next kernel may start (it will prefetch its own weights)
    {   // first pass over the ring: weights do not depend on the previous kernel, so this runs ahead of the wait
        const int n_first = min(p.n_stages, work.ng * work.tpg);
        for (int q = warp; q < n_first; q += NCW)
            tile_issue(p, work, work.tpg == 2 ? q >> 1 : q, work.tpg == 2 ? q & 1 : 0, ring + (size_t)q * p.stage_bytes,
                       &sh.full[q], &sh.issued[q], 0u, lane);
    }
    if (p.norm_w) {     // weights, not activations: warm L1 while the previous kernel finishes
        for (int i = threadIdx.x * 32; i < p.K; i += MV_THREADS * 32)
            asm volatile("prefetch.global.L1 [%0];" ::"l"(p.norm_w + i));
    }
    griddep_wait();     // activations written by the previous kernel are visible from here on
    if (threadIdx.x == 0) tl_stamp(p.tl, 1);
    if (p.trigger_late) griddep_launch();
    // plane layout follows the quant type this CTA reads (a gate/up pair has one type: mv_plan)
    const int type0 = (p.mode == MV_GATEUP) ? p.seg[0].type : p.seg[work.seg].type;
    mv_quantize_input<ACT2>(p, sh, smem_raw, 1 << act_layout_of(type0), warp, lane);
    const ActView av = act_view(p, smem_raw);
    const float rs = mv_rms_scale(p, sh);
    int pos = 0, page = 0;      // QKV epilogue inputs: fetched here, consumed only after the first tile (no round trip left in the epilogue)
    if (p.mode == MV_QKV) { pos = *p.pos_ptr; page = p.block_table[pos / p.block_tokens]; }
    if (threadIdx.x == 0) tl_stamp(p.tl, 2);
    switch (type0) {
        case T_Q4_K: consume<T_Q4_K, ACT2>(p, work, sh, ring, av, rs, pos, page, warp, lane); break;
        case T_Q6_K: consume<T_Q6_K, ACT2>(p, work, sh, ring, av, rs, pos, page, warp, lane); break;
        case T_Q8_0: consume<T_Q8_0, ACT2>(p, work, sh, ring, av, rs, pos, page, warp, lane); break;
        case T_Q4_0: consume<T_Q4_0, ACT2>(p, work, sh, ring, av, rs, pos, page, warp, lane); break;
        case T_Q5_K: consume<T_Q5_K, ACT2>(p, work, sh, ring, av, rs, pos, page, warp, lane); break;
    }
    if (p.tl) { __syncthreads(); if (threadIdx.x == 0) tl_stamp(p.tl, 3); }
}

// ---------------------------------------------------------------- host side
static int g_stage_cap = 0;      // B2_MV_STAGES (tuning): read once in mv_init

cudaError_t mv_init() {   // per device: allow > 48 KB dynamic shared memory; read the tuning knob once
    static bool env_read = false;
    if (!env_read) {
        if (const char* s = getenv("B2_MV_STAGES")) g_stage_cap = atoi(s);
        env_read = true;
    }
    cudaError_t e = cudaFuncSetAttribute(matvec_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, MV_SMEM_MAX);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(matvec_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, MV_SMEM_MAX);
}

int mv_plan(MvParams& p, int n_ctas) {
    if (p.K % 256 || p.K <= 0) return -1;
    p.n_slices = (p.K + B2_SLICE - 1) / B2_SLICE;
    if (p.n_slices > NCW) return -2;                       // one warp team per K-slice: K <= 32768
    int max_row = 0;
    for (int s = 0; s < p.nseg; s++) {
        MvSeg& sg = p.seg[s];
        const int be = block_elems(sg.type), bb = block_bytes(sg.type);
        if (!bb || be == 1) return -3;
        sg.row_bytes = p.K / be * bb;
        if (sg.n_rows % 2) return -4;
        max_row = sg.row_bytes > max_row ? sg.row_bytes : max_row;
    }
    if (p.mode == MV_GATEUP) {
        if (p.nseg != 2 || p.seg[0].n_rows != p.seg[1].n_rows) return -6;
        if (p.seg[0].type != p.seg[1].type) return -8;     // a mixed pair is run unfused by the caller
    }
    // activation block: [plane a][plane b][K/8 floats of packed scales]
    p.act_plane = p.n_slices * B2_SLICE;
    const int act_block = (2 * p.act_plane + p.K / 2 + 127) / 128 * 128;
    p.act_off = MV_SHARED_HDR;
    p.ring_off = MV_SHARED_HDR + act_block;
    // a tile = a row pair (ONE copy of 2 * row_bytes); single rows when pairs of very long rows would leave < 4 stages
    p.nrows = 2;
    p.stage_bytes = (2 * max_row + 32 + 127) / 128 * 128;  // + slack: the copy starts / ends on 16-byte boundaries
    int S = (MV_SMEM_MAX - p.ring_off) / p.stage_bytes;
    if (S < 4 && (p.mode == MV_STORE || p.mode == MV_RESID)) {
        p.nrows = 1;
        p.stage_bytes = (max_row + 32 + 127) / 128 * 128;
        S = (MV_SMEM_MAX - p.ring_off) / p.stage_bytes;
    }
    if (S > MV_MAX_STAGES) S = MV_MAX_STAGES;
    if (g_stage_cap >= 2 && g_stage_cap < S) S = g_stage_cap;
    if (S < 2) return -5;
    p.n_stages = S;
    p.inv_stages = 0xFFFFFFFFu / (unsigned)S + 1u;
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

int mv_ctas_per_sm() { return MV_CTAS_PER_SM; }
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
