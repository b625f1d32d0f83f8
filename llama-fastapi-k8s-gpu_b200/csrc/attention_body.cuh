// Flash-decode attention: shared-memory layout and the per-CTA body (see attention.cu for the design
// notes). Shared by the stand-alone kernel (attention.cu) and the persistent decode kernel (mega.cu).
#pragma once
#include "kernels.h"
#include "ptx.cuh"
#include <cooperative_groups.h>
#include <math.h>

namespace cg = cooperative_groups;

// B2_TIMELINE=1 diagnosis: where timeline slot 2 is taken (1 q in registers, 2 first tile landed, 3 first tile done,
// 4 all tiles done, 5 warps merged = before the split merge), and optionally for which CTA only (KV head * n_splits + split)
#ifndef B2_ATTN_STAMP
#define B2_ATTN_STAMP 5
#endif
#ifndef B2_ATTN_ONE_CTA
#define B2_ATTN_ONE_CTA -1
#endif

namespace b2 {

constexpr int AT_WT = 8;       // positions of a tile served by one warp (the score reduction is written for 4 heads x 8 positions = 32 values)
constexpr int AT_GMAX = 8;     // query heads per KV head
constexpr int AT_MAX_PAGES = 512;
constexpr int AT_BAR = 2;      // named barrier used by the attention threads
constexpr int AT_STAGE = 64;   // floats per head in the merge staging area
constexpr int AT_NST = 3;      // K/V tile buffers (two tiles in flight behind the one being read)
// Positions per tile: 128 with up to 4 query heads per KV head (NP = 1 pass; 16 warps), 64 with up to 8 (NP = 2; 8 warps:
// the second pass doubles the registers per thread). One warp per 8 positions either way.
__host__ __device__ constexpr int at_tt(int np) { return np == 1 ? 128 : 64; }
__host__ __device__ constexpr int at_threads(int np) { return at_tt(np) / AT_WT * 32; }

template <int HD, int TT>
struct __align__(16) AttnShared {
    __half k[AT_NST][TT][HD];          // after the last tile: the warps' un-normalised outputs, float [warps][G][HD]
    __half v[AT_NST][TT][HD];
    float part[AT_GMAX * HD + 8];      // cluster merge: slice [r * per, (r+1) * per) = what split r sent us (its un-normalised outputs of OUR elements)
    float part_ml[8][AT_GMAX][2];      // cluster merge: (running max, running sum) of every head from every split
    float sc[4][AT_GMAX][AT_STAGE];    // staging: the warps' (max, sum) per head, then the global-scratch merge weights
    float m_run[AT_GMAX], l_run[AT_GMAX];
    int pages[AT_MAX_PAGES];
    unsigned int ticket;
};

// Sum 32 per-lane values over the 32 lanes so that lane i ends up with the total of value i (transposed butterfly:
// 31 shuffles instead of 32 x 5).
__device__ __forceinline__ float lanes_transpose_sum(float (&v)[32], int lane) {
    const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4, h2 = lane & 2, h1 = lane & 1;
    float a[16], b[8], c[4], d[2];
#pragma unroll
    for (int i = 0; i < 16; i++) a[i] = (h16 ? v[i + 16] : v[i]) + __shfl_xor_sync(0xffffffffu, h16 ? v[i] : v[i + 16], 16);
#pragma unroll
    for (int i = 0; i < 8; i++) b[i] = (h8 ? a[i + 8] : a[i]) + __shfl_xor_sync(0xffffffffu, h8 ? a[i] : a[i + 8], 8);
#pragma unroll
    for (int i = 0; i < 4; i++) c[i] = (h4 ? b[i + 4] : b[i]) + __shfl_xor_sync(0xffffffffu, h4 ? b[i] : b[i + 4], 4);
#pragma unroll
    for (int i = 0; i < 2; i++) d[i] = (h2 ? c[i + 2] : c[i]) + __shfl_xor_sync(0xffffffffu, h2 ? c[i] : c[i + 2], 2);
    return (h1 ? d[1] : d[0]) + __shfl_xor_sync(0xffffffffu, h1 ? d[0] : d[1], 1);
}

// The whole per-CTA job: split `sp` of KV head `h`. Runs on AT_THREADS threads (tid 0..255) that
// synchronise on named barrier AT_BAR. `full` are two mbarriers that live across calls; `use` counts how often
// each has completed (the caller keeps it; every thread holds the same value).
//
// Inside a tile every warp is a flash-decode of its own: warp w serves positions 8w .. 8w+7 of the tile, a lane
// owns head_dim / 32 adjacent dims. Scores of 4 heads x 8 positions are 32 per-lane partial dot products, summed
// over the lanes with one transposed butterfly (lane i gets score i); max / sum over the 8 positions of a head are
// three shuffles inside an 8-lane group; the probabilities go back out by shuffle for o += p . v. Running max, sum
// and output stay in registers across tiles; there is no block-wide barrier between the phases (the first version
// had three per tile and took ~2.5 us per 64-position tile; measured inside the decode chain). The eight warps'
// results are merged once, through shared memory, before the splits are.
//
// `early`: the body is entered BEFORE griddepcontrol.wait (programmatic dependent launch). Position and page table are
// already valid then (the kernel that writes q and this token's K/V row lets us start only after ITS dependency wait,
// MvParams::trigger_late), so the K/V tiles of the older positions are requested first and the wait happens inside,
// just before q and the newest row are needed: the copies overlap the tail of the QKV kernel.
// NP = passes of 4 query heads (G <= 4: 1, else 2) is a compile-time constant and every warp runs every step of a tile,
// masked, so that the shuffles sit in straight-line code: inside `if (warp has positions)` / `if (4 * pass < G)` the
// compiler could not prove convergence and turned each shuffle into a WARPSYNC.COLLECTIVE / branch / ENDCOLLECTIVE
// sequence (127 of them; a 64-position tile took 2.1 us instead of ~0.5).
template <int HD, bool CLUSTER, int NP>
__device__ __forceinline__ void attn_decode_body(const AttnParams& p, AttnShared<HD, at_tt(NP)>& sh, uint64_t* full, uint32_t (&use)[AT_NST],
                                                 int sp, int h, int tid, bool early = false) {
    constexpr int AT_TT = at_tt(NP), AT_THREADS = at_threads(NP), AT_WARPS = AT_THREADS / 32;
    static_assert(AT_WARPS * 4 * NP <= AT_TT, "warp outputs alias the K tiles");
    static_assert(AT_WARPS * AT_GMAX * 2 <= AT_GMAX * AT_STAGE, "warp (max, sum) staging fits sc[0]");
    constexpr int DPL = HD / 32;                  // dims per lane
    const int warp = tid >> 5, lane = tid & 31;
    const int G = p.n_head / p.n_kv;
    const int bt = p.block_tokens;
    const bool rec_tl = tid == 0 && (B2_ATTN_ONE_CTA < 0 || h * p.n_splits + sp == B2_ATTN_ONE_CTA);

    // one round trip: position and page table
    const int pos = *p.pos_ptr;
    for (int i = tid; i < p.max_pages && i < AT_MAX_PAGES; i += AT_THREADS) sh.pages[i] = p.block_table[i];
    // never index past the page ids that were staged above, whatever the caller's position says
    const int ctx = min(pos + 1, min(p.max_pages, AT_MAX_PAGES) * bt);
    int chunk = (ctx + p.n_splits - 1) / p.n_splits;
    chunk = (chunk + 15) & ~15;
    const int p0 = min(ctx, sp * chunk), p1 = min(ctx, p0 + chunk);
    const int ntile = (p1 - p0 + AT_TT - 1) / AT_TT;
    named_bar_sync(AT_BAR, AT_THREADS);

    // thread 0: arm the tile's barrier for ALL its bytes, copy the positions [lo, hi) of it (whole page runs)
    auto copy_rows = [&](int st, int t0, int lo, int hi) {
        int a = lo;
        while (a < hi) {
            int pg = a / bt, in = a % bt;
            int b = min(hi, (pg + 1) * bt);
            int page = sh.pages[pg];
            size_t off = (((size_t)page * p.n_kv + h) * bt + in) * HD;
            uint32_t bytes = (uint32_t)(b - a) * HD * 2;
            tma_load_1d(&sh.k[st][a - t0][0], p.k_cache + off, bytes, &full[st]);
            tma_load_1d(&sh.v[st][a - t0][0], p.v_cache + off, bytes, &full[st]);
            a = b;
        }
    };
    auto issue = [&](int ti, int excl) {      // positions >= excl are left for a later copy_rows on the same barrier
        int st = ti % AT_NST;
        int t0 = p0 + ti * AT_TT, t1 = min(p1, t0 + AT_TT);
        mbar_arrive_expect_tx(&full[st], (uint32_t)(t1 - t0) * HD * 2 * 2);
        copy_rows(st, t0, t0, min(t1, excl));
    };
    const int n_early = min(ntile, AT_NST - 1);      // tiles requested up front (with `early`: before the dependency wait)
    if (tid == 0)
        for (int ti = 0; ti < n_early; ti++) issue(ti, early ? pos : 0x7fffffff);
    if (early) {
        griddep_wait();     // q and the newest K/V row (position pos) are complete from here on
        if (rec_tl) tl_stamp(p.tl, 1);
        if (tid == 0 && pos >= p0 && pos < min(p1, p0 + n_early * AT_TT)) {
            const int ti = (pos - p0) / AT_TT;
            copy_rows(ti % AT_NST, p0 + ti * AT_TT, pos, pos + 1);
        }
    }

    // this lane's dims of every query head, scaled (written by other SMs: L2)
    constexpr int NG = 4 * NP;                    // head slots in registers
    float qr[NG][DPL];
#pragma unroll
    for (int g = 0; g < NG; g++) {
#pragma unroll
        for (int d = 0; d < DPL; d++) qr[g][d] = 0.f;
        if (g < G) {
            const float* qp = p.q + ((size_t)h * G + g) * HD + DPL * lane;
            if (DPL == 4) { const float4 t = __ldcg(reinterpret_cast<const float4*>(qp)); qr[g][0] = t.x; qr[g][1] = t.y; qr[g][2] = t.z; qr[g][DPL - 1] = t.w; }
            else { const float2 t = __ldcg(reinterpret_cast<const float2*>(qp)); qr[g][0] = t.x; qr[g][1] = t.y; }
#pragma unroll
            for (int d = 0; d < DPL; d++) qr[g][d] *= p.scale * 1.4426950408889634f;     // scores, maxima and merges live in the log2 domain (exp2f)
        }
    }
    if (B2_ATTN_STAMP == 1 && rec_tl && qr[0][0] != 12345.f) tl_stamp(p.tl, 2);
    float o[NG][DPL];
#pragma unroll
    for (int g = 0; g < NG; g++)
#pragma unroll
        for (int d = 0; d < DPL; d++) o[g][d] = 0.f;
    float m_run[NP], l_run[NP];                   // of head 4 * pass + lane / 8 (the same in the 8 lanes of a group)
#pragma unroll
    for (int pass = 0; pass < NP; pass++) { m_run[pass] = -INFINITY; l_run[pass] = 0.f; }
    const int gg_l = lane >> 3, jj_l = lane & 7;

    // rows of a tile as floats: this lane's DPL dims of the warp's 8 positions (rows past the tile's end read the last valid row: finite data, weight 0)
    auto load_rows = [&](const __half (*tile)[HD], int j0, int tl, float (&f)[AT_WT][DPL]) {
#pragma unroll
        for (int jj = 0; jj < AT_WT; jj++) {
            const __half* r = &tile[min(j0 + jj, tl - 1)][DPL * lane];
            if (DPL == 4) {
                const uint2 raw = *reinterpret_cast<const uint2*>(r);
                const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&raw.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
                f[jj][0] = a.x; f[jj][1] = a.y; f[jj][2] = b.x; f[jj][DPL - 1] = b.y;
            } else {
                const float2 a = __half22float2(*reinterpret_cast<const __half2*>(r));
                f[jj][0] = a.x; f[jj][1] = a.y;
            }
        }
    };

    for (int ti = 0; ti < ntile; ti++) {
        const int st = ti % AT_NST;
        const int t0 = p0 + ti * AT_TT, tl = min(p1, t0 + AT_TT) - t0;
        if (tid == 0 && ti + AT_NST - 1 < ntile && ti + AT_NST - 1 >= n_early) issue(ti + AT_NST - 1, 0x7fffffff);   // into the buffer tile ti - 1 left (barrier below)
        mbar_wait(&full[st], use[st] & 1, p.err, 0x20001u, (unsigned)ti);
        use[st]++;
        if (B2_ATTN_STAMP == 2 && rec_tl && ti == 0) tl_stamp(p.tl, 2);
        const int j0 = warp * AT_WT;
        {
            float rows[AT_WT][DPL];
            load_rows(sh.k[st], j0, tl, rows);
            float s[NP];
#pragma unroll
            for (int pass = 0; pass < NP; pass++) {
                float v[32];
#pragma unroll
                for (int gg = 0; gg < 4; gg++)
#pragma unroll
                    for (int jj = 0; jj < AT_WT; jj++) {
                        float t = 0.f;
#pragma unroll
                        for (int d = 0; d < DPL; d++) t += qr[4 * pass + gg][d] * rows[jj][d];
                        v[gg * 8 + jj] = t;
                    }
                const float tot = lanes_transpose_sum(v, lane);
                s[pass] = (j0 + jj_l < tl && 4 * pass + gg_l < G) ? tot : -INFINITY;
            }
            load_rows(sh.v[st], j0, tl, rows);
#pragma unroll
            for (int pass = 0; pass < NP; pass++) {
                float mx = s[pass];
                mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
                mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
                mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
                const float m_new = fmaxf(m_run[pass], mx);
                const float pj = (s[pass] == -INFINITY) ? 0.f : exp2f(s[pass] - m_new);
                float sum = pj;
                sum += __shfl_xor_sync(0xffffffffu, sum, 4);
                sum += __shfl_xor_sync(0xffffffffu, sum, 2);
                sum += __shfl_xor_sync(0xffffffffu, sum, 1);
                const float alpha = (m_run[pass] == -INFINITY) ? 0.f : exp2f(m_run[pass] - m_new);
                l_run[pass] = l_run[pass] * alpha + sum;
                m_run[pass] = m_new;
#pragma unroll
                for (int gg = 0; gg < 4; gg++) {
                    const float al = __shfl_sync(0xffffffffu, alpha, gg * 8);
#pragma unroll
                    for (int d = 0; d < DPL; d++) o[4 * pass + gg][d] *= al;
                }
#pragma unroll
                for (int jj = 0; jj < AT_WT; jj++)
#pragma unroll
                    for (int gg = 0; gg < 4; gg++) {
                        const float pv = __shfl_sync(0xffffffffu, pj, gg * 8 + jj);
#pragma unroll
                        for (int d = 0; d < DPL; d++) o[4 * pass + gg][d] += pv * rows[jj][d];
                    }
            }
        }
        fence_proxy_async();                   // this thread's reads of the tile before the next TMA (async proxy) rewrites it
        named_bar_sync(AT_BAR, AT_THREADS);    // the tile buffer is free again
        if (B2_ATTN_STAMP == 3 && rec_tl && ti == 0) tl_stamp(p.tl, 2);
    }
    if (B2_ATTN_STAMP == 4 && rec_tl) tl_stamp(p.tl, 2);

    // ---------------------------------------------------------------- merge the warps
    // phase ownership of the merges: a pair of output dims for a set of heads
    constexpr int NSET3 = AT_THREADS / (HD / 2);
    constexpr int HG3 = (AT_GMAX + NSET3 - 1) / NSET3;
    const int dpair = tid % (HD / 2), set3 = tid / (HD / 2);
    const int hg3 = (G + NSET3 - 1) / NSET3;          // heads per thread set
    float acc[HG3][2];
    {
        float* wo = reinterpret_cast<float*>(&sh.k[0][0][0]);       // [AT_WARPS][G][HD]; every tile has been consumed (barrier above)
        float* wm = &sh.sc[0][0][0];                                // [AT_WARPS][AT_GMAX]
        float* wl = wm + AT_WARPS * AT_GMAX;
#pragma unroll
        for (int g = 0; g < NG; g++) {
            if (g < G) {
                float* dst = wo + ((size_t)warp * G + g) * HD + DPL * lane;
                if (DPL == 4) *reinterpret_cast<float4*>(dst) = make_float4(o[g][0], o[g][1], o[g][2], o[g][DPL - 1]);
                else *reinterpret_cast<float2*>(dst) = make_float2(o[g][0], o[g][1]);
            }
        }
        if (jj_l == 0) {
#pragma unroll
            for (int pass = 0; pass < NP; pass++) {
                const int g = 4 * pass + gg_l;
                if (g < G) { wm[warp * AT_GMAX + g] = m_run[pass]; wl[warp * AT_GMAX + g] = l_run[pass]; }
            }
        }
        named_bar_sync(AT_BAR, AT_THREADS);
        const int nwu = ntile > 1 ? AT_WARPS : (p1 - p0 + AT_WT - 1) / AT_WT;      // warps that had positions (the others hold -inf / 0)
#pragma unroll
        for (int i = 0; i < HG3; i++) {
            acc[i][0] = acc[i][1] = 0.f;
            const int g = set3 * hg3 + i;
            if (i < hg3 && g < G) {
                float M = -INFINITY;
                for (int w = 0; w < nwu; w++) M = fmaxf(M, wm[w * AT_GMAX + g]);
                float l = 0.f;
                for (int w = 0; w < nwu; w++) {
                    const float mw = wm[w * AT_GMAX + g];
                    const float wgt = (mw == -INFINITY) ? 0.f : exp2f(mw - M);
                    const float2 ov = *reinterpret_cast<const float2*>(wo + ((size_t)w * G + g) * HD + 2 * dpair);
                    acc[i][0] += wgt * ov.x; acc[i][1] += wgt * ov.y;
                    l += wgt * wl[w * AT_GMAX + g];
                }
                if (dpair == 0) { sh.m_run[g] = M; sh.l_run[g] = l; }
            }
        }
        named_bar_sync(AT_BAR, AT_THREADS);     // m_run / l_run of the CTA are published; the staging in sc[] is free
    }

    // ---------------------------------------------------------------- merge the splits
    if (B2_ATTN_STAMP == 5 && rec_tl) tl_stamp(p.tl, 2);
    if (p.n_splits == 1) {
#pragma unroll
        for (int i = 0; i < HG3; i++) {
            int g = set3 * hg3 + i;
            if (i < hg3 && g < G) {
                float inv = 1.0f / sh.l_run[g];
                float* o = p.out + ((size_t)h * G + g) * HD + 2 * dpair;
                o[0] = acc[i][0] * inv; o[1] = acc[i][1] * inv;
            }
        }
        return;
    }
    if (CLUSTER) {
        // Every CTA of the cluster merges a slice of the G * HD outputs. Each split PUSHES its contribution to the
        // owner of every element (remote shared-memory stores), then ONE cluster barrier, then the owners merge from
        // their own shared memory: no remote loads, and nobody has to wait for its peers before leaving.
        cg::cluster_group cluster = cg::this_cluster();
        const int ns = p.n_splits;
        const int per = (G * HD + ns - 1) / ns;           // outputs merged by each CTA of the cluster
        const int rank = (int)cluster.block_rank();
#pragma unroll
        for (int i = 0; i < HG3; i++) {
            int g = set3 * hg3 + i;
            if (i < hg3 && g < G) {
#pragma unroll
                for (int k = 0; k < 2; k++) {
                    const int e = g * HD + 2 * dpair + k, r = e / per;
                    AttnShared<HD, AT_TT>* rs = cluster.map_shared_rank(&sh, r);
                    rs->part[rank * per + (e - r * per)] = acc[i][k];
                }
            }
        }
        for (int i = tid; i < G * ns; i += AT_THREADS) {
            const int g = i / ns, r = i % ns;
            AttnShared<HD, AT_TT>* rs = cluster.map_shared_rank(&sh, r);
            rs->part_ml[rank][g][0] = sh.m_run[g];
            rs->part_ml[rank][g][1] = sh.l_run[g];
        }
        cluster.sync();                                   // every split's contribution has landed in its owner's smem
        for (int e = rank * per + tid; e < min(G * HD, (rank + 1) * per); e += AT_THREADS) {
            const int g = e / HD;
            float M = -INFINITY;
#pragma unroll
            for (int r = 0; r < 8; r++) if (r < ns) M = fmaxf(M, sh.part_ml[r][g][0]);
            float num = 0.f, den = 0.f;
#pragma unroll
            for (int r = 0; r < 8; r++) {
                if (r < ns) {
                    const float ms = sh.part_ml[r][g][0];
                    const float w = (ms == -INFINITY) ? 0.f : exp2f(ms - M);
                    num += w * sh.part[r * per + (e - rank * per)]; den += w * sh.part_ml[r][g][1];
                }
            }
            p.out[(size_t)h * G * HD + e] = num / den;
        }
        return;
    }
    // ---- global-scratch merge (n_splits > 8)
    const int rec = HD + 2;
#pragma unroll
    for (int i = 0; i < HG3; i++) {
        int g = set3 * hg3 + i;
        if (i < hg3 && g < G) {
            float* r = p.scratch + (((size_t)h * p.n_splits + sp) * G + g) * rec;
            r[2 * dpair] = acc[i][0]; r[2 * dpair + 1] = acc[i][1];
            if (dpair == 0) { r[HD] = sh.m_run[g]; r[HD + 1] = sh.l_run[g]; }
        }
    }
    __threadfence();
    named_bar_sync(AT_BAR, AT_THREADS);
    if (tid == 0) sh.ticket = atomicAdd(&p.counters[h], 1u);
    named_bar_sync(AT_BAR, AT_THREADS);
    if (sh.ticket != (unsigned)(p.n_splits - 1)) return;
    __threadfence();    // last CTA of this KV head: every split's partial is visible
    // stage (m, l) of every split in shared memory with independent loads, then merge
    float* wsp = &sh.sc[0][0][0];                         // [G][n_splits] weights
    float* msp = &sh.sc[1][0][0];
    float* lsp = &sh.sc[2][0][0];
    for (int i = tid; i < G * p.n_splits; i += AT_THREADS) {
        int g = i / p.n_splits, s2 = i % p.n_splits;
        const float* r = p.scratch + (((size_t)h * p.n_splits + s2) * G + g) * rec;
        msp[i] = __ldcg(r + HD);
        lsp[i] = __ldcg(r + HD + 1);
    }
    named_bar_sync(AT_BAR, AT_THREADS);
    if (tid < G) {
        float M = -INFINITY, den = 0.f;
        for (int s2 = 0; s2 < p.n_splits; s2++) M = fmaxf(M, msp[tid * p.n_splits + s2]);
        for (int s2 = 0; s2 < p.n_splits; s2++) {
            float ms = msp[tid * p.n_splits + s2];
            float w = (ms == -INFINITY) ? 0.f : exp2f(ms - M);
            wsp[tid * p.n_splits + s2] = w;
            den += w * lsp[tid * p.n_splits + s2];
        }
        sh.l_run[tid] = 1.0f / den;      // (this CTA's own sum went to the scratch record above)
    }
    named_bar_sync(AT_BAR, AT_THREADS);
#pragma unroll
    for (int i = 0; i < HG3; i++) {
        int g = set3 * hg3 + i;
        if (i < hg3 && g < G) {
            const float* base = p.scratch + ((size_t)h * p.n_splits * G + g) * rec + 2 * dpair;
            float n0 = 0.f, n1 = 0.f;
#pragma unroll 8
            for (int s2 = 0; s2 < p.n_splits; s2++) {
                float2 a2 = __ldcg(reinterpret_cast<const float2*>(base + (size_t)s2 * G * rec));
                float w = wsp[g * p.n_splits + s2];
                n0 += w * a2.x; n1 += w * a2.y;
            }
            float inv = sh.l_run[g];
            float* o = p.out + ((size_t)h * G + g) * HD + 2 * dpair;
            o[0] = n0 * inv; o[1] = n1 * inv;
        }
    }
    if (tid == 0) p.counters[h] = 0;    // ready for the next launch
}


}  // namespace b2
