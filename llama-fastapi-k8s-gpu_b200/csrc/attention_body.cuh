// Flash-decode attention: shared-memory layout and the per-CTA body (see attention.cu for the design
// notes). Shared by the stand-alone kernel (attention.cu) and the persistent decode kernel (mega.cu).
#pragma once
#include "kernels.h"
#include "ptx.cuh"
#include <cooperative_groups.h>
#include <math.h>

namespace cg = cooperative_groups;

namespace b2 {

constexpr int AT_THREADS = 256;
constexpr int AT_TT = 64;      // positions per tile
constexpr int AT_GMAX = 8;     // query heads per KV head
constexpr int AT_MAX_PAGES = 512;
constexpr int AT_BAR = 2;      // named barrier used by the 256 attention threads

template <int HD>
struct __align__(16) AttnShared {
    __half k[2][AT_TT][HD];
    __half v[2][AT_TT][HD];
    float q[AT_GMAX][HD];
    float part[AT_GMAX * HD + 8];      // cluster merge: slice [r * per, (r+1) * per) = what split r sent us (its un-normalised outputs of OUR elements)
    float part_ml[8][AT_GMAX][2];      // cluster merge: (running max, running sum) of every head from every split
    float sc[4][AT_GMAX][AT_TT];       // per head_dim-quarter partial scores; sc[0] then holds p
    float m_run[AT_GMAX], l_run[AT_GMAX], alpha[AT_GMAX];
    int pages[AT_MAX_PAGES];
    unsigned int ticket;
};


// The whole per-CTA job: split `sp` of KV head `h`. Runs on AT_THREADS threads (tid 0..255) that
// synchronise on named barrier AT_BAR, so it can be a kernel of its own or a phase of the persistent
// decode kernel (mega.cu). `full` are two mbarriers that live across calls; `use` counts how often
// each has completed (the caller keeps it; every thread holds the same value).
// `early`: the body is entered BEFORE griddepcontrol.wait (programmatic dependent launch). Position and page table are
// already valid then (the kernel that writes q and this token's K/V row lets us start only after ITS dependency wait,
// MvParams::trigger_late), so the K/V tiles of the older positions are requested first and the wait happens inside,
// just before q and the newest row are needed: the copies overlap the tail of the QKV kernel.
template <int HD, bool CLUSTER>
__device__ __forceinline__ void attn_decode_body(const AttnParams& p, AttnShared<HD>& sh, uint64_t* full, uint32_t (&use)[2],
                                                 int sp, int h, int tid, bool early = false) {
    const int warp = tid >> 5, lane = tid & 31;
    const int G = p.n_head / p.n_kv;
    const int bt = p.block_tokens;

    if (tid < AT_GMAX) { sh.m_run[tid] = -INFINITY; sh.l_run[tid] = 0.f; sh.alpha[tid] = 0.f; }

    // one round trip: position and page table (and q, unless it is still being written)
    const int pos = *p.pos_ptr;
    for (int i = tid; i < p.max_pages && i < AT_MAX_PAGES; i += AT_THREADS) sh.pages[i] = p.block_table[i];
    if (!early)
        for (int i = tid; i < G * HD; i += AT_THREADS) sh.q[i / HD][i % HD] = __ldcg(p.q + (size_t)h * G * HD + i);   // written by other SMs
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
        int st = ti & 1;
        int t0 = p0 + ti * AT_TT, t1 = min(p1, t0 + AT_TT);
        mbar_arrive_expect_tx(&full[st], (uint32_t)(t1 - t0) * HD * 2 * 2);
        copy_rows(st, t0, t0, min(t1, excl));
    };
    const int n_early = early ? min(ntile, 2) : min(ntile, 1);      // tiles requested up front
    if (tid == 0)
        for (int ti = 0; ti < n_early; ti++) issue(ti, early ? pos : 0x7fffffff);
    if (early) {
        griddep_wait();     // q and the newest K/V row (position pos) are complete from here on
        if (tid == 0) tl_stamp(p.tl, 1);
        for (int i = tid; i < G * HD; i += AT_THREADS) sh.q[i / HD][i % HD] = __ldcg(p.q + (size_t)h * G * HD + i);
        if (tid == 0 && pos >= p0 && pos < min(p1, p0 + n_early * AT_TT)) {
            const int ti = (pos - p0) / AT_TT;
            copy_rows(ti & 1, p0 + ti * AT_TT, pos, pos + 1);
        }
        named_bar_sync(AT_BAR, AT_THREADS);
    }

    // phase-3 ownership: a pair of output dims for a set of heads
    constexpr int NSET3 = AT_THREADS / (HD / 2);
    constexpr int HG3 = (AT_GMAX + NSET3 - 1) / NSET3;
    const int dpair = tid % (HD / 2), set3 = tid / (HD / 2);
    const int hg3 = (G + NSET3 - 1) / NSET3;          // heads per thread set
    float acc[HG3][2];
#pragma unroll
    for (int i = 0; i < HG3; i++) acc[i][0] = acc[i][1] = 0.f;

    for (int ti = 0; ti < ntile; ti++) {
        const int st = ti & 1;
        const int t0 = p0 + ti * AT_TT, tl = min(p1, t0 + AT_TT) - t0;
        if (tid == 0 && ti + 1 < ntile && ti + 1 >= n_early) issue(ti + 1, 0x7fffffff);
        mbar_wait(&full[st], use[st] & 1, p.err, 0x20001u, (unsigned)ti);
        use[st]++;

        // ---- phase 1: partial scores over a quarter of head_dim, all G heads
        {
            const int j = tid & 63, dq = tid >> 6;
            constexpr int CPQ = HD / 32;                  // 16-byte chunks per quarter
            float s[AT_GMAX];
#pragma unroll
            for (int g = 0; g < AT_GMAX; g++) s[g] = 0.f;
            if (j < tl) {
                const uint4* kr = reinterpret_cast<const uint4*>(&sh.k[st][j][0]);
#pragma unroll
                for (int c = 0; c < CPQ; c++) {
                    int cc = dq * CPQ + ((c + j) & (CPQ - 1));   // rotate inside the quarter: fewer bank conflicts
                    uint4 kv = kr[cc];
                    const __half2* k2 = reinterpret_cast<const __half2*>(&kv);
                    float2 f0 = __half22float2(k2[0]), f1 = __half22float2(k2[1]);
                    float2 f2 = __half22float2(k2[2]), f3 = __half22float2(k2[3]);
#pragma unroll
                    for (int g = 0; g < AT_GMAX; g++) {
                        if (g < G) {
                            const float4* q4 = reinterpret_cast<const float4*>(&sh.q[g][cc * 8]);
                            float4 qa = q4[0], qb = q4[1];
                            s[g] += f0.x * qa.x + f0.y * qa.y + f1.x * qa.z + f1.y * qa.w
                                  + f2.x * qb.x + f2.y * qb.y + f3.x * qb.z + f3.y * qb.w;
                        }
                    }
                }
            }
#pragma unroll
            for (int g = 0; g < AT_GMAX; g++)
                if (g < G) sh.sc[dq][g][j] = s[g];
        }
        named_bar_sync(AT_BAR, AT_THREADS);
        // ---- phase 2: online softmax, one warp per head
        for (int g = warp; g < G; g += AT_THREADS / 32) {
            float v0 = (sh.sc[0][g][lane] + sh.sc[1][g][lane]) + (sh.sc[2][g][lane] + sh.sc[3][g][lane]);
            float v1 = (sh.sc[0][g][lane + 32] + sh.sc[1][g][lane + 32]) + (sh.sc[2][g][lane + 32] + sh.sc[3][g][lane + 32]);
            v0 = (lane < tl) ? v0 * p.scale : -INFINITY;
            v1 = (lane + 32 < tl) ? v1 * p.scale : -INFINITY;
            float mx = warp_max(fmaxf(v0, v1));
            float m_old = sh.m_run[g], m_new = fmaxf(m_old, mx);
            float e0 = (v0 == -INFINITY) ? 0.f : expf(v0 - m_new);
            float e1 = (v1 == -INFINITY) ? 0.f : expf(v1 - m_new);
            float sum = warp_sum(e0 + e1);
            sh.sc[0][g][lane] = e0; sh.sc[0][g][lane + 32] = e1;
            if (lane == 0) {
                float al = (m_old == -INFINITY) ? 0.f : expf(m_old - m_new);
                sh.alpha[g] = al;
                sh.l_run[g] = sh.l_run[g] * al + sum;
                sh.m_run[g] = m_new;
            }
        }
        named_bar_sync(AT_BAR, AT_THREADS);
        // ---- phase 3: o += p . v
#pragma unroll
        for (int i = 0; i < HG3; i++) {
            int g = set3 * hg3 + i;
            if (i < hg3 && g < G) { float al = sh.alpha[g]; acc[i][0] *= al; acc[i][1] *= al; }
        }
#pragma unroll 4
        for (int j = 0; j < tl; j++) {
            float2 vv = __half22float2(*reinterpret_cast<const __half2*>(&sh.v[st][j][2 * dpair]));
#pragma unroll
            for (int i = 0; i < HG3; i++) {
                int g = set3 * hg3 + i;
                if (i < hg3 && g < G) { float pj = sh.sc[0][g][j]; acc[i][0] += pj * vv.x; acc[i][1] += pj * vv.y; }
            }
        }
        fence_proxy_async();                   // this thread's reads of the tile before the next TMA (async proxy) rewrites it
        named_bar_sync(AT_BAR, AT_THREADS);    // tile buffers and sc[] are free again
    }

    // ---------------------------------------------------------------- merge the splits
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
                    AttnShared<HD>* rs = cluster.map_shared_rank(&sh, r);
                    rs->part[rank * per + (e - r * per)] = acc[i][k];
                }
            }
        }
        for (int i = tid; i < G * ns; i += AT_THREADS) {
            const int g = i / ns, r = i % ns;
            AttnShared<HD>* rs = cluster.map_shared_rank(&sh, r);
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
                    const float w = (ms == -INFINITY) ? 0.f : expf(ms - M);
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
            float w = (ms == -INFINITY) ? 0.f : expf(ms - M);
            wsp[tid * p.n_splits + s2] = w;
            den += w * lsp[tid * p.n_splits + s2];
        }
        sh.alpha[tid] = 1.0f / den;
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
            float inv = sh.alpha[g];
            float* o = p.out + ((size_t)h * G + g) * HD + 2 * dpair;
            o[0] = n0 * inv; o[1] = n1 * inv;
        }
    }
    if (tid == 0) p.counters[h] = 0;    // ready for the next launch
}


}  // namespace b2
