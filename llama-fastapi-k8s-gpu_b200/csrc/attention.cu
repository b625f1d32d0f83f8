// Split-K flash-decode attention over the paged f16 KV cache (one new token per sequence).
// Replaces ggml-cuda's mul_mat_vec_p021/nc (K.q), soft_max_f32 and the P.V mat-mul that the
// reference launches separately per layer, on a paged cache instead of
// llama.cpp's contiguous one (H7).
//
// Grid = (n_splits, n_kv_heads). A CTA owns one KV head and one slice of the context and serves
// all G = n_head / n_kv query heads of that KV head from a single read of K and V (GQA-aware).
// KV pages are [n_kv][block_tokens][head_dim] f16, so the tokens of one page for one KV head are
// one contiguous run: thread 0 brings 64-token tiles into shared memory with 1-D TMA bulk copies
// (double buffered, mbarrier complete_tx) while the other threads work on the previous tile:
//   phase 1  scores s[g][j] = scale * q_g . k_j          (thread <-> position, rotated 16 B reads)
//   phase 2  online softmax per head (warp per head)      running max / sum, rescale factor
//   phase 3  o_g += p[g][j] * v_j                         (thread <-> pair of output dims)
// Partials (m, l, o) go to scratch; the last CTA of a KV head to finish (atomic ticket) merges the
// splits and writes the float32 attention output, so no second kernel is needed.
//
// Roofline: HBM; algorithmic bytes per launch = 2 * n_kv * head_dim * 2 B * (pos + 1).
#include "kernels.h"
#include "ptx.cuh"
#include <math.h>

namespace b2 {

constexpr int AT_THREADS = 128;
constexpr int AT_TT = 64;      // positions per tile
constexpr int AT_GMAX = 8;     // query heads per KV head

template <int HD>
struct __align__(16) AttnShared {
    __half k[2][AT_TT][HD];
    __half v[2][AT_TT][HD];
    float q[AT_GMAX][HD];
    float sc[AT_GMAX][AT_TT];
    float m_run[AT_GMAX], l_run[AT_GMAX], alpha[AT_GMAX];
    uint64_t full[2];
    unsigned int ticket;
};

size_t attn_scratch_floats(int n_head, int n_kv, int head_dim, int n_splits) {
    (void)n_kv;
    return (size_t)n_head * n_splits * (head_dim + 2);
}

template <int HD>
__global__ void __launch_bounds__(AT_THREADS) attn_decode_kernel(const AttnParams p) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    AttnShared<HD>& sh = *reinterpret_cast<AttnShared<HD>*>(smem_raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int sp = blockIdx.x, h = blockIdx.y;
    const int G = p.n_head / p.n_kv;
    const int bt = p.block_tokens;

    if (tid == 0) { mbar_init(&sh.full[0], 1); mbar_init(&sh.full[1], 1); fence_mbar_init(); }
    if (tid < AT_GMAX) { sh.m_run[tid] = -INFINITY; sh.l_run[tid] = 0.f; sh.alpha[tid] = 0.f; }
    __syncthreads();
    griddep_launch();
    griddep_wait();     // q and this token's K/V come from the QKV kernel just before us

    const int ctx = *p.pos_ptr + 1;
    int chunk = (ctx + p.n_splits - 1) / p.n_splits;
    chunk = (chunk + 15) & ~15;
    const int p0 = min(ctx, sp * chunk), p1 = min(ctx, p0 + chunk);
    const int ntile = (p1 - p0 + AT_TT - 1) / AT_TT;

    for (int i = tid; i < G * HD; i += AT_THREADS) sh.q[i / HD][i % HD] = p.q[(size_t)h * G * HD + i];

    auto issue = [&](int ti) {      // thread 0 only
        int st = ti & 1;
        int t0 = p0 + ti * AT_TT, t1 = min(p1, t0 + AT_TT);
        mbar_arrive_expect_tx(&sh.full[st], (uint32_t)(t1 - t0) * HD * 2 * 2);
        int a = t0;
        while (a < t1) {
            int pg = a / bt, in = a % bt;
            int b = min(t1, (pg + 1) * bt);
            int page = p.block_table[pg];
            size_t off = (((size_t)page * p.n_kv + h) * bt + in) * HD;
            uint32_t bytes = (uint32_t)(b - a) * HD * 2;
            tma_load_1d(&sh.k[st][a - t0][0], p.k_cache + off, bytes, &sh.full[st]);
            tma_load_1d(&sh.v[st][a - t0][0], p.v_cache + off, bytes, &sh.full[st]);
            a = b;
        }
    };
    if (tid == 0 && ntile > 0) issue(0);

    // phase-3 ownership: a pair of output dims for a set of heads
    constexpr int NSET3 = AT_THREADS / (HD / 2);
    constexpr int HG3 = (AT_GMAX + NSET3 - 1) / NSET3;
    const int dpair = tid % (HD / 2), set3 = tid / (HD / 2);
    const int hg3 = (G + NSET3 - 1) / NSET3;          // heads per thread set
    float acc[HG3][2];
#pragma unroll
    for (int i = 0; i < HG3; i++) acc[i][0] = acc[i][1] = 0.f;
    __syncthreads();

    for (int ti = 0; ti < ntile; ti++) {
        const int st = ti & 1;
        const int t0 = p0 + ti * AT_TT, tl = min(p1, t0 + AT_TT) - t0;
        if (tid == 0 && ti + 1 < ntile) issue(ti + 1);
        mbar_wait(&sh.full[st], (ti >> 1) & 1, p.err, 0x20001u);

        // ---- phase 1: scores
        {
            constexpr int HGS = AT_GMAX / 2;
            const int hgs = (G + 1) >> 1;               // heads per thread half
            const int j = tid & 63, hs = tid >> 6;
            float s[HGS];
#pragma unroll
            for (int i = 0; i < HGS; i++) s[i] = 0.f;
            if (j < tl) {
                const uint4* kr = reinterpret_cast<const uint4*>(&sh.k[st][j][0]);
#pragma unroll 4
                for (int c = 0; c < HD / 8; c++) {
                    int cc = (c + j) & (HD / 8 - 1);      // rotate: conflict-free 16 B reads
                    uint4 kv = kr[cc];
                    const __half2* k2 = reinterpret_cast<const __half2*>(&kv);
                    float2 f0 = __half22float2(k2[0]), f1 = __half22float2(k2[1]);
                    float2 f2 = __half22float2(k2[2]), f3 = __half22float2(k2[3]);
#pragma unroll
                    for (int i = 0; i < HGS; i++) {
                        int g = hs * hgs + i;
                        if (i < hgs && g < G) {
                            const float4* q4 = reinterpret_cast<const float4*>(&sh.q[g][cc * 8]);
                            float4 qa = q4[0], qb = q4[1];
                            s[i] += f0.x * qa.x + f0.y * qa.y + f1.x * qa.z + f1.y * qa.w
                                  + f2.x * qb.x + f2.y * qb.y + f3.x * qb.z + f3.y * qb.w;
                        }
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < HGS; i++) {
                int g = hs * hgs + i;
                if (i < hgs && g < G) sh.sc[g][j] = (j < tl) ? s[i] * p.scale : -INFINITY;
            }
        }
        __syncthreads();
        // ---- phase 2: online softmax, one warp per head
        for (int g = warp; g < G; g += AT_THREADS / 32) {
            float v0 = sh.sc[g][lane], v1 = sh.sc[g][lane + 32];
            float mx = warp_max(fmaxf(v0, v1));
            float m_old = sh.m_run[g], m_new = fmaxf(m_old, mx);
            float e0 = (v0 == -INFINITY) ? 0.f : expf(v0 - m_new);
            float e1 = (v1 == -INFINITY) ? 0.f : expf(v1 - m_new);
            float sum = warp_sum(e0 + e1);
            sh.sc[g][lane] = e0; sh.sc[g][lane + 32] = e1;
            if (lane == 0) {
                float al = (m_old == -INFINITY) ? 0.f : expf(m_old - m_new);
                sh.alpha[g] = al;
                sh.l_run[g] = sh.l_run[g] * al + sum;
                sh.m_run[g] = m_new;
            }
        }
        __syncthreads();
        // ---- phase 3: o += p . v
#pragma unroll
        for (int i = 0; i < HG3; i++) {
            int g = set3 * hg3 + i;
            if (i < hg3 && g < G) { float al = sh.alpha[g]; acc[i][0] *= al; acc[i][1] *= al; }
        }
        for (int j = 0; j < tl; j++) {
            float2 vv = __half22float2(*reinterpret_cast<const __half2*>(&sh.v[st][j][2 * dpair]));
#pragma unroll
            for (int i = 0; i < HG3; i++) {
                int g = set3 * hg3 + i;
                if (i < hg3 && g < G) { float pj = sh.sc[g][j]; acc[i][0] += pj * vv.x; acc[i][1] += pj * vv.y; }
            }
        }
        __syncthreads();    // tile buffers and sc[] are free again
    }

    // ---------------------------------------------------------------- write-out
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
    __syncthreads();
    if (tid == 0) sh.ticket = atomicAdd(&p.counters[h], 1u);
    __syncthreads();
    if (sh.ticket != (unsigned)(p.n_splits - 1)) return;
    __threadfence();    // last CTA of this KV head: every split's partial is visible
#pragma unroll
    for (int i = 0; i < HG3; i++) {
        int g = set3 * hg3 + i;
        if (i < hg3 && g < G) {
            const float* base = p.scratch + ((size_t)h * p.n_splits * G + g) * rec;
            float M = -INFINITY;
            for (int s2 = 0; s2 < p.n_splits; s2++) M = fmaxf(M, __ldcg(base + (size_t)s2 * G * rec + HD));
            float n0 = 0.f, n1 = 0.f, den = 0.f;
            for (int s2 = 0; s2 < p.n_splits; s2++) {
                const float* r = base + (size_t)s2 * G * rec;
                float ms = __ldcg(r + HD);
                float w = (ms == -INFINITY) ? 0.f : expf(ms - M);
                n0 += w * __ldcg(r + 2 * dpair); n1 += w * __ldcg(r + 2 * dpair + 1);
                den += w * __ldcg(r + HD + 1);
            }
            float inv = 1.0f / den;
            float* o = p.out + ((size_t)h * G + g) * HD + 2 * dpair;
            o[0] = n0 * inv; o[1] = n1 * inv;
        }
    }
    if (tid == 0) p.counters[h] = 0;    // ready for the next launch
}

cudaError_t attn_init() {
    cudaError_t e = cudaFuncSetAttribute(attn_decode_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)sizeof(AttnShared<64>));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(attn_decode_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)sizeof(AttnShared<128>));
}

cudaError_t attn_decode_launch(const AttnParams& p, cudaStream_t stream, bool pdl) {
    if (p.n_head % p.n_kv || p.n_head / p.n_kv > AT_GMAX) return cudaErrorInvalidValue;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(p.n_splits, p.n_kv); cfg.blockDim = dim3(AT_THREADS); cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    if (p.head_dim == 64) {
        cfg.dynamicSmemBytes = sizeof(AttnShared<64>);
        return cudaLaunchKernelEx(&cfg, attn_decode_kernel<64>, p);
    } else if (p.head_dim == 128) {
        cfg.dynamicSmemBytes = sizeof(AttnShared<128>);
        return cudaLaunchKernelEx(&cfg, attn_decode_kernel<128>, p);
    }
    return cudaErrorInvalidValue;
}

}  // namespace b2
