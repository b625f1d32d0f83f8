// Prompt-processing (prefill) kernels around the tensor-core GEMM (gemm.cu): T tokens of one sequence at once.
// Replaces the batch > 1 side of ggml-cuda that llama.cpp runs for the prompt under create_chat_completion
// (/root/reference/api.py:55-63): get_rows (K1), rms_norm_f32 (K2), rope_f32 + the KV-cache copy (K6/K7),
// the attention mat-muls + soft_max (K8-K10 at batch T), silu/mul (K11).
//
// f16 activation matrices that feed the GEMM are written in its operand order [K/8][T_pad][8]; the GEMM's
// outputs arrive as n_parts K-split partial sums that the consumer adds in a fixed order.
//   embed_rows      x[t][:]      = dequantised row tokens[t] of the embedding table          (f32)
//   resid_norm_half x[t] += sum(parts); xn[t] = f16( x[t] / sqrt(mean(x[t]^2) + eps) * w )
//   rope_store      sum(parts) -> q rotated f16 [T][n_head*hd]; k rotated, v -> paged f16 KV cache
//   attn_prefill    causal flash attention over the paged cache, mma.sync m16n8k16 (f16 in, f32 accumulate),
//                   one warp per 16 queries of one head, the G heads of a KV group share the K/V tiles
//   silu_mul_half   h[t][:]      = f16( silu(gate) * up )
//   prefill_finish  pushes the prompt tokens into the penalty history and leaves token / position of the
//                   last prompt token in the slot registers
#include "kernels.h"
#include "ptx.cuh"
#include "quant_blocks.cuh"
#include <cuda_fp16.h>

namespace b2 {

// ---------------------------------------------------------------- embedding rows
__global__ void embed_rows_kernel(int type, const uint8_t* table, int K, const int* tokens, float* x) {
    griddep_launch();
    griddep_wait();
    const int t = blockIdx.x;
    const int be = block_elems(type), bb = block_bytes(type);
    const uint8_t* row = table + (size_t)tokens[t] * (K / be) * bb;
    for (int i = threadIdx.x; i < K; i += blockDim.x)
        x[(size_t)t * K + i] = dequant_elem(type, row + (size_t)(i / be) * bb, i % be);
}
cudaError_t embed_rows_launch(int type, const uint8_t* table, int K, const int* tokens, int T, float* x, cudaStream_t s) {
    return launch_chain(embed_rows_kernel, dim3(T), dim3(256), 0, s, type, table, K, tokens, x);
}

__device__ __forceinline__ size_t op_index(int t, int k, int T_pad) { return ((size_t)(k >> 3) * T_pad + t) * 8 + (k & 7); }

// ---------------------------------------------------------------- residual add + RMSNorm -> f16
constexpr int RN_THREADS = 1024;
__global__ void __launch_bounds__(RN_THREADS) resid_norm_half_kernel(float* x, const float* parts, int n_parts, long part_stride,
                                                                     const float* w, float eps, int K, int T_pad, __half* xn) {
    griddep_launch();
    griddep_wait();
    __shared__ float red[RN_THREADS / 32];
    const int t = blockIdx.x;
    float* xr = x + (size_t)t * K;
    // a thread's float4s stay in registers between the sum of squares and the normalisation (K <= 2 * 4 * RN_THREADS = 8192;
    // longer rows re-read what they wrote)
    constexpr int KEEP = 2;
    float4 keep[KEEP];
    float ss = 0.f;
    auto sum_parts = [&](int i) -> float4 {                             // float4 per thread: the split sums are independent loads
        float4 v = *reinterpret_cast<const float4*>(xr + 4 * i);
        const float* pp = parts + (size_t)t * K + 4 * i;
#pragma unroll 4
        for (int s = 0; s < n_parts; s++) {
            const float4 q = *reinterpret_cast<const float4*>(pp + (size_t)s * part_stride);
            v.x += q.x; v.y += q.y; v.z += q.z; v.w += q.w;
        }
        if (n_parts) *reinterpret_cast<float4*>(xr + 4 * i) = v;
        ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
        return v;
    };
#pragma unroll
    for (int it = 0; it < KEEP; it++) {
        const int i = threadIdx.x + it * RN_THREADS;
        keep[it] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < K / 4) keep[it] = sum_parts(i);
    }
    for (int i = threadIdx.x + KEEP * RN_THREADS; i < K / 4; i += RN_THREADS) sum_parts(i);
    if (!w) return;
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
#pragma unroll
    for (int i = 0; i < RN_THREADS / 32; i++) tot += red[i];
    const float r = rsqrtf(tot / (float)K + eps);
    auto emit = [&](int i, const float4& a) {                           // 4 consecutive k = half of a 16-byte run in operand order
        const float4 wa = *reinterpret_cast<const float4*>(w + 4 * i);
        const __half2 h0 = __floats2half2_rn(a.x * r * wa.x, a.y * r * wa.y), h1 = __floats2half2_rn(a.z * r * wa.z, a.w * r * wa.w);
        *reinterpret_cast<uint2*>(xn + ((size_t)(i >> 1) * T_pad + t) * 8 + 4 * (i & 1)) =
            make_uint2(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1));
    };
#pragma unroll
    for (int it = 0; it < KEEP; it++) {
        const int i = threadIdx.x + it * RN_THREADS;
        if (i < K / 4) emit(i, keep[it]);
    }
    for (int i = threadIdx.x + KEEP * RN_THREADS; i < K / 4; i += RN_THREADS) emit(i, *reinterpret_cast<const float4*>(xr + 4 * i));
}
cudaError_t resid_norm_half_launch(float* x, const float* parts, int n_parts, long part_stride, const float* w, float eps,
                                   int K, int T, int T_pad, __half* xn, cudaStream_t s) {
    return launch_chain(resid_norm_half_kernel, dim3(T), dim3(RN_THREADS), 0, s, x, parts, n_parts, part_stride, w, eps, K, T_pad, xn);
}

// ---------------------------------------------------------------- RoPE + KV store
// parts: [T][n_head*hd | n_kv*hd | n_kv*hd] f32. Pair (2i, 2i+1) rotated by pos * freq[i] (ggml "norm" RoPE).
__global__ void rope_store_kernel(const float* parts, int n_parts, long part_stride, int T, int pos0, int n_head, int n_kv,
                                  int hd, int bt, const int* block_table, const float* rope_freq, __half* qh,
                                  __half* k_cache, __half* v_cache) {
    griddep_launch();
    griddep_wait();
    const int t = blockIdx.x, pos = pos0 + t;
    const int nq = n_head * hd, nk = n_kv * hd, row_len = nq + 2 * nk;
    const float* r = parts + (size_t)t * row_len;
    const int page = block_table[pos / bt], slot = pos % bt;
    for (int pi = threadIdx.x; pi < row_len / 2; pi += blockDim.x) {
        const int e = 2 * pi;
        float v0 = 0.f, v1 = 0.f;
#pragma unroll 4
        for (int s = 0; s < n_parts; s++) {
            const float2 v = *reinterpret_cast<const float2*>(r + (size_t)s * part_stride + e);
            v0 += v.x; v1 += v.y;
        }
        if (e < nq + nk) {
            const int i = (e % hd) >> 1;
            float sn, cs;
            sincosf((float)pos * rope_freq[i], &sn, &cs);
            const float a = v0 * cs - v1 * sn, b = v0 * sn + v1 * cs;
            v0 = a; v1 = b;
        }
        const __half2 h = __floats2half2_rn(v0, v1);
        if (e < nq) {
            *reinterpret_cast<__half2*>(qh + (size_t)t * nq + e) = h;
        } else {
            const int ee = e < nq + nk ? e - nq : e - nq - nk;
            const int kvh = ee / hd, d = ee % hd;
            const size_t idx = (((size_t)page * n_kv + kvh) * bt + slot) * hd + d;
            *reinterpret_cast<__half2*>((e < nq + nk ? k_cache : v_cache) + idx) = h;
        }
    }
}
cudaError_t rope_store_launch(const float* parts, int n_parts, long part_stride, int T, int pos0, int n_head, int n_kv, int hd,
                              int bt, const int* block_table, const float* rope_freq, __half* qh, __half* k_cache,
                              __half* v_cache, cudaStream_t s) {
    return launch_chain(rope_store_kernel, dim3(T), dim3(1024), 0, s, parts, n_parts, part_stride, T, pos0, n_head, n_kv, hd, bt, block_table, rope_freq, qh, k_cache, v_cache);
}

// ---------------------------------------------------------------- SiLU(gate) * up -> f16
__global__ void silu_mul_half_kernel(const float* parts, int n_parts, long part_stride, int F, int T, int T_pad, __half* h) {
    griddep_launch();
    griddep_wait();
    // thread <-> (8 consecutive f, token): one 16-byte store in operand order
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int t = (int)(i % T), ch = (int)(i / T);
    if (ch >= F / 8) return;
    float g[8] = {0, 0, 0, 0, 0, 0, 0, 0}, u[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int s = 0; s < n_parts; s++) {
        const float* row = parts + (size_t)s * part_stride + (size_t)t * 2 * F + ch * 8;
        const float4 g0 = *reinterpret_cast<const float4*>(row), g1 = *reinterpret_cast<const float4*>(row + 4);
        const float4 u0 = *reinterpret_cast<const float4*>(row + F), u1 = *reinterpret_cast<const float4*>(row + F + 4);
        g[0] += g0.x; g[1] += g0.y; g[2] += g0.z; g[3] += g0.w; g[4] += g1.x; g[5] += g1.y; g[6] += g1.z; g[7] += g1.w;
        u[0] += u0.x; u[1] += u0.y; u[2] += u0.z; u[3] += u0.w; u[4] += u1.x; u[5] += u1.y; u[6] += u1.z; u[7] += u1.w;
    }
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const float a = g[2 * j] / (1.f + __expf(-g[2 * j])) * u[2 * j], b = g[2 * j + 1] / (1.f + __expf(-g[2 * j + 1])) * u[2 * j + 1];
        const __half2 hh = __floats2half2_rn(a, b);
        o[j] = *reinterpret_cast<const uint32_t*>(&hh);
    }
    *reinterpret_cast<uint4*>(h + ((size_t)ch * T_pad + t) * 8) = make_uint4(o[0], o[1], o[2], o[3]);
}
cudaError_t silu_mul_half_launch(const float* parts, int n_parts, long part_stride, int F, int T, int T_pad, __half* h, cudaStream_t s) {
    const long n = (long)T * (F / 8);
    return launch_chain(silu_mul_half_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, s, parts, n_parts, part_stride, F, T, T_pad, h);
}

// ---------------------------------------------------------------- batched decode step: one token per slot
__global__ void embed_batch_kernel(int type, const uint8_t* table, int K, const SlotDev* tab, const BatchIdx b, float* x) {
    griddep_launch();
    griddep_wait();
    const int t = blockIdx.x;
    const int tok = tab[b.slot[t]].ints[0];
    const int be = block_elems(type), bb = block_bytes(type);
    const uint8_t* row = table + (size_t)tok * (K / be) * bb;
    for (int i = threadIdx.x; i < K; i += blockDim.x)
        x[(size_t)t * K + i] = dequant_elem(type, row + (size_t)(i / be) * bb, i % be);
}
cudaError_t embed_batch_launch(int type, const uint8_t* table, int K, const SlotDev* tab, const BatchIdx& b, float* x, cudaStream_t s) {
    return launch_chain(embed_batch_kernel, dim3(b.n), dim3(256), 0, s, type, table, K, tab, b, x);
}

// like rope_store, but row t belongs to slot b.slot[t] at that slot's own position; q stays float32 for flash-decode
__global__ void rope_store_batch_kernel(const float* parts, int n_parts, long part_stride, const SlotDev* tab, const BatchIdx b,
                                        int n_head, int n_kv, int hd, int bt, const float* rope_freq, float* q_out,
                                        __half* k_cache, __half* v_cache) {
    griddep_launch();
    griddep_wait();
    const int t = blockIdx.x;
    const SlotDev sd = tab[b.slot[t]];
    const int pos = sd.ints[1];
    const int nq = n_head * hd, nk = n_kv * hd, row_len = nq + 2 * nk;
    const float* r = parts + (size_t)t * row_len;
    const int page = sd.block_table[pos / bt], slot = pos % bt;
    for (int pi = threadIdx.x; pi < row_len / 2; pi += blockDim.x) {
        const int e = 2 * pi;
        float v0 = 0.f, v1 = 0.f;
#pragma unroll 4
        for (int s = 0; s < n_parts; s++) {
            const float2 v = *reinterpret_cast<const float2*>(r + (size_t)s * part_stride + e);
            v0 += v.x; v1 += v.y;
        }
        if (e < nq + nk) {
            const int i = (e % hd) >> 1;
            float sn, cs;
            sincosf((float)pos * rope_freq[i], &sn, &cs);
            const float a = v0 * cs - v1 * sn, c = v0 * sn + v1 * cs;
            v0 = a; v1 = c;
        }
        if (e < nq) {
            *reinterpret_cast<float2*>(q_out + (size_t)t * nq + e) = make_float2(v0, v1);
        } else {
            const int ee = e < nq + nk ? e - nq : e - nq - nk;
            const int kvh = ee / hd, d = ee % hd;
            const size_t idx = (((size_t)page * n_kv + kvh) * bt + slot) * hd + d;
            *reinterpret_cast<__half2*>((e < nq + nk ? k_cache : v_cache) + idx) = __floats2half2_rn(v0, v1);
        }
    }
}
cudaError_t rope_store_batch_launch(const float* parts, int n_parts, long part_stride, const SlotDev* tab, const BatchIdx& b,
                                    int n_head, int n_kv, int hd, int bt, const float* rope_freq, float* q_out,
                                    __half* k_cache, __half* v_cache, cudaStream_t s) {
    return launch_chain(rope_store_batch_kernel, dim3(b.n), dim3(1024), 0, s, parts, n_parts, part_stride, tab, b, n_head, n_kv, hd, bt, rope_freq, q_out, k_cache, v_cache);
}

// ---------------------------------------------------------------- slot registers after a prompt chunk
__global__ void prefill_finish_kernel(const int* prompt, int n, int pos_last, int* ints, SamplerState* st) {
    griddep_launch();
    griddep_wait();
    int nh = st->n_hist;
    for (int i = 0; i < n; i++) { st->history[nh & (HISTORY_LEN - 1)] = prompt[i]; nh++; }
    st->n_hist = nh;
    ints[0] = prompt[n - 1];
    ints[1] = pos_last;
}
cudaError_t prefill_finish_launch(const int* prompt, int n, int pos_last, int* ints, SamplerState* st, cudaStream_t s) {
    return launch_chain(prefill_finish_kernel, dim3(1), dim3(1), 0, s, prompt, n, pos_last, ints, st);
}

// ---------------------------------------------------------------- causal attention for T queries
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldmatrix_x2_trans(uint32_t& r0, uint32_t& r1, const void* p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0,%1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(smem_u32(p)));
}
__device__ __forceinline__ uint32_t pack_half2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

constexpr int PA_KT = 64;        // keys per tile

// grid (query tiles, n_kv); block = G * QW warps: warp w -> head kvh*G + w % G, queries 16 * (w / G) .. +15 of the tile
template <int HD>
__global__ void __launch_bounds__(256) attn_prefill_kernel(const __half* qh, const __half* k_cache, const __half* v_cache,
                                                           const int* block_table, int T, int T_pad, int pos0, int n_head,
                                                           int n_kv, int bt, int QW, float scale_log2, __half* out) {
    griddep_launch();
    griddep_wait();
    constexpr int LD = HD + 8;                                  // padded row: conflict-free fragment loads
    extern __shared__ __align__(16) uint8_t smem_raw[];
    __half* ks = reinterpret_cast<__half*>(smem_raw);           // [PA_KT][LD]
    __half* vs = ks + PA_KT * LD;
    const int G = n_head / n_kv;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, tq = lane & 3;
    const int kvh = blockIdx.y, head = kvh * G + warp % G;
    const int q0 = blockIdx.x * 16 * QW + 16 * (warp / G);      // first query (chunk-relative) of this warp
    const int nq = n_head * HD;
    const bool active = q0 < T;

    // Q fragments: rows q0+g and q0+g+8 (clamped; rows >= T are dropped at the end)
    uint32_t qa[HD / 16][4];
    {
        const int r0 = min(q0 + g, T - 1), r1 = min(q0 + g + 8, T - 1);
        const __half* p0 = qh + (size_t)r0 * nq + head * HD, * p1 = qh + (size_t)r1 * nq + head * HD;
#pragma unroll
        for (int kk = 0; kk < HD / 16; kk++) {
            qa[kk][0] = *reinterpret_cast<const uint32_t*>(p0 + kk * 16 + 2 * tq);
            qa[kk][1] = *reinterpret_cast<const uint32_t*>(p1 + kk * 16 + 2 * tq);
            qa[kk][2] = *reinterpret_cast<const uint32_t*>(p0 + kk * 16 + 8 + 2 * tq);
            qa[kk][3] = *reinterpret_cast<const uint32_t*>(p1 + kk * 16 + 8 + 2 * tq);
        }
    }
    float o[HD / 8][4];
#pragma unroll
    for (int i = 0; i < HD / 8; i++) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;   // running max / sum of rows g, g+8 (log2 domain)

    // keys needed by the CTA: positions 0 .. pos0 + (last query of the tile)
    const int last_q = min(blockIdx.x * 16 * QW + 16 * QW - 1, T - 1);
    const int n_keys = pos0 + last_q + 1;
    const int my_last = pos0 + min(q0 + 15, T - 1);             // last key this warp needs
    for (int k0 = 0; k0 < n_keys; k0 += PA_KT) {
        __syncthreads();                                        // previous tile fully consumed
        for (int i = threadIdx.x; i < PA_KT * (HD / 8); i += blockDim.x) {
            const int j = i / (HD / 8), ch = i % (HD / 8);
            const int pos = k0 + j;
            uint4 kv4 = make_uint4(0, 0, 0, 0), vv4 = kv4;
            if (pos < n_keys) {
                const size_t base = (((size_t)block_table[pos / bt] * n_kv + kvh) * bt + pos % bt) * HD + ch * 8;
                kv4 = *reinterpret_cast<const uint4*>(k_cache + base);
                vv4 = *reinterpret_cast<const uint4*>(v_cache + base);
            }
            *reinterpret_cast<uint4*>(ks + j * LD + ch * 8) = kv4;
            *reinterpret_cast<uint4*>(vs + j * LD + ch * 8) = vv4;
        }
        __syncthreads();
        if (!active || k0 > my_last) continue;
        // ---- S = Q K^T for 64 keys: 8 n-tiles of 8 keys
        float sc[PA_KT / 8][4];
#pragma unroll
        for (int nt = 0; nt < PA_KT / 8; nt++) {
            sc[nt][0] = sc[nt][1] = sc[nt][2] = sc[nt][3] = 0.f;
            const __half* kr = ks + (nt * 8 + g) * LD + 2 * tq;
#pragma unroll
            for (int kk = 0; kk < HD / 16; kk++) {
                const uint32_t b0 = *reinterpret_cast<const uint32_t*>(kr + kk * 16);
                const uint32_t b1 = *reinterpret_cast<const uint32_t*>(kr + kk * 16 + 8);
                mma16816(sc[nt], qa[kk], b0, b1);
            }
        }
        // ---- causal mask + online softmax (rows g and g+8; columns 2tq, 2tq+1 of every n-tile)
        const int p_row0 = pos0 + q0 + g, p_row1 = p_row0 + 8;
        float mx0 = m0, mx1 = m1;
#pragma unroll
        for (int nt = 0; nt < PA_KT / 8; nt++) {
            const int kp = k0 + nt * 8 + 2 * tq;
            sc[nt][0] = kp <= p_row0 ? sc[nt][0] * scale_log2 : -INFINITY;
            sc[nt][1] = kp + 1 <= p_row0 ? sc[nt][1] * scale_log2 : -INFINITY;
            sc[nt][2] = kp <= p_row1 ? sc[nt][2] * scale_log2 : -INFINITY;
            sc[nt][3] = kp + 1 <= p_row1 ? sc[nt][3] * scale_log2 : -INFINITY;
            mx0 = fmaxf(mx0, fmaxf(sc[nt][0], sc[nt][1]));
            mx1 = fmaxf(mx1, fmaxf(sc[nt][2], sc[nt][3]));
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        // key 0 is visible to every query, so after the first tile the maxima are finite
        const float f0 = exp2f(m0 - mx0), f1 = exp2f(m1 - mx1);
        m0 = mx0; m1 = mx1;
        float s0 = 0.f, s1 = 0.f;
        uint32_t pa[PA_KT / 16][4];
#pragma unroll
        for (int nt = 0; nt < PA_KT / 8; nt++) {
            const float e0 = exp2f(sc[nt][0] - mx0), e1 = exp2f(sc[nt][1] - mx0);
            const float e2 = exp2f(sc[nt][2] - mx1), e3 = exp2f(sc[nt][3] - mx1);
            s0 += e0 + e1; s1 += e2 + e3;
            pa[nt >> 1][(nt & 1) * 2 + 0] = pack_half2(e0, e1);
            pa[nt >> 1][(nt & 1) * 2 + 1] = pack_half2(e2, e3);
        }
        l0 = l0 * f0 + s0; l1 = l1 * f1 + s1;
#pragma unroll
        for (int i = 0; i < HD / 8; i++) { o[i][0] *= f0; o[i][1] *= f0; o[i][2] *= f1; o[i][3] *= f1; }
        // ---- O += P V : k = keys (4 k-tiles of 16), n = head dims (HD / 8 n-tiles)
#pragma unroll
        for (int kt = 0; kt < PA_KT / 16; kt++) {
#pragma unroll
            for (int nt = 0; nt < HD / 8; nt++) {
                uint32_t b0, b1;
                // 16 keys x 8 dims, transposed on load: lanes 0-15 give the row addresses
                ldmatrix_x2_trans(b0, b1, vs + (kt * 16 + (lane & 15)) * LD + nt * 8);
                mma16816(o[nt], pa[kt], b0, b1);
            }
        }
    }
    if (!active) return;
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = 1.f / l0, i1 = 1.f / l1;
    const int r0 = q0 + g, r1 = q0 + g + 8;
#pragma unroll
    for (int nt = 0; nt < HD / 8; nt++) {
        const int d = head * HD + nt * 8 + 2 * tq;      // column of the [T][n_head*hd] matrix, stored in operand order
        if (r0 < T) *reinterpret_cast<uint32_t*>(out + op_index(r0, d, T_pad)) = pack_half2(o[nt][0] * i0, o[nt][1] * i0);
        if (r1 < T) *reinterpret_cast<uint32_t*>(out + op_index(r1, d, T_pad)) = pack_half2(o[nt][2] * i1, o[nt][3] * i1);
    }
}

cudaError_t attn_prefill_launch(const __half* qh, const __half* k_cache, const __half* v_cache, const int* block_table,
                                int T, int T_pad, int pos0, int n_head, int n_kv, int hd, int bt, __half* out, cudaStream_t s) {
    if (n_kv <= 0 || n_head % n_kv) return cudaErrorInvalidValue;
    const int G = n_head / n_kv;
    if (G > 8) return cudaErrorInvalidValue;
    const int QW = G >= 4 ? 1 : 4 / G;
    const int q_tile = 16 * QW;
    dim3 grid((T + q_tile - 1) / q_tile, n_kv), block(32 * G * QW);
    const float scale_log2 = 1.4426950408889634f / sqrtf((float)hd);
    const size_t smem = (size_t)2 * PA_KT * (hd + 8) * 2;
    if (hd == 64) return launch_chain(attn_prefill_kernel<64>, grid, block, smem, s, qh, k_cache, v_cache, block_table, T, T_pad, pos0, n_head, n_kv, bt, QW, scale_log2, out);
    if (hd == 128) return launch_chain(attn_prefill_kernel<128>, grid, block, smem, s, qh, k_cache, v_cache, block_table, T, T_pad, pos0, n_head, n_kv, bt, QW, scale_log2, out);
    return cudaErrorInvalidValue;
}

}  // namespace b2
