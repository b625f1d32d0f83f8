// Flash-decode attention on the tensor cores, for the cases where the scalar kernel (attention.cu) is bound by
// instruction issue rather than latency: several sequences per step, and long contexts.
// Same job as attn_decode_kernel (K.q, soft_max, P.V of llama.cpp's decode step behind
// /root/reference/api.py:55-63) on the same paged f16 KV cache and the same split-K scratch protocol.
//
// Grid (n_splits, n_kv_heads, sequences), 4 warps. The MMA's 16 rows are the G <= 8 query heads of one KV head
// (rows G..15 are zero padding): one m16n8k16 covers 16 heads x 8 positions x 16 dims, ~12x fewer instructions per
// K/V byte than the scalar kernel. K/V tiles of 64 positions go global -> shared with cp.async (16 B per request,
// double buffered) into rows padded by 16 B, so the fragment loads (LDS.32 for K, ldmatrix.trans for V) are
// conflict-free; pages are looked up per position, so any kv_block_tokens works. Inside a tile warp w owns
// positions 16w..16w+15 end to end (scores, online softmax, P.V) -- no block barrier between the phases -- and the
// four per-warp (max, sum, output) triples are merged once at the end; splits are merged by the last CTA of the
// KV head (atomic ticket), in the log2 domain.
// q is rounded to f16 for the MMA (as llama.cpp's own flash-attention kernels do); accumulation is f32.
#include "kernels.h"
#include "ptx.cuh"
#include <cuda_fp16.h>
#include <math.h>

namespace b2 {

constexpr int AM_THREADS = 128;
constexpr int AM_TT = 64;           // positions per tile (16 per warp)

__device__ __forceinline__ void am_mma(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void am_ldmatrix_x2_trans(uint32_t& r0, uint32_t& r1, const void* p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0,%1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(smem_u32(p)));
}
__device__ __forceinline__ uint32_t am_pack(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ void am_cp16(void* dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
// one output element of sequence z: float vector, or f16 straight into the next GEMM's operand layout (no conversion kernel)
__device__ __forceinline__ void am_store_out(const AttnParams& p, int z, int col, float v) {
    if (p.out_h) p.out_h[((size_t)(col >> 3) * p.out_T_pad + z) * 8 + (col & 7)] = __float2half_rn(v);
    else p.out[col] = v;
}
__device__ __forceinline__ void am_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void am_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int HD>
__global__ void __launch_bounds__(AM_THREADS) attn_decode_mma_kernel(const AttnParams p0, const SlotDev* tab, const BatchIdx b,
                                                                     float scale_log2) {
    constexpr int LD = HD + 8;                                  // padded row (halves)
    extern __shared__ __align__(16) uint8_t smem_raw[];
    __half* ks = reinterpret_cast<__half*>(smem_raw);           // [2][AM_TT][LD]
    __half* vs = ks + 2 * AM_TT * LD;                           // [2][AM_TT][LD]
    __shared__ unsigned int s_ticket;
    griddep_launch();
    griddep_wait();                                             // q, this step's K/V and the slot registers come from earlier kernels

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, tq = lane & 3;
    const int sp = blockIdx.x, h = blockIdx.y, z = blockIdx.z;
    AttnParams p = p0;
    const int G = p.n_head / p.n_kv;
    const size_t qd = (size_t)p.n_head * HD;
    if (tab) {
        const SlotDev sd = tab[b.slot[z]];
        p.q = p0.q + z * qd; p.out = p0.out + z * qd;
        p.pos_ptr = sd.ints + 1; p.block_table = sd.block_table;
        p.scratch = p0.scratch + (size_t)z * p.n_head * p.n_splits * (HD + 2);
        p.counters = p0.counters + z * p.n_kv;
    }
    const int ctx = *p.pos_ptr + 1;
    const int chunk = (ctx + p.n_splits - 1) / p.n_splits;
    const int q0 = min(ctx, sp * chunk), q1 = min(ctx, q0 + chunk);      // positions [q0, q1) belong to this split
    const int ntile = (q1 - q0 + AM_TT - 1) / AM_TT;
    const int bt = p.block_tokens;

    // ---- q fragments: row g = head h*G + g (rows >= G and rows 8..15 are zero)
    uint32_t qa[HD / 16][4];
    {
        const float* qr = p.q + ((size_t)h * G + min(g, G - 1)) * HD;
#pragma unroll
        for (int kk = 0; kk < HD / 16; kk++) {
            float2 a = __ldcg(reinterpret_cast<const float2*>(qr + kk * 16 + 2 * tq));
            float2 c = __ldcg(reinterpret_cast<const float2*>(qr + kk * 16 + 8 + 2 * tq));
            qa[kk][0] = g < G ? am_pack(a.x, a.y) : 0u;
            qa[kk][2] = g < G ? am_pack(c.x, c.y) : 0u;
            qa[kk][1] = 0u; qa[kk][3] = 0u;
        }
    }
    float o[HD / 8][4];
#pragma unroll
    for (int i = 0; i < HD / 8; i++) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;                        // row g of this warp's positions (l: this thread's share)

    auto issue = [&](int ti) {                                   // all threads: tile ti -> buffer ti & 1
        const int st = ti & 1, t0 = q0 + ti * AM_TT;
        __half* kd = ks + st * AM_TT * LD;
        __half* vd = vs + st * AM_TT * LD;
        constexpr int CPR = HD / 8;                              // 16-byte chunks per row
        for (int i = tid; i < AM_TT * CPR; i += AM_THREADS) {
            const int j = i / CPR, c = i % CPR, pos = t0 + j;
            if (pos < q1) {
                const size_t off = (((size_t)p.block_table[pos / bt] * p.n_kv + h) * bt + pos % bt) * HD + c * 8;
                am_cp16(kd + j * LD + c * 8, p.k_cache + off);
                am_cp16(vd + j * LD + c * 8, p.v_cache + off);
            } else {                                             // 0 * garbage must stay 0 in P.V
                *reinterpret_cast<uint4*>(kd + j * LD + c * 8) = make_uint4(0, 0, 0, 0);
                *reinterpret_cast<uint4*>(vd + j * LD + c * 8) = make_uint4(0, 0, 0, 0);
            }
        }
        am_commit();
    };
    if (ntile > 0) issue(0);
    for (int ti = 0; ti < ntile; ti++) {
        if (ti + 1 < ntile) { issue(ti + 1); am_wait<1>(); } else { am_wait<0>(); }
        __syncthreads();                                         // tile ti is in shared memory for every warp
        const int st = ti & 1, w0 = q0 + ti * AM_TT + 16 * warp; // first position of this warp's slice
        if (w0 < q1) {
            const __half* kb = ks + (st * AM_TT + 16 * warp) * LD;
            const __half* vb = vs + (st * AM_TT + 16 * warp) * LD;
            float sc[2][4];
#pragma unroll
            for (int nt = 0; nt < 2; nt++) {
                sc[nt][0] = sc[nt][1] = sc[nt][2] = sc[nt][3] = 0.f;
                const __half* kr = kb + (nt * 8 + g) * LD + 2 * tq;
#pragma unroll
                for (int kk = 0; kk < HD / 16; kk++)
                    am_mma(sc[nt], qa[kk], *reinterpret_cast<const uint32_t*>(kr + kk * 16),
                           *reinterpret_cast<const uint32_t*>(kr + kk * 16 + 8));
            }
            // ---- mask + online softmax of row g (log2 domain); columns 2tq, 2tq+1 of each 8-position group
            float mx = m_run;
#pragma unroll
            for (int nt = 0; nt < 2; nt++) {
                const int pz = w0 + nt * 8 + 2 * tq;
                sc[nt][0] = pz < q1 ? sc[nt][0] * scale_log2 : -INFINITY;
                sc[nt][1] = pz + 1 < q1 ? sc[nt][1] * scale_log2 : -INFINITY;
                mx = fmaxf(mx, fmaxf(sc[nt][0], sc[nt][1]));
            }
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));   // finite: position w0 is valid for every row
            const float f = exp2f(m_run - mx);
            m_run = mx;
            uint32_t pa[4];
            float e0 = exp2f(sc[0][0] - mx), e1 = exp2f(sc[0][1] - mx), e2 = exp2f(sc[1][0] - mx), e3 = exp2f(sc[1][1] - mx);
            l_run = l_run * f + (e0 + e1) + (e2 + e3);
            pa[0] = am_pack(e0, e1); pa[2] = am_pack(e2, e3); pa[1] = 0u; pa[3] = 0u;
#pragma unroll
            for (int nt = 0; nt < HD / 8; nt++) {
                o[nt][0] *= f; o[nt][1] *= f;
                uint32_t b0, b1;
                am_ldmatrix_x2_trans(b0, b1, vb + (lane & 15) * LD + nt * 8);
                am_mma(o[nt], pa, b0, b1);
            }
        }
        __syncthreads();                                         // everyone is done with buffer st before it is refilled
    }
    // ---- merge the four warps (their position slices are disjoint): (m, l, o) per head row, in the tile buffers
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 1);
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 2);
    float* wo = reinterpret_cast<float*>(smem_raw);              // [4][8][HD]
    float* wm = wo + 4 * 8 * HD;                                 // [4][8]
    float* wl = wm + 32;                                         // [4][8]
    if (g < G) {
#pragma unroll
        for (int nt = 0; nt < HD / 8; nt++)
            *reinterpret_cast<float2*>(wo + (warp * 8 + g) * HD + nt * 8 + 2 * tq) = make_float2(o[nt][0], o[nt][1]);
        if (tq == 0) { wm[warp * 8 + g] = m_run; wl[warp * 8 + g] = l_run; }
    }
    __syncthreads();
    const int rec = HD + 2;
    const bool single = p.n_splits == 1;
    for (int e = tid; e < G * HD; e += AM_THREADS) {
        const int gg = e / HD, d = e % HD;
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < 4; w++) M = fmaxf(M, wm[w * 8 + gg]);
        float num = 0.f, den = 0.f;
#pragma unroll
        for (int w = 0; w < 4; w++) {
            const float mw = wm[w * 8 + gg];
            const float wt = mw == -INFINITY ? 0.f : exp2f(mw - M);
            num += wt * wo[(w * 8 + gg) * HD + d];
            den += wt * wl[w * 8 + gg];
        }
        if (single) {
            am_store_out(p, z, (h * G + gg) * HD + d, num / den);
        } else {
            float* r = p.scratch + (((size_t)h * p.n_splits + sp) * G + gg) * rec;
            r[d] = num;
            if (d == 0) { r[HD] = M; r[HD + 1] = den; }          // an empty split leaves M = -inf, den = 0
        }
    }
    if (single) return;
    // ---- the last CTA of this KV head to get here merges the splits
    __threadfence();
    __syncthreads();
    if (tid == 0) s_ticket = atomicAdd(&p.counters[h], 1u);
    __syncthreads();
    if (s_ticket != (unsigned)(p.n_splits - 1)) return;
    __threadfence();
    // (m, l) of every split with independent loads, weights once per (head, split), then the outputs
    float* msp = reinterpret_cast<float*>(smem_raw);             // [G][n_splits]; the tile buffers are free again
    float* lsp = msp + 8 * p.n_splits;
    float* wsp = lsp + 8 * p.n_splits;
    float* inv = wsp + 8 * p.n_splits;                           // [G]
    for (int i = tid; i < G * p.n_splits; i += AM_THREADS) {
        const int gg = i / p.n_splits, s2 = i % p.n_splits;
        const float* r = p.scratch + (((size_t)h * p.n_splits + s2) * G + gg) * rec;
        msp[i] = __ldcg(r + HD);
        lsp[i] = __ldcg(r + HD + 1);
    }
    __syncthreads();
    if (tid < G) {
        float M = -INFINITY, den = 0.f;
        for (int s2 = 0; s2 < p.n_splits; s2++) M = fmaxf(M, msp[tid * p.n_splits + s2]);
        for (int s2 = 0; s2 < p.n_splits; s2++) {
            const float ms = msp[tid * p.n_splits + s2];
            const float wt = ms == -INFINITY ? 0.f : exp2f(ms - M);
            wsp[tid * p.n_splits + s2] = wt;
            den += wt * lsp[tid * p.n_splits + s2];
        }
        inv[tid] = 1.0f / den;
    }
    __syncthreads();
    for (int e = tid; e < G * HD; e += AM_THREADS) {
        const int gg = e / HD, d = e % HD;
        const float* base = p.scratch + ((size_t)h * p.n_splits * G + gg) * rec + d;
        float num = 0.f;
#pragma unroll 8
        for (int s2 = 0; s2 < p.n_splits; s2++) num += wsp[gg * p.n_splits + s2] * __ldcg(base + (size_t)s2 * G * rec);
        am_store_out(p, z, (h * G + gg) * HD + d, num * inv[gg]);
    }
    if (tid == 0) p.counters[h] = 0;                             // ready for the next launch
}

static size_t am_smem_bytes(int hd) { return (size_t)2 * 2 * AM_TT * (hd + 8) * 2; }

cudaError_t attn_mma_init() {
    cudaError_t e = cudaFuncSetAttribute(attn_decode_mma_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)am_smem_bytes(64));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(attn_decode_mma_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)am_smem_bytes(128));
}

// tab == nullptr: one sequence described by p itself (q, out, pos_ptr, block_table); else b.n sequences, q / out are
// [b.n][n_head*head_dim] and scratch / counters hold b.n times the single-sequence size.
cudaError_t attn_decode_mma_launch(const AttnParams& p, const SlotDev* tab, const BatchIdx* b, cudaStream_t stream, bool pdl) {
    const int G = p.n_kv > 0 ? p.n_head / p.n_kv : 0;
    if (p.n_kv <= 0 || p.n_head % p.n_kv || G > 8 || p.n_splits < 1 || p.n_splits > 256) return cudaErrorInvalidValue;
    if (tab && (!b || b->n < 1 || b->n > BATCH_MAX)) return cudaErrorInvalidValue;
    BatchIdx none = {};
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(p.n_splits, p.n_kv, tab ? b->n : 1); cfg.blockDim = dim3(AM_THREADS); cfg.stream = stream;
    cfg.dynamicSmemBytes = am_smem_bytes(p.head_dim);
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = (pdl || g_chain_pdl) ? 1 : 0;
    const float scale_log2 = p.scale * 1.4426950408889634f;
    if (p.head_dim == 64) return cudaLaunchKernelEx(&cfg, attn_decode_mma_kernel<64>, p, tab, tab ? *b : none, scale_log2);
    if (p.head_dim == 128) return cudaLaunchKernelEx(&cfg, attn_decode_mma_kernel<128>, p, tab, tab ? *b : none, scale_log2);
    return cudaErrorInvalidValue;
}

}  // namespace b2
