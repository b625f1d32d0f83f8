// Prefill / batched-decode GEMM on the 5th-generation tensor cores:
//     Y[t][n] (+)= sum_k X[t][k] * W[n][k]        X: f16 activations [T][K], W: GGUF-quantised [N][K]
// Replaces ggml-cuda's "dequantise the whole matrix to an f16 scratch copy, then cublasGemmEx" prefill
// path (convert.cu + cuBLAS) that llama-cpp-python runs for the prompt under
// create_chat_completion (/root/reference/api.py:55-63). No f16 copy of W ever exists in HBM: the quant
// blocks are dequantised straight into the shared-memory operand tile of the MMA.
//
// One CTA owns 128 weight rows (the MMA "M" dimension = TMEM lanes) and ALL T <= 512 tokens (the MMA "N"
// dimension = TMEM columns), so every weight is read and dequantised exactly once per call:
//   per 64-wide K block, double buffered:
//     * thread r dequantises the 64 weights of row r into the A tile      (128 x 64 f16, 16 KB)
//     * all threads copy the activations of the block into the B tile     (T x 64 f16)
//       both tiles use the K-major no-swizzle canonical layout: 8x8 core matrices, LBO = stride between
//       core matrices along K, SBO = stride between 8-row groups
//     * fence.proxy.async + barrier, then ONE thread issues 4 (x2 if T > 256) tcgen05.mma
//       (kind::f16, M = 128, N = T, K = 16) accumulating into TMEM, and tcgen05.commit arrives on the
//       mbarrier that frees the stage
//   epilogue: each warp reads its 32 TMEM lanes with tcgen05.ld (32x32b.x8) and stores / accumulates
//   into the float32 output, token-major.
// Roofline: tensor pipe for large T; for T <= ~64 the dequantisation (CUDA cores) and HBM bound it.
#include "kernels.h"
#include "ptx.cuh"
#include "quant_blocks.cuh"
#include <cuda_fp16.h>

namespace b2 {

constexpr int GM_THREADS = 128;
constexpr int GM_BM = 128;          // weight rows per CTA
constexpr int GM_BK = 64;           // K per stage

__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float (&v)[8]) {
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 8; i++) v[i] = __uint_as_float(r[i]);
}

// shared-memory matrix descriptor, K-major, no swizzle: start, LBO (K direction), SBO (M/N direction)
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((addr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;                  // descriptor version (sm_100)
    return d;                                // layout_type = 0: no swizzle
}
// instruction descriptor: D = f32, A = B = f16, both K-major, M x N
__device__ __forceinline__ uint32_t instr_desc(int M, int N) {
    return (1u << 4) | (0u << 7) | (0u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ---------------------------------------------------------------- dequantise 64 weights of one row
// k0 = first element of the block (multiple of 64). Writes 8 uint4 (8 x f16 each) through `emit(c, v)`.
struct H8 { uint32_t w[4]; };
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

template <typename Emit>
__device__ __forceinline__ void dequant_row64(int type, const uint8_t* row, int k0, Emit emit) {
    float v[8];
    if (type == T_Q4_K || type == T_Q5_K) {
        const int bb = type == T_Q4_K ? 144 : 176;
        const uint8_t* blk = row + (k0 >> 8) * bb;
        const int c = (k0 & 255) >> 6;                                 // 64-chunk: sub-blocks 2c (low nibbles), 2c+1 (high)
        U4 h = ld128(blk);
        const uint8_t* qs = blk + (type == T_Q4_K ? 16 : 48) + 32 * c;
        U4 q0 = ld128(qs), q1 = ld128(qs + 16);
        uint32_t qw[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
        uint32_t hw[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (type == T_Q5_K) { U4 a = ld128(blk + 16), b = ld128(blk + 32); hw[0] = a.x; hw[1] = a.y; hw[2] = a.z; hw[3] = a.w; hw[4] = b.x; hw[5] = b.y; hw[6] = b.z; hw[7] = b.w; }
        uint32_t sh = (c & 1) * 16, w0 = h.y >> sh, w1 = h.z >> sh, w2 = h.w >> sh, scp, mp;
        if (c < 2) { scp = w0 & 0x3F3Fu; mp = w1 & 0x3F3Fu; }
        else { scp = (w2 & 0x0F0Fu) | ((w0 >> 2) & 0x3030u); mp = ((w2 >> 4) & 0x0F0Fu) | ((w1 >> 2) & 0x3030u); }
        const float d = h2f(h.x & 0xffffu), dmin = h2f(h.x >> 16);
        const float dl[2] = {d * (float)(scp & 0xffu), d * (float)(scp >> 8)};
        const float ml[2] = {dmin * (float)(mp & 0xffu), dmin * (float)(mp >> 8)};
#pragma unroll
        for (int ch = 0; ch < 8; ch++) {                               // 8 output chunks of 8 elements
            const int hi = ch >> 2;                                    // chunks 0-3: low nibbles, 4-7: high nibbles
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const int l = (ch & 3) * 8 + i;                        // byte index within the 32-byte run
                uint32_t q = (qw[l >> 2] >> (8 * (l & 3) + 4 * hi)) & 15u;
                if (type == T_Q5_K) q |= ((hw[l >> 2] >> (8 * (l & 3) + 2 * c + hi)) & 1u) << 4;
                v[i] = dl[hi] * (float)q - ml[hi];
            }
            H8 o; o.w[0] = pack_h2(v[0], v[1]); o.w[1] = pack_h2(v[2], v[3]); o.w[2] = pack_h2(v[4], v[5]); o.w[3] = pack_h2(v[6], v[7]);
            emit(ch, o);
        }
    } else if (type == T_Q6_K) {
        const uint8_t* blk = row + (k0 >> 8) * 210;
        const int e0 = k0 & 255, hf = e0 >> 7, nib = (e0 >> 6) & 1;   // 64 elements = one nibble plane of one half
        uint32_t ql[16], qh[8], scw[1];
        ld_words_u2<16>(blk + 64 * hf, ql);
        ld_words_u2<8>(blk + 128 + 32 * hf, qh);
        ld_words_u2<1>(blk + 192 + 8 * hf + 4 * nib, scw);
        const float d = h2f(ld16(blk + 208));
#pragma unroll
        for (int ch = 0; ch < 8; ch++) {
            const float dl = d * (float)(int8_t)(scw[0] >> (8 * (ch >> 1)));
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const int e = ch * 8 + i;                              // 0..63 within the plane
                const int l = e & 31, kq = 2 * nib + (e >> 5);
                uint32_t lo = (ql[e >> 2] >> (8 * (e & 3) + 4 * nib)) & 15u;
                uint32_t hi2 = (qh[l >> 2] >> (8 * (l & 3) + 2 * kq)) & 3u;
                v[i] = dl * (float)((int)(lo | (hi2 << 4)) - 32);
            }
            H8 o; o.w[0] = pack_h2(v[0], v[1]); o.w[1] = pack_h2(v[2], v[3]); o.w[2] = pack_h2(v[4], v[5]); o.w[3] = pack_h2(v[6], v[7]);
            emit(ch, o);
        }
    } else if (type == T_Q8_0) {
        const uint8_t* p = row + (k0 >> 5) * 34;                       // two 34-byte blocks, 4-byte aligned
        uint32_t W[17];
#pragma unroll
        for (int i = 0; i < 17; i++) W[i] = ld32(p + 4 * i);
        const float d0 = h2f(W[0] & 0xffffu), d1 = h2f(W[8] >> 16);
#pragma unroll
        for (int ch = 0; ch < 8; ch++) {
            const float dd = ch < 4 ? d0 : d1;
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const int j = (ch & 3) * 8 + i;                        // int8 index within the block
                uint32_t word = ch < 4 ? fshr(W[j >> 2], W[(j >> 2) + 1], 16) : W[9 + (j >> 2)];
                v[i] = dd * (float)(int8_t)(word >> (8 * (j & 3)));
            }
            H8 o; o.w[0] = pack_h2(v[0], v[1]); o.w[1] = pack_h2(v[2], v[3]); o.w[2] = pack_h2(v[4], v[5]); o.w[3] = pack_h2(v[6], v[7]);
            emit(ch, o);
        }
    } else if (type == T_Q4_0) {
        const uint8_t* p = row + (k0 >> 5) * 18;                       // two 18-byte blocks, 4-byte aligned
        uint32_t W[9];
#pragma unroll
        for (int i = 0; i < 9; i++) W[i] = ld32(p + 4 * i);
        const float d0 = h2f(W[0] & 0xffffu), d1 = h2f(W[4] >> 16);
#pragma unroll
        for (int ch = 0; ch < 8; ch++) {
            const int b = ch >> 2, part = ch & 3;                      // block, 8-element part: 0,1 low nibbles; 2,3 high
            const float dd = b ? d1 : d0;
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const int j = (part & 1) * 8 + i;                      // qs byte 0..15
                uint32_t word = b ? W[5 + (j >> 2)] : fshr(W[j >> 2], W[(j >> 2) + 1], 16);
                uint32_t q = (word >> (8 * (j & 3) + 4 * (part >> 1))) & 15u;
                v[i] = dd * (float)((int)q - 8);
            }
            H8 o; o.w[0] = pack_h2(v[0], v[1]); o.w[1] = pack_h2(v[2], v[3]); o.w[2] = pack_h2(v[4], v[5]); o.w[3] = pack_h2(v[6], v[7]);
            emit(ch, o);
        }
    }
}

struct __align__(16) GemmShared {
    uint64_t mma_done[2];
    uint32_t tmem_base;
    uint32_t pad[3];
};

__global__ void __launch_bounds__(GM_THREADS) gemm_q_kernel(const GemmParams p) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    GemmShared& sh = *reinterpret_cast<GemmShared*>(smem_raw);
    const int tid = threadIdx.x, warp = tid >> 5;
    const int T = p.T_pad;
    const uint32_t a_bytes = GM_BM * GM_BK * 2, b_bytes = (uint32_t)(T + 1) * 16 * (GM_BK / 8);   // B chunks padded by 16 B: conflict-free stores
    uint8_t* stage0 = smem_raw + 128;
    const uint32_t stage_bytes = a_bytes + b_bytes;
    uint32_t ncols = 32;
    while ((int)ncols < T) ncols <<= 1;

    if (tid == 0) { mbar_init(&sh.mma_done[0], 1); mbar_init(&sh.mma_done[1], 1); fence_mbar_init(); }
    if (warp == 0) tmem_alloc(&sh.tmem_base, ncols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = sh.tmem_base;

    const int m0 = blockIdx.x * GM_BM;
    const int my_row = min(m0 + tid, p.N - 1);                        // rows past N are computed and dropped
    const uint8_t* wrow = p.w + (size_t)my_row * p.row_bytes;
    const uint32_t idesc_lo = instr_desc(GM_BM, T > 256 ? 256 : T);
    const uint32_t idesc_hi = T > 256 ? instr_desc(GM_BM, T - 256) : 0u;
    // K-major, no swizzle: element (r, k) of a tile with R rows sits at (k/8) * (R*16) + r*16 + (k%8)*2
    const uint32_t lbo_a = GM_BM * 16, lbo_b = (uint32_t)(T + 1) * 16, sbo = 128;
    const int n_kb = p.K / GM_BK;

    for (int kb = 0; kb < n_kb; kb++) {
        const int s = kb & 1;
        uint8_t* sa = stage0 + (size_t)s * stage_bytes;
        uint8_t* sb = sa + a_bytes;
        if (kb >= 2) mbar_wait(&sh.mma_done[s], ((kb >> 1) - 1) & 1, p.err, 0x40001u, (unsigned)kb);   // MMAs of kb-2 done with this stage
        // ---- A tile: this thread's weight row, 64 values -> 8 core-matrix rows
        dequant_row64(p.type, wrow, kb * GM_BK, [&](int ch, const H8& o) {
            *reinterpret_cast<uint4*>(sa + ch * lbo_a + tid * 16) = make_uint4(o.w[0], o.w[1], o.w[2], o.w[3]);
        });
        // ---- B tile: activations of the block, T rows x 8 chunks of 16 bytes
        for (int i = tid; i < T * 8; i += GM_THREADS) {
            const int t = i >> 3, ch = i & 7;
            const uint4 v = *reinterpret_cast<const uint4*>(p.x + (size_t)t * p.K + kb * GM_BK + ch * 8);
            *reinterpret_cast<uint4*>(sb + ch * lbo_b + t * 16) = v;
        }
        fence_proxy_async();            // generic-proxy smem writes -> visible to the tensor core (async proxy)
        __syncthreads();
        if (tid == 0) {
            tc_fence_after();
            const uint32_t a_addr = smem_u32(sa), b_addr = smem_u32(sb);
#pragma unroll
            for (int ks = 0; ks < GM_BK / 16; ks++) {
                const uint64_t da = smem_desc(a_addr + ks * 2 * lbo_a, lbo_a, sbo);
                const uint64_t db = smem_desc(b_addr + ks * 2 * lbo_b, lbo_b, sbo);
                const uint32_t acc = (kb > 0 || ks > 0) ? 1u : 0u;
                umma_f16(tmem, da, db, idesc_lo, acc);
                if (T > 256) {
                    const uint64_t db2 = smem_desc(b_addr + ks * 2 * lbo_b + 256 * 16, lbo_b, sbo);
                    umma_f16(tmem + 256, da, db2, idesc_hi, acc);
                }
            }
            umma_commit(&sh.mma_done[s]);   // arrives when the MMAs above have finished reading the stage
        }
    }
    // ---- wait for the last MMAs, then read the accumulators back
    {
        const int last = n_kb - 1;
        mbar_wait(&sh.mma_done[last & 1], (last >> 1) & 1, p.err, 0x40002u, (unsigned)last);
        if (n_kb >= 2) mbar_wait(&sh.mma_done[(last - 1) & 1], ((last - 1) >> 1) & 1, p.err, 0x40003u, (unsigned)last);
    }
    tc_fence_after();
    const int row = m0 + tid;                                          // TMEM lane = tid (warp w owns lanes 32w..32w+31)
    const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);
    for (int c0 = 0; c0 < T; c0 += 8) {
        float v[8];
        tmem_ld8(lane_addr + c0, v);
        if (row < p.N) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const int t = c0 + i;
                if (t < p.T) {
                    float* y = p.y + (size_t)t * p.ldy + row;
                    *y = p.accumulate ? *y + v[i] : v[i];
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, ncols);
}

size_t gemm_smem_bytes(int T_pad) { return 128 + 2 * ((size_t)GM_BM * GM_BK * 2 + (size_t)(T_pad + 1) * 16 * (GM_BK / 8)); }

cudaError_t gemm_init() {
    return cudaFuncSetAttribute(gemm_q_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gemm_smem_bytes(512));
}

cudaError_t gemm_launch(const GemmParams& p, cudaStream_t stream) {
    if (p.T_pad < 16 || p.T_pad > 512 || p.T_pad % 16 || p.T > p.T_pad || p.K % GM_BK || p.K % 256) return cudaErrorInvalidValue;
    if (p.T_pad > 256 && (p.T_pad - 256) % 16) return cudaErrorInvalidValue;
    int be = block_elems(p.type);
    if (be == 1 || !block_bytes(p.type)) return cudaErrorInvalidValue;
    const int grid = (p.N + GM_BM - 1) / GM_BM;
    gemm_q_kernel<<<grid, GM_THREADS, gemm_smem_bytes(p.T_pad), stream>>>(p);
    return cudaGetLastError();
}

__global__ void to_half_kernel(const float* x, __half* y, long n) {
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] = __float2half_rn(x[i]);
}
cudaError_t to_half_launch(const float* x, __half* y, long n, cudaStream_t s) {
    to_half_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(x, y, n);
    return cudaGetLastError();
}

}  // namespace b2
