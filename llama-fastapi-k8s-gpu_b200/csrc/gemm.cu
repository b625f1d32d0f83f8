// Prefill / batched-decode GEMM on the 5th-generation tensor cores:
//     P[s][t][n] = sum_{k in split s} X[t][k] * W[n][k]      X: f16 activations, W: GGUF-quantised [N][K]
// Replaces ggml-cuda's "dequantise the whole matrix to an f16 scratch copy, then cublasGemmEx" prompt
// path (convert.cu + cuBLAS) that llama-cpp-python runs under create_chat_completion
// (/root/reference/api.py:55-63), and its batched mul_mat_q path for several sequences per step. No f16 copy
// of W ever exists in HBM: quant blocks go from HBM to registers to the shared-memory operand tile of the MMA.
//
// Work decomposition. A CTA owns one tile of 128 weight rows (MMA "M" = TMEM lanes) of one of up to three
// matrices that share the input (q|k|v, gate|up), ALL T <= 512 tokens (MMA "N" = TMEM columns) and one of
// n_splits contiguous K ranges; split s writes its partial sums to y + s * split_stride and the consumer
// kernel (RoPE/KV store, residual+RMSNorm, SiLU*mul: prefill.cu) adds the splits in a fixed order, so results
// are deterministic. Splitting K is what fills 148 SMs when N is only 1024-4096 rows.
//
// Pipeline per CTA, S stages of 64 K-elements (A tile 16 KB + B tile T*128 B):
//   warps 0-3 (thread r <-> weight row r): raw quant bytes of the NEXT block are already in registers
//       (software prefetch) while the current block is expanded to f16 with byte-permute + half2 arithmetic
//       (exact small integers 1024+q via byte-permute; scales applied without losing bits: ~2-2.5 instr/weight)
//       and stored as 8 core-matrix rows of the A tile; fence.proxy.async; arrive on a_full[stage].
//   warp 4, one thread: the B tile of a block is ONE 1-D TMA bulk copy (cp.async.bulk + mbarrier complete_tx):
//       activations live in HBM in the MMA's own K-major core-matrix order [K/8][T][8] (the producing kernels
//       write that layout), so the tile is contiguous. Waits a_full/b_full, issues 4 (x2 when T > 256)
//       tcgen05.mma kind::f16 M=128 N=T K=16 accumulating in TMEM; tcgen05.commit arrives on empty[stage].
//   epilogue: warps 0-3 read their 32 TMEM lanes (tcgen05.ld 32x32b.x16) and store float32, token-major.
// Shared-memory operand layout: K-major, no swizzle: 8x8 f16 core matrices (8 rows x 16 B), LBO = byte stride
// between core matrices along K, SBO = stride between 8-row groups (128 B).
// Roofline: tensor pipe for large T; for T <= ~64 HBM (weights read once per call) and the dequantisation.
#include "kernels.h"
#include "ptx.cuh"
#include "quant_blocks.cuh"
#include <cuda.h>
#include <cuda_fp16.h>
#include <stdlib.h>
#include <algorithm>

namespace b2 {

constexpr int GM_BM = 128;          // weight rows per CTA
constexpr int GM_BK = 64;           // K per stage
constexpr int GM_HDR = 512;         // barriers at the start of dynamic shared memory
constexpr int GM_MAX_STAGES = 6;     // A-tile ring
constexpr int GM_MAX_BSTAGES = 8;    // activation-tile ring
constexpr int GM_RAW_STAGES = 2;    // raw weight spans in flight (TMA), at most
constexpr int GM_SPAN_WORDS = 68;   // largest 256-weight span: Q8_0, 272 bytes

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
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; i++) v[i] = __uint_as_float(r[i]);
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
__device__ __forceinline__ uint64_t desc_join(uint32_t lo, uint32_t hi) {
    uint64_t d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(hi));
    return d;
}
// instruction descriptor: D = f32, A = B = f16, both K-major, M x N
__device__ __forceinline__ uint32_t instr_desc(int M, int N) {
    return (1u << 4) | (0u << 7) | (0u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ---------------------------------------------------------------- raw block bytes -> f16
// One K block = 64 consecutive weights of one row. RAW_WORDS covers the largest format (Q6_K).
constexpr int RAW_WORDS = 26;
struct RawKB { uint32_t w[RAW_WORDS]; };

__device__ __forceinline__ void load_raw(int type, const uint8_t* row, int k0, RawKB& r) {
    if (type == T_Q4_K) {
        const uint8_t* blk = row + (k0 >> 8) * 144;
        const U4 h = ld128(blk), a = ld128(blk + 16 + 32 * ((k0 & 255) >> 6)), b = ld128(blk + 32 + 32 * ((k0 & 255) >> 6));
        r.w[0] = h.x; r.w[1] = h.y; r.w[2] = h.z; r.w[3] = h.w;
        r.w[4] = a.x; r.w[5] = a.y; r.w[6] = a.z; r.w[7] = a.w; r.w[8] = b.x; r.w[9] = b.y; r.w[10] = b.z; r.w[11] = b.w;
    } else if (type == T_Q5_K) {
        const uint8_t* blk = row + (k0 >> 8) * 176;
        const uint8_t* qs = blk + 48 + 32 * ((k0 & 255) >> 6);
        const U4 h = ld128(blk), a = ld128(qs), b = ld128(qs + 16), c = ld128(blk + 16), d = ld128(blk + 32);
        r.w[0] = h.x; r.w[1] = h.y; r.w[2] = h.z; r.w[3] = h.w;
        r.w[4] = a.x; r.w[5] = a.y; r.w[6] = a.z; r.w[7] = a.w; r.w[8] = b.x; r.w[9] = b.y; r.w[10] = b.z; r.w[11] = b.w;
        r.w[12] = c.x; r.w[13] = c.y; r.w[14] = c.z; r.w[15] = c.w; r.w[16] = d.x; r.w[17] = d.y; r.w[18] = d.z; r.w[19] = d.w;
    } else if (type == T_Q6_K) {
        const uint8_t* blk = row + (k0 >> 8) * 210;
        const int e0 = k0 & 255, hf = e0 >> 7, nib = (e0 >> 6) & 1;
        uint32_t ql[16], qh[8], sc[1];
        ld_words_u2<16>(blk + 64 * hf, ql);
        ld_words_u2<8>(blk + 128 + 32 * hf, qh);
        ld_words_u2<1>(blk + 192 + 8 * hf + 4 * nib, sc);
#pragma unroll
        for (int i = 0; i < 16; i++) r.w[i] = ql[i];
#pragma unroll
        for (int i = 0; i < 8; i++) r.w[16 + i] = qh[i];
        r.w[24] = sc[0];
        r.w[25] = ld16(blk + 208);
    } else if (type == T_Q8_0) {
        const uint8_t* p = row + (k0 >> 5) * 34;                       // two 34-byte blocks, 4-byte aligned
#pragma unroll
        for (int i = 0; i < 17; i++) r.w[i] = ld32(p + 4 * i);
    } else {                                                           // Q4_0: two 18-byte blocks, 4-byte aligned
        const uint8_t* p = row + (k0 >> 5) * 18;
#pragma unroll
        for (int i = 0; i < 9; i++) r.w[i] = ld32(p + 4 * i);
    }
}

// four unsigned bytes -> two half2 holding (1024 + byte) exactly: 0x64xx is the f16 with value 1024 + xx
__device__ __forceinline__ void bytes_to_h2(uint32_t b, __half2& lo, __half2& hi) {
    const uint32_t l = __byte_perm(b, 0x64646464u, 0x4140), h = __byte_perm(b, 0x64646464u, 0x4342);
    lo = *reinterpret_cast<const __half2*>(&l);
    hi = *reinterpret_cast<const __half2*>(&h);
}
__device__ __forceinline__ uint32_t h2_bits(__half2 h) { return *reinterpret_cast<uint32_t*>(&h); }

// Expanding 8 byte-valued weights (two words) to one 16-byte core-matrix row. All three variants round the
// weight to f16 once (plus one more rounding where a tiny correction term is added): the scales themselves
// never lose bits, so a whole sub-block is not biased by a rounded scale.
// K-quants with a minimum (Q4_K, Q5_K): w = d*(sc*q) - dmin*m. sc*q <= 63*31 is an exact f16 integer:
//   t = (1024+q)*sc - 1024*sc (exact), v = t*d + mh, w = v + ml   with mh + ml = -dmin*m to ~22 bits
__device__ __forceinline__ uint4 expand8_k(uint32_t b0, uint32_t b1, __half2 sc, __half2 nsc, __half2 d, __half2 mh, __half2 ml) {
    __half2 p0, p1, p2, p3;
    bytes_to_h2(b0, p0, p1);
    bytes_to_h2(b1, p2, p3);
    uint4 o;
    o.x = h2_bits(__hadd2(__hfma2(__hfma2(p0, sc, nsc), d, mh), ml));
    o.y = h2_bits(__hadd2(__hfma2(__hfma2(p1, sc, nsc), d, mh), ml));
    o.z = h2_bits(__hadd2(__hfma2(__hfma2(p2, sc, nsc), d, mh), ml));
    o.w = h2_bits(__hadd2(__hfma2(__hfma2(p3, sc, nsc), d, mh), ml));
    return o;
}
// scale that is not an f16 (Q6_K: d * int8): s = s_hi + s_lo, w = q*s_hi + q*s_lo
__device__ __forceinline__ uint4 expand8_s(uint32_t b0, uint32_t b1, __half2 off, __half2 s_hi, __half2 s_lo) {
    __half2 p0, p1, p2, p3;
    bytes_to_h2(b0, p0, p1);
    bytes_to_h2(b1, p2, p3);
    p0 = __hsub2(p0, off); p1 = __hsub2(p1, off); p2 = __hsub2(p2, off); p3 = __hsub2(p3, off);
    uint4 o;
    o.x = h2_bits(__hfma2(p0, s_hi, __hmul2(p0, s_lo)));
    o.y = h2_bits(__hfma2(p1, s_hi, __hmul2(p1, s_lo)));
    o.z = h2_bits(__hfma2(p2, s_hi, __hmul2(p2, s_lo)));
    o.w = h2_bits(__hfma2(p3, s_hi, __hmul2(p3, s_lo)));
    return o;
}
// f16 scale (Q8_0, Q4_0): w = q * d, one rounding
__device__ __forceinline__ uint4 expand8_d(uint32_t b0, uint32_t b1, __half2 off, __half2 d) {
    __half2 p0, p1, p2, p3;
    bytes_to_h2(b0, p0, p1);
    bytes_to_h2(b1, p2, p3);
    uint4 o;
    o.x = h2_bits(__hmul2(__hsub2(p0, off), d));
    o.y = h2_bits(__hmul2(__hsub2(p1, off), d));
    o.z = h2_bits(__hmul2(__hsub2(p2, off), d));
    o.w = h2_bits(__hmul2(__hsub2(p3, off), d));
    return o;
}
__device__ __forceinline__ void split_f16(float v, __half2& hi, __half2& lo) {
    const __half h = __float2half_rn(v);
    hi = __half2half2(h);
    lo = __float2half2_rn(v - __half2float(h));
}
__device__ __forceinline__ __half2 h2_from_bits16(uint32_t b) { return __half2half2(__ushort_as_half((unsigned short)b)); }

// dst: this thread's 16-byte slot in chunk 0 of the A tile; chunk ch is at dst + ch * lbo
__device__ __forceinline__ void dequant_store(int type, const RawKB& r, int k0, uint8_t* dst, uint32_t lbo) {
    auto put = [&](int ch, const uint4& v) { *reinterpret_cast<uint4*>(dst + ch * lbo) = v; };
    if (type == T_Q4_K || type == T_Q5_K) {
        const int c = (k0 & 255) >> 6;                                 // 64-chunk: sub-blocks 2c (low nibbles), 2c+1 (high)
        const uint32_t sh = (c & 1) * 16, w0 = r.w[1] >> sh, w1 = r.w[2] >> sh, w2 = r.w[3] >> sh;
        uint32_t scp, mp;
        if (c < 2) { scp = w0 & 0x3F3Fu; mp = w1 & 0x3F3Fu; }
        else { scp = (w2 & 0x0F0Fu) | ((w0 >> 2) & 0x3030u); mp = ((w2 >> 4) & 0x0F0Fu) | ((w1 >> 2) & 0x3030u); }
        const __half2 d = h2_from_bits16(r.w[0] & 0xffffu);
        const float dmin = h2f(r.w[0] >> 16);
        const float sc0 = (float)(scp & 0xffu), sc1 = (float)(scp >> 8);
        const __half2 s0 = __float2half2_rn(sc0), s1 = __float2half2_rn(sc1);
        const __half2 n0 = __float2half2_rn(-1024.f * sc0), n1 = __float2half2_rn(-1024.f * sc1);
        __half2 mh0, ml0, mh1, ml1;
        split_f16(-dmin * (float)(mp & 0xffu), mh0, ml0);
        split_f16(-dmin * (float)(mp >> 8), mh1, ml1);
#pragma unroll
        for (int j = 0; j < 4; j++) {                                  // 8 bytes -> 8 low-nibble + 8 high-nibble weights
            uint32_t a = r.w[4 + 2 * j], b = r.w[5 + 2 * j];
            uint32_t lo0 = a & 0x0F0F0F0Fu, lo1 = b & 0x0F0F0F0Fu, hi0 = (a >> 4) & 0x0F0F0F0Fu, hi1 = (b >> 4) & 0x0F0F0F0Fu;
            if (type == T_Q5_K) {                                      // fifth bit: bit 2c (low) / 2c+1 (high) of qh byte l
                const uint32_t ha = r.w[12 + 2 * j] >> (2 * c), hb = r.w[13 + 2 * j] >> (2 * c);
                lo0 |= (ha & 0x01010101u) << 4; lo1 |= (hb & 0x01010101u) << 4;
                hi0 |= ((ha >> 1) & 0x01010101u) << 4; hi1 |= ((hb >> 1) & 0x01010101u) << 4;
            }
            put(j, expand8_k(lo0, lo1, s0, n0, d, mh0, ml0));
            put(4 + j, expand8_k(hi0, hi1, s1, n1, d, mh1, ml1));
        }
    } else if (type == T_Q6_K) {
        const int nib = ((k0 & 255) >> 6) & 1;                         // which nibble plane of the 128-element half
        const float d = h2f(r.w[25] & 0xffffu);
        const __half2 off = __float2half2_rn(1024.f + 32.f);
#pragma unroll
        for (int ch = 0; ch < 8; ch++) {                               // elements e = 8ch..8ch+7 of the plane
            __half2 s_hi, s_lo;
            split_f16(d * (float)(int8_t)(r.w[24] >> (8 * (ch >> 1))), s_hi, s_lo);
            const int kq = 2 * nib + (ch >> 2);                        // 2-bit field of the qh byte
            uint32_t b[2];
#pragma unroll
            for (int i = 0; i < 2; i++) {
                const uint32_t lo = (r.w[2 * ch + i] >> (4 * nib)) & 0x0F0F0F0Fu;
                const uint32_t hi = (r.w[16 + 2 * (ch & 3) + i] >> (2 * kq)) & 0x03030303u;
                b[i] = lo | (hi << 4);
            }
            put(ch, expand8_s(b[0], b[1], off, s_hi, s_lo));
        }
    } else if (type == T_Q8_0) {
        const __half2 d0 = h2_from_bits16(r.w[0] & 0xffffu), d1 = h2_from_bits16(r.w[8] >> 16);
        const __half2 off = __float2half2_rn(1024.f + 128.f);
#pragma unroll
        for (int ch = 0; ch < 8; ch++) {
            uint32_t b0, b1;
            if (ch < 4) { b0 = fshr(r.w[2 * ch], r.w[2 * ch + 1], 16); b1 = fshr(r.w[2 * ch + 1], r.w[2 * ch + 2], 16); }
            else { b0 = r.w[9 + 2 * (ch - 4)]; b1 = r.w[10 + 2 * (ch - 4)]; }
            // signed byte + 128 is an unsigned byte
            put(ch, expand8_d(b0 ^ 0x80808080u, b1 ^ 0x80808080u, off, ch < 4 ? d0 : d1));
        }
    } else {                                                           // Q4_0
        const __half2 d0 = h2_from_bits16(r.w[0] & 0xffffu), d1 = h2_from_bits16(r.w[4] >> 16);
        const __half2 off = __float2half2_rn(1024.f + 8.f);
#pragma unroll
        for (int blk = 0; blk < 2; blk++) {
#pragma unroll
            for (int j = 0; j < 2; j++) {                              // qs bytes 8j..8j+7: low nibbles = elems 8j.., high = 16+8j..
                uint32_t a, b;
                if (blk == 0) { a = fshr(r.w[2 * j], r.w[2 * j + 1], 16); b = fshr(r.w[2 * j + 1], r.w[2 * j + 2], 16); }
                else { a = r.w[5 + 2 * j]; b = r.w[6 + 2 * j]; }
                put(blk * 4 + j, expand8_d(a & 0x0F0F0F0Fu, b & 0x0F0F0F0Fu, off, blk ? d1 : d0));
                put(blk * 4 + 2 + j, expand8_d((a >> 4) & 0x0F0F0F0Fu, (b >> 4) & 0x0F0F0F0Fu, off, blk ? d1 : d0));
            }
        }
    }
}

// Block formats without a hand-written expansion (Q4_1, Q5_0, Q5_1, Q2_K, Q3_K): element-by-element dequantisation of the
// 64 weights of a K block straight from global memory, one f16 rounding per weight like the fast paths. Slow per weight
// (byte loads), but it keeps prompts and batched steps of such models on the tensor cores instead of token by token.
constexpr int T_GENERIC = -1;
__device__ __forceinline__ void dequant_store_generic(int type, const uint8_t* row, int k0, uint8_t* dst, uint32_t lbo) {
    const int be = block_elems(type), bb = block_bytes(type);
#pragma unroll 1
    for (int ch = 0; ch < 8; ch++) {
        uint32_t h[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int e0 = k0 + 8 * ch + 2 * j, e1 = e0 + 1;
            const float a = dequant_elem(type, row + (size_t)(e0 / be) * bb, e0 % be);
            const float b = dequant_elem(type, row + (size_t)(e1 / be) * bb, e1 % be);
            const __half2 v = __floats2half2_rn(a, b);
            h[j] = *reinterpret_cast<const uint32_t*>(&v);
        }
        *reinterpret_cast<uint4*>(dst + ch * lbo) = make_uint4(h[0], h[1], h[2], h[3]);
    }
}

struct GemmBars {
    uint64_t a_full[GM_MAX_STAGES], empty[GM_MAX_STAGES];
    uint64_t b_full[GM_MAX_BSTAGES], b_empty[GM_MAX_BSTAGES];
    uint64_t raw_full[GM_RAW_STAGES], raw_empty[GM_RAW_STAGES];
    uint32_t tmem_base;
};
static_assert(sizeof(GemmBars) <= GM_HDR, "barrier header");

// (tma_load_2d is in ptx.cuh: box (inner = one 256-weight span of a row, outer = 128 rows) -> dense [128][box bytes] in smem)

// words of one 64-weight block inside the 256-weight span W (registers; c is a compile-time constant after unrolling)
template <int C>
__device__ __forceinline__ void span_block(int type, const uint32_t (&W)[GM_SPAN_WORDS], RawKB& r) {
    if (type == T_Q4_K) {
#pragma unroll
        for (int i = 0; i < 4; i++) r.w[i] = W[i];
#pragma unroll
        for (int i = 0; i < 8; i++) r.w[4 + i] = W[4 + 8 * C + i];
    } else if (type == T_Q5_K) {
#pragma unroll
        for (int i = 0; i < 4; i++) r.w[i] = W[i];
#pragma unroll
        for (int i = 0; i < 8; i++) { r.w[4 + i] = W[12 + 8 * C + i]; r.w[12 + i] = W[4 + i]; }
    } else if (type == T_Q6_K) {
        constexpr int hf = C >> 1, nib = C & 1;
#pragma unroll
        for (int i = 0; i < 16; i++) r.w[i] = W[16 * hf + i];
#pragma unroll
        for (int i = 0; i < 8; i++) r.w[16 + i] = W[32 + 8 * hf + i];
        r.w[24] = W[48 + 2 * hf + nib];
        r.w[25] = W[52];
    } else if (type == T_Q8_0) {
#pragma unroll
        for (int i = 0; i < 17; i++) r.w[i] = W[17 * C + i];
    } else {
#pragma unroll
        for (int i = 0; i < 9; i++) r.w[i] = W[9 * C + i];
    }
}

struct DeqCtx {
    GemmBars* sh;
    uint8_t* stage0;
    const uint8_t* raw0;
    const uint8_t* wrow;
    int S, NR, tid, n, n_sp, sp0, kb0, use_tma, span_bytes, type;
    uint32_t raw_stage_bytes, box, a_bytes, lbo_a;      // weight tile: rows * 128 B per stage, rows * 16 B between K chunks
    unsigned int* err;
    long long* dbg;
};

template <int TYPE>
__device__ __noinline__ void deq_loop(const DeqCtx cx) {
    GemmBars& sh = *cx.sh;
    const uint32_t a_bytes = cx.a_bytes, lbo_a = cx.lbo_a;
    int st = 0;                 // A stage of the next block and how often the ring has wrapped
    uint32_t wraps = 0;
    auto emit = [&](const RawKB& r, int i) {               // block i of this CTA's range -> A tile of its stage
        if (cx.dbg && i < 64) cx.dbg[i * 8 + 0] = clock64();
        if (wraps > 0) mbar_wait(&sh.empty[st], (wraps - 1) & 1, cx.err, 0x40001u, (unsigned)i);   // MMAs that read the stage are done
        if (cx.dbg && i < 64) cx.dbg[i * 8 + 1] = clock64();
        if (TYPE == T_GENERIC) dequant_store_generic(cx.type, cx.wrow, (cx.kb0 + i) * GM_BK, cx.stage0 + (size_t)st * a_bytes + cx.tid * 16, lbo_a);
        else dequant_store(TYPE, r, (i & 3) * GM_BK, cx.stage0 + (size_t)st * a_bytes + cx.tid * 16, lbo_a);
        fence_proxy_async();            // generic-proxy writes -> visible to the tensor core (async proxy)
        mbar_arrive(&sh.a_full[st]);
        if (cx.dbg && i < 64) cx.dbg[i * 8 + 2] = clock64();
        if (++st == cx.S) { st = 0; wraps++; }
    };
    if (TYPE == T_GENERIC) {
        RawKB none;
        for (int i = 0; i < cx.n; i++) emit(none, i);
    } else if (cx.use_tma) {
        // raw spans arrive in shared memory by TMA; this thread's row is one box row. The box starts at the 16-byte
        // boundary below the span (TMA needs it); only Q6_K spans (210 B) can be off it.
        constexpr int span = TYPE == T_Q4_K ? 144 : TYPE == T_Q5_K ? 176 : TYPE == T_Q6_K ? 210 : TYPE == T_Q8_0 ? 272 : 144;
        constexpr int n16 = TYPE == T_Q6_K ? 15 : (span + 15) / 16;
        int rs = 0;
        uint32_t rwraps = 0;
        for (int sp = 0; sp < cx.n_sp; sp++) {
            if (cx.dbg && sp < 16) cx.dbg[sp * 4 * 8 + 6] = clock64();
            mbar_wait(&sh.raw_full[rs], rwraps & 1, cx.err, 0x40006u, (unsigned)sp);
            if (cx.dbg && sp < 16) cx.dbg[sp * 4 * 8 + 7] = clock64();
            const uint8_t* slot = cx.raw0 + (size_t)rs * cx.raw_stage_bytes + (size_t)cx.tid * cx.box;
            uint32_t W[GM_SPAN_WORDS];
#pragma unroll
            for (int j = 0; j < n16; j++) {
                const uint4 v = *reinterpret_cast<const uint4*>(slot + 16 * j);
                W[4 * j] = v.x; W[4 * j + 1] = v.y; W[4 * j + 2] = v.z; W[4 * j + 3] = v.w;
            }
            fence_proxy_async();                           // these generic-proxy reads before the TMA (async proxy) rewrite
            mbar_arrive(&sh.raw_empty[rs]);                // the span is in registers: the stage can be refilled
            if (++rs == cx.NR) { rs = 0; rwraps++; }
            if (TYPE == T_Q6_K) {                          // shift the 53 words of the span down by its misalignment (0,2,..14 bytes)
                const uint32_t mis = ((uint32_t)(cx.sp0 + sp) * 210u) & 15u, sh16 = (mis & 2u) * 8u;
                switch (mis >> 2) {
                    case 0:
#pragma unroll
                        for (int i = 0; i < 53; i++) W[i] = fshr(W[i], W[i + 1], sh16);
                        break;
                    case 1:
#pragma unroll
                        for (int i = 0; i < 53; i++) W[i] = fshr(W[i + 1], W[i + 2], sh16);
                        break;
                    case 2:
#pragma unroll
                        for (int i = 0; i < 53; i++) W[i] = fshr(W[i + 2], W[i + 3], sh16);
                        break;
                    default:
#pragma unroll
                        for (int i = 0; i < 53; i++) W[i] = fshr(W[i + 3], W[i + 4], sh16);
                        break;
                }
            }
            RawKB r;
            span_block<0>(TYPE, W, r); emit(r, sp * 4 + 0);
            span_block<1>(TYPE, W, r); emit(r, sp * 4 + 1);
            span_block<2>(TYPE, W, r); emit(r, sp * 4 + 2);
            span_block<3>(TYPE, W, r); emit(r, sp * 4 + 3);
        }
    } else {
        // rows that are not 16-byte multiples (Q6_K with K % 2048 != 0): direct loads, next block prefetched in registers
        RawKB cur, nxt;
        load_raw(TYPE, cx.wrow, cx.kb0 * GM_BK, cur);
        for (int i = 0; i < cx.n; i++) {
            if (i + 1 < cx.n) load_raw(TYPE, cx.wrow, (cx.kb0 + i + 1) * GM_BK, nxt);
            emit(cur, i);
            cur = nxt;
        }
    }
}

// ROWS weight rows per CTA, ROWS / 32 dequantising warps + 3 service warps.
// SWAP = false (T > 128): weights are the MMA's A operand (M = 128 rows = TMEM lanes), tokens its N (TMEM columns).
// SWAP = true  (T <= 128): tokens are the A operand (M = 128, rows past T are don't-care), 256 weight rows are the N:
//   with few tokens an instruction costs ~130 cycles whatever N is (measured), so each MMA should cover as many
//   weight rows as an instruction can: 256.
template <int ROWS, bool SWAP>
__global__ void __launch_bounds__((ROWS / 32 + 3) * 32) gemm_q_kernel(const __grid_constant__ GemmParams p) {
    constexpr int DW = ROWS / 32;          // dequantising warps; then: MMA warp, weight producer, activation producer
    extern __shared__ __align__(128) uint8_t smem_raw[];
    GemmBars& sh = *reinterpret_cast<GemmBars*>(smem_raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int T = p.T_pad, S = p.n_stages, SB = p.n_bstages;
    const uint32_t a_bytes = ROWS * GM_BK * 2, b_bytes = (uint32_t)T * GM_BK * 2;
    uint8_t* stage0 = smem_raw + GM_HDR;                               // A tiles: S x 16 KB
    uint8_t* bst0 = stage0 + (size_t)S * a_bytes;                      // activation tiles: SB x T*128 B
    uint8_t* raw0 = bst0 + (size_t)SB * b_bytes;                       // raw weight spans: n_raw x [128][box]
    uint32_t ncols = 32;
    while ((int)ncols < (SWAP ? ROWS : T)) ncols <<= 1;

    // ---- which tile / K range (in 256-weight spans = 4 blocks)
    const int split = blockIdx.x % p.n_splits, tile_g = blockIdx.x / p.n_splits;
    int seg = 0;
    while (seg + 1 < p.nseg && tile_g >= p.seg[seg + 1].tile0) seg++;
    const GemmSeg& sg = p.seg[seg];
    const int m0 = (tile_g - sg.tile0) * ROWS;
    const int n_sp_all = p.K / 256;
    const int sp0 = (int)((long)n_sp_all * split / p.n_splits), sp1 = (int)((long)n_sp_all * (split + 1) / p.n_splits);
    const int n_sp = sp1 - sp0, kb0 = sp0 * 4, n = n_sp * 4, NR = p.n_raw;
    const uint32_t box = (uint32_t)sg.box_bytes, raw_stage_bytes = ROWS * box;

    if (warp == DW) {
        if (lane == 0) {
            for (int i = 0; i < S; i++) { mbar_init(&sh.a_full[i], ROWS); mbar_init(&sh.empty[i], 1); }
            for (int i = 0; i < SB; i++) { mbar_init(&sh.b_full[i], 1); mbar_init(&sh.b_empty[i], 1); }
            for (int i = 0; i < GM_RAW_STAGES; i++) { mbar_init(&sh.raw_full[i], 1); mbar_init(&sh.raw_empty[i], ROWS); }
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc(&sh.tmem_base, ncols);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    // Programmatic dependent launch: the next kernel of the chain may start now. In THIS kernel only the threads that
    // touch the previous kernel's data wait for it (the activation producer; the epilogue, which overwrites the planes
    // the previous consumer read): weights stream and dequantise while the predecessor is still running.
    griddep_launch();
    const uint32_t tmem = sh.tmem_base;
    // K-major, no swizzle: element (r, k) of a tile with R rows sits at (k/8) * (R*16) + r*16 + (k%8)*2
    const uint32_t lbo_a = ROWS * 16, lbo_b = (uint32_t)T * 16, sbo = 128;

    if (warp < DW) {
        // ================= dequantising threads: thread <-> weight row. One instantiation per quant type keeps the
        // loop a CTA actually runs small enough for the instruction cache (all five inlined: ~100 KB of SASS).
        DeqCtx cx;
        cx.sh = &sh; cx.stage0 = stage0; cx.raw0 = raw0; cx.S = S; cx.NR = NR; cx.raw_stage_bytes = raw_stage_bytes; cx.box = box;
        cx.tid = tid; cx.n = n; cx.n_sp = n_sp; cx.sp0 = sp0; cx.kb0 = kb0; cx.err = p.err; cx.a_bytes = a_bytes; cx.lbo_a = lbo_a;
        cx.dbg = (p.dbg && blockIdx.x == 0 && tid == 0) ? p.dbg : nullptr;
        cx.use_tma = sg.use_tma; cx.span_bytes = sg.span_bytes; cx.type = sg.type;
        cx.wrow = sg.w + (size_t)min(m0 + tid, sg.n_rows - 1) * sg.row_bytes;      // rows past N are computed and dropped
        switch (sg.type) {
            case T_Q4_K: deq_loop<T_Q4_K>(cx); break;
            case T_Q5_K: deq_loop<T_Q5_K>(cx); break;
            case T_Q6_K: deq_loop<T_Q6_K>(cx); break;
            case T_Q8_0: deq_loop<T_Q8_0>(cx); break;
            case T_Q4_0: deq_loop<T_Q4_0>(cx); break;
            default: deq_loop<T_GENERIC>(cx); break;
        }
        // ================= epilogue
        const int last = n - 1;
        mbar_wait(&sh.empty[last % S], (last / S) & 1, p.err, 0x40002u, (unsigned)last);
        tc_fence_after();
        griddep_wait();
        if (!SWAP) {                               // TMEM lane = weight row, columns = tokens
            const int row = m0 + tid;
            const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);
            float* ybase = p.y + (size_t)split * p.split_stride + sg.out_off + row;
            for (int c0 = 0; c0 < T; c0 += 16) {
                float v[16];
                tmem_ld16(lane_addr + c0, v);
                if (row < sg.n_rows) {
#pragma unroll
                    for (int j = 0; j < 16; j++)
                        if (c0 + j < p.T) ybase[(size_t)(c0 + j) * p.ldy] = v[j];
                }
            }
        } else if (warp < 4 && warp * 32 < p.T) {  // TMEM lane = token (warps 0-3 own the 128 lanes), columns = weight rows
            const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);
            float* yrow = p.y + (size_t)split * p.split_stride + (size_t)tid * p.ldy + sg.out_off + m0;
            const int n_valid = min(ROWS, sg.n_rows - m0);
            for (int c0 = 0; c0 < ROWS; c0 += 16) {
                float v[16];
                tmem_ld16(lane_addr + c0, v);      // warp-wide: every lane takes part, only tokens < T store
                if (tid < p.T) {
                    if (c0 + 16 <= n_valid) {
#pragma unroll
                        for (int j = 0; j < 16; j += 4) *reinterpret_cast<float4*>(yrow + c0 + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                    } else {
#pragma unroll
                        for (int j = 0; j < 16; j++) if (c0 + j < n_valid) yrow[c0 + j] = v[j];
                    }
                }
            }
        }
    } else if (warp == DW && lane == 0) {
        // ================= MMA thread
        const uint32_t idesc_lo = SWAP ? instr_desc(128, ROWS) : instr_desc(ROWS, T > 256 ? 256 : T);
        const uint32_t idesc_hi = (!SWAP && T > 256) ? instr_desc(ROWS, T - 256) : 0u;
        int s = 0, sb = 0;
        uint32_t use = 0, useb = 0;
        const uint32_t a_lo0 = (smem_u32(stage0) >> 4) | ((lbo_a >> 4) << 16), b_lo0 = (smem_u32(bst0) >> 4) | ((lbo_b >> 4) << 16);
        const uint32_t desc_hi = (sbo >> 4) | (1u << 14);          // SBO, descriptor version 1, no swizzle
        for (int i = 0; i < n; i++) {
            mbar_wait(&sh.b_full[sb], useb & 1, p.err, 0x40003u, (unsigned)i);
            if (p.dbg && blockIdx.x == 0 && i < 64) p.dbg[i * 8 + 3] = clock64();
            mbar_wait(&sh.a_full[s], use & 1, p.err, 0x40004u, (unsigned)i);
            if (p.dbg && blockIdx.x == 0 && i < 64) p.dbg[i * 8 + 4] = clock64();
            tc_fence_after();
            // descriptors differ only in the 14-bit start-address field (units of 16 B): add, do not rebuild
            const uint32_t a_lo = a_lo0 + (uint32_t)s * (a_bytes >> 4), b_lo = b_lo0 + (uint32_t)sb * (b_bytes >> 4);
#pragma unroll
            for (int ks = 0; ks < GM_BK / 16; ks++) {
                const uint64_t da = desc_join(a_lo + ks * (2 * lbo_a >> 4), desc_hi);
                const uint64_t db = desc_join(b_lo + ks * (2 * lbo_b >> 4), desc_hi);
                const uint32_t acc = (i > 0 || ks > 0) ? 1u : 0u;
                if (SWAP) {
                    umma_f16(tmem, db, da, idesc_lo, acc);          // A = tokens (rows past T read whatever follows: dropped), B = weights
                } else {
                    umma_f16(tmem, da, db, idesc_lo, acc);
                    if (T > 256) umma_f16(tmem + 256, da, desc_join(b_lo + ks * (2 * lbo_b >> 4) + 256, desc_hi), idesc_hi, acc);
                }
            }
            umma_commit(&sh.empty[s]);      // both arrive when the MMAs above have finished reading their tiles
            umma_commit(&sh.b_empty[sb]);
            if (p.dbg && blockIdx.x == 0 && i < 64) p.dbg[i * 8 + 5] = clock64();
            if (++s == S) { s = 0; use++; }
            if (++sb == SB) { sb = 0; useb++; }
        }
    } else if (warp == DW + 2 && lane == 0) {
        // ================= activation producer: the tile of a block is one contiguous 1-D bulk copy (operand order in HBM)
        const uint8_t* xb = reinterpret_cast<const uint8_t*>(p.x);
        int sb = 0;
        uint32_t useb = 0;
        griddep_wait();                 // the activations come from the previous kernel
        for (int i = 0; i < n; i++) {
            if (useb > 0) mbar_wait(&sh.b_empty[sb], (useb - 1) & 1, p.err, 0x40005u, (unsigned)i);
            mbar_arrive_expect_tx(&sh.b_full[sb], b_bytes);
            tma_load_1d(bst0 + (size_t)sb * b_bytes, xb + (size_t)(kb0 + i) * b_bytes, b_bytes, &sh.b_full[sb]);
            if (++sb == SB) { sb = 0; useb++; }
        }
    } else if (warp == DW + 1 && lane == 0 && sg.use_tma) {
        // ================= weight producer: one 2-D TMA box (128 rows x one 256-weight span) per stage
        const void* tmap = &sg.tmap;
        int rs = 0;
        uint32_t ruse = 0;
        for (int sp = 0; sp < n_sp; sp++, rs = (rs + 1 == NR ? 0 : rs + 1), ruse += (rs == 0)) {
            if (ruse > 0) mbar_wait(&sh.raw_empty[rs], (ruse - 1) & 1, p.err, 0x40007u, (unsigned)sp);
            mbar_arrive_expect_tx(&sh.raw_full[rs], raw_stage_bytes);
            tma_load_2d(raw0 + (size_t)rs * raw_stage_bytes, tmap, (int)((((size_t)(sp0 + sp) * sg.span_bytes) & ~(size_t)15) >> 1), m0, &sh.raw_full[rs]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == DW) tmem_dealloc(tmem, ncols);
}

// ---------------------------------------------------------------- host side
static int gemm_rows(const GemmParams& p) { return p.swap ? 256 : GM_BM; }
static size_t gemm_raw_bytes(const GemmParams& p) {
    size_t m = 0;
    for (int s = 0; s < p.nseg; s++) if (p.seg[s].use_tma) m = std::max(m, (size_t)p.n_raw * gemm_rows(p) * p.seg[s].box_bytes);
    return m;
}
size_t gemm_smem_bytes(const GemmParams& p) {
    // swap: the MMA reads 128 token rows per K chunk whatever T is; the rows past the tile must still be shared memory
    return GM_HDR + (size_t)p.n_stages * gemm_rows(p) * GM_BK * 2 + (size_t)p.n_bstages * p.T_pad * GM_BK * 2 + gemm_raw_bytes(p) +
           (p.swap ? 4096 : 0);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;

cudaError_t gemm_init() {
    if (!g_encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (e != cudaSuccess) return e;
        if (qres != cudaDriverEntryPointSuccess || !fn) return cudaErrorNotSupported;
        g_encode = (EncodeTiledFn)fn;
    }
    cudaError_t e = cudaFuncSetAttribute(gemm_q_kernel<GM_BM, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(gemm_q_kernel<256, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
}

// Fill tile ranges, tensor maps, the stage count and the K split. n_sms: SMs to fill; max_splits: what the partial
// buffer allows.
int gemm_plan(GemmParams& p, int n_sms, int max_splits) {
    static_assert(sizeof(CUtensorMap) == sizeof(p.seg[0].tmap), "tensor map storage");
    if (p.nseg < 1 || p.nseg > 3 || p.K <= 0 || p.K % 256) return -1;
    if (p.T < 1 || p.T > 512) return -2;
    p.T_pad = (p.T + 15) / 16 * 16;
    p.swap = (p.T_pad <= 128 && !getenv("B2_GEMM_NOSWAP")) ? 1 : 0;
    const int ROWS = gemm_rows(p);
    int tiles = 0;
    const bool no_tma = getenv("B2_GEMM_NO_TMA") != nullptr;
    for (int s = 0; s < p.nseg; s++) {
        GemmSeg& g = p.seg[s];
        const int be = block_elems(g.type), bb = block_bytes(g.type);
        if (be == 1 || !bb || g.n_rows <= 0) return -3;
        g.row_bytes = p.K / be * bb;
        g.tile0 = tiles;
        g.n_tiles = (g.n_rows + ROWS - 1) / ROWS;
        tiles += g.n_tiles;
        g.span_bytes = 256 / be * bb;                                  // one 256-weight span of a row
        // box row = span rounded up to 16 B, and an odd multiple of 16 B so 128-bit shared loads of 8 rows hit 8 bank groups
        int box = (g.span_bytes + 15) / 16 * 16;
        if ((box / 16) % 2 == 0 || g.type == T_Q6_K) box = g.type == T_Q6_K ? 240 : box + 16;   // Q6_K: room for the 14-byte misalignment
        g.box_bytes = box;
        g.use_tma = 0;
        const bool fast_type = g.type == T_Q4_0 || g.type == T_Q8_0 || g.type == T_Q4_K || g.type == T_Q5_K || g.type == T_Q6_K;   // the others: dequant_store_generic
        if (fast_type && !no_tma && g_encode && g.row_bytes % 16 == 0 && ((uintptr_t)g.w & 15) == 0 && (g.type != T_Q6_K || !getenv("B2_GEMM_Q6K_DIRECT"))) {
            const cuuint64_t dims[2] = {(cuuint64_t)g.row_bytes / 2, (cuuint64_t)g.n_rows};
            const cuuint64_t strides[1] = {(cuuint64_t)g.row_bytes};
            const cuuint32_t boxd[2] = {(cuuint32_t)box / 2, (cuuint32_t)ROWS};
            const cuuint32_t estr[2] = {1, 1};
            CUresult r = g_encode(reinterpret_cast<CUtensorMap*>(g.tmap), CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, (void*)g.w, dims, strides,
                                  boxd, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            g.use_tma = r == CUDA_SUCCESS ? 1 : 0;
        }
    }
    p.n_tiles = tiles;
    const size_t a_st = (size_t)ROWS * GM_BK * 2, b_st = (size_t)p.T_pad * GM_BK * 2;
    const size_t cap = 227 * 1024 - 1024;
    // activation ring: deep when the tiles are small (their TMA latency must not gate the MMAs)
    int SB = p.T_pad <= 32 ? 8 : p.T_pad <= 64 ? 6 : p.T_pad <= 128 ? 4 : p.T_pad <= 256 ? 3 : 2;
    int S = p.swap ? 3 : (p.T_pad <= 64 ? 2 : 3);
    int per_sm = (!p.swap && p.T_pad <= 64) ? 2 : 1;   // 128-row variant at small T: two CTAs per SM
    if (const char* v = getenv("B2_GEMM_PER_SM")) if (atoi(v) > 0) per_sm = atoi(v);
    if (const char* v = getenv("B2_GEMM_STAGES")) if (atoi(v) > 1) S = atoi(v);
    if (const char* v = getenv("B2_GEMM_BSTAGES")) if (atoi(v) > 1) SB = atoi(v);
    if (S > GM_MAX_STAGES) S = GM_MAX_STAGES;
    if (SB > GM_MAX_BSTAGES) SB = GM_MAX_BSTAGES;
    p.n_raw = GM_RAW_STAGES;
    if (const char* v = getenv("B2_GEMM_RAW")) if (atoi(v) == 1) p.n_raw = 1;
    const size_t budget = per_sm > 1 ? (228 * 1024) / per_sm - 1024 : cap;
    auto total = [&]() { return GM_HDR + S * a_st + SB * b_st + gemm_raw_bytes(p) + (p.swap ? 4096 : 0); };
    if (total() > budget && p.n_raw > 1) p.n_raw = 1;
    while (total() > budget && SB > 2) SB--;
    while (total() > budget && S > 2) S--;
    if (total() > budget && per_sm > 1) per_sm = 1;
    if (total() > cap) { for (int s = 0; s < p.nseg; s++) p.seg[s].use_tma = 0; }
    if (total() > cap) return -4;
    p.n_stages = S; p.n_bstages = SB;
    const int slots = n_sms * per_sm;
    const int n_sp = p.K / 256;
    // choose the K split that minimises (waves) x (blocks per CTA + fixed cost), fixed cost ~ epilogue + ramp
    const int fixed = 6 + p.T_pad / 32;
    long best_cost = -1;
    int best = 1;
    for (int s = 1; s <= max_splits && s <= 16 && s <= n_sp; s++) {
        const long waves = ((long)tiles * s + slots - 1) / slots;
        const long cost = waves * (4 * ((n_sp + s - 1) / s) + fixed);
        if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = s; }
    }
    p.n_splits = best;
    return 0;
}

cudaError_t gemm_launch(const GemmParams& p, cudaStream_t stream) {
    if (p.n_stages < 2 || p.n_stages > GM_MAX_STAGES || p.n_bstages < 2 || p.n_bstages > GM_MAX_BSTAGES || p.n_splits < 1 ||
        p.n_tiles < 1) return cudaErrorInvalidValue;
    if (p.T_pad < 16 || p.T_pad > 512 || p.T_pad % 16 || p.T > p.T_pad || p.n_splits > p.K / 256) return cudaErrorInvalidValue;
    if (p.swap) return launch_chain(gemm_q_kernel<256, true>, dim3(p.n_tiles * p.n_splits), dim3((256 / 32 + 3) * 32), gemm_smem_bytes(p), stream, p);
    return launch_chain(gemm_q_kernel<GM_BM, false>, dim3(p.n_tiles * p.n_splits), dim3((GM_BM / 32 + 3) * 32), gemm_smem_bytes(p), stream, p);
}

// x [T][K] float32 -> f16 in the MMA operand order [K/8][T_pad][8] (rows >= T zero)
__global__ void to_half_blocked_kernel(const float* x, int T, int T_pad, int K, __half* y) {
    griddep_launch();
    griddep_wait();
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;      // one output element
    if (i >= (long)T_pad * K) return;
    const int e = (int)(i & 7);
    const long r = i >> 3;
    const int t = (int)(r % T_pad), ch = (int)(r / T_pad);
    y[i] = t < T ? __float2half_rn(x[(size_t)t * K + ch * 8 + e]) : __float2half_rn(0.f);
}
cudaError_t to_half_blocked_launch(const float* x, int T, int T_pad, int K, __half* y, cudaStream_t s) {
    const long n = (long)T_pad * K;
    return launch_chain(to_half_blocked_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, s, x, T, T_pad, K, y);
}

}  // namespace b2
