// ggml quant-block arithmetic shared by every kernel (and by the host self-test).
//
// Two families of routines:
//  * dequant_elem<T>(blk, e): the value of element e of one block, in float32 with a separate
//    multiply and subtract -- bit-identical to gguf-py's numpy dequantize (the dequant parity
//    bar is bit-exact). Used by the dequantise / embedding-row kernels.
//  * chunk_dot<T>(row, u, act): integer dot product of one 64-weight chunk of a quantised row
//    with a quantised activation chunk, scaled back to float. This is the inner loop of the
//    decode mat-vec: weights stay packed (nibbles/bytes), activations are two int8 planes
//    (x ~= dx * (a + b/252), ~16 bits), products go through dp4a, so the only floating-point
//    work is one fused scale per 16- or 32-element group.
//
// Block layouts: see oracle/ggml_blocks.py (restated from gguf-py quants.py) -- replaces
// ggml-cuda's dequantize_block_* / vec_dot_*_q8_1 that the reference reaches through
// llama-cpp-python (SURVEY.md 2.3 K1,K4,K5).
//
// Everything here compiles for the host too (B2_HOST_EMU) so tests/host_selftest.cpp can check
// the bit-twiddling against the oracle without a GPU.
#pragma once
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__) && !defined(B2_HOST_EMU)
#include <cuda_fp16.h>
#define B2_HD __host__ __device__ __forceinline__
#define B2_D __device__ __forceinline__
#else
#define B2_HD inline
#define B2_D inline
#endif

namespace b2 {

enum GgmlType { T_F32 = 0, T_F16 = 1, T_Q4_0 = 2, T_Q4_1 = 3, T_Q5_0 = 6, T_Q5_1 = 7, T_Q8_0 = 8, T_Q2_K = 10, T_Q3_K = 11,
                T_Q4_K = 12, T_Q5_K = 13, T_Q6_K = 14 };

B2_HD int block_elems(int t) {
    return (t == T_F32 || t == T_F16) ? 1 : (t == T_Q4_0 || t == T_Q8_0 || t == T_Q4_1 || t == T_Q5_0 || t == T_Q5_1) ? 32 : 256;
}
B2_HD int block_bytes(int t) {
    switch (t) {
        case T_F32: return 4;
        case T_F16: return 2;
        case T_Q4_0: return 18;
        case T_Q4_1: return 20;
        case T_Q5_0: return 22;
        case T_Q5_1: return 24;
        case T_Q8_0: return 34;
        case T_Q2_K: return 84;
        case T_Q3_K: return 110;
        case T_Q4_K: return 144;
        case T_Q5_K: return 176;
        case T_Q6_K: return 210;
        default: return 0;
    }
}

// ---------------------------------------------------------------- primitives
#if defined(__CUDA_ARCH__) && !defined(B2_HOST_EMU)
B2_D float h2f(uint32_t h) { return __half2float(__ushort_as_half((unsigned short)h)); }
B2_D float u2f(uint32_t u) { return __uint_as_float(u); }
B2_D int dp4a(int a, int b, int c) { return __dp4a(a, b, c); }
B2_D int dp4a_us(uint32_t a, int b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
B2_D uint32_t fshr(uint32_t lo, uint32_t hi, uint32_t s) { return __funnelshift_r(lo, hi, s); }
B2_D float fmul(float a, float b) { return __fmul_rn(a, b); }
B2_D float fsub(float a, float b) { return __fsub_rn(a, b); }
B2_D uint32_t ld32(const uint8_t* p) { return *reinterpret_cast<const uint32_t*>(p); }
B2_D uint32_t ld16(const uint8_t* p) { return *reinterpret_cast<const uint16_t*>(p); }
struct U4 { uint32_t x, y, z, w; };
B2_D U4 ld128(const uint8_t* p) {
    uint4 v = *reinterpret_cast<const uint4*>(p);
    U4 r; r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w; return r;
}
#else
inline float h2f(uint32_t h) {
    uint32_t s = (h >> 15) & 1, e = (h >> 10) & 31, m = h & 1023, o;
    if (e == 0) {
        if (m == 0) o = s << 31;
        else { int k = 0; while (!(m & 1024)) { m <<= 1; k++; } m &= 1023; o = (s << 31) | ((uint32_t)(113 - k) << 23) | (m << 13); }
    } else if (e == 31) o = (s << 31) | 0x7f800000u | (m << 13);
    else o = (s << 31) | ((e + 112) << 23) | (m << 13);
    float f; memcpy(&f, &o, 4); return f;
}
inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
inline int dp4a(int a, int b, int c) {
    for (int i = 0; i < 4; i++) c += (int)(int8_t)(a >> (8 * i)) * (int)(int8_t)(b >> (8 * i));
    return c;
}
inline int dp4a_us(uint32_t a, int b, int c) {
    for (int i = 0; i < 4; i++) c += (int)(uint8_t)(a >> (8 * i)) * (int)(int8_t)(b >> (8 * i));
    return c;
}
inline uint32_t fshr(uint32_t lo, uint32_t hi, uint32_t s) {
    s &= 31; return s ? (lo >> s) | (hi << (32 - s)) : lo;
}
inline float fmul(float a, float b) { volatile float r = a * b; return r; }
inline float fsub(float a, float b) { volatile float r = a - b; return r; }
inline uint32_t ld32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }
inline uint32_t ld16(const uint8_t* p) { uint16_t v; memcpy(&v, p, 2); return v; }
struct U4 { uint32_t x, y, z, w; };
inline U4 ld128(const uint8_t* p) { U4 r; memcpy(&r, p, 16); return r; }
#endif

// load n 32-bit words starting at a 2-byte-aligned address (aligned loads + funnel shift)
template <int N>
B2_D void ld_words_u2(const uint8_t* p, uint32_t (&w)[N]) {
    uint32_t mis = (uint32_t)(reinterpret_cast<uintptr_t>(p) & 3);
    const uint8_t* pa = p - mis;          // pointer arithmetic keeps the (shared) address space
    uint32_t sh = mis * 8;                // 0 or 16
    uint32_t prev = ld32(pa);
#pragma unroll
    for (int i = 0; i < N; i++) {
        uint32_t next = ld32(pa + 4 * (i + 1));
        w[i] = fshr(prev, next, sh);
        prev = next;
    }
}

// 6-bit (scale, min) of sub-block `is` from the 12-byte K-quant scale field
B2_HD void k_scale_min(const uint8_t* s, int is, uint32_t& sc, uint32_t& m) {
    if (is < 4) {
        sc = s[is] & 63; m = s[is + 4] & 63;
    } else {
        sc = (s[is + 4] & 15) | ((s[is - 4] >> 6) << 4);
        m = (s[is + 4] >> 4) | ((s[is] >> 6) << 4);
    }
}

// ---------------------------------------------------------------- exact element dequant
// blk points at one block (any alignment: byte loads only), e = element index in the block.
B2_D float dequant_elem(int type, const uint8_t* blk, int e) {
    switch (type) {
        case T_F32: { uint32_t v = blk[0] | (blk[1] << 8) | (blk[2] << 16) | ((uint32_t)blk[3] << 24); float f; memcpy(&f, &v, 4); return f; }
        case T_F16: return h2f(blk[0] | (blk[1] << 8));
        case T_Q4_0: {
            float d = h2f(blk[0] | (blk[1] << 8));
            int q = (e < 16) ? (blk[2 + e] & 15) : (blk[2 + e - 16] >> 4);
            return fmul(d, (float)(q - 8));
        }
        case T_Q8_0: {
            float d = h2f(blk[0] | (blk[1] << 8));
            return fmul((float)(int8_t)blk[2 + e], d);
        }
        case T_Q4_1: {
            float d = h2f(blk[0] | (blk[1] << 8)), m = h2f(blk[2] | (blk[3] << 8));
            int q = (e < 16) ? (blk[4 + e] & 15) : (blk[4 + e - 16] >> 4);
            return fsub(fmul(d, (float)q), -m);
        }
        case T_Q5_0:
        case T_Q5_1: {
            const int o = type == T_Q5_0 ? 2 : 4;
            float d = h2f(blk[0] | (blk[1] << 8));
            uint32_t qh = blk[o] | (blk[o + 1] << 8) | (blk[o + 2] << 16) | ((uint32_t)blk[o + 3] << 24);
            int q = (e < 16) ? (blk[o + 4 + e] & 15) : (blk[o + 4 + e - 16] >> 4);
            q |= (int)((qh >> e) & 1u) << 4;
            if (type == T_Q5_0) return fmul(d, (float)(q - 16));
            return fsub(fmul(d, (float)q), -h2f(blk[2] | (blk[3] << 8)));
        }
        case T_Q2_K: {
            float d = h2f(blk[80] | (blk[81] << 8)), dmin = h2f(blk[82] | (blk[83] << 8));
            int sc = blk[e >> 4];
            int q = (blk[16 + 32 * (e >> 7) + (e & 31)] >> (2 * ((e >> 5) & 3))) & 3;
            float dl = fmul(d, (float)(sc & 15)), ml = fmul(dmin, (float)(sc >> 4));
            return fsub(fmul(dl, (float)q), ml);
        }
        case T_Q3_K: {
            float d = h2f(blk[108] | (blk[109] << 8));
            int is = e >> 4;
            int lsc = (blk[96 + (is & 7)] >> (4 * (is >> 3))) & 15, hsc = (blk[104 + (is & 3)] >> (2 * (is >> 2))) & 3;
            int scale = (lsc | (hsc << 4)) - 32;
            int ql = (blk[32 + 32 * (e >> 7) + (e & 31)] >> (2 * ((e >> 5) & 3))) & 3;
            int hb = (blk[e & 31] >> (e >> 5)) & 1;
            return fmul(fmul(d, (float)scale), (float)(ql - (hb ? 0 : 4)));
        }
        case T_Q4_K:
        case T_Q5_K: {
            float d = h2f(blk[0] | (blk[1] << 8)), dmin = h2f(blk[2] | (blk[3] << 8));
            int is = e >> 5, l = e & 31;
            uint32_t sc, m;
            k_scale_min(blk + 4, is, sc, m);
            const uint8_t* qs = blk + (type == T_Q4_K ? 16 : 48);
            int byte = qs[(is >> 1) * 32 + l];
            int q = (is & 1) ? (byte >> 4) : (byte & 15);
            if (type == T_Q5_K) q |= ((blk[16 + l] >> is) & 1) << 4;
            float dl = fmul(d, (float)sc), ml = fmul(dmin, (float)m);
            return fsub(fmul(dl, (float)q), ml);
        }
        case T_Q6_K: {
            int hf = e >> 7, ep = e & 127, nib = ep >> 6, kq = ep >> 5, l = ep & 31;
            int lo = blk[64 * hf + (ep & 63)];
            lo = nib ? (lo >> 4) : (lo & 15);
            int hi = (blk[128 + 32 * hf + l] >> (2 * kq)) & 3;
            int q = (lo | (hi << 4)) - 32;
            float d = h2f(blk[208] | (blk[209] << 8));
            float dl = fmul(d, (float)(int8_t)blk[192 + (e >> 4)]);
            return fmul(dl, (float)q);
        }
    }
    return 0.f;
}

// ---------------------------------------------------------------- quantised activations
// The mat-vec input vector lives in shared memory as two int8 planes per 32-element block,
//   x[i] ~= dx[i/32] * (a[i] + b[i] / ACT_RES)        |a| <= 127, |b| <= 126   (~16 bits)
// plus the exact float sum of x over every 16-element group (for the "-min", "-8", "-32" terms).
#define B2_ACT_RES 252.0f
#define B2_SLICE 2048            // elements of K one warp step covers (32 lanes x 64)
struct ActView {
    const uint8_t* a;            // [K] int8
    const uint8_t* b;            // [K] int8
    // per-block scales, packed so a lane fetches what it needs with one or two 16-byte loads:
    //   layout 0: f[4c .. 4c+3] = {dx'(2c), dx'(2c+1), sum32(2c), sum32(2c+1)} for the 64-element chunk c
    //   layout 1: f[blk] = dx'(blk);  g[act_s16p_index(u)] = 32 * sum16(u)
    //   layout 0 + 16-element sums (Q2_K, Q3_K): additionally g[u] = sum16(u), plain order
    // dx' = dx (one plane) or dx / 252 (two planes: the scale of the folded integer 252 a + b)
    const float* f;
    const float* g;
};
// index of 16-element unit u in the layout-1 sum array: lane (j, hf, qt) reads its four kq values as one float4
B2_HD int act_s16p_index(int u) { return (u & ~15) + 8 * ((u >> 3) & 1) + 4 * (u & 1) + ((u >> 1) & 3); }
// What one lane keeps in registers for one K-slice: 64 activations (16+16 words) and their scales.
// Q4_0/Q8_0/Q4_K/Q5_K: the contiguous chunk [64*lane, 64*lane+64): f[0..1] = dx of its two
// 32-blocks, f[2..3] = float sums of the two 32-blocks.
// Q6_K: lane (j, hf, qt) = (lane>>2, (lane>>1)&1, lane&1) owns the four 16-element runs
// 256j + 128hf + 32kq + 16qt (kq = 0..3): a[4kq..4kq+3]; f[kq] = dx, f[4+kq] = sum of the run.
struct LaneAct {
    int a[16];
    int b[16];
    float f[8];
};

// Quantise 32 floats t[0..31] (already multiplied by any norm weight) into words a[8], b[8].
// Returns amax; dx = amax / 127.
B2_D float quantize32(const float* t, int* a, int* b, float& sum_lo, float& sum_hi) {
    float amax = 0.f, s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int i = 0; i < 32; i++) {
        float v = t[i];
        amax = fmaxf(amax, fabsf(v));
        if (i < 16) s0 += v; else s1 += v;
    }
    sum_lo = s0; sum_hi = s1;
    float id = amax > 0.f ? 127.0f / amax : 0.f;
    float dxl = amax / 127.0f;
    float id2 = id * B2_ACT_RES;
    const float MAGIC = 12582912.0f;  // 1.5 * 2^23: float add rounds to nearest integer
#pragma unroll
    for (int w = 0; w < 8; w++) {
        uint32_t aw = 0, bw = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            float v = t[4 * w + j];
            float ra = v * id + MAGIC;
            float af = ra - MAGIC;
            float r = v - af * dxl;
            float rb = r * id2 + MAGIC;
            uint32_t ia, ib;
            memcpy(&ia, &ra, 4); memcpy(&ib, &rb, 4);
            aw |= (ia & 0xffu) << (8 * j);
            bw |= (ib & 0xffu) << (8 * j);
        }
        a[w] = (int)aw; b[w] = (int)bw;
    }
    return amax;
}

// Two activation planes fold into one integer: x.w ~= dx/252 * (252 A + B); LaneAct.f carries dx/252.
template <bool ACT2>
B2_D int fold(int A, int B) { return ACT2 ? A * (int)B2_ACT_RES + B : A; }

// ---------------------------------------------------------------- activation layout
// The int8 planes are stored in 16-byte units, permuted inside each 2048-element slice so that the
// 32 lanes of a warp read 32 consecutive units (conflict-free LDS.128):
//   layout 0 (Q4_0/Q8_0/Q4_K/Q5_K, lane = 64-element chunk): unit u -> (u & 3) * 32 + (u >> 2)
//   layout 1 (Q6_K, lane = (j, hf, qt)):  u = 16j + 8hf + 2kq + qt -> kq * 32 + (4j + 2hf + qt)
B2_HD int act_layout_of(int type) { return type == T_Q6_K ? 1 : 0; }
B2_HD int act_unit_pos(int layout, int u /* 16-byte unit within the slice, 0..127 */) {
    if (layout == 0) return (u & 3) * 32 + (u >> 2);
    int j = u >> 4, hf = (u >> 3) & 1, kq = (u >> 1) & 3, qt = u & 1;
    return kq * 32 + 4 * j + 2 * hf + qt;
}

// ---------------------------------------------------------------- per-lane activation load
template <int T, bool ACT2>
B2_D void load_act(LaneAct& r, const ActView& v, int e0 /* first element of the slice */, int lane) {
    const uint8_t* pa = v.a + e0 + 16 * lane;
    const uint8_t* pb = v.b + e0 + 16 * lane;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        U4 wa = ld128(pa + 512 * i);
        r.a[4 * i] = (int)wa.x; r.a[4 * i + 1] = (int)wa.y; r.a[4 * i + 2] = (int)wa.z; r.a[4 * i + 3] = (int)wa.w;
        if (ACT2) {
            U4 wb = ld128(pb + 512 * i);
            r.b[4 * i] = (int)wb.x; r.b[4 * i + 1] = (int)wb.y; r.b[4 * i + 2] = (int)wb.z; r.b[4 * i + 3] = (int)wb.w;
        }
    }
    if (T == T_Q6_K) {
        const int b0 = (e0 >> 5) + 8 * (lane >> 2) + 4 * ((lane >> 1) & 1);      // first of the lane's four 32-blocks
        const int u0 = (e0 >> 4) + 16 * (lane >> 2) + 8 * ((lane >> 1) & 1) + 4 * (lane & 1);
        U4 d4 = ld128(reinterpret_cast<const uint8_t*>(v.f + b0));
        U4 s4 = ld128(reinterpret_cast<const uint8_t*>(v.g + u0));
        r.f[0] = u2f(d4.x); r.f[1] = u2f(d4.y); r.f[2] = u2f(d4.z); r.f[3] = u2f(d4.w);
        r.f[4] = u2f(s4.x); r.f[5] = u2f(s4.y); r.f[6] = u2f(s4.z); r.f[7] = u2f(s4.w);
    } else {
        U4 f4 = ld128(reinterpret_cast<const uint8_t*>(v.f + 4 * ((e0 >> 6) + lane)));
        r.f[0] = u2f(f4.x); r.f[1] = u2f(f4.y); r.f[2] = u2f(f4.z); r.f[3] = u2f(f4.w);
        if (T == T_Q2_K || T == T_Q3_K) {       // sums of the chunk's four 16-element runs (plain order in g)
            U4 s4 = ld128(reinterpret_cast<const uint8_t*>(v.g + 4 * ((e0 >> 6) + lane)));
            r.f[4] = u2f(s4.x); r.f[5] = u2f(s4.y); r.f[6] = u2f(s4.z); r.f[7] = u2f(s4.w);
        }
    }
}

// ---------------------------------------------------------------- per-lane row dot products
// `row` = first byte of this row's K-slice (shared memory on the GPU). Returns the lane's share of
// the dot product of that slice with the activations in `act`.

template <bool ACT2>
B2_D float row_dot_q4k(const uint8_t* row, int lane, const LaneAct& act) {
    int j = lane >> 2, c = lane & 3;
    const uint8_t* blk = row + j * 144;
    U4 h = ld128(blk);
    U4 q0 = ld128(blk + 16 + 32 * c), q1 = ld128(blk + 32 + 32 * c);
    uint32_t qw[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
    uint32_t sh = (c & 1) * 16;
    uint32_t w0 = h.y >> sh, w1 = h.z >> sh, w2 = h.w >> sh;
    // both layouts of the 6-bit fields, then a select: lanes of one warp differ in c, a branch would diverge
    const uint32_t sc_a = w0 & 0x3F3Fu, m_a = w1 & 0x3F3Fu;
    const uint32_t sc_b = (w2 & 0x0F0Fu) | ((w0 >> 2) & 0x3030u), m_b = ((w2 >> 4) & 0x0F0Fu) | ((w1 >> 2) & 0x3030u);
    const uint32_t scp = c < 2 ? sc_a : sc_b, mp = c < 2 ? m_a : m_b;
    float d = h2f(h.x & 0xffffu), dmin = h2f(h.x >> 16);
    int Alo = 0, Blo = 0, Ahi = 0, Bhi = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        int lo = (int)(qw[k] & 0x0F0F0F0Fu);
        uint32_t hi16 = qw[k] & 0xF0F0F0F0u;          // 16 * high nibble, as unsigned bytes
        Alo = dp4a(lo, act.a[k], Alo);
        Ahi = dp4a_us(hi16, act.a[8 + k], Ahi);
        if (ACT2) { Blo = dp4a(lo, act.b[k], Blo); Bhi = dp4a_us(hi16, act.b[8 + k], Bhi); }
    }
    // integer fold: T = 252 A + B, then the 6-bit sub-block scale, one int->float per sub-block
    int U0 = fold<ACT2>(Alo, Blo) * (int)(scp & 0xffu);
    int U1 = (fold<ACT2>(Ahi, Bhi) >> 4) * (int)(scp >> 8);       // >> 4 exact: hi sums are multiples of 16
    float F = (float)U0 * act.f[0] + (float)U1 * act.f[1];
    float M = (float)(mp & 0xffu) * act.f[2] + (float)(mp >> 8) * act.f[3];
    return d * F - dmin * M;
}

template <bool ACT2>
B2_D float row_dot_q5k(const uint8_t* row, int lane, const LaneAct& act) {
    int j = lane >> 2, c = lane & 3;
    const uint8_t* blk = row + j * 176;
    U4 h = ld128(blk);
    U4 h0 = ld128(blk + 16), h1 = ld128(blk + 32);
    U4 q0 = ld128(blk + 48 + 32 * c), q1 = ld128(blk + 64 + 32 * c);
    uint32_t qw[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
    uint32_t hw[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
    uint32_t sh = (c & 1) * 16;
    uint32_t w0 = h.y >> sh, w1 = h.z >> sh, w2 = h.w >> sh;
    // both layouts of the 6-bit fields, then a select: lanes of one warp differ in c, a branch would diverge
    const uint32_t sc_a = w0 & 0x3F3Fu, m_a = w1 & 0x3F3Fu;
    const uint32_t sc_b = (w2 & 0x0F0Fu) | ((w0 >> 2) & 0x3030u), m_b = ((w2 >> 4) & 0x0F0Fu) | ((w1 >> 2) & 0x3030u);
    const uint32_t scp = c < 2 ? sc_a : sc_b, mp = c < 2 ? m_a : m_b;
    float d = h2f(h.x & 0xffffu), dmin = h2f(h.x >> 16);
    int Alo = 0, Blo = 0, Ahi = 0, Bhi = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        uint32_t hb = hw[k] >> (2 * c);
        int lo = (int)((qw[k] & 0x0F0F0F0Fu) | ((hb & 0x01010101u) << 4));
        int hi = (int)(((qw[k] >> 4) & 0x0F0F0F0Fu) | (((hb >> 1) & 0x01010101u) << 4));
        Alo = dp4a(lo, act.a[k], Alo);
        Ahi = dp4a(hi, act.a[8 + k], Ahi);
        if (ACT2) { Blo = dp4a(lo, act.b[k], Blo); Bhi = dp4a(hi, act.b[8 + k], Bhi); }
    }
    int U0 = fold<ACT2>(Alo, Blo) * (int)(scp & 0xffu);
    int U1 = fold<ACT2>(Ahi, Bhi) * (int)(scp >> 8);
    float F = (float)U0 * act.f[0] + (float)U1 * act.f[1];
    float M = (float)(mp & 0xffu) * act.f[2] + (float)(mp >> 8) * act.f[3];
    return d * F - dmin * M;
}

template <bool ACT2>
B2_D float row_dot_q6k(const uint8_t* row, int lane, const LaneAct& act) {
    int j = lane >> 2, hf = (lane >> 1) & 1, qt = lane & 1;
    const uint8_t* blk = row + j * 210;                // 2-byte aligned only
    uint32_t wlA[4], wlB[4], wh[4], scw[2];
    ld_words_u2<4>(blk + 64 * hf + 16 * qt, wlA);       // ql[l .. l+15],      l = 16 qt
    ld_words_u2<4>(blk + 64 * hf + 32 + 16 * qt, wlB);  // ql[32+l .. 32+l+15]
    ld_words_u2<4>(blk + 128 + 32 * hf + 16 * qt, wh);  // qh[l .. l+15]
    ld_words_u2<2>(blk + 192 + 8 * hf, scw);            // the 8 int8 scales of this half
    float d = h2f(ld16(blk + 208));
    int A[4] = {0, 0, 0, 0}, B[4] = {0, 0, 0, 0};
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int q0 = (int)((wlA[i] & 0x0F0F0F0Fu) + 16u * (wh[i] & 0x03030303u));
        int q1 = (int)((wlB[i] & 0x0F0F0F0Fu) + 16u * ((wh[i] >> 2) & 0x03030303u));
        int q2 = (int)(((wlA[i] >> 4) & 0x0F0F0F0Fu) + 16u * ((wh[i] >> 4) & 0x03030303u));
        int q3 = (int)(((wlB[i] >> 4) & 0x0F0F0F0Fu) + 16u * ((wh[i] >> 6) & 0x03030303u));
        A[0] = dp4a(q0, act.a[i], A[0]); A[1] = dp4a(q1, act.a[4 + i], A[1]);
        A[2] = dp4a(q2, act.a[8 + i], A[2]); A[3] = dp4a(q3, act.a[12 + i], A[3]);
        if (ACT2) {
            B[0] = dp4a(q0, act.b[i], B[0]); B[1] = dp4a(q1, act.b[4 + i], B[1]);
            B[2] = dp4a(q2, act.b[8 + i], B[2]); B[3] = dp4a(q3, act.b[12 + i], B[3]);
        }
    }
    float r = 0.f;
#pragma unroll
    for (int kq = 0; kq < 4; kq++) {
        int si = 2 * kq + qt;                           // scale index within the half
        float sc = (float)(int8_t)((si < 4 ? scw[0] : scw[1]) >> (8 * (si & 3)));
        r += sc * ((float)fold<ACT2>(A[kq], B[kq]) * act.f[kq] - act.f[4 + kq]);
    }
    return d * r;
}

template <bool ACT2>
B2_D float row_dot_q4_0(const uint8_t* row, int lane, const LaneAct& act) {
    const uint8_t* p = row + 36 * lane;   // two 18-byte blocks; 4-byte aligned
    uint32_t W[9];
#pragma unroll
    for (int i = 0; i < 9; i++) W[i] = ld32(p + 4 * i);
    float d0 = h2f(W[0] & 0xffffu), d1 = h2f(W[4] >> 16);
    int A0 = 0, B0 = 0, A1 = 0, B1 = 0, A0h = 0, B0h = 0, A1h = 0, B1h = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        uint32_t q0 = fshr(W[k], W[k + 1], 16), q1 = W[5 + k];
        int lo0 = (int)(q0 & 0x0F0F0F0Fu), lo1 = (int)(q1 & 0x0F0F0F0Fu);
        uint32_t hi0 = q0 & 0xF0F0F0F0u, hi1 = q1 & 0xF0F0F0F0u;
        A0 = dp4a(lo0, act.a[k], A0); A0h = dp4a_us(hi0, act.a[4 + k], A0h);
        A1 = dp4a(lo1, act.a[8 + k], A1); A1h = dp4a_us(hi1, act.a[12 + k], A1h);
        if (ACT2) {
            B0 = dp4a(lo0, act.b[k], B0); B0h = dp4a_us(hi0, act.b[4 + k], B0h);
            B1 = dp4a(lo1, act.b[8 + k], B1); B1h = dp4a_us(hi1, act.b[12 + k], B1h);
        }
    }
    A0 += A0h >> 4; B0 += B0h >> 4; A1 += A1h >> 4; B1 += B1h >> 4;
    return d0 * ((float)fold<ACT2>(A0, B0) * act.f[0] - 8.0f * act.f[2]) + d1 * ((float)fold<ACT2>(A1, B1) * act.f[1] - 8.0f * act.f[3]);
}

template <bool ACT2>
B2_D float row_dot_q8_0(const uint8_t* row, int lane, const LaneAct& act) {
    const uint8_t* p = row + 68 * lane;   // two 34-byte blocks; 4-byte aligned
    uint32_t W[17];
#pragma unroll
    for (int i = 0; i < 17; i++) W[i] = ld32(p + 4 * i);
    float d0 = h2f(W[0] & 0xffffu), d1 = h2f(W[8] >> 16);
    int A0 = 0, B0 = 0, A1 = 0, B1 = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        int q0 = (int)fshr(W[k], W[k + 1], 16), q1 = (int)W[9 + k];
        A0 = dp4a(q0, act.a[k], A0);
        A1 = dp4a(q1, act.a[8 + k], A1);
        if (ACT2) { B0 = dp4a(q0, act.b[k], B0); B1 = dp4a(q1, act.b[8 + k], B1); }
    }
    return d0 * ((float)fold<ACT2>(A0, B0) * act.f[0]) + d1 * ((float)fold<ACT2>(A1, B1) * act.f[1]);
}

// bit i of x (i < 4) -> bit 0 of byte i
B2_D uint32_t spread4(uint32_t x) { return ((x & 0xFu) * 0x00204081u) & 0x01010101u; }

template <bool ACT2>
B2_D float row_dot_q4_1(const uint8_t* row, int lane, const LaneAct& act) {
    const uint8_t* p = row + 40 * lane;   // two 20-byte blocks; 8-byte aligned
    uint32_t W[10];
#pragma unroll
    for (int i = 0; i < 10; i++) W[i] = ld32(p + 4 * i);
    float d0 = h2f(W[0] & 0xffffu), m0 = h2f(W[0] >> 16), d1 = h2f(W[5] & 0xffffu), m1 = h2f(W[5] >> 16);
    int A0 = 0, B0 = 0, A1 = 0, B1 = 0, A0h = 0, B0h = 0, A1h = 0, B1h = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        uint32_t q0 = W[1 + k], q1 = W[6 + k];
        int lo0 = (int)(q0 & 0x0F0F0F0Fu), lo1 = (int)(q1 & 0x0F0F0F0Fu);
        uint32_t hi0 = q0 & 0xF0F0F0F0u, hi1 = q1 & 0xF0F0F0F0u;
        A0 = dp4a(lo0, act.a[k], A0); A0h = dp4a_us(hi0, act.a[4 + k], A0h);
        A1 = dp4a(lo1, act.a[8 + k], A1); A1h = dp4a_us(hi1, act.a[12 + k], A1h);
        if (ACT2) {
            B0 = dp4a(lo0, act.b[k], B0); B0h = dp4a_us(hi0, act.b[4 + k], B0h);
            B1 = dp4a(lo1, act.b[8 + k], B1); B1h = dp4a_us(hi1, act.b[12 + k], B1h);
        }
    }
    A0 += A0h >> 4; B0 += B0h >> 4; A1 += A1h >> 4; B1 += B1h >> 4;
    return d0 * ((float)fold<ACT2>(A0, B0) * act.f[0]) + m0 * act.f[2] + d1 * ((float)fold<ACT2>(A1, B1) * act.f[1]) + m1 * act.f[3];
}

// Q5_0 (OFF = 2: d | qh | qs) and Q5_1 (OFF = 4: d | m | qh | qs): two blocks per lane, 4-byte aligned
template <bool ACT2, bool HAS_MIN>
B2_D float row_dot_q5_01(const uint8_t* row, int lane, const LaneAct& act) {
    constexpr int BB = HAS_MIN ? 24 : 22, NW = 2 * BB / 4;
    const uint8_t* p = row + 2 * BB * lane;
    uint32_t W[NW];
#pragma unroll
    for (int i = 0; i < NW; i++) W[i] = ld32(p + 4 * i);
    uint32_t qh0, qh1, qs0[4], qs1[4];
    float d0, d1, m0 = 0.f, m1 = 0.f;
    if (HAS_MIN) {
        d0 = h2f(W[0] & 0xffffu); m0 = h2f(W[0] >> 16); qh0 = W[1];
        d1 = h2f(W[6] & 0xffffu); m1 = h2f(W[6] >> 16); qh1 = W[7];
#pragma unroll
        for (int k = 0; k < 4; k++) { qs0[k] = W[2 + k]; qs1[k] = W[8 + k]; }
    } else {
        d0 = h2f(W[0] & 0xffffu); qh0 = fshr(W[0], W[1], 16);
        d1 = h2f(W[5] >> 16); qh1 = W[6];
#pragma unroll
        for (int k = 0; k < 4; k++) { qs0[k] = fshr(W[1 + k], W[2 + k], 16); qs1[k] = W[7 + k]; }
    }
    int A0 = 0, B0 = 0, A1 = 0, B1 = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        int lo0 = (int)((qs0[k] & 0x0F0F0F0Fu) | (spread4(qh0 >> (4 * k)) << 4));
        int hi0 = (int)(((qs0[k] >> 4) & 0x0F0F0F0Fu) | (spread4(qh0 >> (16 + 4 * k)) << 4));
        int lo1 = (int)((qs1[k] & 0x0F0F0F0Fu) | (spread4(qh1 >> (4 * k)) << 4));
        int hi1 = (int)(((qs1[k] >> 4) & 0x0F0F0F0Fu) | (spread4(qh1 >> (16 + 4 * k)) << 4));
        A0 = dp4a(lo0, act.a[k], A0); A0 = dp4a(hi0, act.a[4 + k], A0);
        A1 = dp4a(lo1, act.a[8 + k], A1); A1 = dp4a(hi1, act.a[12 + k], A1);
        if (ACT2) {
            B0 = dp4a(lo0, act.b[k], B0); B0 = dp4a(hi0, act.b[4 + k], B0);
            B1 = dp4a(lo1, act.b[8 + k], B1); B1 = dp4a(hi1, act.b[12 + k], B1);
        }
    }
    const float F0 = (float)fold<ACT2>(A0, B0) * act.f[0], F1 = (float)fold<ACT2>(A1, B1) * act.f[1];
    if (HAS_MIN) return d0 * F0 + m0 * act.f[2] + d1 * F1 + m1 * act.f[3];
    return d0 * (F0 - 16.0f * act.f[2]) + d1 * (F1 - 16.0f * act.f[3]);
}

// Q2_K / Q3_K: lane (j, c) = (lane >> 2, lane & 3) owns elements 64c .. 64c+63 of 256-block j = four 16-element sub-blocks;
// the 2-bit planes of 32 bytes of qs: bytes 0-15 / 16-31 at shift 4(c&1) are sub-blocks 0 / 1, at shift 4(c&1)+2 sub-blocks 2 / 3
template <bool ACT2>
B2_D float row_dot_q2k(const uint8_t* row, int lane, const LaneAct& act) {
    const int j = lane >> 2, c = lane & 3;
    const uint8_t* blk = row + j * 84;                  // 4-byte aligned
    const uint32_t scw = ld32(blk + 4 * c), dd = ld32(blk + 80);
    uint32_t W[8];
#pragma unroll
    for (int i = 0; i < 8; i++) W[i] = ld32(blk + 16 + 32 * (c >> 1) + 4 * i);
    const uint32_t s0 = 4 * (c & 1);
    int A[4] = {0, 0, 0, 0}, B[4] = {0, 0, 0, 0};
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int q0 = (int)((W[k] >> s0) & 0x03030303u), q2 = (int)((W[k] >> (s0 + 2)) & 0x03030303u);
        const int q1 = (int)((W[4 + k] >> s0) & 0x03030303u), q3 = (int)((W[4 + k] >> (s0 + 2)) & 0x03030303u);
        A[0] = dp4a(q0, act.a[k], A[0]); A[1] = dp4a(q1, act.a[4 + k], A[1]);
        A[2] = dp4a(q2, act.a[8 + k], A[2]); A[3] = dp4a(q3, act.a[12 + k], A[3]);
        if (ACT2) {
            B[0] = dp4a(q0, act.b[k], B[0]); B[1] = dp4a(q1, act.b[4 + k], B[1]);
            B[2] = dp4a(q2, act.b[8 + k], B[2]); B[3] = dp4a(q3, act.b[12 + k], B[3]);
        }
    }
    int U[4];
    float M = 0.f;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        U[i] = fold<ACT2>(A[i], B[i]) * (int)((scw >> (8 * i)) & 15u);
        M += (float)((scw >> (8 * i + 4)) & 15u) * act.f[4 + i];
    }
    const float F = (float)(U[0] + U[1]) * act.f[0] + (float)(U[2] + U[3]) * act.f[1];
    return h2f(dd & 0xffffu) * F - h2f(dd >> 16) * M;
}

template <bool ACT2>
B2_D float row_dot_q3k(const uint8_t* row, int lane, const LaneAct& act) {
    const int j = lane >> 2, c = lane & 3;
    const uint8_t* blk = row + j * 110;                 // 2-byte aligned only
    uint32_t H[8], W[8], S[3];
    ld_words_u2<8>(blk, H);
    ld_words_u2<8>(blk + 32 + 32 * (c >> 1), W);
    ld_words_u2<3>(blk + 96, S);
    const float d = h2f(ld16(blk + 108));
    // the lane's four 6-bit scales (+32) as the bytes of one word
    const uint32_t scw = ((((c & 1) ? S[1] : S[0]) >> (4 * (c >> 1))) & 0x0F0F0F0Fu) | (((S[2] >> (2 * c)) & 0x03030303u) << 4);
    const uint32_t s0 = 4 * (c & 1), hb = 2 * c;
    int A[4] = {0, 0, 0, 0}, B[4] = {0, 0, 0, 0};
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int q0 = (int)(((W[k] >> s0) & 0x03030303u) | (((H[k] >> hb) & 0x01010101u) << 2));
        const int q2 = (int)(((W[k] >> (s0 + 2)) & 0x03030303u) | (((H[k] >> (hb + 1)) & 0x01010101u) << 2));
        const int q1 = (int)(((W[4 + k] >> s0) & 0x03030303u) | (((H[4 + k] >> hb) & 0x01010101u) << 2));
        const int q3 = (int)(((W[4 + k] >> (s0 + 2)) & 0x03030303u) | (((H[4 + k] >> (hb + 1)) & 0x01010101u) << 2));
        A[0] = dp4a(q0, act.a[k], A[0]); A[1] = dp4a(q1, act.a[4 + k], A[1]);
        A[2] = dp4a(q2, act.a[8 + k], A[2]); A[3] = dp4a(q3, act.a[12 + k], A[3]);
        if (ACT2) {
            B[0] = dp4a(q0, act.b[k], B[0]); B[1] = dp4a(q1, act.b[4 + k], B[1]);
            B[2] = dp4a(q2, act.b[8 + k], B[2]); B[3] = dp4a(q3, act.b[12 + k], B[3]);
        }
    }
    float r = 0.f;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const float sc = (float)((int)((scw >> (8 * i)) & 63u) - 32);
        r += sc * ((float)fold<ACT2>(A[i], B[i]) * act.f[i >> 1] - 4.0f * act.f[4 + i]);     // sum (u - 4) x = sum u x - 4 sum x
    }
    return d * r;
}

template <int T, bool ACT2>
B2_D float row_dot(const uint8_t* row, int lane, const LaneAct& act) {
    if (T == T_Q4_1) return row_dot_q4_1<ACT2>(row, lane, act);
    if (T == T_Q5_0) return row_dot_q5_01<ACT2, false>(row, lane, act);
    if (T == T_Q5_1) return row_dot_q5_01<ACT2, true>(row, lane, act);
    if (T == T_Q2_K) return row_dot_q2k<ACT2>(row, lane, act);
    if (T == T_Q3_K) return row_dot_q3k<ACT2>(row, lane, act);
    if (T == T_Q4_K) return row_dot_q4k<ACT2>(row, lane, act);
    if (T == T_Q6_K) return row_dot_q6k<ACT2>(row, lane, act);
    if (T == T_Q5_K) return row_dot_q5k<ACT2>(row, lane, act);
    if (T == T_Q4_0) return row_dot_q4_0<ACT2>(row, lane, act);
    if (T == T_Q8_0) return row_dot_q8_0<ACT2>(row, lane, act);
    return 0.f;
}

// bytes of one full K-slice (2048 elements) of a row
B2_HD int slice_bytes(int t) { return B2_SLICE / block_elems(t) * block_bytes(t); }

}  // namespace b2
