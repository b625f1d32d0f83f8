/* ORACLE / CPU BASELINE (test infrastructure, not product code).
 *
 * C restatement of the CPU path the reference would run for one decode step with
 * `Llama(..., n_gpu_layers=0)`: llama-cpp-python 0.2.77 -> llama.cpp llm_build_llama on ggml's
 * CPU backend (un-vendored upstream of /root/reference/api.py:24-28,55-63; SURVEY.md 2.2 H6,
 * 8(d) "CPU baseline"). PARITY UNPINNED BY THE REFERENCE (the source is not in this image);
 * the arithmetic below follows ggml's published CPU scheme: activations are quantised to int8
 * (Q8_0 blocks of 32 for Q4_0/Q8_0 weights, Q8_K blocks of 256 with 16-element partial sums
 * for K-quants), dot products are integer with a float fix-up per (sub-)block, everything else
 * is float32, the KV cache is float16. tests/test_cpu_ref.py checks it against the float32
 * oracle (oracle/llama_ref.py) within the int8-activation noise.
 *
 * Only bench.py (cpu_baseline leg and --impl reference) and tests/ load this. Built with
 *   gcc -O3 -march=native -fopenmp -shared -fPIC oracle/cpu_ref.c -o oracle/_build/libcpu_ref.so -lm
 * by oracle/build_cpu_ref.py on the machine that runs it.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#if defined(__AVX512VNNI__) && defined(__AVX512VL__) && defined(__AVX512BW__) && defined(__AVX2__)
#include <immintrin.h>
#define CR_VNNI 1      /* integer dots on vpdpbusd, as ggml's x86 CPU kernels do them (SURVEY.md 8(d) "CPU baseline") */
#else
#define CR_VNNI 0
#endif
static int g_simd = CR_VNNI;   /* cr_set_simd(0) forces the scalar loops (tests compare the two) */
static int g_exact = 0;        /* cr_set_exact(1): float activations against dequantised weights, double accumulation: not what
                                  the reference computes (it quantises activations to int8), but a sharp checker for deep models,
                                  where the int8 noise of 100+ mat-vecs adds up to ~10% of the logit spread */

enum { T_F32 = 0, T_F16 = 1, T_Q4_0 = 2, T_Q4_1 = 3, T_Q5_0 = 6, T_Q5_1 = 7, T_Q8_0 = 8, T_Q2_K = 10, T_Q3_K = 11,
       T_Q4_K = 12, T_Q5_K = 13, T_Q6_K = 14 };

typedef struct {
    const uint8_t* data;
    int32_t type;
    int32_t pad;
    int64_t ne0, ne1;
} cr_tensor;

typedef struct {
    int32_t n_layer, n_embd, n_head, n_head_kv, head_dim, n_ff, n_vocab, n_ctx;
    float rms_eps, rope_base;
    cr_tensor tok_embd, out_norm, output;
    cr_tensor* layers; /* 9 per layer: attn_norm wq wk wv wo ffn_norm wgate wup wdown */
} cr_model;

typedef struct {
    cr_model m;
    uint16_t *k, *v;   /* [layer][pos][n_kv*head_dim] f16 */
    float *x, *h, *q, *kk, *vv, *att, *g, *u, *logits, *freq, *scores;
    int8_t* q8;        /* quantised activations */
    float* q8d;        /* per-block scales */
    int16_t* q8s;      /* Q8_K bsums */
    int n_past;
} cr_ctx;

static inline float h2f(uint16_t h) {
    uint32_t s = (h >> 15) & 1, e = (h >> 10) & 31, m = h & 1023, o;
    if (e == 0) {
        if (m == 0) o = s << 31;
        else { int k = 0; while (!(m & 1024)) { m <<= 1; k++; } m &= 1023; o = (s << 31) | ((uint32_t)(113 - k) << 23) | (m << 13); }
    } else if (e == 31) o = (s << 31) | 0x7f800000u | (m << 13);
    else o = (s << 31) | ((e + 112) << 23) | (m << 13);
    float f; memcpy(&f, &o, 4); return f;
}
static inline uint16_t f2h(float f) {   /* round to nearest even */
    uint32_t x; memcpy(&x, &f, 4);
    uint32_t s = (x >> 16) & 0x8000; int32_t e = ((x >> 23) & 255) - 127 + 15; uint32_t m = x & 0x7fffff;
    if (((x >> 23) & 255) == 255) return (uint16_t)(s | 0x7c00 | (m ? 0x200 : 0));
    if (e >= 31) return (uint16_t)(s | 0x7c00);
    if (e <= 0) {
        if (e < -10) return (uint16_t)s;
        m |= 0x800000; int sh = 14 - e;
        uint32_t r = m >> sh, rem = m & ((1u << sh) - 1), half = 1u << (sh - 1);
        if (rem > half || (rem == half && (r & 1))) r++;
        return (uint16_t)(s | r);
    }
    uint32_t r = (uint32_t)(e << 10) | (m >> 13), rem = m & 0x1fff;
    if (rem > 0x1000 || (rem == 0x1000 && (r & 1))) r++;
    return (uint16_t)(s | r);
}

static void k_scale_min(const uint8_t* s, int is, int* sc, int* m) {
    if (is < 4) { *sc = s[is] & 63; *m = s[is + 4] & 63; }
    else { *sc = (s[is + 4] & 15) | ((s[is - 4] >> 6) << 4); *m = (s[is + 4] >> 4) | ((s[is] >> 6) << 4); }
}

static int blk_elems(int t) { return (t == T_F32 || t == T_F16) ? 1 : (t == T_Q4_0 || t == T_Q8_0 || t == T_Q4_1 || t == T_Q5_0 || t == T_Q5_1) ? 32 : 256; }
static int blk_bytes(int t) {
    switch (t) { case T_F32: return 4; case T_F16: return 2; case T_Q4_0: return 18; case T_Q8_0: return 34;
                 case T_Q4_K: return 144; case T_Q5_K: return 176; case T_Q6_K: return 210;
                 case T_Q4_1: return 20; case T_Q5_0: return 22; case T_Q5_1: return 24; case T_Q2_K: return 84; case T_Q3_K: return 110; }
    return 0;
}

/* dequantise one block (block_elems(type) values) */
static void dequant_block(int type, const uint8_t* p, float* o) {
    switch (type) {
        case T_F32: memcpy(o, p, 4); break;
        case T_F16: { uint16_t hh; memcpy(&hh, p, 2); o[0] = h2f(hh); } break;
        case T_Q4_0: { uint16_t dh; memcpy(&dh, p, 2); float d = h2f(dh);
            for (int j = 0; j < 16; j++) { o[j] = d * ((p[2 + j] & 15) - 8); o[16 + j] = d * ((p[2 + j] >> 4) - 8); } } break;
        case T_Q8_0: { uint16_t dh; memcpy(&dh, p, 2); float d = h2f(dh);
            for (int j = 0; j < 32; j++) o[j] = d * (int8_t)p[2 + j]; } break;
        case T_Q4_K: case T_Q5_K: {
            uint16_t dh, mh; memcpy(&dh, p, 2); memcpy(&mh, p + 2, 2); float d = h2f(dh), dm = h2f(mh);
            const uint8_t* qs = p + (type == T_Q4_K ? 16 : 48);
            for (int is = 0; is < 8; is++) { int sc, m; k_scale_min(p + 4, is, &sc, &m);
                for (int l = 0; l < 32; l++) { int by = qs[(is >> 1) * 32 + l]; int q = (is & 1) ? by >> 4 : by & 15;
                    if (type == T_Q5_K) q |= ((p[16 + l] >> is) & 1) << 4;
                    o[is * 32 + l] = d * sc * q - dm * m; } } } break;
        /* the block formats below follow gguf-py's dequantisers (Q4_1, Q5_0, Q5_1, Q2_K, Q3_K in gguf/quants.py) */
        case T_Q4_1: { uint16_t dh, mh; memcpy(&dh, p, 2); memcpy(&mh, p + 2, 2); float d = h2f(dh), m = h2f(mh);
            for (int j = 0; j < 16; j++) { o[j] = d * (p[4 + j] & 15) + m; o[16 + j] = d * (p[4 + j] >> 4) + m; } } break;
        case T_Q5_0: case T_Q5_1: { const int of = type == T_Q5_0 ? 2 : 4; uint16_t dh, mh = 0; memcpy(&dh, p, 2);
            if (type == T_Q5_1) memcpy(&mh, p + 2, 2);
            float d = h2f(dh), m = type == T_Q5_1 ? h2f(mh) : 0.f; uint32_t qh; memcpy(&qh, p + of, 4);
            for (int e = 0; e < 32; e++) { int q = e < 16 ? (p[of + 4 + e] & 15) : (p[of + 4 + e - 16] >> 4); q |= (int)((qh >> e) & 1u) << 4;
                o[e] = type == T_Q5_0 ? d * (q - 16) : d * q + m; } } break;
        case T_Q2_K: { uint16_t dh, mh; memcpy(&dh, p + 80, 2); memcpy(&mh, p + 82, 2); float d = h2f(dh), dm = h2f(mh);
            for (int e = 0; e < 256; e++) { int sc = p[e >> 4]; int q = (p[16 + 32 * (e >> 7) + (e & 31)] >> (2 * ((e >> 5) & 3))) & 3;
                o[e] = d * (sc & 15) * q - dm * (sc >> 4); } } break;
        case T_Q3_K: { uint16_t dh; memcpy(&dh, p + 108, 2); float d = h2f(dh);
            for (int e = 0; e < 256; e++) { int is = e >> 4;
                int sc = (((p[96 + (is & 7)] >> (4 * (is >> 3))) & 15) | (((p[104 + (is & 3)] >> (2 * (is >> 2))) & 3) << 4)) - 32;
                int ql = (p[32 + 32 * (e >> 7) + (e & 31)] >> (2 * ((e >> 5) & 3))) & 3; int hb = (p[e & 31] >> (e >> 5)) & 1;
                o[e] = d * sc * (ql - (hb ? 0 : 4)); } } break;
        case T_Q6_K: { uint16_t dh; memcpy(&dh, p + 208, 2); float d = h2f(dh);
            for (int e = 0; e < 256; e++) { int hf = e >> 7, ep = e & 127, nib = ep >> 6, kq = ep >> 5, l = ep & 31;
                int lo = p[64 * hf + (ep & 63)]; lo = nib ? lo >> 4 : lo & 15;
                int hi = (p[128 + 32 * hf + l] >> (2 * kq)) & 3;
                o[e] = d * (int8_t)p[192 + (e >> 4)] * ((lo | (hi << 4)) - 32); } } break;
    }
}

/* dequantise one row (embedding lookup) */
static void dequant_row(const cr_tensor* t, int64_t row, float* out) {
    int K = (int)t->ne0, be = blk_elems(t->type), bb = blk_bytes(t->type);
    const uint8_t* r = t->data + row * (int64_t)(K / be * bb);
    for (int b = 0; b < K / be; b++) dequant_block(t->type, r + (size_t)b * bb, out + (size_t)b * be);
}

/* ggml quantize_row_q8_0: blocks of 32, d = amax/127 */
static void quant_q8_0(const float* x, int K, int8_t* q, float* d) {
    for (int b = 0; b < K / 32; b++) {
        float amax = 0; for (int i = 0; i < 32; i++) { float a = fabsf(x[b * 32 + i]); if (a > amax) amax = a; }
        float dd = amax / 127.0f, id = dd ? 1.0f / dd : 0.0f;
        d[b] = h2f(f2h(dd));   /* ggml stores d as f16 */
        for (int i = 0; i < 32; i++) q[b * 32 + i] = (int8_t)roundf(x[b * 32 + i] * id);
    }
}
/* ggml quantize_row_q8_K: blocks of 256, iscale = -128/max, bsums per 16 */
static void quant_q8_K(const float* x, int K, int8_t* q, float* d, int16_t* bs) {
    for (int b = 0; b < K / 256; b++) {
        float max = 0, amax = 0;
        for (int i = 0; i < 256; i++) { float v = x[b * 256 + i], a = fabsf(v); if (a > amax) { amax = a; max = v; } }
        if (!amax) { d[b] = 0; memset(q + b * 256, 0, 256); memset(bs + b * 16, 0, 32); continue; }
        float iscale = -128.f / max;
        for (int i = 0; i < 256; i++) { int v = (int)lrintf(iscale * x[b * 256 + i]); q[b * 256 + i] = (int8_t)(v > 127 ? 127 : v); }
        for (int j = 0; j < 16; j++) { int s = 0; for (int i = 0; i < 16; i++) s += q[b * 256 + j * 16 + i]; bs[b * 16 + j] = (int16_t)s; }
        d[b] = 1.0f / iscale;
    }
}

#if CR_VNNI
/* Integer block dots on AVX-512 VNNI (vpdpbusd: 4 x u8*s8 -> s32 per lane). Same integers as the scalar loops below. */
static inline int hsum256(__m256i v) {
    __m128i s = _mm_add_epi32(_mm256_castsi256_si128(v), _mm256_extracti128_si256(v, 1));
    s = _mm_add_epi32(s, _mm_shuffle_epi32(s, 0x4e));
    s = _mm_add_epi32(s, _mm_shuffle_epi32(s, 0xb1));
    return _mm_cvtsi128_si32(s);
}
/* Q4_0 block (16 nibble bytes) x 32 int8: sum (q - 8) * a */
static inline int dot_q4_0_vnni(const uint8_t* qs, const int8_t* a) {
    const __m128i b = _mm_loadu_si128((const __m128i*)qs), m4 = _mm_set1_epi8(15);
    const __m256i q = _mm256_set_m128i(_mm_and_si128(_mm_srli_epi16(b, 4), m4), _mm_and_si128(b, m4));   /* [0..15 | 16..31] */
    const __m256i av = _mm256_loadu_si256((const __m256i*)a);
    const __m256i qa = _mm256_dpbusd_epi32(_mm256_setzero_si256(), q, av);
    const __m256i sa = _mm256_dpbusd_epi32(_mm256_setzero_si256(), _mm256_set1_epi8(8), av);              /* 8 * sum a */
    return hsum256(_mm256_sub_epi32(qa, sa));
}
/* Q8_0 block (32 int8) x 32 int8 */
static inline int dot_q8_0_vnni(const int8_t* w, const int8_t* a) {
    const __m256i wv = _mm256_loadu_si256((const __m256i*)w), av = _mm256_loadu_si256((const __m256i*)a);
    return hsum256(_mm256_dpbusd_epi32(_mm256_setzero_si256(), _mm256_sign_epi8(wv, wv), _mm256_sign_epi8(av, wv)));
}
/* Q4_K super-block: sum_is sc[is] * sum_l q[is][l] * a[is*32+l]  (qs: 128 nibble bytes, sc: the 8 unpacked 6-bit scales) */
static inline int dot_q4_K_vnni(const uint8_t* qs, const int8_t* a, const int* sc) {
    const __m256i m4 = _mm256_set1_epi8(15);
    __m512i acc = _mm512_setzero_si512();
    for (int c = 0; c < 4; c++) {
        const __m256i b = _mm256_loadu_si256((const __m256i*)(qs + 32 * c));
        const __m512i q = _mm512_inserti64x4(_mm512_castsi256_si512(_mm256_and_si256(b, m4)),
                                             _mm256_and_si256(_mm256_srli_epi16(b, 4), m4), 1);     /* sub-block 2c | sub-block 2c+1 */
        const __m512i p = _mm512_dpbusd_epi32(_mm512_setzero_si512(), q, _mm512_loadu_si512((const void*)(a + 64 * c)));
        const __m512i scv = _mm512_inserti64x4(_mm512_castsi256_si512(_mm256_set1_epi32(sc[2 * c])), _mm256_set1_epi32(sc[2 * c + 1]), 1);
        acc = _mm512_add_epi32(acc, _mm512_mullo_epi32(p, scv));
    }
    return _mm512_reduce_add_epi32(acc);
}
/* Q6_K super-block: sum over the 16 groups of 16: sc[g] * sum_l (q - 32) * a */
static inline int dot_q6_K_vnni(const uint8_t* p, const int8_t* a) {
    const __m256i m4 = _mm256_set1_epi8(15), m2 = _mm256_set1_epi8(0x30);
    const __m512i expand = _mm512_set_epi32(3, 3, 3, 3, 2, 2, 2, 2, 1, 1, 1, 1, 0, 0, 0, 0);
    const __m512i sc_all = _mm512_cvtepi8_epi32(_mm_loadu_si128((const __m128i*)(p + 192)));        /* 16 int8 scales */
    __m512i acc = _mm512_setzero_si512();
    for (int hf = 0; hf < 2; hf++) {
        const __m256i l0 = _mm256_loadu_si256((const __m256i*)(p + 64 * hf)), l1 = _mm256_loadu_si256((const __m256i*)(p + 64 * hf + 32));
        const __m256i qh = _mm256_loadu_si256((const __m256i*)(p + 128 + 32 * hf));
        const __m256i q1 = _mm256_or_si256(_mm256_and_si256(l0, m4), _mm256_and_si256(_mm256_slli_epi16(qh, 4), m2));
        const __m256i q2 = _mm256_or_si256(_mm256_and_si256(l1, m4), _mm256_and_si256(_mm256_slli_epi16(qh, 2), m2));
        const __m256i q3 = _mm256_or_si256(_mm256_and_si256(_mm256_srli_epi16(l0, 4), m4), _mm256_and_si256(qh, m2));
        const __m256i q4 = _mm256_or_si256(_mm256_and_si256(_mm256_srli_epi16(l1, 4), m4), _mm256_and_si256(_mm256_srli_epi16(qh, 2), m2));
        for (int k = 0; k < 2; k++) {          /* elements 64k .. 64k+63 of this half: groups 4k .. 4k+3 */
            const __m512i q = _mm512_inserti64x4(_mm512_castsi256_si512(k ? q3 : q1), k ? q4 : q2, 1);
            const __m512i av = _mm512_loadu_si512((const void*)(a + 128 * hf + 64 * k));
            __m512i pr = _mm512_dpbusd_epi32(_mm512_setzero_si512(), q, av);
            pr = _mm512_sub_epi32(pr, _mm512_dpbusd_epi32(_mm512_setzero_si512(), _mm512_set1_epi8(32), av));   /* - 32 * a */
            const __m512i idx = _mm512_add_epi32(expand, _mm512_set1_epi32(8 * hf + 4 * k));
            acc = _mm512_add_epi32(acc, _mm512_mullo_epi32(pr, _mm512_permutexvar_epi32(idx, sc_all)));
        }
    }
    return _mm512_reduce_add_epi32(acc);
}
#endif

/* y[n] = W[n,:] . x  with x pre-quantised; rows split across OpenMP threads */
static void matvec(const cr_tensor* w, const int8_t* q8, const float* q8d, const int16_t* q8s, const float* xf, float* y) {
    const int K = (int)w->ne0, N = (int)w->ne1, t = w->type;
    const int64_t rb = (int64_t)K / blk_elems(t) * blk_bytes(t);
    if (g_exact) {
        const int be = blk_elems(t), bb = blk_bytes(t);
#pragma omp parallel for schedule(static)
        for (int n = 0; n < N; n++) {
            const uint8_t* r = w->data + n * rb; double acc = 0; float tmp[256];
            for (int b = 0; b < K / be; b++) { dequant_block(t, r + (size_t)b * bb, tmp);
                double s = 0; for (int i = 0; i < be; i++) s += (double)tmp[i] * xf[b * be + i];
                acc += s; }
            y[n] = (float)acc;
        }
        return;
    }
#pragma omp parallel for schedule(static)
    for (int n = 0; n < N; n++) {
        const uint8_t* r = w->data + n * rb; float acc = 0.f;
        if (t == T_Q4_0) {
            for (int b = 0; b < K / 32; b++) { const uint8_t* p = r + b * 18; uint16_t dh; memcpy(&dh, p, 2);
                const int8_t* a = q8 + b * 32; int s = 0;
#if CR_VNNI
                if (g_simd) s = dot_q4_0_vnni(p + 2, a); else
#endif
                for (int j = 0; j < 16; j++) s += ((p[2 + j] & 15) - 8) * a[j] + ((p[2 + j] >> 4) - 8) * a[16 + j];
                acc += h2f(dh) * q8d[b] * s; }
        } else if (t == T_Q8_0) {
            for (int b = 0; b < K / 32; b++) { const uint8_t* p = r + b * 34; uint16_t dh; memcpy(&dh, p, 2);
                const int8_t* a = q8 + b * 32; const int8_t* wq = (const int8_t*)(p + 2); int s = 0;
#if CR_VNNI
                if (g_simd) s = dot_q8_0_vnni(wq, a); else
#endif
                for (int j = 0; j < 32; j++) s += wq[j] * a[j];
                acc += h2f(dh) * q8d[b] * s; }
        } else if (t == T_Q4_K || t == T_Q5_K) {
            const int bb = blk_bytes(t);
            for (int b = 0; b < K / 256; b++) { const uint8_t* p = r + b * bb; uint16_t dh, mh; memcpy(&dh, p, 2); memcpy(&mh, p + 2, 2);
                const uint8_t* qs = p + (t == T_Q4_K ? 16 : 48); const int8_t* a = q8 + b * 256; const int16_t* bs = q8s + b * 16;
                int sumi = 0, summ = 0;
#if CR_VNNI
                if (g_simd && t == T_Q4_K) {
                    int scs[8];
                    for (int is = 0; is < 8; is++) { int m; k_scale_min(p + 4, is, &scs[is], &m); summ += m * (bs[2 * is] + bs[2 * is + 1]); }
                    sumi = dot_q4_K_vnni(qs, a, scs);
                } else
#endif
                for (int is = 0; is < 8; is++) { int sc, m; k_scale_min(p + 4, is, &sc, &m);
                    const uint8_t* qq = qs + (is >> 1) * 32; int s = 0;
                    if (t == T_Q4_K) { if (is & 1) for (int l = 0; l < 32; l++) s += (qq[l] >> 4) * a[is * 32 + l];
                                       else for (int l = 0; l < 32; l++) s += (qq[l] & 15) * a[is * 32 + l]; }
                    else for (int l = 0; l < 32; l++) { int q = (is & 1) ? qq[l] >> 4 : qq[l] & 15; q |= ((p[16 + l] >> is) & 1) << 4; s += q * a[is * 32 + l]; }
                    sumi += sc * s; summ += m * (bs[2 * is] + bs[2 * is + 1]); }
                acc += h2f(dh) * q8d[b] * sumi - h2f(mh) * q8d[b] * summ; }
        } else if (t == T_Q6_K) {
            for (int b = 0; b < K / 256; b++) { const uint8_t* p = r + b * 210; uint16_t dh; memcpy(&dh, p + 208, 2);
                const int8_t* a = q8 + b * 256; const int8_t* sc = (const int8_t*)(p + 192); int sumi = 0;
#if CR_VNNI
                if (g_simd) sumi = dot_q6_K_vnni(p, a); else
#endif
                for (int hf = 0; hf < 2; hf++) { const uint8_t* ql = p + 64 * hf; const uint8_t* qh = p + 128 + 32 * hf; const int8_t* aa = a + 128 * hf;
                    int s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                    for (int l = 0; l < 32; l++) { int g = l >> 4;
                        s[g]     += (((ql[l] & 15) | (((qh[l] >> 0) & 3) << 4)) - 32) * aa[l];
                        s[2 + g] += (((ql[l + 32] & 15) | (((qh[l] >> 2) & 3) << 4)) - 32) * aa[l + 32];
                        s[4 + g] += (((ql[l] >> 4) | (((qh[l] >> 4) & 3) << 4)) - 32) * aa[l + 64];
                        s[6 + g] += (((ql[l + 32] >> 4) | (((qh[l] >> 6) & 3) << 4)) - 32) * aa[l + 96]; }
                    for (int g = 0; g < 8; g++) sumi += sc[8 * hf + g] * s[g]; }
                acc += h2f(dh) * q8d[b] * sumi; }
        } else {        /* Q4_1, Q5_0, Q5_1, Q2_K, Q3_K: no integer kernel here, float dot on the dequantised block */
            const int be = blk_elems(t), bb = blk_bytes(t); float tmp[256];
            for (int b = 0; b < K / be; b++) { dequant_block(t, r + (size_t)b * bb, tmp);
                float s = 0.f; for (int i = 0; i < be; i++) s += tmp[i] * xf[b * be + i];
                acc += s; }
        }
        y[n] = acc;
    }
}

static void quantize_for(const cr_tensor* w, const float* x, cr_ctx* c) {
    int K = (int)w->ne0;
    if (w->type == T_Q4_0 || w->type == T_Q8_0) quant_q8_0(x, K, c->q8, c->q8d);
    else quant_q8_K(x, K, c->q8, c->q8d, c->q8s);
}

static void rmsnorm(const float* x, const cr_tensor* w, float eps, int K, float* y) {
    double ss = 0; for (int i = 0; i < K; i++) ss += (double)x[i] * x[i];
    float rs = 1.0f / sqrtf((float)(ss / K) + eps); const float* ww = (const float*)w->data;
    for (int i = 0; i < K; i++) y[i] = x[i] * rs * ww[i];
}

static void rope(float* v, int n_heads, int hd, int pos, const float* freq) {
    for (int h = 0; h < n_heads; h++) for (int i = 0; i < hd / 2; i++) {
        float ang = (float)pos * freq[i], c = cosf(ang), s = sinf(ang);
        float a = v[h * hd + 2 * i], b = v[h * hd + 2 * i + 1];
        v[h * hd + 2 * i] = a * c - b * s; v[h * hd + 2 * i + 1] = a * s + b * c; }
}

cr_ctx* cr_create(const cr_model* m) {
    cr_ctx* c = (cr_ctx*)calloc(1, sizeof(cr_ctx));
    c->m = *m;
    size_t L = m->n_layer;
    c->m.layers = (cr_tensor*)malloc(sizeof(cr_tensor) * 9 * L);
    memcpy(c->m.layers, m->layers, sizeof(cr_tensor) * 9 * L);
    int kvd = m->n_head_kv * m->head_dim, qd = m->n_head * m->head_dim;
    int maxk = m->n_ff > m->n_embd ? m->n_ff : m->n_embd; if (qd > maxk) maxk = qd;
    c->k = (uint16_t*)calloc((size_t)L * m->n_ctx * kvd, 2); c->v = (uint16_t*)calloc((size_t)L * m->n_ctx * kvd, 2);
    c->x = malloc(4 * (size_t)m->n_embd); c->h = malloc(4 * (size_t)maxk); c->q = malloc(4 * (size_t)qd);
    c->kk = malloc(4 * (size_t)kvd); c->vv = malloc(4 * (size_t)kvd); c->att = malloc(4 * (size_t)qd);
    c->g = malloc(4 * (size_t)m->n_ff); c->u = malloc(4 * (size_t)m->n_ff); c->logits = malloc(4 * (size_t)m->n_vocab);
    c->scores = malloc(4 * (size_t)m->n_ctx * m->n_head);
    c->q8 = malloc((size_t)maxk); c->q8d = malloc(4 * (size_t)(maxk / 32 + 1)); c->q8s = malloc(2 * (size_t)(maxk / 16 + 1));
    c->freq = malloc(4 * (size_t)m->head_dim / 2);
    float scale = powf(m->rope_base, -2.0f / (float)m->head_dim), f = 1.0f;
    for (int i = 0; i < m->head_dim / 2; i++) { c->freq[i] = f; f *= scale; }
    return c;
}
void cr_destroy(cr_ctx* c) {
    if (!c) return;
    free(c->m.layers); free(c->k); free(c->v); free(c->x); free(c->h); free(c->q); free(c->kk); free(c->vv); free(c->att);
    free(c->g); free(c->u); free(c->logits); free(c->scores); free(c->q8); free(c->q8d); free(c->q8s); free(c->freq); free(c);
}
void cr_reset(cr_ctx* c) { c->n_past = 0; }
/* OpenMP thread count for the mat-vecs and attention (torchrun exports OMP_NUM_THREADS=1, and the variable is only read
 * when the OpenMP runtime starts: the bench sets the count through this call instead) */
void cr_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
int cr_simd(void) { return g_simd; }          /* 1: AVX-512 VNNI integer dots, 0: scalar loops */
int cr_has_vnni(void) { return CR_VNNI; }
void cr_set_exact(int on) { g_exact = on; }   /* 1: float activations x dequantised weights (checker mode, not the timed baseline) */
void cr_set_simd(int on) { g_simd = on && CR_VNNI; }
int cr_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* forward one token at position n_past; logits (if want_logits) in the ctx buffer */
const float* cr_eval(cr_ctx* c, int token, int want_logits) {
    const cr_model* m = &c->m;
    const int d = m->n_embd, hd = m->head_dim, kvd = m->n_head_kv * hd, qd = m->n_head * hd, G = m->n_head / m->n_head_kv;
    const int pos = c->n_past;
    if (pos >= m->n_ctx || token < 0 || token >= m->n_vocab) return NULL;
    dequant_row(&m->tok_embd, token, c->x);
    for (int l = 0; l < m->n_layer; l++) {
        const cr_tensor* T = m->layers + 9 * (size_t)l;
        rmsnorm(c->x, &T[0], m->rms_eps, d, c->h);
        quantize_for(&T[1], c->h, c); matvec(&T[1], c->q8, c->q8d, c->q8s, c->h, c->q);
        if (T[2].type != T[1].type) quantize_for(&T[2], c->h, c);
        matvec(&T[2], c->q8, c->q8d, c->q8s, c->h, c->kk);
        if (T[3].type != T[2].type) quantize_for(&T[3], c->h, c);
        matvec(&T[3], c->q8, c->q8d, c->q8s, c->h, c->vv);
        rope(c->q, m->n_head, hd, pos, c->freq); rope(c->kk, m->n_head_kv, hd, pos, c->freq);
        uint16_t* kc = c->k + ((size_t)l * m->n_ctx + pos) * kvd; uint16_t* vc = c->v + ((size_t)l * m->n_ctx + pos) * kvd;
        for (int i = 0; i < kvd; i++) { kc[i] = f2h(c->kk[i]); vc[i] = f2h(c->vv[i]); }
        const float sc = 1.0f / sqrtf((float)hd);
#pragma omp parallel for schedule(static)
        for (int h = 0; h < m->n_head; h++) {
            const int kh = h / G; float* s = c->scores + (size_t)h * m->n_ctx; const float* qh = c->q + h * hd; float mx = -INFINITY;
            for (int t = 0; t <= pos; t++) { const uint16_t* kr = c->k + ((size_t)l * m->n_ctx + t) * kvd + kh * hd; float a = 0;
                for (int i = 0; i < hd; i++) a += qh[i] * h2f(kr[i]);
                s[t] = a * sc; if (s[t] > mx) mx = s[t]; }
            float sum = 0; for (int t = 0; t <= pos; t++) { s[t] = expf(s[t] - mx); sum += s[t]; }
            float* o = c->att + h * hd; for (int i = 0; i < hd; i++) o[i] = 0;
            for (int t = 0; t <= pos; t++) { const uint16_t* vr = c->v + ((size_t)l * m->n_ctx + t) * kvd + kh * hd; float p = s[t] / sum;
                for (int i = 0; i < hd; i++) o[i] += p * h2f(vr[i]); }
        }
        quantize_for(&T[4], c->att, c); matvec(&T[4], c->q8, c->q8d, c->q8s, c->att, c->h);
        for (int i = 0; i < d; i++) c->x[i] += c->h[i];
        rmsnorm(c->x, &T[5], m->rms_eps, d, c->h);
        quantize_for(&T[6], c->h, c); matvec(&T[6], c->q8, c->q8d, c->q8s, c->h, c->g);
        if (T[7].type != T[6].type) quantize_for(&T[7], c->h, c);
        matvec(&T[7], c->q8, c->q8d, c->q8s, c->h, c->u);
        for (int i = 0; i < m->n_ff; i++) c->g[i] = c->g[i] / (1.0f + expf(-c->g[i])) * c->u[i];
        quantize_for(&T[8], c->g, c); matvec(&T[8], c->q8, c->q8d, c->q8s, c->g, c->h);
        for (int i = 0; i < d; i++) c->x[i] += c->h[i];
    }
    c->n_past++;
    if (!want_logits) return c->logits;
    rmsnorm(c->x, &m->out_norm, m->rms_eps, d, c->h);
    quantize_for(&m->output, c->h, c); matvec(&m->output, c->q8, c->q8d, c->q8s, c->h, c->logits);
    return c->logits;
}
