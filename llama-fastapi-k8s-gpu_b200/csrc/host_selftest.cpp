// Host build of the block arithmetic in quant_blocks.cuh (same source, B2_HOST_EMU), exported
// for tests/test_host_selftest.py: lets the CPU test-suite check the nibble/bit-plane
// unpacking and the two-plane int8 activation scheme against the oracle without a GPU.
#define B2_HOST_EMU 1
#include <math.h>
#include "quant_blocks.cuh"

template <int T>
static float host_row(const uint8_t* rowp, int K, const b2::ActView& v, bool act2) {
    // what one warp does for one row: loop over K-slices, 32 lanes each
    const int sb = b2::slice_bytes(T);
    float acc = 0.f;
    for (int s = 0; s * B2_SLICE < K; s++) {
        int se = K - s * B2_SLICE < B2_SLICE ? K - s * B2_SLICE : B2_SLICE;
        for (int lane = 0; lane < 32; lane++) {
            if (lane * 64 >= se) continue;
            b2::LaneAct la;
            if (act2) { b2::load_act<T, true>(la, v, s * B2_SLICE, lane); acc += b2::row_dot<T, true>(rowp + (size_t)s * sb, lane, la); }
            else { b2::load_act<T, false>(la, v, s * B2_SLICE, lane); acc += b2::row_dot<T, false>(rowp + (size_t)s * sb, lane, la); }
        }
    }
    return acc;
}

extern "C" {

// dequantise n_elems elements (whole blocks) with the exact element routine
int b2_host_dequant(int type, const uint8_t* raw, long n_elems, float* out) {
    int be = b2::block_elems(type), bb = b2::block_bytes(type);
    if (!bb || n_elems % be) return -1;
    for (long i = 0; i < n_elems; i++) out[i] = b2::dequant_elem(type, raw + (i / be) * bb, (int)(i % be));
    return 0;
}

// y[r] = W[r,:] . x through quantize32 + load_act + row_dot (what the GPU lanes do)
int b2_host_matvec(int type, const uint8_t* w, int n_rows, int K, const float* x, float* y, int act2,
                   int misalign) {
    if (K % 256) return -1;
    long row_bytes = (long)K / b2::block_elems(type) * b2::block_bytes(type);
    uint8_t* buf = new uint8_t[row_bytes + 64];
    memset(buf, 0xA5, row_bytes + 64);
    // activation planes (16-byte aligned like the shared-memory arrays)
    const int Kp = (K + B2_SLICE - 1) / B2_SLICE * B2_SLICE;   // planes padded to whole slices
    uint8_t* abuf = new uint8_t[2 * Kp + 64];
    memset(abuf, 0, 2 * Kp + 64);
    uint8_t* aa = (uint8_t*)(((uintptr_t)abuf + 15) & ~(uintptr_t)15);
    uint8_t* ab = aa + Kp;
    // packed per-block scales exactly as the mat-vec prologue stores them (quant_blocks.cuh ActView)
    float* fbuf = new float[K / 16 + K / 16 + 16];
    float* f = (float*)(((uintptr_t)fbuf + 15) & ~(uintptr_t)15);
    float* g = f + K / 16;
    const float fs = act2 ? 1.0f / B2_ACT_RES : 1.0f;
    const int layout = b2::act_layout_of(type);
    for (int b = 0; b < K / 32; b++) {
        int qa[8], qb[8]; float lo, hi;
        float amax = b2::quantize32(x + 32 * b, qa, qb, lo, hi);
        for (int h2 = 0; h2 < 2; h2++) {       // two 16-byte units per 32-block, permuted inside the slice
            int u = 2 * b + h2, sl = u / 128, pos = b2::act_unit_pos(layout, u % 128);
            memcpy(aa + sl * B2_SLICE + 16 * pos, qa + 4 * h2, 16);
            memcpy(ab + sl * B2_SLICE + 16 * pos, qb + 4 * h2, 16);
        }
        if (layout == 0) {
            f[4 * (b >> 1) + (b & 1)] = amax / 127.0f * fs;
            f[4 * (b >> 1) + 2 + (b & 1)] = lo + hi;
            if (type == b2::T_Q2_K || type == b2::T_Q3_K) { g[2 * b] = lo; g[2 * b + 1] = hi; }   // 16-element sums, plain order
        } else {
            f[b] = amax / 127.0f * fs;
            g[b2::act_s16p_index(2 * b)] = 32.0f * lo;
            g[b2::act_s16p_index(2 * b + 1)] = 32.0f * hi;
        }
    }
    b2::ActView v; v.a = aa; v.b = ab; v.f = f; v.g = g;
    uintptr_t base = ((uintptr_t)buf + 15) & ~(uintptr_t)15;
    uint8_t* rowp = (uint8_t*)base + misalign;
    int rc = 0;
    for (int r = 0; r < n_rows; r++) {
        memcpy(rowp, w + r * row_bytes, row_bytes);
        switch (type) {
            case b2::T_Q4_0: y[r] = host_row<b2::T_Q4_0>(rowp, K, v, act2); break;
            case b2::T_Q8_0: y[r] = host_row<b2::T_Q8_0>(rowp, K, v, act2); break;
            case b2::T_Q4_K: y[r] = host_row<b2::T_Q4_K>(rowp, K, v, act2); break;
            case b2::T_Q5_K: y[r] = host_row<b2::T_Q5_K>(rowp, K, v, act2); break;
            case b2::T_Q6_K: y[r] = host_row<b2::T_Q6_K>(rowp, K, v, act2); break;
            case b2::T_Q4_1: y[r] = host_row<b2::T_Q4_1>(rowp, K, v, act2); break;
            case b2::T_Q5_0: y[r] = host_row<b2::T_Q5_0>(rowp, K, v, act2); break;
            case b2::T_Q5_1: y[r] = host_row<b2::T_Q5_1>(rowp, K, v, act2); break;
            case b2::T_Q2_K: y[r] = host_row<b2::T_Q2_K>(rowp, K, v, act2); break;
            case b2::T_Q3_K: y[r] = host_row<b2::T_Q3_K>(rowp, K, v, act2); break;
            default: rc = -2;
        }
    }
    delete[] buf; delete[] abuf; delete[] fbuf;
    return rc;
}

// reconstruct the quantised activations (for checking the ~16-bit claim)
int b2_host_act_roundtrip(const float* x, int n, float* xr, int act2) {
    if (n % 32) return -1;
    for (int i = 0; i < n; i += 32) {
        int a[8], b[8]; float lo, hi;
        float amax = b2::quantize32(x + i, a, b, lo, hi);
        float dx = amax / 127.0f;
        for (int k = 0; k < 32; k++) {
            int av = (int8_t)(a[k / 4] >> (8 * (k % 4))), bv = (int8_t)(b[k / 4] >> (8 * (k % 4)));
            xr[i + k] = act2 ? dx * ((float)av + (float)bv / B2_ACT_RES) : dx * (float)av;
        }
    }
    return 0;
}
}
