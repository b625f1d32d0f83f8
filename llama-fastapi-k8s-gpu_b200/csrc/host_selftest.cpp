// Host build of the block arithmetic in quant_blocks.cuh (same source, B2_HOST_EMU), exported
// for tests/test_host_selftest.py: lets the CPU test-suite check the nibble/bit-plane
// unpacking and the two-plane int8 activation scheme against the oracle without a GPU.
#define B2_HOST_EMU 1
#include <math.h>
#include "quant_blocks.cuh"

extern "C" {

// dequantise n_elems elements (whole blocks) with the exact element routine
int b2_host_dequant(int type, const uint8_t* raw, long n_elems, float* out) {
    int be = b2::block_elems(type), bb = b2::block_bytes(type);
    if (!bb || n_elems % be) return -1;
    for (long i = 0; i < n_elems; i++) out[i] = b2::dequant_elem(type, raw + (i / be) * bb, (int)(i % be));
    return 0;
}

// y[r] = W[r,:] . x through quantize32 + chunk_dot, one 64-chunk at a time (what one GPU lane does)
int b2_host_matvec(int type, const uint8_t* w, int n_rows, int K, const float* x, float* y, int act2,
                   int misalign) {
    if (K % 256) return -1;
    long row_bytes = (long)K / b2::block_elems(type) * b2::block_bytes(type);
    // copy each row into a scratch buffer at a chosen alignment (tests the 2-byte-aligned path)
    uint8_t* buf = new uint8_t[row_bytes + 64];
    memset(buf, 0xA5, row_bytes + 64);
    int nchunk = K / 64;
    b2::ActChunk* acts = new b2::ActChunk[nchunk];
    for (int u = 0; u < nchunk; u++) {
        for (int h = 0; h < 2; h++) {
            float lo, hi;
            float amax = b2::quantize32(x + 64 * u + 32 * h, acts[u].a + 8 * h, acts[u].b + 8 * h, lo, hi);
            acts[u].dx[h] = amax / 127.0f;
            acts[u].s16[2 * h] = lo; acts[u].s16[2 * h + 1] = hi;
        }
    }
    uintptr_t base = ((uintptr_t)buf + 15) & ~(uintptr_t)15;
    uint8_t* rowp = (uint8_t*)base + misalign;
    for (int r = 0; r < n_rows; r++) {
        memcpy(rowp, w + r * row_bytes, row_bytes);
        float acc = 0.f;
        for (int u = 0; u < nchunk; u++)
            acc += act2 ? b2::chunk_dot<true>(type, rowp, u, acts[u]) : b2::chunk_dot<false>(type, rowp, u, acts[u]);
        y[r] = acc;
    }
    delete[] acts; delete[] buf;
    return 0;
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
