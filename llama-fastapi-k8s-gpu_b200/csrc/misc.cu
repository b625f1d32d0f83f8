// Small kernels: block dequantisation (bit-exact vs gguf-py), embedding-row gather, RMSNorm.
// Replace ggml-cuda's dequantize_block_* (convert.cu), the host-side get_rows of token_embd and
// rms_norm_f32 + mul that the reference path uses.
#include "kernels.h"
#include "ptx.cuh"
#include "quant_blocks.cuh"
#include <mutex>

namespace b2 {

thread_local bool g_chain_pdl = false;

unsigned int* kernel_error_word() {
    static unsigned int* w = nullptr;
    static std::once_flag once;               // first use may come from any host thread
    std::call_once(once, []() {
        unsigned int* p = nullptr;
        if (cudaHostAlloc(&p, 8 * sizeof(unsigned int), cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) return;
        for (int i = 0; i < 8; i++) p[i] = 0;
        w = p;
    });
    return w;
}

__global__ void dequantize_kernel(int type, const uint8_t* __restrict__ raw, long n, float* __restrict__ out) {
    const int be = block_elems(type), bb = block_bytes(type);
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
        out[i] = dequant_elem(type, raw + (i / be) * bb, (int)(i % be));
}

cudaError_t dequantize_launch(int type, const uint8_t* raw, long n_elems, float* out, cudaStream_t s) {
    if (!block_bytes(type) || n_elems % block_elems(type)) return cudaErrorInvalidValue;
    if (n_elems == 0) return cudaSuccess;
    long blocks = (n_elems + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    dequantize_kernel<<<(unsigned)blocks, 256, 0, s>>>(type, raw, n_elems, out);
    return cudaGetLastError();
}

__global__ void embed_row_kernel(int type, const uint8_t* __restrict__ table, int K, const int* token_ptr,
                                 float* __restrict__ x) {
    griddep_launch();
    griddep_wait();                       // the token id comes from the previous step's sampler
    const int be = block_elems(type), bb = block_bytes(type);
    const uint8_t* row = table + (size_t)(*token_ptr) * ((size_t)K / be * bb);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < K; i += gridDim.x * blockDim.x)
        x[i] = dequant_elem(type, row + (size_t)(i / be) * bb, i % be);
}

cudaError_t embed_row_launch(int type, const uint8_t* table, int K, const int* token_ptr, float* x,
                             cudaStream_t s, bool pdl) {
    if (!block_bytes(type) || K % block_elems(type)) return cudaErrorInvalidValue;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((K + 255) / 256); cfg.blockDim = dim3(256); cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, embed_row_kernel, type, table, K, token_ptr, x);
}

// h[i] = silu(g[i]) * h[i]: the un-fused tail of a gate/up pair whose two matrices have different quant types
__global__ void silu_mul_kernel(const float* __restrict__ g, float* __restrict__ h, int n) {
    griddep_launch();
    griddep_wait();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float a = g[i];
        h[i] = a / (1.0f + __expf(-a)) * h[i];
    }
}

cudaError_t silu_mul_launch(const float* g, float* h, int n, cudaStream_t s, bool pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((n + 255) / 256); cfg.blockDim = dim3(256); cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, silu_mul_kernel, g, h, n);
}

// y[r,:] = x[r,:] * rsqrt(mean(x[r,:]^2) + eps) * w      (one CTA per row)
__global__ void rmsnorm_kernel(const float* __restrict__ x, const float* __restrict__ w, float eps, int K,
                               float* __restrict__ y) {
    __shared__ float red[32];
    const float* xr = x + (size_t)blockIdx.x * K;
    float* yr = y + (size_t)blockIdx.x * K;
    float ss = 0.f;
    for (int i = threadIdx.x; i < K; i += blockDim.x) { float v = xr[i]; ss += v * v; }
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    float tot = 0.f;
    for (int i = 0; i < (int)(blockDim.x >> 5); i++) tot += red[i];
    float rs = rsqrtf(tot / (float)K + eps);
    for (int i = threadIdx.x; i < K; i += blockDim.x) yr[i] = xr[i] * rs * w[i];
}

cudaError_t rmsnorm_launch(const float* x, const float* w, float eps, int K, int rows, float* y, cudaStream_t s) {
    if (rows <= 0) return cudaSuccess;
    rmsnorm_kernel<<<rows, 256, 0, s>>>(x, w, eps, K, y);
    return cudaGetLastError();
}

}  // namespace b2
