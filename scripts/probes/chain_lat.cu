// Probe: latency of the first cp.async.bulk of a kernel that runs right behind a heavy kernel in the same stream
// (plain stream order, or programmatic dependent launch with griddepcontrol.wait before the copy).
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
// K1: mode 1 = plain-load streaming, 2 = bulk-copy streaming, 3 = arithmetic spin
__global__ void __launch_bounds__(512, 1) k1(const uint4* src, size_t n16_per_cta, int mode, unsigned* sink, int trigger) {
    extern __shared__ __align__(128) uint8_t buf[];
    __shared__ uint64_t bar[2];
    if (trigger) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    unsigned acc = 0;
    const uint4* p = src + (size_t)blockIdx.x * n16_per_cta;
    if (mode == 1) {
        for (size_t i = threadIdx.x; i < n16_per_cta; i += 512) { uint4 v = __ldcg(p + i); acc ^= v.x ^ v.w; }
    } else if (mode == 2) {
        if (threadIdx.x == 0) {
            for (int s = 0; s < 2; s++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[s])));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            const uint32_t chunk = 32768; const size_t n = n16_per_cta * 16 / chunk;
            for (size_t c = 0; c < n + 1; c++) {
                if (c < n) {
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar[c & 1])), "r"(chunk) : "memory");
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 ::"r"(smem_u32(buf + (c & 1) * chunk)), "l"((const uint8_t*)p + c * chunk), "r"(chunk), "r"(smem_u32(&bar[c & 1])) : "memory");
                }
                if (c > 0) {
                    uint32_t ok = 0; const size_t d = c - 1;
                    while (!ok) asm volatile("{\n\t.reg .pred q;\n\tmbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2;\n\tselp.u32 %0, 1, 0, q;\n\t}"
                                             : "=r"(ok) : "r"(smem_u32(&bar[d & 1])), "r"((unsigned)((d >> 1) & 1)) : "memory");
                }
            }
        }
    } else if (mode == 3) {
        long long t0 = clock64();
        while (clock64() - t0 < 20000) acc += 1;
    }
    if (acc == 0x12345678u) *sink = acc;
}
__global__ void __launch_bounds__(256) k2(const uint8_t* src, size_t stride, int bytes, long long* out, int wait_first) {
    extern __shared__ __align__(128) uint8_t buf[];
    __shared__ uint64_t bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (wait_first) asm volatile("griddepcontrol.wait;" ::: "memory");
    if (threadIdx.x == 0) {
        const uint8_t* s = src + (size_t)blockIdx.x * stride;
        long long t0 = clock64();
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(2 * bytes) : "memory");
        for (int k = 0; k < 2; k++)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(smem_u32(buf + k * 16384)), "l"(s + k * (stride / 2)), "r"(bytes), "r"(smem_u32(&bar)) : "memory");
        uint32_t ok = 0;
        while (!ok) asm volatile("{\n\t.reg .pred q;\n\tmbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2;\n\tselp.u32 %0, 1, 0, q;\n\t}"
                                 : "=r"(ok) : "r"(smem_u32(&bar)), "r"(0) : "memory");
        out[blockIdx.x] = clock64() - t0;
    }
}
int main() {
    const size_t per_cta = 448 * 1024;                    // K1 streams 148 * 448 KB = 66 MB per launch, from a 4 GB region (fresh bytes every launch)
    const size_t region = (size_t)4 << 30;
    uint8_t* w; cudaMalloc(&w, region); cudaMemset(w, 1, region);
    uint8_t* kv; const size_t stride = 1 << 20; cudaMalloc(&kv, 64 * stride * 64); cudaMemset(kv, 2, 64 * stride * 64);
    long long* out; cudaMalloc(&out, 64 * 8); unsigned* sink; cudaMalloc(&sink, 4);
    cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, 82 * 1024);
    cudaStream_t st; cudaStreamCreate(&st);
    for (int mode : {0, 1, 2, 3})
        for (int pdl : {0, 1}) {
            double sum = 0, mx = 0; int n = 0;
            for (int it = 0; it < 24; it++) {
                const uint4* src = (const uint4*)(w + ((size_t)it * 148 * per_cta) % (region - 148 * per_cta));
                if (mode) k1<<<148, 512, 200 * 1024, st>>>(src, per_cta / 16, mode, sink, pdl);
                cudaLaunchConfig_t cfg = {};
                cfg.gridDim = dim3(64); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = 82 * 1024; cfg.stream = st;
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
                cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
                cudaLaunchKernelEx(&cfg, k2, (const uint8_t*)(kv + (size_t)(it % 64) * 64 * stride), stride, 4864, out, pdl);
                cudaStreamSynchronize(st);
                std::vector<long long> h(64); cudaMemcpy(h.data(), out, 64 * 8, cudaMemcpyDeviceToHost);
                if (it >= 4) for (auto v : h) { sum += v; mx = v > mx ? v : mx; n++; }
            }
            printf("K1 mode %d (%s), %s: first bulk copy of K2 lands after %7.0f cycles on average, %7.0f at worst\n", mode,
                   mode == 0 ? "none" : mode == 1 ? "ldg stream" : mode == 2 ? "bulk stream" : "spin", pdl ? "PDL + wait" : "stream order", sum / n, mx);
        }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
