// Probe: latency of cp.async.bulk (global -> shared, mbarrier complete_tx) as seen by the issuing thread.
// Per CTA: thread 0 issues `ncopy` copies of `bytes` each, one after the other (each waited for), and records the
// cycles of every one; the first copy of a freshly started CTA is reported separately from the later ones.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void probe(const uint8_t* src, size_t stride, int bytes, int ncopy, long long* out, int flush_words, unsigned* flush) {
    extern __shared__ __align__(128) uint8_t buf[];
    __shared__ uint64_t bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 0; i < ncopy; i++) {
            const uint8_t* s = src + ((size_t)blockIdx.x * ncopy + i) * stride;
            long long t0 = clock64();
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(smem_u32(buf)), "l"(s), "r"(bytes), "r"(smem_u32(&bar)) : "memory");
            uint32_t ok = 0;
            while (!ok)
                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                             : "=r"(ok) : "r"(smem_u32(&bar)), "r"(i & 1) : "memory");
            long long t1 = clock64();
            out[blockIdx.x * ncopy + i] = t1 - t0;
        }
    }
}
// read one word of every 2 MB page of a large region: evicts the translation caches
__global__ void thrash_tlb(const unsigned* p, size_t n_pages, unsigned* sink) {
    unsigned acc = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n_pages; i += (size_t)gridDim.x * blockDim.x)
        acc += p[i * (2u << 20) / 4 + (i % 64) * 32];
    if (acc == 0xdeadbeef) *sink = acc;
}
__global__ void flush_l2(unsigned* p, size_t n) { for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = (unsigned)i; }
int main() {
    const int ncopy = 4;
    const size_t stride = 1 << 20;        // copies far apart: different lines / pages
    uint8_t* src; cudaMalloc(&src, (size_t)148 * ncopy * stride + (1 << 20)); cudaMemset(src, 1, (size_t)148 * ncopy * stride);
    unsigned* fl; size_t fn = (size_t)256 << 20; cudaMalloc(&fl, fn);
    long long* out; cudaMallocManaged(&out, sizeof(long long) * 148 * ncopy);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    unsigned* big = nullptr; size_t big_pages = 30000;        // ~60 GB
    if (cudaMalloc(&big, big_pages * (2u << 20)) != cudaSuccess) { big_pages = 8000; cudaGetLastError(); cudaMalloc(&big, big_pages * (2u << 20)); }
    printf("thrash region: %zu pages of 2 MB\n", big_pages);
    for (int grid : {1, 64})
        for (int bytes : {4096, 16384})
            for (int cold = 0; cold < 3; cold++) {
                double first = 0, later = 0; int reps = 5;
                for (int r = 0; r < reps; r++) {
                    if (cold) flush_l2<<<1184, 256>>>(fl, fn / 4);
                    if (cold == 2) thrash_tlb<<<592, 256>>>(big, big_pages, fl);
                    probe<<<grid, 128, 64 * 1024>>>(src, stride, bytes, ncopy, out, 0, nullptr);
                    cudaDeviceSynchronize();
                    double f = 0, l = 0;
                    for (int b = 0; b < grid; b++) { f += out[b * ncopy]; for (int i = 1; i < ncopy; i++) l += out[b * ncopy + i]; }
                    first += f / grid; later += l / (grid * (ncopy - 1));
                }
                printf("grid %3d bytes %5d %s: first copy %7.0f cyc, later %7.0f cyc (clock %d kHz nominal)\n", grid, bytes, cold == 2 ? "cold+tlb" : cold ? "cold" : "warm",
                       first / reps, later / reps, clk);
            }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
