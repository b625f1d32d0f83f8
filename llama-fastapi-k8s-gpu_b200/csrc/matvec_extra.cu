// The mat-vec kernel (matvec_body.cuh) instantiated for launches that contain a block format outside llama.cpp's common
// mixes (Q4_1, Q5_0, Q5_1, Q2_K, Q3_K). A translation unit of its own: with all ten formats next to each other the main
// kernel's code came out 1-2 % slower per token on Q4_K_M (measured on the same box, twice).
#include "matvec_body.cuh"

namespace b2 {

cudaError_t mv_extra_init(int smem_max) {
    cudaError_t e = cudaFuncSetAttribute(matvec_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(matvec_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max);
}

cudaError_t mv_extra_launch(cudaLaunchConfig_t* cfg, const MvParams& p, bool act2) {
    return act2 ? cudaLaunchKernelEx(cfg, matvec_kernel<true, true>, p) : cudaLaunchKernelEx(cfg, matvec_kernel<false, true>, p);
}

}  // namespace b2
