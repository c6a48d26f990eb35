// i7_sqp_inst.cu — one instance of the CTA-per-problem fused kernel; compiled once per (I7_T, I7_NT).
#include "i7_launch.h"

#ifndef I7_T
#error "compile with -DI7_T=double|float -DI7_NT=128|384 -DI7_SPILL=false|true"
#endif

namespace i7 {

template <>
cudaError_t launch_sqp_cta<I7_T, I7_NT, I7_SPILL>(const Robot<I7_T>& rob, const SolveParams<I7_T>& P, int smem_bytes,
                                        cudaStream_t s) {
    sqp_kernel<I7_T, I7_NT, I7_SPILL><<<P.batch, I7_NT, smem_bytes, s>>>(rob, P);
    return cudaGetLastError();
}

template <>
cudaError_t smem_cap_sqp_cta<I7_T, I7_NT, I7_SPILL>(int optin_bytes, int* cap) {
    cudaFuncAttributes fa;
    cudaError_t e = cudaFuncGetAttributes(&fa, sqp_kernel<I7_T, I7_NT, I7_SPILL>);
    if (e != cudaSuccess) return e;
    *cap = optin_bytes - (int)fa.sharedSizeBytes;
    return cudaFuncSetAttribute(sqp_kernel<I7_T, I7_NT, I7_SPILL>, cudaFuncAttributeMaxDynamicSharedMemorySize, *cap);
}

}  // namespace i7
