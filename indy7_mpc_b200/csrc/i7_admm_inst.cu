// i7_admm_inst.cu — the ADMM QP kernel of qp_mode "osqp" (fp64 only: sigma = 1e-6 against rho = 100 leaves no room in fp32).
#include "i7_launch.h"
#include "i7_admm.cuh"

namespace i7 {

cudaError_t launch_admm(const AdmmParams<double>& P, int smem_bytes, cudaStream_t s) {
    admm_kernel<double><<<P.batch, ADMM_NT, smem_bytes, s>>>(P);
    return cudaGetLastError();
}

cudaError_t smem_cap_admm(int optin_bytes, int* cap) {
    cudaFuncAttributes fa;
    cudaError_t e = cudaFuncGetAttributes(&fa, admm_kernel<double>);
    if (e != cudaSuccess) return e;
    *cap = optin_bytes - (int)fa.sharedSizeBytes;
    return cudaFuncSetAttribute(admm_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, *cap);
}

int admm_smem_elems(int N) { return admm_smem_len(N, N <= ADMM_BLOCKS_IN_SMEM_MAX_N); }
int admm_fac_stride() { return FAC_STRIDE; }

}  // namespace i7
