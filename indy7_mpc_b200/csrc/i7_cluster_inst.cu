// i7_cluster_inst.cu — one instance of the cluster-per-problem (latency) kernel; compiled once per I7_T.
#include "i7_cluster.cuh"
#include "i7_launch.h"

#if !defined(I7_T) || !defined(I7_CL)
#error "compile with -DI7_T=double|float -DI7_CL=8|4|2"
#endif

namespace i7 {

static_assert(CL_NT == kClusterThreads, "i7_launch.h: kClusterThreads out of date");

template <>
cudaError_t launch_sqp_cluster<I7_T, I7_CL>(const Robot<I7_T>& rob, const SolveParams<I7_T>& P, I7_T* xbuf, int smem_bytes,
                                     cudaStream_t s) {
    sqp_cluster_kernel<I7_T, I7_CL><<<P.batch * I7_CL, CL_NT, smem_bytes, s>>>(rob, P, xbuf);
    return cudaGetLastError();
}

template <>
cudaError_t smem_cap_sqp_cluster<I7_T, I7_CL>(int optin_bytes, int* cap) {
    cudaFuncAttributes fa;
    cudaError_t e = cudaFuncGetAttributes(&fa, sqp_cluster_kernel<I7_T, I7_CL>);
    if (e != cudaSuccess) return e;
    *cap = optin_bytes - (int)fa.sharedSizeBytes;
    return cudaFuncSetAttribute(sqp_cluster_kernel<I7_T, I7_CL>, cudaFuncAttributeMaxDynamicSharedMemorySize, *cap);
}

// how many clusters of this kernel the device can hold at once (GPC granularity: fewer than SMs / 8 on B200)
template <>
cudaError_t max_active_sqp_clusters<I7_T, I7_CL>(int smem_bytes, int* n) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(I7_CL, 1, 1);
    cfg.blockDim = dim3(CL_NT, 1, 1);
    cfg.dynamicSmemBytes = (size_t)smem_bytes;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = I7_CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaOccupancyMaxActiveClusters(n, sqp_cluster_kernel<I7_T, I7_CL>, &cfg);
}

#ifdef I7_CLUSTER_DEFINE_LAYOUT
int cluster_smem_len(int N, int CL) { return make_clayout(N, CL).total; }
int cluster_xbuf_len(int N) { return xbuf_size(N); }
#endif

}  // namespace i7
