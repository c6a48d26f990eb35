// i7_launch.h — host-side launch entry points of the fused kernels.  Each (arithmetic type, CTA size) instance of
// the big kernels is compiled in its own translation unit (i7_sqp_inst.cu / i7_cluster_inst.cu with -DI7_T / -DI7_NT)
// so that the instances build in parallel; i7_api.cu only sees these declarations.
#pragma once
#include "i7_sqp.cuh"

namespace i7 {

// CTA-per-problem kernel (throughput form): grid = batch
template <typename T, int NT, bool SPILL> cudaError_t launch_sqp_cta(const Robot<T>& rob, const SolveParams<T>& P, int smem_bytes,
                                                         cudaStream_t s);
// raises the dynamic shared-memory cap of that instance to everything the SM offers; returns the cap
template <typename T, int NT, bool SPILL> cudaError_t smem_cap_sqp_cta(int optin_bytes, int* cap);

// cluster-per-problem kernel (latency form): grid = batch x CL CTAs (CL = 8, 4 or 2); xbuf = batch x cluster_xbuf_len(N)
// exchange buffer
template <typename T, int CL> cudaError_t launch_sqp_cluster(const Robot<T>& rob, const SolveParams<T>& P, T* xbuf, int smem_bytes,
                                                     cudaStream_t s);
template <typename T, int CL> cudaError_t smem_cap_sqp_cluster(int optin_bytes, int* cap);
template <typename T, int CL> cudaError_t max_active_sqp_clusters(int smem_bytes, int* n);     // co-resident clusters
int cluster_smem_len(int N, int CL);     // shared-memory elements per CTA of the cluster kernel
int cluster_xbuf_len(int N);     // exchange-buffer elements per problem
constexpr int kClusterSize = 8;
constexpr int kClusterThreads = 384;     // = CL_NT of i7_cluster.cuh (checked in i7_cluster_inst.cu)


// ADMM QP kernel of qp_mode "osqp" (i7_admm.cuh; fp64): grid = batch
struct AdmmSettings;
template <typename T> struct AdmmParams;
cudaError_t launch_admm(const AdmmParams<double>& P, int smem_bytes, cudaStream_t s);
cudaError_t smem_cap_admm(int optin_bytes, int* cap);
int admm_smem_elems(int N);
int admm_fac_stride();

}  // namespace i7
