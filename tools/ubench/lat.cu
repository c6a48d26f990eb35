// Latency microbenchmarks for design decisions (single warp unless noted): dependent DFMA/FFMA chain, fp64 and fp32
// reciprocal / division / sqrt, shared-memory and generic loads, __syncthreads at several CTA sizes, shuffles.
#include <cstdio>
#include <cuda_runtime.h>
#define N 4096
template <typename T> __global__ void k_fma(T* out, T a, T b) {
    T x = out[0];
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) x = x * a + b;
    long long t1 = clock64();
    out[0] = x; ((long long*)out)[4] = t1 - t0;
}
template <typename T> __global__ void k_rcp(T* out) {
    T x = out[0];
    long long t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < N; ++i) x = T(1) / x + T(0.5);
    long long t1 = clock64();
    out[0] = x; ((long long*)out)[4] = t1 - t0;
}
__global__ void k_rcp_fast(double* out) {     // fp32 seed + 2 Newton steps in fp64
    double x = out[0];
    long long t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < N; ++i) {
        double r = (double)__frcp_rn((float)x);
        double e = fma(-x, r, 1.0); r = fma(r, e, r);
        e = fma(-x, r, 1.0); r = fma(r, e, r);
        x = r + 0.5;
    }
    long long t1 = clock64();
    out[0] = x; ((long long*)out)[4] = t1 - t0;
}
__global__ void k_rcp_approx(double* out) {   // rcp.approx.ftz.f64 + 2 Newton steps
    double x = out[0];
    long long t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < N; ++i) {
        double r; asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
        double e = fma(-x, r, 1.0); r = fma(r, e, r);
        e = fma(-x, r, 1.0); r = fma(r, e, r);
        x = r + 0.5;
    }
    long long t1 = clock64();
    out[0] = x; ((long long*)out)[4] = t1 - t0;
}
template <typename T> __global__ void k_sqrt(T* out) {
    T x = out[0];
    long long t0 = clock64();
#pragma unroll 4
    for (int i = 0; i < N; ++i) x = sqrt(x) + T(1.5);
    long long t1 = clock64();
    out[0] = x; ((long long*)out)[4] = t1 - t0;
}
__global__ void k_sincos(double* out) {
    double x = out[0];
    long long t0 = clock64();
    for (int i = 0; i < 256; ++i) { double s, c; sincos(x, &s, &c); x = s + c; }
    long long t1 = clock64();
    out[0] = x; ((long long*)out)[4] = t1 - t0;
}
__global__ void k_lds(double* out, int generic) {
    __shared__ double sm[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = (double)((i * 7 + 1) % 1024);
    __syncthreads();
    double* p = generic ? (double*)__cvta_generic_to_global(out) : sm;   // dummy to keep 'generic' opaque
    volatile double* base = sm;
    int idx = threadIdx.x;
    long long t0 = clock64();
    if (!generic) {
#pragma unroll 8
        for (int i = 0; i < N; ++i) idx = (int)sm[idx & 1023];
    } else {
        double* g = sm;                    // generic pointer into shared memory
        asm volatile("" : "+l"(g));
#pragma unroll 8
        for (int i = 0; i < N; ++i) idx = (int)g[idx & 1023];
    }
    long long t1 = clock64();
    out[0] = idx + (p == nullptr) + base[0]; ((long long*)out)[4] = t1 - t0;
}
__global__ void k_ldg(double* out, const double* buf) {
    int idx = threadIdx.x;
    long long t0 = clock64();
#pragma unroll 8
    for (int i = 0; i < N; ++i) idx = (int)buf[idx & 1023];
    long long t1 = clock64();
    out[0] = idx; ((long long*)out)[4] = t1 - t0;
}
__global__ void k_sync(double* out) {
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) __syncthreads();
    long long t1 = clock64();
    if (threadIdx.x == 0) ((long long*)out)[4] = t1 - t0;
}
__global__ void k_syncwarp_shfl(double* out) {
    double x = out[0] + threadIdx.x;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) x += __shfl_xor_sync(0xffffffffu, x, 1);
    long long t1 = clock64();
    out[threadIdx.x + 8] = x; if (threadIdx.x == 0) ((long long*)out)[4] = t1 - t0;
}
int main() {
    double* d; cudaMalloc(&d, 4096); double* buf; cudaMalloc(&buf, 8192);
    double hb[1024]; for (int i = 0; i < 1024; ++i) hb[i] = (double)((i * 7 + 1) % 1024);
    cudaMemcpy(buf, hb, sizeof hb, cudaMemcpyHostToDevice);
    auto rd = [&](const char* name, int n) { long long h[8]; cudaDeviceSynchronize(); cudaMemcpy(h, d, 64, cudaMemcpyDeviceToHost);
        printf("%-28s %8.2f cycles/op\n", name, (double)h[4] / n); double one = 1.25; cudaMemcpy(d, &one, 8, cudaMemcpyHostToDevice); };
    double one = 1.25; cudaMemcpy(d, &one, 8, cudaMemcpyHostToDevice);
    for (int rep = 0; rep < 2; ++rep) {
        k_fma<double><<<1, 32>>>(d, 0.999, 0.001); rd("DFMA dependent", N);
        { float f = 1.25f; cudaMemcpy(d, &f, 4, cudaMemcpyHostToDevice); }
        k_fma<float><<<1, 32>>>((float*)d, 0.999f, 0.001f); rd("FFMA dependent", N);
        k_rcp<double><<<1, 32>>>(d); rd("1/x fp64 (+add)", N);
        { float f = 1.25f; cudaMemcpy(d, &f, 4, cudaMemcpyHostToDevice); }
        k_rcp<float><<<1, 32>>>((float*)d); rd("1/x fp32 (+add)", N);
        k_rcp_fast<<<1, 32>>>(d); rd("rcp f32 seed + 2 Newton", N);
        k_rcp_approx<<<1, 32>>>(d); rd("rcp.approx.f64 + 2 Newton", N);
        k_sqrt<double><<<1, 32>>>(d); rd("sqrt fp64 (+add)", N);
        k_sincos<<<1, 32>>>(d); rd("sincos fp64 (+add)", 256);
        k_lds<<<1, 32>>>(d, 0); rd("LDS.64 dependent (+cvt)", N);
        k_lds<<<1, 32>>>(d, 1); rd("generic LD->smem dep (+cvt)", N);
        k_ldg<<<1, 32>>>(d, buf); rd("LDG L1-hit dependent (+cvt)", N);
        k_sync<<<1, 32>>>(d); rd("__syncthreads 1 warp", N);
        k_sync<<<1, 128>>>(d); rd("__syncthreads 4 warps", N);
        k_sync<<<1, 384>>>(d); rd("__syncthreads 12 warps", N);
        k_syncwarp_shfl<<<1, 32>>>(d); rd("shfl f64 + add dependent", N);
    }
    return 0;
}
