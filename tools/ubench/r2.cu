// Microbenchmark of the serial core of one backward-sweep stage: 6x6 SPD solve with one right-hand side per thread
// (13 threads of warp 0), the operands read from / written to shared memory, repeated over 31 dependent stages.
// Variants: V0 library division (as in spd6_factor), V1 rcp.approx + 2 Newton steps, V2 also skips the barrier.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../indy7_mpc_b200/csrc/i7_rbd.cuh"
using namespace i7;
#define AUGW 19
__device__ __forceinline__ double rcp_fast(double x) {
    double r; asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    double e = fma(-x, r, 1.0); r = fma(r, e, r);
    e = fma(-x, r, 1.0); r = fma(r, e, r);
    return r;
}
template <int V> __device__ __forceinline__ void factor(double (&A)[6][6], double (&dinv)[6]) {
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        dinv[j] = V == 0 ? 1.0 / A[j][j] : rcp_fast(A[j][j]);
        double col[6];
#pragma unroll
        for (int i = j + 1; i < 6; ++i) col[i] = A[i][j];
#pragma unroll
        for (int i = j + 1; i < 6; ++i) {
            const double l = col[i] * dinv[j];
#pragma unroll
            for (int k = j + 1; k <= i; ++k) A[i][k] -= l * col[k];
            A[i][j] = l;
        }
    }
}
template <int V, bool BAR> __global__ void k_r2(double* out, int stages) {
    __shared__ double aug[2][6 * AUGW];
    __shared__ double X[6 * 13];
    const int tid = threadIdx.x;
    for (int t = tid; t < 2 * 6 * AUGW; t += blockDim.x) {
        const int e = t % (6 * AUGW), r = e / AUGW, c = e % AUGW;
        (&aug[0][0])[t] = (c < 6 ? (r == c ? 4.0 + r : 0.3 / (1 + r + c)) : 0.1 * (c - r));
    }
    __syncthreads();
    long long t0 = clock64();
    int cur = 0;
    for (int s = 0; s < stages; ++s) {
        if (tid < 13) {
            double A[6][6], dinv[6], b[6];
#pragma unroll
            for (int i = 0; i < 6; ++i)
#pragma unroll
                for (int q = 0; q <= i; ++q) A[i][q] = aug[cur][AUGW * i + q];
#pragma unroll
            for (int i = 0; i < 6; ++i) b[i] = aug[cur][AUGW * i + 6 + tid];
            factor<V>(A, dinv);
            spd6_solve_factored<double>(A, dinv, b);
#pragma unroll
            for (int i = 0; i < 6; ++i) X[13 * i + tid] = -b[i];
        }
        if (BAR) __syncthreads(); else __syncwarp();
        // dependent follow-up: next aug depends on X (stand-in for R3)
        if (tid < 114) {
            const int r = tid / AUGW, c = tid % AUGW;
            double v = aug[cur][tid];
            if (c < 6) v += 1e-3 * X[13 * r + c] * X[13 * c + r];
            else if (c < 19) v += 1e-3 * X[13 * r + (c - 6)];
            aug[cur ^ 1][tid] = v;
        }
        if (BAR) __syncthreads(); else __syncwarp();
        cur ^= 1;
    }
    long long t1 = clock64();
    if (tid == 0) { out[0] = aug[cur][0]; ((long long*)out)[1] = t1 - t0; }
}
template <int V, bool BAR> void run(const char* name, int threads) {
    double* d; cudaMalloc(&d, 64);
    long long h[2];
    for (int rep = 0; rep < 2; ++rep) {
        k_r2<V, BAR><<<1, threads>>>(d, 31);
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    }
    printf("%-44s threads %4d: %7.1f cycles/stage (%s)\n", name, threads, (double)h[1] / 31, cudaGetErrorString(cudaGetLastError()));
    cudaFree(d);
}
int main() {
    run<0, true>("V0 1/x, __syncthreads", 128);
    run<0, true>("V0 1/x, __syncthreads", 384);
    run<1, true>("V1 rcp.approx+2N, __syncthreads", 128);
    run<1, true>("V1 rcp.approx+2N, __syncthreads", 384);
    run<1, false>("V1 rcp.approx+2N, single warp syncwarp", 32);
    run<0, false>("V0 1/x, single warp syncwarp", 32);
    return 0;
}
