// i7_admm.cuh — qp_mode "osqp": the QP of one SQP pass solved the way the reference solves it, by OSQP's ADMM
// (src/osqp_solver.py:39-41,137-143; Stellato et al. 2020, Alg. 1) at the reference's tolerances (eps_abs = eps_rel =
// 1e-3, check every 25 iterations, rho = 0.1 x 1e3 on the equality rows, sigma = 1e-6, alpha = 1.6) with a per-problem
// warm start (x, y) kept in the handle, and — SURVEY.md section 8 f-4 — optional box rows lo <= z_i <= hi on every
// q_k, v_k (k >= 1) and u_k (description/indy7.urdf:203-238).  The tests hold it against a CPU restatement of the same iteration.
//
// OSQP factorises the quasi-definite KKT matrix [[P + sigma I, A'], [A, -1/rho]] with a sparse LDL'.  Here the
// multipliers are eliminated (nu = rho (A x~ - z) + y), which leaves the SPD system
//        (P + sigma I + A' rho A) x~ = sigma x - q + A' (rho z - y),        z~ = A x~
// whose matrix is block tridiagonal over the knots (18 x 18 blocks [x_k u_k], coupling only through x_{k+1}).  One
// CTA per problem:
//   factor   once per QP: block Cholesky in stage order.  Per stage, one warp factorises the 18 x 18 pivot block and
//            inverts its triangular factor in registers (lane i = row i / column i, operands exchanged by shuffles);
//            all threads form the products  G_k = L_k^-1 E_k,  M_k = L_k^-1[:, x] G_{k-1}',  F_k = L_k^-T G_k.
//            (explicit INVERSES OF THE TRIANGULAR FACTORS keep the accuracy of a Cholesky solve — 1e-10 against the
//            reference form on these problems, cond ~ 4e9 — where explicit inverses of the pivot blocks lose 1e-6)
//   iterate  right-hand side and z~ = A x~ in parallel over rows; the solve is two parallel triangular products
//            (a = L^-1 r, c = L^-T w) around two serial chains of ONE small matrix-vector product per stage
//            (w_k = a_k - M_k w_{k-1} forward, x_k = c_k - F_k x_{k+1} backward) run by one warp with the stage matrices
//            prefetched into registers one stage ahead.
#pragma once
#include "i7_sqp.cuh"

namespace i7 {

struct AdmmSettings {
    double rho_eq, rho_x, rho_u, sigma, alpha, eps_abs, eps_rel;
    int max_iter, check, limits;
    double qmin[6], qmax[6], vmax[6], umax[6];
};

template <typename T>
struct AdmmParams {
    const T* blk;     // batch x N x QB_STRIDE          (MODE_EXPORT of the fused kernel)
    const T* xu;      // batch x (18N-6)                 iterate the QP was linearised at (the linear cost term needs it)
    const T* x0;      // batch x 12
    T* wx;            // batch x 18N                     warm start / final x   (z order, padded to 18 per knot)
    T* wy;            // batch x 12N                     duals of the equality rows
    T* wyl;           // batch x 18N                     duals of the limit rows (z order; unused without limits)
    T* z_out;         // batch x (18N-6)                 = x
    T* y_out;         // optional, batch x 12N
    T* info;          // batch x 4: iterations, primal residual, dual residual, status (0 solved, 1 max_iter)
    T* fac;           // workspace: batch x N x FAC_STRIDE
    const int* done;  // optional skip flags
    long long* prof;  // optional: per problem 32 cycle counters (debugging aid): [0] factorisation, [1] vectors, [2..10] steps 1-9 summed over the iterations
    int N, batch;
    T dt;
    AdmmSettings S;
};

constexpr int ADMM_NT = 128;     // two CTAs per SM (registers: the in-register 18 x 18 factorisation), three where shared memory allows
constexpr int FAC_LINV = 0, FAC_MT = 324, FAC_FT = 648, FAC_STRIDE = 864;
// shared memory: the vectors (156 N), the stage scratch of the factorisation, and — while it fits — the QP blocks
// (QB_STRIDE N), which every iteration reads twice (A x~ and A' t)
__host__ __device__ inline int admm_smem_len(int N, bool blocks) { return 156 * N + 2 * 324 + 2 * 216 + 108 + 16 + 64 + (blocks ? QB_STRIDE * N : 0); }
constexpr int ADMM_BLOCKS_IN_SMEM_MAX_N = 64;

// element (r, c) of G^ = [Gq Gv Hm] (6 x 18) in a QB block
template <typename T> I7_DEV T ghat(const T* blk, int r, int c) {
    return c < 6 ? blk[QB_GQ + 6 * r + c] : (c < 12 ? blk[QB_GV + 6 * r + c - 6] : blk[QB_HM + 6 * r + c - 12]);
}

template <typename T>
__global__ void __launch_bounds__(ADMM_NT, 2) admm_kernel(const __grid_constant__ AdmmParams<T> P) {
    constexpr int NT = ADMM_NT;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* sm = reinterpret_cast<T*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, N = P.N;
    const int nz = 18 * N - 6, nr = 12 * N, nv = 18 * N;
    const AdmmSettings& S = P.S;
    const T rho = T(S.rho_eq), sigma = T(S.sigma), al = T(S.alpha), dt = P.dt;
    // vectors (z order padded to 18 per knot: index 18 k + e; rows: 12 k + e)
    T* s_x = sm;                 // 18N
    T* s_q = s_x + nv;           // 18N
    T* s_b1 = s_q + nv;          // 18N
    T* s_b2 = s_b1 + nv;         // 18N
    T* s_yl = s_b2 + nv;         // 18N
    T* s_zl = s_yl + nv;         // 18N
    T* s_y = s_zl + nv;          // 12N
    T* s_l = s_y + nr;           // 12N
    T* s_z = s_l + nr;           // 12N
    T* s_t = s_z + nr;           // 12N
    T* s_D = s_t + nr;           // 324  pivot block (row-major 18 x 18, lower part used)
    T* s_X = s_D + 324;          // 324  inverse of its Cholesky factor (row-major, zeros above the diagonal)
    T* s_G = s_X + 324;          // 216  G_{k-1} (18 x 12 row-major)
    T* s_Gn = s_G + 216;         // 216  G_k
    T* s_Gh = s_Gn + 216;        // 108  G^_k (6 x 18)
    T* s_ck = s_Gh + 108;        // 16   cost terms of the knot
    T* s_red = s_ck + 16;        // 64   reduction scratch
    T* s_blk = s_red + 64;       // QB_STRIDE N  the QP blocks (short horizons only)
    const bool blk_smem = N <= ADMM_BLOCKS_IN_SMEM_MAX_N;
    __shared__ int s_stop;

    for (int b = blockIdx.x; b < P.batch; b += gridDim.x) {
        if (P.done && P.done[b]) continue;
        const T* blk = P.blk + (size_t)b * N * QB_STRIDE;
        if (blk_smem) {
            for (int t = tid; t < N * QB_STRIDE; t += NT) s_blk[t] = blk[t];
            blk = s_blk;
            __syncthreads();
        }
        T* fac = P.fac + (size_t)b * N * FAC_STRIDE;
        const T* xu = P.xu + (size_t)b * nz;
        const bool lim = S.limits != 0;
        auto rho_l = [&](int e) { return e < 12 ? T(S.rho_x) : T(S.rho_u); };
        auto lo_of = [&](int e) { return e < 6 ? T(S.qmin[e]) : (e < 12 ? T(-S.vmax[e - 6]) : T(-S.umax[e - 12])); };
        auto hi_of = [&](int e) { return e < 6 ? T(S.qmax[e]) : (e < 12 ? T(S.vmax[e - 6]) : T(S.umax[e - 12])); };

        long long* prof = P.prof ? P.prof + (size_t)b * 32 : nullptr;
        long long pacc[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, ptt = prof ? clock64() : 0;
// every thread keeps the counters (a branch on tid == 0 here would leave warp 0 diverged in front of its shuffle chains)
#define I7_ACC(i) do { if (prof) { const long long n_ = clock64(); pacc[i] += n_ - ptt; ptt = n_; } } while (0)
        // ---------------------------------------------------------------- factorisation ----------------
        for (int k = 0; k < N; ++k) {
            const T* bk = blk + (size_t)k * QB_STRIDE;
            if (tid < 108) s_Gh[tid] = k < N - 1 ? ghat<T>(bk, tid / 18, tid % 18) : T(0);
            if (tid >= 112 && tid < 112 + 9) s_ck[tid - 112] = bk[QB_COST + tid - 112];
            __syncthreads();
            // pivot block D = H_kk - G_{k-1}' G_{k-1} (x x part), lower triangle
            for (int e = tid; e < 324; e += NT) {
                const int a = e / 18, c = e - 18 * a;
                if (c > a) continue;
                T v = T(0);
                if (k < N - 1) {
                    T acc = T(0);
#pragma unroll
                    for (int r = 0; r < 6; ++r) acc += s_Gh[18 * r + a] * s_Gh[18 * r + c];
                    if (a < 12 && (a % 6) == (c % 6)) acc += (a < 6 ? T(1) : dt) * (c < 6 ? T(1) : dt);
                    v = rho * acc;
                }
                if (a < 12 && a == c) v += rho;                                   // the -x_k of row block k
                if (a < 6) v += s_ck[6] * s_ck[a] * s_ck[c];
                else if (a == c) v += a < 12 ? s_ck[7] : s_ck[8];
                if (a == c) v += sigma + ((lim && (k > 0 || a >= 12)) ? rho_l(a) : T(0));
                if (k == N - 1 && a >= 12) v = (a == c) ? T(1) : T(0);             // padding variables
                if (k > 0 && a < 12) {
                    T acc = T(0);
#pragma unroll
                    for (int i = 0; i < 18; ++i) acc += s_G[12 * i + a] * s_G[12 * i + c];
                    v -= acc;
                }
                s_D[e] = v;
            }
            __syncthreads();
            // Cholesky D = L L' and X = L^-1 by warp 0, in registers: lane i owns row i of L and column i of X
            if (warp == 0) {
                __syncwarp();
                T A[18], X[18], rr[18];
#pragma unroll
                for (int c = 0; c < 18; ++c) A[c] = (lane < 18 && c <= lane) ? s_D[18 * lane + c] : T(0);
#pragma unroll
                for (int j = 0; j < 18; ++j) {
                    const T d = __shfl_sync(0xffffffffu, A[j], j);
                    const T r = rsqrt(d);
                    rr[j] = r;
                    const T lij = A[j] * r;                    // lane j: sqrt(d); lanes i > j: L[i][j]
                    A[j] = lij;
#pragma unroll
                    for (int c = j + 1; c < 18; ++c) {
                        const T lcj = __shfl_sync(0xffffffffu, lij, c);
                        A[c] -= lij * lcj;                      // meaningful for lanes i >= c
                    }
                }
#pragma unroll
                for (int i = 0; i < 18; ++i) {
                    T acc = (i == lane) ? T(1) : T(0);
#pragma unroll
                    for (int c = 0; c < i; ++c) {
                        const T lic = __shfl_sync(0xffffffffu, A[c], i);     // L[i][c]
                        acc -= lic * X[c];
                    }
                    X[i] = (i >= lane) ? acc * rr[i] : T(0);
                }
                if (lane < 18) {
#pragma unroll
                    for (int i = 0; i < 18; ++i) s_X[18 * i + lane] = X[i];
                }
            }
            __syncthreads();
            // G_k = X E_k with E_k = -rho [A_k B_k]' ; M_k = X[:, :12] G_{k-1}' ; L^-1 to the workspace
            T* fk = fac + (size_t)k * FAC_STRIDE;
            for (int e = tid; e < 324; e += NT) fk[FAC_LINV + e] = s_X[e];
            if (k < N - 1) {
                for (int e = tid; e < 216; e += NT) {
                    const int i = e / 12, c = e - 12 * i;
                    T acc = T(0);
                    for (int a = 0; a <= i; ++a) {
                        const T eac = c < 6 ? (a == c ? T(1) : (a == c + 6 ? dt : T(0))) : s_Gh[18 * (c - 6) + a];
                        acc += s_X[18 * i + a] * eac;
                    }
                    s_Gn[e] = -rho * acc;
                }
            }
            if (k > 0) {
                for (int e = tid; e < 324; e += NT) {
                    const int i = e / 18, j = e - 18 * i;
                    T acc = T(0);
#pragma unroll
                    for (int c = 0; c < 12; ++c) acc += s_X[18 * i + c] * s_G[12 * j + c];
                    fk[FAC_MT + 18 * j + i] = acc;
                }
            }
            __syncthreads();
            // F_k = X' G_k
            if (k < N - 1) {
                for (int e = tid; e < 216; e += NT) {
                    const int i = e / 12, j = e - 12 * i;
                    T acc = T(0);
                    for (int r = i; r < 18; ++r) acc += s_X[18 * r + i] * s_Gn[12 * r + j];
                    fk[FAC_FT + 18 * j + i] = acc;
                }
                for (int e = tid; e < 216; e += NT) s_G[e] = s_Gn[e];   // nobody reads G_{k-1} any more
            }
            __syncthreads();
        }

        I7_ACC(0);
        // ---------------------------------------------------------------- vectors ----------------------
        for (int i = tid; i < nv; i += NT) {
            const int k = i / 18, e = i - 18 * k;
            const bool pad = (k == N - 1 && e >= 12);
            const T* ck = blk + (size_t)k * QB_STRIDE + QB_COST;
            s_q[i] = pad ? T(0) : (e < 6 ? ck[6] * ck[e] : (e < 12 ? ck[7] * xu[i] : ck[8] * xu[i]));
            s_x[i] = pad ? T(0) : P.wx[(size_t)b * nv + i];
            s_yl[i] = (lim && !pad && i >= 12) ? P.wyl[(size_t)b * nv + i] : T(0);
        }
        for (int r = tid; r < nr; r += NT) {
            const int k = r / 12, e = r - 12 * k;
            s_l[r] = k == 0 ? -P.x0[(size_t)b * 12 + e] : (e < 6 ? T(0) : -blk[(size_t)(k - 1) * QB_STRIDE + QB_CV + e - 6]);
            s_y[r] = P.wy[(size_t)b * nr + r];
        }
        __syncthreads();
        // A v for a z-ordered vector v (row r)
        auto a_row = [&](const T* v, int r) -> T {
            const int kk = r / 12, e = r - 12 * kk;
            if (kk == 0) return -v[e];
            const int k = kk - 1;
            const T* vk = v + 18 * k;
            const T* vn = v + 18 * kk;
            if (e < 6) return vk[e] + dt * vk[6 + e] - vn[e];
            const T* bk = blk + (size_t)k * QB_STRIDE;
            T a0 = -vn[e], a1 = T(0), a2 = T(0);
#pragma unroll
            for (int c = 0; c < 6; ++c) {
                a0 += bk[QB_GQ + 6 * (e - 6) + c] * vk[c];
                a1 += bk[QB_GV + 6 * (e - 6) + c] * vk[6 + c];
                a2 += bk[QB_HM + 6 * (e - 6) + c] * vk[12 + c];
            }
            return a0 + a1 + a2;
        };
        // (A' w)[i] for a row-ordered vector w (variable i = 18 k + e, not a padding variable)
        auto at_var = [&](const T* w, int i) -> T {
            const int k = i / 18, e = i - 18 * k;
            T acc = e < 12 ? -w[12 * k + e] : T(0);
            if (k < N - 1) {
                const T* wq = w + 12 * (k + 1);
                const T* wv = wq + 6;
                const T* bk = blk + (size_t)k * QB_STRIDE;
                if (e < 6) acc += wq[e]; else if (e < 12) acc += dt * wq[e - 6];
                T a1 = T(0);
#pragma unroll
                for (int r = 0; r < 6; ++r) a1 += ghat<T>(bk, r, e) * wv[r];
                acc += a1;
            }
            return acc;
        };
        // warm start: z = A x (OSQP), limit rows z_l = x
        for (int r = tid; r < nr; r += NT) s_z[r] = a_row(s_x, r);
        for (int i = tid; i < nv; i += NT) s_zl[i] = s_x[i];
        __syncthreads();

        I7_ACC(1);
        int it = 0, status = 1;
        T rp = T(0), rd = T(0);
        while (it < S.max_iter) {
            ++it;
            // (1) t = rho z - y
            for (int r = tid; r < nr; r += NT) s_t[r] = rho * s_z[r] - s_y[r];
            __syncthreads();
            I7_ACC(2);
            // (2) right-hand side
            for (int i = tid; i < nv; i += NT) {
                const int k = i / 18, e = i - 18 * k;
                T v = T(0);
                if (!(k == N - 1 && e >= 12)) {
                    v = sigma * s_x[i] - s_q[i] + at_var(s_t, i);
                    if (lim && i >= 12) v += rho_l(e) * s_zl[i] - s_yl[i];
                }
                s_b1[i] = v;
            }
            __syncthreads();
            I7_ACC(3);
            // (3) a = L^-1 r
            for (int i = tid; i < nv; i += NT) {
                const int k = i / 18, e = i - 18 * k;
                // full 18-term rows (the stored factor has explicit zeros above the diagonal): 18 independent loads in
                // flight instead of a dependent load-FMA chain of data-dependent length
                const T* Li = fac + (size_t)k * FAC_STRIDE + FAC_LINV + 18 * e;
                const T* rk = s_b1 + 18 * k;
                T lv[18];
#pragma unroll
                for (int c = 0; c < 18; ++c) lv[c] = Li[c];
                T a0 = T(0), a1 = T(0), a2 = T(0);
#pragma unroll
                for (int c = 0; c < 18; c += 3) { a0 += lv[c] * rk[c]; a1 += lv[c + 1] * rk[c + 1]; a2 += lv[c + 2] * rk[c + 2]; }
                s_b2[i] = a0 + (a1 + a2);
            }
            __syncthreads();
            I7_ACC(4);
            // (4) forward chain  w_k = a_k - M_k w_{k-1}   (warp 0; lane i = component i).  The stage matrices are held
            //     in three register buffers used in rotation (static renaming, no moves): the one consumed two stages from
            //     now is being fetched while the current one is applied (an L2 round trip is ~2.5 stages of this chain)
            if (warp == 0) {
                const int li = lane < 18 ? lane : 0;
                T mA[18], mB[18], mC[18];
                T w = s_b2[li];
                auto load = [&](T (&m)[18], int k) {
                    const T* Mk = fac + (size_t)(k < N ? k : N - 1) * FAC_STRIDE + FAC_MT + li;
#pragma unroll
                    for (int j = 0; j < 18; ++j) m[j] = Mk[18 * j];
                };
                auto stage = [&](const T (&m)[18], int k) {
                    __syncwarp();
                    T a0 = s_b2[18 * k + li], a1 = T(0), a2 = T(0);
#pragma unroll
                    for (int j = 0; j < 18; j += 3) {
                        a0 -= m[j] * __shfl_sync(0xffffffffu, w, j);
                        a1 -= m[j + 1] * __shfl_sync(0xffffffffu, w, j + 1);
                        a2 -= m[j + 2] * __shfl_sync(0xffffffffu, w, j + 2);
                    }
                    w = a0 + (a1 + a2);
                    if (lane < 18) s_b2[18 * k + lane] = w;
                };
                load(mA, 1); load(mB, 2);
                for (int k = 1; k < N; k += 3) {
                    load(mC, k + 2); stage(mA, k);
                    if (k + 1 >= N) break;
                    load(mA, k + 3); stage(mB, k + 1);
                    if (k + 2 >= N) break;
                    load(mB, k + 4); stage(mC, k + 2);
                }
            }
            __syncthreads();
            I7_ACC(5);
            // (5) c = L^-T w
            for (int i = tid; i < nv; i += NT) {
                const int k = i / 18, e = i - 18 * k;
                const T* Lk = fac + (size_t)k * FAC_STRIDE + FAC_LINV + e;
                const T* wk = s_b2 + 18 * k;
                T lv[18];
#pragma unroll
                for (int r = 0; r < 18; ++r) lv[r] = Lk[18 * r];
                T a0 = T(0), a1 = T(0), a2 = T(0);
#pragma unroll
                for (int r = 0; r < 18; r += 3) { a0 += lv[r] * wk[r]; a1 += lv[r + 1] * wk[r + 1]; a2 += lv[r + 2] * wk[r + 2]; }
                s_b1[i] = a0 + (a1 + a2);
            }
            __syncthreads();
            I7_ACC(6);
            // (6) backward chain  x_k = c_k - F_k x_{k+1}[:12]   (warp 0), same rotation of three register buffers
            if (warp == 0) {
                const int li = lane < 18 ? lane : 0;
                T fA[12], fB[12], fC[12];
                T xv = s_b1[18 * (N - 1) + li];
                auto load = [&](T (&f)[12], int k) {
                    const T* Fk = fac + (size_t)(k > 0 ? k : 0) * FAC_STRIDE + FAC_FT + li;
#pragma unroll
                    for (int j = 0; j < 12; ++j) f[j] = Fk[18 * j];
                };
                auto stage = [&](const T (&f)[12], int k) {
                    __syncwarp();
                    T a0 = s_b1[18 * k + li], a1 = T(0), a2 = T(0);
#pragma unroll
                    for (int j = 0; j < 12; j += 3) {
                        a0 -= f[j] * __shfl_sync(0xffffffffu, xv, j);
                        a1 -= f[j + 1] * __shfl_sync(0xffffffffu, xv, j + 1);
                        a2 -= f[j + 2] * __shfl_sync(0xffffffffu, xv, j + 2);
                    }
                    xv = a0 + (a1 + a2);
                    if (lane < 18) s_b1[18 * k + lane] = xv;
                };
                load(fA, N - 2); load(fB, N - 3);
                for (int k = N - 2; k >= 0; k -= 3) {
                    load(fC, k - 2); stage(fA, k);
                    if (k - 1 < 0) break;
                    load(fA, k - 3); stage(fB, k - 1);
                    if (k - 2 < 0) break;
                    load(fB, k - 4); stage(fC, k - 2);
                }
            }
            __syncthreads();
            I7_ACC(7);
            // (7) z~ = A x~ and the updates of the equality rows; (8) x and the limit rows
            for (int r = tid; r < nr; r += NT) {
                const T zt = a_row(s_b1, r);
                const T zr = al * zt + (T(1) - al) * s_z[r];
                s_y[r] += rho * (zr - s_l[r]);
                s_z[r] = s_l[r];
            }
            for (int i = tid; i < nv; i += NT) {
                const int k = i / 18, e = i - 18 * k;
                if (k == N - 1 && e >= 12) continue;
                const T xt = s_b1[i];
                if (lim && i >= 12) {
                    const T rl = rho_l(e);
                    const T zr = al * xt + (T(1) - al) * s_zl[i];
                    T zn = zr + s_yl[i] / rl;
                    const T lo = lo_of(e), hi = hi_of(e);
                    zn = zn < lo ? lo : (zn > hi ? hi : zn);
                    s_yl[i] += rl * (zr - zn);
                    s_zl[i] = zn;
                }
                s_x[i] = al * xt + (T(1) - al) * s_x[i];
            }
            __syncthreads();
            I7_ACC(8);
            // (9) termination check (OSQP: every `check` iterations)
            if (it % S.check == 0) {
                T m_rp = T(0), m_ax = T(0), m_z = T(0), m_rd = T(0), m_px = T(0), m_aty = T(0), m_q = T(0);
                for (int r = tid; r < nr; r += NT) {
                    const T ax = a_row(s_x, r);
                    m_rp = fmax(m_rp, fabs(ax - s_l[r])); m_ax = fmax(m_ax, fabs(ax)); m_z = fmax(m_z, fabs(s_l[r]));
                }
                for (int i = tid; i < nv; i += NT) {
                    const int k = i / 18, e = i - 18 * k;
                    if (k == N - 1 && e >= 12) continue;
                    const T* ck = blk + (size_t)k * QB_STRIDE + QB_COST;
                    T px;
                    if (e < 6) {
                        T jx = T(0);
#pragma unroll
                        for (int c = 0; c < 6; ++c) jx += ck[c] * s_x[18 * k + c];
                        px = ck[6] * ck[e] * jx;
                    } else px = (e < 12 ? ck[7] : ck[8]) * s_x[i];
                    T aty = at_var(s_y, i);
                    if (lim && i >= 12) {
                        aty += s_yl[i];
                        m_rp = fmax(m_rp, fabs(s_x[i] - s_zl[i])); m_ax = fmax(m_ax, fabs(s_x[i])); m_z = fmax(m_z, fabs(s_zl[i]));
                    }
                    m_rd = fmax(m_rd, fabs(px + s_q[i] + aty));
                    m_px = fmax(m_px, fabs(px)); m_aty = fmax(m_aty, fabs(aty)); m_q = fmax(m_q, fabs(s_q[i]));
                }
                T vals[7] = {m_rp, m_ax, m_z, m_rd, m_px, m_aty, m_q};
#pragma unroll
                for (int q = 0; q < 7; ++q) {
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) vals[q] = fmax(vals[q], __shfl_xor_sync(0xffffffffu, vals[q], o));
                    if (lane == 0) s_red[8 * warp + q] = vals[q];
                }
                __syncthreads();
                if (tid == 0) {
                    T v[7];
                    for (int q = 0; q < 7; ++q) { v[q] = T(0); for (int w = 0; w < NT / 32; ++w) v[q] = fmax(v[q], s_red[8 * w + q]); }
                    const T ep = T(S.eps_abs) + T(S.eps_rel) * fmax(v[1], v[2]);
                    const T ed = T(S.eps_abs) + T(S.eps_rel) * fmax(fmax(v[4], v[5]), v[6]);
                    s_red[56] = v[0]; s_red[57] = v[3];
                    s_stop = (v[0] <= ep && v[3] <= ed) ? 1 : 0;
                }
                __syncthreads();
                rp = s_red[56]; rd = s_red[57];
                const int stop = s_stop;
                __syncthreads();
                I7_ACC(9);
                if (stop) { status = 0; break; }
            }
        }
        if (prof && tid == 0) for (int i = 0; i < 12; ++i) prof[i] = pacc[i];
#undef I7_ACC
        // ---------------------------------------------------------------- results ----------------------
        for (int i = tid; i < nv; i += NT) {
            P.wx[(size_t)b * nv + i] = s_x[i];
            if (lim) P.wyl[(size_t)b * nv + i] = s_yl[i];
            if (i < nz) P.z_out[(size_t)b * nz + i] = s_x[i];
        }
        for (int r = tid; r < nr; r += NT) {
            P.wy[(size_t)b * nr + r] = s_y[r];
            if (P.y_out) P.y_out[(size_t)b * nr + r] = s_y[r];
        }
        if (tid == 0 && P.info) {
            T* o = P.info + (size_t)b * 4;
            o[0] = T(it); o[1] = rp; o[2] = rd; o[3] = T(status);
        }
        __syncthreads();
    }
}

}  // namespace i7
