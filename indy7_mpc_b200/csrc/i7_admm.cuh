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
//            (w_k = a_k - M_k w_{k-1} forward, x_k = c_k - F_k x_{k+1} backward) run by one warp, the stage matrices
//            staged through a shared-memory ring by the other warps (see ADMM_CG below).
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

constexpr int ADMM_NT = 128;     // two CTAs per SM (registers: the in-register 18 x 18 factorisation; a cap for three spills, measured 1.9x slower)
// factor workspace per knot: M (18 x 18, FAC_MT) and F (18 x 12, FAC_FT) in planes of column PAIRS — element (i, c) at
// ((c / 2) * 18 + i) * 2 + c % 2 — so that lane i of the serial chains reads its row as 16-byte vectors (one warp's
// instruction count bounds a stage) AND the 18 lanes of a load touch 288 contiguous bytes (coalesced).  L^-1 is kept
// twice, as planes of COLUMN pairs (FAC_LA, same formula, read by a = L^-1 r) and of ROW pairs (FAC_LB: (r, i) at
// ((r / 2) * 18 + i) * 2 + r % 2, read by c = L^-T w): either product is nine coalesced 16-byte loads
constexpr int FAC_LA = 0, FAC_LB = 324, FAC_MT = 648, FAC_FT = 972, FAC_STRIDE = 1188;
// serial chains: the stage matrices are staged through shared memory in groups of ADMM_CG stages, double-buffered: while
// warp 0 works through group g the other warps (idle during the chain) copy group g + 1 from the L2 workspace.  Measured
// alternatives: rows prefetched from L2 into registers two stages ahead (warp 0 waits for them: 285 k / 216 k cycles per
// QP for the forward / backward chain against 238 k / 209 k), a three-slot ring of three stages filled two groups ahead with
// cp.async (286 k / 228 k: the larger shared-memory footprint costs more than the deeper prefetch gains).
constexpr int ADMM_CG = 4, ADMM_SLOTS = 2;
// shared memory: the vectors (156 N), the stage scratch of the factorisation, the ring of the chains and — while it fits —
// the QP blocks (QB_STRIDE N), which every iteration reads twice (A x~ and A' t)
constexpr int ADMM_BLOCKS_IN_SMEM_MAX_N = 64;
__host__ __device__ inline int admm_smem_len(int N, bool blocks) {
    return 156 * N + 2 * 324 + 2 * 216 + 108 + 16 + 64 + ADMM_SLOTS * ADMM_CG * 324 + (blocks ? QB_STRIDE * N : 0);
}

// element (r, c) of G^ = [Gq Gv Hm] (6 x 18) in a QB block
template <typename T> I7_DEV T ghat(const T* blk, int r, int c) {
    return c < 6 ? blk[QB_GQ + 6 * r + c] : (c < 12 ? blk[QB_GV + 6 * r + c - 6] : blk[QB_HM + 6 * r + c - 12]);
}

template <typename T>
__global__ void __launch_bounds__(ADMM_NT, 2) admm_kernel(const __grid_constant__ AdmmParams<T> P) {
    constexpr int NT = ADMM_NT;
    static_assert(sizeof(T) == 8, "the ADMM kernel is fp64 only (16-byte vector loads of pairs of doubles)");
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
    T* s_ring = s_red + 64;      // ADMM_SLOTS x ADMM_CG x 324  stage matrices of the serial chains, double-buffered by group
    T* s_blk = s_ring + ADMM_SLOTS * ADMM_CG * 324;   // QB_STRIDE N  the QP blocks (short horizons only)
    const bool blk_smem = N <= ADMM_BLOCKS_IN_SMEM_MAX_N;
    __shared__ int s_stop;
    __shared__ unsigned short s_tri[171];       // (row << 8 | column) of the lower-triangular entries of an 18 x 18 block
    for (int e = tid; e < 324; e += NT) {
        const int a = e / 18, c = e - 18 * a;
        if (c <= a) s_tri[(a * (a + 1)) / 2 + c] = (unsigned short)((a << 8) | c);
    }
    __syncthreads();

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
        long long pacc[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, ptt = prof ? clock64() : 0, pt2 = 0;
#define I7_SUB(i) do { if (prof) { const long long n_ = clock64(); pacc[i] += n_ - pt2; pt2 = n_; } } while (0)
// every thread keeps the counters (a branch on tid == 0 here would leave warp 0 diverged in front of its shuffle chains)
#define I7_ACC(i) do { if (prof) { const long long n_ = clock64(); pacc[i] += n_ - ptt; ptt = n_; } } while (0)
        // ---------------------------------------------------------------- factorisation ----------------
        for (int k = 0; k < N; ++k) {
            const T* bk = blk + (size_t)k * QB_STRIDE;
            if (prof) pt2 = clock64();
            if (tid < 108) s_Gh[tid] = k < N - 1 ? ghat<T>(bk, tid / 18, tid % 18) : T(0);
            if (tid >= 112 && tid < 112 + 9) s_ck[tid - 112] = bk[QB_COST + tid - 112];
            __syncthreads();
            // pivot block D = H_kk - G_{k-1}' G_{k-1} (x x part), lower triangle
            for (int e0 = tid; e0 < 171; e0 += NT) {
                const int a = s_tri[e0] >> 8, c = s_tri[e0] & 255;         // lower triangle only: 171 tasks, not 324
                const int e = 18 * a + c;
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
            I7_SUB(12);
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
            I7_SUB(13);
            // G_k = X E_k with E_k = -rho [A_k B_k]' ; M_k = X[:, :12] G_{k-1}' ; L^-1 to the workspace
            T* fk = fac + (size_t)k * FAC_STRIDE;
            for (int e = tid; e < 324; e += NT) {
                const int i = e / 18, c = e - 18 * i;               // s_X[e] = L^-1 (i, c)
                fk[FAC_LA + ((c >> 1) * 18 + i) * 2 + (c & 1)] = s_X[e];
                fk[FAC_LB + ((i >> 1) * 18 + c) * 2 + (i & 1)] = s_X[e];
            }
            if (k < N - 1) {
                for (int e = tid; e < 216; e += NT) {
                    const int i = e / 12, c = e - 12 * i;
                    // rows c < 6 of [A_k B_k] are [I, dt I, 0]; the others are G^; L^-1 carries explicit zeros above its
                    // diagonal, so the dot products run over all 18 terms with every load independent of the others
                    T acc;
                    if (c < 6) acc = s_X[18 * i + c] + dt * s_X[18 * i + c + 6];
                    else {
                        T a0 = T(0), a1 = T(0), a2 = T(0);
#pragma unroll
                        for (int a = 0; a < 18; a += 3) {
                            a0 += s_X[18 * i + a] * s_Gh[18 * (c - 6) + a];
                            a1 += s_X[18 * i + a + 1] * s_Gh[18 * (c - 6) + a + 1];
                            a2 += s_X[18 * i + a + 2] * s_Gh[18 * (c - 6) + a + 2];
                        }
                        acc = a0 + (a1 + a2);
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
                    fk[FAC_MT + ((j >> 1) * 18 + i) * 2 + (j & 1)] = acc;
                }
            }
            __syncthreads();
            I7_SUB(14);
            // F_k = X' G_k
            if (k < N - 1) {
                for (int e = tid; e < 216; e += NT) {
                    const int i = e / 12, j = e - 12 * i;
                    T b0 = T(0), b1 = T(0), b2 = T(0);
#pragma unroll
                    for (int r = 0; r < 18; r += 3) {
                        b0 += s_X[18 * r + i] * s_Gn[12 * r + j];
                        b1 += s_X[18 * (r + 1) + i] * s_Gn[12 * (r + 1) + j];
                        b2 += s_X[18 * (r + 2) + i] * s_Gn[12 * (r + 2) + j];
                    }
                    const T acc = b0 + (b1 + b2);
                    fk[FAC_FT + ((j >> 1) * 18 + i) * 2 + (j & 1)] = acc;
                }
                for (int e = tid; e < 216; e += NT) s_G[e] = s_Gn[e];   // nobody reads G_{k-1} any more
            }
            __syncthreads();
            I7_SUB(15);
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
                const double2* Li = reinterpret_cast<const double2*>(fac + (size_t)k * FAC_STRIDE + FAC_LA) + e;
                const double2* rk = reinterpret_cast<const double2*>(s_b1 + 18 * k);
                double2 lv[9];
#pragma unroll
                for (int c = 0; c < 9; ++c) lv[c] = Li[18 * c];
                T a0 = T(0), a1 = T(0), a2 = T(0);
#pragma unroll
                for (int c = 0; c < 9; c += 3) {
                    const double2 r0 = rk[c], r1 = rk[c + 1], r2 = rk[c + 2];
                    a0 += lv[c].x * r0.x; a0 += lv[c].y * r0.y;
                    a1 += lv[c + 1].x * r1.x; a1 += lv[c + 1].y * r1.y;
                    a2 += lv[c + 2].x * r2.x; a2 += lv[c + 2].y * r2.y;
                }
                s_b2[i] = a0 + (a1 + a2);
            }
            __syncthreads();
            I7_ACC(4);
            // (4) forward chain  w_k = a_k - M_k w_{k-1}   (warp 0; lane i = component i).  One warp's instruction count
            //     bounds a stage: a lane reads its row of M_k and w_{k-1} as 16-byte vectors from SHARED memory.  The
            //     matrices get there in groups of ADMM_CG stages: while warp 0 works through group g, the other warps copy
            //     group g + 1 into the other half of the ring — every load of a filler thread issued before its first store
            //     (one L2 round trip per group, not one per element)
            {
                const int nst = N - 1, ngrp = (nst + ADMM_CG - 1) / ADMM_CG;
                auto fill = [&](int g, int t0, int nthr) {
                    const int k0 = 1 + g * ADMM_CG;
                    const int ns = N - k0 < ADMM_CG ? N - k0 : ADMM_CG;
                    double2* dst = reinterpret_cast<double2*>(s_ring + (g & 1) * ADMM_CG * 324);
                    constexpr int R = (ADMM_CG * 162 + (NT - 32) - 1) / (NT - 32);
                    double2 v[R];
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        const int e = (tid - t0) + r * nthr;
                        if (e < ns * 162) {
                            const int st = e / 162, q = e - 162 * st;
                            v[r] = reinterpret_cast<const double2*>(fac + (size_t)(k0 + st) * FAC_STRIDE + FAC_MT)[q];
                        }
                    }
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        const int e = (tid - t0) + r * nthr;
                        if (e < ns * 162) dst[e] = v[r];
                    }
                };
                fill(0, 0, NT);
                __syncthreads();
                for (int g = 0; g < ngrp; ++g) {
                    if (warp == 0) {
                        const int li = lane < 18 ? lane : 0;
                        const int k0 = 1 + g * ADMM_CG;
                        for (int st = 0; st < ADMM_CG && k0 + st < N; ++st) {
                            const int k = k0 + st;
                            const double2* Mr = reinterpret_cast<const double2*>(s_ring + (g & 1) * ADMM_CG * 324 + 324 * st) + li;
                            __syncwarp();                         // w_{k-1} written by the lanes below is visible
                            const double2* wp = reinterpret_cast<const double2*>(s_b2 + 18 * (k - 1));
                            T a0 = s_b2[18 * k + li], a1 = T(0), a2 = T(0);
#pragma unroll
                            for (int j = 0; j < 9; j += 3) {
                                const double2 m0 = Mr[18 * j], m1 = Mr[18 * (j + 1)], m2 = Mr[18 * (j + 2)];
                                const double2 w0 = wp[j], w1 = wp[j + 1], w2 = wp[j + 2];
                                a0 -= m0.x * w0.x; a0 -= m0.y * w0.y;
                                a1 -= m1.x * w1.x; a1 -= m1.y * w1.y;
                                a2 -= m2.x * w2.x; a2 -= m2.y * w2.y;
                            }
                            if (lane < 18) s_b2[18 * k + lane] = a0 + (a1 + a2);
                        }
                    } else if (g + 1 < ngrp) {
                        fill(g + 1, 32, NT - 32);
                    }
                    __syncthreads();
                }
            }
            __syncthreads();
            I7_ACC(5);
            // (5) c = L^-T w
            for (int i = tid; i < nv; i += NT) {
                const int k = i / 18, e = i - 18 * k;
                const double2* Lk = reinterpret_cast<const double2*>(fac + (size_t)k * FAC_STRIDE + FAC_LB) + e;
                const double2* wk = reinterpret_cast<const double2*>(s_b2 + 18 * k);
                double2 lv[9];
#pragma unroll
                for (int r = 0; r < 9; ++r) lv[r] = Lk[18 * r];
                T a0 = T(0), a1 = T(0), a2 = T(0);
#pragma unroll
                for (int r = 0; r < 9; r += 3) {
                    const double2 w0 = wk[r], w1 = wk[r + 1], w2 = wk[r + 2];
                    a0 += lv[r].x * w0.x; a0 += lv[r].y * w0.y;
                    a1 += lv[r + 1].x * w1.x; a1 += lv[r + 1].y * w1.y;
                    a2 += lv[r + 2].x * w2.x; a2 += lv[r + 2].y * w2.y;
                }
                s_b1[i] = a0 + (a1 + a2);
            }
            __syncthreads();
            I7_ACC(6);
            // (6) backward chain  x_k = c_k - F_k x_{k+1}[:12]   (warp 0), same scheme: groups of stages from the top
            {
                const int nst = N - 1, ngrp = (nst + ADMM_CG - 1) / ADMM_CG;
                auto fill = [&](int g, int t0, int nthr) {
                    const int k0 = N - 2 - g * ADMM_CG;             // stages k0, k0 - 1, ...
                    const int ns = k0 + 1 < ADMM_CG ? k0 + 1 : ADMM_CG;
                    double2* dst = reinterpret_cast<double2*>(s_ring + (g & 1) * ADMM_CG * 324);
                    constexpr int R = (ADMM_CG * 108 + (NT - 32) - 1) / (NT - 32);
                    double2 v[R];
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        const int e = (tid - t0) + r * nthr;
                        if (e < ns * 108) {
                            const int st = e / 108, q = e - 108 * st;
                            v[r] = reinterpret_cast<const double2*>(fac + (size_t)(k0 - st) * FAC_STRIDE + FAC_FT)[q];
                        }
                    }
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        const int e = (tid - t0) + r * nthr;
                        if (e < ns * 108) dst[e] = v[r];
                    }
                };
                fill(0, 0, NT);
                __syncthreads();
                for (int g = 0; g < ngrp; ++g) {
                    if (warp == 0) {
                        const int li = lane < 18 ? lane : 0;
                        const int k0 = N - 2 - g * ADMM_CG;
                        for (int st = 0; st < ADMM_CG && k0 - st >= 0; ++st) {
                            const int k = k0 - st;
                            const double2* Fr = reinterpret_cast<const double2*>(s_ring + (g & 1) * ADMM_CG * 324 + 216 * st) + li;
                            __syncwarp();
                            const double2* xp = reinterpret_cast<const double2*>(s_b1 + 18 * (k + 1));
                            T a0 = s_b1[18 * k + li], a1 = T(0), a2 = T(0);
#pragma unroll
                            for (int j = 0; j < 6; j += 3) {
                                const double2 f0 = Fr[18 * j], f1 = Fr[18 * (j + 1)], f2 = Fr[18 * (j + 2)];
                                const double2 x0 = xp[j], x1 = xp[j + 1], x2 = xp[j + 2];
                                a0 -= f0.x * x0.x; a0 -= f0.y * x0.y;
                                a1 -= f1.x * x1.x; a1 -= f1.y * x1.y;
                                a2 -= f2.x * x2.x; a2 -= f2.y * x2.y;
                            }
                            if (lane < 18) s_b1[18 * k + lane] = a0 + (a1 + a2);
                        }
                    } else if (g + 1 < ngrp) {
                        fill(g + 1, 32, NT - 32);
                    }
                    __syncthreads();
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
        if (prof && tid == 0) for (int i = 0; i < 16; ++i) prof[i] = pacc[i];
#undef I7_ACC
#undef I7_SUB
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
