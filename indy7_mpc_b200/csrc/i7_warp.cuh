// i7_warp.cuh — throughput form of the fused SQP-MPC solve: ONE WARP per independent problem, lane = knot.
//
// The CTA-per-problem kernel (i7_sqp.cuh) spreads (knot, column) tasks over 128..384 threads and pays a block
// barrier between every sub-phase; with thousands of independent problems that parallelism is not needed.
// Here every lane owns one knot (knots lane, lane+32, ... for N > 32) and runs that knot's whole linearisation
// serially in registers — no barrier, no shared-memory traffic, no redundant tasks — and a warp needs so little
// shared memory (~24 KB fp64 at N=32) that 8+ problems are resident per SM, which is what keeps the FP64 pipe fed
// while other warps sit in their latency-bound backward recursion.
//
//   setup     lane = knot: FK/cost terms, M (CRBA columns), bias, LDL', qdd, 12 analytic dID columns -> G, and the
//             stage-local products S, SG, d, L, hx; the per-knot record goes to a global (L2-resident) scratch
//   riccati   the 32 lanes cooperate on one stage at a time; the record of stage k-1/k-2 is prefetched
//             global -> shared with cp.async (3-deep ring) while stage k is eliminated; gains go back to the scratch
//   roll-out  forward over the gains (prefetched the same way), then controls with lane = knot
//   merit     lane = knot, candidates evaluated in the reference's order with early exit (src/osqp_sqp.py:59-72)
#pragma once
#include "i7_sqp.cuh"

namespace i7 {

// per-knot record in the global scratch; the part the backward sweep needs comes first (multiple of 16 bytes)
constexpr int WR_L = 0;       // 78  upper triangle of G'SG
constexpr int WR_SG = 78;     // 72  S G
constexpr int WR_S = 150;     // 36  S = R_k Y'Y   (MODE_DUMP: H = dt M^-1)
constexpr int WR_D = 186;     // 6   d = S cv - R Y u
constexpr int WR_HX = 192;    // 12  gx + G'd
constexpr int WR_RIC = 204;   //     elements staged for the backward sweep
constexpr int WR_GQ = 204;    // 36
constexpr int WR_GV = 240;    // 36
constexpr int WR_CV = 276;    // 6
constexpr int WR_Y = 282;     // 36  Y = M / dt
constexpr int WR_STRIDE = 320;
constexpr int WG_STRIDE = 80; // gains of one stage: Kw (6 x 12 row-major), kw (6), 2 pad
constexpr int WSTG = 208;     // one slot of the staging ring (>= WR_RIC, >= WG_STRIDE, multiple of 2)

struct WarpRecLayout {        // offsets for dump_qp
    static constexpr int GQ = WR_GQ, GV = WR_GV, CV = WR_CV, S = WR_S, STRIDE = WR_STRIDE;
};

struct WLayout { int xu, z, goals, x0, cost, work, stage, total; };
__host__ __device__ inline WLayout make_wlayout(int N) {
    WLayout L;
    int o = 0;
    L.xu = o; o += 18 * N;
    L.z = o; o += 18 * N;
    L.goals = o; o += 3 * N + (N & 1);
    L.x0 = o; o += 12;
    L.cost = o; o += COST_STRIDE * N + (N & 1);
    L.work = o; o += WK_SIZE;
    L.stage = o; o += 3 * WSTG;
    L.total = o;
    return L;
}
__host__ __device__ inline size_t warp_scratch_elems(int N) { return (size_t)N * (WR_STRIDE + WG_STRIDE); }

template <typename T>
struct WPtrs {
    T *xu, *z, *goals, *x0, *cost, *work, *stage;
    T *rec, *gains;           // global scratch of this CTA
};

I7_DEV void cp_async16(void* smem_dst, const void* gsrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gsrc) : "memory");
}
I7_DEV void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int NKEEP> I7_DEV void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(NKEEP) : "memory"); }

// copy `nelem` elements (a multiple of 16 bytes) global -> shared, all 32 lanes
template <typename T> I7_DEV void stage_copy(T* dst, const T* src, int nelem, int lane) {
    constexpr int PER = 16 / (int)sizeof(T);
    const int nch = nelem / PER;
    for (int c = lane; c < nch; c += 32) cp_async16(dst + PER * c, src + PER * c);
}

template <typename T> I7_DEV T warp_sum(T a) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    return a;
}

// -----------------------------------------------------------------------------------------------------------
// setup of ONE knot by ONE lane (src/osqp_solver.py:70-135): cost terms into shared memory, the dynamics record
// into the global scratch.
template <typename T>
__device__ __noinline__ void wsetup_knot(const Robot<T>& rob, const SolveParams<T>& P, const T* xk, const T* gk,
                                         const T* fext, int k, bool for_dump, T* ck, T* rec) {
    const int N = P.N;
    const T dt = P.dt;
    T cs[12];
#pragma unroll
    for (int i = 0; i < 6; ++i) sincos_t<T>(xk[i], &cs[6 + i], &cs[i]);
    T Rk;
    {   // end-effector error, j = J'e and the (regularised) weights of this knot
        T pee[3], zs[18], ps[18];
        fk_ee<T>(rob, cs, pee, zs, ps);
        const T e[3] = {pee[0] - gk[0], pee[1] - gk[1], pee[2] - gk[2]};
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            const T d[3] = {pee[0] - ps[3 * i], pee[1] - ps[3 * i + 1], pee[2] - ps[3 * i + 2]};
            T c[3];
            cross3(zs + 3 * i, d, c);
            ck[i] = e[0] * c[0] + e[1] * c[1] + e[2] * c[2];
        }
        const T nrm = sqrt_t<T>(e[0] * e[0] + e[1] * e[1] + e[2] * e[2]);
        const T sc = P.regularize ? T(1) / (nrm + P.eps) : T(1);
        ck[6] = (k == N - 1) ? P.QN : T(1);
        ck[7] = P.dQ * sc;
        Rk = P.R * sc;
        ck[8] = Rk;
    }
    if (k >= N - 1) return;
    T qd[6], u[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) { qd[i] = xk[6 + i]; u[i] = xk[12 + i]; }
    // mass matrix (lower triangle kept twice: Ms stays, A is factored in place), bias, forward dynamics
    T A[6][6], Ms[6][6], dinv[6], qdd[6];
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        T out[6];
        crba_column<T>(rob, cs, j, out);
#pragma unroll
        for (int r = j; r < 6; ++r) { A[r][j] = out[r]; Ms[r][j] = out[r]; Ms[j][r] = out[r]; }
    }
    {
        T bias[6];
        rnea_bias<T>(rob, cs, qd, fext, bias);
#pragma unroll
        for (int r = 0; r < 6; ++r) qdd[r] = u[r] - bias[r];
    }
    spd6_factor<T>(A, dinv);
    spd6_solve_factored<T>(A, dinv, qdd);
    // G = [dt dqdd/dq | I + dt dqdd/dv] = -dt M^-1 dID (+ I), one column at a time
    T G[72];
#pragma unroll 1
    for (int j = 0; j < 6; ++j) {
        T y[6];
        rnea_dcolumn<T, true>(rob, cs, qd, qdd, fext, j, y);
        spd6_solve_factored<T>(A, dinv, y);
#pragma unroll
        for (int r = 0; r < 6; ++r) { const T g = -dt * y[r]; G[12 * r + j] = g; rec[WR_GQ + 6 * r + j] = g; }
    }
#pragma unroll 1
    for (int j = 0; j < 6; ++j) {
        T y[6];
        rnea_dcolumn<T, false>(rob, cs, qd, qdd, nullptr, j, y);
        spd6_solve_factored<T>(A, dinv, y);
#pragma unroll
        for (int r = 0; r < 6; ++r) {
            const T g = -dt * y[r] + (r == j ? T(1) : T(0));
            G[12 * r + 6 + j] = g; rec[WR_GV + 6 * r + j] = g;
        }
    }
    // cv = v + dt qdd - Gq q - Gv v - dt M^-1 u
    T cv[6];
    {
        T mu[6];
#pragma unroll
        for (int r = 0; r < 6; ++r) mu[r] = u[r];
        spd6_solve_factored<T>(A, dinv, mu);
#pragma unroll
        for (int r = 0; r < 6; ++r) {
            T a0 = qd[r] + dt * qdd[r], a1 = T(0);
#pragma unroll
            for (int s = 0; s < 6; ++s) { a0 -= G[12 * r + s] * xk[s]; a1 -= G[12 * r + 6 + s] * qd[s]; }
            cv[r] = a0 + a1 - dt * mu[r];
            rec[WR_CV + r] = cv[r];
        }
    }
    if (for_dump) {                   // B_k = [0; dt M^-1] wanted explicitly
#pragma unroll 1
        for (int c = 0; c < 6; ++c) {
            T b[6];
#pragma unroll
            for (int r = 0; r < 6; ++r) b[r] = (r == c) ? T(1) : T(0);
            spd6_solve_factored<T>(A, dinv, b);
#pragma unroll
            for (int r = 0; r < 6; ++r) rec[WR_S + 6 * r + c] = dt * b[r];
        }
        return;
    }
    // Y = M / dt,  S = R_k Y'Y (symmetric),  d = S cv - R_k Y u
    T S[6][6], d[6];
    {
        const T coef = Rk / (dt * dt);
#pragma unroll
        for (int r = 0; r < 6; ++r)
#pragma unroll
            for (int c = 0; c <= r; ++c) {
                T acc = T(0);
#pragma unroll
                for (int s = 0; s < 6; ++s) acc += Ms[s][r] * Ms[s][c];
                S[r][c] = coef * acc; S[c][r] = S[r][c];
            }
#pragma unroll
        for (int r = 0; r < 6; ++r) {
            T a0 = T(0), a1 = T(0);
#pragma unroll
            for (int s = 0; s < 6; ++s) { a0 += S[r][s] * cv[s]; a1 += Ms[r][s] * u[s]; }
            d[r] = a0 - (Rk / dt) * a1;
            rec[WR_D + r] = d[r];
#pragma unroll
            for (int c = 0; c < 6; ++c) { rec[WR_S + 6 * r + c] = S[r][c]; rec[WR_Y + 6 * r + c] = Ms[r][c] / dt; }
        }
    }
    // SG = S G (column by column, kept in SGl), hx = gx + G'd, L = G'(SG) upper triangle
    T SGl[72];
#pragma unroll 1
    for (int c = 0; c < 12; ++c) {
        T g[6];
#pragma unroll
        for (int s = 0; s < 6; ++s) g[s] = G[12 * s + c];
        T hx = c < 6 ? ck[6] * ck[c] : ck[7] * xk[c];
#pragma unroll
        for (int r = 0; r < 6; ++r) {
            T acc = T(0);
#pragma unroll
            for (int s = 0; s < 6; ++s) acc += S[r][s] * g[s];
            SGl[12 * r + c] = acc; rec[WR_SG + 12 * r + c] = acc;
            hx += g[r] * d[r];
        }
        rec[WR_HX + c] = hx;
    }
    {
        int e = 0;
#pragma unroll 1
        for (int r = 0; r < 12; ++r) {
            T g[6];
#pragma unroll
            for (int s = 0; s < 6; ++s) g[s] = G[12 * s + r];
#pragma unroll 1
            for (int c = r; c < 12; ++c) {
                T acc = T(0);
#pragma unroll
                for (int s = 0; s < 6; ++s) acc += g[s] * SGl[12 * s + c];
                rec[WR_L + e] = acc; ++e;
            }
        }
    }
}

// -----------------------------------------------------------------------------------------------------------
// Backward block recursion + forward roll-out by one warp (same algebra as phase_riccati in i7_sqp.cuh).
// Lane tasks of the value update: lanes own up to three entries of the upper triangle of the 12x12 value matrix
// (t = lane, lane+32, lane+64 < 78); lanes 14..25 additionally own one entry of the value gradient.
template <typename T>
__device__ void wriccati(const SolveParams<T>& P, const WPtrs<T>& M, int lane) {
    const int N = P.N;
    const T dt = P.dt;
    T* wk = M.work;
    const T* xu = M.xu;
    int tr[3], tc[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const int t = lane + 32 * i;
        tr[i] = t < 78 ? c_tri_r[t] : 0;
        tc[i] = t < 78 ? c_tri_c[t] : 0;
    }
    const int nown = lane < 14 ? 3 : 2;
    const int pr = lane - 14;                       // value-gradient entry of this lane (0..11) or out of range
    // prologue of the record ring: records N-2 and N-3
    stage_copy<T>(M.stage + WSTG * ((N - 2) % 3), M.rec + (size_t)WR_STRIDE * (N - 2), WR_RIC, lane);
    cp_async_commit();
    if (N >= 3) stage_copy<T>(M.stage + WSTG * ((N - 3) % 3), M.rec + (size_t)WR_STRIDE * (N - 3), WR_RIC, lane);
    cp_async_commit();
    cp_async_wait<1>();
    __syncwarp();
    int cur = 0;
    {   // terminal value function: P_N = blkdiag(QN j j', dQ_N I), p_N = [QN j; dQ_N v]; aug of stage N-2
        const T* cN = M.cost + COST_STRIDE * (N - 1);
        const T* xN = xu + 18 * (N - 1);
        const T* kr = M.stage + WSTG * ((N - 2) % 3);
        T* aug = wk + WK_AUG;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            if (i < nown) {
                const int r = tr[i], c = tc[i];
                T v = T(0);
                if (c < 6) v = cN[6] * cN[r] * cN[c];
                else if (r == c) v = cN[7];
                wk[WK_P + 12 * r + c] = v;
                wk[WK_P + 12 * c + r] = v;
                if (r < 6 && c >= 6) {
                    const int a = c - 6;
                    aug[AUGW * a + 6 + r] = v - kr[WR_SG + 12 * a + r];
                    aug[AUGW * a + 12 + r] = dt * v - kr[WR_SG + 12 * a + 6 + r];
                } else if (r >= 6) {
                    const int a = r - 6, b = c - 6;
                    aug[AUGW * a + b] = v + kr[WR_S + 6 * a + b];
                    if (a != b) aug[AUGW * b + a] = v + kr[WR_S + 6 * b + a];
                }
            }
        }
        if (pr >= 0 && pr < 12) {
            const T v = pr < 6 ? cN[6] * cN[pr] : cN[7] * xN[pr];
            wk[WK_PV + pr] = v;
            if (pr >= 6) aug[AUGW * (pr - 6) + 18] = v - kr[WR_D + pr - 6];
        }
    }
    __syncwarp();
    for (int k = N - 2; k >= 0; --k) {
        if (k >= 2) stage_copy<T>(M.stage + WSTG * ((k - 2) % 3), M.rec + (size_t)WR_STRIDE * (k - 2), WR_RIC, lane);
        cp_async_commit();
        const T* aug = wk + WK_AUG + 114 * cur;
        T* X = wk + WK_X;
        {   // X = Hww^-1 [Hwx | hw]: lane t < 13 solves right-hand side t (the factorisation is redundant per lane)
            const int t = lane < 13 ? lane : 12;
            T A[6][6], dinv[6], b[6];
#pragma unroll
            for (int r = 0; r < 6; ++r)
#pragma unroll
                for (int q = 0; q <= r; ++q) A[r][q] = aug[AUGW * r + q];
#pragma unroll
            for (int r = 0; r < 6; ++r) b[r] = aug[AUGW * r + 6 + t];
            spd6_factor<T>(A, dinv);
            spd6_solve_factored<T>(A, dinv, b);
            if (lane < 13) {
                T* gk = M.gains + (size_t)WG_STRIDE * k;
#pragma unroll
                for (int r = 0; r < 6; ++r) {
                    X[13 * r + t] = b[r];
                    if (t < 12) gk[12 * r + t] = -b[r]; else gk[72 + r] = -b[r];      // w = Kw x + kw
                }
            }
        }
        cp_async_wait<1>();                         // record k-1 has landed (record k landed a stage earlier)
        __syncwarp();
        {   // P = Qx + L + T'PT - Hwx'X,  p = hx + [pq; dt pq] - Hwx' x13;  aug of stage k-1
            const T* Pm = wk + WK_P + 144 * cur;
            const T* pv = wk + WK_PV + 12 * cur;
            T* Pn = wk + WK_P + 144 * (cur ^ 1);
            T* pn = wk + WK_PV + 12 * (cur ^ 1);
            T* augn = wk + WK_AUG + 114 * (cur ^ 1);
            const T* ck = M.cost + COST_STRIDE * k;
            const T* kr = M.stage + WSTG * (k % 3);
            const T* krn = M.stage + WSTG * ((k + 2) % 3);        // == (k - 1) % 3 for k >= 1
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                if (i < nown) {
                    const int r = tr[i], c = tc[i];
                    const int ra = r < 6 ? r : r - 6, ca = c < 6 ? c : c - 6;
                    T acc = kr[WR_L + lane + 32 * i] + Pm[12 * ra + ca] * ((r < 6 ? T(1) : dt) * (c < 6 ? T(1) : dt));
                    if (c < 6) acc += ck[6] * ck[r] * ck[c];
                    else if (r == c) acc += ck[7];
#pragma unroll
                    for (int s = 0; s < 6; ++s) acc -= aug[AUGW * s + 6 + r] * X[13 * s + c];
                    Pn[12 * r + c] = acc;
                    Pn[12 * c + r] = acc;
                    if (k > 0) {
                        if (r < 6 && c >= 6) {
                            const int a = c - 6;
                            augn[AUGW * a + 6 + r] = acc - krn[WR_SG + 12 * a + r];
                            augn[AUGW * a + 12 + r] = dt * acc - krn[WR_SG + 12 * a + 6 + r];
                        } else if (r >= 6) {
                            const int a = r - 6, b = c - 6;
                            augn[AUGW * a + b] = acc + krn[WR_S + 6 * a + b];
                            if (a != b) augn[AUGW * b + a] = acc + krn[WR_S + 6 * b + a];
                        }
                    }
                }
            }
            if (pr >= 0 && pr < 12) {
                T acc = kr[WR_HX + pr] + (pr < 6 ? pv[pr] : dt * pv[pr - 6]);
#pragma unroll
                for (int s = 0; s < 6; ++s) acc -= aug[AUGW * s + 6 + pr] * X[13 * s + 12];
                pn[pr] = acc;
                if (k > 0 && pr >= 6) augn[AUGW * (pr - 6) + 18] = acc - krn[WR_D + pr - 6];
            }
        }
        __syncwarp();
        cur ^= 1;
    }
    cp_async_wait<0>();
    __threadfence();                                // gains written by lanes < 13 -> visible to the cp.async reads
    __syncwarp();
    // forward roll-out: x_0 = xs, v_{k+1} = Kw_k x_k + kw_k, q_{k+1} = q_k + dt v_k; gains prefetched 2 stages ahead
    T* z = M.z;
    if (lane < 12) z[lane] = M.x0[lane];
    stage_copy<T>(M.stage, M.gains, WG_STRIDE, lane);
    cp_async_commit();
    if (N >= 3) stage_copy<T>(M.stage + WSTG, M.gains + WG_STRIDE, WG_STRIDE, lane);
    cp_async_commit();
    for (int k = 0; k < N - 1; ++k) {
        if (k + 2 < N - 1) stage_copy<T>(M.stage + WSTG * ((k + 2) % 3), M.gains + (size_t)WG_STRIDE * (k + 2), WG_STRIDE, lane);
        cp_async_commit();
        cp_async_wait<2>();
        __syncwarp();
        const T* gk = M.stage + WSTG * (k % 3);
        T* zk = z + 18 * k;
        if (lane < 6) {
            T a0 = gk[72 + lane], a1 = T(0);
#pragma unroll
            for (int c = 0; c < 6; ++c) {
                a0 += gk[12 * lane + c] * zk[c];
                a1 += gk[12 * lane + 6 + c] * zk[6 + c];
            }
            zk[18 + 6 + lane] = a0 + a1;
        } else if (lane < 12) {
            const int r = lane - 6;
            zk[18 + r] = zk[r] + dt * zk[6 + r];
        }
        __syncwarp();
    }
    cp_async_wait<0>();
    // controls, lane = knot: u_k = Y_k (v_{k+1} - Gq q_k - Gv v_k - cv_k)
    for (int k = lane; k < N - 1; k += 32) {
        const T* kr = M.rec + (size_t)WR_STRIDE * k;
        const T* zk = z + 18 * k;
        T tmp[6];
#pragma unroll
        for (int r = 0; r < 6; ++r) {
            T a0 = zk[18 + 6 + r] - kr[WR_CV + r], a1 = T(0);
#pragma unroll
            for (int s = 0; s < 6; ++s) {
                a0 -= kr[WR_GQ + 6 * r + s] * zk[s];
                a1 -= kr[WR_GV + 6 * r + s] * zk[6 + s];
            }
            tmp[r] = a0 + a1;
        }
#pragma unroll
        for (int r = 0; r < 6; ++r) {
            T acc = T(0);
#pragma unroll
            for (int s = 0; s < 6; ++s) acc += kr[WR_Y + 6 * r + s] * tmp[s];
            z[18 * k + 12 + r] = acc;
        }
    }
    __syncwarp();
}

// -----------------------------------------------------------------------------------------------------------
// Merit terms of the point xa + alpha (xb - xa), lane = knot (src/osqp_sqp.py:13-47): every lane returns the four
// sums {qcost, vcost, ucost, integrator_err} over all knots (lane-strided partial sums, then an xor tree — the same
// order as merit_reduce in i7_sqp.cuh).
template <typename T>
__device__ __noinline__ void wmerit_eval(const Robot<T>& rob, const SolveParams<T>& P, const T* goals, const T* xa,
                                         const T* xb, const T* fext, T alpha, int lane, T* sums) {
    const int N = P.N;
    T s0 = T(0), s1 = T(0), s2 = T(0), s3 = T(0);
    for (int k = lane; k < N; k += 32) {
        const bool dyn = k < N - 1;
        T q[6], qd[6], u[6];
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            q[i] = xa[18 * k + i] + alpha * (xb[18 * k + i] - xa[18 * k + i]);
            qd[i] = xa[18 * k + 6 + i] + alpha * (xb[18 * k + 6 + i] - xa[18 * k + 6 + i]);
            u[i] = dyn ? xa[18 * k + 12 + i] + alpha * (xb[18 * k + 12 + i] - xa[18 * k + 12 + i]) : T(0);
        }
        T cs[12], pee[3];
#pragma unroll
        for (int i = 0; i < 6; ++i) sincos_t<T>(q[i], &cs[6 + i], &cs[i]);
        fk_ee_only<T>(rob, cs, pee);
        const T* goal = goals + 3 * k;
        const T e[3] = {pee[0] - goal[0], pee[1] - goal[1], pee[2] - goal[2]};
        const T Qk = (k == N - 1) ? P.QN : T(1);
        s0 += Qk * (e[0] * e[0] + e[1] * e[1] + e[2] * e[2]);
        T vv = T(0), uu = T(0);
#pragma unroll
        for (int i = 0; i < 6; ++i) { vv += qd[i] * qd[i]; uu += u[i] * u[i]; }
        s1 += P.dQ * vv;
        s2 += P.R * uu;
        if (dyn) {
            T qdd[6];
            aba<T, false>(rob, cs, qd, u, fext, qdd, nullptr, nullptr, nullptr);
            T qe = T(0), ve = T(0);
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                const T xnq = xa[18 * (k + 1) + i] + alpha * (xb[18 * (k + 1) + i] - xa[18 * (k + 1) + i]);
                const T xnv = xa[18 * (k + 1) + 6 + i] + alpha * (xb[18 * (k + 1) + 6 + i] - xa[18 * (k + 1) + 6 + i]);
                const T dq = q[i] + qd[i] * P.dt - xnq;
                const T dv = qd[i] + qdd[i] * P.dt - xnv;
                qe += dq * dq; ve += dv * dv;
            }
            s3 += sqrt_t<T>(qe) + sqrt_t<T>(ve);
        }
    }
    sums[0] = warp_sum<T>(s0); sums[1] = warp_sum<T>(s1); sums[2] = warp_sum<T>(s2); sums[3] = warp_sum<T>(s3);
}

template <typename T>
I7_DEV T wmerit_total(const SolveParams<T>& P, const T* xa, const T* xb, T alpha, const T* sums) {
    T d2 = T(0);
    for (int i = 0; i < 12; ++i) { const T d = alpha * (xb[i] - xa[i]); d2 += d * d; }
    return sums[0] + sums[1] + sums[2] + P.mu * (sums[3] + sqrt_t<T>(d2));
}

// -----------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(32, (sizeof(T) == 8 ? 8 : 12))
sqp_warp_kernel(const __grid_constant__ Robot<T> rob, const __grid_constant__ SolveParams<T> P, unsigned int* work_counter) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* sm = reinterpret_cast<T*>(smem_raw);
    const int lane = threadIdx.x, N = P.N;
    const int nz = 18 * N - 6;
    const WLayout L = make_wlayout(N);
    WPtrs<T> M;
    M.xu = sm + L.xu; M.z = sm + L.z; M.goals = sm + L.goals; M.x0 = sm + L.x0; M.cost = sm + L.cost;
    M.work = sm + L.work; M.stage = sm + L.stage;
    M.rec = P.spill_ws + (size_t)blockIdx.x * warp_scratch_elems(N);
    M.gains = M.rec + (size_t)N * WR_STRIDE;
    __shared__ T s_fext[6];

    int b = blockIdx.x;
    while (b < P.batch) {
        // ---- load inputs ----
        for (int t = lane; t < nz; t += 32) M.xu[t] = P.xu_in[(size_t)b * nz + t];
        if (lane < 6) M.xu[nz + lane] = T(0);                          // padded u_{N-1}
        for (int t = lane; t < 3 * N; t += 32) {
            const int k = t / 3, a = t - 3 * k;
            M.goals[t] = P.goals[(size_t)b * N * P.goal_stride + k * P.goal_stride + a];
        }
        if (lane < 12) M.x0[lane] = P.x0 ? P.x0[(size_t)b * 12 + lane] : T(0);
        if (lane < 6) s_fext[lane] = P.fext ? P.fext[(size_t)b * 6 + lane] : T(0);
        if (P.mode == MODE_LINESEARCH || P.mode == MODE_MERIT) {
            for (int t = lane; t < nz; t += 32) M.z[t] = P.xu_full[(size_t)b * nz + t];
            if (lane < 6) M.z[nz + lane] = T(0);
        }
        __syncwarp();
        const T* fext = P.fext ? s_fext : nullptr;
        long long* prof = P.prof ? P.prof + (size_t)b * 32 : nullptr;

        if (P.mode == MODE_MERIT) {
            T sums[4];
            wmerit_eval<T>(rob, P, M.goals, M.xu, M.xu, fext, T(0), lane, sums);
            if (lane == 0) {
                T d2 = T(0);
                for (int i = 0; i < 12; ++i) { const T d = M.xu[i] - M.z[i]; d2 += d * d; }
                T* o = P.merits_out + (size_t)b * 6;
                o[0] = sums[0]; o[1] = sums[1]; o[2] = sums[2]; o[3] = sums[3]; o[4] = sqrt_t<T>(d2);
                o[5] = sums[0] + sums[1] + sums[2] + P.mu * (sums[3] + o[4]);
            }
        } else {
            int qp_iters = 0, n_ls = 0, n_steps = 0, status = 0;
            T alphas[2] = {T(0), T(0)}, steps[2] = {T(0), T(0)};
            T merit_final = T(0), merit_first = T(0);
            for (int qp = 0; qp < P.max_qp; ++qp) {
                qp_iters = qp + 1;
                if (prof && lane == 0) prof[5 * qp + 0] = clock64();
                if (P.mode != MODE_LINESEARCH) {
                    for (int k = lane; k < N; k += 32)
                        wsetup_knot<T>(rob, P, M.xu + 18 * k, M.goals + 3 * k, fext, k, P.mode == MODE_DUMP,
                                       M.cost + COST_STRIDE * k, M.rec + (size_t)WR_STRIDE * k);
                    __threadfence();                    // records -> visible to the other lanes' cp.async reads
                    __syncwarp();
                    if (prof && lane == 0) prof[5 * qp + 1] = clock64();
                    if (P.mode == MODE_DUMP) {
                        Ptrs<T> D;
                        D.xu = M.xu; D.x0 = M.x0; D.cost = M.cost; D.kr = M.rec;
                        dump_qp<T, WarpRecLayout>(P, D, b);
                        break;
                    }
                    wriccati<T>(P, M, lane);
                    if (prof && lane == 0) prof[5 * qp + 2] = clock64();
                    if (P.mode == MODE_QP) {
                        for (int t = lane; t < nz; t += 32) P.z_out[(size_t)b * nz + t] = M.z[t];
                        break;
                    }
                }
                // ---- line search (src/osqp_sqp.py:49-74): first alpha in decreasing order with merit <= base ----
                T alpha = T(0), merit_acc, base;
                {
                    const bool all = (P.mode == MODE_LINESEARCH);
                    T sums[4];
                    wmerit_eval<T>(rob, P, M.goals, M.xu, M.z, fext, T(0), lane, sums);
                    base = wmerit_total<T>(P, M.xu, M.z, T(0), sums);
                    merit_acc = base;
                    if (!(base == base)) status = 2;
                    if (all && lane == 0) P.merits_out[(size_t)b * 9] = base;
                    for (int c = 1; c < 9; ++c) {
                        const T a = T(c_alphas[c]);
                        wmerit_eval<T>(rob, P, M.goals, M.xu, M.z, fext, a, lane, sums);
                        const T m = wmerit_total<T>(P, M.xu, M.z, a, sums);
                        if (all && lane == 0) P.merits_out[(size_t)b * 9 + c] = m;
                        if (alpha == T(0) && m <= base) { alpha = a; merit_acc = m; if (!all) break; }
                    }
                }
                if (prof && lane == 0) prof[5 * qp + 3] = clock64();
                merit_final = merit_acc;
                if (qp == 0) merit_first = base;
                if (qp < 2) alphas[qp] = alpha;
                ++n_ls;
                if (P.mode == MODE_LINESEARCH) {
                    if (lane == 0) P.alpha_out[b] = alpha;
                    break;
                }
                if (alpha == T(0)) {
                    // src/osqp_sqp.py:81-82 `continue`: the next pass would re-solve the identical QP and, with an
                    // exact QP solve, reach the same alpha = 0; the remaining passes are skipped but still counted.
                    if (status == 0) status = 1;
                    for (int q2 = qp + 1; q2 < P.max_qp; ++q2) { if (q2 < 2) alphas[q2] = T(0); ++n_ls; qp_iters = q2 + 1; }
                    break;
                }
                // XU += alpha (z - XU); step norm
                T loc = T(0);
                for (int t = lane; t < nz; t += 32) {
                    const T st = alpha * (M.z[t] - M.xu[t]);
                    M.xu[t] += st;
                    loc += st * st;
                }
                const T stepsize = sqrt_t<T>(warp_sum<T>(loc));
                __syncwarp();
                if (prof && lane == 0) prof[5 * qp + 4] = clock64();
                if (n_steps < 2) steps[n_steps] = stepsize;
                ++n_steps;
                if (stepsize < P.step_tol) break;
            }
            if (P.mode == MODE_SOLVE) {
                for (int t = lane; t < nz; t += 32) P.xu_out[(size_t)b * nz + t] = M.xu[t];
                if (P.stats && lane == 0) {
                    T* st = P.stats + (size_t)b * 12;
                    st[0] = T(qp_iters); st[1] = T(n_ls); st[2] = T(n_steps); st[3] = T(status);
                    st[4] = alphas[0]; st[5] = alphas[1]; st[6] = steps[0]; st[7] = steps[1];
                    st[8] = merit_final; st[9] = merit_first; st[10] = T(0); st[11] = T(0);
                }
            }
        }
        __syncwarp();
        // next problem: the first gridDim.x problems are assigned statically, the rest through a work counter
        unsigned int nxt = 0;
        if (lane == 0) nxt = atomicAdd(work_counter, 1u);
        b = (int)gridDim.x + (int)__shfl_sync(0xffffffffu, nxt, 0);
    }
}

}  // namespace i7
