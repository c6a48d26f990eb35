// i7_sqp.cuh — the fused SQP-MPC solve kernel: one CTA per independent problem, everything between the
// input trajectory and the output trajectory stays in shared memory / registers.
//
// Phases per SQP iteration (reference call stack: src/osqp_sqp.py:76-93):
//   setup     per-knot FK + cost terms, mass-matrix columns (CRBA by unit-acceleration RNEA), bias torques,
//             M^-1 / forward dynamics by a 6x6 LDL', analytic dID columns, -M^-1 dID, and every stage-local
//             product the QP recursion needs — all parallel over (knot, column) tasks    (src/osqp_solver.py:70-135)
//   riccati   exact solve of the equality-constrained QP  min 1/2 z'Pz + g'z  s.t.  Az = l  by a backward
//             block recursion over the knots — a block LDL' of the block-tridiagonal KKT system in stage
//             order.  B_k = [0; H_k] with H_k = dt M^-1 invertible, so the free variable of stage k is taken to
//             be w = v_{k+1} (u_k = H_k^-1 (w - G_k x_k - c_k)): the sequential part of a stage is ONE 6x6
//             SPD solve with 13 right-hand sides plus a rank-6 update of the 12x12 value matrix; everything
//             else is precomputed per knot in parallel.  Forward roll-out, then u_k for all knots in parallel.
//                                                                                        (src/osqp_solver.py:137-143)
//   merit     line-search candidates x knots in parallel, evaluated lazily in two waves  (src/osqp_sqp.py:13-74)
//   update    first-accept alpha selection, XU += alpha (z - XU), statistics              (src/osqp_sqp.py:80-92)
#pragma once
#include "i7_wf.cuh"
#include "i7_xchg.cuh"

namespace i7 {

enum Mode : int {
    MODE_SOLVE = 0,       // full sqp()
    MODE_QP = 1,          // z = setup_and_solve_qp(xu).x
    MODE_DUMP = 2,        // reference-ordered Adata / Pdata / g / l
    MODE_LINESEARCH = 3,  // alpha + nine merits for given (xu, xu_full)
    MODE_MERIT = 4,       // merit terms of xu against xu_ref (passed as xu_full)
    // the two halves of an SQP pass around an external QP solver (qp_mode "osqp": i7_admm.cuh):
    MODE_EXPORT = 5,      // linearisation + cost terms of the pass, as compact per-knot blocks (QB_*)
    MODE_STEP = 6         // line search on (xu, xu_full = QP solution), update, statistics of pass `qp_index`
};
// per-knot block of MODE_EXPORT: A_k = [[I, dt I], [Gq, Gv]], B_k = [0; Hm], c_k = [0; cv]; cost = j[6], Q_k, dQ_k, R_k
constexpr int QB_GQ = 0, QB_GV = 36, QB_CV = 72, QB_HM = 78, QB_COST = 114, QB_STRIDE = 124;

// Epilogue of MODE_SOLVE (optional): the CTA that finishes the LAST problem of the launch packs [costs | best_cost,
// best_index, XU_best], stores it into every peer's exchange buffer over NVLink, waits for the other ranks and writes
// the overall winner — selection and multi-GPU exchange without a second launch (i7_solve_select).
struct XchgArgs {
    unsigned char* const* peers;       // exchange buffers of all ranks (device copy of the pointer table)
    unsigned long long seq, timeout_ns;
    void* out;                         // [best_cost, global index, XU_best]
    int* err;
    unsigned* ticket;                  // problems finished so far in this launch (reset by the last one)
    long long* wait_ns;                // optional: time spent waiting for the peers, summed over the launches
    int rank, world, index_offset, enabled;
};

template <typename T>
struct SolveParams {
    const T* xu_in;       // batch x (18N-6)
    const T* x0;          // batch x 12
    const T* goals;       // batch x (N * goal_stride)
    const T* fext;        // batch x 6, Featherstone order [n; f], or nullptr
    const T* xu_full;     // MODE_LINESEARCH: full step; MODE_MERIT: reference iterate
    T* xu_out;            // batch x (18N-6)
    T* stats;             // batch x 12 or nullptr
    T* z_out;             // MODE_QP
    T* y_out;             // MODE_QP, optional: duals of the 12N equality rows
    T* Adata; T* Pdata; T* g; T* l;   // MODE_DUMP (each may be nullptr)
    T* alpha_out; T* merits_out;      // MODE_LINESEARCH (9 per problem) / MODE_MERIT (6 per problem)
    long long* prof;      // optional: per problem 32 clock64() stamps (phase boundaries), debugging aid
    T* spill_ws;          // gridDim.x x Layout::reg_size workspace when the knot records live in global memory
    T* blk_out;           // MODE_EXPORT: batch x N x QB_STRIDE
    const T* qp_info;     // MODE_STEP, optional: batch x 4 record of the QP solver ([0] = its iterations)
    int* done;            // MODE_EXPORT / MODE_STEP, optional: per problem 1 = the SQP loop has ended (skipped; set by MODE_STEP)
    int qp_index;         // MODE_STEP: index of the pass (0 .. max_qp-1)
    XchgArgs xchg;        // MODE_SOLVE epilogue (enabled != 0)
    int x0_into_xu;       // != 0: the warm start's x_0 is overwritten by x0 when the inputs are loaded (XU[:12] = xcur,
                          // src/osqp_mpc.py:69, done here instead of by the caller / i7_shift)
    int N, batch, goal_stride, mode, max_qp, regularize;
    T dt, dQ, R, QN, eps, mu, step_tol;
};

// ---- per-knot record (stride odd => conflict-free when consecutive threads own consecutive knots) ----
constexpr int KR_GQ = 0;      // 36  dt dqdd/dq                        (A_k lower-left)
constexpr int KR_GV = 36;     // 36  I + dt dqdd/dv                    (A_k lower-right)
constexpr int KR_CV = 72;     // 6   affine term of the velocity rows  (c_k = [0; cv])
constexpr int KR_Y = 78;      // 36  M during setup, then Y = H^-1 = M / dt
constexpr int KR_S = 114;     // 36  S = R_k Y'Y   (MODE_DUMP: H = dt M^-1)
constexpr int KR_L = 150;     // 78  upper triangle of L = G'SG + Qx; setup scratch before that
constexpr int KR_SG = 228;    // 72  S G ; after the stage is eliminated: feedback gain Kw (6x12)
constexpr int KR_D = 300;     // 6   d = S cv - R Y u ; after elimination: feed-forward kw
constexpr int KR_HX = 306;    // 12  gx + G'd
constexpr int KR_STRIDE = 319;
constexpr int COST_STRIDE = 9;                       // j[6], Q, dQ, R
// merit scratch per evaluated point (aliases the knot records, dead after the QP solve)
constexpr int MP_STRIDE = WP_STRIDE;      // 97: axes S, world inertias, bias (i7_wf.cuh: wf_point_forward_dynamics)
// Riccati work area (first in the layouts: 16-byte aligned)
constexpr int WK_SIZE = 296;   // two sweep-state blocks (SB_*), see riccati_sweep

// SPILL form (long horizons: the records of all knots do not fit an SM): the records live in a global workspace, but
// they are BUILT in a shared-memory tile of TILE_K knots at a time (the setup makes ~10 passes over a record and uses it
// as scratch: from L2 that costs 3.6x per knot) and copied out once; the same tile is the scratch of the line-search
// points (MP_STRIDE numbers each), evaluated TILE_PTS at a time.
constexpr int TILE_K = 20;      // 6380 numbers: 65 line-search points, two windows of a 128-knot horizon
struct Layout {
    int xu, z, goals, x0, cost, work, ls, reg, reg_size, tile, total;
};
__host__ __device__ inline Layout make_layout(int N, bool spill) {
    Layout L;
    int o = 0;
    L.work = o; o += WK_SIZE;
    L.xu = o; o += 18 * N;
    L.goals = o; o += 3 * N;
    L.x0 = o; o += 12;
    L.cost = o; o += COST_STRIDE * N;
    L.z = o; o += 18 * N;            // z and ls are adjacent: during setup (both dead) they hold the aux side array
    // line-search terms of up to 5 candidates per wave; SPILL form: one candidate at a time (its waves are serial anyway:
    // the points of ONE candidate already exceed the scratch tile), so only the 6N of phase_controls / 4N terms remain
    L.ls = o; o += (spill ? (6 * N > 400 ? 6 * N : 400) : (4 * 5 * N > 400 ? 4 * 5 * N : 400)) + 16;
    L.reg = o;
    L.reg_size = KR_STRIDE * N;
    L.tile = o;
    if (!spill) o += L.reg_size;
    else o += TILE_K * KR_STRIDE;
    L.total = o;
    return L;
}

template <typename T>
struct Ptrs {
    T *xu, *z, *goals, *x0, *cost, *work, *ls, *kr, *aux;
    T* pair;            // latency form only: 432 numbers per knot for the per-(column, link) derivative terms, or nullptr
    int merit_cap;      // candidates per merit wave that fit the scratch region
};

// the knots a CTA works on in the per-knot phases: [k0, k0 + nc) for cost terms, [k0, k0 + nd) for dynamics;
// the whole horizon (0, N-1, N) when one CTA owns the problem.  Records / cost terms are indexed window-relative.
struct Win { int k0, nd, nc; };

static __device__ __constant__ double c_alphas[9] = {0.0, 1.0, 0.5, 0.25, 0.125, 0.0625, 0.03125, 0.015625, 0.0078125};
// -----------------------------------------------------------------------------------------------------------
// setup: linearised dynamics, cost terms and all stage-local QP products at the current iterate.
// The dynamics part works in world-frame coordinates (i7_rbd.cuh, wf_*): every wave below is a set of independent
// (knot, link) / (knot, column) tasks with a dependency depth of a few dozen FMAs.
// Scratch of a knot while its record is being built (dead before S / SG / d / L / hx are written):
constexpr int WS_CS = KR_GQ;    // 12  cos / sin of q                 (raw dID columns land in GQ / GV afterwards)
constexpr int WS_PEE = KR_GQ + 12;  // 3  position of the last joint = spatial origin of the knot
constexpr int WS_S = 114;       // 36  world joint axes S_l
constexpr int WS_V = 150;       // 36  link velocities v_l
constexpr int WS_A = 186;       // 36  link accelerations a_l
constexpr int WS_F = 222;       // 36  link forces f_l
constexpr int WS_I = 258;       // 54  world inertias (IO[6], H[3]) per link          -> 312 <= KR_STRIDE
constexpr int AX_L = 0, AX_DINV = 15, AX_MU = 21, AUX_STRIDE = 28;   // LDL' of M, M^-1 u  (side array, per knot)
static_assert(WS_I + 54 <= KR_STRIDE, "setup scratch overflows the knot record");

template <typename T, int NT>
__device__ void phase_setup(const Robot<T>& rob, const SolveParams<T>& P, const Ptrs<T>& M, const Win W, const T* fext,
                            bool for_dump, long long* prof) {
    const int tid = threadIdx.x;
    int pi = 16;
#define I7_STAMP() do { if (prof && tid == 0) prof[pi++] = clock64(); } while (0)
    I7_STAMP();
    const int N = P.N, nk = W.nd, nc = W.nc;      // dynamics knots / cost knots of this CTA's window
    const T* xu = M.xu + 18 * W.k0;               // window-relative: knot index k below is local, records are local
    const T* goals = M.goals + 3 * W.k0;
    const T dt = P.dt, rdt = T(1) / P.dt;
    // task index = major * np + knot.  With a whole horizon per CTA the knots are padded to whole warps so that the
    // lanes of a warp share the major index (link | column: uniform trip counts and predicates); with a few knots
    // per CTA (cluster form) the tasks are packed densely instead — idle lanes would only waste FP64 issue slots
    const int np = nk > 24 ? (nk + 31) & ~31 : nk;       // (a 20-knot tile padded to 32 would idle 3 lanes in 8)
    // W0: cos / sin of every joint angle
    for (int t = tid; t < 6 * nc; t += NT) {
        const int i = t / nc, k = t - i * nc;
        T s, c;
        sincos_t<T>(xu[18 * k + i], &s, &c);
        T* kr = M.kr + KR_STRIDE * k;
        kr[WS_CS + i] = c; kr[WS_CS + 6 + i] = s;
    }
    __syncthreads();
    // W0b: per knot: position of the last joint (the spatial origin of the knot, and eepos of the cost)
    for (int k = tid; k < nc; k += NT) {
        T* kr = M.kr + KR_STRIDE * k;
        T pee[3];
        wf_tip<T>(rob, kr + WS_CS, pee);
        kr[WS_PEE] = pee[0]; kr[WS_PEE + 1] = pee[1]; kr[WS_PEE + 2] = pee[2];
    }
    __syncthreads();
    I7_STAMP();
    // W1: per (knot, link): world axis, velocity, acceleration at qdd = 0, world inertia, force
    for (int t = tid; t < 6 * np; t += NT) {
        const int l = t / np, k = t - l * np;
        if (k >= nk) continue;
        T* kr = M.kr + KR_STRIDE * k;
        wf_link<T>(rob, kr + WS_CS, xu + 18 * k + 6, fext, kr + WS_PEE, l, kr + WS_S + 6 * l, kr + WS_V + 6 * l,
                   kr + WS_A + 6 * l, kr + WS_I + 9 * l, kr + WS_F + 6 * l);
    }
    // the last knot of the horizon has a cost but no dynamics: only its joint axes are needed (for J'e below)
    if (W.k0 + nc == N) {
        for (int l = tid; l < 6; l += NT) {
            T* kr = M.kr + KR_STRIDE * (nc - 1);
            T dummy_v[6], dummy_a[6], dummy_I[9], dummy_f[6];
            wf_link<T>(rob, kr + WS_CS, xu + 18 * (nc - 1) + 6, nullptr, kr + WS_PEE, l, kr + WS_S + 6 * l, dummy_v, dummy_a,
                       dummy_I, dummy_f);
        }
    }
    __syncthreads();
    I7_STAMP();
    // W2: per (knot, j): column j of M (rows <= j, mirrored) and the bias torque b_j (parked in the cv slot);
    //     per cost knot: j = J'e (column i of the translational Jacobian is the linear part of S_i) and the weights
    for (int t = tid; t < 6 * np + nc; t += NT) {
        if (t < 6 * np) {
            const int j = t / np, k = t - j * np;
            if (k >= nk) continue;
            T* kr = M.kr + KR_STRIDE * k;
            T bj;
            wf_mass_column<T>(rob, kr + WS_S, kr + WS_I, kr + WS_F, j, kr + KR_Y, &bj);
            kr[KR_CV + j] = bj;
        } else {
            const int k = t - 6 * np;
            const T* kr = M.kr + KR_STRIDE * k;
            const T* gk = goals + 3 * k;
            const T e[3] = {kr[WS_PEE] - gk[0], kr[WS_PEE + 1] - gk[1], kr[WS_PEE + 2] - gk[2]};
            T* ck = M.cost + COST_STRIDE * k;
#pragma unroll
            for (int i = 0; i < 6; ++i)
                ck[i] = e[0] * kr[WS_S + 6 * i + 3] + e[1] * kr[WS_S + 6 * i + 4] + e[2] * kr[WS_S + 6 * i + 5];
            const T nrm = sqrt_t<T>(e[0] * e[0] + e[1] * e[1] + e[2] * e[2]);
            const T sc = P.regularize ? T(1) / (nrm + P.eps) : T(1);
            ck[6] = (W.k0 + k == N - 1) ? P.QN : T(1);
            ck[7] = P.dQ * sc;
            ck[8] = P.R * sc;
        }
    }
    __syncthreads();
    I7_STAMP();
    // W3: per knot: LDL' of M (kept), qdd = M^-1 (u - b) (into the cv slot), M^-1 u
    for (int k = tid; k < nk; k += NT) {
        T* kr = M.kr + KR_STRIDE * k;
        T* ax = M.aux + AUX_STRIDE * k;
        T A[6][6], dinv[6], b[6], u[6];
#pragma unroll
        for (int r = 0; r < 6; ++r)
#pragma unroll
            for (int q = 0; q <= r; ++q) A[r][q] = kr[KR_Y + 6 * r + q];
#pragma unroll
        for (int r = 0; r < 6; ++r) { u[r] = xu[18 * k + 12 + r]; b[r] = u[r] - kr[KR_CV + r]; }
        spd6_factor<T>(A, dinv);
        spd6_solve_factored<T>(A, dinv, b);
        spd6_solve_factored<T>(A, dinv, u);
        int o = 0;
#pragma unroll
        for (int r = 1; r < 6; ++r)
#pragma unroll
            for (int q = 0; q < r; ++q) ax[AX_L + o++] = A[r][q];
#pragma unroll
        for (int r = 0; r < 6; ++r) { ax[AX_DINV + r] = dinv[r]; ax[AX_MU + r] = u[r]; kr[KR_CV + r] = b[r]; }
    }
    __syncthreads();
    I7_STAMP();
    // W4: per (knot, link): fold qdd into the link accelerations / forces
    for (int t = tid; t < 6 * np; t += NT) {
        const int l = t / np, k = t - l * np;
        if (k >= nk) continue;
        T* kr = M.kr + KR_STRIDE * k;
        wf_add_qdd<T>(rob, kr + WS_S, kr + KR_CV, l, kr + WS_I + 9 * l, kr + WS_A + 6 * l, kr + WS_F + 6 * l);
    }
    __syncthreads();
    I7_STAMP();
    // W5: 12 analytic inverse-dynamics derivative columns per knot (raw d tau/dq, d tau/dqd into the G slots).
    if (M.pair) {
        // latency form: one task per (column, knot, link) term, then one per (column, knot, row) — with few knots per
        // CTA this fills the lanes that whole-column tasks would leave idle
        const int slots = (6 * nk + 31) & ~31;
        for (int t = tid; t < 12 * slots; t += NT) {
            const int col = t / slots, rem = t - col * slots;
            const int k = rem / 6, l = rem - 6 * k;
            const int j = col < 6 ? col : col - 6;
            if (k >= nk || l < j) continue;
            const T* kr = M.kr + KR_STRIDE * k;
            T* pr = M.pair + ((col * nk + k) * 6 + l) * 6;
            if (col < 6) wf_dpair<T, true>(rob, kr + WS_S, kr + WS_V, kr + WS_A, kr + WS_I, j, l, pr);
            else wf_dpair<T, false>(rob, kr + WS_S, kr + WS_V, kr + WS_A, kr + WS_I, j, l, pr);
        }
        __syncthreads();
        for (int t = tid; t < 12 * slots; t += NT) {
            const int col = t / slots, rem = t - col * slots;
            const int k = rem / 6, i = rem - 6 * k;
            const int j = col < 6 ? col : col - 6;
            if (k >= nk) continue;
            T* kr = M.kr + KR_STRIDE * k;
            const T* pr = M.pair + (col * nk + k) * 36;
            if (col < 6) kr[KR_GQ + 6 * i + j] = wf_dgather<T, true>(kr + WS_S, kr + WS_F, pr, j, i);
            else kr[KR_GV + 6 * i + j] = wf_dgather<T, false>(kr + WS_S, kr + WS_F, pr, j, i);
        }
    } else {
    // (whole warps per column here even for a 17..24-knot tile: with mixed columns in a warp the longest one sets the pace)
    const int np5 = nk > 16 ? (nk + 31) & ~31 : nk;
    for (int t = tid; t < 12 * np5; t += NT) {
        // a column of joint j costs ~ (6 - j) link terms and a warp owns whole columns (slots w, w + NT/32, ...):
        // the order below gives every warp of a 256-thread CTA at most 6 units of work (10 in natural order)
        const int slot = t / np5, k = t - slot * np5;
        // (128-thread CTA, four warps: q0 q3 q5 | v0 v3 v5 | q1 q2 q4 | v1 v2 v4 — 10, 10, 11, 11 units against 12 with
        // the 256-thread order)
        const int col = NT == 128 ? (int)((0xA4B582937160ull >> (4 * slot)) & 15)       // q0 v0 q1 v1 q3 v3 q2 v2 q5 v5 q4 v4
                                  : (int)((0xB9A471605382ull >> (4 * slot)) & 15);      // q2 v2 q3 q5 q0 v0 q1 v1 q4 v4 v3 v5
        if (k >= nk) continue;
        T* kr = M.kr + KR_STRIDE * k;
        // the raw columns land in the GQ / GV slots, which hold cos / sin and the origin until here: every column
        // task of a knot has read what it needs from the link records, not from those slots
        if (col < 6) wf_dcolumn<T, true>(rob, kr + WS_S, kr + WS_V, kr + WS_A, kr + WS_F, kr + WS_I, col, kr + KR_GQ + col, 6);
        else wf_dcolumn<T, false>(rob, kr + WS_S, kr + WS_V, kr + WS_A, kr + WS_F, kr + WS_I, col - 6, kr + KR_GV + (col - 6), 6);
    }
    }
    __syncthreads();
    I7_STAMP();
    // W6: G = -dt M^-1 dID (+ I on the velocity block), in place, one column per task, with the stored LDL'
    for (int t = tid; t < 12 * nk; t += NT) {
        const int col = t / nk, k = t - col * nk;
        T* kr = M.kr + KR_STRIDE * k;
        const T* ax = M.aux + AUX_STRIDE * k;
        T* G = kr + (col < 6 ? KR_GQ : KR_GV);
        const int c = col < 6 ? col : col - 6;
        T A[6][6], dinv[6], y[6];
        int o = 0;
#pragma unroll
        for (int r = 1; r < 6; ++r)
#pragma unroll
            for (int q = 0; q < r; ++q) A[r][q] = ax[AX_L + o++];
#pragma unroll
        for (int r = 0; r < 6; ++r) { dinv[r] = ax[AX_DINV + r]; y[r] = G[6 * r + c]; }
        spd6_solve_factored<T>(A, dinv, y);
#pragma unroll
        for (int r = 0; r < 6; ++r) y[r] = -dt * y[r];
        if (col >= 6) y[c] += T(1);
#pragma unroll
        for (int r = 0; r < 6; ++r) G[6 * r + c] = y[r];
    }
    __syncthreads();
    I7_STAMP();
    // W7: cv = v + dt qdd - Gq q - Gv v - dt M^-1 u  (row tasks);  S = R Y'Y with Y = M / dt, one ROW of S per task
    //     (six independent accumulators) — dump mode: a column of H = dt M^-1 instead
    for (int t = tid; t < 12 * np; t += NT) {
        const int e = t / np, k = t - e * np;
        if (k >= nk) continue;
        T* kr = M.kr + KR_STRIDE * k;
        const T* ax = M.aux + AUX_STRIDE * k;
        const T* xk = xu + 18 * k;
        if (e < 6) {
            const int r = e;
            T a0 = xk[6 + r] + dt * (kr[KR_CV + r] - ax[AX_MU + r]), a1 = T(0);
#pragma unroll
            for (int s = 0; s < 6; ++s) {
                a0 -= kr[KR_GQ + 6 * r + s] * xk[s];
                a1 -= kr[KR_GV + 6 * r + s] * xk[6 + s];
            }
            kr[KR_CV + r] = a0 + a1;
        } else {
            const int r = e - 6;
            if (for_dump) {             // column r of H = dt M^-1 with the stored factor
                T A[6][6], dinv[6], y[6];
                int o = 0;
#pragma unroll
                for (int i = 1; i < 6; ++i)
#pragma unroll
                    for (int q = 0; q < i; ++q) A[i][q] = ax[AX_L + o++];
#pragma unroll
                for (int i = 0; i < 6; ++i) { dinv[i] = ax[AX_DINV + i]; y[i] = (i == r) ? T(1) : T(0); }
                spd6_solve_factored<T>(A, dinv, y);
#pragma unroll
                for (int i = 0; i < 6; ++i) kr[KR_S + 6 * i + r] = dt * y[i];
            } else {
                T acc[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
#pragma unroll
                for (int s = 0; s < 6; ++s) {
                    const T ysr = kr[KR_Y + 6 * s + r];
#pragma unroll
                    for (int c = 0; c < 6; ++c) acc[c] += ysr * kr[KR_Y + 6 * s + c];
                }
                const T sc = M.cost[COST_STRIDE * k + 8] * (rdt * rdt);
#pragma unroll
                for (int c = 0; c < 6; ++c) kr[KR_S + 6 * r + c] = sc * acc[c];
            }
        }
    }
    __syncthreads();
    I7_STAMP();
    if (for_dump) return;
    // W8: one COLUMN of SG = S G per task (12 per knot), plus d = S cv - (R/dt) M u as a 13th.  With a whole horizon
    //     per CTA there are several rounds of these tasks per thread and the wave is bound by shared-memory loads
    //     (42 per 36 FMAs): TWO columns per task then share every load of S (48 loads per 72 FMAs)
    const bool wide = nk > 16;
    if (wide) {
        for (int t = tid; t < 6 * np; t += NT) {
            const int cp = t / np, k = t - cp * np;
            if (k >= nk) continue;
            T* kr = M.kr + KR_STRIDE * k;
            const int c0 = 2 * cp;
            const T* G = kr + (c0 < 6 ? KR_GQ + c0 : KR_GV + c0 - 6);
            T g0[6], g1[6], a0[6] = {T(0), T(0), T(0), T(0), T(0), T(0)}, a1[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
#pragma unroll
            for (int s = 0; s < 6; ++s) { g0[s] = G[6 * s]; g1[s] = G[6 * s + 1]; }
#pragma unroll
            for (int s = 0; s < 6; ++s)
#pragma unroll
                for (int r = 0; r < 6; ++r) {
                    const T srs = kr[KR_S + 6 * r + s];
                    a0[r] += srs * g0[s]; a1[r] += srs * g1[s];
                }
#pragma unroll
            for (int r = 0; r < 6; ++r) { kr[KR_SG + 12 * r + c0] = a0[r]; kr[KR_SG + 12 * r + c0 + 1] = a1[r]; }
        }
    }
    for (int t = tid + (wide ? 12 * np : 0); t < 13 * np; t += NT) {
        const int c = t / np, k = t - c * np;
        if (k >= nk) continue;
        T* kr = M.kr + KR_STRIDE * k;
        T g[6], acc[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
        if (c < 12) {
            const T* G = kr + (c < 6 ? KR_GQ + c : KR_GV + c - 6);
#pragma unroll
            for (int s = 0; s < 6; ++s) g[s] = G[6 * s];
        } else {
#pragma unroll
            for (int s = 0; s < 6; ++s) g[s] = kr[KR_CV + s];
        }
#pragma unroll
        for (int s = 0; s < 6; ++s)
#pragma unroll
            for (int r = 0; r < 6; ++r) acc[r] += kr[KR_S + 6 * r + s] * g[s];
        if (c < 12) {
#pragma unroll
            for (int r = 0; r < 6; ++r) kr[KR_SG + 12 * r + c] = acc[r];
        } else {
            const T* xk = xu + 18 * k;
            const T sc = M.cost[COST_STRIDE * k + 8] * rdt;
            T mu[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
#pragma unroll
            for (int s = 0; s < 6; ++s) {
                const T us = xk[12 + s];
#pragma unroll
                for (int r = 0; r < 6; ++r) mu[r] += kr[KR_Y + 6 * r + s] * us;
            }
#pragma unroll
            for (int r = 0; r < 6; ++r) kr[KR_D + r] = acc[r] - sc * mu[r];
        }
    }
    __syncthreads();
    I7_STAMP();
    // W9: half a COLUMN c of L = G'(SG) + Qx per task (rows 6h .. 6h+5 of the upper triangle, r <= c; overwrites the
    //     setup scratch); the h = 0 tasks also do hx[c] = gx[c] + G[:, c] . d and, for c < 6, row c of Y = M / dt
    //     Whole horizon per CTA: two columns per task (9 tasks per knot), sharing every load of G.
    if (wide) {
        for (int t = tid; t < 9 * np; t += NT) {
            const int tp = t / np, k = t - tp * np;
            if (k >= nk) continue;
            const int hf = tp >= 6 ? 1 : 0;
            const int c0 = hf ? 6 + 2 * (tp - 6) : 2 * tp;
            T* kr = M.kr + KR_STRIDE * k;
            const T* ck = M.cost + COST_STRIDE * k;
            T sg0[6], sg1[6], l0[6], l1[6];
#pragma unroll
            for (int s = 0; s < 6; ++s) { sg0[s] = kr[KR_SG + 12 * s + c0]; sg1[s] = kr[KR_SG + 12 * s + c0 + 1]; }
            const T* Gh = kr + (hf ? KR_GV : KR_GQ);
#pragma unroll
            for (int rr = 0; rr < 6; ++rr) {
                T b0 = T(0), b1 = T(0);
#pragma unroll
                for (int s = 0; s < 6; ++s) { const T g = Gh[6 * s + rr]; b0 += g * sg0[s]; b1 += g * sg1[s]; }
                l0[rr] = b0; l1[rr] = b1;
            }
            // h (hx terms) and the scaling of Y read G, d, Y only: before the stores below, which reuse setup scratch
            T h0 = T(0), h1 = T(0);
            if (!hf) {
                const T* Gc = kr + (c0 < 6 ? KR_GQ + c0 : KR_GV + c0 - 6);
                h0 = c0 < 6 ? ck[6] * ck[c0] : ck[7] * xu[18 * k + c0];
                h1 = c0 < 6 ? ck[6] * ck[c0 + 1] : ck[7] * xu[18 * k + c0 + 1];
#pragma unroll
                for (int s = 0; s < 6; ++s) { const T dd = kr[KR_D + s]; h0 += Gc[6 * s] * dd; h1 += Gc[6 * s + 1] * dd; }
            }
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const int c = c0 + q;
                const T qc = c < 6 ? ck[6] * ck[c] : T(0);
#pragma unroll
                for (int rr = 0; rr < 6; ++rr) {
                    const int r = 6 * hf + rr;
                    if (r <= c) {
                        T v = q ? l1[rr] : l0[rr];
                        if (c < 6) v += qc * ck[r];
                        else if (r == c) v += ck[7];
                        kr[KR_L + r * 12 - (r * (r - 1)) / 2 + (c - r)] = v;
                    }
                }
                if (!hf) {
                    kr[KR_HX + c] = q ? h1 : h0;
                    if (c < 6) {
#pragma unroll
                        for (int s = 0; s < 6; ++s) kr[KR_Y + 6 * c + s] = kr[KR_Y + 6 * c + s] * rdt;
                    }
                }
            }
        }
    }
    for (int t = tid + (wide ? 24 * np : 0); t < 24 * np; t += NT) {
        const int ch = t / np, k = t - ch * np;
        if (k >= nk) continue;
        const int c = ch >> 1, hf = ch & 1;
        if (hf && c < 6) continue;                       // rows 6..11 of a column c < 6 lie below the diagonal
        T* kr = M.kr + KR_STRIDE * k;
        const T* ck = M.cost + COST_STRIDE * k;
        T sg[6];
#pragma unroll
        for (int s = 0; s < 6; ++s) sg[s] = kr[KR_SG + 12 * s + c];
        const T* Gh = kr + (hf ? KR_GV : KR_GQ);         // rows r = 6 hf + rr  <->  columns rr of Gq | Gv
        T lcol[6];
#pragma unroll
        for (int rr = 0; rr < 6; ++rr) {
            T acc = T(0);
#pragma unroll
            for (int s = 0; s < 6; ++s) acc += Gh[6 * s + rr] * sg[s];
            lcol[rr] = acc;
        }
        const T qc = c < 6 ? ck[6] * ck[c] : T(0);      // + Qx: rank-1 position block, dQ on the velocity diagonal
#pragma unroll
        for (int rr = 0; rr < 6; ++rr) {
            const int r = 6 * hf + rr;
            if (r <= c) {
                T v = lcol[rr];
                if (c < 6) v += qc * ck[r];
                else if (r == c) v += ck[7];
                kr[KR_L + r * 12 - (r * (r - 1)) / 2 + (c - r)] = v;
            }
        }
        if (!hf) {
            const T* Gc = kr + (c < 6 ? KR_GQ + c : KR_GV + c - 6);
            T h = c < 6 ? ck[6] * ck[c] : ck[7] * xu[18 * k + c];
#pragma unroll
            for (int s = 0; s < 6; ++s) h += Gc[6 * s] * kr[KR_D + s];
            kr[KR_HX + c] = h;
            if (c < 6) {
#pragma unroll
                for (int s = 0; s < 6; ++s) kr[KR_Y + 6 * c + s] = kr[KR_Y + 6 * c + s] * rdt;
            }
        }
    }
    __syncthreads();
    I7_STAMP();
#undef I7_STAMP
}

// -----------------------------------------------------------------------------------------------------------
// Stage k of the backward sweep solves  Hww X = [Hwx | hw]  (6 x 6, 13 right-hand sides) built from the value
// function (P, p) of stage k+1:   Hww = P_vv + S,  Hwx = [P_qv' - SG_q, dt P_qv' - SG_v],  hw = p_v - d.
// Double-buffered sweep state in the work area (one block per parity of the stage):
constexpr int SB_P = 0;       // 36  qq block of the value matrix (the only block other entries read back)
constexpr int SB_PV = 36;     // 12  value gradient (q part)
constexpr int SB_H = 48;      // 22  Hww, packed lower triangle row-major (21) + pad
constexpr int SB_R = 70;      // 78  right-hand sides, column-major: R[c][s] = SB_R + 6 c + s  (c = 12: hw)
constexpr int SB_SIZE = 148;  // even: both blocks 16-byte aligned
static_assert(2 * SB_SIZE <= WK_SIZE, "work area too small for the sweep state");
constexpr int GAIN_LEN = 78;  // gains of a stage, column-major like the right-hand sides: K[c][s] = 6 c + s, c = 12: kw

// shared-memory base of the dynamic allocation, typed: every function re-derives it from the symbol so that the
// compiler can prove the address space (LDS / STS instead of generic accesses) even across non-inlined calls
template <typename T> I7_DEV T* smem_base() {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    return reinterpret_cast<T*>(smem_raw);
}

// Where the sweep finds the stage data and keeps the gains.
//   RicLocal   one CTA owns the problem: per-knot records in its shared memory (SPILL: in its global workspace);
//              the gains overwrite the SG / d slots of the record once the stage is eliminated
//   RicGlobal  cluster form: the owners of the knots exported [L | SG | S | d | hx | cost] to an L2-resident
//              exchange buffer; the lead CTA streams it (ld.global.cg, one stage ahead) and keeps the gains in its
//              own shared memory
constexpr int RIC_L = 0, RIC_SG = 78, RIC_S = 150, RIC_D = 186, RIC_HX = 192, RIC_COST = 204, RIC_STRIDE = 216;
constexpr int GAIN_STRIDE = 80;
template <typename T, bool SPILL> struct RicLocal {
    int o_kr, o_cost; T* krg;
    static constexpr int oL = KR_L, oSG = KR_SG, oS = KR_S, oD = KR_D, oHX = KR_HX;
    I7_DEV const T* rec(int k) const { return (SPILL ? krg : smem_base<T>() + o_kr) + KR_STRIDE * k; }
    I7_DEV const T* ck(int k) const { return smem_base<T>() + o_cost + COST_STRIDE * k; }
    I7_DEV T* gain(int k) const { return (SPILL ? krg : smem_base<T>() + o_kr) + KR_STRIDE * k + KR_SG; }
    static I7_DEV T ld(const T* p) { return *p; }
    static constexpr bool kStageGains = SPILL;
    int o_stage, stage_len;      // SPILL: shared-memory region the gains are staged through for the roll-out
};
template <typename T> struct RicGlobal {
    const T* ric; int o_gains;
    static constexpr int oL = RIC_L, oSG = RIC_SG, oS = RIC_S, oD = RIC_D, oHX = RIC_HX;
    I7_DEV const T* rec(int k) const { return ric + RIC_STRIDE * k; }
    I7_DEV const T* ck(int k) const { return ric + RIC_STRIDE * k + RIC_COST; }
    I7_DEV T* gain(int k) const { return smem_base<T>() + o_gains + GAIN_STRIDE * k; }
    static I7_DEV T ld(const T* p) { return __ldcg(p); }
    static constexpr bool kStageGains = false;
};
static_assert(KR_D == KR_SG + 72, "the gains (Kw | kw) must be contiguous in the record");

// reciprocal on the serial path: hardware seed + two Newton steps (a few ulp; 10+ cycles shorter than the IEEE divide)
template <typename T> I7_DEV T rcp_serial(T x) { return T(1) / x; }
template <> I7_DEV double rcp_serial<double>(double x) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    double e = fma(-x, r, 1.0); r = fma(r, e, r);
    e = fma(-x, r, 1.0); r = fma(r, e, r);
    return r;
}
template <typename T> I7_DEV void spd6_factor_serial(T (&A)[6][6], T (&dinv)[6]) {
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        dinv[j] = rcp_serial<T>(A[j][j]);
        T col[6];
#pragma unroll
        for (int i = j + 1; i < 6; ++i) col[i] = A[i][j];
#pragma unroll
        for (int i = j + 1; i < 6; ++i) {
            const T l = col[i] * dinv[j];
#pragma unroll
            for (int k = j + 1; k <= i; ++k) A[i][k] -= l * col[k];
            A[i][j] = l;
        }
    }
}

// x = H^-1 b for a 6x6 SPD matrix given as packed lower triangle (row-major, 21 numbers), b overwritten by x.
// Generic version: LDL' in registers (6 dependent reciprocals on the critical path).
template <typename T> I7_DEV void spd6_solve_packed(const T* Hp, T* b) {
    T A[6][6], dinv[6];
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int q = 0; q <= i; ++q) A[i][q] = Hp[(i * (i + 1)) / 2 + q];
    spd6_factor_serial<T>(A, dinv);
    spd6_solve_factored<T>(A, dinv, b);
}
// fp64: block elimination with 3x3 blocks H = [A B'; B D] through cofactor matrices, everything scaled by the
// determinants so that only ONE reciprocal sits on the dependency chain (and one beside it):
//   C = cof(A), dA = det A;  Ss = dA D - (B C) B'  (= dA x Schur complement);  Cs = cof(Ss), dS = det Ss
//   x2 = Cs (dA b2 - (B C) b1) / dS;   x1 = C (b1 - B' x2) / dA
// Depth ~22 dependent FMAs + 1 reciprocal (~235 cycles) against ~690 for the LDL' chain; the conditioning of the
// stage matrices (<= 1e7) leaves 1e-9 relative accuracy in fp64.  Not used in fp32 (range and accuracy).
template <> I7_DEV void spd6_solve_packed<double>(const double* H, double* b) {
    const double a00 = H[0], a10 = H[1], a11 = H[2], a20 = H[3], a21 = H[4], a22 = H[5];
    const double B[3][3] = {{H[6], H[7], H[8]}, {H[10], H[11], H[12]}, {H[15], H[16], H[17]}};
    const double d33 = H[9], d43 = H[13], d44 = H[14], d53 = H[18], d54 = H[19], d55 = H[20];
    // cofactor matrix of A (symmetric)
    double C[3][3];
    C[0][0] = a11 * a22 - a21 * a21;
    C[0][1] = a21 * a20 - a10 * a22;
    C[0][2] = a10 * a21 - a11 * a20;
    C[1][1] = a00 * a22 - a20 * a20;
    C[1][2] = a10 * a20 - a00 * a21;
    C[2][2] = a00 * a11 - a10 * a10;
    C[1][0] = C[0][1]; C[2][0] = C[0][2]; C[2][1] = C[1][2];
    const double dA = a00 * C[0][0] + a10 * C[0][1] + a20 * C[0][2];
    const double rdA = rcp_serial<double>(dA);                 // off the critical path
    // W = B C
    double W[3][3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) W[i][j] = B[i][0] * C[0][j] + B[i][1] * C[1][j] + B[i][2] * C[2][j];
    // Ss = dA D - W B'  (symmetric: lower part)
    const double s00 = dA * d33 - (W[0][0] * B[0][0] + W[0][1] * B[0][1] + W[0][2] * B[0][2]);
    const double s10 = dA * d43 - (W[1][0] * B[0][0] + W[1][1] * B[0][1] + W[1][2] * B[0][2]);
    const double s11 = dA * d44 - (W[1][0] * B[1][0] + W[1][1] * B[1][1] + W[1][2] * B[1][2]);
    const double s20 = dA * d53 - (W[2][0] * B[0][0] + W[2][1] * B[0][1] + W[2][2] * B[0][2]);
    const double s21 = dA * d54 - (W[2][0] * B[1][0] + W[2][1] * B[1][1] + W[2][2] * B[1][2]);
    const double s22 = dA * d55 - (W[2][0] * B[2][0] + W[2][1] * B[2][1] + W[2][2] * B[2][2]);
    double Cs[3][3];
    Cs[0][0] = s11 * s22 - s21 * s21;
    Cs[0][1] = s21 * s20 - s10 * s22;
    Cs[0][2] = s10 * s21 - s11 * s20;
    Cs[1][1] = s00 * s22 - s20 * s20;
    Cs[1][2] = s10 * s20 - s00 * s21;
    Cs[2][2] = s00 * s11 - s10 * s10;
    Cs[1][0] = Cs[0][1]; Cs[2][0] = Cs[0][2]; Cs[2][1] = Cs[1][2];
    const double dS = s00 * Cs[0][0] + s10 * Cs[0][1] + s20 * Cs[0][2];
    const double rdS = rcp_serial<double>(dS);
    // r2s = dA b2 - W b1 ; x2 = Cs r2s / dS
    double r2[3], x2[3], t1[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) r2[i] = dA * b[3 + i] - (W[i][0] * b[0] + W[i][1] * b[1] + W[i][2] * b[2]);
#pragma unroll
    for (int i = 0; i < 3; ++i) x2[i] = (Cs[i][0] * r2[0] + Cs[i][1] * r2[1] + Cs[i][2] * r2[2]) * rdS;
    // x1 = C (b1 - B' x2) / dA
#pragma unroll
    for (int j = 0; j < 3; ++j) t1[j] = b[j] - (B[0][j] * x2[0] + B[1][j] * x2[1] + B[2][j] * x2[2]);
#pragma unroll
    for (int i = 0; i < 3; ++i) b[i] = (C[i][0] * t1[0] + C[i][1] * t1[1] + C[i][2] * t1[2]) * rdA;
#pragma unroll
    for (int i = 0; i < 3; ++i) b[3 + i] = x2[i];
}

// riccati: backward recursion for the gains, forward roll-out for z.  x0 = measured state (l[:12] = -xs).
//
// Thread roles are fixed for the whole recursion and decoded once into small per-lane "programs" (offsets and
// scales), so the stage loop itself is branch-light and free of index arithmetic:
//   warp 0, lanes 0..12   R2: X = Hww^-1 [Hwx | hw], one right-hand side each (LDL' in registers); the negated
//                         column is stored as the gain column K[c][:] / kw
//   warp 1                the 21 qq entries of the new value matrix and p_q          (stay in the P / PV slots)
//   warp 2                the 36 qv entries: each feeds two entries of the next Hwx
//   warp 3                the 21 vv entries (-> next Hww, packed) and p_v (-> next hw)
// Every entry is  base + Hwx[:, r] . K[c][:]  with base = (L + Qx)[r][c] + (T'PT)[r][c] (resp. hx + T'p), 6-term dot
// products over contiguous columns; the stage-local operands (L + Qx, S, SG, d of the next stage) are fetched one
// stage ahead while the 6x6 solves run.
template <typename T, int NT, typename Src>
__device__ __noinline__ void riccati_sweep(const SolveParams<T>& P, int o_work, int o_z, int o_x0, int o_xu,
                                           const Src src, long long* prof) {
    static_assert(NT >= 128, "riccati_sweep needs at least 4 warps");
    const int tid = threadIdx.x;
    long long acc_r2 = 0, acc_r3 = 0, tt = 0;
    const int N = P.N;
    const T dt = P.dt;
    T* sm = smem_base<T>();
    T* wk = sm + o_work;
    T* z = sm + o_z;
    const T* x0 = sm + o_x0;
    const T* xu = sm + o_xu;
    // ---- lane programs (decoded once) ----
    const int warp = tid >> 5, lane = tid & 31;
    int r = 0, c = 0;
    bool act = false, is_p = false;
    if (warp == 1) {                                   // qq: 21 entries r <= c < 6, then p_q
        if (lane < 21) { int rem = lane, len = 6; while (rem >= len) { rem -= len; --len; ++r; } c = r + rem; act = true; }
        else if (lane < 27) { r = lane - 21; is_p = true; act = true; }
    } else if (warp == 2) {                            // qv: r < 6 <= c   (36 entries -> lanes 0..31 and warp 3 lanes 27..30)
        if (lane < 32) { r = lane / 6; c = 6 + lane % 6; act = true; }
    } else if (warp == 3) {                            // vv: 21 entries 6 <= r <= c, then p_v, then the last 4 qv entries
        if (lane < 21) { int rem = lane, len = 6; while (rem >= len) { rem -= len; --len; ++r; } c = r + rem; r += 6; c += 6; act = true; }
        else if (lane < 27) { r = 6 + lane - 21; is_p = true; act = true; }
        else if (lane < 31) { const int e = 32 + lane - 27; r = e / 6; c = 6 + e % 6; act = true; }
    }
    const int ra = r < 6 ? r : r - 6, ca = c < 6 ? c : c - 6;
    const bool is_P = act && !is_p;
    // entry index e78 of (r, c) in the row-major upper triangle of a 12 x 12 matrix (layout of L)
    const int e78 = r * 12 - (r * (r - 1)) / 2 + (c - r);
    const int oA = SB_R + 6 * r;                                   // Hwx[:, r], contiguous
    const int oB = is_P ? 6 * c : 72;                              // K[c][:] resp. kw
    const int oOld = is_P ? SB_P + 6 * ra + ca : SB_PV + ra;       // (T'PT) resp. (T'p) operand
    const T tfac = (r < 6 ? T(1) : dt) * ((is_P && c >= 6) ? dt : T(1));
    const int oL = is_P ? Src::oL + e78 : Src::oHX + r;            // base term in the record of the stage
    int on0 = -1, on1 = -1, st0 = -1, st1 = -1;                    // next-stage operands (record) / stores (sweep block)
    T m1 = T(1), sg = T(1);
    if (is_P && c < 6) {
        st0 = SB_P + 6 * r + c;
        if (r != c) st1 = SB_P + 6 * c + r;
    } else if (is_P && r < 6) {
        const int i = c - 6;
        st0 = SB_R + 6 * r + i; st1 = SB_R + 6 * (6 + r) + i; m1 = dt; sg = T(-1);
        on0 = Src::oSG + 12 * i + r; on1 = Src::oSG + 12 * i + 6 + r;
    } else if (is_P) {
        const int i = r - 6, j = c - 6;
        st0 = SB_H + (j * (j + 1)) / 2 + i;
        on0 = Src::oS + 6 * i + j;
    } else if (is_p && r < 6) {
        st0 = SB_PV + r;
    } else if (is_p) {
        st0 = SB_R + 72 + (r - 6); sg = T(-1);
        on0 = Src::oD + (r - 6);
    }
    // store an entry of the value function into the block of the next stage
    auto emit = [&](T* blk, T acc, T n0, T n1) {
        if (st0 >= 0) blk[st0] = acc + sg * n0;
        if (st1 >= 0) blk[st1] = m1 * acc + sg * n1;
    };
    // stage-local operands, fetched ONE STAGE AHEAD (they may sit in L2 when the records were exported by other CTAs)
    auto fetch = [&](int k, T& fL, T& f0, T& f1) {
        fL = T(0); f0 = T(0); f1 = T(0);
        if (k < 0 || !act) return;
        fL = Src::ld(src.rec(k) + oL);
        if (k > 0) {
            const T* rn = src.rec(k - 1);
            if (on0 >= 0) f0 = Src::ld(rn + on0);
            if (on1 >= 0) f1 = Src::ld(rn + on1);
        }
    };
    int cur = 0;
    // terminal value function: P_N = blkdiag(QN j j', dQ_N I), p_N = [QN j; dQ_N v]; block 0 <- operands of stage N-2
    if (act) {
        const T* cN = src.ck(N - 1);
        const T* xN = xu + 18 * (N - 1);
        const T* rn = src.rec(N - 2);
        const T n0 = on0 >= 0 ? Src::ld(rn + on0) : T(0), n1 = on1 >= 0 ? Src::ld(rn + on1) : T(0);
        T v = T(0);
        if (is_P) {
            if (c < 6) v = Src::ld(cN + 6) * Src::ld(cN + r) * Src::ld(cN + c);
            else if (r == c) v = Src::ld(cN + 7);
        } else {
            v = r < 6 ? Src::ld(cN + 6) * Src::ld(cN + r) : Src::ld(cN + 7) * xN[r];
        }
        emit(wk, v, n0, n1);
    }
    T cL, cn0, cn1;
    fetch(N - 2, cL, cn0, cn1);
    __syncthreads();
    if (prof && tid == 0) prof[14] = clock64();            // end of the terminal-stage initialisation
    for (int k = N - 2; k >= 0; --k) {
        T* G = src.gain(k);
        T* blk = wk + SB_SIZE * cur;
        T* blkn = wk + SB_SIZE * (cur ^ 1);
        if (prof && tid == 0) tt = clock64();
        T pL, pn0, pn1;
        fetch(k - 1, pL, pn0, pn1);            // consumed in the next iteration
        T base = T(0);
        if (tid < 13) {
            // R2: one right-hand side of X = Hww^-1 [Hwx | hw]
            T b[6];
#pragma unroll
            for (int i = 0; i < 6; ++i) b[i] = blk[SB_R + 6 * tid + i];
            spd6_solve_packed<T>(blk + SB_H, b);
            // feedback / feed-forward of w = v_{k+1}: Kw = -X[:, :12], kw = -X[:, 12]
#pragma unroll
            for (int i = 0; i < 6; ++i) G[6 * tid + i] = -b[i];
        } else if (act) {
            base = cL + blk[oOld] * tfac;
        }
        __syncthreads();
        if (prof && tid == 0) { const long long n = clock64(); acc_r2 += n - tt; tt = n; }
        // R3: P = Qx + L + T'aT - Hwx'X, p = hx + [pq; dt pq] - Hwx' x13 (X = -K); operands of stage k-1
        if (act) {
            T a0 = base, a1 = T(0);
            const T* hb = blk + oA;
            const T* kb = G + oB;
#pragma unroll
            for (int s = 0; s < 3; ++s) {
                a0 += hb[s] * kb[s];
                a1 += hb[s + 3] * kb[s + 3];
            }
            emit(blkn, a0 + a1, cn0, cn1);
        }
        __syncthreads();
        if (prof && tid == 0) acc_r3 += clock64() - tt;
        cur ^= 1;
        cL = pL; cn0 = pn0; cn1 = pn1;
    }
    if (prof && tid == 0) { prof[10] = acc_r2; prof[11] = acc_r3; tt = clock64(); }
    // forward roll-out by warp 0: x_0 = xs, v_{k+1} = Kw_k x_k + kw_k, q_{k+1} = q_k + dt v_k.  The state lives in
    // registers, replicated in every lane; lane r (mod 6) evaluates row r of the gain and the six results are
    // exchanged with shuffles, so the serial chain per stage is 3 FMAs + 2 adds + one shuffle (no shared-memory
    // round trip); the gain row of the next stage is fetched while the current one is applied.
    {
        const int row = tid % 6;
        T x[12], my_q = x0[row], my_v = x0[6 + row];
#pragma unroll
        for (int i = 0; i < 12; ++i) x[i] = x0[i];
        if (tid < 12) z[tid] = x0[tid];
        // stages [ka, kb) by warp 0; gp(k) = gains of stage k; the gain row of the next stage is fetched while the
        // current one is applied
        auto roll = [&](int ka, int kb, auto gp) {
            // the warp may arrive here still split by the loops in front of it (lanes with different trip counts): without
            // re-convergence every shuffle below takes the divergent slow path (measured: 600 instead of 225 cycles / stage)
            __syncwarp();
            T g[12], ff;
            {
                const T* G = gp(ka);
#pragma unroll
                for (int i = 0; i < 12; ++i) g[i] = G[6 * i + row];
                ff = G[72 + row];
            }
            for (int k = ka; k < kb; ++k) {
                T gn[12], ffn;
                {
                    const T* G = gp(k + 1 < kb ? k + 1 : k);
#pragma unroll
                    for (int i = 0; i < 12; ++i) gn[i] = G[6 * i + row];
                    ffn = G[72 + row];
                }
                T a0 = ff, a1 = T(0), a2 = T(0), a3 = T(0);
#pragma unroll
                for (int i = 0; i < 3; ++i) {
                    a0 += g[i] * x[i];
                    a1 += g[3 + i] * x[3 + i];
                    a2 += g[6 + i] * x[6 + i];
                    a3 += g[9 + i] * x[9 + i];
                }
                const T vn = (a0 + a1) + (a2 + a3);
                my_q = my_q + dt * my_v;
                my_v = vn;
                if (tid < 6) { z[18 * (k + 1) + row] = my_q; z[18 * (k + 1) + 6 + row] = vn; }
                __syncwarp();
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    x[i] = x[i] + dt * x[6 + i];
                    x[6 + i] = __shfl_sync(0xffffffffu, vn, i);
                }
#pragma unroll
                for (int i = 0; i < 12; ++i) g[i] = gn[i];
                ff = ffn;
            }
        };
        if constexpr (Src::kStageGains) {
            // the gains sit in the global workspace: an L2 round trip per stage would dominate this serial chain, so all
            // threads copy them into the (idle) shared-memory tile, a chunk of stages at a time, and the chain runs there
            T* ring = sm + src.o_stage;
            const int cap = src.stage_len / GAIN_LEN;
            for (int c0 = 0; c0 < N - 1; c0 += cap) {
                const int nst = N - 1 - c0 < cap ? N - 1 - c0 : cap;
                __syncthreads();
                // a warp per stage, three independent loads per lane, four stages in flight: the copy is bound by the
                // L2 round trip, so what matters is how many loads are outstanding (one element per thread per iteration
                // with a division in the address left them serialised: 34 k cycles for the two chunks of N = 128)
#pragma unroll 4
                for (int k = tid >> 5; k < nst; k += NT / 32) {
                    const T* gk = src.gain(c0 + k);
                    const int l0 = tid & 31;
                    const T v0 = gk[l0], v1 = gk[l0 + 32], v2 = l0 + 64 < GAIN_LEN ? gk[l0 + 64] : T(0);
                    T* rk = ring + GAIN_LEN * k;
                    rk[l0] = v0; rk[l0 + 32] = v1;
                    if (l0 + 64 < GAIN_LEN) rk[l0 + 64] = v2;
                }
                __syncthreads();
                if (tid < 32) roll(c0, c0 + nst, [&](int k) { return (const T*)(ring + GAIN_LEN * (k - c0)); });
            }
        } else {
            if (tid < 32) roll(0, N - 1, [&](int k) { return (const T*)src.gain(k); });
        }
    }
    __syncthreads();
    if (prof && tid == 0) { prof[12] = clock64() - tt; tt = clock64(); }
    (void)tt;
}

// controls of the window's knots: u_k = Y_k (v_{k+1} - Gq q_k - Gv v_k - cv_k); needs the x part of z for knots k, k+1.
// zu_out (cluster form): the controls are also published to the exchange buffer, 6 per knot.
template <typename T, int NT, bool SPILL>
__device__ __noinline__ void phase_controls(const SolveParams<T>& P, int o_z, int o_ls, int o_kr, T* krg, const Win W,
                                            T* zu_out) {
    const int tid = threadIdx.x;
    const int nk = W.nd;
    T* sm = smem_base<T>();
    T* z = sm + o_z + 18 * W.k0;
    T* tmp = sm + o_ls;
    const T* krb = SPILL ? krg : sm + o_kr;
    for (int t = tid; t < 6 * nk; t += NT) {
        const int rr = t / nk, k = t - rr * nk;
        const T* kr = krb + KR_STRIDE * k;
        const T* zk = z + 18 * k;
        T a0 = zk[18 + 6 + rr] - kr[KR_CV + rr], a1 = T(0);
#pragma unroll
        for (int s = 0; s < 6; ++s) {
            a0 -= kr[KR_GQ + 6 * rr + s] * zk[s];
            a1 -= kr[KR_GV + 6 * rr + s] * zk[6 + s];
        }
        tmp[6 * k + rr] = a0 + a1;
    }
    __syncthreads();
    for (int t = tid; t < 6 * nk; t += NT) {
        const int rr = t / nk, k = t - rr * nk;
        const T* kr = krb + KR_STRIDE * k;
        T acc = T(0);
#pragma unroll
        for (int s = 0; s < 6; ++s) acc += kr[KR_Y + 6 * rr + s] * tmp[6 * k + s];
        z[18 * k + 12 + rr] = acc;
        if (zu_out) zu_out[6 * (W.k0 + k) + rr] = acc;
    }
    __syncthreads();
}

// Duals of the equality rows at the QP solution z, in the row order of the reference's A (src/osqp_solver.py:83-101:
// block 0 = -x_0, block k+1 = A_k x_k + B_k u_k - x_{k+1}).  Stationarity in x_j reads
// P_j x_j + g_j - y_j + A_j' y_{j+1} = 0, so y_j = P_j z_j + g_j + A_j' y_{j+1} (y_N = 0): the co-states of the backward
// sweep, recovered here by one more pass over the knots (warp 0, lane c = component of y_j; debugging / parity entry).
template <typename T>
__device__ void qp_duals(const SolveParams<T>& P, const T* z, const T* xu, const T* cost, const T* krb, int kstride, int oGQ,
                         int oGV, T* y_out) {
    if (threadIdx.x >= 32) return;
    const int lane = threadIdx.x, c = lane < 12 ? lane : 0, N = P.N;
    const int ca = c < 6 ? c : c - 6;
    T y = T(0);
    for (int j = N - 1; j >= 0; --j) {
        const T* ck = cost + COST_STRIDE * j;
        const T* zj = z + 18 * j;
        T own;
        if (c < 6) {
            T jz = T(1);
            for (int i = 0; i < 6; ++i) jz += ck[i] * zj[i];
            own = ck[6] * ck[c] * jz;                       // Q_k j (j'z) + Q_k j
        } else {
            own = ck[7] * (zj[c] + xu[18 * j + c]);         // dQ_k z_v + dQ_k v_bar
        }
        T acc = T(0);
        const T yq = __shfl_sync(0xffffffffu, y, ca);       // y_{j+1}, q part, component ca
        if (j < N - 1) {
            const T* G = krb + (size_t)kstride * j + (c < 6 ? oGQ : oGV);
            acc = c < 6 ? yq : P.dt * yq;
#pragma unroll
            for (int r = 0; r < 6; ++r) acc += G[6 * r + ca] * __shfl_sync(0xffffffffu, y, 6 + r);
        } else {
#pragma unroll
            for (int r = 0; r < 6; ++r) (void)__shfl_sync(0xffffffffu, y, 6 + r);
        }
        y = own + acc;
        if (lane < 12) y_out[12 * j + lane] = y;
    }
}

// -----------------------------------------------------------------------------------------------------------
// fixed-order reduction helper: sum of vals[0..n) by one warp (lane-strided partial sums + xor tree)
template <typename T> I7_DEV T warp_sum_strided(const T* vals, int n, int lane) {
    T s = T(0);
    for (int i = lane; i < n; i += 32) s += vals[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    return s;
}

// Merit terms of the candidates c0 .. c0+nc-1 (index into c_alphas; 0 = base) of the line  xa + alpha (xb - xa) at the
// knots of the window W:  comps[4 * ((cbase + ci) * N + k) + {0,1,2,3}] = {qcost, vcost, ucost, integrator_err} of knot k
// (k = horizon index; src/osqp_sqp.py:13-47).  One thread per point: kinematic chain, link records (world frame),
// M and b from composite inertias, one 6x6 LDL' solve.  scr: MP_STRIDE numbers of scratch per point.
template <typename T, int NT>
__device__ void merit_points(const Robot<T>& rob, const SolveParams<T>& P, const Ptrs<T>& M, const Win W, T* scr,
                             const T* xa_, const T* xb_, const T* fext, int c0, int nc, T* comps, int cbase) {
    const int tid = threadIdx.x, N = P.N, nk = W.nd, nw = W.nc;
    const int npts = nc * nw;
    const T* xa = xa_ + 18 * W.k0;
    const T* xb = xb_ + 18 * W.k0;
    // M0: cos / sin of the joint angles of every point (parked in the point's scratch, see wf_point_forward_dynamics)
    for (int t = tid; t < 6 * npts; t += NT) {
        const int i = t / npts, p = t - i * npts;
        const int ci = p / nw, k = p - ci * nw;
        const T alpha = T(c_alphas[c0 + ci]);
        const T x = xa[18 * k + i] + alpha * (xb[18 * k + i] - xa[18 * k + i]);
        T sn, cs;
        sincos_t<T>(x, &sn, &cs);
        T* sp = scr + MP_STRIDE * p;
        sp[wp_cs(i)] = cs; sp[wp_cs(i) + 1] = sn;
    }
    __syncthreads();
    // M1: per point
    for (int p = tid; p < npts; p += NT) {
        const int ci = p / nw, k = p - ci * nw;
        const int kg = W.k0 + k;
        const T alpha = T(c_alphas[c0 + ci]);
        T* sp = scr + MP_STRIDE * p;
        T pee[3];
        wf_tip_scr<T>(rob, sp, pee);
        const T* goal = M.goals + 3 * kg;
        const T e[3] = {pee[0] - goal[0], pee[1] - goal[1], pee[2] - goal[2]};
        const T Qk = (kg == N - 1) ? P.QN : T(1);
        T* out = comps + 4 * ((cbase + ci) * N + kg);
        out[0] = Qk * (e[0] * e[0] + e[1] * e[1] + e[2] * e[2]);
        const bool has_u = (kg < N - 1);                    // the last knot carries no control in XU
        T qv[6], u[6];
        T vv = T(0), uu = T(0);
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            qv[i] = xa[18 * k + 6 + i] + alpha * (xb[18 * k + 6 + i] - xa[18 * k + 6 + i]);
            u[i] = has_u ? xa[18 * k + 12 + i] + alpha * (xb[18 * k + 12 + i] - xa[18 * k + 12 + i]) : T(0);
            vv += qv[i] * qv[i]; uu += u[i] * u[i];
        }
        out[1] = P.dQ * vv;
        out[2] = P.R * uu;
        T ie = T(0);
        if (k < nk) {
            T qdd[6];
            wf_point_forward_dynamics<T>(rob, sp, qv, u, fext, pee, qdd);
            T qe = T(0), ve = T(0);
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                const T q = xa[18 * k + i] + alpha * (xb[18 * k + i] - xa[18 * k + i]);
                const T xnq = xa[18 * (k + 1) + i] + alpha * (xb[18 * (k + 1) + i] - xa[18 * (k + 1) + i]);
                const T xnv = xa[18 * (k + 1) + 6 + i] + alpha * (xb[18 * (k + 1) + 6 + i] - xa[18 * (k + 1) + 6 + i]);
                const T dq = q + qv[i] * P.dt - xnq;
                const T dv = qv[i] + qdd[i] * P.dt - xnv;
                qe += dq * dq; ve += dv * dv;
            }
            ie = sqrt_t<T>(qe) + sqrt_t<T>(ve);
        }
        out[3] = ie;
    }
    __syncthreads();
}

// merits[c] for c in [c0, c0 + nc): sum over all N knots (fixed order) + initial-state violation term.
// CG: comps live in the global exchange buffer written by other CTAs of the cluster (read with ld.global.cg).
template <typename T, int NT, bool CG>
__device__ void merit_reduce(const SolveParams<T>& P, const T* xa, const T* xb, int c0, int nc, const T* comps, int cbase,
                             T* merits, T* comps_out /* 4 sums of candidate c0 or nullptr */) {
    const int tid = threadIdx.x, N = P.N;
    const int warp = tid >> 5, lane = tid & 31;
    for (int ci = warp; ci < nc; ci += NT / 32) {
        T s[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            T a = T(0);
            for (int k = lane; k < N; k += 32) a += CG ? __ldcg(comps + 4 * ((cbase + ci) * N + k) + j) : comps[4 * ((cbase + ci) * N + k) + j];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
            s[j] = a;
        }
        if (lane == 0) {
            const T alpha = T(c_alphas[c0 + ci]);
            T d2 = T(0);
            for (int i = 0; i < 12; ++i) { const T d = alpha * (xb[i] - xa[i]); d2 += d * d; }
            merits[c0 + ci] = s[0] + s[1] + s[2] + P.mu * (s[3] + sqrt_t<T>(d2));
            if (comps_out && ci == 0) { comps_out[0] = s[0]; comps_out[1] = s[1]; comps_out[2] = s[2]; comps_out[3] = s[3]; }
        }
    }
    __syncthreads();
}

// -----------------------------------------------------------------------------------------------------------
struct BlockRecLayout {       // record offsets of the CTA-per-problem kernel, for dump_qp
    static constexpr int GQ = KR_GQ, GV = KR_GV, CV = KR_CV, S = KR_S, STRIDE = KR_STRIDE;
};
template <typename T, typename Lay> __device__ void dump_qp(const SolveParams<T>& P, const Ptrs<T>& M, int b) {
    const int N = P.N, tid = threadIdx.x, NT = blockDim.x;
    const T* xu = M.xu;
    const int nA = 360 * (N - 1) + 144, nP = 27 * N + 6 * (N - 1), nz = 18 * N - 6, nl = 12 * N;
    if (P.Adata) {
        T* A = P.Adata + (size_t)b * nA;
        for (int t = tid; t < nA; t += NT) {
            T v;
            const int k = t / 360, e = t - 360 * k;
            if (k >= N - 1) {
                const int e2 = t - 360 * (N - 1);
                v = (e2 / 12 == e2 % 12) ? T(-1) : T(0);
            } else if (e < 288) {
                const int c = e / 24, r = e - 24 * c;
                const T* kr = M.kr + (size_t)Lay::STRIDE * k;
                if (r < 12) v = (r == c) ? T(-1) : T(0);
                else {
                    const int rr = r - 12;
                    if (rr < 6) v = (c < 6) ? (rr == c ? T(1) : T(0)) : (rr == c - 6 ? P.dt : T(0));
                    else v = (c < 6) ? kr[Lay::GQ + 6 * (rr - 6) + c] : kr[Lay::GV + 6 * (rr - 6) + c - 6];
                }
            } else {
                const int e2 = e - 288, c = e2 / 12, r = e2 - 12 * c;
                const T* kr = M.kr + (size_t)Lay::STRIDE * k;
                v = (r < 6) ? T(0) : kr[Lay::S + 6 * (r - 6) + c];      // H = dt Minv (dump mode of phase_setup)
            }
            A[t] = v;
        }
    }
    if (P.l) {
        T* l = P.l + (size_t)b * nl;
        for (int t = tid; t < nl; t += NT) {
            const int k = t / 12, r = t - 12 * k;
            T v;
            if (k == 0) v = -M.x0[r];
            else v = (r < 6) ? T(0) : -M.kr[(size_t)Lay::STRIDE * (k - 1) + Lay::CV + r - 6];
            l[t] = v;
        }
    }
    if (P.Pdata) {
        T* Pd = P.Pdata + (size_t)b * nP;
        for (int t = tid; t < nP; t += NT) {
            int k = t / 33, e = t - 33 * k;        // 33 values per knot except the last (27)
            const T* ck = M.cost + COST_STRIDE * k;
            T v;
            if (e < 21) {
                int r = 0, ee = e;
                while (ee > r) { ee -= r + 1; ++r; }
                v = ck[6] * ck[r] * ck[ee];
            } else if (e < 27) v = ck[7];
            else v = ck[8];
            Pd[t] = v;
        }
    }
    if (P.g) {
        T* g = P.g + (size_t)b * nz;
        for (int t = tid; t < nz; t += NT) {
            const int k = t / 18, e = t - 18 * k;
            const T* ck = M.cost + COST_STRIDE * k;
            T v;
            if (e < 6) v = ck[6] * ck[e];
            else if (e < 12) v = ck[7] * xu[t];
            else v = ck[8] * xu[t];
            g[t] = v;
        }
    }
}

// -----------------------------------------------------------------------------------------------------------
template <typename T, int NT, bool SPILL>
// resident CTAs per SM the register budget is sized for: 2 (shared memory allows no more in fp64); fp32 records are half
// the size, so four 128-thread CTAs fit
__global__ void __launch_bounds__(NT, NT == 256 ? 2 : ((NT == 128 && sizeof(T) == 4) ? 4 : 1)) sqp_kernel(const __grid_constant__ Robot<T> rob, const __grid_constant__ SolveParams<T> P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* sm = reinterpret_cast<T*>(smem_raw);
    const int tid = threadIdx.x, N = P.N;
    const int nz = 18 * N - 6;
    const Layout L = make_layout(N, SPILL);
    Ptrs<T> M;
    M.xu = sm + L.xu; M.z = sm + L.z; M.goals = sm + L.goals; M.x0 = sm + L.x0; M.cost = sm + L.cost;
    M.work = sm + L.work; M.ls = sm + L.ls;
    T* const krg = SPILL ? P.spill_ws + (size_t)blockIdx.x * L.reg_size : nullptr;
    M.kr = SPILL ? krg : sm + L.reg;
    M.aux = M.z; M.pair = nullptr;
    {
        int cap = L.reg_size / (MP_STRIDE * N);
        M.merit_cap = SPILL ? 1 : (cap > 5 ? 5 : (cap < 1 ? 1 : cap));
    }
    __shared__ T s_merits[9];
    __shared__ T s_alpha;
    __shared__ T s_merit_acc;
    __shared__ T s_fext[6];
    __shared__ int s_flag;
    const Win W{0, N - 1, N};

    for (int b = blockIdx.x; b < P.batch; b += gridDim.x) {
        if (P.done && P.done[b]) continue;                           // multi-launch SQP loop: this problem has converged
        // ---- load inputs (every global load issued before the first store: one memory round trip, not one per
        //      pass of a load-store loop; the inputs may sit in mapped HOST memory, zero-copy path) ----
        {
            const T* xin = P.xu_in + (size_t)b * nz;
            for (int base = 0; base < nz; base += 6 * NT) {
                T r[6];
#pragma unroll
                for (int q = 0; q < 6; ++q) { const int t = base + q * NT + tid; r[q] = t < nz ? xin[t] : T(0); }
#pragma unroll
                for (int q = 0; q < 6; ++q) { const int t = base + q * NT + tid; if (t < nz) M.xu[t] = r[q]; }
            }
        }
        if (tid < 6) M.xu[nz + tid] = T(0);                          // padded u_{N-1}
        for (int t = tid; t < 3 * N; t += NT) {
            const int k = t / 3, a = t - 3 * k;
            M.goals[t] = P.goals[(size_t)b * N * P.goal_stride + k * P.goal_stride + a];
        }
        if (tid < 12) M.x0[tid] = P.x0 ? P.x0[(size_t)b * 12 + tid] : T(0);
        if (P.x0_into_xu && P.x0 && tid < 12) M.xu[tid] = P.x0[(size_t)b * 12 + tid];      // same thread as the load of xu[tid] above
        if (tid < 6) s_fext[tid] = P.fext ? P.fext[(size_t)b * 6 + tid] : T(0);
        if (P.mode == MODE_LINESEARCH || P.mode == MODE_MERIT || P.mode == MODE_STEP) {
            for (int t = tid; t < nz; t += NT) M.z[t] = P.xu_full[(size_t)b * nz + t];
            if (tid < 6) M.z[nz + tid] = T(0);
        }
        __syncthreads();
        const T* fext = P.fext ? s_fext : nullptr;

        if (P.mode == MODE_MERIT) {
            // merit terms of xu; the initial-state term is taken against the reference iterate z
            T* comps = M.ls;
            if constexpr (SPILL) {
                constexpr int TILE_PTS = (TILE_K * KR_STRIDE) / MP_STRIDE;
                for (int k0 = 0; k0 < N; k0 += TILE_PTS) {
                    const Win Wm{k0, N - 1 - k0 < TILE_PTS ? N - 1 - k0 : TILE_PTS, N - k0 < TILE_PTS ? N - k0 : TILE_PTS};
                    merit_points<T, NT>(rob, P, M, Wm, sm + L.tile, M.xu, M.xu, fext, 0, 1, comps, 0);
                }
            } else
            merit_points<T, NT>(rob, P, M, W, M.kr, M.xu, M.xu, fext, 0, 1, comps, 0);
            merit_reduce<T, NT, false>(P, M.xu, M.xu, 0, 1, comps, 0, s_merits, M.work);
            if (tid == 0) {
                T d2 = T(0);
                for (int i = 0; i < 12; ++i) { const T d = M.xu[i] - M.z[i]; d2 += d * d; }
                T* o = P.merits_out + (size_t)b * 6;
                const T* part = M.work;
                o[0] = part[0]; o[1] = part[1]; o[2] = part[2]; o[3] = part[3]; o[4] = sqrt_t<T>(d2);
                o[5] = part[0] + part[1] + part[2] + P.mu * (part[3] + o[4]);
            }
            __syncthreads();
            continue;
        }

        int qp_iters = 0, n_ls = 0, n_steps = 0, status = 0;
        T alphas[2] = {T(0), T(0)}, steps[2] = {T(0), T(0)};
        T merit_final = T(0), merit_first = T(0), qp_its[2] = {T(0), T(0)};
        const bool step_mode = (P.mode == MODE_STEP);
        const int qoff = step_mode ? P.qp_index : 0;                 // MODE_STEP runs ONE pass: number qp_index
        if (step_mode && qoff > 0 && P.stats) {                      // the record so far (written by the previous pass)
            const T* st = P.stats + (size_t)b * 12;
            n_ls = (int)st[1]; n_steps = (int)st[2]; status = (int)st[3];
            alphas[0] = st[4]; steps[0] = st[6]; steps[1] = st[7]; merit_first = st[9]; qp_its[0] = st[10];
        }
        if (step_mode && P.qp_info && qoff < 2) qp_its[qoff] = P.qp_info[(size_t)b * 4];
        for (int qp = 0; qp < (step_mode ? 1 : P.max_qp); ++qp) {
            qp_iters = qp + qoff + 1;
            if (P.prof && tid == 0) P.prof[(size_t)b * 32 + 5 * qp + 0] = clock64();
            if (P.mode != MODE_LINESEARCH && !step_mode) {
                if constexpr (SPILL) {
                    for (int k0 = 0; k0 < N; k0 += TILE_K) {
                        const Win Wt{k0, N - 1 - k0 < TILE_K ? N - 1 - k0 : TILE_K, N - k0 < TILE_K ? N - k0 : TILE_K};
                        Ptrs<T> Mt = M;
                        Mt.kr = sm + L.tile; Mt.cost = M.cost + COST_STRIDE * k0;
                        phase_setup<T, NT>(rob, P, Mt, Wt, fext, P.mode == MODE_DUMP || P.mode == MODE_EXPORT,
                                           (P.prof && qp == 0 && k0 == 0) ? P.prof + (size_t)b * 32 : nullptr);
                        T* dst = krg + (size_t)KR_STRIDE * k0;
                        for (int t = tid; t < Wt.nc * KR_STRIDE; t += NT) dst[t] = Mt.kr[t];
                        __syncthreads();
                    }
                } else {
                    phase_setup<T, NT>(rob, P, M, W, fext, P.mode == MODE_DUMP || P.mode == MODE_EXPORT, (P.prof && qp == 0) ? P.prof + (size_t)b * 32 : nullptr);
                }
                if (P.prof && tid == 0) P.prof[(size_t)b * 32 + 5 * qp + 1] = clock64();
                if (P.mode == MODE_DUMP) { dump_qp<T, BlockRecLayout>(P, M, b); break; }
                if (P.mode == MODE_EXPORT) {
                    T* o = P.blk_out + (size_t)b * N * QB_STRIDE;
                    for (int t = tid; t < N * QB_STRIDE; t += NT) {
                        const int k = t / QB_STRIDE, e = t - QB_STRIDE * k;
                        const T* kr = M.kr + (size_t)KR_STRIDE * k;
                        T v = T(0);
                        if (e >= QB_COST) { if (e < QB_COST + COST_STRIDE) v = M.cost[COST_STRIDE * k + e - QB_COST]; }
                        else if (k < N - 1) v = e < QB_GV ? kr[KR_GQ + e] : (e < QB_CV ? kr[KR_GV + e - QB_GV] : (e < QB_HM ? kr[KR_CV + e - QB_CV] : kr[KR_S + e - QB_HM]));
                        o[t] = v;
                    }
                    break;
                }
                {
                    long long* pr = (P.prof && qp == 0) ? P.prof + (size_t)b * 32 : nullptr;
                    long long tc = 0;
                    riccati_sweep<T, NT, RicLocal<T, SPILL>>(P, L.work, L.z, L.x0, L.xu, RicLocal<T, SPILL>{L.reg, L.cost, krg, L.tile, TILE_K * KR_STRIDE}, pr);
                    if (pr && tid == 0) tc = clock64();
                    phase_controls<T, NT, SPILL>(P, L.z, L.ls, L.reg, krg, W, nullptr);
                    if (pr && tid == 0) pr[13] = clock64() - tc;
                }
                if (P.prof && tid == 0) P.prof[(size_t)b * 32 + 5 * qp + 2] = clock64();
                if (P.mode == MODE_QP) {
                    for (int t = tid; t < nz; t += NT) P.z_out[(size_t)b * nz + t] = M.z[t];
                    if (P.y_out) qp_duals<T>(P, M.z, M.xu, M.cost, M.kr, KR_STRIDE, KR_GQ, KR_GV, P.y_out + (size_t)b * 12 * N);
                    break;
                }
            }
            // ---- line search: candidates evaluated in waves, stopping at the first wave that holds an accepted
            //      alpha (src/osqp_sqp.py:59-72 takes the first alpha, in decreasing order, with merit <= base) ----
            {
                const bool all = (P.mode == MODE_LINESEARCH);     // the debug entry point reports all nine merits
                int c0 = 0;
                if (tid == 0) { s_alpha = T(0); s_flag = 0; }
                while (c0 < 9) {
                    int nc = 9 - c0 < M.merit_cap ? 9 - c0 : M.merit_cap;
                    if constexpr (SPILL) {
                        constexpr int TILE_PTS = (TILE_K * KR_STRIDE) / MP_STRIDE;
                        const int nwin = (N + TILE_PTS - 1) / TILE_PTS, wlen = (N + nwin - 1) / nwin;    // equal windows
                        for (int ci = 0; ci < nc; ++ci)
                            for (int k0 = 0; k0 < N; k0 += wlen) {
                                const Win Wm{k0, N - 1 - k0 < wlen ? N - 1 - k0 : wlen, N - k0 < wlen ? N - k0 : wlen};
                                merit_points<T, NT>(rob, P, M, Wm, sm + L.tile, M.xu, M.z, fext, c0 + ci, 1, M.ls, ci);
                            }
                    } else {
                        merit_points<T, NT>(rob, P, M, W, M.kr, M.xu, M.z, fext, c0, nc, M.ls, 0);
                    }
                    merit_reduce<T, NT, false>(P, M.xu, M.z, c0, nc, M.ls, 0, s_merits, nullptr);
                    if (tid == 0 && s_alpha == T(0)) {
                        if (!(s_merits[0] == s_merits[0])) s_flag = 1;
                        for (int c = (c0 == 0 ? 1 : c0); c < c0 + nc; ++c)
                            if (s_merits[c] <= s_merits[0]) { s_alpha = T(c_alphas[c]); s_merit_acc = s_merits[c]; break; }
                    }
                    __syncthreads();
                    c0 += nc;
                    if (!all && s_alpha != T(0)) break;
                }
                if (tid == 0 && s_alpha == T(0)) s_merit_acc = s_merits[0];
                __syncthreads();
            }
            if (P.prof && tid == 0) P.prof[(size_t)b * 32 + 5 * qp + 3] = clock64();
            const T alpha = s_alpha;
            merit_final = s_merit_acc;
            if (qp + qoff == 0) merit_first = s_merits[0];
            if (s_flag) status = 2;
            if (qp + qoff < 2) alphas[qp + qoff] = alpha;
            ++n_ls;
            if (P.mode == MODE_LINESEARCH) {
                if (tid == 0) P.alpha_out[b] = alpha;
                if (tid < 9) P.merits_out[(size_t)b * 9 + tid] = s_merits[tid];
                break;
            }
            if (alpha == T(0)) {
                // src/osqp_sqp.py:81-82 `continue`: the next pass would re-solve the identical QP and, with an
                // exact QP solve, reach the same alpha = 0; the remaining passes are skipped but still counted.
                // (MODE_STEP: the QP solver is iterative and warm-started, the next pass really re-solves, as the reference does)
                if (status == 0) status = 1;
                if (!step_mode)
                    for (int q2 = qp + 1; q2 < P.max_qp; ++q2) { if (q2 < 2) alphas[q2] = T(0); ++n_ls; qp_iters = q2 + 1; }
                break;
            }
            __syncthreads();
            // XU += alpha (z - XU); step norm
            T* part = M.ls;
            T loc = T(0);
            for (int t = tid; t < nz; t += NT) {
                const T st = alpha * (M.z[t] - M.xu[t]);
                M.xu[t] += st;
                loc += st * st;
            }
            part[tid] = loc;
            __syncthreads();
            if (tid < 32) {
                T s = warp_sum_strided<T>(part, NT, tid);
                if (tid == 0) s_merit_acc = sqrt_t<T>(s);
            }
            __syncthreads();
            const T stepsize = s_merit_acc;
            if (P.prof && tid == 0) P.prof[(size_t)b * 32 + 5 * qp + 4] = clock64();
            if (n_steps < 2) steps[n_steps] = stepsize;
            ++n_steps;
            __syncthreads();
            if (stepsize < P.step_tol) {
                if (step_mode && P.done && tid == 0) P.done[b] = 1;
                break;
            }
        }
        if (P.mode == MODE_SOLVE || step_mode) {
            for (int t = tid; t < nz; t += NT) P.xu_out[(size_t)b * nz + t] = M.xu[t];
            if (P.stats && tid == 0) {
                T* st = P.stats + (size_t)b * 12;
                st[0] = T(qp_iters); st[1] = T(n_ls); st[2] = T(n_steps); st[3] = T(status);
                st[4] = alphas[0]; st[5] = alphas[1]; st[6] = steps[0]; st[7] = steps[1];
                st[8] = merit_final; st[9] = merit_first; st[10] = qp_its[0]; st[11] = qp_its[1];
            }
        }
        __syncthreads();
        if (P.mode == MODE_SOLVE && P.xchg.enabled) {
            // last-problem-done: every CTA publishes its results (fence), takes a ticket; the holder of the last ticket
            // sees all of them (fence after the atomic) and runs the selection + exchange
            __threadfence();
            __syncthreads();
            if (tid == 0) s_flag = (atomicAdd(P.xchg.ticket, 1u) == (unsigned)P.batch - 1u) ? 1 : 0;
            __syncthreads();
            if (s_flag) {
                if (tid == 0) *P.xchg.ticket = 0u;
                __threadfence();
                pack_exchange_reduce<T>(P.batch, nz, P.stats, 12, 8, P.xu_out, P.xchg.index_offset, P.xchg.rank, P.xchg.world,
                                        P.xchg.peers, P.xchg.seq, P.xchg.timeout_ns, static_cast<T*>(P.xchg.out), P.xchg.err,
                                        P.xchg.wait_ns);
            }
            __syncthreads();
        }
    }
}

}  // namespace i7
