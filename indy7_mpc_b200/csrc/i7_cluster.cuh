// i7_cluster.cuh — latency form of the fused SQP-MPC solve: one thread-block CLUSTER of 8 CTAs (8 SMs) per problem.
//
// With one or a handful of problems the CTA-per-problem kernel leaves 147 SMs idle while one SM grinds through the
// per-knot phases at the rate of its FP64 pipe.  Here the knots of a problem are dealt out to the 8 CTAs of a
// cluster (contiguous windows of ceil(N/8) knots): linearisation, cost terms and the stage-local QP products
// (setup), the controls and every line-search candidate (merit) run on 8 SMs at once; only the inherently serial
// backward sweep + roll-out runs on the lead CTA.  The CTAs meet at three cluster barriers per SQP iteration and
// exchange data through a small L2-resident buffer (barrier.cluster release/acquire orders the global accesses;
// cross-CTA reads use ld.global.cg):
//     setup  -> export [L | SG | S | d | hx | cost] per knot ............ barrier 1
//     lead CTA: backward sweep (streams the exported records) + roll-out -> publish the states x_k ... barrier 2
//     all: controls u_k of the own knots (published) + merit terms of all 9 candidates at the own knots .. barrier 3
//     all (redundantly, bit-identical): sum the merit terms, pick alpha, update XU, step norm, stop test
// Same arithmetic per knot / per stage as the CTA-per-problem kernel (same device functions), reference call
// stack src/osqp_sqp.py:76-93.
#pragma once
#include <cooperative_groups.h>

#include "i7_sqp.cuh"

namespace i7 {

// CTAs per cluster = per problem: template parameter CL in {8 (portable maximum), 4, 2}; the smaller clusters serve
// batches that no longer fit the device as clusters of 8 (batch 16 / 32 / 64)
constexpr int CL_NT = 384;     // threads per CTA

__host__ __device__ inline int cluster_merit_cap(int chunk) { return chunk <= 8 ? 8 : (chunk <= 16 ? 4 : 2); }
// Line-search points of the latency forms: instead of one thread per point walking the whole chain (a 20 k-cycle
// dependency chain), the dynamics of a point are split into the task waves of the setup — (point, joint) sin / cos,
// (point) tip, (point, link) link records, (point, column) mass-matrix column + bias, (point) 6x6 solve + merit terms —
// about half the latency; needs PR_STRIDE numbers of scratch per point instead of MP_STRIDE.
constexpr int PR_CS = 0, PR_PEE = 12, PR_QD = 15, PR_U = 21, PR_S = 27, PR_F = 63, PR_I = 99, PR_M = 153, PR_B = 189, PR_STRIDE = 197;
struct CLayout { int xu, z, goals, x0, cost, work, ls, kr, aux, pair, gains, total, chunk, par_merit; };
__host__ __device__ inline CLayout make_clayout(int N, int CL) {
    CLayout L;
    L.chunk = (N + CL - 1) / CL;
    // knot records, later the merit scratch (as many line-search candidates per wave as fit: 8 up to N = 32)
    const int mcap = cluster_merit_cap(L.chunk);
    // task-parallel line-search points while a wave of them (candidates x knots of the window, 6 tasks each) fits the
    // CTA's threads and its scratch stays moderate
    // (and the whole layout stays within an SM's shared memory: long horizons keep the one-thread-per-point form)
    const int fixed = WK_SIZE + 36 * N + 3 * N + 16 + COST_STRIDE * L.chunk + CL_NT + 16 + AUX_STRIDE * L.chunk +
                      (L.chunk <= 8 ? 432 * L.chunk : 0) + GAIN_STRIDE * N + 8;
    L.par_merit = (6 * mcap * L.chunk <= CL_NT && fixed + mcap * L.chunk * PR_STRIDE <= 27500) ? 1 : 0;
    const int mw = L.par_merit ? mcap * PR_STRIDE : mcap * MP_STRIDE;
    const int recw = KR_STRIDE > mw ? KR_STRIDE : mw;
    int o = 0;
    L.work = o; o += WK_SIZE;
    L.xu = o; o += 18 * N;
    L.z = o; o += 18 * N;
    L.goals = o; o += 3 * N + (N & 1);
    L.x0 = o; o += 12;
    L.cost = o; o += COST_STRIDE * L.chunk + ((COST_STRIDE * L.chunk) & 1);
    L.ls = o; o += CL_NT + 16;
    L.kr = o; o += recw * L.chunk + ((recw * L.chunk) & 1);
    L.aux = o; o += AUX_STRIDE * L.chunk;
    L.pair = -1;
    if (L.chunk <= 8) { L.pair = o; o += 432 * L.chunk; }     // per-(column, link) derivative terms (latency form)
    L.gains = o; o += GAIN_STRIDE * N;
    L.total = o;
    return L;
}
// exchange buffer of one problem (global, L2-resident): ric | zx | zu | mcomps
__host__ __device__ inline int xbuf_ric(int) { return 0; }
__host__ __device__ inline int xbuf_zx(int N) { return RIC_STRIDE * N; }
__host__ __device__ inline int xbuf_zu(int N) { return RIC_STRIDE * N + 12 * N; }
__host__ __device__ inline int xbuf_mc(int N) { return RIC_STRIDE * N + 18 * N; }
__host__ __device__ inline int xbuf_size(int N) { return RIC_STRIDE * N + 18 * N + 36 * N; }

// Merit terms of candidates c0 .. c0+nc-1 at the knots of window W, task-parallel form (same outputs as merit_points).
template <typename T, int NT>
__device__ void merit_points_par(const Robot<T>& rob, const SolveParams<T>& P, const Ptrs<T>& M, const Win W, T* scr,
                                 const T* xa_, const T* xb_, const T* fext, int c0, int nc, T* comps, int cbase) {
    const int tid = threadIdx.x, N = P.N, nk = W.nd, nw = W.nc;
    const int npts = nc * nw;
    const T* xa = xa_ + 18 * W.k0;
    const T* xb = xb_ + 18 * W.k0;
    // P0: per (point, i): interpolated q_i -> cos / sin; qd_i, u_i (u = 0 at the last knot of the horizon)
    for (int t = tid; t < 6 * npts; t += NT) {
        const int i = t / npts, p = t - i * npts;
        const int ci = p / nw, k = p - ci * nw;
        const T alpha = T(c_alphas[c0 + ci]);
        const T* a = xa + 18 * k;
        const T* b = xb + 18 * k;
        T sn, cs;
        sincos_t<T>(a[i] + alpha * (b[i] - a[i]), &sn, &cs);
        T* pr = scr + PR_STRIDE * p;
        pr[PR_CS + i] = cs; pr[PR_CS + 6 + i] = sn;
        pr[PR_QD + i] = a[6 + i] + alpha * (b[6 + i] - a[6 + i]);
        pr[PR_U + i] = (W.k0 + k < N - 1) ? a[12 + i] + alpha * (b[12 + i] - a[12 + i]) : T(0);
    }
    __syncthreads();
    // P1: per point: position of the last joint (spatial origin of the point; eepos of the cost)
    for (int p = tid; p < npts; p += NT) {
        T* pr = scr + PR_STRIDE * p;
        T pee[3];
        wf_tip<T>(rob, pr + PR_CS, pee);
        pr[PR_PEE] = pee[0]; pr[PR_PEE + 1] = pee[1]; pr[PR_PEE + 2] = pee[2];
    }
    __syncthreads();
    // P2: per (point, link): axis, world inertia, force at qdd = 0 (velocity / acceleration stay in registers)
    for (int t = tid; t < 6 * npts; t += NT) {
        const int l = t / npts, p = t - l * npts;
        const int k = p % nw;
        if (k >= nk) continue;
        T* pr = scr + PR_STRIDE * p;
        T v[6], a[6];
        wf_link<T>(rob, pr + PR_CS, pr + PR_QD, fext, pr + PR_PEE, l, pr + PR_S + 6 * l, v, a, pr + PR_I + 9 * l, pr + PR_F + 6 * l);
    }
    __syncthreads();
    // P3: per (point, j): column j of M and the bias torque b_j
    for (int t = tid; t < 6 * npts; t += NT) {
        const int j = t / npts, p = t - j * npts;
        const int k = p % nw;
        if (k >= nk) continue;
        T* pr = scr + PR_STRIDE * p;
        T bj;
        wf_mass_column<T>(rob, pr + PR_S, pr + PR_I, pr + PR_F, j, pr + PR_M, &bj);
        pr[PR_B + j] = bj;
    }
    __syncthreads();
    // P4: per point: qdd = M^-1 (u - b), merit terms (src/osqp_sqp.py:13-47)
    for (int p = tid; p < npts; p += NT) {
        const int ci = p / nw, k = p - ci * nw;
        const int kg = W.k0 + k;
        const T alpha = T(c_alphas[c0 + ci]);
        const T* pr = scr + PR_STRIDE * p;
        const T* goal = M.goals + 3 * kg;
        const T e[3] = {pr[PR_PEE] - goal[0], pr[PR_PEE + 1] - goal[1], pr[PR_PEE + 2] - goal[2]};
        const T Qk = (kg == N - 1) ? P.QN : T(1);
        T* out = comps + 4 * ((cbase + ci) * N + kg);
        out[0] = Qk * (e[0] * e[0] + e[1] * e[1] + e[2] * e[2]);
        T vv = T(0), uu = T(0);
#pragma unroll
        for (int i = 0; i < 6; ++i) { vv += pr[PR_QD + i] * pr[PR_QD + i]; uu += pr[PR_U + i] * pr[PR_U + i]; }
        out[1] = P.dQ * vv;
        out[2] = P.R * uu;
        T ie = T(0);
        if (k < nk) {
            T A[6][6], dinv[6], qdd[6];
#pragma unroll
            for (int r = 0; r < 6; ++r)
#pragma unroll
                for (int q = 0; q <= r; ++q) A[r][q] = pr[PR_M + 6 * r + q];
#pragma unroll
            for (int r = 0; r < 6; ++r) qdd[r] = pr[PR_U + r] - pr[PR_B + r];
            spd6_factor<T>(A, dinv);
            spd6_solve_factored<T>(A, dinv, qdd);
            const T* a = xa + 18 * k;
            const T* b = xb + 18 * k;
            T qe = T(0), ve = T(0);
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                const T q = a[i] + alpha * (b[i] - a[i]);
                const T xnq = a[18 + i] + alpha * (b[18 + i] - a[18 + i]);
                const T xnv = a[24 + i] + alpha * (b[24 + i] - a[24 + i]);
                const T dq = q + pr[PR_QD + i] * P.dt - xnq;
                const T dv = pr[PR_QD + i] + qdd[i] * P.dt - xnv;
                qe += dq * dq; ve += dv * dv;
            }
            ie = sqrt_t<T>(qe) + sqrt_t<T>(ve);
        }
        out[3] = ie;
    }
    __syncthreads();
}

template <typename T, int CL>
__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(CL_NT)
sqp_cluster_kernel(const __grid_constant__ Robot<T> rob, const __grid_constant__ SolveParams<T> P, T* xbuf_all) {
    constexpr int NT = CL_NT;
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* sm = reinterpret_cast<T*>(smem_raw);
    const int tid = threadIdx.x, N = P.N;
    const int rank = (int)cluster.block_rank();
    const int b = blockIdx.x / CL;
    const int nz = 18 * N - 6;
    const CLayout L = make_clayout(N, CL);
    Ptrs<T> M;
    M.xu = sm + L.xu; M.z = sm + L.z; M.goals = sm + L.goals; M.x0 = sm + L.x0; M.cost = sm + L.cost;
    M.work = sm + L.work; M.ls = sm + L.ls; M.kr = sm + L.kr; M.aux = sm + L.aux; M.pair = L.pair >= 0 ? sm + L.pair : nullptr; M.merit_cap = cluster_merit_cap(L.chunk);
    T* xbuf = xbuf_all + (size_t)b * xbuf_size(N);
    T* ric = xbuf + xbuf_ric(N);
    T* zx = xbuf + xbuf_zx(N);
    T* zu = xbuf + xbuf_zu(N);
    T* mcomps = xbuf + xbuf_mc(N);
    Win W;
    W.k0 = rank * L.chunk;
    W.nc = N - W.k0 < 0 ? 0 : (N - W.k0 > L.chunk ? L.chunk : N - W.k0);
    W.nd = N - 1 - W.k0 < 0 ? 0 : (N - 1 - W.k0 > L.chunk ? L.chunk : N - 1 - W.k0);
    __shared__ T s_merits[9];
    __shared__ T s_alpha;
    __shared__ T s_merit_acc;
    __shared__ T s_fext[6];
    __shared__ int s_flag;
    long long* prof = (P.prof && rank == 0) ? P.prof + (size_t)b * 32 : nullptr;

    // ---- load inputs (every CTA keeps the whole iterate; it is 4.6 KB).  All global loads are issued before the first
    //      use: the inputs may sit in mapped HOST memory (small host-buffer solves), where a dependent load-store
    //      sequence per element would cost one PCIe round trip each ----
    {
        const T* gin = P.goals + (size_t)b * N * P.goal_stride;
        T rg[2] = {T(0), T(0)};
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int t = tid + q * NT;
            if (t < 3 * N) { const int k = t / 3; rg[q] = gin[k * P.goal_stride + (t - 3 * k)]; }
        }
        const T rx = tid < 12 ? P.x0[(size_t)b * 12 + tid] : T(0);
        const T rf = (tid < 6 && P.fext) ? P.fext[(size_t)b * 6 + tid] : T(0);
        const T* xin = P.xu_in + (size_t)b * nz;
        for (int base = 0; base < nz; base += 4 * NT) {
            T r[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) { const int t = base + q * NT + tid; r[q] = t < nz ? xin[t] : T(0); }
#pragma unroll
            for (int q = 0; q < 4; ++q) { const int t = base + q * NT + tid; if (t < nz) M.xu[t] = r[q]; }
        }
#pragma unroll
        for (int q = 0; q < 2; ++q) { const int t = tid + q * NT; if (t < 3 * N) M.goals[t] = rg[q]; }
        for (int t = tid + 2 * NT; t < 3 * N; t += NT) {            // horizons beyond 2 NT / 3 knots
            const int k = t / 3;
            M.goals[t] = gin[k * P.goal_stride + (t - 3 * k)];
        }
        if (tid < 12) M.x0[tid] = rx;
        if (P.x0_into_xu && tid < 12) M.xu[tid] = rx;               // same thread as the store of xu[tid] above
        if (tid < 6) s_fext[tid] = rf;
        if (tid < 6) M.xu[nz + tid] = T(0);                          // padded u_{N-1}
        for (int t = tid; t < 18 * N; t += NT) M.z[t] = T(0);
    }
    __syncthreads();
    const T* fext = P.fext ? s_fext : nullptr;

    int qp_iters = 0, n_ls = 0, n_steps = 0, status = 0;
    T alphas[2] = {T(0), T(0)}, steps[2] = {T(0), T(0)};
    T merit_final = T(0), merit_first = T(0);
    for (int qp = 0; qp < P.max_qp; ++qp) {
        qp_iters = qp + 1;
        if (prof && tid == 0) prof[5 * qp + 0] = clock64();
        // ---- setup of the own knots; export what the backward sweep needs ----
        phase_setup<T, NT>(rob, P, M, W, fext, false, qp == 0 ? prof : nullptr);
        for (int t = tid; t < RIC_STRIDE * W.nc; t += NT) {
            const int k = t / RIC_STRIDE, e = t - RIC_STRIDE * k;
            const T* kr = M.kr + KR_STRIDE * k;
            T v = T(0);
            if (e >= RIC_COST) { if (e < RIC_COST + COST_STRIDE) v = M.cost[COST_STRIDE * k + e - RIC_COST]; }
            else if (k < W.nd) {
                if (e < RIC_SG) v = kr[KR_L + e];
                else if (e < RIC_S) v = kr[KR_SG + e - RIC_SG];
                else if (e < RIC_D) v = kr[KR_S + e - RIC_S];
                else if (e < RIC_HX) v = kr[KR_D + e - RIC_D];
                else v = kr[KR_HX + e - RIC_HX];
            }
            ric[RIC_STRIDE * (W.k0 + k) + e] = v;
        }
        cluster.sync();                                                             // barrier 1
        if (prof && tid == 0) prof[5 * qp + 1] = clock64();
        // ---- lead CTA: exact QP solve (backward sweep + roll-out), publish the states ----
        if (rank == 0) {
            riccati_sweep<T, NT, RicGlobal<T>>(P, L.work, L.z, L.x0, L.xu, RicGlobal<T>{ric, L.gains}, qp == 0 ? prof : nullptr);
            if (prof && qp == 0 && tid == 0) prof[27] = clock64();
            for (int t = tid; t < 12 * N; t += NT) {
                const int k = t / 12, i = t - 12 * k;
                zx[t] = M.z[18 * k + i];
            }
        }
        if (prof && qp == 0 && tid == 0) prof[28] = clock64();
        cluster.sync();                                                             // barrier 2
        if (prof && qp == 0 && tid == 0) prof[29] = clock64();
        if (rank != 0) {
            for (int t = tid; t < 12 * N; t += NT) {
                const int k = t / 12, i = t - 12 * k;
                M.z[18 * k + i] = __ldcg(zx + t);
            }
        }
        __syncthreads();
        if (prof && qp == 0 && tid == 0) prof[30] = clock64();
        phase_controls<T, NT, false>(P, L.z, L.ls, L.kr, nullptr, W, zu);
        if (prof && tid == 0) prof[5 * qp + 2] = clock64();
        // ---- line search: candidates 0..7 (base, alpha = 1 .. 1/64) at the own knots — with 4 knots per CTA that is
        //      exactly one warp per task type; the last candidate (1/128) only if none of those is accepted ----
#pragma unroll 1      // the body is 24 KB of code: unrolled by four it evicts itself from the instruction cache
        for (int c0 = 0; c0 < 8; c0 += M.merit_cap) {
            const int ncw = 8 - c0 < M.merit_cap ? 8 - c0 : M.merit_cap;
            if (L.par_merit) merit_points_par<T, NT>(rob, P, M, W, M.kr, M.xu, M.z, fext, c0, ncw, mcomps, c0);
            else merit_points<T, NT>(rob, P, M, W, M.kr, M.xu, M.z, fext, c0, ncw, mcomps, c0);
        }
        cluster.sync();                                                             // barrier 3
        for (int t = tid; t < 6 * (N - 1); t += NT) {
            const int k = t / 6, i = t - 6 * k;
            if (k < W.k0 || k >= W.k0 + W.nd) M.z[18 * k + 12 + i] = __ldcg(zu + t);
        }
        if (tid == 0) { s_alpha = T(0); s_flag = 0; }
        merit_reduce<T, NT, true>(P, M.xu, M.z, 0, 8, mcomps, 0, s_merits, nullptr);
        if (tid == 0) {
            if (!(s_merits[0] == s_merits[0])) s_flag = 1;
            s_merit_acc = s_merits[0];
            for (int c = 1; c < 8; ++c)
                if (s_merits[c] <= s_merits[0]) { s_alpha = T(c_alphas[c]); s_merit_acc = s_merits[c]; break; }
        }
        __syncthreads();
        if (s_alpha == T(0)) {          // identical decision in every CTA of the cluster (same data, same arithmetic)
            if (L.par_merit) merit_points_par<T, NT>(rob, P, M, W, M.kr, M.xu, M.z, fext, 8, 1, mcomps, 8);
            else merit_points<T, NT>(rob, P, M, W, M.kr, M.xu, M.z, fext, 8, 1, mcomps, 8);
            cluster.sync();                                                         // barrier 3b (rare)
            merit_reduce<T, NT, true>(P, M.xu, M.z, 8, 1, mcomps, 8, s_merits, nullptr);
            if (tid == 0 && s_merits[8] <= s_merits[0]) { s_alpha = T(c_alphas[8]); s_merit_acc = s_merits[8]; }
            __syncthreads();
        }
        if (prof && tid == 0) prof[5 * qp + 3] = clock64();
        const T alpha = s_alpha;
        merit_final = s_merit_acc;
        if (qp == 0) merit_first = s_merits[0];
        if (s_flag) status = 2;
        if (qp < 2) alphas[qp] = alpha;
        ++n_ls;
        if (alpha == T(0)) {
            // src/osqp_sqp.py:81-82 `continue`: the remaining passes would re-solve the identical QP; counted, skipped
            if (status == 0) status = 1;
            for (int q2 = qp + 1; q2 < P.max_qp; ++q2) { if (q2 < 2) alphas[q2] = T(0); ++n_ls; qp_iters = q2 + 1; }
            break;
        }
        // XU += alpha (z - XU); step norm (every CTA, identically)
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
        if (prof && tid == 0) prof[5 * qp + 4] = clock64();
        if (n_steps < 2) steps[n_steps] = stepsize;
        ++n_steps;
        __syncthreads();
        if (stepsize < P.step_tol) break;
    }
    // ---- outputs: every CTA writes a slice of the trajectory, the lead CTA the statistics ----
    for (int t = rank * NT + tid; t < nz; t += CL * NT) P.xu_out[(size_t)b * nz + t] = M.xu[t];
    if (P.stats && rank == 0 && tid == 0) {
        T* st = P.stats + (size_t)b * 12;
        st[0] = T(qp_iters); st[1] = T(n_ls); st[2] = T(n_steps); st[3] = T(status);
        st[4] = alphas[0]; st[5] = alphas[1]; st[6] = steps[0]; st[7] = steps[1];
        st[8] = merit_final; st[9] = merit_first; st[10] = T(0); st[11] = T(0);
    }
}

}  // namespace i7
