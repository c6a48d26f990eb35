// i7_wf.cuh — rigid-body dynamics of the 6-link chain in WORLD-FRAME coordinates, written as ROLLED loops.
//
// Spatial vectors [angular; linear] are expressed in world axes at ONE fixed spatial origin.  In these coordinates
// the recursions over the links lose their 6x6 transforms:
//     v_l = sum_{k<=l} S_k qd_k,   a_l = a_base + sum_{k<=l} (S_k qdd_k + v_k x S_k qd_k),   F_i = sum_{l>=i} f_l,
//     tau_i = S_i . F_i,   M_ij = S_i . (sum_{l>=max(i,j)} I_l) S_j,
// and the partial derivatives of inverse dynamics have closed forms per (joint j, link l >= j) pair, so every
// quantity is a short prefix / suffix sum of independent terms.  Replaces pin.computeABADerivatives / pin.aba
// (src/osqp_solver.py:71, src/osqp_sqp.py:40); the derivative recursion is the world-frame one of Carpentier &
// Mansard 2018, re-derived so that each (knot, link) and (knot, column) is an independent task.
//
// Two deliberate choices:
//  * the spatial origin is the position of the LAST joint (a fixed point of space for the purpose of every
//    derivative): it keeps the parallel-axis terms of the light wrist links small, which is what fp32 needs, and it
//    makes the translational end-effector Jacobian column i equal to the linear part of S_i;
//  * every loop over the links is a real loop (#pragma unroll 1) over per-link data in shared memory / the constant
//    bank: the fused kernels execute each phase once per SQP iteration, so fully unrolled bodies are instruction-
//    cache misses from start to end (ncu: stall_no_instruction was the top stall of the unrolled version).
// World inertia of a link: IW = (IO xx xy xz yy yz zz about the origin, first moment H[3]); the mass is rob.m[l].
#pragma once
#include "i7_rbd.cuh"

namespace i7 {

// o = v x m (spatial motion cross product)
template <typename T> I7_DEV void crm6(const T* v, const T* m, T* o) {
    cross3(v, m, o);
    cross3(v, m + 3, o + 3);
    cross3_acc(v + 3, m, o + 3);
}
// o = I x for a world inertia (IO, H, ms):  [IO w + H x v ; ms v - H x w]
template <typename T> I7_DEV void winertia_mul(const T* IW, T ms, const T* x, T* o) {
    const T* H = IW + 6;
    o[0] = IW[0] * x[0] + IW[1] * x[1] + IW[2] * x[2] + (H[1] * x[5] - H[2] * x[4]);
    o[1] = IW[1] * x[0] + IW[3] * x[1] + IW[4] * x[2] + (H[2] * x[3] - H[0] * x[5]);
    o[2] = IW[2] * x[0] + IW[4] * x[1] + IW[5] * x[2] + (H[0] * x[4] - H[1] * x[3]);
    o[3] = ms * x[3] - (H[1] * x[2] - H[2] * x[1]);
    o[4] = ms * x[4] - (H[2] * x[0] - H[0] * x[2]);
    o[5] = ms * x[5] - (H[0] * x[1] - H[1] * x[0]);
}

// one step of the kinematic chain: (Rc, pc) of the parent of joint i -> of joint frame i.  cs = {cos q, sin q}.
template <typename T> I7_DEV void wf_chain_step(const Robot<T>& rob, int i, const T* cs, T* Rc, T* pc) {
    T d[3];
    rot(Rc, rob.p[i], d);
    pc[0] += d[0]; pc[1] += d[1]; pc[2] += d[2];
    T E[9], Rn[9];
    joint_E(rob, i, cs[i], cs[6 + i], E);
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c)
            Rn[3 * r + c] = Rc[3 * r] * E[3 * c] + Rc[3 * r + 1] * E[3 * c + 1] + Rc[3 * r + 2] * E[3 * c + 2];
#pragma unroll
    for (int r = 0; r < 9; ++r) Rc[r] = Rn[r];
}
// position of the last joint frame (= eepos of the reference, src/osqp_solver.py:146-148)
template <typename T> I7_DEV void wf_tip(const Robot<T>& rob, const T* cs, T* pee) {
    T Rc[9] = {T(1), T(0), T(0), T(0), T(1), T(0), T(0), T(0), T(1)};
    T pc[3] = {T(0), T(0), T(0)};
#pragma unroll 1
    for (int i = 0; i < I7_NJ; ++i) wf_chain_step(rob, i, cs, Rc, pc);
    pee[0] = pc[0]; pee[1] = pc[1]; pee[2] = pc[2];
}

// inertia of link l about the spatial origin from (Ib about the link origin, h = m c, m); Rw, p = pose of the link
// frame relative to the spatial origin
template <typename T> I7_DEV void wf_inertia(const Robot<T>& rob, int l, const T* Rw, const T* p, T* IW) {
    const T* Ib = rob.Ib[l]; const T* h = rob.h[l]; const T ms = rob.m[l];
    const T I3[9] = {Ib[0], Ib[1], Ib[2], Ib[1], Ib[3], Ib[4], Ib[2], Ib[4], Ib[5]};
    T B[9];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c)
            B[3 * r + c] = Rw[3 * r] * I3[c] + Rw[3 * r + 1] * I3[3 + c] + Rw[3 * r + 2] * I3[6 + c];
    T hw[3];
    rot(Rw, h, hw);
    const T pp = p[0] * p[0] + p[1] * p[1] + p[2] * p[2];
    const T ph = p[0] * hw[0] + p[1] * hw[1] + p[2] * hw[2];
    const T dg = ms * pp + T(2) * ph;
    const int rr[6] = {0, 0, 0, 1, 1, 2}, cc[6] = {0, 1, 2, 1, 2, 2};
#pragma unroll
    for (int e = 0; e < 6; ++e) {
        const int r = rr[e], c = cc[e];
        T x = B[3 * r] * Rw[3 * c] + B[3 * r + 1] * Rw[3 * c + 1] + B[3 * r + 2] * Rw[3 * c + 2];
        x -= ms * p[r] * p[c] + p[r] * hw[c] + hw[r] * p[c];
        if (r == c) x += dg;
        IW[e] = x;
    }
#pragma unroll
    for (int r = 0; r < 3; ++r) IW[6 + r] = hw[r] + ms * p[r];
}

// Task (knot, link l): walk the chain to link l accumulating the velocity and the acceleration at qdd = 0 (gravity
// as base acceleration); outputs the link's axis S_l, v_l, a0_l, inertia IW_l (9 numbers) and force
// f0_l = I_l a0_l + v_l x* I_l v_l (minus the external wrench on the last link, given in link coordinates).
// cs, qd: arrays indexed by joint (shared memory); pee: position of the last joint = spatial origin.
template <typename T>
I7_DEV void wf_link(const Robot<T>& rob, const T* cs, const T* qd, const T* fext, const T* pee, int l, T* S_o, T* v_o,
                    T* a_o, T* IW_o, T* f_o) {
    T Rc[9] = {T(1), T(0), T(0), T(0), T(1), T(0), T(0), T(0), T(1)};
    T pc[3] = {T(0), T(0), T(0)};
    T p[3] = {T(0), T(0), T(0)};
    T v[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
    T a[6] = {T(0), T(0), T(0), -rob.g[0], -rob.g[1], -rob.g[2]};
    T S[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
#pragma unroll 1
    for (int i = 0; i <= l; ++i) {
        wf_chain_step(rob, i, cs, Rc, pc);
        p[0] = pc[0] - pee[0]; p[1] = pc[1] - pee[1]; p[2] = pc[2] - pee[2];
        S[0] = Rc[2]; S[1] = Rc[5]; S[2] = Rc[8];
        cross3(p, S, S + 3);
        const T qi = qd[i];
        T vj[6], c6[6];
#pragma unroll
        for (int r = 0; r < 6; ++r) { vj[r] = S[r] * qi; v[r] += vj[r]; }
        crm6(v, vj, c6);
#pragma unroll
        for (int r = 0; r < 6; ++r) a[r] += c6[r];
    }
    T IW[9];
    wf_inertia(rob, l, Rc, p, IW);
    const T ms = rob.m[l];
    T Ia[6], Iv[6];
    winertia_mul(IW, ms, a, Ia);
    winertia_mul(IW, ms, v, Iv);
    crf<T, true>(v, Iv, Ia);
    if (l == I7_NJ - 1 && fext) {
        T nw[3], fw[3];
        rot(Rc, fext, nw);              // the spatial origin IS the last joint's origin: no moment arm
        rot(Rc, fext + 3, fw);
#pragma unroll
        for (int r = 0; r < 3; ++r) { Ia[r] -= nw[r]; Ia[3 + r] -= fw[r]; }
    }
#pragma unroll
    for (int r = 0; r < 6; ++r) { S_o[r] = S[r]; v_o[r] = v[r]; a_o[r] = a[r]; f_o[r] = Ia[r]; }
#pragma unroll
    for (int r = 0; r < 9; ++r) IW_o[r] = IW[r];
}

// Task (point): the same for ALL links in one pass (line-search points: one thread per point); S (36), IW (54),
// f0 (36) go to the point's scratch, v / a stay in registers.
template <typename T>
I7_DEV void wf_links_all(const Robot<T>& rob, const T* cs, const T* qd, const T* fext, const T* pee, T* S_o, T* IW_o,
                         T* f_o) {
    T Rc[9] = {T(1), T(0), T(0), T(0), T(1), T(0), T(0), T(0), T(1)};
    T pc[3] = {T(0), T(0), T(0)};
    T v[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
    T a[6] = {T(0), T(0), T(0), -rob.g[0], -rob.g[1], -rob.g[2]};
#pragma unroll 1
    for (int i = 0; i < I7_NJ; ++i) {
        wf_chain_step(rob, i, cs, Rc, pc);
        const T p[3] = {pc[0] - pee[0], pc[1] - pee[1], pc[2] - pee[2]};
        T S[6] = {Rc[2], Rc[5], Rc[8], T(0), T(0), T(0)};
        cross3(p, S, S + 3);
        const T qi = qd[i];
        T vj[6], c6[6];
#pragma unroll
        for (int r = 0; r < 6; ++r) { vj[r] = S[r] * qi; v[r] += vj[r]; }
        crm6(v, vj, c6);
#pragma unroll
        for (int r = 0; r < 6; ++r) a[r] += c6[r];
        T IW[9];
        wf_inertia(rob, i, Rc, p, IW);
        const T ms = rob.m[i];
        T Ia[6], Iv[6];
        winertia_mul(IW, ms, a, Ia);
        winertia_mul(IW, ms, v, Iv);
        crf<T, true>(v, Iv, Ia);
        if (i == I7_NJ - 1 && fext) {
            T nw[3], fw[3];
            rot(Rc, fext, nw);
            rot(Rc, fext + 3, fw);
#pragma unroll
            for (int r = 0; r < 3; ++r) { Ia[r] -= nw[r]; Ia[3 + r] -= fw[r]; }
        }
#pragma unroll
        for (int r = 0; r < 6; ++r) { S_o[6 * i + r] = S[r]; f_o[6 * i + r] = Ia[r]; }
#pragma unroll
        for (int r = 0; r < 9; ++r) IW_o[9 * i + r] = IW[r];
    }
}

// Task (knot, j): column j of the joint-space inertia matrix (rows i <= j, mirrored into the 6x6 array Mm) and the
// bias torque b_j, from the per-link records:  M_ij = S_i . (sum_{l>=j} I_l) S_j,  b_j = S_j . sum_{l>=j} f0_l.
template <typename T>
I7_DEV void wf_mass_column(const Robot<T>& rob, const T* S /*36*/, const T* IW /*54*/, const T* f0 /*36*/, int j,
                           T* Mm /*36*/, T* bj) {
    T IC[9] = {T(0), T(0), T(0), T(0), T(0), T(0), T(0), T(0), T(0)};
    T F[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
    T mc = T(0);
#pragma unroll 1
    for (int l = I7_NJ - 1; l >= j; --l) {
#pragma unroll
        for (int r = 0; r < 9; ++r) IC[r] += IW[9 * l + r];
#pragma unroll
        for (int r = 0; r < 6; ++r) F[r] += f0[6 * l + r];
        mc += rob.m[l];
    }
    T Sj[6], y[6];
#pragma unroll
    for (int r = 0; r < 6; ++r) Sj[r] = S[6 * j + r];
    winertia_mul(IC, mc, Sj, y);
    T b = T(0);
#pragma unroll
    for (int r = 0; r < 6; ++r) b += Sj[r] * F[r];
    *bj = b;
#pragma unroll 1
    for (int i = 0; i <= j; ++i) {
        T d = T(0);
#pragma unroll
        for (int r = 0; r < 6; ++r) d += S[6 * i + r] * y[r];
        Mm[6 * i + j] = d; Mm[6 * j + i] = d;
    }
}

// Task (knot, l): fold the solved accelerations into the link record:  a_l += sum_{k<=l} S_k qdd_k,
// f_l += I_l (that sum).
template <typename T>
I7_DEV void wf_add_qdd(const Robot<T>& rob, const T* S /*36*/, const T* qdd, int l, const T* IW_l, T* a_l, T* f_l) {
    T s[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
#pragma unroll 1
    for (int k = 0; k <= l; ++k) {
        const T qk = qdd[k];
#pragma unroll
        for (int r = 0; r < 6; ++r) s[r] += S[6 * k + r] * qk;
    }
    T Is[6];
    winertia_mul(IW_l, rob.m[l], s, Is);
#pragma unroll
    for (int r = 0; r < 6; ++r) { a_l[r] += s[r]; f_l[r] += Is[r]; }
}

// Column j of d tau/dq (WRT_Q) or d tau/dqd at (q, qd, qdd), from the per-link records S, v, a, f (at the solved
// qdd), IW.  For every link l >= j
//     rest_l = I_l da_l + dv x* (I_l v_l) + v_l x* (I_l dv),      da_l = a0 + B x (v_l - v_j)
//   d/dq_j :  dv = -S_j x v_j,  a0 = -S_j x a_j,  B = v_j x S_j    (the rigid rotation S_j x* f_l of the whole
//             sub-chain cancels against d S_i/dq_j for i > j and is added once, as S_j x* F_j, for i <= j)
//   d/dqd_j:  dv = S_j,         a0 = v_j x S_j,   B = S_j
//   out_i = S_i . sum_{l >= max(i,j)} rest_l  (+ S_i . (S_j x* F_j) for i <= j, WRT_Q)
template <typename T> struct DSeed { T Sj[6], vj[6], B[6], dv[6], a0[6]; };
template <typename T, bool WRT_Q>
I7_DEV void wf_dseed(const T* S, const T* V, const T* A, int j, DSeed<T>& d) {
    T vjS[6];
#pragma unroll
    for (int r = 0; r < 6; ++r) { d.Sj[r] = S[6 * j + r]; d.vj[r] = V[6 * j + r]; }
    crm6(d.vj, d.Sj, vjS);
    if (WRT_Q) {
        T aj[6];
#pragma unroll
        for (int r = 0; r < 6; ++r) aj[r] = A[6 * j + r];
        crm6(d.Sj, d.vj, d.dv);
        crm6(d.Sj, aj, d.a0);
#pragma unroll
        for (int r = 0; r < 6; ++r) { d.dv[r] = -d.dv[r]; d.a0[r] = -d.a0[r]; d.B[r] = vjS[r]; }
    } else {
#pragma unroll
        for (int r = 0; r < 6; ++r) { d.dv[r] = d.Sj[r]; d.a0[r] = vjS[r]; d.B[r] = d.Sj[r]; }
    }
}
// rest += I_l da_l + dv x* (I_l v_l) + v_l x* (I_l dv)   for one link l >= j
template <typename T>
I7_DEV void wf_drest(const Robot<T>& rob, const T* V, const T* IW, int l, const DSeed<T>& d, T* rest) {
    const T ms = rob.m[l];
    const T* IWl = IW + 9 * l;
    T vl[6], W[6], da[6], t[6];
#pragma unroll
    for (int r = 0; r < 6; ++r) { vl[r] = V[6 * l + r]; W[r] = vl[r] - d.vj[r]; }
    crm6(d.B, W, da);
#pragma unroll
    for (int r = 0; r < 6; ++r) da[r] += d.a0[r];
    winertia_mul(IWl, ms, da, t);
#pragma unroll
    for (int r = 0; r < 6; ++r) rest[r] += t[r];
    winertia_mul(IWl, ms, vl, t);
    crf<T, true>(d.dv, t, rest);
    winertia_mul(IWl, ms, d.dv, t);
    crf<T, true>(vl, t, rest);
}
// whole column: out[stride * i] = d tau_i / d (q | qd)_j
template <typename T, bool WRT_Q>
I7_DEV void wf_dcolumn(const Robot<T>& rob, const T* S, const T* V, const T* A, const T* F_, const T* IW, int j, T* out,
                       int stride) {
    DSeed<T> d;
    wf_dseed<T, WRT_Q>(S, V, A, j, d);
    T rest[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
    T Fs[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
#pragma unroll 1
    for (int l = I7_NJ - 1; l >= j; --l) {
        wf_drest<T>(rob, V, IW, l, d, rest);
        if (WRT_Q) {
#pragma unroll
            for (int r = 0; r < 6; ++r) Fs[r] += F_[6 * l + r];
        }
        if (l > j) {
            T dd = T(0);
#pragma unroll
            for (int r = 0; r < 6; ++r) dd += S[6 * l + r] * rest[r];
            out[stride * l] = dd;
        }
    }
    if (WRT_Q) crf<T, true>(d.Sj, Fs, rest);
#pragma unroll 1
    for (int i = 0; i <= j; ++i) {
        T dd = T(0);
#pragma unroll
        for (int r = 0; r < 6; ++r) dd += S[6 * i + r] * rest[r];
        out[stride * i] = dd;
    }
}

// Finer split of the same column (latency form): wf_dpair is the (j, l) term on its own, wf_dgather sums the
// terms of links >= max(i, j) for output row i (same order of additions as wf_dcolumn: from the tip inwards).
template <typename T, bool WRT_Q>
I7_DEV void wf_dpair(const Robot<T>& rob, const T* S, const T* V, const T* A, const T* IW, int j, int l, T* rest6) {
    DSeed<T> d;
    wf_dseed<T, WRT_Q>(S, V, A, j, d);
    T rest[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
    wf_drest<T>(rob, V, IW, l, d, rest);
#pragma unroll
    for (int r = 0; r < 6; ++r) rest6[r] = rest[r];
}
template <typename T, bool WRT_Q>
I7_DEV T wf_dgather(const T* S, const T* F_, const T* pairs /* [l][6] of this (knot, column) */, int j, int i) {
    const int i0 = i > j ? i : j;
    T rest[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
#pragma unroll 1
    for (int l = I7_NJ - 1; l >= i0; --l) {
#pragma unroll
        for (int r = 0; r < 6; ++r) rest[r] += pairs[6 * l + r];
    }
    if (WRT_Q && i <= j) {
        T Fs[6] = {T(0), T(0), T(0), T(0), T(0), T(0)}, Sj[6];
#pragma unroll 1
        for (int l = I7_NJ - 1; l >= j; --l) {
#pragma unroll
            for (int r = 0; r < 6; ++r) Fs[r] += F_[6 * l + r];
        }
#pragma unroll
        for (int r = 0; r < 6; ++r) Sj[r] = S[6 * j + r];
        crf<T, true>(Sj, Fs, rest);
    }
    T dd = T(0);
#pragma unroll
    for (int r = 0; r < 6; ++r) dd += S[6 * i + r] * rest[r];
    return dd;
}

// Forward dynamics of one point from its link records (one thread): qdd = M^-1 (u - b) with
// M_ij = S_i . IC_j S_j (composite inertias accumulated from the tip), b_j = S_j . F_j.  u: 6 torques.
template <typename T>
I7_DEV void wf_forward_dynamics(const Robot<T>& rob, const T* S, const T* IW, const T* f0, const T* u, T* qdd) {
    T A[6][6], dinv[6];
    T IC[9] = {T(0), T(0), T(0), T(0), T(0), T(0), T(0), T(0), T(0)};
    T F[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
    T mc = T(0);
#pragma unroll
    for (int j = I7_NJ - 1; j >= 0; --j) {
#pragma unroll
        for (int r = 0; r < 9; ++r) IC[r] += IW[9 * j + r];
#pragma unroll
        for (int r = 0; r < 6; ++r) F[r] += f0[6 * j + r];
        mc += rob.m[j];
        T Sj[6], y[6];
#pragma unroll
        for (int r = 0; r < 6; ++r) Sj[r] = S[6 * j + r];
        winertia_mul(IC, mc, Sj, y);
        T b = T(0);
#pragma unroll
        for (int r = 0; r < 6; ++r) b += Sj[r] * F[r];
        qdd[j] = u[j] - b;
#pragma unroll
        for (int i = 0; i <= j; ++i) {
            T d = T(0);
#pragma unroll
            for (int r = 0; r < 6; ++r) d += S[6 * i + r] * y[r];
            A[j][i] = d;
        }
    }
    spd6_factor<T>(A, dinv);
    spd6_solve_factored<T>(A, dinv, qdd);
}

// ---------------------------------------------------------------------------------------------------------
// Forward dynamics of one point by ONE thread with 96 numbers of scratch (line-search points: the smaller the
// scratch, the more candidates fit one wave).  Same quantities as wf_links_all + wf_forward_dynamics, but the bias
// b_j = sum_{l >= j} S_j . f0_l is accumulated while the chain is walked, so the link forces are never stored.
// scr: [S 36 | IW 54 | b 6]; on entry cos q_i, sin q_i sit in scr[wp_cs(i)], scr[wp_cs(i) + 1], slots of the
// inertias of links 4 and 5 that are first written after the last use of those cos / sin.
constexpr int WP_S = 0, WP_I = 36, WP_B = 90, WP_STRIDE = 97;      // odd stride: conflict-free across points
I7_DEV int wp_cs(int i) { return WP_I + 36 + 2 * i + (i >= 2 ? 5 : 0); }

template <typename T> I7_DEV void wf_chain_step_cs(const Robot<T>& rob, int i, T c, T s, T* Rc, T* pc) {
    T d[3];
    rot(Rc, rob.p[i], d);
    pc[0] += d[0]; pc[1] += d[1]; pc[2] += d[2];
    T E[9], Rn[9];
    joint_E(rob, i, c, s, E);
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int cc = 0; cc < 3; ++cc)
            Rn[3 * r + cc] = Rc[3 * r] * E[3 * cc] + Rc[3 * r + 1] * E[3 * cc + 1] + Rc[3 * r + 2] * E[3 * cc + 2];
#pragma unroll
    for (int r = 0; r < 9; ++r) Rc[r] = Rn[r];
}
// position of the last joint frame, cos / sin read from the point's scratch
template <typename T> I7_DEV void wf_tip_scr(const Robot<T>& rob, const T* scr, T* pee) {
    T Rc[9] = {T(1), T(0), T(0), T(0), T(1), T(0), T(0), T(0), T(1)};
    T pc[3] = {T(0), T(0), T(0)};
#pragma unroll 1
    for (int i = 0; i < I7_NJ; ++i) wf_chain_step_cs(rob, i, scr[wp_cs(i)], scr[wp_cs(i) + 1], Rc, pc);
    pee[0] = pc[0]; pee[1] = pc[1]; pee[2] = pc[2];
}

template <typename T>
I7_DEV void wf_point_forward_dynamics(const Robot<T>& rob, T* scr, const T* qd, const T* u, const T* fext, const T* pee,
                                      T* qdd) {
    T* S = scr + WP_S; T* IWs = scr + WP_I; T* bb = scr + WP_B;
    {
        T Rc[9] = {T(1), T(0), T(0), T(0), T(1), T(0), T(0), T(0), T(1)};
        T pc[3] = {T(0), T(0), T(0)};
        T v[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
        T a[6] = {T(0), T(0), T(0), -rob.g[0], -rob.g[1], -rob.g[2]};
#pragma unroll 1
        for (int l = 0; l < I7_NJ; ++l) {
            wf_chain_step_cs(rob, l, scr[wp_cs(l)], scr[wp_cs(l) + 1], Rc, pc);
            const T p[3] = {pc[0] - pee[0], pc[1] - pee[1], pc[2] - pee[2]};
            T Sl[6] = {Rc[2], Rc[5], Rc[8], T(0), T(0), T(0)};
            cross3(p, Sl, Sl + 3);
            T ql = T(0);
#pragma unroll
            for (int r = 0; r < 6; ++r) if (r == l) ql = qd[r];          // qd lives in registers
            T vj[6], c6[6];
#pragma unroll
            for (int r = 0; r < 6; ++r) { vj[r] = Sl[r] * ql; v[r] += vj[r]; S[6 * l + r] = Sl[r]; }
            crm6(v, vj, c6);
#pragma unroll
            for (int r = 0; r < 6; ++r) a[r] += c6[r];
            T IW[9];
            wf_inertia(rob, l, Rc, p, IW);
            const T ms = rob.m[l];
            T f[6], Iv[6];
            winertia_mul(IW, ms, a, f);
            winertia_mul(IW, ms, v, Iv);
            crf<T, true>(v, Iv, f);
            if (l == I7_NJ - 1 && fext) {
                T nw[3], fw[3];
                rot(Rc, fext, nw);
                rot(Rc, fext + 3, fw);
#pragma unroll
                for (int r = 0; r < 3; ++r) { f[r] -= nw[r]; f[3 + r] -= fw[r]; }
            }
#pragma unroll
            for (int r = 0; r < 9; ++r) IWs[9 * l + r] = IW[r];
            // b_j += S_j . f_l for every joint j <= l that carries link l
            T bl = T(0);
#pragma unroll
            for (int r = 0; r < 6; ++r) bl += Sl[r] * f[r];
            bb[l] = bl;
#pragma unroll 1
            for (int j = 0; j < l; ++j) {
                T d = T(0);
#pragma unroll
                for (int r = 0; r < 6; ++r) d += S[6 * j + r] * f[r];
                bb[j] += d;
            }
        }
    }
    T A[6][6], dinv[6];
    T IC[9] = {T(0), T(0), T(0), T(0), T(0), T(0), T(0), T(0), T(0)};
    T mc = T(0);
#pragma unroll
    for (int j = I7_NJ - 1; j >= 0; --j) {
#pragma unroll
        for (int r = 0; r < 9; ++r) IC[r] += IWs[9 * j + r];
        mc += rob.m[j];
        T Sj[6], y[6];
#pragma unroll
        for (int r = 0; r < 6; ++r) Sj[r] = S[6 * j + r];
        winertia_mul(IC, mc, Sj, y);
        qdd[j] = u[j] - bb[j];
#pragma unroll
        for (int i = 0; i <= j; ++i) {
            T d = T(0);
#pragma unroll
            for (int r = 0; r < 6; ++r) d += S[6 * i + r] * y[r];
            A[j][i] = d;
        }
    }
    spd6_factor<T>(A, dinv);
    spd6_solve_factored<T>(A, dinv, qdd);
}

}  // namespace i7
