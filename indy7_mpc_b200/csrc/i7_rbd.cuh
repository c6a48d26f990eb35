// i7_rbd.cuh — rigid-body kinematics/dynamics of a 6-link revolute-z serial chain, device code for sm_100a.
//
// Replaces the Pinocchio calls of the reference hot path (src/osqp_solver.py:71,77,147-154;
// src/osqp_sqp.py:40-41; src/utils.py:5-17).  Featherstone spatial algebra, [angular; linear] ordering,
// link-local coordinates, S = e_z(angular) for every joint.  Everything is templated on the arithmetic type
// (double / float), fully unrolled over the 6 links so that all small arrays live in registers, and reads
// the robot constants from the kernel-parameter constant bank (uniform loads).
#pragma once
#include <cuda_runtime.h>

#define I7_DEV __device__ __forceinline__
#define I7_NJ 6

namespace i7 {

template <typename T>
struct Robot {
    T R[I7_NJ][9];    // joint placement rotation (row-major), parent <- joint frame
    T p[I7_NJ][3];    // joint placement translation
    T m[I7_NJ];       // mass
    T h[I7_NJ][3];    // m * com
    T Ib[I7_NJ][6];   // rotational inertia about the link-frame origin: xx xy xz yy yz zz
    T g[3];           // world gravity
};

template <typename T> I7_DEV void sincos_t(T x, T* s, T* c);
template <> I7_DEV void sincos_t<double>(double x, double* s, double* c) { sincos(x, s, c); }
template <> I7_DEV void sincos_t<float>(float x, float* s, float* c) { sincosf(x, s, c); }
template <typename T> I7_DEV T sqrt_t(T x);
template <> I7_DEV double sqrt_t<double>(double x) { return sqrt(x); }
template <> I7_DEV float sqrt_t<float>(float x) { return sqrtf(x); }

// o = a x b
template <typename T> I7_DEV void cross3(const T* a, const T* b, T* o) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}
// o += a x b
template <typename T> I7_DEV void cross3_acc(const T* a, const T* b, T* o) {
    o[0] += a[1] * b[2] - a[2] * b[1];
    o[1] += a[2] * b[0] - a[0] * b[2];
    o[2] += a[0] * b[1] - a[1] * b[0];
}

// E = (R_i Rz(q))^T : parent coordinates -> link-i coordinates
template <typename T> I7_DEV void joint_E(const Robot<T>& rob, int i, T c, T s, T* E) {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const T r0 = rob.R[i][3 * a + 0], r1 = rob.R[i][3 * a + 1], r2 = rob.R[i][3 * a + 2];
        E[0 + a] = r0 * c + r1 * s;
        E[3 + a] = r1 * c - r0 * s;
        E[6 + a] = r2;
    }
}

// o = E x
template <typename T> I7_DEV void rot(const T* E, const T* x, T* o) {
#pragma unroll
    for (int a = 0; a < 3; ++a) o[a] = E[3 * a] * x[0] + E[3 * a + 1] * x[1] + E[3 * a + 2] * x[2];
}
// o = E^T x
template <typename T> I7_DEV void rot_t(const T* E, const T* x, T* o) {
#pragma unroll
    for (int a = 0; a < 3; ++a) o[a] = E[a] * x[0] + E[3 + a] * x[1] + E[6 + a] * x[2];
}

// child <- parent motion transform: [E w ; E (v - r x w)]
template <typename T> I7_DEV void xmot(const T* E, const T* r, const T* m, T* o) {
    T t[3];
    t[0] = m[3] - (r[1] * m[2] - r[2] * m[1]);
    t[1] = m[4] - (r[2] * m[0] - r[0] * m[2]);
    t[2] = m[5] - (r[0] * m[1] - r[1] * m[0]);
    rot(E, m, o);
    rot(E, t, o + 3);
}
// parent <- child force transform X^T f: [E^T n + r x (E^T l) ; E^T l]
template <typename T> I7_DEV void xfrc_t(const T* E, const T* r, const T* f, T* o) {
    T n[3];
    rot_t(E, f, n);
    rot_t(E, f + 3, o + 3);
    o[0] = n[0] + (r[1] * o[5] - r[2] * o[4]);
    o[1] = n[1] + (r[2] * o[3] - r[0] * o[5]);
    o[2] = n[2] + (r[0] * o[4] - r[1] * o[3]);
}

// o = I_i m for a motion vector m = [w; v]:  n = Ib w + h x v,  f = m v - h x w
template <typename T> I7_DEV void inertia_mul(const Robot<T>& rob, int i, const T* m, T* o) {
    const T* Ib = rob.Ib[i];
    const T* h = rob.h[i];
    const T ms = rob.m[i];
    o[0] = Ib[0] * m[0] + Ib[1] * m[1] + Ib[2] * m[2] + (h[1] * m[5] - h[2] * m[4]);
    o[1] = Ib[1] * m[0] + Ib[3] * m[1] + Ib[4] * m[2] + (h[2] * m[3] - h[0] * m[5]);
    o[2] = Ib[2] * m[0] + Ib[4] * m[1] + Ib[5] * m[2] + (h[0] * m[4] - h[1] * m[3]);
    o[3] = ms * m[3] - (h[1] * m[2] - h[2] * m[1]);
    o[4] = ms * m[4] - (h[2] * m[0] - h[0] * m[2]);
    o[5] = ms * m[5] - (h[0] * m[1] - h[1] * m[0]);
}

// o = v x (S qd) = qd * [w_y, -w_x, 0, v_y, -v_x, 0]
template <typename T> I7_DEV void crm_S(const T* v, T qd, T* o) {
    o[0] = v[1] * qd; o[1] = -v[0] * qd; o[2] = T(0);
    o[3] = v[4] * qd; o[4] = -v[3] * qd; o[5] = T(0);
}
// o += v x (S qd)
template <typename T> I7_DEV void crm_S_acc(const T* v, T qd, T* o) {
    o[0] += v[1] * qd; o[1] -= v[0] * qd;
    o[3] += v[4] * qd; o[4] -= v[3] * qd;
}
// o (+)= v x* f  (spatial force cross product)
template <typename T, bool ACC> I7_DEV void crf(const T* v, const T* f, T* o) {
    T n[3], l[3];
    cross3(v, f, n);
    cross3_acc(v + 3, f + 3, n);
    cross3(v, f + 3, l);
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        if (ACC) { o[a] += n[a]; o[3 + a] += l[a]; } else { o[a] = n[a]; o[3 + a] = l[a]; }
    }
}

// ---------------------------------------------------------------------------------------------------------
// Forward kinematics of the last joint frame + cost-gradient direction j = J^T e  (src/osqp_solver.py:111-120,
// 146-155).  cs = {cos q_0..5, sin q_0..5}.  J[:, i] = z_i x (p_ee - p_i).
// If J_out != nullptr writes the 3x6 Jacobian row-major.
template <typename T>
I7_DEV void fk_ee(const Robot<T>& rob, const T* cs, T* pee, T* zs /*18*/, T* ps /*18*/) {
    T Rw[9] = {T(1), T(0), T(0), T(0), T(1), T(0), T(0), T(0), T(1)};   // world <- current frame
    T p[3] = {T(0), T(0), T(0)};
#pragma unroll
    for (int i = 0; i < I7_NJ; ++i) {
        T d[3];
        rot(Rw, rob.p[i], d);
        p[0] += d[0]; p[1] += d[1]; p[2] += d[2];
        // Rw <- Rw * (R_i Rz(q_i));  (R_i Rz) = E^T
        T E[9];
        joint_E(rob, i, cs[i], cs[6 + i], E);
        T Rn[9];
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
            for (int b = 0; b < 3; ++b)
                Rn[3 * a + b] = Rw[3 * a] * E[3 * b] + Rw[3 * a + 1] * E[3 * b + 1] + Rw[3 * a + 2] * E[3 * b + 2];
#pragma unroll
        for (int a = 0; a < 9; ++a) Rw[a] = Rn[a];
        if (zs) {
            zs[3 * i] = Rw[2]; zs[3 * i + 1] = Rw[5]; zs[3 * i + 2] = Rw[8];
            ps[3 * i] = p[0]; ps[3 * i + 1] = p[1]; ps[3 * i + 2] = p[2];
        }
    }
    pee[0] = p[0]; pee[1] = p[1]; pee[2] = p[2];
}

template <typename T> I7_DEV void fk_ee_only(const Robot<T>& rob, const T* cs, T* pee) {
    fk_ee<T>(rob, cs, pee, nullptr, nullptr);
}

// 3x6 translational Jacobian (row-major) from the per-joint axes / origins
template <typename T> I7_DEV void ee_jacobian(const T* pee, const T* zs, const T* ps, T* J) {
#pragma unroll
    for (int i = 0; i < I7_NJ; ++i) {
        T d[3] = {pee[0] - ps[3 * i], pee[1] - ps[3 * i + 1], pee[2] - ps[3 * i + 2]};
        T c[3];
        cross3(zs + 3 * i, d, c);
        J[i] = c[0]; J[6 + i] = c[1]; J[12 + i] = c[2];
    }
}

// ---------------------------------------------------------------------------------------------------------
// Articulated-body algorithm (Featherstone RBDA Table 7.1).  fext: Featherstone-ordered local wrench on the
// last link or nullptr.  Optional outputs for the derivative phase (strided shared-memory records):
//   U_out[6*i + r], D_out[i]  (articulated U_i = IA_i S, D_i = S^T U_i)
//   F_out[6*i + r]            total spatial force across joint i (RNEA composite force at qdd = result)
template <typename T, bool WITH_AUX>
I7_DEV void aba(const Robot<T>& rob, const T* cs, const T* qd, const T* tau, const T* fext, T* qdd,
                T* U_out, T* D_out, T* F_out) {
    T vel[I7_NJ][6];
    // pass 1: velocities
#pragma unroll
    for (int i = 0; i < I7_NJ; ++i) {
        if (i == 0) {
#pragma unroll
            for (int a = 0; a < 6; ++a) vel[0][a] = T(0);
            vel[0][2] = qd[0];
        } else {
            T E[9];
            joint_E(rob, i, cs[i], cs[6 + i], E);
            xmot(E, rob.p[i], vel[i - 1], vel[i]);
            vel[i][2] += qd[i];
        }
    }
    // pass 2: articulated inertias and bias forces, tip to base
    T IA[36], pA[6];
    T Us[I7_NJ][6], Ds[I7_NJ], us[I7_NJ];
#pragma unroll
    for (int a = 0; a < 36; ++a) IA[a] = T(0);
#pragma unroll
    for (int a = 0; a < 6; ++a) pA[a] = T(0);
#pragma unroll
    for (int i = I7_NJ - 1; i >= 0; --i) {
        // add the link's own inertia and bias force
        {
            const T* Ib = rob.Ib[i]; const T* h = rob.h[i]; const T ms = rob.m[i];
            IA[0] += Ib[0]; IA[1] += Ib[1]; IA[2] += Ib[2];
            IA[6] += Ib[1]; IA[7] += Ib[3]; IA[8] += Ib[4];
            IA[12] += Ib[2]; IA[13] += Ib[4]; IA[14] += Ib[5];
            // upper-right = skew(h), lower-left = skew(h)^T
            IA[4] -= h[2]; IA[5] += h[1]; IA[9] += h[2]; IA[11] -= h[0]; IA[15] -= h[1]; IA[16] += h[0];
            IA[19] += h[2]; IA[20] -= h[1]; IA[24] -= h[2]; IA[26] += h[0]; IA[30] += h[1]; IA[31] -= h[0];
            IA[21] += ms; IA[28] += ms; IA[35] += ms;
            T Iv[6];
            inertia_mul(rob, i, vel[i], Iv);
            crf<T, true>(vel[i], Iv, pA);
            if (i == I7_NJ - 1 && fext) {
#pragma unroll
                for (int a = 0; a < 6; ++a) pA[a] -= fext[a];
            }
        }
#pragma unroll
        for (int a = 0; a < 6; ++a) Us[i][a] = IA[6 * a + 2];
        Ds[i] = Us[i][2];
        us[i] = tau[i] - pA[2];
        if (WITH_AUX) {
#pragma unroll
            for (int a = 0; a < 6; ++a) U_out[6 * i + a] = Us[i][a];
            D_out[i] = Ds[i];
        }
        if (i > 0) {
            const T dinv = T(1) / Ds[i];
            T c[6];
            crm_S(vel[i], qd[i], c);
            // Ia = IA - U U^T / D ;  pa = pA + Ia c + U u / D
            T pa[6];
#pragma unroll
            for (int a = 0; a < 6; ++a) {
                const T ua = Us[i][a] * dinv;
#pragma unroll
                for (int b = 0; b < 6; ++b) IA[6 * a + b] -= ua * Us[i][b];
            }
#pragma unroll
            for (int a = 0; a < 6; ++a) {
                T s = pA[a] + Us[i][a] * (us[i] * dinv);
#pragma unroll
                for (int b = 0; b < 6; ++b) s += IA[6 * a + b] * c[b];
                pa[a] = s;
            }
            // parent accumulators: IA_parent = X^T Ia X, pA_parent = X^T pa  (own inertia added next iteration)
            T E[9];
            joint_E(rob, i, cs[i], cs[6 + i], E);
            T tmp[36];
#pragma unroll
            for (int col = 0; col < 6; ++col) {       // tmp = X^T Ia  (transform each column)
                T f[6], o[6];
#pragma unroll
                for (int a = 0; a < 6; ++a) f[a] = IA[6 * a + col];
                xfrc_t(E, rob.p[i], f, o);
#pragma unroll
                for (int a = 0; a < 6; ++a) tmp[6 * a + col] = o[a];
            }
#pragma unroll
            for (int row = 0; row < 6; ++row) {       // IA = tmp X  (transform each row; Ia symmetric)
                T o[6];
                xfrc_t(E, rob.p[i], tmp + 6 * row, o);
#pragma unroll
                for (int a = 0; a < 6; ++a) IA[6 * row + a] = o[a];
            }
            T o[6];
            xfrc_t(E, rob.p[i], pa, o);
#pragma unroll
            for (int a = 0; a < 6; ++a) pA[a] = o[a];
        }
    }
    // pass 3: accelerations, base to tip
    T acc[6] = {T(0), T(0), T(0), -rob.g[0], -rob.g[1], -rob.g[2]};
    T accs[I7_NJ][6];
#pragma unroll
    for (int i = 0; i < I7_NJ; ++i) {
        T E[9], ap[6];
        joint_E(rob, i, cs[i], cs[6 + i], E);
        xmot(E, rob.p[i], acc, ap);
        crm_S_acc(vel[i], qd[i], ap);
        T s = us[i];
#pragma unroll
        for (int a = 0; a < 6; ++a) s -= Us[i][a] * ap[a];
        const T qi = s / Ds[i];
        qdd[i] = qi;
#pragma unroll
        for (int a = 0; a < 6; ++a) acc[a] = ap[a];
        acc[2] += qi;
        if (WITH_AUX) {
#pragma unroll
            for (int a = 0; a < 6; ++a) accs[i][a] = acc[a];
        }
    }
    if (WITH_AUX) {
        // RNEA backward pass at the solved acceleration: total force across each joint
        T F[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
#pragma unroll
        for (int i = I7_NJ - 1; i >= 0; --i) {
            T Ia[6], Iv[6];
            inertia_mul(rob, i, accs[i], Ia);
            inertia_mul(rob, i, vel[i], Iv);
#pragma unroll
            for (int a = 0; a < 6; ++a) F[a] += Ia[a];
            crf<T, true>(vel[i], Iv, F);
            if (i == I7_NJ - 1 && fext) {
#pragma unroll
                for (int a = 0; a < 6; ++a) F[a] -= fext[a];
            }
#pragma unroll
            for (int a = 0; a < 6; ++a) F_out[6 * i + a] = F[a];
            if (i > 0) {
                T E[9], o[6];
                joint_E(rob, i, cs[i], cs[6 + i], E);
                xfrc_t(E, rob.p[i], F, o);
#pragma unroll
                for (int a = 0; a < 6; ++a) F[a] = o[a];
            }
        }
    }
}

// Column `col` of M^{-1}: ABA with zero velocity, zero gravity and unit torque on joint `col`, re-using the
// articulated U_i, D_i of the full ABA pass.  out[r] = Minv[r][col].
template <typename T>
I7_DEV void minv_column(const Robot<T>& rob, const T* cs, const T* U, const T* D, int col, T* out) {
    T pA[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
    T us[I7_NJ];
#pragma unroll
    for (int i = I7_NJ - 1; i >= 0; --i) {
        us[i] = (i == col ? T(1) : T(0)) - pA[2];
        if (i > 0) {
            const T f = us[i] / D[i];
            T pa[6];
#pragma unroll
            for (int a = 0; a < 6; ++a) pa[a] = pA[a] + U[6 * i + a] * f;
            T E[9];
            joint_E(rob, i, cs[i], cs[6 + i], E);
            xfrc_t(E, rob.p[i], pa, pA);
        }
    }
    T acc[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
#pragma unroll
    for (int i = 0; i < I7_NJ; ++i) {
        T E[9], ap[6];
        joint_E(rob, i, cs[i], cs[6 + i], E);
        xmot(E, rob.p[i], acc, ap);
        T s = us[i];
#pragma unroll
        for (int a = 0; a < 6; ++a) s -= U[6 * i + a] * ap[a];
        const T qi = s / D[i];
        out[i] = qi;
#pragma unroll
        for (int a = 0; a < 6; ++a) acc[a] = ap[a];
        acc[2] += qi;
    }
}

// Column j of the joint-space inertia matrix: M[:, j] = ID(q, qd = 0, qdd = e_j) without gravity (one RNEA pass
// with unit acceleration on joint j; links below j stay at rest).  out[i] = M[i][j].
template <typename T>
I7_DEV void crba_column(const Robot<T>& rob, const T* cs, int j, T* out) {
    T a[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
    T f[I7_NJ][6];
#pragma unroll
    for (int i = 0; i < I7_NJ; ++i) {
        if (i == j) {
#pragma unroll
            for (int k = 0; k < 6; ++k) a[k] = T(0);
            a[2] = T(1);
        } else if (i > j) {
            T E[9], t[6];
            joint_E(rob, i, cs[i], cs[6 + i], E);
            xmot(E, rob.p[i], a, t);
#pragma unroll
            for (int k = 0; k < 6; ++k) a[k] = t[k];
        }
        if (i >= j) inertia_mul(rob, i, a, f[i]);
    }
    T F[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
#pragma unroll
    for (int i = I7_NJ - 1; i >= 0; --i) {
        if (i >= j) {
#pragma unroll
            for (int k = 0; k < 6; ++k) F[k] += f[i][k];
        }
        out[i] = F[2];
        if (i > 0) {
            T E[9], o[6];
            joint_E(rob, i, cs[i], cs[6 + i], E);
            xfrc_t(E, rob.p[i], F, o);
#pragma unroll
            for (int k = 0; k < 6; ++k) F[k] = o[k];
        }
    }
}

// Bias torques b = ID(q, qd, qdd = 0) with gravity and the optional local wrench on the last link.
template <typename T>
I7_DEV void rnea_bias(const Robot<T>& rob, const T* cs, const T* qd, const T* fext, T* out) {
    T v[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
    T a[6] = {T(0), T(0), T(0), -rob.g[0], -rob.g[1], -rob.g[2]};
    T f[I7_NJ][6];
#pragma unroll
    for (int i = 0; i < I7_NJ; ++i) {
        T E[9], t[6];
        joint_E(rob, i, cs[i], cs[6 + i], E);
        xmot(E, rob.p[i], a, t);
#pragma unroll
        for (int k = 0; k < 6; ++k) a[k] = t[k];
        if (i == 0) {
            v[2] = qd[0];
        } else {
            xmot(E, rob.p[i], v, t);
#pragma unroll
            for (int k = 0; k < 6; ++k) v[k] = t[k];
            v[2] += qd[i];
        }
        crm_S_acc(v, qd[i], a);
        T Iv[6];
        inertia_mul(rob, i, a, f[i]);
        inertia_mul(rob, i, v, Iv);
        crf<T, true>(v, Iv, f[i]);
    }
    if (fext) {
#pragma unroll
        for (int k = 0; k < 6; ++k) f[I7_NJ - 1][k] -= fext[k];
    }
    T F[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
#pragma unroll
    for (int i = I7_NJ - 1; i >= 0; --i) {
#pragma unroll
        for (int k = 0; k < 6; ++k) F[k] += f[i][k];
        out[i] = F[2];
        if (i > 0) {
            T E[9], o[6];
            joint_E(rob, i, cs[i], cs[6 + i], E);
            xfrc_t(E, rob.p[i], F, o);
#pragma unroll
            for (int k = 0; k < 6; ++k) F[k] = o[k];
        }
    }
}

// Right-looking LDL' of a 6x6 SPD matrix held in registers (A is destroyed: unit-lower L below the diagonal),
// dinv[j] = 1 / D_j.  Dependency depth per pivot: one reciprocal, one multiply, one FMA.
template <typename T> I7_DEV void spd6_factor(T (&A)[6][6], T (&dinv)[6]) {
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        dinv[j] = T(1) / A[j][j];
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
template <typename T> I7_DEV void spd6_solve_factored(const T (&A)[6][6], const T (&dinv)[6], T* b) {
#pragma unroll
    for (int j = 0; j < 5; ++j)
#pragma unroll
        for (int i = j + 1; i < 6; ++i) b[i] -= A[i][j] * b[j];
#pragma unroll
    for (int i = 0; i < 6; ++i) b[i] *= dinv[i];
#pragma unroll
    for (int j = 5; j > 0; --j)
#pragma unroll
        for (int i = 0; i < j; ++i) b[i] -= A[j][i] * b[j];
}

// Column j of d tau/d q (WRT_Q) or d tau/d qd of inverse dynamics evaluated at (q, qd, qdd): forward-mode
// chain rule in link coordinates (the recursion of Carpentier & Mansard 2018 in local frames).  fext: local wrench
// on the last link (enters the total joint force of the WRT_Q seed) or nullptr.  out[i] = d tau_i / d (q|qd)_j.
template <typename T, bool WRT_Q>
I7_DEV void rnea_dcolumn(const Robot<T>& rob, const T* cs, const T* qd, const T* qdd, const T* fext, int j, T* out) {
    T v[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
    T a[6] = {T(0), T(0), T(0), -rob.g[0], -rob.g[1], -rob.g[2]};
    T dv[6], da[6];
    T df[I7_NJ][6];
    T fn[I7_NJ][6];          // nominal link forces (WRT_Q only): needed for S x* F_j
#pragma unroll
    for (int i = 0; i < I7_NJ; ++i) {
        T E[9], xa[6];
        joint_E(rob, i, cs[i], cs[6 + i], E);
        xmot(E, rob.p[i], a, xa);                 // X a_parent
        if (i == 0) {
            v[2] = qd[0];
        } else {
            T tv[6];
            xmot(E, rob.p[i], v, tv);
#pragma unroll
            for (int k = 0; k < 6; ++k) v[k] = tv[k];
            v[2] += qd[i];
        }
#pragma unroll
        for (int k = 0; k < 6; ++k) a[k] = xa[k];
        a[2] += qdd[i];
        crm_S_acc(v, qd[i], a);
        if (i == j) {
            // seeds:  d/dq_j: dv = v_j x S, da = (X a_parent) x S + dv x S qd_j ;  d/dqd_j: dv = S, da = v_j x S
            if (WRT_Q) {
                crm_S(v, T(1), dv);
                crm_S(xa, T(1), da);
                crm_S_acc(dv, qd[i], da);
            } else {
#pragma unroll
                for (int k = 0; k < 6; ++k) dv[k] = T(0);
                dv[2] = T(1);
                crm_S(v, T(1), da);
            }
        } else if (i > j) {
            T t1[6], t2[6];
            xmot(E, rob.p[i], dv, t1);
            xmot(E, rob.p[i], da, t2);
#pragma unroll
            for (int k = 0; k < 6; ++k) { dv[k] = t1[k]; da[k] = t2[k]; }
            crm_S_acc(dv, qd[i], da);
        }
        if (i >= j) {
            T Iv[6], Idv[6];
            inertia_mul(rob, i, da, df[i]);
            inertia_mul(rob, i, v, Iv);
            inertia_mul(rob, i, dv, Idv);
            crf<T, true>(dv, Iv, df[i]);
            crf<T, true>(v, Idv, df[i]);
            if (WRT_Q) {
                inertia_mul(rob, i, a, fn[i]);
                crf<T, true>(v, Iv, fn[i]);
                if (i == I7_NJ - 1 && fext) {
#pragma unroll
                    for (int k = 0; k < 6; ++k) fn[i][k] -= fext[k];
                }
            }
        }
    }
    T dF[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
    T F[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
#pragma unroll
    for (int i = I7_NJ - 1; i >= 0; --i) {
        if (i >= j) {
#pragma unroll
            for (int k = 0; k < 6; ++k) dF[k] += df[i][k];
            if (WRT_Q) {
#pragma unroll
                for (int k = 0; k < 6; ++k) F[k] += fn[i][k];
            }
        }
        out[i] = dF[2];
        if (WRT_Q && i == j) {
            // d(X_j^T F_j)/dq_j = X_j^T (dF_j + S x* F_j),  S x* F = [-n_y, n_x, 0, -f_y, f_x, 0]
            dF[0] -= F[1]; dF[1] += F[0];
            dF[3] -= F[4]; dF[4] += F[3];
        }
        if (i > 0) {
            T E[9], o[6];
            joint_E(rob, i, cs[i], cs[6 + i], E);
            xfrc_t(E, rob.p[i], dF, o);
#pragma unroll
            for (int k = 0; k < 6; ++k) dF[k] = o[k];
            if (WRT_Q && i > j) {
                xfrc_t(E, rob.p[i], F, o);
#pragma unroll
                for (int k = 0; k < 6; ++k) F[k] = o[k];
            }
        }
    }
}

// =========================================================================================================
// World-frame formulation (spatial vectors [angular; linear] expressed in world axes at ONE fixed spatial origin).
// In these coordinates the recursions over the links lose their transforms:  v_l = sum_{k<=l} S_k qd_k,
// a_l = a_base + sum_{k<=l} (S_k qdd_k + v_k x S_k qd_k),  F_i = sum_{l>=i} f_l,  tau_i = S_i . F_i,
// M_ij = S_i . (sum_{l>=max(i,j)} I_l) S_j,  and the partial derivatives of inverse dynamics have closed forms per
// (joint j, link l >= j) pair — so every quantity is a short prefix / suffix sum of independent terms instead of a
// 6-deep chain of 6x6 transforms.  Replaces pin.computeABADerivatives (src/osqp_solver.py:71); the recursion is the
// world-frame one of Carpentier & Mansard 2018, re-derived so that each (knot, link) and (knot, column) is an
// independent task.  World inertia of a link: IO (xx xy xz yy yz zz about the world origin) and the first moment H;
// the mass is the model constant.
// =========================================================================================================

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

// Task (knot, link l): forward kinematics of the whole chain (joint axes z_i and origins p_i), then — with the
// spatial origin placed at the last joint's position p_ref (a fixed point of space for the purpose of every
// derivative; it keeps the parallel-axis terms of the light wrist links small, which is what fp32 needs) — the
// velocity and the acceleration at qdd = 0 of link l (gravity as base acceleration), the link's axis S_l, inertia
// IW_l (9 numbers) and force f0_l = I_l a0_l + v_l x* I_l v_l (minus the external wrench on the last link, given in
// link coordinates).
template <typename T>
I7_DEV void wf_link(const Robot<T>& rob, const T* cs, const T* qd, const T* fext, int l, T* S_o, T* v_o, T* a_o,
                    T* IW_o, T* f_o) {
    T Rc[9] = {T(1), T(0), T(0), T(0), T(1), T(0), T(0), T(0), T(1)};
    T Rw[9], zs[I7_NJ][3], ps[I7_NJ][3];
    T pc[3] = {T(0), T(0), T(0)};
#pragma unroll
    for (int i = 0; i < I7_NJ; ++i) {
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
        zs[i][0] = Rc[2]; zs[i][1] = Rc[5]; zs[i][2] = Rc[8];
        ps[i][0] = pc[0]; ps[i][1] = pc[1]; ps[i][2] = pc[2];
        if (i == l) {
#pragma unroll
            for (int r = 0; r < 9; ++r) Rw[r] = Rc[r];
        }
    }
    T p[3] = {T(0), T(0), T(0)};
    T v[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
    T a[6] = {T(0), T(0), T(0), -rob.g[0], -rob.g[1], -rob.g[2]};
    T S[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
#pragma unroll
    for (int i = 0; i < I7_NJ; ++i) {
        if (i <= l) {
            p[0] = ps[i][0] - pc[0]; p[1] = ps[i][1] - pc[1]; p[2] = ps[i][2] - pc[2];
            S[0] = zs[i][0]; S[1] = zs[i][1]; S[2] = zs[i][2];
            cross3(p, S, S + 3);
            T vj[6], c6[6];
#pragma unroll
            for (int r = 0; r < 6; ++r) { vj[r] = S[r] * qd[i]; v[r] += vj[r]; }
            crm6(v, vj, c6);
#pragma unroll
            for (int r = 0; r < 6; ++r) a[r] += c6[r];
        }
    }
    // inertia about the spatial origin from (Ib about the link origin, h = m c, m) of link l; p = link origin
    const T* Ib = rob.Ib[l]; const T* h = rob.h[l]; const T ms = rob.m[l];
    T IW[9];
    {
        // B = Rw Ib (Ib symmetric), IO = B Rw^T
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
    T Ia[6], Iv[6];
    winertia_mul(IW, ms, a, Ia);
    winertia_mul(IW, ms, v, Iv);
    crf<T, true>(v, Iv, Ia);
    if (l == I7_NJ - 1 && fext) {
        T nw[3], fw[3];
        rot(Rw, fext, nw);              // the spatial origin IS the last joint's origin: no moment arm
        rot(Rw, fext + 3, fw);
#pragma unroll
        for (int r = 0; r < 3; ++r) { Ia[r] -= nw[r]; Ia[3 + r] -= fw[r]; }
    }
#pragma unroll
    for (int r = 0; r < 6; ++r) { S_o[r] = S[r]; v_o[r] = v[r]; a_o[r] = a[r]; f_o[r] = Ia[r]; }
#pragma unroll
    for (int r = 0; r < 9; ++r) IW_o[r] = IW[r];
}

// Task (knot, j): column j of the joint-space inertia matrix for rows i <= j and the bias torque b_j, from the
// per-link records:  M_ij = S_i . (sum_{l>=j} I_l) S_j,  b_j = S_j . sum_{l>=j} f0_l.   Mcol[i] for i <= j.
template <typename T>
I7_DEV void wf_mass_column(const Robot<T>& rob, const T* S /*36*/, const T* IW /*54*/, const T* f0 /*36*/, int j,
                           T* Mcol, T* bj) {
    T IC[9] = {T(0), T(0), T(0), T(0), T(0), T(0), T(0), T(0), T(0)};
    T F[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
    T mc = T(0);
#pragma unroll
    for (int l = I7_NJ - 1; l >= 0; --l) {
        if (l >= j) {
#pragma unroll
            for (int r = 0; r < 9; ++r) IC[r] += IW[9 * l + r];
#pragma unroll
            for (int r = 0; r < 6; ++r) F[r] += f0[6 * l + r];
            mc += rob.m[l];
        }
    }
    T Sj[6], y[6];
#pragma unroll
    for (int r = 0; r < 6; ++r) Sj[r] = S[6 * j + r];
    winertia_mul(IC, mc, Sj, y);
    T b = T(0);
#pragma unroll
    for (int r = 0; r < 6; ++r) b += Sj[r] * F[r];
    *bj = b;
#pragma unroll
    for (int i = 0; i < I7_NJ; ++i) {
        T d = T(0);
        if (i <= j) {
#pragma unroll
            for (int r = 0; r < 6; ++r) d += S[6 * i + r] * y[r];
        }
        Mcol[i] = d;
    }
}

// Task (knot, l): fold the solved accelerations into the link record:  a_l += sum_{k<=l} S_k qdd_k,
// f_l += I_l (that sum).
template <typename T>
I7_DEV void wf_add_qdd(const Robot<T>& rob, const T* S /*36*/, const T* qdd, int l, const T* IW_l, T* a_l, T* f_l) {
    T s[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
#pragma unroll
    for (int k = 0; k < I7_NJ; ++k) {
        if (k <= l) {
#pragma unroll
            for (int r = 0; r < 6; ++r) s[r] += S[6 * k + r] * qdd[k];
        }
    }
    T Is[6];
    winertia_mul(IW_l, rob.m[l], s, Is);
#pragma unroll
    for (int r = 0; r < 6; ++r) { a_l[r] += s[r]; f_l[r] += Is[r]; }
}

// Task (knot, j, kind): column j of d tau/dq (WRT_Q) or d tau/dqd at (q, qd, qdd), from the per-link records
// S, v, a, f (at the solved qdd), IW.  For every link l >= j
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
template <typename T, bool WRT_Q>
I7_DEV void wf_dcolumn(const Robot<T>& rob, const T* S, const T* V, const T* A, const T* F_, const T* IW, int j, T* out) {
    DSeed<T> d;
    wf_dseed<T, WRT_Q>(S, V, A, j, d);
    T rest[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
    T Fs[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
#pragma unroll
    for (int l = I7_NJ - 1; l >= 0; --l) {
        if (l >= j) {
            wf_drest<T>(rob, V, IW, l, d, rest);
            if (WRT_Q) {
#pragma unroll
                for (int r = 0; r < 6; ++r) Fs[r] += F_[6 * l + r];
            }
            if (l > j) {
                T dd = T(0);
#pragma unroll
                for (int r = 0; r < 6; ++r) dd += S[6 * l + r] * rest[r];
                out[l] = dd;
            }
        }
    }
    if (WRT_Q) crf<T, true>(d.Sj, Fs, rest);
#pragma unroll
    for (int i = 0; i < I7_NJ; ++i) {
        if (i <= j) {
            T dd = T(0);
#pragma unroll
            for (int r = 0; r < 6; ++r) dd += S[6 * i + r] * rest[r];
            out[i] = dd;
        }
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
    T Fs[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
#pragma unroll
    for (int l = I7_NJ - 1; l >= 0; --l) {
        if (l >= i0) {
#pragma unroll
            for (int r = 0; r < 6; ++r) rest[r] += pairs[6 * l + r];
        }
        if (WRT_Q && l >= j) {
#pragma unroll
            for (int r = 0; r < 6; ++r) Fs[r] += F_[6 * l + r];
        }
    }
    if (WRT_Q && i <= j) {
        T Sj[6];
#pragma unroll
        for (int r = 0; r < 6; ++r) Sj[r] = S[6 * j + r];
        crf<T, true>(Sj, Fs, rest);
    }
    T dd = T(0);
#pragma unroll
    for (int r = 0; r < 6; ++r) dd += S[6 * i + r] * rest[r];
    return dd;
}

}  // namespace i7
