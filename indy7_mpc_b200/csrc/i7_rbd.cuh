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


}  // namespace i7
