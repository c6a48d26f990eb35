/* oracle.c — plain-C (fp64) restatement of the reference SQP-MPC path.  TEST / BASELINE INFRASTRUCTURE ONLY.
 *
 * Follows, function by function, the reference files (paths relative to the reference checkout):
 *   src/osqp_solver.py:70-81   compute_dynamics_jacobians   -> linearize_knot()
 *   src/osqp_solver.py:83-101  update_constraint_matrix     -> assemble()  (block form of the same A, l)
 *   src/osqp_solver.py:103-135 update_cost_matrix           -> assemble()  (block form of the same P, g)
 *   src/osqp_solver.py:137-143 setup_and_solve_qp           -> solve_qp()  (OSQP-style ADMM or refined KKT solve)
 *   src/osqp_solver.py:146-155 eepos / d_eepos              -> fk_jac()
 *   src/osqp_sqp.py:13-47      eepos_cost / integrator_err  -> merit()
 *   src/osqp_sqp.py:49-74      linesearch                   -> linesearch()
 *   src/osqp_sqp.py:76-93      sqp                          -> orc_sqp()
 *   src/utils.py:3-18          rk4                          -> orc_rk4()
 * and restates the published algorithms of the two un-vendored third-party libraries that carry the arithmetic:
 * Pinocchio (ABA / RNEA derivatives, Featherstone 2008; Carpentier & Mansard 2018) and OSQP (ADMM with a
 * quasi-definite KKT factorisation, Stellato et al. 2020; defaults rho=0.1 x 1e3 on equality rows, sigma=1e-6,
 * alpha=1.6, eps_abs=eps_rel=1e-3, check_termination=25, warm start; Ruiz scaling and the wall-clock-driven
 * adaptive rho are not restated).  The KKT matrix is factorised by a banded LDL' in the stage-interleaved
 * ordering [lambda_k, x_k, u_k] (half bandwidth 41), the dense-banded analogue of OSQP's QDLDL.
 *
 * This file is the timed CPU baseline of bench.py ("cpu_baseline.kind" = "port") and is cross-checked against
 * the numpy oracle in tests/test_oracle_c.py.  The product never links or loads it.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define NJ 6
#define NX 12
#define NXU 18

typedef struct {
    double R[NJ][9];   /* joint placement rotation (row-major) */
    double p[NJ][3];   /* joint placement translation */
    double I[NJ][36];  /* 6x6 spatial inertia about the link origin, [angular; linear], row-major */
    double g[3];
} orc_model;

typedef struct {
    int N;
    double dt, dQ, R, QN, eps;
    int regularize;
    int backend;       /* 0: refined KKT solve (exact), 1: OSQP-style ADMM */
    double mu, step_tol;
    int max_qp;
    /* extension: box limits of description/indy7.urdf:203-238 as inequality rows
     * lo <= z_i <= hi on every q_k, v_k (k >= 1) and u_k; ADMM back-end only.  rho of the state rows = rho_x (0.1, OSQP's
     * rho for inequality rows; equality rows carry 1e3 x 0.1), of the effort rows rho_u (1e-3: without OSQP's Ruiz
     * scaling a rho far above the 1e-5 curvature of the control cost stalls the iteration). max_iter: 0 = 4000. */
    int limits;
    double qmin[NJ], qmax[NJ], vmax[NJ], umax[NJ];
    double rho_x, rho_u;
    int max_iter;
} orc_cfg;

/* ---------------------------------------------------------------- spatial algebra ---------------------- */
static void cross(const double* a, const double* b, double* o) {
    o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}
static void joint_E(const orc_model* m, int i, double q, double* E) {
    double c = cos(q), s = sin(q);
    for (int a = 0; a < 3; ++a) {
        E[a] = m->R[i][3 * a] * c + m->R[i][3 * a + 1] * s;
        E[3 + a] = -m->R[i][3 * a] * s + m->R[i][3 * a + 1] * c;
        E[6 + a] = m->R[i][3 * a + 2];
    }
}
static void mv3(const double* E, const double* x, double* o) {
    for (int a = 0; a < 3; ++a) o[a] = E[3 * a] * x[0] + E[3 * a + 1] * x[1] + E[3 * a + 2] * x[2];
}
static void mtv3(const double* E, const double* x, double* o) {
    for (int a = 0; a < 3; ++a) o[a] = E[a] * x[0] + E[3 + a] * x[1] + E[6 + a] * x[2];
}
static void xmot(const double* E, const double* r, const double* v, double* o) {
    double t[3], c[3];
    cross(r, v, c);
    for (int a = 0; a < 3; ++a) t[a] = v[3 + a] - c[a];
    mv3(E, v, o); mv3(E, t, o + 3);
}
static void xfrc_t(const double* E, const double* r, const double* f, double* o) {
    double n[3], c[3];
    mtv3(E, f, n); mtv3(E, f + 3, o + 3);
    cross(r, o + 3, c);
    for (int a = 0; a < 3; ++a) o[a] = n[a] + c[a];
}
static void crm(const double* v, const double* m, double* o) {
    double a[3], b[3];
    cross(v, m, o); cross(v, m + 3, a); cross(v + 3, m, b);
    for (int i = 0; i < 3; ++i) o[3 + i] = a[i] + b[i];
}
static void crf(const double* v, const double* f, double* o) {
    double a[3], b[3];
    cross(v, f, a); cross(v + 3, f + 3, b); cross(v, f + 3, o + 3);
    for (int i = 0; i < 3; ++i) o[i] = a[i] + b[i];
}
static void mv6(const double* M, const double* x, double* o) {
    for (int a = 0; a < 6; ++a) { double s = 0; for (int b = 0; b < 6; ++b) s += M[6 * a + b] * x[b]; o[a] = s; }
}

/* FK of the last joint frame and translational LOCAL_WORLD_ALIGNED Jacobian (3x6 row-major) */
static void fk_jac(const orc_model* m, const double* q, double* pee, double* J) {
    double Rw[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, p[3] = {0, 0, 0}, zs[NJ][3], ps[NJ][3];
    for (int i = 0; i < NJ; ++i) {
        double d[3], E[9], Rn[9];
        mv3(Rw, m->p[i], d);
        for (int a = 0; a < 3; ++a) p[a] += d[a];
        joint_E(m, i, q[i], E);
        for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) {
            double s = 0; for (int k = 0; k < 3; ++k) s += Rw[3 * a + k] * E[3 * b + k]; Rn[3 * a + b] = s; }
        memcpy(Rw, Rn, sizeof Rw);
        for (int a = 0; a < 3; ++a) { zs[i][a] = Rw[3 * a + 2]; ps[i][a] = p[a]; }
    }
    for (int a = 0; a < 3; ++a) pee[a] = p[a];
    if (J) for (int i = 0; i < NJ; ++i) {
        double d[3], c[3];
        for (int a = 0; a < 3; ++a) d[a] = p[a] - ps[i][a];
        cross(zs[i], d, c);
        for (int a = 0; a < 3; ++a) J[6 * a + i] = c[a];
    }
}

typedef struct { double E[NJ][9], v[NJ][6], a[NJ][6], F[NJ][6]; } rnea_aux;

/* ABA (Featherstone RBDA Table 7.1).  fext: Featherstone-ordered local wrench on the last link or NULL. */
static void aba(const orc_model* m, const double* q, const double* qd, const double* tau, const double* fext,
                double* qdd, rnea_aux* aux) {
    double E[NJ][9], v[NJ][6], c[NJ][6], IA[NJ][36], pA[NJ][6], U[NJ][6], D[NJ], u[NJ];
    for (int i = 0; i < NJ; ++i) {
        joint_E(m, i, q[i], E[i]);
        if (i == 0) memset(v[0], 0, sizeof v[0]); else xmot(E[i], m->p[i], v[i - 1], v[i]);
        v[i][2] += qd[i];
        double vj[6] = {0, 0, qd[i], 0, 0, 0}, Iv[6];
        crm(v[i], vj, c[i]);
        memcpy(IA[i], m->I[i], sizeof IA[i]);
        mv6(m->I[i], v[i], Iv);
        crf(v[i], Iv, pA[i]);
        if (fext && i == NJ - 1) for (int a = 0; a < 6; ++a) pA[i][a] -= fext[a];
    }
    for (int i = NJ - 1; i >= 0; --i) {
        for (int a = 0; a < 6; ++a) U[i][a] = IA[i][6 * a + 2];
        D[i] = U[i][2];
        u[i] = tau[i] - pA[i][2];
        if (i > 0) {
            double Ia[36], pa[6], t[6], tmp[36];
            for (int a = 0; a < 6; ++a) for (int b = 0; b < 6; ++b) Ia[6 * a + b] = IA[i][6 * a + b] - U[i][a] * U[i][b] / D[i];
            mv6(Ia, c[i], t);
            for (int a = 0; a < 6; ++a) pa[a] = pA[i][a] + t[a] + U[i][a] * u[i] / D[i];
            for (int col = 0; col < 6; ++col) {
                double f[6], o[6];
                for (int a = 0; a < 6; ++a) f[a] = Ia[6 * a + col];
                xfrc_t(E[i], m->p[i], f, o);
                for (int a = 0; a < 6; ++a) tmp[6 * a + col] = o[a];
            }
            for (int row = 0; row < 6; ++row) {
                double o[6];
                xfrc_t(E[i], m->p[i], tmp + 6 * row, o);
                for (int a = 0; a < 6; ++a) IA[i - 1][6 * row + a] += o[a];
            }
            xfrc_t(E[i], m->p[i], pa, t);
            for (int a = 0; a < 6; ++a) pA[i - 1][a] += t[a];
        }
    }
    double acc[6] = {0, 0, 0, -m->g[0], -m->g[1], -m->g[2]}, accs[NJ][6];
    for (int i = 0; i < NJ; ++i) {
        double ap[6], s = u[i];
        xmot(E[i], m->p[i], acc, ap);
        for (int a = 0; a < 6; ++a) { ap[a] += c[i][a]; s -= U[i][a] * ap[a]; }
        qdd[i] = s / D[i];
        memcpy(acc, ap, sizeof acc);
        acc[2] += qdd[i];
        memcpy(accs[i], acc, sizeof acc);
    }
    if (aux) {
        double F[6] = {0, 0, 0, 0, 0, 0};
        for (int i = NJ - 1; i >= 0; --i) {
            double Ia[6], Iv[6], b[6];
            mv6(m->I[i], accs[i], Ia); mv6(m->I[i], v[i], Iv); crf(v[i], Iv, b);
            for (int a = 0; a < 6; ++a) F[a] += Ia[a] + b[a];
            if (fext && i == NJ - 1) for (int a = 0; a < 6; ++a) F[a] -= fext[a];
            memcpy(aux->F[i], F, sizeof F);
            if (i > 0) { double o[6]; xfrc_t(E[i], m->p[i], F, o); memcpy(F, o, sizeof F); }
        }
        memcpy(aux->E, E, sizeof E); memcpy(aux->v, v, sizeof v); memcpy(aux->a, accs, sizeof accs);
    }
}

/* joint-space inertia by CRBA, then its inverse by Cholesky */
static void minv(const orc_model* m, const rnea_aux* aux, double* Mi) {
    double IC[NJ][36], M[36];
    for (int i = 0; i < NJ; ++i) memcpy(IC[i], m->I[i], sizeof IC[i]);
    for (int i = NJ - 1; i > 0; --i) {
        double tmp[36];
        for (int col = 0; col < 6; ++col) {
            double f[6], o[6];
            for (int a = 0; a < 6; ++a) f[a] = IC[i][6 * a + col];
            xfrc_t(aux->E[i], m->p[i], f, o);
            for (int a = 0; a < 6; ++a) tmp[6 * a + col] = o[a];
        }
        for (int row = 0; row < 6; ++row) {
            double o[6];
            xfrc_t(aux->E[i], m->p[i], tmp + 6 * row, o);
            for (int a = 0; a < 6; ++a) IC[i - 1][6 * row + a] += o[a];
        }
    }
    for (int i = 0; i < NJ; ++i) {
        double F[6];
        for (int a = 0; a < 6; ++a) F[a] = IC[i][6 * a + 2];
        M[6 * i + i] = F[2];
        for (int j = i; j > 0; --j) {
            double o[6];
            xfrc_t(aux->E[j], m->p[j], F, o);
            memcpy(F, o, sizeof F);
            M[6 * i + j - 1] = M[6 * (j - 1) + i] = F[2];
        }
    }
    double L[36] = {0};
    for (int j = 0; j < 6; ++j) {
        double d = M[6 * j + j];
        for (int k = 0; k < j; ++k) d -= L[6 * j + k] * L[6 * j + k];
        L[6 * j + j] = sqrt(d);
        for (int i = j + 1; i < 6; ++i) {
            double s = M[6 * i + j];
            for (int k = 0; k < j; ++k) s -= L[6 * i + k] * L[6 * j + k];
            L[6 * i + j] = s / L[6 * j + j];
        }
    }
    for (int c = 0; c < 6; ++c) {
        double y[6];
        for (int i = 0; i < 6; ++i) {
            double s = (i == c) ? 1.0 : 0.0;
            for (int k = 0; k < i; ++k) s -= L[6 * i + k] * y[k];
            y[i] = s / L[6 * i + i];
        }
        for (int i = 5; i >= 0; --i) {
            double s = y[i];
            for (int k = i + 1; k < 6; ++k) s -= L[6 * k + i] * Mi[6 * k + c];
            Mi[6 * i + c] = s / L[6 * i + i];
        }
    }
}

/* d tau/dq, d tau/dqd of RNEA at (q, qd, qdd) by forward-mode chain rule in link coordinates */
static void rnea_derivs(const orc_model* m, const double* qd, const double* qdd, const rnea_aux* aux, double* dq, double* dv) {
    const double S[6] = {0, 0, 1, 0, 0, 0};
    for (int j = 0; j < NJ; ++j) {
        for (int wrt = 0; wrt < 2; ++wrt) {
            double dvel[6], dacc[6], df[NJ][6], t[6], t2[6];
            if (wrt == 0) {
                double Xa[6], vS[6];
                crm(aux->v[j], S, vS);
                for (int a = 0; a < 6; ++a) Xa[a] = aux->a[j][a] - S[a] * qdd[j] - vS[a] * qd[j];
                memcpy(dvel, vS, sizeof vS);
                crm(Xa, S, dacc); crm(vS, S, t);
                for (int a = 0; a < 6; ++a) dacc[a] += t[a] * qd[j];
            } else {
                memcpy(dvel, S, sizeof S);
                crm(aux->v[j], S, dacc);
            }
            for (int i = j; i < NJ; ++i) {
                if (i > j) {
                    xmot(aux->E[i], m->p[i], dvel, t); memcpy(dvel, t, sizeof t);
                    xmot(aux->E[i], m->p[i], dacc, t); crm(dvel, S, t2);
                    for (int a = 0; a < 6; ++a) dacc[a] = t[a] + t2[a] * qd[i];
                }
                double Iv[6], Idv[6], b1[6], b2[6];
                mv6(m->I[i], dacc, df[i]); mv6(m->I[i], aux->v[i], Iv); mv6(m->I[i], dvel, Idv);
                crf(dvel, Iv, b1); crf(aux->v[i], Idv, b2);
                for (int a = 0; a < 6; ++a) df[i][a] += b1[a] + b2[a];
            }
            double dF[6] = {0, 0, 0, 0, 0, 0};
            double* out = wrt == 0 ? dq : dv;
            for (int i = NJ - 1; i >= 0; --i) {
                if (i >= j) for (int a = 0; a < 6; ++a) dF[a] += df[i][a];
                out[6 * i + j] = dF[2];
                if (i == j && wrt == 0) { crf(S, aux->F[i], t); for (int a = 0; a < 6; ++a) dF[a] += t[a]; }
                if (i > 0) { xfrc_t(aux->E[i], m->p[i], dF, t); memcpy(dF, t, sizeof t); }
            }
        }
    }
}

/* ---------------------------------------------------------------- QP data in block form ---------------- */
typedef struct {
    int N;
    double *Gq, *Gv, *H, *cv;      /* (N-1) x 36/36/36/6 : A_k = [[I, dt I],[Gq, Gv]], B_k = [0; H], c_k = [0; cv] */
    double *j, *Qk, *dQk, *Rk;     /* N x 6, N, N, N */
    /* ADMM state (warm start across calls, like the OSQP object held by the reference solver) */
    double *x, *y;
    double *yl, *zl;               /* duals / auxiliary variables of the n - 12 limit rows (z index 12 + i) */
    double rp, rd;                 /* residuals at the last termination check */
    int n, mrows;
    /* banded workspace */
    double *band, *rhs, *sol, *perm_work;
    int nk, bw;
} orc_work;

static void linearize_knot(const orc_model* m, const orc_cfg* c, const double* xk, const double* fext, double* Gq,
                           double* Gv, double* H, double* cv) {
    double qdd[6], Mi[36], dq[36], dv[36];
    rnea_aux aux;
    aba(m, xk, xk + 6, xk + 12, fext, qdd, &aux);
    minv(m, &aux, Mi);
    rnea_derivs(m, xk + 6, qdd, &aux, dq, dv);
    for (int r = 0; r < 6; ++r) for (int cc = 0; cc < 6; ++cc) {
        double sq = 0, sv = 0;
        for (int k = 0; k < 6; ++k) { sq -= Mi[6 * r + k] * dq[6 * k + cc]; sv -= Mi[6 * r + k] * dv[6 * k + cc]; }
        Gq[6 * r + cc] = sq * c->dt;
        Gv[6 * r + cc] = sv * c->dt + (r == cc ? 1.0 : 0.0);
        H[6 * r + cc] = Mi[6 * r + cc] * c->dt;
    }
    for (int r = 0; r < 6; ++r) {
        double s = xk[6 + r] + qdd[r] * c->dt;
        for (int k = 0; k < 6; ++k) s -= Gq[6 * r + k] * xk[k] + Gv[6 * r + k] * xk[6 + k] + H[6 * r + k] * xk[12 + k];
        cv[r] = s;
    }
}

static void assemble(const orc_model* m, const orc_cfg* c, const double* xu, const double* goals, const double* fext, orc_work* w) {
    const int N = c->N;
    for (int k = 0; k < N; ++k) {
        const double* xk = xu + NXU * k;
        double pee[3], J[18], e[3];
        fk_jac(m, xk, pee, J);
        for (int a = 0; a < 3; ++a) e[a] = pee[a] - goals[3 * k + a];
        double nrm = sqrt(e[0] * e[0] + e[1] * e[1] + e[2] * e[2]);
        double sc = c->regularize ? 1.0 / (fabs(nrm) + c->eps) : 1.0;
        for (int i = 0; i < 6; ++i) w->j[6 * k + i] = e[0] * J[i] + e[1] * J[6 + i] + e[2] * J[12 + i];
        w->Qk[k] = (k == N - 1) ? c->QN : 1.0;
        w->dQk[k] = c->dQ * sc;
        w->Rk[k] = c->R * sc;
        if (k < N - 1) linearize_knot(m, c, xk, fext, w->Gq + 36 * k, w->Gv + 36 * k, w->H + 36 * k, w->cv + 6 * k);
    }
}

/* z index helpers */
static inline int zx(int k) { return NXU * k; }
static inline int zu(int k) { return NXU * k + NX; }

/* y = P z ; P block-diagonal: Qk j j' on q, dQk on v, Rk on u */
static void P_mul(const orc_work* w, const double* z, double* y) {
    const int N = w->N;
    for (int k = 0; k < N; ++k) {
        const double* jk = w->j + 6 * k;
        double d = 0;
        for (int i = 0; i < 6; ++i) d += jk[i] * z[zx(k) + i];
        for (int i = 0; i < 6; ++i) {
            y[zx(k) + i] = w->Qk[k] * jk[i] * d;
            y[zx(k) + 6 + i] = w->dQk[k] * z[zx(k) + 6 + i];
            if (k < N - 1) y[zu(k) + i] = w->Rk[k] * z[zu(k) + i];
        }
    }
}
/* r = A z  (rows: -x_0 ; A_k x_k + B_k u_k - x_{k+1}) */
static void A_mul(const orc_work* w, double dt, const double* z, double* r) {
    const int N = w->N;
    for (int i = 0; i < NX; ++i) r[i] = -z[i];
    for (int k = 0; k < N - 1; ++k) {
        const double *Gq = w->Gq + 36 * k, *Gv = w->Gv + 36 * k, *H = w->H + 36 * k;
        const double *x = z + zx(k), *u = z + zu(k), *xn = z + zx(k + 1);
        double* rr = r + NX * (k + 1);
        for (int i = 0; i < 6; ++i) {
            rr[i] = x[i] + dt * x[6 + i] - xn[i];
            double s = -xn[6 + i];
            for (int c = 0; c < 6; ++c) s += Gq[6 * i + c] * x[c] + Gv[6 * i + c] * x[6 + c] + H[6 * i + c] * u[c];
            rr[6 + i] = s;
        }
    }
}
/* z = A' y */
static void At_mul(const orc_work* w, double dt, const double* y, double* z) {
    const int N = w->N;
    const int n = NXU * N - 6;
    memset(z, 0, sizeof(double) * n);
    for (int i = 0; i < NX; ++i) z[i] -= y[i];
    for (int k = 0; k < N - 1; ++k) {
        const double *Gq = w->Gq + 36 * k, *Gv = w->Gv + 36 * k, *H = w->H + 36 * k;
        const double* yy = y + NX * (k + 1);
        double *x = z + zx(k), *u = z + zu(k), *xn = z + zx(k + 1);
        for (int i = 0; i < 6; ++i) {
            x[i] += yy[i]; x[6 + i] += dt * yy[i]; xn[i] -= yy[i]; xn[6 + i] -= yy[6 + i];
            for (int c = 0; c < 6; ++c) {
                x[c] += Gq[6 * i + c] * yy[6 + i]; x[6 + c] += Gv[6 * i + c] * yy[6 + i]; u[c] += H[6 * i + c] * yy[6 + i];
            }
        }
    }
}
static void g_vec(const orc_work* w, const double* xu, double* g) {
    const int N = w->N;
    for (int k = 0; k < N; ++k) for (int i = 0; i < 6; ++i) {
        g[zx(k) + i] = w->Qk[k] * w->j[6 * k + i];
        g[zx(k) + 6 + i] = w->dQk[k] * xu[zx(k) + 6 + i];
        if (k < N - 1) g[zu(k) + i] = w->Rk[k] * xu[zu(k) + i];
    }
}
static void l_vec(const orc_work* w, const double* xs, double* l) {
    const int N = w->N;
    for (int i = 0; i < NX; ++i) l[i] = -xs[i];
    for (int k = 0; k < N - 1; ++k) for (int i = 0; i < 6; ++i) { l[NX * (k + 1) + i] = 0.0; l[NX * (k + 1) + 6 + i] = -w->cv[6 * k + i]; }
}

/* ---- banded symmetric LDL' of K = [[P + sigma I, A'], [A, -delta I]] in the ordering [lam_k, x_k, u_k] ---- */
#define BW 30
static inline int pos_lam(int k) { return 30 * k; }
static inline int pos_x(int k) { return 30 * k + 12; }
static inline int pos_u(int k) { return 30 * k + 24; }
static inline void band_add(double* band, int i, int j, double v) {   /* lower band storage: band[i][i-j] */
    if (i < j) { int t = i; i = j; j = t; }
    band[(size_t)i * (BW + 1) + (i - j)] += v;
}
static void kkt_build(const orc_work* w, double dt, double sigma, double delta, double rho_x, double rho_u, double* band) {
    const int N = w->N, nk = 30 * N - 6;
    memset(band, 0, sizeof(double) * (size_t)nk * (BW + 1));
    for (int k = 0; k < N; ++k) {
        const double* jk = w->j + 6 * k;
        for (int a = 0; a < 6; ++a) {
            for (int b = 0; b <= a; ++b) band_add(band, pos_x(k) + a, pos_x(k) + b, w->Qk[k] * jk[a] * jk[b]);
            /* limit rows are identity rows: eliminating their multipliers adds rho_lim to the diagonal (x_0 has none) */
            band_add(band, pos_x(k) + a, pos_x(k) + a, sigma + (k > 0 ? rho_x : 0.0));
            band_add(band, pos_x(k) + 6 + a, pos_x(k) + 6 + a, w->dQk[k] + sigma + (k > 0 ? rho_x : 0.0));
            if (k < N - 1) band_add(band, pos_u(k) + a, pos_u(k) + a, w->Rk[k] + sigma + rho_u);
        }
        for (int a = 0; a < NX; ++a) band_add(band, pos_lam(k) + a, pos_lam(k) + a, -delta);
    }
    for (int a = 0; a < NX; ++a) band_add(band, pos_lam(0) + a, pos_x(0) + a, -1.0);
    for (int k = 0; k < N - 1; ++k) {
        const double *Gq = w->Gq + 36 * k, *Gv = w->Gv + 36 * k, *H = w->H + 36 * k;
        const int lr = pos_lam(k + 1);
        for (int i = 0; i < 6; ++i) {
            band_add(band, lr + i, pos_x(k) + i, 1.0);
            band_add(band, lr + i, pos_x(k) + 6 + i, dt);
            band_add(band, lr + i, pos_x(k + 1) + i, -1.0);
            band_add(band, lr + 6 + i, pos_x(k + 1) + 6 + i, -1.0);
            for (int c = 0; c < 6; ++c) {
                band_add(band, lr + 6 + i, pos_x(k) + c, Gq[6 * i + c]);
                band_add(band, lr + 6 + i, pos_x(k) + 6 + c, Gv[6 * i + c]);
                band_add(band, lr + 6 + i, pos_u(k) + c, H[6 * i + c]);
            }
        }
    }
}
static void band_factor(double* band, int n) {      /* in place: unit L below the diagonal, D on it */
    for (int j = 0; j < n; ++j) {
        double d = band[(size_t)j * (BW + 1)];
        int k0 = j - BW < 0 ? 0 : j - BW;
        for (int k = k0; k < j; ++k) { double l = band[(size_t)j * (BW + 1) + (j - k)]; d -= l * l * band[(size_t)k * (BW + 1)]; }
        band[(size_t)j * (BW + 1)] = d;
        int imax = j + BW < n - 1 ? j + BW : n - 1;
        for (int i = j + 1; i <= imax; ++i) {
            double s = band[(size_t)i * (BW + 1) + (i - j)];
            int kk0 = i - BW < 0 ? 0 : i - BW; if (kk0 < k0) kk0 = k0;
            for (int k = kk0; k < j; ++k)
                s -= band[(size_t)i * (BW + 1) + (i - k)] * band[(size_t)j * (BW + 1) + (j - k)] * band[(size_t)k * (BW + 1)];
            band[(size_t)i * (BW + 1) + (i - j)] = s / d;
        }
    }
}
static void band_solve(const double* band, int n, double* b) {
    for (int i = 0; i < n; ++i) {
        int k0 = i - BW < 0 ? 0 : i - BW; double s = b[i];
        for (int k = k0; k < i; ++k) s -= band[(size_t)i * (BW + 1) + (i - k)] * b[k];
        b[i] = s;
    }
    for (int i = 0; i < n; ++i) b[i] /= band[(size_t)i * (BW + 1)];
    for (int i = n - 1; i >= 0; --i) {
        int kmax = i + BW < n - 1 ? i + BW : n - 1; double s = b[i];
        for (int k = i + 1; k <= kmax; ++k) s -= band[(size_t)k * (BW + 1) + (k - i)] * b[k];
        b[i] = s;
    }
}
/* scatter (z-part, y-part) <-> permuted KKT vector */
static void to_perm(int N, const double* zp, const double* yp, double* b) {
    for (int k = 0; k < N; ++k) {
        memcpy(b + pos_lam(k), yp + NX * k, sizeof(double) * NX);
        memcpy(b + pos_x(k), zp + zx(k), sizeof(double) * NX);
        if (k < N - 1) memcpy(b + pos_u(k), zp + zu(k), sizeof(double) * 6);
    }
}
static void from_perm(int N, const double* b, double* zp, double* yp) {
    for (int k = 0; k < N; ++k) {
        memcpy(yp + NX * k, b + pos_lam(k), sizeof(double) * NX);
        memcpy(zp + zx(k), b + pos_x(k), sizeof(double) * NX);
        if (k < N - 1) memcpy(zp + zu(k), b + pos_u(k), sizeof(double) * 6);
    }
}
static double inf_norm(const double* a, int n) { double m = 0; for (int i = 0; i < n; ++i) { double v = fabs(a[i]); if (v > m) m = v; } return m; }

/* returns the number of ADMM iterations (backend 1) or refinement steps (backend 0) */
static int solve_qp(const orc_cfg* c, orc_work* w, const double* xu, const double* xs, double* zout) {
    const int N = c->N, n = NXU * N - 6, mr = NX * N, nk = n + mr;
    double *g = (double*)malloc(sizeof(double) * (size_t)(6 * n + 6 * mr + 2 * nk));
    double *l = g + n, *t1 = l + mr, *t2 = t1 + n, *t3 = t2 + n, *r1 = t3 + n, *r2 = r1 + mr, *b = r2 + mr, *zt = b + nk, *nu = zt + n;
    double *xprev = nu + mr, *Ax = xprev + n;   /* sized within the 6n + 6mr pool */
    g_vec(w, xu, g); l_vec(w, xs, l);
    int iters = 0;
    if (c->backend == 0) {
        /* [[P + d I, A'], [A, -d I]] with d = 1e-8, iterative refinement on the true KKT residual */
        const double d = 1e-8;
        kkt_build(w, c->dt, d, d, 0.0, 0.0, w->band);
        band_factor(w->band, nk);
        memset(zt, 0, sizeof(double) * n); memset(nu, 0, sizeof(double) * mr);
        for (iters = 0; iters < 30; ++iters) {
            /* residual: r_z = -g - P z - A' y ; r_y = l - A z */
            P_mul(w, zt, t1); At_mul(w, c->dt, nu, t2); A_mul(w, c->dt, zt, r1);
            for (int i = 0; i < n; ++i) t3[i] = -g[i] - t1[i] - t2[i];
            for (int i = 0; i < mr; ++i) r2[i] = l[i] - r1[i];
            if (iters > 0 && inf_norm(t3, n) < 1e-11 * (1 + inf_norm(g, n)) && inf_norm(r2, mr) < 1e-12 * (1 + inf_norm(l, mr))) break;
            to_perm(N, t3, r2, b);
            band_solve(w->band, nk, b);
            from_perm(N, b, t1, r1);
            for (int i = 0; i < n; ++i) zt[i] += t1[i];
            for (int i = 0; i < mr; ++i) nu[i] += r1[i];
        }
        memcpy(zout, zt, sizeof(double) * n);
    } else {
        const double rho = 0.1 * 1e3, sigma = 1e-6, alpha = 1.6, eps_abs = 1e-3, eps_rel = 1e-3;
        const int lim = c->limits, nl = lim ? n - NX : 0;
        const double rho_x = lim ? c->rho_x : 0.0, rho_u = lim ? c->rho_u : 0.0;
        const int max_iter = c->max_iter > 0 ? c->max_iter : 4000;
        /* osqp.update(Px) and osqp.update(Ax) each refactorise (src/osqp_solver.py:140-141) */
        kkt_build(w, c->dt, sigma, 1.0 / rho, rho_x, rho_u, w->band); band_factor(w->band, nk);
        kkt_build(w, c->dt, sigma, 1.0 / rho, rho_x, rho_u, w->band); band_factor(w->band, nk);
        double *x = w->x, *y = w->y, *yl = w->yl, *zl = w->zl;
        double* z = r2;                     /* z = projection onto {l} = l after the first iteration */
        A_mul(w, c->dt, x, z);              /* OSQP warm start: z = A x */
        for (int i = 0; i < nl; ++i) zl[i] = x[NX + i];
        w->rp = w->rd = 0;
        for (iters = 1; iters <= max_iter; ++iters) {
            for (int i = 0; i < n; ++i) t1[i] = sigma * x[i] - g[i];
            /* the limit rows [I; -1/rho_l] are eliminated by hand: nu_l = rho_l (x~ - z_l) + y_l */
            for (int i = 0; i < nl; ++i) t1[NX + i] += ((NX + i) % NXU < NX ? rho_x : rho_u) * zl[i] - yl[i];
            for (int i = 0; i < mr; ++i) r1[i] = z[i] - y[i] / rho;
            to_perm(N, t1, r1, b);
            band_solve(w->band, nk, b);
            from_perm(N, b, zt, nu);
            for (int i = 0; i < n; ++i) x[i] = alpha * zt[i] + (1 - alpha) * x[i];
            for (int i = 0; i < mr; ++i) {
                double ztil = z[i] + (nu[i] - y[i]) / rho;
                double zr = alpha * ztil + (1 - alpha) * z[i];
                y[i] += rho * (zr - l[i]);
                z[i] = l[i];
            }
            for (int i = 0; i < nl; ++i) {
                const int e = (NX + i) % NXU;
                const double rho_l = e < NX ? rho_x : rho_u;
                const double lo = e < 6 ? c->qmin[e] : (e < 12 ? -c->vmax[e - 6] : -c->umax[e - 12]);
                const double hi = e < 6 ? c->qmax[e] : (e < 12 ? c->vmax[e - 6] : c->umax[e - 12]);
                double zr = alpha * zt[NX + i] + (1 - alpha) * zl[i];          /* z~_l = x~_l */
                double zn = zr + yl[i] / rho_l;
                zn = zn < lo ? lo : (zn > hi ? hi : zn);
                yl[i] += rho_l * (zr - zn);
                zl[i] = zn;
            }
            if (iters % 25 == 0) {
                A_mul(w, c->dt, x, Ax); P_mul(w, x, t1); At_mul(w, c->dt, y, t2);
                for (int i = 0; i < nl; ++i) t2[NX + i] += yl[i];
                double rp = 0, rd = 0, nAx = inf_norm(Ax, mr), nz = inf_norm(l, mr);
                for (int i = 0; i < mr; ++i) { double v = fabs(Ax[i] - l[i]); if (v > rp) rp = v; }
                for (int i = 0; i < nl; ++i) {
                    double v = fabs(x[NX + i] - zl[i]); if (v > rp) rp = v;
                    v = fabs(x[NX + i]); if (v > nAx) nAx = v;
                    v = fabs(zl[i]); if (v > nz) nz = v;
                }
                for (int i = 0; i < n; ++i) { double v = fabs(t1[i] + g[i] + t2[i]); if (v > rd) rd = v; }
                double ep = eps_abs + eps_rel * fmax(nAx, nz);
                double ed = eps_abs + eps_rel * fmax(fmax(inf_norm(t1, n), inf_norm(t2, n)), inf_norm(g, n));
                w->rp = rp; w->rd = rd;
                if (rp <= ep && rd <= ed) break;
            }
        }
        memcpy(zout, x, sizeof(double) * n);
    }
    free(g);
    if (c->backend == 1 && iters > (c->max_iter > 0 ? c->max_iter : 4000)) iters = c->max_iter > 0 ? c->max_iter : 4000;
    return iters;
}

/* ---------------------------------------------------------------- merit / line search / sqp ------------- */
static double merit(const orc_model* m, const orc_cfg* c, const double* xu, const double* xref, const double* goals, const double* fext, double* comps) {
    const int N = c->N;
    double qc = 0, vc = 0, uc = 0, ie = 0;
    for (int k = 0; k < N; ++k) {
        const double* xk = xu + NXU * k;
        double pee[3], e2 = 0, vv = 0, uu = 0;
        fk_jac(m, xk, pee, NULL);
        for (int a = 0; a < 3; ++a) { double e = pee[a] - goals[3 * k + a]; e2 += e * e; }
        qc += ((k == N - 1) ? c->QN : 1.0) * e2;
        for (int i = 0; i < 6; ++i) vv += xk[6 + i] * xk[6 + i];
        vc += c->dQ * vv;
        if (k < N - 1) {
            for (int i = 0; i < 6; ++i) uu += xk[12 + i] * xk[12 + i];
            uc += c->R * uu;
            double qdd[6], qe = 0, ve = 0;
            aba(m, xk, xk + 6, xk + 12, fext, qdd, NULL);
            for (int i = 0; i < 6; ++i) {
                double dq = xk[i] + xk[6 + i] * c->dt - xk[NXU + i], dv = xk[6 + i] + qdd[i] * c->dt - xk[NXU + 6 + i];
                qe += dq * dq; ve += dv * dv;
            }
            ie += sqrt(qe) + sqrt(ve);
        }
    }
    double d2 = 0;
    for (int i = 0; i < NX; ++i) { double d = xu[i] - xref[i]; d2 += d * d; }
    if (comps) { comps[0] = qc; comps[1] = vc; comps[2] = uc; comps[3] = ie; comps[4] = sqrt(d2); }
    return qc + vc + uc + c->mu * (ie + sqrt(d2));
}

static double linesearch(const orc_model* m, const orc_cfg* c, const double* xu, const double* z, const double* goals, const double* fext, double* tmp) {
    const int n = NXU * c->N - 6;
    const double base = merit(m, c, xu, xu, goals, fext, NULL);
    double alpha = 1.0;
    for (int t = 0; t < 8; ++t, alpha *= 0.5) {
        for (int i = 0; i < n; ++i) tmp[i] = xu[i] + alpha * (z[i] - xu[i]);
        if (merit(m, c, tmp, xu, goals, fext, NULL) <= base) return alpha;
    }
    return 0.0;
}

/* ---------------------------------------------------------------- public C entry points ----------------- */
orc_work* orc_work_new(int N) {
    orc_work* w = (orc_work*)calloc(1, sizeof(orc_work));
    w->N = N; w->n = NXU * N - 6; w->mrows = NX * N; w->nk = w->n + w->mrows; w->bw = BW;
    w->Gq = (double*)calloc((size_t)36 * N, sizeof(double)); w->Gv = (double*)calloc((size_t)36 * N, sizeof(double));
    w->H = (double*)calloc((size_t)36 * N, sizeof(double)); w->cv = (double*)calloc((size_t)6 * N, sizeof(double));
    w->j = (double*)calloc((size_t)6 * N, sizeof(double)); w->Qk = (double*)calloc(N, sizeof(double));
    w->dQk = (double*)calloc(N, sizeof(double)); w->Rk = (double*)calloc(N, sizeof(double));
    w->x = (double*)calloc(w->n, sizeof(double)); w->y = (double*)calloc(w->mrows, sizeof(double));
    w->yl = (double*)calloc(w->n, sizeof(double)); w->zl = (double*)calloc(w->n, sizeof(double));
    w->band = (double*)calloc((size_t)w->nk * (BW + 1), sizeof(double));
    return w;
}
void orc_work_free(orc_work* w) {
    if (!w) return;
    free(w->Gq); free(w->Gv); free(w->H); free(w->cv); free(w->j); free(w->Qk); free(w->dQk); free(w->Rk);
    free(w->x); free(w->y); free(w->yl); free(w->zl); free(w->band); free(w);
}
void orc_work_reset(orc_work* w) {
    memset(w->x, 0, sizeof(double) * w->n); memset(w->y, 0, sizeof(double) * w->mrows);
    memset(w->yl, 0, sizeof(double) * w->n); memset(w->zl, 0, sizeof(double) * w->n);
}
/* ADMM state of a work area (parity tests): x (n), y (12N equality rows), yl (n - 12 limit rows), residuals of the last check */
void orc_work_state(const orc_work* w, double* x, double* y, double* yl, double* res2) {
    if (x) memcpy(x, w->x, sizeof(double) * w->n);
    if (y) memcpy(y, w->y, sizeof(double) * w->mrows);
    if (yl) memcpy(yl, w->yl, sizeof(double) * (w->n - NX));
    if (res2) { res2[0] = w->rp; res2[1] = w->rd; }
}

static void pin_to_featherstone(const double* f, double* o) { for (int a = 0; a < 3; ++a) { o[a] = f[3 + a]; o[3 + a] = f[a]; } }

/* z = setup_and_solve_qp(xu, xs, goals).x ; returns QP iterations */
int orc_solve_qp(const orc_model* m, const orc_cfg* c, orc_work* w, const double* xu, const double* xs, const double* goals,
                 const double* fext_pin, double* z) {
    double fx[6];
    if (fext_pin) pin_to_featherstone(fext_pin, fx);
    double* pad = (double*)calloc((size_t)NXU * c->N, sizeof(double));
    memcpy(pad, xu, sizeof(double) * (NXU * c->N - 6));
    assemble(m, c, pad, goals, fext_pin ? fx : NULL, w);
    int it = solve_qp(c, w, pad, xs, z);
    free(pad);
    return it;
}

/* xu_out = sqp(xs, goals, xu_in); stats[12] laid out like the product's record ([8], [9]: QP-solver iterations of the
 * two passes, [10] their sum); returns total QP iterations */
int orc_sqp(const orc_model* m, const orc_cfg* c, orc_work* w, const double* xs, const double* goals, const double* xu_in,
            const double* fext_pin, double* xu_out, double* stats) {
    const int N = c->N, n = NXU * N - 6;
    double fx[6];
    if (fext_pin) pin_to_featherstone(fext_pin, fx);
    const double* fext = fext_pin ? fx : NULL;
    double* xu = (double*)calloc((size_t)3 * NXU * N, sizeof(double));
    double *z = xu + NXU * N, *tmp = z + NXU * N;
    memcpy(xu, xu_in, sizeof(double) * n);
    int qp, n_ls = 0, n_steps = 0, status = 0, total_it = 0;
    double alphas[2] = {0, 0}, steps[2] = {0, 0}, its[2] = {0, 0};
    for (qp = 0; qp < c->max_qp; ++qp) {
        assemble(m, c, xu, goals, fext, w);
        const int it = solve_qp(c, w, xu, xs, z);
        if (qp < 2) its[qp] = it;
        total_it += it;
        double alpha = linesearch(m, c, xu, z, goals, fext, tmp);
        if (n_ls < 2) alphas[n_ls] = alpha;
        ++n_ls;
        if (alpha == 0.0) { status = 1; continue; }
        double s2 = 0;
        for (int i = 0; i < n; ++i) { double st = alpha * (z[i] - xu[i]); xu[i] += st; s2 += st * st; }
        if (n_steps < 2) steps[n_steps] = sqrt(s2);
        ++n_steps;
        if (sqrt(s2) < c->step_tol) { ++qp; break; }
    }
    if (qp > c->max_qp) qp = c->max_qp;
    memcpy(xu_out, xu, sizeof(double) * n);
    if (stats) {
        stats[0] = qp; stats[1] = n_ls; stats[2] = n_steps; stats[3] = status;
        stats[4] = alphas[0]; stats[5] = alphas[1]; stats[6] = steps[0]; stats[7] = steps[1];
        stats[8] = its[0]; stats[9] = its[1]; stats[10] = total_it; stats[11] = 0;
    }
    free(xu);
    return total_it;
}

/* batch of independent problems, OpenMP across problems; works: one per problem (warm-start state) */
int orc_sqp_batch(const orc_model* m, const orc_cfg* c, orc_work** works, int batch, const double* xs, const double* goals,
                  const double* xu_in, const double* fext_pin, double* xu_out, double* stats, int nthreads) {
    const int n = NXU * c->N - 6;
    long total = 0;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : total)
    for (int b = 0; b < batch; ++b)
        total += orc_sqp(m, c, works[b], xs + (size_t)NX * b, goals + (size_t)3 * c->N * b, xu_in + (size_t)n * b,
                         fext_pin ? fext_pin + 6 * (size_t)b : NULL, xu_out + (size_t)n * b, stats ? stats + 12 * (size_t)b : NULL);
    return (int)total;
}

void orc_dynamics(const orc_model* m, const double* q, const double* v, const double* tau, const double* fext_pin, double* qdd,
                  double* dq, double* dv, double* Mi) {
    double fx[6], rq[36], rv[36];
    rnea_aux aux;
    if (fext_pin) pin_to_featherstone(fext_pin, fx);
    aba(m, q, v, tau, fext_pin ? fx : NULL, qdd, &aux);
    minv(m, &aux, Mi);
    rnea_derivs(m, v, qdd, &aux, rq, rv);
    for (int r = 0; r < 6; ++r) for (int c = 0; c < 6; ++c) {
        double sq = 0, sv = 0;
        for (int k = 0; k < 6; ++k) { sq -= Mi[6 * r + k] * rq[6 * k + c]; sv -= Mi[6 * r + k] * rv[6 * k + c]; }
        dq[6 * r + c] = sq; dv[6 * r + c] = sv;
    }
}

void orc_eepos(const orc_model* m, const double* q, double* p, double* J) { fk_jac(m, q, p, J); }

void orc_rk4(const orc_model* m, const double* q, const double* v, const double* u, double dt, const double* fext_pin, double* qn, double* vn) {
    double fx[6], k1v[6], k2v[6], k3v[6], k4v[6], k2q[6], k3q[6], k4q[6], qt[6];
    if (fext_pin) pin_to_featherstone(fext_pin, fx);
    const double* f = fext_pin ? fx : NULL;
    aba(m, q, v, u, f, k1v, NULL);
    for (int a = 0; a < 6; ++a) { qt[a] = q[a] + v[a] * dt / 2; k2q[a] = v[a] + k1v[a] * dt / 2; }
    aba(m, qt, k2q, u, f, k2v, NULL);
    for (int a = 0; a < 6; ++a) { qt[a] = q[a] + k2q[a] * dt / 2; k3q[a] = v[a] + k2v[a] * dt / 2; }
    aba(m, qt, k3q, u, f, k3v, NULL);
    for (int a = 0; a < 6; ++a) { qt[a] = q[a] + k3q[a] * dt; k4q[a] = v[a] + k3v[a] * dt; }
    aba(m, qt, k4q, u, f, k4v, NULL);
    for (int a = 0; a < 6; ++a) {
        vn[a] = v[a] + (dt / 6) * (k1v[a] + 2 * k2v[a] + 2 * k3v[a] + k4v[a]);
        qn[a] = q[a] + (v[a] + 2 * k2q[a] + 2 * k3q[a] + k4q[a]) / 6 * dt;
    }
}

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
