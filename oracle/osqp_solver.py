"""CPU restatement of the reference QP builder/solver ``OSQPSolver`` (src/osqp_solver.py:6-155).
TEST INFRASTRUCTURE — never imported by the product package.

Same constructor, attributes and methods as the reference class.  Pinocchio calls are replaced by
``oracle.rbd``; the OSQP call is replaced by one of two back-ends chosen with ``qp_backend``:

  'kkt'   exact solution of the equality-constrained QP by a sparse LU of the KKT system
          [[P, A'], [A, 0]] [z; y] = [-g; l]  (what OSQP converges to; the parity oracle);
  'admm'  a restatement of OSQP's published ADMM (Stellato et al. 2020, Alg. 1) with the defaults the
          reference relies on (src/osqp_solver.py:39-41 passes only verbose=False): rho=0.1 (x1e3 on
          equality rows), sigma=1e-6, alpha=1.6, eps_abs=eps_rel=1e-3, max_iter=4000,
          check_termination=25, Ruiz scaling (10 passes), warm start across calls, no polish.
          OSQP's wall-clock-driven adaptive-rho schedule is not reproducible and is left out
          ("QP-iteration parity unpinned").  ``scaling=0`` gives the unscaled iteration
          that oracle/c/oracle.c and the product's "osqp" QP mode run.

Extension (not reference behaviour: src/osqp_solver.py:41 passes l == u): with
``limits=(qmin, qmax, vmax, umax)`` (description/indy7.urdf:203-238) the QP gets the box rows
lo <= z_i <= hi on every q_k, v_k (k >= 1) and u_k (identity rows below the 12N dynamics rows; rho = rho_x on the
state rows, rho_u on the effort rows — see oracle/c/oracle.c);
ADMM back-end only.

The CSC value orders of ``Adata`` / ``Pdata`` follow the reference exactly so the
product's debug dump (``i7_dump_qp``) can be compared entry by entry.
"""
import numpy as np
from scipy.sparse import bmat, csc_matrix, triu, eye as speye, diags
from scipy.sparse.linalg import splu

from . import rbd


class QPResult:
    def __init__(self, x, y, iters, status, r_prim=0.0, r_dual=0.0):
        self.x, self.y = x, y
        self.info = type("info", (), dict(iter=iters, status=status, pri_res=r_prim, dua_res=r_dual))()


class OSQPSolver:
    def __init__(self, model, dt=0.01, N=32, dQ_cost=0.01, R_cost=1e-5, QN_cost=100, regularize=True, eps=1,
                 qp_backend='kkt', f_ext=None, limits=None, admm_scaling=10, rho_x=0.1, rho_u=1e-3, max_iter=4000):
        self.model = model
        self.data = model.createData()
        self.N, self.dt = N, dt
        self.nq, self.nv = model.nq, model.nv
        self.nx = self.nq + self.nv
        self.nu = len(model.joints) - 1                      # src/osqp_solver.py:20
        self.nxu = self.nx + self.nu
        self.traj_len = self.nxu * N - self.nu               # src/osqp_solver.py:22
        self.dQ_cost, self.R_cost, self.QN_cost = dQ_cost, R_cost, QN_cost
        self.regularize, self.eps = regularize, eps
        self.qp_backend = qp_backend
        self.f_ext = f_ext                                    # extension: local wrench on the last link
        self.limits = None if limits is None else tuple(np.asarray(v, float) for v in limits)
        if self.limits is not None and qp_backend != 'admm':
            raise ValueError("limits need qp_backend='admm' (the exact KKT back-end is equality-only)")
        self.admm_scaling, self.rho_x, self.rho_u, self.max_iter = admm_scaling, rho_x, rho_u, max_iter
        self.A = self.initialize_A()
        self.P = self.initialize_P()
        self.l = np.zeros(N * self.nx)
        self.g = np.zeros(self.traj_len)
        self.Pdata = np.zeros(self.P.nnz)
        self.Adata = np.zeros(self.A.nnz)
        self._admm = None

    # ---- sparsity templates (src/osqp_solver.py:48-68) -------------------------------------------------
    def initialize_P(self):
        blk = np.eye(self.nxu)
        blk[:self.nq, :self.nq] = 1.0
        bd = np.kron(np.eye(self.N), blk)[:-self.nu, :-self.nu]
        return csc_matrix(triu(csc_matrix(bd)), shape=(self.traj_len, self.traj_len))

    def initialize_A(self):
        nx, nu, N = self.nx, self.nu, self.N
        rows = [[np.ones((nx, nx))] + [None] * (2 * N)]
        for i in range(N - 1):
            row = [None] * (2 * N + 1)
            row[2 * i], row[2 * i + 1], row[2 * i + 2] = np.ones((nx, nx)), 2 * np.ones((nx, nu)), -np.ones((nx, nx))
            rows.append(row)
        # the reference pads every row to 2N+1 block columns; the last (all-None) block column has no width
        rows = [r[:2 * N - 1] for r in rows]
        return bmat(rows, format='csc')

    # ---- per-knot quantities, vectorised over the horizon ------------------------------------------------
    def _split(self, xu):
        xu = np.asarray(xu, float)
        full = np.concatenate([xu, np.zeros(self.nu)]).reshape(self.N, self.nxu)
        return full[:, :self.nq], full[:, self.nq:self.nx], full[:-1, self.nx:]

    def linearize(self, xu):
        """A_k (N-1,12,12), B_k (N-1,12,6), c_k (N-1,12) of src/osqp_solver.py:70-81."""
        q, v, u = self._split(xu)
        q, v = q[:-1], v[:-1]
        dq, dv, du, a = rbd.aba_derivatives(self.model, q, v, u, f_ext=self.f_ext)
        n, dt = self.nq, self.dt
        Ak = np.zeros((self.N - 1, self.nx, self.nx))
        Ak[:, :n, :n] = np.eye(n); Ak[:, :n, n:] = dt * np.eye(n)
        Ak[:, n:, :n] = dq * dt; Ak[:, n:, n:] = dv * dt + np.eye(n)
        Bk = np.zeros((self.N - 1, self.nx, self.nu))
        Bk[:, n:, :] = du * dt
        xnext = np.concatenate([q + v * dt, v + a * dt], axis=1)
        xcur = np.concatenate([q, v], axis=1)
        ck = xnext - np.einsum('kij,kj->ki', Ak, xcur) - np.einsum('kij,kj->ki', Bk, u)
        return Ak, Bk, ck

    def update_constraint_matrix(self, xu, xs):
        Ak, Bk, ck = self.linearize(xu)
        nx = self.nx
        self.l[:nx] = -np.asarray(xs, float)
        stack = np.concatenate([np.broadcast_to(-np.eye(nx), Ak.shape), Ak], axis=1)       # (N-1,24,12)
        a_part = np.swapaxes(stack, 1, 2).reshape(self.N - 1, -1)                           # column-major
        b_part = np.swapaxes(Bk, 1, 2).reshape(self.N - 1, -1)
        body = np.concatenate([a_part, b_part], axis=1).reshape(-1)
        self.Adata[:body.size] = body
        self.Adata[body.size:] = -np.eye(nx).reshape(-1)
        self.l[nx:] = -ck.reshape(-1)

    def cost_terms(self, XU, eepos_g):
        """Per-knot (j_k (N,6), Q_k, dQ_k, R_k (N,)) of src/osqp_solver.py:110-120."""
        q, v, u = self._split(XU)
        p, J = rbd.d_eepos(self.model, q)
        err = p - np.asarray(eepos_g, float).reshape(self.N, 3)
        scale = 1.0 / (np.linalg.norm(err, axis=1) + self.eps) if self.regularize else np.ones(self.N)
        Qk = np.ones(self.N); Qk[-1] = self.QN_cost
        return np.einsum('ki,kij->kj', err, J), Qk, self.dQ_cost * scale, self.R_cost * scale

    def update_cost_matrix(self, XU, eepos_g):
        q, v, u = self._split(XU)
        j, Qk, dQk, Rk = self.cost_terms(XU, eepos_g)
        il = np.tril_indices(self.nq)
        pind = 0
        for k in range(self.N):
            s = k * self.nxu
            self.g[s:s + self.nq] = Qk[k] * j[k]
            self.g[s + self.nq:s + self.nx] = dQk[k] * v[k]
            pos = Qk[k] * np.outer(j[k], j[k])[il]
            self.Pdata[pind:pind + pos.size] = pos; pind += pos.size
            self.Pdata[pind:pind + self.nv] = dQk[k]; pind += self.nv
            if k < self.N - 1:
                self.Pdata[pind:pind + self.nu] = Rk[k]; pind += self.nu
                self.g[s + self.nx:s + self.nxu] = Rk[k] * u[k]

    # ---- QP solve (src/osqp_solver.py:137-143) -----------------------------------------------------------
    def matrices(self):
        P = csc_matrix((self.Pdata, self.P.indices, self.P.indptr), shape=self.P.shape)
        A = csc_matrix((self.Adata, self.A.indices, self.A.indptr), shape=self.A.shape)
        return P, A

    def setup_and_solve_qp(self, xu, xs, eepos_g):
        self.update_constraint_matrix(xu, xs)
        self.update_cost_matrix(xu, eepos_g)
        Pu, A = self.matrices()
        P = Pu + triu(Pu, 1).T
        if self.qp_backend == 'kkt':
            return solve_kkt(P, self.g, A, self.l)
        lo = hi = self.l
        if self.limits is not None:                          # box rows on z[12:]
            qmin, qmax, vmax, umax = self.limits
            blo = np.tile(np.concatenate([qmin, -vmax, -umax]), self.N)[:self.traj_len][self.nx:]
            bhi = np.tile(np.concatenate([qmax, vmax, umax]), self.N)[:self.traj_len][self.nx:]
            sel = speye(self.traj_len, format='csc')[self.nx:, :]
            A = bmat([[A], [sel]], format='csc')
            lo, hi = np.concatenate([self.l, blo]), np.concatenate([self.l, bhi])
            rho_rows = np.tile(np.concatenate([np.full(self.nx, self.rho_x), np.full(self.nu, self.rho_u)]), self.N)[:self.traj_len][self.nx:]
        else:
            rho_rows = np.zeros(0)
        if self._admm is None:
            self._admm = OSQPADMM(self.traj_len, A.shape[0], scaling=self.admm_scaling, max_iter=self.max_iter)
        return self._admm.solve(P, self.g, A, lo, hi, rho_ineq=rho_rows)

    # ---- end effector (src/osqp_solver.py:146-155) -------------------------------------------------------
    def eepos(self, q):
        return rbd.eepos(self.model, np.asarray(q, float))

    def d_eepos(self, q):
        return rbd.d_eepos(self.model, np.asarray(q, float))


def solve_kkt(P, g, A, l):
    n, m = P.shape[0], A.shape[0]
    K = bmat([[P, A.T], [A, None]], format='csc')
    sol = splu(K).solve(np.concatenate([-g, l]))
    # one step of iterative refinement keeps the answer at rounding level (KKT cond ~ 2e6)
    r = np.concatenate([-g, l]) - K @ sol
    sol = sol + splu(K).solve(r)
    return QPResult(sol[:n], sol[n:], 1, 'solved')


class OSQPADMM:
    """OSQP's ADMM for l <= Ax <= u, default settings, warm-started between calls.  rho per row: rho on inequality
    rows, rho x 1e3 on equality rows (l == u), as OSQP sets it."""

    def __init__(self, n, m, rho=0.1, sigma=1e-6, alpha=1.6, eps_abs=1e-3, eps_rel=1e-3, max_iter=4000,
                 check_termination=25, scaling=10, rho_eq_factor=1e3):
        self.n, self.m = n, m
        self.rho, self.rho_eq_factor = rho, rho_eq_factor     # RHO_EQ_OVER_RHO_INEQ = 1e3
        self.sigma, self.alpha = sigma, alpha
        self.eps_abs, self.eps_rel = eps_abs, eps_rel
        self.max_iter, self.check = max_iter, check_termination
        self.scaling = scaling
        self.x = np.zeros(n); self.y = np.zeros(m)

    def _ruiz(self, P, A, q):
        n, m = self.n, self.m
        D, E = np.ones(n), np.ones(m)
        c = 1.0
        for _ in range(self.scaling):
            Ps = diags(D) @ P @ diags(D)
            As = diags(E) @ A @ diags(D)
            cn = np.maximum(abs(Ps).max(axis=0).toarray().ravel(), abs(As).max(axis=0).toarray().ravel())
            rn = abs(As).max(axis=1).toarray().ravel()
            cn[cn < 1e-4] = 1.0; rn[rn < 1e-4] = 1.0
            D = D / np.sqrt(np.minimum(cn, 1e4)); E = E / np.sqrt(np.minimum(rn, 1e4))
            Ps = diags(D) @ P @ diags(D)
            qs = D * q
            pn = abs(Ps).max(axis=0).toarray().ravel().mean()
            qn = np.abs(qs).max()
            cost = max(pn, qn if qn > 1e-4 else 1.0)
            cost = min(max(cost, 1e-4), 1e4)
            c = 1.0 / cost  # OSQP accumulates c across passes on the already-scaled data; one-shot here
        return D, E, c

    def solve(self, P, q, A, l, u=None, rho_ineq=None):
        n, m = self.n, self.m
        u = l if u is None else u
        if self.scaling:
            D, E, c = self._ruiz(P, A, q)
        else:
            D, E, c = np.ones(n), np.ones(m), 1.0
        Ps = (diags(D) @ P @ diags(D) * c).tocsc()
        As = (diags(E) @ A @ diags(D)).tocsc()
        qs, ls, us = c * D * q, E * l, E * u
        rho = np.where(l == u, self.rho * self.rho_eq_factor, self.rho)
        if rho_ineq is not None and len(rho_ineq):               # per-row rho of the trailing inequality rows
            rho[-len(rho_ineq):] = rho_ineq
        K = bmat([[Ps + self.sigma * speye(n), As.T], [As, -diags(1.0 / rho)]], format='csc')
        lu = splu(K)
        x, y = self.x / D, self.y * c / E      # warm start in scaled variables
        z = As @ x                              # OSQP warm start: z = A x
        al = self.alpha
        it, status = 0, 'maximum iterations reached'
        rp = rd = np.inf
        while it < self.max_iter:
            it += 1
            sol = lu.solve(np.concatenate([self.sigma * x - qs, z - y / rho]))
            xt, nu = sol[:n], sol[n:]
            zt = z + (nu - y) / rho
            x = al * xt + (1 - al) * x
            zr = al * zt + (1 - al) * z
            znew = np.clip(zr + y / rho, ls, us)       # projection onto [l, u]
            y = y + rho * (zr - znew)
            z = znew
            if it % self.check == 0:
                xu_, yu_, zu_ = D * x, E * y / c, z / E
                Ax, Px, Aty = A @ xu_, P @ xu_, A.T @ yu_
                rp = np.abs(Ax - zu_).max()
                rd = np.abs(Px + q + Aty).max()
                ep = self.eps_abs + self.eps_rel * max(np.abs(Ax).max(), np.abs(zu_).max())
                ed = self.eps_abs + self.eps_rel * max(np.abs(Px).max(), np.abs(Aty).max(), np.abs(q).max())
                if rp <= ep and rd <= ed:
                    status = 'solved'
                    break
        self.x, self.y = D * x, E * y / c
        return QPResult(self.x.copy(), self.y.copy(), it, status, rp, rd)
