"""``OSQPSolver`` with the reference's constructor, attributes and methods (src/osqp_solver.py:6-155), backed by
the sm_100a kernels through the C ABI.  No Pinocchio, no OSQP, no CPU fallback.

By default (``qp_mode='exact'``) the QP of one SQP iteration is solved exactly on the GPU (block recursion over the
knots), so the object returned by ``setup_and_solve_qp`` carries the solution OSQP converges to and ``.info.status`` is
always 'solved'.  ``qp_mode='osqp'`` runs the reference's own regime instead — OSQP's ADMM at eps_abs = eps_rel = 1e-3,
warm-started between calls like the OSQP object of src/osqp_solver.py:39-41 — and is what ``limits=True`` (URDF joint
position / velocity / effort limits as inequality rows; not in the reference QP) needs.
"""
import numpy as np

from .engine import Engine
from .model import RobotModel


class QPResult:
    def __init__(self, x, y=None, iters=1, status="solved", pri_res=0.0, dua_res=0.0):
        self.x = x
        self.y = y
        self.info = type("info", (), dict(status=status, iter=iters, pri_res=pri_res, dua_res=dua_res))()


class OSQPSolver:
    def __init__(self, model, dt=0.01, N=32, dQ_cost=0.01, R_cost=1e-5, QN_cost=100, regularize=True, eps=1,
                 dtype="fp64", device=0, qp_mode="exact", limits=False, admm=None):
        if not isinstance(model, RobotModel):
            raise TypeError("model must come from indy7_mpc_b200.load_robot_model")
        self.model = model
        self.data = model.createData()
        self.N, self.dt = N, dt
        self.nq, self.nv = model.nq, model.nv
        self.nx = self.nq + self.nv
        self.nu = len(model.joints) - 1
        self.nxu = self.nx + self.nu
        self.traj_len = self.nxu * self.N - self.nu
        self.dQ_cost, self.R_cost, self.QN_cost = dQ_cost, R_cost, QN_cost
        self.regularize, self.eps = regularize, eps
        self.engine = Engine(model, N=N, dt=dt, dQ_cost=dQ_cost, R_cost=R_cost, QN_cost=QN_cost,
                             regularize=regularize, eps=eps, batch=1, dtype=dtype, device=device, qp_mode=qp_mode,
                             limits=limits, admm=admm)
        self.qp_mode, self.limits = qp_mode, bool(limits)
        self.l = np.zeros(self.N * self.nx)
        self.g = np.zeros(self.traj_len)
        self.Pdata = np.zeros(27 * N + 6 * (N - 1))
        self.Adata = np.zeros(360 * (N - 1) + 144)

    def _sync_weights(self):
        self.engine.set_weights(self.dQ_cost, self.R_cost, self.QN_cost, self.regularize, self.eps)

    def _check(self, xu, xs, eepos_g):
        xu, xs, eepos_g = (np.asarray(a, float).reshape(-1) for a in (xu, xs, eepos_g))
        if xu.size != self.traj_len or xs.size != self.nx or eepos_g.size != 3 * self.N:
            raise ValueError(f"expected xu ({self.traj_len},), xs ({self.nx},), eepos_g ({3 * self.N},)")
        return xu, xs, eepos_g

    def update_constraint_matrix(self, xu, xs):
        """Fills ``Adata`` and ``l`` in the reference's CSC value order (src/osqp_solver.py:83-101)."""
        self._sync_weights()
        A, _, _, l = self.engine.dump_qp(xu, xs, np.zeros(3 * self.N))
        self.Adata[:], self.l[:] = A[0], l[0]

    def update_cost_matrix(self, XU, eepos_g):
        """Fills ``Pdata`` and ``g`` (src/osqp_solver.py:103-135)."""
        self._sync_weights()
        _, Pd, g, _ = self.engine.dump_qp(XU, np.asarray(XU, float).reshape(-1)[:self.nx], eepos_g)
        self.Pdata[:], self.g[:] = Pd[0], g[0]

    def setup_and_solve_qp(self, xu, xs, eepos_g):
        xu, xs, eepos_g = self._check(xu, xs, eepos_g)
        self._sync_weights()
        z, y = self.engine.solve_qp(xu, xs, eepos_g, want_duals=True)
        if self.qp_mode == "osqp":
            _, _, yl, info = self.engine.qp_state(1)
            yy = np.concatenate([np.asarray(y[0], float), yl[0]]) if self.limits else np.asarray(y[0], float)
            return QPResult(np.asarray(z[0], float), yy, int(info[0, 0]),
                            "solved" if info[0, 3] == 0 else "maximum iterations reached", float(info[0, 1]), float(info[0, 2]))
        return QPResult(np.asarray(z[0], float), np.asarray(y[0], float))

    def eepos(self, q):
        return np.asarray(self.engine.eepos(np.asarray(q, float).reshape(1, 6))[0], float)

    def d_eepos(self, q):
        p, J = self.engine.eepos(np.asarray(q, float).reshape(1, 6), jacobian=True)
        return np.asarray(p[0], float), np.asarray(J[0], float)
