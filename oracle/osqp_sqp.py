"""CPU restatement of the reference SQP loop and merit line search ``SQP_OSQP`` (src/osqp_sqp.py:4-96).
TEST INFRASTRUCTURE — never imported by the product package.

Quirks reproduced literally: no 1/2 in the merit; un-regularised dQ/R weights in the
merit; integrator error sums un-squared 2-norms; the initial-state term compares with the previous iterate;
alpha == 0 -> ``continue`` (the identical QP is solved again); ``qp_iters`` records ``qp+1`` after the loop.
"""
import numpy as np

from . import rbd

ALPHAS = np.array([1.0, 0.5, 0.25, 0.125, 0.0625, 0.03125, 0.015625, 0.0078125])   # src/osqp_sqp.py:58


class SQP_OSQP:
    def __init__(self, solver, stats=None):
        self.solver = solver
        self.stats = stats or {
            'qp_iters': {'values': [], 'unit': '', 'multiplier': 1},
            'linesearch_alphas': {'values': [], 'unit': '', 'multiplier': 1},
            'sqp_stepsizes': {'values': [], 'unit': '', 'multiplier': 1},
        }
        self.last_merits = None   # (base, [merit per tried alpha]) of the latest line search; test aid

    def eepos_cost(self, eepos_goals, XU):                                     # src/osqp_sqp.py:13-30
        s = self.solver
        q, v, u = s._split(XU)
        err = rbd.eepos(s.model, q) - np.asarray(eepos_goals, float).reshape(s.N, 3)
        Qk = np.ones(s.N); Qk[-1] = s.QN_cost
        qcost = float(np.sum(Qk * np.sum(err * err, axis=1)))
        vcost = float(s.dQ_cost * np.sum(v * v))
        ucost = float(s.R_cost * np.sum(u * u))
        return qcost, vcost, ucost

    def integrator_err(self, XU):                                               # src/osqp_sqp.py:32-47
        s = self.solver
        q, v, u = s._split(XU)
        a = rbd.aba(s.model, q[:-1], v[:-1], u, f_ext=s.f_ext)
        qerr = q[:-1] + v[:-1] * s.dt - q[1:]
        verr = v[:-1] + a * s.dt - v[1:]
        return float(np.sum(np.linalg.norm(qerr, axis=1) + np.linalg.norm(verr, axis=1)))

    def merit(self, XU, XU_ref, eepos_goals, mu=10.0):
        nx = self.solver.nx
        qc, vc, uc = self.eepos_cost(eepos_goals, XU)
        cv = self.integrator_err(XU) + np.linalg.norm(XU[:nx] - XU_ref[:nx])
        return qc + vc + uc + mu * cv

    def linesearch(self, XU, XU_fullstep, eepos_goals):                         # src/osqp_sqp.py:49-74
        base = self.merit(XU, XU, eepos_goals)
        diff = XU_fullstep - XU
        tried = []
        alpha_acc = 0.0
        for alpha in ALPHAS:
            m = self.merit(XU + alpha * diff, XU, eepos_goals)
            tried.append(m)
            if m <= base:
                alpha_acc = float(alpha)
                break
        self.last_merits = (base, tried)
        self.stats['linesearch_alphas']['values'].append(alpha_acc)
        return alpha_acc

    def sqp(self, xcur, eepos_goals, XU):                                       # src/osqp_sqp.py:76-93
        XU = np.array(XU, float)
        for qp in range(2):
            sol = self.solver.setup_and_solve_qp(XU, xcur, eepos_goals)
            alpha = self.linesearch(XU, sol.x, eepos_goals)
            if alpha == 0.0:
                continue
            step = alpha * (sol.x - XU)
            XU = XU + step
            stepsize = float(np.linalg.norm(step))
            self.stats['sqp_stepsizes']['values'].append(stepsize)
            if stepsize < 1e-3:
                break
        self.stats['qp_iters']['values'].append(qp + 1)
        return XU

    def get_stats(self):
        return self.stats
