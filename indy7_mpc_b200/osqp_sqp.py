"""``SQP_OSQP`` with the reference's interface (src/osqp_sqp.py:4-96); one fused kernel launch per ``sqp`` call."""
import numpy as np


class SQP_OSQP:
    def __init__(self, solver, stats=None):
        self.solver = solver
        self.stats = stats or {
            'qp_iters': {'values': [], 'unit': '', 'multiplier': 1},
            'linesearch_alphas': {'values': [], 'unit': '', 'multiplier': 1},
            'sqp_stepsizes': {'values': [], 'unit': '', 'multiplier': 1},
        }
        self.last_record = None

    def eepos_cost(self, eepos_goals, XU):
        self.solver._sync_weights()
        m = self.solver.engine.merit(XU, XU, eepos_goals)[0]
        return float(m[0]), float(m[1]), float(m[2])

    def integrator_err(self, XU):
        self.solver._sync_weights()
        return float(self.solver.engine.merit(XU, XU, np.zeros(3 * self.solver.N))[0][3])

    def linesearch(self, XU, XU_fullstep, eepos_goals):
        self.solver._sync_weights()
        alpha, merits = self.solver.engine.linesearch(XU, XU_fullstep, eepos_goals)
        self.last_merits = np.asarray(merits[0], float)
        self.stats['linesearch_alphas']['values'].append(float(alpha[0]))
        return float(alpha[0])

    def sqp(self, xcur, eepos_goals, XU):
        s = self.solver
        XU = np.asarray(XU, float).reshape(-1)
        xcur = np.asarray(xcur, float).reshape(-1)
        eepos_goals = np.asarray(eepos_goals, float).reshape(-1)
        if XU.size != s.traj_len or xcur.size != s.nx or eepos_goals.size != 3 * s.N:
            raise ValueError(f"expected XU ({s.traj_len},), xcur ({s.nx},), eepos_goals ({3 * s.N},)")
        s._sync_weights()
        out, st = s.engine.solve(XU, xcur, eepos_goals)
        rec = np.asarray(st[0], float)
        self.last_record = rec
        n_ls, n_steps = int(rec[1]), int(rec[2])
        self.stats['linesearch_alphas']['values'].extend(float(a) for a in rec[4:4 + n_ls])
        self.stats['sqp_stepsizes']['values'].extend(float(a) for a in rec[6:6 + n_steps])
        self.stats['qp_iters']['values'].append(int(rec[0]))
        # status 2 = a non-finite merit was met: every comparison of the line search is then false, alpha = 0 and XU
        # comes back unchanged — what the reference loop does as well (src/osqp_sqp.py:66-72,81-82); the status stays
        # readable in `last_record[3]`
        return np.asarray(out[0], float)

    def get_stats(self):
        return self.stats
