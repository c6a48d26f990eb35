"""Receding-horizon driver with the reference's interface (``MPC_OSQP``, src/osqp_mpc.py:5-71) on the CUDA
backend: ``sqp`` is one fused kernel, the plant step is ``i7_sim_forward`` (rk4, zero external wrench)."""
import numpy as np

from .utils import rk4


class MPC_OSQP:
    def __init__(self, model, sqp_optimizer, solver):
        self.model = model
        self.model.gravity.linear = np.array([0.0, 0.0, -9.81])
        self.sqp_optimizer = sqp_optimizer
        self.solver = solver
        self.xpath = []
        self.goal_distances = []

    def run_mpc(self, xstart, endpoints, num_steps=500, verbose=False):
        s = self.solver
        nq, nx, nu = s.nq, s.nx, s.nu
        stride = nx + nu
        xcur = np.array(xstart, float)
        which = 0
        goal = np.tile(endpoints[which], s.N)
        XU = self.sqp_optimizer.sqp(xcur, goal, np.zeros(s.N * stride - nu))
        for _ in range(num_steps):
            dist = float(np.linalg.norm(s.eepos(xcur[:nq]) - goal[:3]))
            if dist < 1e-1:
                if verbose:
                    print('switching goals')
                which = (which + 1) % len(endpoints)
                goal = np.tile(endpoints[which], s.N)
            if verbose:
                print(dist)
            self.goal_distances.append(dist)
            if dist > 1.1:
                if verbose:
                    print("breaking on big goal dist")
                break
            xu_new = self.sqp_optimizer.sqp(xcur, goal, XU)
            remaining, taken = 0.01, 0
            while remaining > 0:
                h = min(remaining, s.dt)
                control = XU[taken * stride + nx:(taken + 1) * stride]     # control of the OLD trajectory (osqp_mpc.py:55)
                qn, vn = rk4(self.model, s.data, xcur[:nq], xcur[nq:nx], control, h)
                xcur = np.concatenate([qn, vn])
                if h > 0.5 * s.dt:
                    taken += 1
                remaining -= h
                self.xpath.append(xcur[:nq])
            if taken > 0:
                XU[:-(taken * stride) or len(XU)] = xu_new[taken * stride:]
            XU[:nx] = xcur
            XU[-nx:] = np.concatenate([np.ones(nq), np.zeros(nq)])
        return self.xpath
