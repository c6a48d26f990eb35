"""CPU restatement of the reference receding-horizon driver ``MPC_OSQP.run_mpc`` (src/osqp_mpc.py:14-71)
with the plant integrator ``rk4`` (src/utils.py:3-18, zero external wrench — the stored notebook output
predates the 7th ``f_ext`` argument).  TEST INFRASTRUCTURE.

Quirks kept: the plant is driven with the control of the OLD trajectory (src/osqp_mpc.py:55); the last knot
is reset to [1]*6+[0]*6 after the shift (src/osqp_mpc.py:70); goals switch below 0.1 and the loop aborts
above 1.1 (src/osqp_mpc.py:34-43).
"""
import numpy as np

from . import rbd


class MPC_OSQP:
    def __init__(self, model, sqp_optimizer, solver):
        self.model = model
        self.model.gravity = np.array([0.0, 0.0, -9.81])
        self.sqp_optimizer, self.solver = sqp_optimizer, solver
        self.xpath = []
        self.goal_distances = []

    def run_mpc(self, xstart, endpoints, num_steps=500):
        s = self.solver
        nq, nx, nu = s.nq, s.nx, s.nu
        xcur = np.array(xstart, float)
        ei = 0
        goal = np.tile(endpoints[ei], s.N)
        XU = np.zeros(s.N * (nx + nu) - nu)
        XU = self.sqp_optimizer.sqp(xcur, goal, XU)
        for _ in range(num_steps):
            dist = float(np.linalg.norm(s.eepos(xcur[:nq]) - goal[:3]))
            if dist < 1e-1:
                ei = (ei + 1) % len(endpoints)
                goal = np.tile(endpoints[ei], s.N)
            self.goal_distances.append(dist)
            if dist > 1.1:
                break
            xu_new = self.sqp_optimizer.sqp(xcur, goal, XU)
            sim_time, sim_steps = 0.01, 0
            while sim_time > 0:
                h = min(sim_time, s.dt)
                control = XU[sim_steps * (nx + nu) + nx:(sim_steps + 1) * (nx + nu)]
                qn, vn = rbd.rk4(self.model, xcur[:nq], xcur[nq:nx], control, h)
                xcur = np.concatenate([qn, vn])
                if h > 0.5 * s.dt:
                    sim_steps += 1
                sim_time -= h
                self.xpath.append(xcur[:nq])
            if sim_steps > 0:
                XU[:-(sim_steps) * (nx + nu) or len(XU)] = xu_new[sim_steps * (nx + nu):]
            XU[:nx] = xcur
            XU[-nx:] = np.concatenate([np.ones(nq), np.zeros(nq)])
        return self.xpath
