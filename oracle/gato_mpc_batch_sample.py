"""TEST INFRASTRUCTURE (CPU oracle) — restatement of the reference's batched sampled-wrench MPC driver
src/gato_mpc_batch_sample.py:80-300 (``MPC_GATO_Batch_Sample.run_mpc`` / ``find_best_trajectory_id``) with the
per-lane Python loops of the reference kept as loops.

The external ``bindings.batch_sqp`` solver the reference calls (:13-29,:57; binary absent) is
stood in for by the C restatement of the osqp_sqp path (oracle/c, exact-KKT back-end) applied lane by lane with the
lane's wrench, its float32 output rounding included (``SQPSolverfloat``).  Parity of this file is pinned only through
the pieces it is made of (FK, ABA, rk4, sqp — see tests/test_oracle_golden.py); the reference stores no output of
this driver for the osqp back-end: "parity unpinned" for the closed loop as a whole.
"""
import numpy as np

from . import rbd
from .cport import COracle


class MPC_GATO_Batch_Sample:
    def __init__(self, model, N=32, dt=0.01, batch_size=4, constant_f_ext=None, resample_f_ext=False, f_ext_std=1.0,
                 f_ext_resample_std=2.0):
        self.model, self.N, self.dt, self.batch_size = model, N, dt, batch_size
        self.orc = COracle(model, N=N, dt=dt, backend="kkt", batch=batch_size)
        # src/gato_mpc_batch_sample.py:36-42
        if f_ext_std != 0.0:
            self.f_ext_batch = np.random.normal(0, f_ext_std, (batch_size, 6))
            self.f_ext_batch[:, 3:] = 0.0
            self.f_ext_batch[0] = 0.0
        else:
            self.f_ext_batch = np.zeros((batch_size, 6))
        self.resample_f_ext, self.f_ext_resample_std = resample_f_ext, f_ext_resample_std
        self.constant_f_ext = np.zeros(6)                       # :96-101 local wrench on the last joint, force only
        if constant_f_ext is not None:
            self.constant_f_ext[:3] = np.asarray(constant_f_ext, float)[:3]
        self.xpath, self.best_ids = [], []

    def _solve(self, xcur_batch, goals6, XU_batch):             # :53-66 (solver.solve -> "xu_trajectory", float32)
        B, N = self.batch_size, self.N
        goals3 = goals6.reshape(B, N, 6)[:, :, :3].reshape(B, 3 * N)
        out, _, _ = self.orc.sqp(xcur_batch, goals3, XU_batch, f_ext=self.f_ext_batch)
        return out.astype(np.float32).astype(float)

    def run_mpc(self, xstart, endpoints, sim_dt=0.001, sim_time=5):            # :106-252
        B, N, nq, nx, nu = self.batch_size, self.N, 6, 12, 6
        xcur = np.array(xstart, float)
        xcur_batch = np.tile(xcur, (B, 1))
        endpoint = np.hstack([endpoints[0], np.zeros(3)])
        goal_batch = np.tile(np.tile(endpoint, N), (B, 1))
        stats = {'goal_distances': [], 'control_inputs': []}
        XU = np.zeros(N * (nx + nu) - nu)
        XU[:nx] = xcur
        XU_batch = self._solve(xcur_batch, goal_batch, np.tile(XU, (B, 1)))
        XU_best = XU_batch[0, :].copy()
        goal_idx, t_acc = 0, 0.0
        for i in range(int(sim_time / sim_dt)):
            x_last, u_last = xcur, XU_best[nx:nx + nu].copy()
            rem = sim_dt
            q, v = xcur[:nq], xcur[nq:nx]
            while rem > 0:                                                     # :169-186
                h = min(rem, self.dt - t_acc)
                knot = min(int(t_acc / self.dt), N - 1)
                u = XU_best[knot * (nx + nu) + nx:(knot + 1) * (nx + nu)]
                q, v = rbd.rk4(self.model, q, v, u, h, self.constant_f_ext)
                t_acc += h
                rem -= h
                self.xpath.append(q)
                if abs(t_acc - self.dt) < 1e-10:
                    t_acc = 0.0
            xcur = np.hstack([q, v])
            for j in range(B):                                                 # :189-192 (no shift)
                xcur_batch[j, :] = xcur
                XU_batch[j, :nx] = xcur
                XU_batch[j, nx:] = XU_best[nx:]
            goaldist = np.linalg.norm(rbd.eepos(self.model, xcur[:nq]) - goal_batch[0, :3])
            if goaldist < 0.095 and goal_idx < len(endpoints) - 1:              # :209-218
                goal_idx += 1
                endpoint = np.hstack([endpoints[goal_idx], np.zeros(3)])
                goal_batch = np.tile(np.tile(endpoint, N), (B, 1))
            elif goaldist < 0.095 and goal_idx == len(endpoints) - 1:
                break
            XU_new = self._solve(xcur_batch, goal_batch, XU_batch)             # :223-229
            best_id = self.find_best_trajectory_id(x_last, u_last, xcur, sim_dt)
            self.best_ids.append(best_id)
            XU_best = XU_new[best_id, :].copy()
            if abs((i * sim_dt) % 0.05) < 1e-5:                                # :236-241
                stats['goal_distances'].append(float(round(goaldist, 5)))
                stats['control_inputs'].append(np.round(XU_best[nx:nx + nu].copy(), 5))
            new_force = np.random.normal(self.constant_f_ext[:3], 0.1)          # :243-249
            self.constant_f_ext = np.concatenate([new_force, np.zeros(3)])
        return self.xpath, stats

    def find_best_trajectory_id(self, x_last, u_last, xcur, dt):               # :254-300
        best_error, best_id, f_best = np.inf, None, np.zeros(6)
        p, Rw = rbd.forward_kinematics(self.model, x_last[:6])
        p6, R6 = p[-1], Rw[-1]
        for i in range(self.batch_size):
            f = self.f_ext_batch[i, :]
            # oMi[6].actInv(Force(linear=f[:3], angular=f[3:]))  (:268-277)
            local = np.concatenate([R6.T @ f[:3], R6.T @ (f[3:] - np.cross(p6, f[:3]))])
            q, v = rbd.rk4(self.model, x_last[:6], x_last[6:], u_last, dt, local)
            err = np.linalg.norm(np.hstack([q, v]) - xcur)
            if err < best_error:
                best_error, best_id, f_best = err, i, f.copy()
        if self.resample_f_ext:                                                # :288-296
            self.f_ext_batch = f_best + np.random.normal(0, self.f_ext_resample_std, self.f_ext_batch.shape)
            self.f_ext_batch[:, 3:] = 0.0
            self.f_ext_batch[0] = 0.0
        return best_id

    # ---- notebooks/gato_mpc_indy7_fig8.ipynb cell 2: run_mpc_fig8 / find_best_trajectory_id (sim_forward flavour) ----
    def run_mpc_fig8(self, xstart, fig8_traj, sim_dt=0.001, sim_time=5):
        B, N, nq, nx, nu = self.batch_size, self.N, 6, 12, 6
        fig8_traj = np.asarray(fig8_traj, float)
        xcur = np.array(xstart, float)
        xcur_batch = np.tile(xcur, (B, 1))
        eepos_offset = 0
        eepos_goal = fig8_traj[:6 * N]
        goal_batch = np.tile(eepos_goal, (B, 1))
        stats = {'goal_distances': [], 'control_inputs': []}
        XU = np.zeros(N * (nx + nu) - nu)
        XU[:nx] = xcur
        XU_batch = self._solve(xcur_batch, goal_batch, np.tile(XU, (B, 1)))
        XU_best = XU_batch[0, :].copy()
        t_acc = 0.0
        for i in range(int(sim_time / sim_dt)):
            x_last, u_last = xcur, XU_best[nx:nx + nu].copy()
            rem = sim_dt
            q, v = xcur[:nq], xcur[nq:nx]
            while rem > 0:
                h = min(rem, self.dt - t_acc)
                knot = min(int(t_acc / self.dt), N - 1)
                u = XU_best[knot * (nx + nu) + nx:(knot + 1) * (nx + nu)]
                q, v = rbd.rk4(self.model, q, v, u, h, self.constant_f_ext)     # the 'world' force goes in unconverted
                t_acc += h
                rem -= h
                self.xpath.append(q)
                if abs(t_acc - self.dt) < 1e-10:
                    eepos_offset += 1
                    t_acc = 0.0
            xcur = np.hstack([q, v])
            for j in range(B):
                xcur_batch[j, :] = xcur
                XU_batch[j, :nx] = xcur
                XU_batch[j, nx:] = XU_best[nx:]
            if eepos_offset < len(fig8_traj) / 6 - 6 * N:
                eepos_goal = fig8_traj[6 * eepos_offset:6 * (eepos_offset + N)]
                goal_batch[:, :] = eepos_goal
            else:
                break
            goaldist = np.linalg.norm(rbd.eepos(self.model, xcur[:nq]) - eepos_goal[:3])
            XU_new = self._solve(xcur_batch, goal_batch, XU_batch)
            best_id = self.find_best_trajectory_id_fig8(x_last, u_last, xcur, sim_dt)
            self.best_ids.append(best_id)
            XU_best = XU_new[best_id, :].copy()
            if abs((i * sim_dt) % 0.05) < 1e-5:
                stats['goal_distances'].append(float(round(goaldist, 5)))
                stats['control_inputs'].append(np.round(XU_best[nx:nx + nu].copy(), 5))
        return self.xpath, stats

    def find_best_trajectory_id_fig8(self, x_last, u_last, xcur, dt):
        best_error, best_id = np.inf, None
        for i in range(self.batch_size):                 # solver.sim_forward: lane i under its own (unconverted) wrench
            q, v = rbd.rk4(self.model, x_last[:6], x_last[6:], u_last, dt, self.f_ext_batch[i])
            error = np.linalg.norm(np.hstack([q, v]) - xcur)
            if error <= best_error:
                best_error, best_id = error, i
        if self.resample_f_ext and self.f_ext_resample_std != 0:
            f_best = self.f_ext_batch[best_id, :].copy()
            self.f_ext_batch = np.tile(f_best, (self.batch_size, 1))
            self.f_ext_batch += np.random.normal(0, self.f_ext_resample_std, self.f_ext_batch.shape)
            self.f_ext_batch[:, 3:] = 0.0
            if self.batch_size > 1:
                self.f_ext_batch[1] = f_best
            self.f_ext_batch[0] = 0.0
        return best_id
