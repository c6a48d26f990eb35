"""Batched sampled-wrench MPC driver with the reference's interface (``GATO_Batch_Sample``,
``MPC_GATO_Batch_Sample``; src/gato_mpc_batch_sample.py:8-303) on the CUDA backend.

Every control step: B SQP problems (same state and goals, one external-wrench hypothesis per lane) are solved by
ONE fused kernel launch through the ``batch_sqp`` surface; the plant is the rk4 kernel with the actual wrench; the
hypothesis whose one-step prediction matches the measured state best is selected by ONE batched ``i7_sim_forward``
(all lanes at once instead of the reference's Python loop over copies of the Pinocchio model, :254-284).

Kept from the reference, on purpose: the solution is NOT shifted between steps (:193-201 is commented out there),
lane 0 always carries the zero wrench (:40,:296), the plant wrench random-walks with sigma 0.1 on its force part
after every step (:243-249), hypotheses are world-frame wrenches re-expressed in the last joint frame at the
previous state (:268-277), goals switch at 0.095 m (:209-218), logging every 0.05 s (:236-241).
Differences: ``constant_f_ext=None`` means a zero plant wrench (the reference raises TypeError at :243 in that
case); printing is off unless ``verbose=True``.
"""
import time

import numpy as np

from . import batch_sqp
from .engine import Engine
from .utils import rk4


class GATO_Batch_Sample:
    def __init__(self, N=32, dt=0.01, batch_size=4, stats=None, f_ext_std=1.0, verbose=False):
        self.N, self.dt = N, dt
        name = f"SQPSolverfloat_{batch_size}"
        if not hasattr(batch_sqp, name):
            raise ValueError(f"Batch size {batch_size} not supported")
        self.batch_size = batch_size
        self.solver = getattr(batch_sqp, name)()
        self.stats = stats or {
            'solve_time': {'values': [], 'unit': 'us', 'multiplier': 1},
            'sqp_iters': {'values': [], 'unit': '', 'multiplier': 1},
            'pcg_iters': {'values': [], 'unit': '', 'multiplier': 1},
            "step_size": {"values": [], "unit": "", "multiplier": 1},
        }
        if f_ext_std != 0.0:
            self.f_ext_batch = np.random.normal(0, f_ext_std, (self.batch_size, 6))
            self.f_ext_batch[:, 3:] = 0.0
            self.f_ext_batch[0] = 0.0
        else:
            self.f_ext_batch = np.zeros((self.batch_size, 6))
        if verbose:
            print("f_ext_batch:")
            print(self.f_ext_batch)
        self.solver.set_external_wrench_batch(self.f_ext_batch)

    def solve(self, xcur_batch, eepos_goals_batch, XU_batch):
        result = self.solver.solve(XU_batch, self.dt, xcur_batch, eepos_goals_batch)
        self.stats['solve_time']['values'].append(result["solve_time_us"])
        self.stats['sqp_iters']['values'].append(result["sqp_iterations"])
        for s in result["pcg_stats"]:
            self.stats['pcg_iters']['values'].append(s["pcg_iterations"])
        for s in result["line_search_stats"]:
            self.stats["step_size"]["values"].append(s["step_size"])
        return result["xu_trajectory"]

    def reset(self):
        self.solver.reset()

    def get_stats(self):
        return self.stats


class MPC_GATO_Batch_Sample:
    def __init__(self, model, N=32, dt=0.01, batch_size=4, constant_f_ext=None, resample_f_ext=False, f_ext_std=1.0,
                 f_ext_resample_std=2.0):
        self.model = model
        self.model.gravity.linear = np.array([0.0, 0.0, -9.81])
        self.data = model.createData()
        self.solver = GATO_Batch_Sample(N, dt, batch_size, f_ext_std=f_ext_std)
        self.xpath = []
        self.nq, self.nv = model.nq, model.nv
        self.nx = self.nq + self.nv
        self.nu = len(model.joints) - 1
        self.resample_f_ext = resample_f_ext
        self.f_ext_resample_std = f_ext_resample_std
        self.actual_f_ext = np.zeros(6)
        # plant wrench on the last joint, local frame, force part only (pin.Force(constant_f_ext[:3], 0), :96-101)
        self.constant_f_ext = np.zeros(6)
        if constant_f_ext is not None:
            self.constant_f_ext[:3] = np.asarray(constant_f_ext, float)[:3]
        self._hyp = Engine(model, N=2, dt=dt, batch=batch_size, dtype="fp64")       # hypothesis roll-outs (B lanes)

    def eepos(self, q):
        return self._hyp.eepos(np.asarray(q, float).reshape(1, 6))[0]

    def run_mpc(self, xstart, endpoints, sim_dt=0.001, sim_time=5, verbose=False):
        S, nx, nu = self.solver, self.nx, self.nu
        B, N, stride = S.batch_size, S.N, self.nx + self.nu
        xcur = np.array(xstart, float)
        xcur_batch = np.tile(xcur, (B, 1))
        endpoint = np.hstack([endpoints[0], np.zeros(3)])
        eepos_goal_batch = np.tile(np.tile(endpoint, N), (B, 1))
        stats = {'solve_times': [], 'goal_distances': [], 'control_inputs': []}
        XU = np.zeros(N * stride - nu)
        XU[:nx] = xcur
        XU_batch = np.array(S.solve(xcur_batch, eepos_goal_batch, np.tile(XU, (B, 1))), float)
        XU_best = XU_batch[0, :].copy()
        current_goal_idx, time_accumulated = 0, 0.0
        num_steps = int(sim_time / sim_dt)
        for i in range(num_steps):
            if verbose and i == num_steps - 1:
                print("Maximum steps reached without convergence")
            x_last = xcur
            u_last = XU_best[nx:nx + nu].copy()
            sim_time_step = sim_dt
            q, v = xcur[:self.nq], xcur[self.nq:nx]
            while sim_time_step > 0:
                timestep = min(sim_time_step, S.dt - time_accumulated)
                knot = min(int(time_accumulated / S.dt), N - 1)
                u = XU_best[knot * stride + nx:(knot + 1) * stride]
                q, v = rk4(self.model, self.data, q, v, u, timestep, self.constant_f_ext)
                time_accumulated += timestep
                sim_time_step -= timestep
                self.xpath.append(q)
                if abs(time_accumulated - S.dt) < 1e-10:
                    time_accumulated = 0.0
            xcur = np.hstack([q, v])
            xcur_batch[:] = xcur
            XU_batch[:, :nx] = xcur
            XU_batch[:, nx:] = XU_best[nx:]
            goaldist = float(np.linalg.norm(self.eepos(xcur[:self.nq]) - eepos_goal_batch[0, :3]))
            if goaldist < 0.095 and current_goal_idx < len(endpoints) - 1:
                current_goal_idx += 1
                endpoint = np.hstack([endpoints[current_goal_idx], np.zeros(3)])
                eepos_goal_batch = np.tile(np.tile(endpoint, N), (B, 1))
                if verbose:
                    print(f"Reached intermediate goal {current_goal_idx}, moving to next goal: {endpoint}")
            elif goaldist < 0.095 and current_goal_idx == len(endpoints) - 1:
                if verbose:
                    print(f"Reached final goal {current_goal_idx}")
                break
            S.reset()
            S.solver.resetRho()
            t0 = time.time()
            XU_batch_new = np.array(S.solve(xcur_batch, eepos_goal_batch, XU_batch), float)
            t1 = time.time()
            best_id = self.find_best_trajectory_id(x_last, u_last, xcur, sim_dt)
            XU_best = XU_batch_new[best_id, :].copy()
            if abs((i * sim_dt) % 0.05) < 1e-5:
                if verbose:
                    print(f"t: {i * sim_dt:8.3f}, goal dist: {goaldist:8.5f}, best_id: {best_id}")
                stats['goal_distances'].append(float(round(goaldist, 5)))
                stats['control_inputs'].append(np.round(XU_best[nx:nx + nu].copy(), 5))
                stats['solve_times'].append(float(round(t1 - t0, 5)))
            new_force = np.random.normal(self.constant_f_ext[:3], 0.1)
            self.constant_f_ext = np.concatenate([new_force, np.zeros(3)])
        return self.xpath, stats

    # ---- figure-8 tracking variant of notebooks/gato_mpc_indy7_fig8.ipynb (cell 2: run_mpc_fig8) ----------------
    def run_mpc_fig8(self, xstart, fig8_traj, sim_dt=0.001, sim_time=5, verbose=False):
        """Tracks ``fig8_traj`` (flat [x y z 0 0 0] per sample, one sample per solver dt): the goal window advances
        by one sample per completed solver step, the plant wrench is ``constant_f_ext`` applied as it is in the last
        joint frame (the notebook appends the 'world' force unconverted), hypotheses are rolled out by the solver's
        own ``sim_forward`` (lane wrenches unconverted as well) and ties go to the LAST best lane (`<=`)."""
        S, nx, nu = self.solver, self.nx, self.nu
        B, N, stride = S.batch_size, S.N, self.nx + self.nu
        fig8_traj = np.asarray(fig8_traj, float)
        xcur = np.array(xstart, float)
        xcur_batch = np.tile(xcur, (B, 1))
        eepos_offset = 0
        eepos_goal = fig8_traj[:6 * N]
        eepos_goal_batch = np.tile(eepos_goal, (B, 1))
        stats = {'solve_times': [], 'goal_distances': [], 'control_inputs': []}
        XU = np.zeros(N * stride - nu)
        XU[:nx] = xcur
        XU_batch = np.array(S.solve(xcur_batch, eepos_goal_batch, np.tile(XU, (B, 1))), float)
        XU_best = XU_batch[0, :].copy()
        time_accumulated = 0.0
        for i in range(int(sim_time / sim_dt)):
            x_last, u_last = xcur, XU_best[nx:nx + nu].copy()
            sim_time_step = sim_dt
            q, v = xcur[:self.nq], xcur[self.nq:nx]
            while sim_time_step > 0:
                timestep = min(sim_time_step, S.dt - time_accumulated)
                knot = min(int(time_accumulated / S.dt), N - 1)
                u = XU_best[knot * stride + nx:(knot + 1) * stride]
                q, v = rk4(self.model, self.data, q, v, u, timestep, self.constant_f_ext)
                time_accumulated += timestep
                sim_time_step -= timestep
                self.xpath.append(q)
                if abs(time_accumulated - S.dt) < 1e-10:
                    eepos_offset += 1
                    time_accumulated = 0.0
            xcur = np.hstack([q, v])
            xcur_batch[:] = xcur
            XU_batch[:, :nx] = xcur
            XU_batch[:, nx:] = XU_best[nx:]
            if eepos_offset < len(fig8_traj) / 6 - 6 * N:
                eepos_goal = fig8_traj[6 * eepos_offset:6 * (eepos_offset + N)]
                eepos_goal_batch[:, :] = eepos_goal
            else:
                break
            goaldist = float(np.linalg.norm(self.eepos(xcur[:self.nq]) - eepos_goal[:3]))
            S.reset()
            S.solver.resetRho()
            t0 = time.time()
            XU_batch_new = np.array(S.solve(xcur_batch, eepos_goal_batch, XU_batch), float)
            t1 = time.time()
            best_id = self.find_best_trajectory_id_fig8(x_last, u_last, xcur, sim_dt)
            XU_best = XU_batch_new[best_id, :].copy()
            if abs((i * sim_dt) % 0.05) < 1e-5:
                if verbose:
                    print(f"t: {i * sim_dt:8.3f}, goal dist: {goaldist:8.5f}, best_id: {best_id}")
                stats['goal_distances'].append(float(round(goaldist, 5)))
                stats['control_inputs'].append(np.round(XU_best[nx:nx + nu].copy(), 5))
                stats['solve_times'].append(float(round(t1 - t0, 5)))
        return self.xpath, stats

    def find_best_trajectory_id_fig8(self, x_last, u_last, xcur, dt):
        """The notebook's selection: one ``sim_forward`` of the solver over all lanes, last minimum wins; resampling
        keeps the best wrench in lane 1 and the zero wrench in lane 0."""
        S = self.solver
        S.solver.set_external_wrench_batch(S.f_ext_batch)
        best_id = S.solver.find_best_idx(x_last, u_last, xcur, float(dt), tie_last=True)   # one launch, `<=` argmin
        if self.resample_f_ext and self.f_ext_resample_std != 0:
            # f = f[best] + noise; torque part 0; lane 1 = f[best]; lane 0 = 0 — on the device (i7_resample_wrench); the
            # normal draws stay np.random's (the reference's stream, seeded on the host)
            noise = np.random.normal(0, self.f_ext_resample_std, S.f_ext_batch.shape)
            S.f_ext_batch = S.solver.resample_f_ext_batch(best_id, noise, keep_lane=1, keep_torque=True)
        return best_id

    def find_best_trajectory_id(self, x_last, u_last, xcur, dt):
        """Lane whose wrench hypothesis best predicts the measured state over one plant step (:254-300): one
        batched rk4 launch with the B world-frame hypotheses re-expressed in the last joint frame at x_last."""
        S = self.solver
        local = self.model.world_wrench_to_last_joint(np.asarray(x_last, float)[:self.nq], S.f_ext_batch)
        self._hyp.set_external_wrench(local)
        # plant step of every lane + error + first minimum (the strict '<' of the reference loop) in one launch
        best_id = self._hyp.find_best_idx(np.asarray(x_last, float), np.asarray(u_last, float), np.asarray(xcur, float),
                                          float(dt), batch=S.batch_size, tie_last=False)
        if self.resample_f_ext:
            f_best = S.f_ext_batch[best_id].copy()
            S.f_ext_batch = f_best + np.random.normal(0, self.f_ext_resample_std, S.f_ext_batch.shape)
            S.f_ext_batch[:, 3:] = 0.0
            S.f_ext_batch[0] = 0.0
            S.solver.set_external_wrench_batch(S.f_ext_batch)
        return best_id
