#!/usr/bin/env python3
"""Closed loop of ControllerTick (the reference controller's tick, gato_controller.py:201-256) against a plant that
carries an unknown external wrench: does the hypothesis selection find it, and does the arm track the figure 8?"""
import argparse, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from indy7_mpc_b200 import Engine, ControllerTick
from indy7_mpc_b200.reference_traj import figure8

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--horizon", type=int, default=32)
ap.add_argument("--ticks", type=int, default=600)
ap.add_argument("--f-actual", type=float, nargs=3, default=[-12.0, 6.0, -9.0])
a = ap.parse_args()
np.random.seed(42)                                                        # gato_controller.py:308
dt, B, N = 0.01, a.batch, a.horizon
ref = figure8(A_x=0.5, A_z=0.55, offset=[0.0, 0.4, 0.45], period=10, dt=dt, cycles=2)
ref = np.concatenate([np.tile(ref[0:6], 200), ref])                       # :322-323
eng = Engine(N=N, dt=dt, batch=B)
plant = Engine(N=N, dt=dt, batch=1)
f_act = np.array([*a.f_actual, 0.0, 0.0, 0.0])
plant.set_external_wrench(f_act[None, :])
ct = ControllerTick(eng, ref, B, f_ext_std=8.0, f_ext_resample_std=0.5)
x = np.array([1.57994932, 0.06313131, -1.18071498, 1.09272369, -0.6255293, -0.01895007, 0, 0, 0, 0, 0, 0.0])
errs, fest = [], []
for t in range(a.ticks):
    u, best = ct.tick(x, dt)
    x = np.asarray(plant.sim_forward(x, u, dt, batch=1), float).reshape(12)
    off = int(ct.ref_traj_offset)
    errs.append(float(np.linalg.norm(np.asarray(eng.eepos(x[None, :6]))[0] - ref[6 * off:6 * off + 3])))
    if t % 100 == 99:
        f = ct.f_ext_batch
        fest.append(f[best, :3].copy())
        print(f"tick {t + 1}: tracking error (last 100 ticks, mean) {np.mean(errs[-100:]):.4f} m, selected lane {best}, its wrench "
              f"{np.round(f[best, :3], 2)}, actual {f_act[:3]}, finite {bool(np.all(np.isfinite(x)))}", flush=True)
