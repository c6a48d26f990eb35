#!/usr/bin/env python3
"""Per-lane solve durations (clock64) and statistics records of consecutive warm ticks — input for studying launch orders."""
import argparse, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from indy7_mpc_b200 import Engine

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=512)
ap.add_argument("--horizon", type=int, default=32)
ap.add_argument("--ticks", type=int, default=12)
ap.add_argument("--out", default="gpurun_out/lane_times.npz")
a = ap.parse_args()
N, B = a.horizon, a.batch
wl = bench.make_workload(4096, N)
wl = {k: (v[:B] if k != "fig" else v) for k, v in wl.items()}
eng = Engine(N=N, batch=B)
eng.set_external_wrench(wl["f_ext"])
eng.profile_phases(True)
xu, x0 = wl["xu"], wl["x0"]
T, S, C0 = [], [], []
for t in range(a.ticks):
    out, st = eng.solve(xu, x0, bench.goals_at(wl["fig"], wl["offsets"], t, N))
    c = eng.profile_phases(True, batch=B).astype(np.int64)
    T.append(c[:, 9] - c[:, 0]); S.append(np.asarray(st, float).copy()); C0.append(c[:, 0].copy())
    out = np.asarray(out, float)
    xu, x0 = bench.shift_host(out, N); xu[:, :12] = x0
np.savez(a.out, T=np.array(T), S=np.array(S), C0=np.array(C0))
T = np.array(T, float)
print("corr of lane durations tick t vs t+1:", [round(float(np.corrcoef(T[i], T[i + 1])[0, 1]), 3) for i in range(4, a.ticks - 1)])
