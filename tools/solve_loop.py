#!/usr/bin/env python3
"""A few warm-started solves of the synthetic workload at a given batch (profiling target for ncu)."""
import argparse, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from indy7_mpc_b200 import Engine

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=1)
ap.add_argument("--horizon", type=int, default=32)
ap.add_argument("--dtype", default="fp64")
ap.add_argument("--ticks", type=int, default=8)
ap.add_argument("--cluster", type=int, default=0)
a = ap.parse_args()
N, B = a.horizon, a.batch
wl = bench.make_workload(B, N)
eng = Engine(N=N, batch=B, dtype=a.dtype, cluster=a.cluster)
eng.set_external_wrench(wl["f_ext"])
xu, x0 = wl["xu"], wl["x0"]
for t in range(a.ticks):
    out, st = eng.solve(xu, x0, bench.goals_at(wl["fig"], wl["offsets"], t, N))
    out = np.asarray(out, float)
    xu, x0 = bench.shift_host(out, N); xu[:, :12] = x0
print(f"B={B} N={N}: last solve {eng.last_solve_time_us():.1f} us, alphas {st[0, 4:6]}")
