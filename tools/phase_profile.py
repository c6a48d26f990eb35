#!/usr/bin/env python3
"""Phase-level clock64 breakdown of the fused SQP kernel (setup / QP / merit / update) for a given batch."""
import argparse, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from indy7_mpc_b200 import Engine

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=1)
ap.add_argument("--threads", type=int, default=0)
ap.add_argument("--horizon", type=int, default=32)
ap.add_argument("--dtype", default="fp64")
ap.add_argument("--ticks", type=int, default=12)
a = ap.parse_args()
N, B = a.horizon, a.batch
wl = bench.make_workload(B, N)
eng = Engine(N=N, batch=B, dtype=a.dtype, threads=a.threads)
eng.set_external_wrench(wl["f_ext"])
eng.profile_phases(True)
xu, x0 = wl["xu"], wl["x0"]
for t in range(a.ticks):
    out, st = eng.solve(xu, x0, bench.goals_at(wl["fig"], wl["offsets"], t, N))
    us = eng.last_solve_time_us()
    out = np.asarray(out, float)
    xu, x0 = bench.shift_host(out, N); xu[:, :12] = x0
c = eng.profile_phases(True, batch=B).astype(float)
names = ["setup", "qp", "merit", "update"]
print(f"B={B} N={N} {a.dtype} threads={a.threads or 'auto'}: kernel {us:.1f} us; per-phase SM cycles (median over problems)")
for i in range(2):
    d = np.diff(c[:, 5 * i:5 * i + 5], axis=1)
    print(f"  qp {i}: " + ", ".join(f"{n} {np.median(d[:, j]):.0f}" for j, n in enumerate(names)) + f"   total {np.median(c[:, 5*i+4]-c[:, 5*i]):.0f}")
print(f"  first QP: 6x6 solves {np.median(c[:,10]):.0f}, value update {np.median(c[:,11]):.0f}, roll-out {np.median(c[:,12]):.0f}, controls {np.median(c[:,13]):.0f}")
print("  first setup waves W0..W9: " + ", ".join(f"{v:.0f}" for v in np.median(np.diff(c[:, 16:27], axis=1), axis=0)))
tot = c[:, 9] - c[:, 0]
print(f"  whole solve (first stamp -> last): median {np.median(tot):.0f} cycles; alphas {st[0,4:6]}")
print("  whole solve percentiles 1/10/50/90/99/max: " + ", ".join(f"{np.percentile(tot, q):.0f}" for q in (1, 10, 50, 90, 99, 100)))
if np.any(c[:, 27] > 0):       # cluster kernel: finer stamps of the first QP phase (lead CTA)
    m = lambda a: float(np.median(a))
    print(f"  first QP (cluster): barrier1->init done {m(c[:,14]-c[:,1]):.0f}, sweep loop+roll-out {m(c[:,27]-c[:,14]):.0f}, "
          f"publish {m(c[:,28]-c[:,27]):.0f}, barrier2 {m(c[:,29]-c[:,28]):.0f}, zx load {m(c[:,30]-c[:,29]):.0f}, "
          f"controls {m(c[:,2]-c[:,30]):.0f}")
