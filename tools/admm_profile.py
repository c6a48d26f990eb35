#!/usr/bin/env python3
"""Cycle breakdown of the ADMM kernel of qp_mode 'osqp' (factorisation / per-iteration steps) for a given batch."""
import argparse, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from indy7_mpc_b200 import Engine

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=1)
ap.add_argument("--horizon", type=int, default=32)
a = ap.parse_args()
N, B = a.horizon, a.batch
wl = bench.make_workload(B, N)
eng = Engine(N=N, batch=B, qp_mode="osqp")
eng.set_external_wrench(wl["f_ext"])
eng.profile_phases(True)
xu, x0 = wl["xu"], wl["x0"]
for t in range(4):
    z = eng.solve_qp(xu, x0, bench.goals_at(wl["fig"], wl["offsets"], t, N))
c = eng.profile_phases(True, batch=B).astype(float)
_, _, _, info = eng.qp_state(B)
names = ["factor", "vectors", "1 t", "2 rhs", "3 L^-1 r", "4 fwd chain", "5 L^-T w", "6 bwd chain", "7-8 A x~ + updates", "9 check"]
print(f"B={B} N={N}: ADMM iterations (median) {np.median(info[:, 0]):.0f}; cycles per QP solve (median over problems):")
for i, n in enumerate(names):
    print(f"  {n:22s} {np.median(c[:, i]):10.0f}")
print(f"  total                  {np.median(c[:, :10].sum(axis=1)):10.0f}")
print("  inside the factorisation: load + pivot block %.0f, Cholesky + triangular inverse (one warp) %.0f, G / M products + stores %.0f, F product %.0f"
      % tuple(np.median(c[:, 12 + i]) for i in range(4)))
