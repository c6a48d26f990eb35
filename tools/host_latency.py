#!/usr/bin/env python3
"""Where the host wall clock of a single solve goes: Engine.solve (Python argument handling + C ABI) against the
bare i7_solve call with prepared pointers, on the warm fig-8 problem.  Prints p50 / p99 in microseconds."""
import argparse
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ticks", type=int, default=400)
    ap.add_argument("--horizon", type=int, default=32)
    args = ap.parse_args()
    from indy7_mpc_b200 import Engine
    from indy7_mpc_b200.reference_traj import figure8
    N = args.horizon
    eng = Engine(N=N, batch=1, dtype="fp64")
    fig = figure8(0.5, 0.55, (0.0, 0.4, 0.45), 10.0, 0.01, cycles=3, stride=3).reshape(-1, 3)
    x = np.concatenate([[0, 0, -1.57, 0, -1.57, 0], np.zeros(6)])[None, :]
    xu = np.zeros((1, 18 * N - 6)); xu[0, :12] = x[0]
    for t in range(30):                                   # warm trajectory
        goals = np.ascontiguousarray(fig[(t + np.arange(N)) % fig.shape[0]].reshape(1, -1))
        out, st = eng.solve(xu, x, goals)
        xu = np.ascontiguousarray(out)
    wall, bare, dev = [], [], []
    out_b, st_b = np.empty_like(xu), np.empty((1, 12))
    ptr = lambda a: C.c_void_p(a.ctypes.data)
    p_xu, p_x, p_g, p_o, p_s = ptr(xu), ptr(x), ptr(goals), ptr(out_b), ptr(st_b)
    for t in range(args.ticks):
        t0 = time.perf_counter()
        eng.solve(xu, x, goals)
        wall.append((time.perf_counter() - t0) * 1e6)
        dev.append(eng.last_solve_time_us() if os.environ.get("I7_TIMING_EVENTS", "1") != "0" else float("nan"))
        t0 = time.perf_counter()
        rc = eng.lib.i7_solve(eng._h, 1, p_xu, p_x, p_g, 3, p_o, p_s, None)
        bare.append((time.perf_counter() - t0) * 1e6)
        assert rc == 0
    q = lambda a: f"p50 {np.percentile(a, 50):7.1f}  p99 {np.percentile(a, 99):7.1f}"
    print(f"Engine.solve wall  {q(wall)}\nbare i7_solve wall {q(bare)}\ndevice (events)    {q(dev)}"
          f"   I7_TIMING_EVENTS={os.environ.get('I7_TIMING_EVENTS', '1')}")


if __name__ == "__main__":
    main()
