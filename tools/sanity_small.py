#!/usr/bin/env python3
"""Small run through every kernel form and entry point (target for `compute-sanitizer --tool memcheck`)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from indy7_mpc_b200 import Engine

rng = np.random.default_rng(0)
for N, B, dtype, cluster, threads in [(8, 1, "fp64", 0, 0), (8, 3, "fp64", 4, 0), (8, 3, "fp64", 2, 0), (8, 3, "fp64", 1, 128),
                                      (8, 3, "fp64", 1, 256), (8, 3, "fp32", 1, 128), (8, 2, "fp32", 8, 0), (32, 2, "fp64", 1, 384),
                                      (64, 2, "fp64", 1, 128), (5, 2, "fp64", 8, 0)]:
    eng = Engine(N=N, batch=B, dtype=dtype, cluster=cluster, threads=threads)
    nz = 18 * N - 6
    x0 = np.concatenate([rng.uniform(-1, 1, (B, 6)), rng.normal(0, 0.2, (B, 6))], axis=1)
    xu = np.zeros((B, nz)); xu[:, :12] = x0
    goals = np.tile(eng.eepos(rng.uniform(-1, 1, (B, 6))), (1, N))
    eng.set_external_wrench(rng.normal(0, 2, (B, 6)))
    out, st = eng.solve(xu, x0, goals)
    out2, _ = eng.solve(np.asarray(out), x0, goals)
    z = eng.solve_qp(xu, x0, goals)
    eng.dump_qp(xu, x0, goals)
    eng.merit(np.asarray(out), xu, goals)
    eng.linesearch(xu, np.asarray(z), goals)
    eng.dynamics(x0[:, :6], x0[:, 6:], rng.normal(0, 1, (B, 6)))
    eng.sim_forward(x0, rng.normal(0, 1, (B, 6)), 0.01)
    p = eng.pack_best(st, out)
    eng.reduce_best(np.asarray(p)[None, :], 1)
    print(N, B, dtype, cluster, threads, eng.kernel_form(), "ok", float(np.abs(np.asarray(out2)).max()))
    eng.close()
os.environ["I7_PIPE_CHUNK"] = "4"
eng = Engine(N=8, batch=11, dtype="fp64")
x0 = np.concatenate([rng.uniform(-1, 1, (11, 6)), np.zeros((11, 6))], axis=1)
xu = np.zeros((11, 138)); xu[:, :12] = x0
out, st = eng.solve(xu, x0, np.tile(eng.eepos(x0[:, :6] * 0.5), (1, 8)))
print("pipeline ok", np.isfinite(np.asarray(out)).all())
