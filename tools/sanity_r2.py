#!/usr/bin/env python3
"""Small run through the round-2 device code (target for `compute-sanitizer --tool memcheck | racecheck`): the tile form of
the CTA kernel (I7_FORCE_SPILL=1 makes short horizons take it), the osqp QP mode with and without limits, the
task-parallel line-search points of the cluster forms, the solve kernel's selection / exchange epilogue, shift and goal
window."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from indy7_mpc_b200 import Engine

rng = np.random.default_rng(0)


def case(N, B, **kw):
    eng = Engine(N=N, batch=B, **kw)
    nz = 18 * N - 6
    x0 = np.concatenate([rng.uniform(-1, 1, (B, 6)), rng.normal(0, 0.2, (B, 6))], axis=1)
    xu = np.zeros((B, nz)); xu[:, :12] = x0
    goals = np.tile(eng.eepos(rng.uniform(-1, 1, (B, 6))), (1, N))
    eng.set_external_wrench(rng.normal(0, 2, (B, 6)))
    return eng, xu, x0, goals


os.environ["I7_FORCE_SPILL"] = "1"
for N, B, dtype in [(8, 2, "fp64"), (23, 2, "fp64"), (8, 2, "fp32")]:
    eng, xu, x0, goals = case(N, B, dtype=dtype, cluster=1, threads=128)
    out, st = eng.solve(xu, x0, goals)
    eng.merit(np.asarray(out), xu, goals)
    print("tile form", N, B, dtype, eng.kernel_form(), st[0, :6])
    eng.close()
del os.environ["I7_FORCE_SPILL"]
for limits in (False, True):
    eng, xu, x0, goals = case(8, 2, qp_mode="osqp", limits=limits, admm=dict(max_iter=50))
    out, st = eng.solve(xu, x0, goals)
    z = eng.solve_qp(xu, x0, goals)
    print("osqp mode, limits", limits, st[0], eng.qp_state(2)[3][0])
    eng.close()
for cl in (8, 4, 2):
    eng, xu, x0, goals = case(8, 2, cluster=cl)
    out, st = eng.solve(xu, x0, goals)
    print("cluster", cl, eng.kernel_form(), st[0, :6])
    eng.close()
eng, xu, x0, goals = case(8, 5, cluster=1)
eng.p2p_connect([eng.p2p_create(1, 0)])
d = lambda a: torch.tensor(a, device="cuda")
o, s, b = eng.solve_select(d(xu), d(x0), d(goals), index_offset=3)
nxt, x0n = eng.shift(o, steps=1, tail="repeat")
g = eng.goal_window(d(rng.uniform(-1, 1, (40, 3))), torch.arange(5, dtype=torch.int32, device="cuda"), tick=3, batch=5)
torch.cuda.synchronize()
print("epilogue selection", b[:2].cpu().numpy(), "shift", tuple(nxt.shape), "goals", tuple(g.shape))
