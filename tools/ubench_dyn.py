#!/usr/bin/env python3
"""Throughput of the per-sample (one thread = one knot, everything serial in registers) dynamics kernels on
device-resident data: forward dynamics + analytic derivatives (i7_dynamics) and the rk4 plant step
(i7_sim_forward).  Gives the cost of the lane-serial form of the SQP setup phase."""
import argparse, ctypes as C, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from indy7_mpc_b200 import Engine

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=4096 * 31)
ap.add_argument("--dtype", default="fp64")
a = ap.parse_args()
n = a.n
eng = Engine(N=32, batch=32768, dtype=a.dtype)
td = torch.float64 if a.dtype == "fp64" else torch.float32
g = torch.Generator(device="cuda").manual_seed(0)
q = (torch.rand(n, 6, device="cuda", dtype=td, generator=g) - 0.5) * 6
v = (torch.rand(n, 6, device="cuda", dtype=td, generator=g) - 0.5) * 4
u = (torch.rand(n, 6, device="cuda", dtype=td, generator=g) - 0.5) * 100
qdd = torch.empty(n, 6, device="cuda", dtype=td)
dq = torch.empty(n, 36, device="cuda", dtype=td); dv = torch.empty_like(dq); mi = torch.empty_like(dq)
p = lambda t: C.c_void_p(t.data_ptr())
s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]


def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        ev[0].record(); fn(); ev[1].record(); torch.cuda.synchronize()
        best = min(best, ev[0].elapsed_time(ev[1]))
    return best


ms = timeit(lambda: eng._check(eng.lib.i7_dynamics(eng._h, n, p(q), p(v), p(u), None, p(qdd), p(dq), p(dv), p(mi), s)))
print(f"i7_dynamics  n={n} {a.dtype}: {ms:.3f} ms  -> {ms * 1e6 / n:.1f} ns/sample")
x = torch.cat([q, v], dim=1).contiguous(); xn = torch.empty_like(x)
nb = min(n, eng.batch)
ms = timeit(lambda: eng._check(eng.lib.i7_sim_forward(eng._h, nb, p(x), p(u), 0.01, 0, p(xn), s)))
print(f"i7_sim_forward (4 ABA) n={nb} {a.dtype}: {ms:.3f} ms -> {ms * 1e6 / nb / 4:.1f} ns/ABA")
