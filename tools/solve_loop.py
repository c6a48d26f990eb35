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
import torch
dev = torch.device("cuda", 0)                    # device-resident buffers: ONE launch per tick (host buffers would go
td = torch.float64 if a.dtype == "fp64" else torch.float32      # through the chunked pipeline of i7_solve)
xu, x0 = torch.tensor(wl["xu"], dtype=td, device=dev), torch.tensor(wl["x0"], dtype=td, device=dev)
for t in range(a.ticks):
    goals = torch.tensor(bench.goals_at(wl["fig"], wl["offsets"], t, N), dtype=td, device=dev)
    out, st = eng.solve(xu, x0, goals)
    xu = torch.cat([out[:, 18:], out[:, -18:]], dim=1).contiguous()
    x0 = out[:, 18:30].contiguous()
    xu[:, :12] = x0
torch.cuda.synchronize()
st = st.cpu().numpy()
print(f"B={B} N={N}: last solve {eng.last_solve_time_us():.1f} us, kernel form {eng.kernel_form()}, alphas {st[0, 4:6]}")
