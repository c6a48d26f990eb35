import sys, time, numpy as np, torch
sys.path.insert(0, '/root/repo')
from indy7_mpc_b200 import Engine
for B in (1, 64, 512, 4096):
    eng = Engine(N=8, batch=B)
    x = torch.zeros(B, 12, dtype=torch.float64, device="cuda"); x[:, :6] = 0.3
    u = torch.ones(B, 6, dtype=torch.float64, device="cuda")
    for _ in range(5): eng.sim_forward(x, u, 0.01)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(50): eng.sim_forward(x, u, 0.01)
    e1.record(); torch.cuda.synchronize()
    xh, uh = x.cpu().numpy(), u.cpu().numpy()
    t0 = time.perf_counter()
    for _ in range(50): eng.sim_forward(xh, uh, 0.01)
    w = (time.perf_counter() - t0) / 50 * 1e6
    print(B, "device us/call", e0.elapsed_time(e1) * 1e3 / 50, "host-buffer wall us/call", w)
