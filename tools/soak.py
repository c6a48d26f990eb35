import sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
import bench
from indy7_mpc_b200 import Engine
B, N = 4096, 32
wl = bench.make_workload(B, N)
for dtype, td in (("fp64", torch.float64), ("fp32", torch.float32)):
    eng = Engine(N=N, batch=B, dtype=dtype); eng.set_external_wrench(wl["f_ext"])
    dev = torch.device("cuda", 0)
    fig = torch.tensor(wl["fig"], dtype=td, device=dev); off = torch.tensor(wl["offsets"], device=dev); ar = torch.arange(N, device=dev)
    xu = torch.tensor(wl["xu"], dtype=td, device=dev); x0 = torch.tensor(wl["x0"], dtype=td, device=dev)
    bad = 0; fails = 0; iters = []
    for t in range(400):
        goals = fig[(off[:, None] + t + ar[None, :]) % fig.shape[0]].reshape(B, 3 * N).contiguous()
        out, st = eng.solve(xu, x0, goals)
        s = st.cpu().numpy()
        bad += int((s[:, 3] == 2).sum()) + int((~np.isfinite(s)).sum()); fails += int((s[:, 3] == 1).sum()); iters.append(s[:, 0].mean())
        xu = torch.cat([out[:, 18:], out[:, -18:]], dim=1).contiguous(); x0 = out[:, 18:30].contiguous(); xu[:, :12] = x0
    err = (out.view(B, -1)[:, :3 * 0 + 12]).abs().max().item()
    # tracking error of the first knot's end effector at the last tick
    p = np.asarray(eng.eepos(out[:, :6].double().cpu().numpy()))
    g = goals[:, :3].double().cpu().numpy()
    print(dtype, "ticks 400: non-finite/status2", bad, "linesearch fails", fails, "mean qp iters", float(np.mean(iters)),
          "median EE tracking error", float(np.median(np.linalg.norm(p - g, axis=1))))
