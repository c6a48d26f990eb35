#!/usr/bin/env python3
"""Device-timed sweep over the BASELINE.json configs that are not the bench.py headline line:
configs[2] (batch 16/32/64, N=32), the per-GPU shares of configs[3] (512..4096 problems) and configs[4]
(N=128, batch 512, fp64 vs fp32, with the fp32 deviation from fp64 on identical inputs).  One JSON line per case.
Warm-started MPC ticks on device tensors, CUDA events on the launch stream, L2 flushed between ticks."""
import argparse, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from indy7_mpc_b200 import Engine

ap = argparse.ArgumentParser()
ap.add_argument("--ticks", type=int, default=30)
ap.add_argument("--warmup", type=int, default=5)
ap.add_argument("--cases", default="1x32,16x32,32x32,64x32,148x32,512x32,1024x32,2048x32,4096x32,512x128")
ap.add_argument("--dtypes", default="fp64,fp32")
ap.add_argument("--threads", type=int, default=0)
ap.add_argument("--qp-mode", default="exact", help="exact | osqp (the reference regime: warm-started ADMM at eps 1e-3)")
ap.add_argument("--limits", action="store_true", help="URDF joint limits as inequality rows (osqp mode)")
ap.add_argument("--cluster", type=int, default=0, help="kernel form: 0 auto, 1 CTA per problem, 2|4|8 clusters (multi-wave when the batch is not co-resident)")
a = ap.parse_args()
dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def run(B, N, dtype):
    tdt = torch.float64 if dtype == "fp64" else torch.float32
    wl = bench.make_workload(B, N)
    eng = Engine(N=N, batch=B, dtype=dtype, threads=a.threads, cluster=a.cluster, qp_mode=a.qp_mode, limits=a.limits)
    eng.set_external_wrench(wl["f_ext"])
    fig = torch.tensor(wl["fig"], dtype=tdt, device=dev)
    off = torch.tensor(wl["offsets"], device=dev)
    ar = torch.arange(N, device=dev)
    xu = torch.tensor(wl["xu"], dtype=tdt, device=dev)
    x0 = torch.tensor(wl["x0"], dtype=tdt, device=dev)
    ms, first = [], None
    for t in range(a.warmup + a.ticks):
        goals = fig[(off[:, None] + t + ar[None, :]) % fig.shape[0]].reshape(B, 3 * N).contiguous()
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out, st = eng.solve(xu, x0, goals)
        e1.record()
        torch.cuda.synchronize()
        if t == 0:
            first = out.double().cpu().numpy()
        if t >= a.warmup:
            ms.append(e0.elapsed_time(e1))
        xu = torch.cat([out[:, 18:], out[:, -18:]], dim=1).contiguous()
        x0 = out[:, 18:30].contiguous()
        xu[:, :12] = x0
    ms = np.array(ms)
    sth = st.cpu().numpy()
    return dict(batch=B, horizon=N, dtype=dtype, qp_mode=a.qp_mode, limits=bool(a.limits), admm_iters_mean=float(sth[:, 10:12].sum(axis=1).mean()), threads=a.threads, cluster=a.cluster, form=list(eng.kernel_form()), ms_p50=float(np.median(ms)), ms_mean=float(ms.mean()),
                ms_p99=float(np.percentile(ms, 99)), solves_per_s=float(B / ms.mean() * 1e3),
                mean_qp_iters=float(sth[:, 0].mean()), nonfinite=int((~np.isfinite(sth)).sum())), first


for case in a.cases.split(","):
    B, N = (int(v) for v in case.split("x"))
    ref = None
    for dtype in a.dtypes.split(","):
        r, first = run(B, N, dtype)
        if dtype == "fp64":
            ref = first
        elif ref is not None:      # cold first tick, identical inputs: fp32 trajectory vs fp64 (north_star: 1e-4)
            r["fp32_vs_fp64_rel"] = float(np.abs(first - ref).max() / np.abs(ref).max())
        print(json.dumps(r), flush=True)
