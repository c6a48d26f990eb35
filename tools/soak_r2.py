#!/usr/bin/env python3
"""Soak of the round-2 paths on one GPU: long closed loops of DeviceTick (local and fused-epilogue selection), the tile
form at N=128, and the osqp mode with limits; counts non-finite records, failed line searches, and tracks the goal."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from indy7_mpc_b200 import Engine, DeviceTick


def run(name, B, N, ticks, mode="local", **kw):
    wl = bench.make_workload(B, N)
    eng = Engine(N=N, batch=B, **kw)
    eng.set_external_wrench(wl["f_ext"])
    if mode == "fused":
        eng.p2p_connect([eng.p2p_create(1, 0)])
    tk = DeviceTick(eng, wl["fig"], wl["offsets"], wl["xu"], wl["x0"], mode=mode)
    bad = fails = 0
    its, admm = [], []
    for t in range(ticks):
        tk.run(t)
        if t % 20 == 19 or t == ticks - 1:
            s = tk.stats.cpu().numpy()
            bad += int((s[:, 3] == 2).sum()) + int((~np.isfinite(s)).sum())
            fails += int((s[:, 3] == 1).sum())
            its.append(s[:, 0].mean()); admm.append(s[:, 10:12].sum(axis=1).mean())
    if mode == "fused":
        eng.p2p_status()
    out = tk.out.double().cpu().numpy()
    p = np.asarray(eng.eepos(out[:, :6]))
    g = tk.goals[:, :3].double().cpu().numpy()
    print(f"{name}: {ticks} ticks x {B} lanes N={N}: non-finite/status2 {bad}, line-search fails (sampled) {fails}, mean qp iters "
          f"{np.mean(its):.2f}, mean ADMM its {np.mean(admm):.1f}, median EE tracking error {np.median(np.linalg.norm(p - g, axis=1)):.4f}, "
          f"best {tk.best[:2].cpu().numpy()}", flush=True)


run("exact, local selection", 4096, 32, 1000)
run("exact, fused epilogue", 4096, 32, 1000, mode="fused")
run("exact fp32", 4096, 32, 400, dtype="fp32")
run("tile form N=128 fp64", 512, 128, 200)
run("tile form N=128 fp32", 512, 128, 200, dtype="fp32")
run("latency forms B=16", 16, 32, 1000)
run("osqp mode", 512, 32, 100, qp_mode="osqp")
run("osqp mode + URDF limits", 512, 32, 100, qp_mode="osqp", limits=True)
