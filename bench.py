#!/usr/bin/env python3
"""Benchmark of the SQP-MPC hot path (BASELINE.json metric: batched SQP-MPC solves/s + p50 single-solve latency).

  python bench.py --gpus N --steps K --warmup W            the B200 arm (one rank per GPU under torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K ...  the reference arm: the CPU restatement of osqp_sqp
                                                            (oracle/c, OSQP-style ADMM) on the host cores

Workload (BASELINE.json configs[3]): a batch of 4096 independent Indy7 N=32 SQP-MPC
problems — sampled start states around the simulator home pose, figure-8 end-effector references at per-lane
offsets, per-lane external-wrench hypotheses — sharded across the ranks (strong scaling: the total batch is fixed).
One "step" = one control tick: every lane runs one sqp() (fused sm_100a kernel), then the per-problem costs and
the best trajectory are selected and, for N > 1, exchanged by ONE kernel over peer memory (i7_p2p_exchange: pack +
NVLink stores + selection; I7_P2P=0: i7_pack_best, ONE NCCL all-gather, i7_reduce_best).
Between ticks (untimed) the lanes advance like an MPC loop: x0 <- x_1 of the solution, trajectory shifted by one
knot, goal window advanced by one sample — so every timed step is a realistic warm-started solve.

End-to-end arm ("e2e"): the same closed loop from page-locked HOST buffers through the public API — every tick the
kernel reads XU / x0 / goals of all lanes from host memory (zero-copy), the selection (N > 1: the exchange kernel) runs
on the GPU and the result record is read back, all inside the wall-clock-timed call; "all_trajectories_to_host" is the
variant that writes every trajectory back and selects on the host.

Timing: CUDA events around every step on the launch stream, L2 flushed (256 MB memset) between steps outside the
events, sum over the K timed steps, MAX over ranks; barrier + synchronize on both sides of the timed loop.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

HOME = np.array([1.57994932, 0.06313131, -1.18071498, 1.09272369, -0.6255293, -0.01895007])   # sim_node.cpp:196
JOINT_LIMIT = 3.0543261909900767                                                              # indy7.urdf:203
METRIC = "sqp_mpc_solves_per_sec"
UNIT = "solves/s"


# ------------------------------------------------------------------------------------------- workload
def make_workload(batch_total, N, seed=1234):
    """SURVEY.md §8d config 4 generator (numpy default_rng(1234))."""
    from indy7_mpc_b200.reference_traj import figure8
    rng = np.random.default_rng(seed)
    x0 = np.zeros((batch_total, 12))
    x0[:, :6] = np.clip(HOME + rng.normal(0, 0.05, (batch_total, 6)), -JOINT_LIMIT, JOINT_LIMIT)
    x0[:, 6:] = rng.normal(0, 0.1, (batch_total, 6))
    f_ext = np.zeros((batch_total, 6))
    f_ext[:, :3] = rng.normal(0, 5.0, (batch_total, 3))
    f_ext[0] = 0.0
    fig = figure8(0.5, 0.55, (0.0, 0.4, 0.45), 10.0, 0.01, cycles=3, stride=3).reshape(-1, 3)     # 3000 samples
    offsets = (37 * np.arange(batch_total)) % 1000
    xu = np.zeros((batch_total, 18 * N - 6))
    xu[:, :12] = x0
    return dict(x0=x0, f_ext=f_ext, fig=fig, offsets=offsets, xu=xu)


def goals_at(fig, offsets, tick, N):
    idx = (offsets[:, None] + tick + np.arange(N)[None, :]) % fig.shape[0]
    return fig[idx].reshape(len(offsets), 3 * N)


def shift_host(out, N):
    nxt = np.concatenate([out[:, 18:], out[:, -18:]], axis=1)
    x0 = out[:, 18:30].copy()
    return nxt, x0


def algorithmic_flops(stats, N):
    """SURVEY.md §8d flop model per solve, evaluated on the measured (n_qp, n_try) of every lane."""
    F_abad, F_aba, F_fk, F_fkjac, F_cost, F_lin, F_mk = 8268, 2067, 400, 500, 100, 432, 60
    F_setup = (N - 1) * (F_abad + F_lin) + N * (F_fkjac + F_cost)
    F_qp = (N - 1) * 8640 + N * 4032 + N * 2700
    F_merit = (N - 1) * F_aba + N * (F_fk + F_mk)
    st = np.asarray(stats, float)
    n_ls = st[:, 1]
    total = 0.0
    for i in range(2):
        a = st[:, 4 + i]
        ran = n_ls > i
        tries = np.where(a > 0, np.round(-np.log2(np.maximum(a, 2.0 ** -7))) + 1, 8)
        total += np.sum(ran * (F_setup + F_qp + (1 + tries) * F_merit))
    return total / len(st)


def algorithmic_bytes(N, esize):
    # in: xu (18N-6) + x0 12 + goals 3N + f_ext 6 ; out: xu (18N-6) + stats 12
    return (2 * (18 * N - 6) + 12 + 3 * N + 6 + 12) * esize


# ------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.samples, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "50", "-i", str(index)], stdout=subprocess.PIPE, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for ts, line in self.samples:
            if ts < t0 - 0.05 or ts > t1 + 0.1:
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except Exception:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------- CPU baseline
def host_cores():
    """Host threads this process may use (the affinity mask, not OMP_NUM_THREADS, which torchrun forces to 1)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_baseline(N, wl, tick_goals, seconds, sample_cap, nthreads=0):
    """The plain-C restatement of osqp_sqp (OSQP-style ADMM) on the host cores, OpenMP across problems."""
    from oracle.cport import COracle
    from oracle.model import load_model
    model = load_model()
    cores = host_cores()
    nthreads = nthreads or cores      # explicit: torchrun exports OMP_NUM_THREADS=1 to every rank
    chunk = max(8, min(sample_cap, 8 * cores))
    orc = COracle(model, N=N, backend="admm", batch=chunk)
    done, t0 = 0, time.perf_counter()
    lo = 0
    B = wl["x0"].shape[0]
    while True:
        idx = (lo + np.arange(chunk)) % B
        orc.reset()
        orc.sqp(wl["x0"][idx], tick_goals[idx], wl["xu"][idx], f_ext=wl["f_ext"][idx], nthreads=nthreads)
        done += chunk; lo += chunk
        el = time.perf_counter() - t0
        if el >= seconds or done >= sample_cap:
            break
    return {"value": done / el, "unit": UNIT, "cores": nthreads, "kind": "port",
            "sample": f"{done} cold-start problems of the same workload (tick 0) in {el:.1f} s, C restatement of "
                      f"osqp_sqp with OSQP-style ADMM (oracle/c), OpenMP over problems"}


def run_reference(args):
    """Reference arm: CPU restatement of the reference path on the host cores (Pinocchio / OSQP are not installable
    here, so oracle/_ref does not exist; kind = 'port')."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle.cport import COracle
    from oracle.model import load_model
    N = args.horizon
    wl = make_workload(args.batch, N)
    model = load_model()
    cores = host_cores()
    per_step = max(8, min(args.batch, 4 * cores))
    orc = COracle(model, N=N, backend="admm", batch=per_step)
    idx = np.arange(per_step)
    xu, x0 = wl["xu"][idx].copy(), wl["x0"][idx].copy()
    times = []
    for t in range(args.warmup + args.steps):
        goals = goals_at(wl["fig"], wl["offsets"][idx], t, N)
        t0 = time.perf_counter()
        out, st, _ = orc.sqp(x0, goals, xu, f_ext=wl["f_ext"][idx], nthreads=cores)
        dt = time.perf_counter() - t0
        if t >= args.warmup:
            times.append(dt)
        xu, x0 = shift_host(out, N)
        xu[:, :12] = x0
    total = float(np.sum(times))
    value = per_step * len(times) / total
    sample = f"{per_step} lanes of the {args.batch}-lane workload per step, warm-started MPC ticks"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"batch {args.batch} sampled Indy7 SQP-MPC problems, N={N} (bounded sample on CPU)",
                   "horizon": N, "batch_total": args.batch, "sample_per_step": per_step},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------- B200 arm
def run_b200(args):
    import torch
    import torch.distributed as dist
    from indy7_mpc_b200 import Engine, Indy7SQP
    from indy7_mpc_b200 import _lib
    from indy7_mpc_b200 import dist as i7d

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # NCCL writes its version / debug lines to stdout; keep stdout for the ONE JSON line of the contract
        os.environ.setdefault("NCCL_DEBUG_FILE", "/tmp/nccl_debug_%h_%p.log")
        dist.init_process_group("nccl", device_id=dev)
    N, K, W = args.horizon, args.steps, args.warmup
    assert args.batch % world == 0, "batch must divide across ranks"
    B = args.batch // world
    esize = 8 if args.dtype == "fp64" else 4
    tdt = torch.float64 if args.dtype == "fp64" else torch.float32
    nz = 18 * N - 6

    wl_all = make_workload(args.batch, N)
    sl = slice(rank * B, (rank + 1) * B)
    wl = {k: (v[sl] if k != "fig" else v) for k, v in wl_all.items()}
    eng = Engine(N=N, batch=B, dtype=args.dtype, device=local, threads=args.threads)
    eng.set_external_wrench(wl["f_ext"])
    fig_d = torch.tensor(wl["fig"], dtype=tdt, device=dev)
    off_d = torch.tensor(wl["offsets"], device=dev)
    ar_d = torch.arange(N, device=dev)

    def goals_dev(t):
        idx = (off_d[:, None] + t + ar_d[None, :]) % fig_d.shape[0]
        return fig_d[idx].reshape(B, 3 * N).contiguous()

    def shift_dev(out):
        nxt = torch.cat([out[:, 18:], out[:, -18:]], dim=1).contiguous()
        return nxt, out[:, 18:30].contiguous()

    gathered = torch.empty(world * (B + 2 + nz), dtype=tdt, device=dev) if world > 1 else None
    # exchange of the per-rank records: ONE kernel over peer memory (NVLink stores + flags), fused with the packing and
    # the selection; I7_P2P=0 selects the NCCL path (pack kernel, all_gather_into_tensor, reduce kernel)
    use_p2p = world > 1 and os.environ.get("I7_P2P", "1") != "0"
    if use_p2p:
        use_p2p = i7d.connect_peers(eng, rank, world)      # False (on every rank) -> the NCCL path

    def exchange(st, out):
        if use_p2p:
            return eng.p2p_exchange(st, out, index_offset=rank * B)
        packed = eng.pack_best(st, out, index_offset=rank * B)
        return i7d.gather_best(packed, B, world, eng.reduce_best, gathered)[0]

    def tick(xu, x0, goals):
        out, st = eng.solve(xu, x0, goals)
        best = exchange(st, out)
        if use_p2p:            # warm-up ticks double as a check against the NCCL path
            packed = eng.pack_best(st, out, index_offset=rank * B)
            ref = i7d.gather_best(packed, B, world, eng.reduce_best, gathered)[0]
            if not torch.equal(best, ref):
                raise RuntimeError("peer-memory exchange disagrees with the NCCL all-gather path")
        return out, st, best

    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
    xu = torch.tensor(wl["xu"], dtype=tdt, device=dev)
    x0 = torch.tensor(wl["x0"], dtype=tdt, device=dev)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]

    # ---- device-resident arm: W warm-up ticks, then K timed ticks ----
    for t in range(W):
        out, st, best = tick(xu, x0, goals_dev(t))
        xu, x0 = shift_dev(out)
        xu[:, :12] = x0
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = ClockSampler(local) if rank == 0 else None
    launches0 = eng.launch_count()
    t_wall0 = time.time()
    last_st = None
    for i in range(K):
        goals = goals_dev(W + i)
        flush.zero_()                                    # L2 flush, outside the timed events
        ev[i][0].record()
        kev[i][0].record()
        out, st = eng.solve(xu, x0, goals)               # dominant kernel timed on its own as well
        kev[i][1].record()
        best = exchange(st, out)                         # world > 1: ONE exchange per tick (peer memory, or NCCL all-gather)
        ev[i][1].record()
        xu, x0 = shift_dev(out)
        xu[:, :12] = x0
        last_st = st
    torch.cuda.synchronize()
    if use_p2p:
        eng.p2p_status()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t_wall1 = time.time()
    launches = eng.launch_count() - launches0
    step_ms = np.array([a.elapsed_time(b) for a, b in ev])
    kern_ms = np.array([a.elapsed_time(b) for a, b in kev])
    total_ms = torch.tensor([step_ms.sum()], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    total_ms = float(total_ms.item())
    st_host = last_st.cpu().numpy()

    # ---- end-to-end arm through the public API with HOST buffers (H2D + D2H inside the timed region) ----
    solver = Indy7SQP(N=N, batch=B, dtype=args.dtype, device=f"cuda:{local}", threads=args.threads)
    solver.set_external_wrench(wl["f_ext"])
    hxu, hx0 = wl["xu"].copy(), wl["x0"].copy()
    npdt = np.float64 if args.dtype == "fp64" else np.float32
    # the caller's buffers are page-locked host memory (the C ABI then transfers without a staging copy)
    pin = lambda shape: torch.empty(shape, dtype=tdt, pin_memory=True)
    p_xu, p_x0, p_goals = pin((B, nz)), pin((B, 12)), pin((B, 3 * N))
    def e2e_loop(select_on_device):
        """W + K closed-loop ticks from HOST buffers; returns the sum of the K timed ticks (max over ranks).
        select_on_device: the trajectories stay on the GPU, the selection (+ exchange for N > 1) runs there and only
        its result comes back; otherwise every trajectory and stats row comes back and the selection runs on the host."""
        hxu, hx0 = wl["xu"].copy(), wl["x0"].copy()
        times = []
        for t in range(W + K):
            p_goals.copy_(torch.from_numpy(goals_at(wl["fig"], wl["offsets"], t, N).astype(npdt)))
            p_xu.copy_(torch.from_numpy(hxu.astype(npdt)))
            p_x0.copy_(torch.from_numpy(hx0.astype(npdt)))
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            if select_on_device:
                out_d, st_d = solver.engine.solve(p_xu, p_x0, p_goals, device_out=True)
                if world > 1 and use_p2p:
                    res_d = solver.engine.p2p_exchange(st_d, out_d, index_offset=rank * B)
                else:
                    res_d = solver.engine.pack_best(st_d, out_d, index_offset=rank * B)
                    if world > 1:
                        res_d = i7d.gather_best(res_d, B, world, solver.engine.reduce_best, gathered)[0]
                rec = res_d.cpu().numpy()                      # the step's result: [costs |] best_cost, index, XU_best
                dt_ = time.perf_counter() - t0
                full = out_d.cpu().numpy().astype(float)       # untimed: the host keeps the warm start of every lane
            else:
                x_traj, u_traj, stats = solver.solve(p_x0, p_goals, xu_init=p_xu)
                cost = np.asarray(stats["cost"], float)        # best-trajectory selection on the host result
                bi = int(np.argmin(np.where(np.isnan(cost), np.inf, cost)))
                rec = np.concatenate([[cost[bi], rank * B + bi], stats["xu"][bi].numpy()]).astype(np.float64)
                if world > 1:
                    g = torch.empty(world * rec.size, dtype=torch.float64, device=dev)
                    dist.all_gather_into_tensor(g, torch.tensor(rec, device=dev))
                    gh = g.cpu().numpy().reshape(world, -1)
                    rec = gh[int(np.argmin(gh[:, 0]))]
                dt_ = time.perf_counter() - t0
                full = stats["xu"].numpy().astype(float)
            if t >= W:
                times.append(dt_)
            hxu, hx0 = shift_host(full, N)
            hxu[:, :12] = hx0
        total = torch.tensor([float(np.sum(times))], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(total, op=dist.ReduceOp.MAX)
        return float(total.item()), int(rec.size)

    if world > 1 and use_p2p:
        use_p2p = i7d.connect_peers(solver.engine, rank, world)
    e2e_total, res_numbers = e2e_loop(True)
    e2e_full_total, _ = e2e_loop(False)
    clk = clocks.stop(t_wall0, time.time()) if clocks else None

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- single-problem latency (BASELINE.json configs[1]): fig-8 tracking, B = 1, warm-started closed loop ----
    lat = single_problem_latency(args, local)

    # ---- roofline of the dominant kernel (sqp_kernel) ----
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
    kern_avg_s = float(kern_ms.mean()) * 1e-3
    bytes_launch = algorithmic_bytes(N, esize) * B
    flops_solve = algorithmic_flops(st_host, N)
    fma_peak = eng.measure_fma_peak()
    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic = tj.get(f"sqp_kernel_B{B}_N{N}_{args.dtype}")
        if traffic is None and f"sqp_kernel_B4096_N{N}_{args.dtype}" in tj:      # per-launch, scaled by problems
            traffic = tj[f"sqp_kernel_B4096_N{N}_{args.dtype}"] * B // 4096
    except Exception:
        pass
    roofline = {
        "bound": "hbm", "achieved": bytes_launch / kern_avg_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
        "frac": bytes_launch / kern_avg_s / 1e9 / hbm_peak, "traffic": traffic, "peak_source": peak_src,
        "kernel": f"sqp_kernel<{args.dtype},{eng_threads(eng, B, args)}>", "kernel_ms": float(kern_ms.mean()),
        "algorithmic_bytes_per_solve": algorithmic_bytes(N, esize),
        "note": "the path is FP-pipe / dependency-latency bound, not HBM bound (SURVEY.md §8d); see 'fp_pipe'",
        "fp_pipe": {"achieved_tflops": flops_solve * B / kern_avg_s / 1e12, "peak_tflops": fma_peak,
                    "frac": flops_solve * B / kern_avg_s / 1e12 / fma_peak if fma_peak else None,
                    "algorithmic_flops_per_solve": flops_solve,
                    "peak_source": "i7_measure_fma_peak (FMA microbenchmark on this GPU, this run)"},
    }

    cpu = None
    if not args.no_cpu_baseline:
        cpu = cpu_baseline(N, wl_all, goals_at(wl_all["fig"], wl_all["offsets"], 0, N), args.cpu_seconds, args.cpu_sample)

    value = args.batch * K / (total_ms * 1e-3)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64" if args.dtype == "fp64" else "f32", "data": "synthetic",
        "config": {"workload": f"batch {args.batch} sampled Indy7 SQP-MPC problems (BASELINE.json configs[3]), N={N}, "
                               f"dt=0.01, warm-started MPC ticks, per-lane wrench hypotheses, best-trajectory selection"
                               + ((" + one peer-memory exchange kernel (NVLink stores, fused pack + gather + selection)"
                                   if use_p2p else " + one NCCL all-gather") if world > 1 else ""),
                   "horizon": N, "batch_total": args.batch, "batch_per_gpu": B, "parallelism": f"batch-shard x{world}",
                   "l2": "flushed between timed steps (256 MB memset outside the events)",
                   "threads_per_cta": eng_threads(eng, B, args)},
        "e2e": {"value": args.batch * K / e2e_total, "unit": UNIT,
                "h2d_bytes_per_step": int(B * (nz + 12 + 3 * N) * esize), "d2h_bytes_per_step": int(res_numbers * esize),
                "ms_per_step": 1e3 * e2e_total / K,
                "api": "Engine.solve(XU, x0, xref: page-locked host tensors, device_out=True) -> selection on the GPU "
                       "(pack_best; N > 1: the exchange kernel) -> .cpu() of the result [costs | best_cost, best_index, XU_best]; "
                       "the kernel reads the host buffers itself (zero-copy), the result read is inside the timed call",
                "all_trajectories_to_host": {
                    "value": args.batch * K / e2e_full_total, "ms_per_step": 1e3 * e2e_full_total / K,
                    "d2h_bytes_per_step": int(B * (nz + 12) * esize),
                    "api": "Indy7SQP.solve(x0, xref, xu_init) with page-locked host tensors: every trajectory and stats "
                           "row written back to host memory, selection on the host (N > 1: + NCCL all-gather of the records)"}},
        "gpu_launches": int(launches),
        "clocks": clk,
        "roofline": roofline,
        "cpu_baseline": cpu,
        "latency": lat,
        "wall_s_timed_loop": t_wall1 - t_wall0,
        "stats_last_tick": {"mean_qp_iters": float(st_host[:, 0].mean()), "linesearch_fail_frac": float((st_host[:, 3] == 1).mean()),
                            "nonfinite": int((st_host[:, 3] == 2).sum())},
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def eng_threads(eng, B, args):
    return eng.kernel_form()[0]


def single_problem_latency(args, local):
    """BASELINE.json configs[1]: fig-8 tracking, N=32, one problem, one B200: p50 / p99 of the device-timed solve and
    of the host wall clock through the public API, warm-started closed loop with the rk4 plant (zero wrench)."""
    import torch
    from indy7_mpc_b200 import Engine
    from indy7_mpc_b200.reference_traj import figure8
    N = args.horizon
    n_ticks = args.latency_ticks
    eng = Engine(N=N, batch=1, dtype=args.dtype, device=local)
    fig = figure8(0.5, 0.55, (0.0, 0.4, 0.45), 10.0, 0.01, cycles=3, stride=3).reshape(-1, 3)
    x = np.concatenate([HOME, np.zeros(6)])
    xu = np.zeros((1, 18 * N - 6)); xu[0, :12] = x
    dev_us, wall_us, qp_it = [], [], []
    npdt = np.float64 if args.dtype == "fp64" else np.float32
    for t in range(n_ticks + 20):
        goals = fig[(t + np.arange(N)) % fig.shape[0]].reshape(1, -1)
        t0 = time.perf_counter()
        out, st = eng.solve(xu.astype(npdt), x.reshape(1, 12).astype(npdt), goals.astype(npdt))
        w = (time.perf_counter() - t0) * 1e6
        d = eng.last_solve_time_us()
        if t >= 20:
            dev_us.append(d); wall_us.append(w); qp_it.append(float(st[0, 0]))
        out = np.asarray(out, float)
        xn = np.asarray(eng.sim_forward(x.astype(npdt), out[0, 12:18].astype(npdt), 0.01), float)[0]
        x = xn
        xu = np.concatenate([out[:, 18:], out[:, -18:]], axis=1)
        xu[0, :12] = x
    p = lambda a, q: float(np.percentile(a, q)) if len(a) else None
    if not dev_us:
        return {"config": "skipped (--latency-ticks 0)"}
    # the same closed loop with DEVICE-resident buffers: the solve kernel alone (the host-buffer path above lets the
    # kernel read / write mapped host memory itself, so its event time includes those PCIe accesses)
    tdt = torch.float64 if args.dtype == "fp64" else torch.float32
    dev = torch.device("cuda", local)
    xd = torch.tensor(np.concatenate([HOME, np.zeros(6)]), dtype=tdt, device=dev).reshape(1, 12)
    xud = torch.zeros(1, 18 * N - 6, dtype=tdt, device=dev); xud[0, :12] = xd[0]
    figd = torch.tensor(fig, dtype=tdt, device=dev)
    res_us = []
    for t in range(n_ticks + 20):
        goals = figd[(t + torch.arange(N, device=dev)) % figd.shape[0]].reshape(1, -1).contiguous()
        out, st = eng.solve(xud, xd, goals)
        if t >= 20:
            res_us.append(eng.last_solve_time_us())
        xd = eng.sim_forward(xd, out[:, 12:18].contiguous(), 0.01)
        xud = torch.cat([out[:, 18:], out[:, -18:]], dim=1).contiguous()
        xud[0, :12] = xd[0]
    res = {"config": f"fig-8 tracking, N={N}, 1 problem, closed loop with rk4 plant, {n_ticks} warm ticks",
           "device_us": {"p50": p(res_us, 50), "p99": p(res_us, 99), "mean": float(np.mean(res_us)),
                         "what": "solve kernel, buffers resident in HBM (CUDA events)"},
           "device_us_host_buffers": {"p50": p(dev_us, 50), "p99": p(dev_us, 99), "mean": float(np.mean(dev_us)),
                                      "what": "solve kernel reading / writing mapped host memory (the host-buffer path)"},
           "host_wall_us": {"p50": p(wall_us, 50), "p99": p(wall_us, 99), "mean": float(np.mean(wall_us)),
                            "what": "Engine.solve with numpy buffers, wall clock"},
           "mean_qp_iters": float(np.mean(qp_it))}
    if not args.no_cpu_baseline:
        res["cpu_1thread_us"] = cpu_single_latency(N, fig, min(n_ticks, 100))
    return res


def cpu_single_latency(N, fig, n_ticks):
    """The same closed loop on ONE host thread: C restatement of osqp_sqp (OSQP-style ADMM, warm-started like the
    reference's persistent OSQP object) + its rk4 plant (single-problem CPU latency)."""
    from oracle.cport import COracle
    from oracle.model import load_model
    orc = COracle(load_model(), N=N, backend="admm", batch=1)
    x = np.concatenate([HOME, np.zeros(6)])
    xu = np.zeros((1, 18 * N - 6)); xu[0, :12] = x
    us = []
    for t in range(n_ticks + 5):
        goals = fig[(t + np.arange(N)) % fig.shape[0]].reshape(1, -1)
        t0 = time.perf_counter()
        out, _, _ = orc.sqp(x.reshape(1, 12), goals, xu, nthreads=1)
        w = (time.perf_counter() - t0) * 1e6
        if t >= 5:
            us.append(w)
        q, v = orc.rk4(x[:6], x[6:], out[0, 12:18], 0.01)
        x = np.concatenate([q, v])
        xu = np.concatenate([out[:, 18:], out[:, -18:]], axis=1)
        xu[0, :12] = x
    return {"p50": float(np.percentile(us, 50)), "p99": float(np.percentile(us, 99)), "ticks": n_ticks,
            "kind": "port (oracle/c, ADMM back-end), 1 thread"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--horizon", type=int, default=32)
    ap.add_argument("--dtype", default="fp64", choices=["fp64", "fp32"])
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--cpu-sample", type=int, default=4096)
    ap.add_argument("--latency-ticks", type=int, default=500)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
