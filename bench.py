#!/usr/bin/env python3
"""Benchmark of the SQP-MPC hot path (BASELINE.json metric: batched SQP-MPC solves/s + p50 single-solve latency).

  python bench.py --gpus N --steps K --warmup W            the B200 arm (one rank per GPU under torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K ...  the reference arm: the reference's CPU loop on the host cores
                                                            (the real src/osqp_sqp.py when pinocchio + osqp import,
                                                            else its C restatement oracle/c with OSQP-style ADMM)

Workload of both arms (BASELINE.json configs[3]): a batch of 4096 independent Indy7 N=32
SQP-MPC problems — sampled start states around the simulator home pose, figure-8 end-effector references at per-lane
offsets, per-lane external-wrench hypotheses — sharded across the ranks (strong scaling: the total batch is fixed).
One "step" = one control tick of ALL lanes: every lane runs one sqp() warm-started from its shifted previous solution,
then the per-problem costs and the best trajectory are selected (N > 1: exchanged between the ranks once per tick).
Between ticks the lanes advance like an MPC loop (i7_shift / i7_goal_window on the device): x0 <- x_1 of the solution,
trajectory moved one knot on, goal window advanced by one sample.

B200 arm, three measurements in one JSON line:
  value   device-resident tick (inputs in HBM when the timed region starts): CUDA events around solve + selection /
          exchange on the launch stream, L2 flushed (256 MB memset) between ticks outside the events, max over ranks
  e2e     the same tick through the public call Indy7SQP.step(x0, xref) with page-locked HOST inputs, wall clock: the
          tick's inputs (measured states, goal windows) are read from host memory, the result (costs, best trajectory,
          the state every lane predicts) is written back to host memory; the warm start stays on the GPU on both
          sides.  'all_trajectories_to_host' = Indy7SQP.solve(x0, xref, xu_init) with every trajectory in and out.
  latency BASELINE.json configs[1]: one problem, fig-8 tracking, closed loop with the rk4 plant
plus 'configs': configs[2] (batch 16/32/64) and configs[4] (N=128, batch 512, fp64 and fp32) with their own CPU figures.
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


def workload_config(args):
    """The `config` object of the JSON line — the same in both arms (the reference arm runs this very workload)."""
    return {"workload": f"batch {args.batch} sampled Indy7 SQP-MPC problems (BASELINE.json configs[3]), N={args.horizon}, dt=0.01, "
                        f"fp64 reference arithmetic, warm-started closed-loop MPC ticks of all lanes, per-lane wrench hypotheses, "
                        f"per-problem costs + best-trajectory selection every tick",
            "horizon": args.horizon, "batch_total": args.batch, "lanes_per_step": args.batch,
            "parallelism": "batch sharded over the ranks (GPU arm) / OpenMP over lanes (CPU arm)",
            "l2": "GPU arm: flushed between timed steps (256 MB memset outside the events)"}


def probe_reference():
    """can the real reference path run on this box?  It needs pinocchio + osqp (absent from
    the image) and the reference's own sources (never shipped to the GPU box)."""
    out = {}
    for mod in ("pinocchio", "osqp"):
        try:
            m = __import__(mod)
            out[mod] = getattr(m, "__version__", "present")
        except Exception:
            out[mod] = "absent"
    src = os.environ.get("I7_REFERENCE_SRC", "")
    out["reference_src"] = src if (src and os.path.exists(os.path.join(src, "osqp_sqp.py"))) else "absent"
    out["runnable"] = all(v != "absent" for v in out.values())
    return out


class CpuLoop:
    """The closed loop of the workload on the host cores: sqp() of every lane per tick, warm-started, then the shift.
    kind 'port' = oracle/c (plain-C restatement of src/osqp_sqp.py + src/osqp_solver.py, OSQP-style ADMM at the
    reference's tolerances, warm-started per lane, OpenMP over lanes); kind 'reference' = the unmodified reference
    classes, one solver object per lane, when pinocchio + osqp + the reference sources are importable."""

    def __init__(self, wl, lanes, N, nthreads, probe=None, limits=None):
        self.wl, self.lanes, self.N, self.nthreads = wl, lanes, N, nthreads
        self.xu, self.x0 = wl["xu"][lanes].copy(), wl["x0"][lanes].copy()
        self.tick = 0
        self.kind = "port"
        if probe and probe.get("runnable") and limits is None:
            self._init_reference(probe["reference_src"])
        else:
            from oracle.cport import COracle
            from oracle.model import load_model
            self.orc = COracle(load_model(), N=N, backend="admm", batch=len(lanes), limits=limits)

    def _init_reference(self, src):            # pragma: no cover - needs pinocchio + osqp
        sys.path.insert(0, src)
        import pinocchio as pin
        from osqp_solver import OSQPSolver
        from osqp_sqp import SQP_OSQP
        urdf = os.path.join(ROOT, "indy7_mpc_b200", "description", "indy7.urdf")
        model = pin.buildModelFromUrdf(urdf)
        model.gravity.linear = np.array([0.0, 0.0, -9.81])
        self.ref = [SQP_OSQP(OSQPSolver(model, N=self.N)) for _ in self.lanes]
        self.kind = "reference"

    def step(self):
        goals = goals_at(self.wl["fig"], self.wl["offsets"][self.lanes], self.tick, self.N)
        t0 = time.perf_counter()
        if self.kind == "port":
            out, _, _ = self.orc.sqp(self.x0, goals, self.xu, f_ext=self.wl["f_ext"][self.lanes], nthreads=self.nthreads)
        else:                                  # pragma: no cover
            out = np.stack([r.sqp(self.x0[i], goals[i], self.xu[i]) for i, r in enumerate(self.ref)])
        dt = time.perf_counter() - t0
        self.xu, self.x0 = shift_host(out, self.N)
        self.xu[:, :12] = self.x0
        self.tick += 1
        return dt


def cpu_baseline(wl, N, seconds, lanes=None, max_ticks=50, warm=2, limits=None):
    """Bounded sample of the SAME workload on the host cores: `warm` untimed ticks (cold -> warm start), then
    warm-started ticks of the given lanes (default: all of them) until `seconds` have been timed."""
    cores = host_cores()
    lanes = np.arange(wl["x0"].shape[0]) if lanes is None else lanes
    probe = probe_reference()
    loop = CpuLoop(wl, lanes, N, cores, probe, limits)
    for _ in range(warm):
        loop.step()
    times = []
    while sum(times) < seconds and len(times) < max_ticks:
        times.append(loop.step())
    el = float(np.sum(times))
    what = ("the unmodified reference classes (pinocchio + osqp)" if loop.kind == "reference" else
            "C restatement of osqp_sqp with OSQP-style ADMM (oracle/c), OpenMP over lanes")
    return {"value": len(lanes) * len(times) / el, "unit": UNIT, "cores": cores if loop.kind == "port" else 1, "kind": loop.kind,
            "sample": f"{len(times)} warm-started ticks x {len(lanes)} lanes of the same workload in {el:.1f} s after "
                      f"{warm} untimed ticks; {what}",
            "probe": {k: probe[k] for k in ("pinocchio", "osqp", "reference_src")}}


def run_reference(args):
    """Reference arm: the reference's CPU loop on the host cores, on the B200 arm's workload and config — ALL lanes of
    the batch per step.  Pinocchio / OSQP are probed at start; without them (this image) the C restatement runs."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    N = args.horizon
    wl = make_workload(args.batch, N)
    cores = host_cores()
    probe = probe_reference()
    loop = CpuLoop(wl, np.arange(args.batch), N, cores, probe)
    times = []
    for t in range(args.warmup + args.steps):
        dt = loop.step()
        if t >= args.warmup:
            times.append(dt)
    total = float(np.sum(times))
    value = args.batch * len(times) / total
    sample = f"all {args.batch} lanes per step, {len(times)} warm-started closed-loop ticks after {args.warmup} warm-up ticks"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores if loop.kind == "port" else 1, "kind": loop.kind,
                         "sample": sample, "probe": {k: probe[k] for k in ("pinocchio", "osqp", "reference_src")}},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------- B200 arm
def device_loop(eng, wl, N, W, K, dtype, dev, exchange=None, flush=None, check=None, mode=None, index_offset=0):
    """W + K device-resident closed-loop ticks (goal window -> solve -> selection/exchange -> shift, all library kernels on
    torch's current stream).  mode 'local' | 'p2p' | 'fused': the bound-buffer runner `DeviceTick` (one C-ABI call per
    stage, ~50 us of host time per tick); mode None: the generic Engine calls with `exchange(stats, out)` (NCCL path).
    Returns per-tick (step_ms, kernel_ms) of the K timed ticks, the last stats."""
    import torch
    from indy7_mpc_b200 import DeviceTick
    tdt = torch.float64 if dtype == "fp64" else torch.float32
    B = wl["x0"].shape[0]
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(K)]
    if mode is not None:
        tk = DeviceTick(eng, wl["fig"], wl["offsets"], wl["xu"], wl["x0"], mode=mode, index_offset=index_offset)
        for t in range(W + K):
            tk.goal(t)
            i = t - W
            if i >= 0:
                if flush is not None:
                    flush.zero_()                        # L2 flush, outside the timed events
                ev[i][0].record()
            tk.solve()
            if i >= 0:
                ev[i][1].record()                        # the dominant kernel on its own (fused: incl. its epilogue)
            tk.select()
            if i >= 0:
                ev[i][2].record()
            elif check is not None:
                check(tk.stats, tk.out, tk.best)
            tk.shift()
        st = tk.stats
    else:
        fig_d = torch.tensor(wl["fig"], dtype=tdt, device=dev)
        off_d = torch.tensor(wl["offsets"], dtype=torch.int32, device=dev)
        xu = torch.tensor(wl["xu"], dtype=tdt, device=dev)
        x0 = torch.tensor(wl["x0"], dtype=tdt, device=dev)
        st = None
        for t in range(W + K):
            goals = eng.goal_window(fig_d, off_d, tick=t, wrap=True, batch=B)
            i = t - W
            if i >= 0:
                if flush is not None:
                    flush.zero_()
                ev[i][0].record()
            out, st = eng.solve(xu, x0, goals)
            if i >= 0:
                ev[i][1].record()
            best = exchange(st, out) if exchange is not None else eng.pack_best(st, out)
            if i >= 0:
                ev[i][2].record()
            elif check is not None:
                check(st, out, best)
            xu, x0 = eng.shift(out, steps=1, tail="repeat")
    torch.cuda.synchronize()
    step_ms = np.array([e[0].elapsed_time(e[2]) for e in ev])
    kern_ms = np.array([e[0].elapsed_time(e[1]) for e in ev])
    return step_ms, kern_ms, st


def fp_roofline(eng, st_host, N, B, kern_ms, dtype, fma_peak):
    """Roofline of the solve kernel against the FP64 / FP32 FMA pipe (arithmetic intensity
    ~215 flop/B, the bound is the CUDA-core FMA pipe): algorithmic flops of the measured (n_qp, n_try) per lane."""
    flops = algorithmic_flops(st_host, N)
    ach = flops * B / (kern_ms * 1e-3) / 1e12
    return {"achieved": ach, "peak": fma_peak, "unit": "TFLOP/s", "frac": ach / fma_peak if fma_peak else None,
            "algorithmic_flops_per_solve": flops}


def controller_tick_latency(eng, wl, B, N, ticks):
    """The reference controller's own tick on configs[2] (GATO_Controller.joint_callback, gato_controller.py:201-256: all
    lanes share state and reference, selection by one-step prediction error, resampling around the selected lane, its
    trajectory handed to every lane): wall clock per `ControllerTick.tick` call — host buffers in, control out."""
    import time
    from indy7_mpc_b200 import ControllerTick
    np.random.seed(42)                                   # gato_controller.py:308
    ref = np.zeros((wl["fig"].shape[0], 6))
    ref[:, :3] = wl["fig"][:, :3]
    ct = ControllerTick(eng, ref.reshape(-1), B, f_ext_std=2.0 if B >= 32 else 20.0, f_ext_resample_std=0.2 if B >= 32 else 1.0)
    x = np.asarray(wl["x0"][0], float).copy()
    t_us = []
    for t in range(ticks + 10):
        t0 = time.perf_counter()
        u, best = ct.tick(x, 0.01)
        t_us.append((time.perf_counter() - t0) * 1e6)
        x = ct.XU_best[18:30].copy()                     # plant = the selected lane's own prediction (timing run)
    t_us = np.array(t_us[10:])
    eng.set_external_wrench(wl["f_ext"])
    return {"host_wall_us": {"p50": float(np.percentile(t_us, 50)), "p99": float(np.percentile(t_us, 99)), "mean": float(t_us.mean())},
            "ticks": int(ticks), "launches_per_tick": 5, "finite": bool(np.all(np.isfinite(ct.XU_best))),
            "what": "ControllerTick.tick: goal window, solve, selection by prediction error, wrench resampling, hand-over "
                    "of the best trajectory; measured state / elapsed time / normal draws from the host, control back on the host"}


def extra_configs(args, local, fma_peaks):
    """BASELINE.json configs[2] (batch 16/32/64, N=32) and configs[4] (N=128, batch 512, fp64 vs fp32) on one GPU:
    device-resident closed-loop ticks, L2 flushed, with the FMA-pipe fraction and the CPU figure of the same lanes."""
    import torch
    from indy7_mpc_b200 import Engine
    dev = torch.device("cuda", local)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
    # the last two run the QP of every pass the way the reference does (qp_mode 'osqp': warm-started ADMM at eps 1e-3),
    # i.e. the SAME algorithm as the CPU port beside them; the very last adds the URDF joint limits as inequality rows
    cases = [("configs[2]", 16, 32, "fp64", "exact", False), ("configs[2]", 32, 32, "fp64", "exact", False),
             ("configs[2]", 64, 32, "fp64", "exact", False),
             ("configs[4]", 512, 128, "fp64", "exact", False), ("configs[4]", 512, 128, "fp32", "exact", False),
             ("configs[3] (per-GPU share at 8 GPUs) in the reference's QP regime", 512, 32, "fp64", "osqp", False),
             ("configs[3] (per-GPU share at 8 GPUs) in the reference's QP regime + joint limits (SURVEY 8 f-4)", 512, 32, "fp64", "osqp", True)]
    out, cpu_cache, first = [], {}, {}
    for name, B, N, dtype, qp_mode, limits in cases:
        wl = make_workload(B, N)
        eng = Engine(N=N, batch=B, dtype=dtype, device=local, qp_mode=qp_mode, limits=limits)
        eng.set_external_wrench(wl["f_ext"])
        step_ms, kern_ms, st = device_loop(eng, wl, N, 5, args.config_ticks, dtype, dev, flush=flush, mode="local")
        sth = st.double().cpu().numpy()
        if dtype not in fma_peaks:
            fma_peaks[dtype] = eng.measure_fma_peak()
        thr, ctas, smem = eng.kernel_form()
        rec = {"config": name, "batch": B, "horizon": N, "dtype": "f64" if dtype == "fp64" else "f32",
               "ms_per_tick": float(np.median(step_ms)), "solves_per_s": float(B / np.mean(step_ms) * 1e3),
               "kernel": f"{thr} threads x {ctas} CTA(s) per problem, {smem} B shared memory",
               "fp_pipe": fp_roofline(eng, sth, N, B, float(kern_ms.mean()), dtype, fma_peaks[dtype]),
               "mean_qp_iters": float(sth[:, 0].mean())}
        if qp_mode == "osqp":
            rec["qp_mode"] = "osqp: OSQP-style ADMM, eps_abs = eps_rel = 1e-3, warm-started per lane" + (", joint position / velocity / effort limits as inequality rows" if limits else "")
            rec["mean_admm_iters_per_solve"] = float(sth[:, 10:12].sum(axis=1).mean())
            rec["fp_pipe"]["note"] = "the flop model counts the direct-method floor of the QP; ADMM iterations are overhead by the survey's definition"
        if N == 128 and qp_mode == "exact":                         # fp32 against fp64 on identical inputs (north_star: 1e-4)
            o, _ = eng.solve(torch.tensor(wl["xu"], dtype=st.dtype, device=dev), torch.tensor(wl["x0"], dtype=st.dtype, device=dev),
                             torch.tensor(goals_at(wl["fig"], wl["offsets"], 0, N), dtype=st.dtype, device=dev))
            first[dtype] = o.double().cpu().numpy()
            if "fp64" in first and dtype == "fp32":
                rec["fp32_vs_fp64_rel"] = float(np.abs(first["fp32"] - first["fp64"]).max() / np.abs(first["fp64"]).max())
        if name == "configs[2]":
            rec["controller_tick"] = controller_tick_latency(eng, wl, B, N, max(args.config_ticks, 20) * 5)
        if not args.no_cpu_baseline:
            key = (B, N, limits)
            if key not in cpu_cache:
                lim = None
                if limits:
                    m = eng.model
                    lim = (m.lowerPositionLimit, m.upperPositionLimit, m.velocityLimit, m.effortLimit)
                cpu_cache[key] = cpu_baseline(wl, N, args.config_cpu_seconds, max_ticks=20, limits=lim)
            rec["cpu_baseline"] = cpu_cache[key]
        out.append(rec)
        del eng
    return out


def run_b200(args):
    import torch
    import torch.distributed as dist
    from indy7_mpc_b200 import Engine, Indy7SQP
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

    gathered = torch.empty(world * (B + 2 + nz), dtype=tdt, device=dev) if world > 1 else None
    # exchange of the per-rank records: ONE kernel over peer memory (NVLink stores + flags), fused with the packing and
    # the selection; I7_P2P=0 selects the NCCL path (pack kernel, all_gather_into_tensor, reduce kernel)
    use_p2p = world > 1 and os.environ.get("I7_P2P", "1") != "0"
    if use_p2p:
        use_p2p = i7d.connect_peers(eng, rank, world)      # False (on every rank) -> the NCCL path

    def nccl_exchange(e):
        def f(st, out):
            packed = e.pack_best(st, out, index_offset=rank * B)
            return i7d.gather_best(packed, B, world, e.reduce_best, gathered)[0]
        return f

    def make_exchange(e, p2p):
        if world == 1:
            return None
        if p2p:
            return lambda st, out: e.p2p_exchange(st, out, index_offset=rank * B)
        return nccl_exchange(e)

    # selection (+ exchange) as the epilogue of the solve kernel — one launch per tick instead of two — whenever the
    # one-CTA-per-problem form runs and the peer buffers are mapped (a single GPU maps its own); I7_FUSED=0: two launches
    # OFF by default: measured on the same box it loses — the selection scan + fences become a serial tail of the big kernel
    # (1 GPU, 4096 problems: 1.975 vs 1.954 ms per tick; 2 GPUs: 1.038 vs 1.030 ms); I7_FUSED=1 selects it
    fused_env = os.environ.get("I7_FUSED", "0")
    fused = fused_env != "0" and eng.kernel_form()[1] == 1 and (use_p2p or world == 1)
    if fused and world == 1:
        eng.p2p_connect([eng.p2p_create(1, 0)])

    def check(st, out, best):          # warm-up ticks double as a check of the peer-memory exchange against NCCL
        if world == 1 and fused:
            if not torch.equal(best, eng.pack_best(st, out, index_offset=rank * B)[B:]):
                raise RuntimeError("fused selection disagrees with i7_pack_best")
        if use_p2p:
            ref = nccl_exchange(eng)(st, out)
            if not torch.equal(best, ref):
                raise RuntimeError("peer-memory exchange disagrees with the NCCL all-gather path")

    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)

    # ---- device-resident arm: W warm-up ticks, then K timed ticks ----
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = ClockSampler(local) if rank == 0 else None
    launches0 = eng.launch_count()
    t_wall0 = time.time()
    if fused:
        eng.exchange_wait_ns(reset=True)
    loop_mode = "fused" if fused else ("p2p" if use_p2p else ("local" if world == 1 else None))
    step_ms, kern_ms, last_st = device_loop(eng, wl, N, W, K, args.dtype, dev, make_exchange(eng, use_p2p), flush, check,
                                            mode=loop_mode, index_offset=rank * B)
    wait_ms = eng.exchange_wait_ns() / 1e6 / (W + K) if fused else 0.0      # per tick, inside the fused kernel
    if use_p2p or fused:
        eng.p2p_status()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t_wall1 = time.time()
    launches = (eng.launch_count() - launches0) * K // (W + K)      # the timed ticks' share (every tick launches alike)
    total_ms = torch.tensor([step_ms.sum()], dtype=torch.float64, device=dev)
    per_rank = torch.tensor([kern_ms.mean(), (step_ms - kern_ms).mean(), wait_ms], dtype=torch.float64, device=dev)
    per_rank_all = [torch.zeros_like(per_rank) for _ in range(world)]
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
        dist.all_gather(per_rank_all, per_rank)
    else:
        per_rank_all = [per_rank]
    total_ms = float(total_ms.item())
    st_host = last_st.double().cpu().numpy()

    # ---- end-to-end arm through the public API with HOST buffers (H2D + D2H inside the timed region) ----
    solver = Indy7SQP(N=N, batch=B, dtype=args.dtype, device=f"cuda:{local}", threads=args.threads)
    solver.set_external_wrench(wl["f_ext"])
    npdt = np.float64 if args.dtype == "fp64" else np.float32
    pin = lambda shape: torch.empty(shape, dtype=tdt).pin_memory()
    p_xu, p_x0, p_goals = pin((B, nz)), pin((B, 12)), pin((B, 3 * N))
    e2e_p2p = use_p2p and i7d.connect_peers(solver.engine, rank, world)
    e2e_exchange = "p2p" if e2e_p2p else make_exchange(solver.engine, e2e_p2p)     # "p2p": the bound fast path of step()

    def e2e_loop(full):
        """W + K closed-loop ticks from HOST buffers, wall clock per tick (max over ranks of the sum).
        full = False: Indy7SQP.step — inputs x0 + goal windows from page-locked host memory, results (costs, best
        trajectory, next states) to host memory, warm start resident on the GPU.
        full = True: Indy7SQP.solve(x0, xref, xu_init) — every warm start in, every trajectory out, selection on the host."""
        solver.reset()
        hxu, hx0 = wl["xu"].copy(), wl["x0"].copy()
        times, h2d, d2h = [], 0, 0
        for t in range(W + K):
            p_goals.copy_(torch.from_numpy(goals_at(wl["fig"], wl["offsets"], t, N).astype(npdt)))
            p_x0.copy_(torch.from_numpy(hx0.astype(npdt)))
            if full:
                p_xu.copy_(torch.from_numpy(hxu.astype(npdt)))
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            if not full:
                res = solver.step(p_x0, p_goals, exchange=e2e_exchange, index_offset=rank * B)
                dt_ = time.perf_counter() - t0
                hx0 = res["x_next"].astype(float)
                h2d, d2h = res["h2d_bytes"], res["d2h_bytes"]
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
                hxu, hx0 = shift_host(stats["xu"].numpy().astype(float), N)
                hxu[:, :12] = hx0
                h2d, d2h = int(B * (nz + 12 + 3 * N) * esize), int(B * (nz + 12) * esize)
            if t >= W:
                times.append(dt_)
        total = torch.tensor([float(np.sum(times))], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(total, op=dist.ReduceOp.MAX)
        return float(total.item()), h2d, d2h

    e2e_total, e2e_h2d, e2e_d2h = e2e_loop(False)
    e2e_full_total, full_h2d, full_d2h = e2e_loop(True)
    if e2e_p2p:
        solver.engine.p2p_status()
    clk = clocks.stop(t_wall0, time.time()) if clocks else None

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- single-problem latency (BASELINE.json configs[1]): fig-8 tracking, B = 1, warm-started closed loop ----
    lat = single_problem_latency(args, local)

    # ---- roofline of the dominant kernel: the FMA pipe of the arithmetic type ----
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
    kern_avg_s = float(kern_ms.mean()) * 1e-3
    bytes_launch = algorithmic_bytes(N, esize) * B
    fma_peaks = {args.dtype: eng.measure_fma_peak()}
    traffic, traffic_src = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        key = f"sqp_kernel_B{B}_N{N}_{args.dtype}"
        if key in tj:
            traffic, traffic_src = tj[key], tj.get("source", "profiles/traffic.json")
        elif f"sqp_kernel_B4096_N{N}_{args.dtype}" in tj:      # per-launch, scaled by problems
            traffic, traffic_src = tj[f"sqp_kernel_B4096_N{N}_{args.dtype}"] * B // 4096, tj.get("source", "profiles/traffic.json") + " (scaled by problems)"
    except Exception:
        pass
    thr, ctas, smem = eng.kernel_form()
    fp = fp_roofline(eng, st_host, N, B, float(kern_ms.mean()), args.dtype, fma_peaks[args.dtype])
    roofline = {
        "bound": "fp64_fma" if args.dtype == "fp64" else "fp32_fma", "achieved": fp["achieved"], "peak": fp["peak"],
        "unit": "TFLOP/s", "frac": fp["frac"], "traffic": traffic, "traffic_source": traffic_src,
        "peak_source": "i7_measure_fma_peak: dependent-chain-free FMA loop on every SM of this GPU, this run "
                       "(MEASURED_PEAKS.json holds no FP64 / FP32 CUDA-core figure)",
        "kernel": f"sqp_kernel<{args.dtype}>: {thr} threads x {ctas} CTA(s) per problem, {smem} B shared memory",
        "kernel_ms": float(kern_ms.mean()), "algorithmic_flops_per_solve": fp["algorithmic_flops_per_solve"],
        "why": "arithmetic intensity ~215 flop/B (SURVEY.md section 8d): the FMA pipe of the CUDA cores bounds the kernel, "
               "neither HBM nor the tensor cores (6x6 / 12x12 fp64 blocks with data-dependent structure)",
        "hbm": {"achieved": bytes_launch / kern_avg_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                "frac": bytes_launch / kern_avg_s / 1e9 / hbm_peak, "peak_source": peak_src,
                "algorithmic_bytes_per_solve": algorithmic_bytes(N, esize)},
    }
    try:        # shared-memory pipe (SURVEY.md section 8d asks for it beside the FMA pipe): wavefronts counted by ncu, time live
        wf = tj.get(f"sqp_kernel_B{B}_N{N}_{args.dtype}_smem_wavefronts")
        if wf:
            props = torch.cuda.get_device_properties(dev)
            sm_hz = props.clock_rate * 1e3 if hasattr(props, "clock_rate") else 1.965e9
            per_clk = wf / (kern_avg_s * sm_hz * props.multi_processor_count)
            roofline["smem"] = {"achieved": per_clk, "peak": 1.0, "unit": "wavefronts (128 B) per cycle per SM", "frac": per_clk,
                                "wavefronts_per_launch": wf, "bank_conflict_share": tj.get(f"sqp_kernel_B{B}_N{N}_{args.dtype}_smem_conflicts", 0) / wf,
                                "source": "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum of the ncu capture named in profiles/traffic.json; "
                                          "duration measured in this run, SM clock = the device's maximum"}
    except Exception:
        pass

    cpu = None
    if not args.no_cpu_baseline:
        cpu = cpu_baseline(wl_all, N, args.cpu_seconds)
    configs = extra_configs(args, local, fma_peaks) if (world == 1 and args.config_ticks > 0) else None

    value = args.batch * K / (total_ms * 1e-3)
    run_cfg = {"batch_per_gpu": B, "parallelism": f"batch-shard x{world}",
               "exchange": (("none (one GPU); selection in the epilogue of the solve kernel" if fused else "none (one GPU)") if world == 1 else
                            ("epilogue of the solve kernel: the CTA that finishes the last problem packs, stores the record into "
                             "every peer over NVLink, waits for the ranks and selects (no second launch)" if fused else
                             ("one peer-memory exchange kernel per tick (NVLink stores, fused pack + gather + selection)"
                              if use_p2p else "pack kernel + one NCCL all-gather + reduce kernel per tick")))}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64" if args.dtype == "fp64" else "f32", "data": "synthetic",
        "config": workload_config(args), "run_config": run_cfg,
        "e2e": {"value": args.batch * K / e2e_total, "unit": UNIT,
                "h2d_bytes_per_step": e2e_h2d, "d2h_bytes_per_step": e2e_d2h,
                "ms_per_step": 1e3 * e2e_total / K,
                "api": "Indy7SQP.step(x0, xref) per tick, wall clock: x0 and the goal windows of all lanes are read from "
                       "page-locked host memory by the kernels (zero-copy), shift + solve + selection (N > 1: exchange) on "
                       "the GPU, [costs | best_cost, best_index, XU_best] and the next state of every lane written back to "
                       "host memory inside the timed call; the warm start of every lane stays on the GPU on both sides",
                "all_trajectories_to_host": {
                    "value": args.batch * K / e2e_full_total, "ms_per_step": 1e3 * e2e_full_total / K,
                    "h2d_bytes_per_step": full_h2d, "d2h_bytes_per_step": full_d2h,
                    "api": "Indy7SQP.solve(x0, xref, xu_init) with page-locked host tensors: every warm start read from and every "
                           "trajectory + stats row written to host memory, selection on the host (N > 1: + NCCL all-gather of the records)"}},
        "gpu_launches": int(launches),
        "clocks": clk,
        "roofline": roofline,
        "cpu_baseline": cpu,
        "latency": lat,
        "configs": configs,
        "per_rank": {"kernel_ms": [float(p[0]) for p in per_rank_all], "selection_exchange_ms": [float(p[1]) for p in per_rank_all],
                     "exchange_wait_ms": [float(p[2]) for p in per_rank_all],
                     "what": ("fused: kernel_ms = solve + selection + exchange in one launch, exchange_wait_ms = the part of it its "
                              "epilogue spent waiting for the other ranks" if fused else
                              "kernel_ms = solve kernel, selection_exchange_ms = the launch(es) after it")},
        "wall_s_timed_loop": t_wall1 - t_wall0,
        "stats_last_tick": {"mean_qp_iters": float(st_host[:, 0].mean()), "linesearch_fail_frac": float((st_host[:, 3] == 1).mean()),
                            "nonfinite": int((st_host[:, 3] == 2).sum())},
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


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
    for t in range(n_ticks + 20):        # every step of the tick is a library kernel: window, solve, plant, shift
        goals = eng.goal_window(figd, None, tick=t, wrap=True, batch=1)
        out, st = eng.solve(xud, xd, goals)
        if t >= 20:
            res_us.append(eng.last_solve_time_us())
        xd = eng.sim_forward(xd, out[:, 12:18].contiguous(), 0.01)
        xud, _ = eng.shift(out, steps=1, tail="repeat", x_cur=xd, want_x0=False)
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
    ap.add_argument("--config-ticks", type=int, default=20, help="timed ticks per extra config (0: skip configs[2], configs[4])")
    ap.add_argument("--config-cpu-seconds", type=float, default=3.0)
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
