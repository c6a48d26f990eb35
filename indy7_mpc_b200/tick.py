"""Device-resident closed-loop MPC tick of a batch of sampled problems with every buffer bound ONCE.

The generic `Engine` calls accept any mix of numpy / torch / host / device arguments and allocate their outputs; a control
loop that runs thousands of ticks on the same buffers pays that flexibility on the host every tick (~175 us for the
four calls of a tick, against ~310 us of GPU work at 512 problems — little head-room when eight ranks share a host).
`DeviceTick` preallocates the buffers, caches their addresses and issues exactly one C-ABI call per stage:

    goal(t)    goal window of every lane out of one reference trajectory      i7_goal_window   (gato_controller.py:210-218)
    solve()    sqp() of every lane, warm-started                              i7_solve         (src/osqp_sqp.py:76-93)
    select()   per-lane costs + best trajectory: local selection, peer-memory exchange, or both fused into solve()
                                                                              i7_pack_best | i7_p2p_exchange | i7_solve_select
    shift()    warm start and x_0 of the next tick                            i7_shift         (src/osqp_mpc.py:65-70)

Results stay on the device: `out` (B, 18N-6), `stats` (B, 12), `best` ([best_cost, global index, XU_best]), `packed`
(local selection: [costs | best...]).  Everything runs on torch's current stream of the engine's device.
"""
import torch

from . import _lib


class DeviceTick:
    def __init__(self, engine, traj, offsets, xu0, x0, mode="local", index_offset=0, wrap=True):
        """mode: 'local' (i7_pack_best), 'p2p' (i7_p2p_exchange; needs connected peers) or 'fused' (i7_solve_select)."""
        if mode not in ("local", "p2p", "fused"):
            raise ValueError("mode must be 'local', 'p2p' or 'fused'")
        e = self.engine = engine
        dev = torch.device("cuda", e.device)
        tdt = e._torch_dtype()
        B, N, nz = int(xu0.shape[0]), e.N, e.nz
        if B > e.batch:
            raise ValueError(f"batch {B} exceeds the engine's maximum batch {e.batch}")
        self.B, self.mode, self.index_offset, self.wrap = B, mode, int(index_offset), int(bool(wrap))
        as_dev = lambda a, dt: torch.as_tensor(a).to(device=dev, dtype=dt).contiguous()
        self.traj = as_dev(traj, tdt)
        self.traj_len = int(self.traj.shape[0])
        self.traj_stride = self.traj.numel() // self.traj_len
        self.offsets = as_dev(offsets, torch.int32)
        self.xu = as_dev(xu0, tdt).reshape(B, nz).clone()
        self.x0 = as_dev(x0, tdt).reshape(B, 12).clone()
        self.goals = torch.empty(B, 3 * N, dtype=tdt, device=dev)
        self.out = torch.empty(B, nz, dtype=tdt, device=dev)
        self.stats = torch.empty(B, _lib.I7_STATS_LEN, dtype=tdt, device=dev)
        self.packed = torch.empty(B + 2 + nz, dtype=tdt, device=dev)
        self.best = self.packed[B:] if mode == "local" else torch.empty(2 + nz, dtype=tdt, device=dev)
        self._p = {k: getattr(self, k).data_ptr() for k in ("traj", "offsets", "xu", "x0", "goals", "out", "stats", "packed", "best")}
        self._dev = dev

    def _stream(self):
        return torch.cuda.current_stream(self._dev).cuda_stream

    def goal(self, tick):
        e, p = self.engine, self._p
        e._check(e.lib.i7_goal_window(e._h, self.B, p["traj"], self.traj_len, self.traj_stride, p["offsets"], int(tick), self.wrap, 3,
                                      p["goals"], self._stream()))

    def solve(self):
        e, p = self.engine, self._p
        if self.mode == "fused":
            e._check(e.lib.i7_solve_select(e._h, self.B, p["xu"], p["x0"], p["goals"], 3, p["out"], p["stats"], self.index_offset,
                                           p["best"], self._stream()))
        else:
            e._check(e.lib.i7_solve(e._h, self.B, p["xu"], p["x0"], p["goals"], 3, p["out"], p["stats"], self._stream()))

    def select(self):
        e, p = self.engine, self._p
        if self.mode == "local":
            e._check(e.lib.i7_pack_best(e._h, self.B, p["stats"], p["out"], self.index_offset, p["packed"], self._stream()))
        elif self.mode == "p2p":
            e._check(e.lib.i7_p2p_exchange(e._h, self.B, p["stats"], p["out"], self.index_offset, p["best"], self._stream()))

    def shift(self):
        e, p = self.engine, self._p
        e._check(e.lib.i7_shift(e._h, self.B, 1, 2, p["out"], None, None, p["xu"], p["x0"], self._stream()))

    def run(self, tick):
        """One whole tick; results in `out`, `stats`, `best`."""
        self.goal(tick)
        self.solve()
        self.select()
        self.shift()

    def capture(self):
        """solve + select + shift recorded ONCE into a CUDA graph on torch's current stream (the library's calls on
        device buffers are capturable: launches and device-to-device copies only).  Returns `replay(tick)`: the goal
        window of the tick (its tick number is a launch argument, so it stays an ordinary launch) followed by one graph
        launch — bit-identical to `run(tick)`, a few microseconds less host time per tick.  Run at least one eager tick
        first (first-use set-up of a kernel must not fall inside the capture)."""
        stream = torch.cuda.current_stream(self._dev)
        stream.synchronize()
        keep = tuple(t.clone() for t in (self.xu, self.x0, self.out, self.stats, self.packed))
        best_keep = self.best.clone()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            self.solve()
            self.select()
            self.shift()
        # capturing executes nothing, but keep the state exactly as it was in case a driver version replays while instantiating
        for dst, src in zip((self.xu, self.x0, self.out, self.stats, self.packed), keep):
            dst.copy_(src)
        self.best.copy_(best_keep)
        self._graph = graph

        def replay(tick):
            self.goal(tick)
            graph.replay()
        return replay


class ControllerTick:
    """`GATO_Controller.joint_callback` (gato_controller.py:201-256) without the ROS plumbing: B wrench hypotheses solve
    the same tracking problem; the hypothesis that best predicted the measured state over the last step is selected, the
    wrench batch is resampled around it, its first control is returned and its trajectory becomes every lane's warm
    start.  One tick is a chain of five launches on one stream, every buffer bound once:

        i7_goal_window     ee_pos_traj_batch[:, :] = ref_traj[6 offset : 6 (offset + N)]          (:214-216)
        i7_solve           XU_batch[:, :12] = x_curr inside the kernel, sqp() of every lane        (:218,224)
        i7_find_best_idx   plant step of every lane from (x_last, u_last), error against x_curr    (:225, :109-118)
        i7_resample_wrench re-centre the wrench batch on the selected lane (lane read on the device; :226, :120-129)
        i7_broadcast_best  XU_best = XU_new[best]; XU_batch[:, :] = XU_best                        (:227,249)

    and ONE device -> host copy of [best index, error | XU_best].  The host supplies what the reference takes from the
    host too: the measured state, the elapsed time and np.random's normal draws (the seeded stream is the parity).
    State the reference keeps (`ref_traj_offset`, `x_last`, `u_last`, `f_ext_batch`) is kept here under the same names.
    """

    def __init__(self, engine, ref_traj, batch_size, f_ext_std=1.0, f_ext_resample_std=0.0, decay=0.97):
        e = self.engine = engine
        if e.batch < batch_size:
            raise ValueError(f"batch {batch_size} exceeds the engine's maximum batch {e.batch}")
        import numpy as np
        self.np = np
        dev = self._dev = torch.device("cuda", e.device)
        tdt = e._torch_dtype()
        B, N, nz = int(batch_size), e.N, e.nz
        self.B, self.N, self.dt, self.decay = B, N, float(e.dt), float(decay)
        # gato_controller.py:77-83: the initial hypotheses (np.random stream of the caller)
        if f_ext_std != 0.0:
            f0 = np.random.normal(0, f_ext_std, (B, 6))
            f0[:, 3:] = 0.0
            f0[0] = 0.0
        else:
            f0 = np.zeros((B, 6))
        self.resample_f_ext = f_ext_resample_std > 0.0
        self.f_ext_resample_std = float(f_ext_resample_std)
        e.set_external_wrench(f0)
        traj = np.asarray(ref_traj, dtype=np.float64).reshape(-1, 6)
        self.traj = torch.as_tensor(traj).to(device=dev, dtype=tdt).contiguous()
        self.traj_len = int(traj.shape[0])
        self.ref_traj_offset = 0.0
        self.x_last = self.u_last = None
        self.xu = torch.zeros(B, nz, dtype=tdt, device=dev)                 # XU_batch
        self.out = torch.empty(B, nz, dtype=tdt, device=dev)
        self.stats = torch.empty(B, _lib.I7_STATS_LEN, dtype=tdt, device=dev)
        self.goals = torch.empty(B, 6 * N, dtype=tdt, device=dev)           # ee_pos_traj_batch
        self.f_dev = torch.as_tensor(f0).to(dev)                            # f_ext_batch: float64 whatever the engine dtype
        self.rec = torch.empty(2 + nz, dtype=tdt, device=dev)               # [best index, error | XU_best]
        self.rec_h = torch.empty(2 + nz, dtype=tdt).pin_memory()
        self.xs_batch = torch.empty(B, 12, dtype=tdt).pin_memory()          # xs_batch (read in place by the solve kernel)
        self.sel_in = torch.empty(30, dtype=tdt).pin_memory()               # [x_last 12 | u_last 6 | x_curr 12]
        self.noise_h = torch.empty(B, 6, dtype=torch.float64).pin_memory()
        self._p = {k: getattr(self, k).data_ptr() for k in ("traj", "xu", "out", "stats", "goals", "f_dev", "rec", "xs_batch",
                                                           "sel_in", "noise_h")}
        self.best_idx = None
        self.XU_best = None

    @property
    def f_ext_batch(self):
        """The wrench hypotheses (B, 6), fetched from the device on demand (the tick itself never needs them on the host)."""
        return self.f_dev.cpu().numpy()

    def tick(self, x_curr, elapsed_time):
        """One control tick.  Returns (u_curr (6,), best_idx)."""
        np, e, p, B = self.np, self.engine, self._p, self.B
        lib, h = e.lib, e._h
        es = self.rec.element_size()
        x_curr = np.asarray(x_curr, dtype=np.float64).reshape(12)
        if self.x_last is None:
            self.x_last, self.u_last = x_curr, np.zeros(6)
        stream = torch.cuda.current_stream(self._dev).cuda_stream
        self.ref_traj_offset += elapsed_time / self.dt
        offset = int(self.ref_traj_offset)
        e._check(lib.i7_goal_window(h, B, p["traj"], self.traj_len, 6, None, offset, 0, 6, p["goals"], stream))
        xs = self.xs_batch.numpy()
        xs[:, :] = x_curr
        sel = self.sel_in.numpy()
        sel[:12], sel[12:18], sel[18:] = self.x_last, self.u_last, x_curr
        if e.qp_mode == "osqp":           # solver.reset(), reset_rho(), resetLambda() (:221-223): the warm start of the ADMM; no state in the exact mode
            e.reset(); e.reset_rho(); e.reset_lambda()
        lib.i7_set_option(h, _lib.I7_OPT_X0_INTO_WARM_START, 1)
        rc = lib.i7_solve(h, B, p["xu"], p["xs_batch"], p["goals"], 6, p["out"], p["stats"], stream)
        lib.i7_set_option(h, _lib.I7_OPT_X0_INTO_WARM_START, 0)
        e._check(rc)
        e._check(lib.i7_find_best_idx(h, B, p["sel_in"], p["sel_in"] + 12 * es, p["sel_in"] + 18 * es, float(elapsed_time), 1,
                                      None, None, p["rec"], stream))
        if self.resample_f_ext:
            self.noise_h.numpy()[:, :] = np.random.normal(0, self.f_ext_resample_std, (B, 6))
            e._check(lib.i7_resample_wrench(h, B, -1, p["rec"], p["f_dev"], p["noise_h"], _lib.I7_KEEP_BEST, _lib.I7_RESAMPLE_APPLY,
                                            self.decay, p["f_dev"], stream))
        e._check(lib.i7_broadcast_best(h, B, p["out"], -1, p["rec"], p["xu"], p["rec"] + 2 * es, stream))
        self.rec_h.copy_(self.rec, non_blocking=True)
        torch.cuda.current_stream(self._dev).synchronize()
        r = self.rec_h.numpy()
        self.best_idx = int(r[0])
        self.XU_best = r[2:].astype(np.float64)
        u_curr = self.XU_best[12:18].copy()
        self.x_last, self.u_last = x_curr, u_curr
        return u_curr, self.best_idx
