"""North-star facade: ``Indy7SQP(urdf, N, weights...).solve(x0, xref) -> (x_traj, u_traj, stats)``.

A thin layer over the same C-ABI call the reference-shaped classes use (``i7_solve`` = ``SQP_OSQP.sqp``,
src/osqp_sqp.py:76-93), batched over independent problems.  Warm start = the previous solution of the same lane
(the reference drivers pass the shifted previous trajectory, src/osqp_mpc.py:46,65-70); the QP solve itself is
direct and carries no hidden state.
"""
import numpy as np

from .engine import Engine, _is_torch
from .model import RobotModel, build_model_from_urdf

try:
    import torch
except Exception:  # pragma: no cover
    torch = None


class Indy7SQP:
    def __init__(self, urdf=None, N=32, dt=0.01, dQ_cost=0.01, R_cost=1e-5, QN_cost=100.0, regularize=True, eps=1.0,
                 dtype="fp64", device="cuda", batch=1, threads=0, cluster=0, qp_mode="exact", limits=False, admm=None):
        if isinstance(urdf, RobotModel):
            self.model = urdf
        else:
            self.model = build_model_from_urdf(urdf) if urdf else build_model_from_urdf()
        if isinstance(device, str):
            if not device.startswith("cuda"):
                raise ValueError("indy7_mpc_b200 runs on CUDA devices only (no CPU fallback)")
            device = int(device.split(":")[1]) if ":" in device else 0
        self.N, self.dt, self.batch, self.dtype = int(N), float(dt), int(batch), dtype
        self.nz = 18 * self.N - 6
        self.engine = Engine(self.model, N=N, dt=dt, dQ_cost=dQ_cost, R_cost=R_cost, QN_cost=QN_cost,
                             regularize=regularize, eps=eps, batch=batch, dtype=dtype, device=device, threads=threads,
                             cluster=cluster, qp_mode=qp_mode, limits=limits, admm=admm)
        self._warm = None
        self._tick_prev = None          # step(): previous solution of every lane, device-resident
        self._tick_bufs = None
        self._bound = None              # step(): buffers and addresses of the bound fast path

    def reset(self):
        self._warm = None
        self._tick_prev = None
        if self._bound is not None:
            self._bound["started"] = False
        self.engine.reset()             # qp_mode 'osqp': also clears the ADMM warm start of every problem slot

    def step(self, x0, xref, exchange=None, index_offset=0):
        """One closed-loop MPC tick of a batch of sampled problems (BASELINE.json configs[3]; the tick of
        gato_controller.py:201-256 with the batch as wrench / start-state hypotheses):

            warm start = previous solution of every lane moved one knot on, x_0 <- x0      (i7_shift; kept on the GPU)
            sqp() of every lane                                                           (i7_solve, one launch)
            per-lane costs + best trajectory                                              (i7_pack_best | `exchange`)

        x0 (B,12) and xref (B,3N | B,6N) are the tick's inputs — host (page-locked tensors are read by the kernels
        themselves) or device; the warm start never leaves the device.  Returns a dict of HOST arrays: 'costs' (B,),
        'best_cost', 'best_index', 'x_traj' (N,12) / 'u_traj' (N-1,6) of the best lane, and 'x_next' (B,12) = the
        state every lane predicts for the next tick.  `exchange`: None = local selection; "p2p" = the peer-memory
        exchange kernel (multi-GPU, peers connected); or a callable `exchange(stats, xu) -> device record` (e.g. pack_best +
        all-gather + reduce_best).  With torch inputs of the engine dtype (page-locked or CUDA) and no callable, the tick
        runs on buffers bound once (see _step_bound)."""
        eng = self.engine
        B = self.batch
        tdt = eng._torch_dtype()
        dev = torch.device("cuda", eng.device)
        if not _is_torch(x0):
            x0 = torch.as_tensor(np.ascontiguousarray(np.asarray(x0, dtype=eng.np_dtype).reshape(B, 12)))
        if not _is_torch(xref):
            xref = torch.as_tensor(np.ascontiguousarray(np.asarray(xref, dtype=eng.np_dtype).reshape(B, -1)))
        fast = (x0.dtype == tdt and xref.dtype == tdt and x0.is_contiguous() and xref.is_contiguous() and
                (x0.is_cuda or x0.is_pinned()) and (xref.is_cuda or xref.is_pinned()) and x0.numel() == 12 * B and
                (exchange is None or exchange == "p2p"))
        if fast:
            return self._step_bound(x0, xref, exchange, index_offset)
        if exchange == "p2p":
            exchange = lambda st, out: eng.p2p_exchange(st, out, index_offset=index_offset)      # noqa: E731
        if self._tick_bufs is None:
            self._tick_bufs = (torch.empty(B + 2 + self.nz, dtype=tdt).pin_memory(), torch.empty(B, 12, dtype=tdt).pin_memory())
        rec_h, x1_h = self._tick_bufs
        if self._tick_prev is None:
            xu = torch.zeros(B, self.nz, dtype=tdt, device=dev)
            xu[:, :12] = x0.to(dev, non_blocking=True)
        else:
            xu, _ = eng.shift(self._tick_prev, steps=1, tail="repeat", x_cur=x0, want_x0=False)
        out, st = eng.solve(xu, x0, xref, device_out=True)
        rec = exchange(st, out) if exchange is not None else eng.pack_best(st, out, index_offset=index_offset)
        n = rec.numel()
        rec_h[:n].copy_(rec, non_blocking=True)
        x1_h.copy_(out[:, 18:30], non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        self._tick_prev = out
        return self._tick_result(rec_h[:n].numpy(), x1_h.numpy(), x0, xref, n)

    def _tick_result(self, r, x_next, x0, xref, n):
        best = r[n - self.nz:]
        full = np.concatenate([best, np.zeros(6, best.dtype)]).reshape(self.N, 18)
        return {'costs': r[:n - 2 - self.nz], 'best_cost': float(r[n - 2 - self.nz]), 'best_index': int(r[n - 1 - self.nz]),
                'x_traj': full[:, :12], 'u_traj': full[:-1, 12:], 'x_next': x_next,
                'h2d_bytes': int((x0.numel() + xref.numel()) * x0.element_size()) if not x0.is_cuda else 0,
                'd2h_bytes': int((n + x_next.size) * r.itemsize)}

    def _step_bound(self, x0, xref, exchange, index_offset):
        """step() with every buffer of the tick bound once: warm start, solution, records and their page-locked host
        mirror are allocated on the first call and their addresses cached.  A tick is three C-ABI calls and ONE copy:
            i7_solve       (x_0 of the warm start <- x0 inside the kernel: I7_OPT_X0_INTO_WARM_START)
            i7_pack_best | i7_p2p_exchange      -> [costs | best_cost, best_index, XU_best]
            i7_shift       the warm start of the NEXT tick and x_next = every lane's predicted next state, at the end of this
                           tick instead of in front of the next solve
            one device -> host copy of [record | x_next]
        (the host issues the solve ~15 us after the call instead of ~80 us through the generic Engine calls — at 512
        problems per GPU that is a fifth of the tick)."""
        from . import _lib
        eng, lib, B, nz = self.engine, self.engine.lib, self.batch, self.nz
        dev = torch.device("cuda", eng.device)
        tb = self._bound
        if tb is None:
            tdt = eng._torch_dtype()
            nrec = B + 2 + nz
            tb = self._bound = {
                "xu": torch.zeros(B, nz, dtype=tdt, device=dev), "out": torch.empty(B, nz, dtype=tdt, device=dev),
                "stats": torch.empty(B, 12, dtype=tdt, device=dev),
                "buf": torch.empty(nrec + 12 * B, dtype=tdt, device=dev),            # [record | x_next]
                "buf_h": torch.empty(nrec + 12 * B, dtype=tdt).pin_memory(),
                "started": False, "nrec": nrec,
            }
            tb["p"] = {k: tb[k].data_ptr() for k in ("xu", "out", "stats", "buf")}
        p = tb["p"]
        stream = torch.cuda.current_stream(dev).cuda_stream
        p_x0, p_ref = x0.data_ptr(), xref.data_ptr()
        gs = xref.numel() // (B * self.N)
        if gs not in (3, 6):
            raise ValueError(f"xref must hold N*3 or N*6 numbers per problem, got {xref.numel() // B} for N={self.N}")
        if self._tick_prev is not None and self._tick_prev is not tb["out"]:
            # the previous tick went through the generic path (other kinds of inputs): take its solution over
            xu_prev, _ = eng.shift(self._tick_prev, steps=1, tail="repeat", want_x0=False)
            tb["xu"].copy_(xu_prev)
            tb["started"] = True
        if not tb["started"]:
            tb["xu"].zero_()
            tb["started"] = True
        lib.i7_set_option(eng._h, _lib.I7_OPT_X0_INTO_WARM_START, 1)        # for this launch only: solve() keeps the
        rc = lib.i7_solve(eng._h, B, p["xu"], p_x0, p_ref, gs, p["out"], p["stats"], stream)     # reference's semantics
        lib.i7_set_option(eng._h, _lib.I7_OPT_X0_INTO_WARM_START, 0)
        eng._check(rc)
        n = 2 + nz if exchange == "p2p" else tb["nrec"]
        es = tb["buf"].element_size()
        if exchange == "p2p":
            eng._check(lib.i7_p2p_exchange(eng._h, B, p["stats"], p["out"], int(index_offset), p["buf"], stream))
        else:
            eng._check(lib.i7_pack_best(eng._h, B, p["stats"], p["out"], int(index_offset), p["buf"], stream))
        # warm start of the next tick (its x_0 is replaced by the measured state inside the next solve) + x_next behind the record
        eng._check(lib.i7_shift(eng._h, B, 1, 2, p["out"], None, None, p["xu"], p["buf"] + n * es, stream))
        m = n + 12 * B
        tb["buf_h"][:m].copy_(tb["buf"][:m], non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        self._tick_prev = tb["out"]
        h = tb["buf_h"].numpy()
        return self._tick_result(h[:n], h[n:m].reshape(B, 12), x0, xref, n)

    def set_external_wrench(self, f_ext):
        self.engine.set_external_wrench(f_ext)

    def _goals(self, xref, B):
        shp = tuple(xref.shape)
        n = int(np.prod(shp))
        per = n // B if n % B == 0 and len(shp) > 1 and shp[0] == B and n // B in (3 * self.N, 6 * self.N) else n
        if per in (3 * self.N, 6 * self.N) and n == per * B:
            return xref.reshape(B, per)
        if n in (3 * self.N, 6 * self.N):          # one reference shared by every problem
            flat = xref.reshape(1, n)
            return flat.expand(B, n) if _is_torch(flat) else np.broadcast_to(flat, (B, n))
        raise ValueError(f"xref must hold N x 3 or N x 6 numbers (optionally with a leading batch dim), got {shp}")

    def solve(self, x0, xref, xu_init=None):
        """x0: (12,) or (B,12); xref: (N,3) | (3N,) | (N,6) | (6N,) optionally batched; returns
        (x_traj (...,N,12), u_traj (...,N-1,6), stats dict).  Inputs are never modified."""
        tens = _is_torch(x0)
        x0a = x0 if tens else np.asarray(x0, float)
        batched = x0a.ndim == 2
        B = x0a.shape[0] if batched else 1
        if x0a.shape[-1] != 12 or B > self.batch:
            raise ValueError(f"x0 must be (12,) or (B<= {self.batch}, 12)")
        x0a = x0a.reshape(B, 12)
        xr = xref if _is_torch(xref) else np.asarray(xref, float)
        goals = self._goals(xr, B)
        if tens and not _is_torch(goals):
            goals = torch.as_tensor(np.ascontiguousarray(goals), device=x0a.device)
        if xu_init is not None:
            xu = xu_init if _is_torch(xu_init) else np.asarray(xu_init, float)
            xu = xu.reshape(B, self.nz)
        elif self._warm is not None and self._warm.shape[0] == B and _is_torch(self._warm) == tens:
            xu = self._warm
        else:
            if tens:
                xu = torch.zeros(B, self.nz, dtype=x0a.dtype, device=x0a.device)
            else:
                xu = np.zeros((B, self.nz))
            xu[:, :12] = x0a
        if tens and goals.is_cuda != x0a.is_cuda:
            goals = goals.to(x0a.device)
        if tens:
            goals = goals.contiguous()
        out, st = self.engine.solve(xu, x0a, goals)
        self._warm = out
        # views into the solver's [x0 u0 x1 u1 ... x_{N-1}] layout (no copy)
        if tens:
            x_traj = torch.as_strided(out, (B, self.N, 12), (self.nz, 18, 1))
            u_traj = torch.as_strided(out, (B, self.N - 1, 6), (self.nz, 18, 1), 12)
        else:
            es = out.itemsize
            x_traj = np.lib.stride_tricks.as_strided(out, (B, self.N, 12), (self.nz * es, 18 * es, es), writeable=False)
            u_traj = np.lib.stride_tricks.as_strided(out[:, 12:], (B, self.N - 1, 6), (self.nz * es, 18 * es, es),
                                                     writeable=False)
        stn = st.cpu().numpy() if _is_torch(st) else st
        stats = {
            'qp_iters': stn[:, 0].astype(int), 'n_linesearch': stn[:, 1].astype(int), 'n_steps': stn[:, 2].astype(int),
            'status': stn[:, 3].astype(int), 'linesearch_alphas': stn[:, 4:6].copy(), 'sqp_stepsizes': stn[:, 6:8].copy(),
            'cost': stn[:, 8].copy(), 'cost_in': stn[:, 9].copy(),
            'xu': out if batched else out[0],
        }
        if not batched:
            x_traj, u_traj = x_traj[0], u_traj[0]
            stats = {k: (v[0] if k != 'xu' else v) for k, v in stats.items()}
        return x_traj, u_traj, stats
