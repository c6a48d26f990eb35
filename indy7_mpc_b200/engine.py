"""Thin Python owner of one ``i7_handle`` (C ABI in include/indy7_b200.h).

Accepts numpy arrays (host) or torch tensors (host or CUDA) and hands raw pointers to the library; the library
detects where a pointer lives.  Outputs follow the inputs: torch CUDA in -> torch CUDA out (no host round trip),
otherwise numpy.
"""
import ctypes as C

import numpy as np

from . import _lib
from .model import RobotModel, build_model_from_urdf

try:  # torch is plumbing (device memory, streams, torch.distributed), never a compute fallback
    import torch
except Exception:  # pragma: no cover
    torch = None


def _is_torch(x):
    return torch is not None and isinstance(x, torch.Tensor)


# the single-problem latency path counts microseconds of host time: ndarray.ctypes.data builds a helper object per
# call (1.7 us) and np.prod of a shape tuple costs 3.5 us; these two do the same in ~0.4 us and ~0.05 us
_from_buffer, _addressof = C.c_char.from_buffer, C.addressof


def _addr(a):
    try:
        return _addressof(_from_buffer(a))
    except (TypeError, ValueError, BufferError):          # read-only, empty or exotic arrays
        return a.ctypes.data


def _numel(x):
    if isinstance(x, np.ndarray):
        return x.size
    if _is_torch(x):
        return x.numel()
    return int(np.prod(np.shape(x)))


class Engine:
    def __init__(self, model=None, N=32, dt=0.01, dQ_cost=0.01, R_cost=1e-5, QN_cost=100.0, regularize=True, eps=1.0,
                 batch=1, dtype="fp64", device=0, max_qp=2, mu=10.0, step_tol=1e-3, threads=0, cluster=0,
                 qp_mode="exact", limits=False, admm=None):
        """qp_mode: 'exact' (block recursion, default) or 'osqp' (the reference's regime: warm-started ADMM at eps 1e-3,
        include/indy7_b200.h: i7_qp_mode).  limits=True adds the URDF position / velocity / effort limits of `model` as
        inequality rows of every QP (needs qp_mode='osqp').  admm: optional dict overriding rho, rho_u, sigma, alpha,
        eps_abs, eps_rel, max_iter, check."""
        self.lib = _lib.load()
        self._pinned = {}
        if model is None:
            model = build_model_from_urdf()
        if not isinstance(model, RobotModel):
            raise TypeError("model must be an indy7_mpc_b200.model.RobotModel (use load_robot_model)")
        if dtype not in ("fp64", "fp32"):
            raise ValueError("dtype must be 'fp64' or 'fp32'")
        if qp_mode not in ("exact", "osqp"):
            raise ValueError("qp_mode must be 'exact' or 'osqp'")
        self.qp_mode, self.limits = qp_mode, bool(limits)
        self.model = model
        self.N, self.batch, self.dtype, self.dt = int(N), int(batch), dtype, float(dt)
        self.np_dtype = np.float64 if dtype == "fp64" else np.float32
        self.nz = 18 * self.N - 6
        if isinstance(device, str):
            device = int(device.split(":")[1]) if ":" in device else 0
        self.device = int(device)
        cfg = _lib.I7Config()
        self.lib.i7_default_config(C.byref(cfg))
        cfg.N, cfg.batch, cfg.dtype, cfg.device = self.N, self.batch, (_lib.I7_F64 if dtype == "fp64" else _lib.I7_F32), self.device
        cfg.dt, cfg.dQ_cost, cfg.R_cost, cfg.QN_cost, cfg.eps = float(dt), float(dQ_cost), float(R_cost), float(QN_cost), float(eps)
        cfg.regularize, cfg.max_qp, cfg.mu, cfg.step_tol, cfg.threads = int(bool(regularize)), int(max_qp), float(mu), float(step_tol), int(threads)
        cfg.cluster = int(cluster)      # 0 auto, 1 one CTA per problem, 2|4|8 cluster of that many CTAs per problem (latency forms)
        cfg.qp_mode, cfg.limits = (_lib.I7_QP_OSQP if qp_mode == "osqp" else _lib.I7_QP_EXACT), int(bool(limits))
        for k, v in (admm or {}).items():
            if k not in ("rho", "rho_u", "sigma", "alpha", "eps_abs", "eps_rel", "max_iter", "check"):
                raise ValueError(f"unknown ADMM setting {k!r}")
            setattr(cfg, "admm_" + k, int(v) if k in ("max_iter", "check") else float(v))
        cfg.model = model.to_c()
        self._h = C.c_void_p()
        rc = self.lib.i7_create(C.byref(cfg), C.byref(self._h))
        if rc != 0:
            msg = self.lib.i7_last_error(None).decode()
            self._h = None
            if rc in (1, 4):
                raise ValueError(f"i7_create: {msg}")
            raise RuntimeError(f"i7_create failed (status {rc}): {msg}")

    def close(self):
        if getattr(self, "_h", None):
            self.lib.i7_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- helpers ----------------------------------------------------------------------------------------
    def _check(self, rc):
        if rc != 0:
            msg = self.lib.i7_last_error(self._h).decode()
            if rc == 1:
                raise ValueError(msg)
            raise RuntimeError(f"indy7_b200 status {rc}: {msg}")

    def _torch_dtype(self):
        return torch.float64 if self.dtype == "fp64" else torch.float32

    def _in(self, x, shape, keep):
        """-> (pointer, object kept alive).  `shape` is the required 2-D shape."""
        if x is None:
            return None
        if _is_torch(x):
            t = x.detach().to(self._torch_dtype()).reshape(shape).contiguous()
            if t.is_cuda and t.device.index != self.device:
                raise ValueError(f"tensor on {t.device}, engine on cuda:{self.device}")
            keep.append(t)
            return t.data_ptr()
        if type(x) is np.ndarray and x.dtype == self.np_dtype and x.flags.c_contiguous and x.size == shape[0] * shape[1]:
            a = x
        else:
            a = np.ascontiguousarray(np.asarray(x, dtype=self.np_dtype).reshape(shape))
        keep.append(a)
        return _addr(a)

    def _out(self, shape, like, keep, tag=""):
        if _is_torch(like) and like.is_cuda:
            t = torch.empty(shape, dtype=self._torch_dtype(), device=like.device)
            keep.append(t)
            return t, t.data_ptr()
        if _is_torch(like) and like.is_pinned():
            # page-locked inputs get page-locked outputs (the kernel then writes them itself, zero-copy).  Page-locking
            # costs milliseconds, so the buffers are owned by the engine: a ring of TWO per call site and shape — a
            # result stays valid until the second-next call of the same kind (long enough to be the warm start of the
            # next one), then it is overwritten.  Callers that keep results longer clone them.
            key = (tag, tuple(shape))
            ring = self._pinned.get(key)
            if ring is None:
                ring = self._pinned[key] = [[torch.empty(shape, dtype=self._torch_dtype(), pin_memory=True) for _ in range(2)], 0]
            t = ring[0][ring[1]]
            ring[1] ^= 1
            keep.append(t)
            return t, t.data_ptr()
        a = np.empty(shape, dtype=self.np_dtype)
        keep.append(a)
        return a, _addr(a)

    @staticmethod
    def _stream(like):
        if _is_torch(like) and like.is_cuda:
            return C.c_void_p(torch.cuda.current_stream(like.device).cuda_stream)
        return None

    def _goal_stride(self, goals, B):
        n = _numel(goals) // B
        if n == 3 * self.N:
            return 3
        if n == 6 * self.N:
            return 6
        raise ValueError(f"goals must hold N*3 or N*6 numbers per problem, got {n} for N={self.N}")

    def _batch_of(self, xu):
        n = _numel(xu)
        if n % self.nz:
            raise ValueError(f"trajectory length must be a multiple of 18N-6={self.nz}, got {n}")
        B = n // self.nz
        if B > self.batch:
            raise ValueError(f"batch {B} exceeds the engine's maximum batch {self.batch}")
        return B

    # ---- C-ABI calls ------------------------------------------------------------------------------------
    def solve(self, xu, x0, goals, want_stats=True, device_out=False):
        """sqp() of every problem.  Outputs live where the inputs live; with device_out=True and host inputs the
        results stay on the GPU (for a selection kernel to follow) and the call returns without waiting for the
        stream — page-locked inputs must then stay untouched until the stream is synchronised."""
        B = self._batch_of(xu)
        gs = self._goal_stride(goals, B)
        keep = []
        p_xu, p_x0, p_g = self._in(xu, (B, self.nz), keep), self._in(x0, (B, 12), keep), self._in(goals, (B, gs * self.N), keep)
        like = xu
        if device_out and not (_is_torch(xu) and xu.is_cuda):
            like = torch.empty(0, dtype=self._torch_dtype(), device=f"cuda:{self.device}")
        out, p_out = self._out((B, self.nz), like, keep, "solve.xu")
        stats, p_st = self._out((B, _lib.I7_STATS_LEN), like, keep, "solve.st") if want_stats else (None, None)
        self._check(self.lib.i7_solve(self._h, B, p_xu, p_x0, p_g, gs, p_out, p_st, self._stream(xu)))
        return out, stats

    def solve_qp(self, xu, x0, goals, want_duals=False):
        """z = setup_and_solve_qp(xu, x0, goals).x; with want_duals also y (B x 12N, duals of the equality rows)."""
        B = self._batch_of(xu)
        gs = self._goal_stride(goals, B)
        keep = []
        p_xu, p_x0, p_g = self._in(xu, (B, self.nz), keep), self._in(x0, (B, 12), keep), self._in(goals, (B, gs * self.N), keep)
        out, p_out = self._out((B, self.nz), xu, keep)
        y, p_y = self._out((B, 12 * self.N), xu, keep) if want_duals else (None, None)
        self._check(self.lib.i7_solve_qp(self._h, B, p_xu, p_x0, p_g, gs, p_out, p_y, self._stream(xu)))
        return (out, y) if want_duals else out

    def qp_state(self, batch=None):
        """qp_mode 'osqp': (x (B, 18N-6), y (B, 12N), y_lim (B, 18N-18), info (B, 4) = [iterations, r_prim, r_dual,
        status]) of the ADMM solver after the latest solve / solve_qp (i7_qp_state)."""
        B = int(batch) if batch is not None else self.batch
        x = np.empty((B, self.nz)); y = np.empty((B, 12 * self.N)); yl = np.empty((B, self.nz - 12)); info = np.empty((B, 4))
        self._check(self.lib.i7_qp_state(self._h, B, _addr(x), _addr(y), _addr(yl), _addr(info), None))
        return x, y, yl, info

    def dump_qp(self, xu, x0, goals):
        B = self._batch_of(xu)
        gs = self._goal_stride(goals, B)
        N = self.N
        keep = []
        p_xu, p_x0, p_g = self._in(xu, (B, self.nz), keep), self._in(x0, (B, 12), keep), self._in(goals, (B, gs * N), keep)
        A, pA = self._out((B, 360 * (N - 1) + 144), None, keep)
        Pd, pP = self._out((B, 27 * N + 6 * (N - 1)), None, keep)
        g, pg = self._out((B, self.nz), None, keep)
        l, pl = self._out((B, 12 * N), None, keep)
        self._check(self.lib.i7_dump_qp(self._h, B, p_xu, p_x0, p_g, gs, pA, pP, pg, pl, None))
        return A, Pd, g, l

    def merit(self, xu, xu_ref, goals):
        B = self._batch_of(xu)
        gs = self._goal_stride(goals, B)
        keep = []
        p_xu, p_ref, p_g = self._in(xu, (B, self.nz), keep), self._in(xu_ref, (B, self.nz), keep), self._in(goals, (B, gs * self.N), keep)
        out, p_out = self._out((B, 6), None, keep)
        self._check(self.lib.i7_merit(self._h, B, p_xu, p_ref, p_g, gs, p_out, None))
        return out

    def linesearch(self, xu, xu_full, goals):
        B = self._batch_of(xu)
        gs = self._goal_stride(goals, B)
        keep = []
        p_xu, p_f, p_g = self._in(xu, (B, self.nz), keep), self._in(xu_full, (B, self.nz), keep), self._in(goals, (B, gs * self.N), keep)
        alpha, p_a = self._out((B,), None, keep)
        merits, p_m = self._out((B, 9), None, keep)
        self._check(self.lib.i7_linesearch(self._h, B, p_xu, p_f, p_g, gs, p_a, p_m, None))
        return alpha, merits

    def dynamics(self, q, v, tau, f_ext=None):
        n = int(np.prod(np.shape(q))) // 6
        keep = []
        pq, pv, pt = self._in(q, (n, 6), keep), self._in(v, (n, 6), keep), self._in(tau, (n, 6), keep)
        pf = self._in(f_ext, (n, 6), keep)
        qdd, p1 = self._out((n, 6), None, keep)
        dq, p2 = self._out((n, 6, 6), None, keep)
        dv, p3 = self._out((n, 6, 6), None, keep)
        mi, p4 = self._out((n, 6, 6), None, keep)
        self._check(self.lib.i7_dynamics(self._h, n, pq, pv, pt, pf, p1, p2, p3, p4, None))
        return qdd, dq, dv, mi

    def eepos(self, q, jacobian=False):
        n = int(np.prod(np.shape(q))) // 6
        keep = []
        pq = self._in(q, (n, 6), keep)
        p, pp = self._out((n, 3), None, keep)
        J, pJ = self._out((n, 3, 6), None, keep) if jacobian else (None, None)
        self._check(self.lib.i7_eepos(self._h, n, pq, pp, pJ, None))
        return (p, J) if jacobian else p

    def sim_forward(self, x, u, dt, batch=None):
        x_arr = x if _is_torch(x) else np.asarray(x)
        rows = int(np.prod(x_arr.shape)) // 12
        bcast = 1 if (rows == 1 and batch not in (None, 1)) else 0
        B = int(batch) if batch is not None else rows
        keep = []
        px, pu = self._in(x, (rows, 12), keep), self._in(u, (rows, 6), keep)
        out, po = self._out((B, 12), x, keep)
        self._check(self.lib.i7_sim_forward(self._h, B, px, pu, float(dt), bcast, po, self._stream(x)))
        return out

    TAILS = {"osqp_mpc": 0, "hold": 1, "repeat": 2}

    def shift(self, xu_new, steps=1, tail="repeat", xu_prev=None, x_cur=None, want_x0=True):
        """Warm start of the next tick (i7_shift): the solution moved `steps` knots towards the present.  tail:
        'osqp_mpc' (src/osqp_mpc.py:65-70, needs xu_prev), 'hold' (trajopt_solver.cuh:153-172), 'repeat' (batch drivers).
        Returns (xu_next, x0_next)."""
        B = self._batch_of(xu_new)
        keep = []
        p_new = self._in(xu_new, (B, self.nz), keep)
        p_prev = self._in(xu_prev, (B, self.nz), keep)
        p_cur = self._in(x_cur, (B, 12), keep)
        out, p_out = self._out((B, self.nz), xu_new, keep, "shift.xu")
        x0, p_x0 = self._out((B, 12), xu_new, keep, "shift.x0") if want_x0 else (None, None)
        self._check(self.lib.i7_shift(self._h, B, int(steps), self.TAILS[tail], p_new, p_prev, p_cur, p_out, p_x0,
                                      self._stream(xu_new)))
        return out, x0

    def goal_window(self, traj, offsets=None, tick=0, wrap=True, out_stride=3, batch=None):
        """Goal windows of `batch` lanes out of one reference trajectory (T x 3 or T x 6), i7_goal_window."""
        B = int(batch) if batch is not None else self.batch
        tlen = int(traj.shape[0])
        tstride = _numel(traj) // tlen
        keep = []
        p_traj = self._in(traj, (tlen, tstride), keep)
        p_off = None
        if offsets is not None:
            if _is_torch(offsets):
                o = offsets.to(torch.int32).contiguous(); keep.append(o); p_off = o.data_ptr()
            else:
                o = np.ascontiguousarray(np.asarray(offsets, dtype=np.int32)); keep.append(o); p_off = o.ctypes.data
        goals, p_g = self._out((B, self.N * out_stride), traj, keep, "goals")
        self._check(self.lib.i7_goal_window(self._h, B, p_traj, tlen, tstride, p_off, int(tick), int(bool(wrap)), int(out_stride),
                                            p_g, self._stream(traj)))
        return goals

    def find_best_idx(self, x_last, u_last, xs, dt, batch=None, tie_last=True, return_errors=False):
        """ONE launch: plant step of every lane under its wrench hypothesis from (x_last, u_last), error against the
        measured state xs, argmin (i7_find_best_idx).  Returns the index (None when no lane qualified), or
        (index, errors) with return_errors."""
        B = int(batch) if batch is not None else self.batch
        keep = []
        px, pu, ps = self._in(x_last, (1, 12), keep), self._in(u_last, (1, 6), keep), self._in(xs, (1, 12), keep)
        res, pr = self._out((2,), x_last, keep)
        errs, pe = self._out((B,), x_last, keep) if return_errors else (None, None)
        self._check(self.lib.i7_find_best_idx(self._h, B, px, pu, ps, float(dt), 1 if tie_last else 0, None, pe, pr,
                                              self._stream(x_last)))
        idx = int(res[0])
        idx = None if idx < 0 else idx
        return (idx, errs) if return_errors else idx

    def find_best_idx_static(self, x_last, u_last, xs, dt, batch=None, tie_last=True, return_errors=False):
        """find_best_idx without allocations inside the library call (i7_find_best_idx_static: scratch in a static
        device array of at most _lib.I7_STATIC_MAX_BATCH lanes, shared by every call on the device — overlapping calls
        race).  No counterpart in the reference.  Same results as find_best_idx."""
        B = int(batch) if batch is not None else self.batch
        keep = []
        px, pu, ps = self._in(x_last, (1, 12), keep), self._in(u_last, (1, 6), keep), self._in(xs, (1, 12), keep)
        res, pr = self._out((2,), x_last, keep)
        errs, pe = self._out((B,), x_last, keep) if return_errors else (None, None)
        self._check(self.lib.i7_find_best_idx_static(self._h, B, px, pu, ps, float(dt), 1 if tie_last else 0, None, pe, pr,
                                                     self._stream(x_last)))
        idx = int(res[0])
        idx = None if idx < 0 else idx
        return (idx, errs) if return_errors else idx

    def find_best_idx_async(self, x_last, u_last, xs, dt, batch=None, tie_last=True):
        """find_best_idx without the read-back: returns the 2-number result [index (-1: no lane qualified), error] in the
        memory of the inputs (a device tensor for device inputs — nothing waits for the GPU), to be chained into
        `resample_wrench(best=...)`."""
        B = int(batch) if batch is not None else self.batch
        keep = []
        px, pu, ps = self._in(x_last, (1, 12), keep), self._in(u_last, (1, 6), keep), self._in(xs, (1, 12), keep)
        res, pr = self._out((2,), x_last, keep)
        self._check(self.lib.i7_find_best_idx(self._h, B, px, pu, ps, float(dt), 1 if tie_last else 0, None, None, pr,
                                              self._stream(x_last)))
        return res

    def resample_wrench(self, f_ext, noise, best, keep_lane=_lib.I7_KEEP_BEST, keep_torque=False, decay=1.0, apply=True,
                        inplace=False):
        """resample_f_ext_batch (gato_controller.py:120-129) on the device (i7_resample_wrench): f = f[best] + noise,
        kept lane = f[best], torque part 0, lane 0 = 0, x decay — bit-identical to the numpy statements.  f_ext, noise:
        (B, 6) float64, numpy or torch (host or device); `noise` are the caller's np.random draws.  best: a lane index, or
        the result of `find_best_idx_async` (read on the device).  keep_lane: lane index, I7_KEEP_BEST, I7_KEEP_NONE
        (None).  apply: the new batch also becomes the engine's wrench batch.  Returns the new batch (same kind of
        array as f_ext; the same tensor with inplace=True)."""
        keep = []

        def f64(x):
            if _is_torch(x):
                t = x.detach().reshape(-1, 6)
                if t.dtype != torch.float64 or not t.is_contiguous():
                    if inplace and x is f_ext:
                        raise ValueError("inplace needs a contiguous float64 tensor")
                    t = t.to(torch.float64).contiguous()
                keep.append(t)
                return t, t.data_ptr()
            a = np.ascontiguousarray(np.asarray(x, dtype=np.float64).reshape(-1, 6))
            keep.append(a)
            return a, _addr(a)
        f, pf = f64(f_ext)
        n, pn = f64(noise)
        B = f.shape[0]
        if tuple(n.shape) != (B, 6):
            raise ValueError(f"noise must be ({B}, 6)")
        if B > self.batch:
            raise ValueError(f"batch {B} exceeds the engine's maximum batch {self.batch}")
        if isinstance(best, (int, np.integer)):
            bi, pb = int(best), None
            if not 0 <= bi < B:
                raise ValueError(f"best lane {bi} outside the batch of {B}")
        else:
            bi, pb = -1, self._in(best, (1, 2), keep)
        if inplace:
            out, po = f, pf
        elif _is_torch(f):
            out = torch.empty_like(f)
            po = out.data_ptr()
        else:
            out = np.empty_like(f)
            po = _addr(out)
        kl = _lib.I7_KEEP_NONE if keep_lane is None else int(keep_lane)
        flags = (_lib.I7_RESAMPLE_KEEP_TORQUE if keep_torque else 0) | (_lib.I7_RESAMPLE_APPLY if apply else 0)
        like = f if (_is_torch(f) and f.is_cuda) else None
        self._check(self.lib.i7_resample_wrench(self._h, B, bi, pb, pf, pn, kl, flags, float(decay), po, self._stream(like)))
        return out

    def broadcast_best(self, xu_new, best, inplace=False):
        """XU_best = xu_new[best]; every lane of the returned batch = XU_best (gato_controller.py:226,249;
        i7_broadcast_best).  best: a lane index or the result of `find_best_idx_async`.  Returns (xu_batch, xu_best)."""
        B = self._batch_of(xu_new)
        keep = []
        p_new = self._in(xu_new, (B, self.nz), keep)
        if isinstance(best, (int, np.integer)):
            bi, pb = int(best), None
            if not 0 <= bi < B:
                raise ValueError(f"best lane {bi} outside the batch of {B}")
        else:
            bi, pb = -1, self._in(best, (1, 2), keep)
        if inplace:
            if not (_is_torch(xu_new) and xu_new.is_cuda and keep[0].data_ptr() == xu_new.data_ptr()):
                raise ValueError("inplace needs a contiguous device tensor of the engine dtype")
            out, p_out = xu_new, p_new
        else:
            out, p_out = self._out((B, self.nz), xu_new, keep, "bcast")
        bo, p_bo = self._out((self.nz,), xu_new, keep, "bcast_best")
        self._check(self.lib.i7_broadcast_best(self._h, B, p_new, bi, pb, p_out, p_bo, self._stream(xu_new)))
        return out, bo

    def pack_best(self, stats, xu, index_offset=0):
        """[costs | best_cost, best_index, XU_best] on the device the inputs live on (i7_pack_best)."""
        B = self._batch_of(xu)
        keep = []
        p_st, p_xu = self._in(stats, (B, _lib.I7_STATS_LEN), keep), self._in(xu, (B, self.nz), keep)
        out, p_out = self._out((B + 2 + self.nz,), xu, keep)
        self._check(self.lib.i7_pack_best(self._h, B, p_st, p_xu, int(index_offset), p_out, self._stream(xu)))
        return out

    def reduce_best(self, gathered, nranks):
        n = int(np.prod(gathered.shape))
        stride = n // nranks
        keep = []
        p_in = self._in(gathered, (nranks, stride), keep)
        out, p_out = self._out((2 + self.nz,), gathered, keep)
        self._check(self.lib.i7_reduce_best(self._h, nranks, stride, p_in, p_out, self._stream(gathered)))
        return out

    # ---- peer-memory exchange between the ranks of one node (i7_p2p_*) ----
    def p2p_create(self, world, rank):
        """Allocates this rank's exchange buffer; returns its 64-byte inter-process handle."""
        buf = C.create_string_buffer(64)
        self._check(self.lib.i7_p2p_create(self._h, int(world), int(rank), buf))
        return buf.raw

    def p2p_connect(self, handles):
        """handles: the handles of all ranks in rank order (list of 64-byte strings)."""
        self._check(self.lib.i7_p2p_connect(self._h, b"".join(handles)))

    def p2p_exchange(self, stats, xu, index_offset=0):
        """ONE kernel: pack the local record, store it into every peer's buffer over NVLink, wait for all ranks, pick
        the winner.  Device tensors only; returns [best_cost, global index, XU_best] on the device."""
        B = self._batch_of(xu)
        keep = []
        p_st, p_xu = self._in(stats, (B, _lib.I7_STATS_LEN), keep), self._in(xu, (B, self.nz), keep)
        out, p_out = self._out((2 + self.nz,), xu, keep)
        self._check(self.lib.i7_p2p_exchange(self._h, B, p_st, p_xu, int(index_offset), p_out, self._stream(xu)))
        return out

    def solve_select(self, xu, x0, goals, index_offset=0):
        """sqp() of every problem AND the best-trajectory selection / multi-GPU exchange in ONE launch (i7_solve_select):
        returns (xu_out, stats, best = [best_cost, global index, XU_best]) on the device.  Device tensors only."""
        B = self._batch_of(xu)
        gs = self._goal_stride(goals, B)
        keep = []
        p_xu, p_x0, p_g = self._in(xu, (B, self.nz), keep), self._in(x0, (B, 12), keep), self._in(goals, (B, gs * self.N), keep)
        out, p_out = self._out((B, self.nz), xu, keep)
        stats, p_st = self._out((B, _lib.I7_STATS_LEN), xu, keep)
        best, p_best = self._out((2 + self.nz,), xu, keep)
        self._check(self.lib.i7_solve_select(self._h, B, p_xu, p_x0, p_g, gs, p_out, p_st, int(index_offset), p_best, self._stream(xu)))
        return out, stats, best

    def exchange_wait_ns(self, reset=True):
        """Nanoseconds the epilogues of solve_select have waited for the other ranks since the last reset."""
        v = C.c_int64()
        self._check(self.lib.i7_exchange_wait_ns(self._h, C.byref(v), int(bool(reset))))
        return int(v.value)

    def p2p_status(self, like=None):
        self._check(self.lib.i7_p2p_status(self._h, self._stream(like) if like is not None else None))

    def p2p_reset(self):
        """Collective (every rank, barrier before and after): clears the time-out state, flags and tick counter."""
        self._check(self.lib.i7_p2p_reset(self._h))

    def measure_fma_peak(self):
        tf = C.c_double()
        self._check(self.lib.i7_measure_fma_peak(self._h, C.byref(tf)))
        return tf.value

    def profile_phases(self, enable=True, batch=None):
        """Enable the phase clocks, or (batch given) fetch the (batch, 32) clock64 stamps of the latest solve."""
        if batch is None:
            self._check(self.lib.i7_profile_phases(self._h, int(enable), None, 0))
            return None
        buf = np.zeros((batch, 32), dtype=np.int64)
        self._check(self.lib.i7_profile_phases(self._h, int(enable), buf.ctypes.data_as(C.POINTER(C.c_int64)), batch))
        return buf

    def set_external_wrench(self, f_ext):
        if f_ext is None:
            self._check(self.lib.i7_set_external_wrench(self._h, None, 0))
            return
        f = np.ascontiguousarray(np.asarray(f_ext, dtype=np.float64).reshape(-1, 6))
        self._check(self.lib.i7_set_external_wrench(self._h, f.ctypes.data_as(C.POINTER(C.c_double)), f.shape[0]))

    def set_option(self, option, value):
        """i7_set_option; option: an I7_OPT_* constant of indy7_mpc_b200._lib."""
        self._check(self.lib.i7_set_option(self._h, int(option), int(value)))

    def set_weights(self, dQ_cost, R_cost, QN_cost, regularize, eps):
        self._check(self.lib.i7_set_weights(self._h, float(dQ_cost), float(R_cost), float(QN_cost), int(bool(regularize)), float(eps)))

    def reset(self):
        self._check(self.lib.i7_reset(self._h))

    def reset_rho(self):
        self._check(self.lib.i7_reset_rho(self._h))

    def reset_lambda(self):
        self._check(self.lib.i7_reset_lambda(self._h))

    def last_solve_time_us(self):
        us = C.c_double()
        self._check(self.lib.i7_last_solve_time_us(self._h, C.byref(us)))
        return us.value

    def kernel_form(self):
        """(threads per CTA, CTAs per problem, shared-memory bytes per CTA) of the i7_solve kernel of this handle."""
        t, c, m = C.c_int32(), C.c_int32(), C.c_int32()
        self._check(self.lib.i7_kernel_form(self._h, C.byref(t), C.byref(c), C.byref(m)))
        return t.value, c.value, m.value

    def launch_count(self):
        return int(self.lib.i7_launch_count(self._h))
