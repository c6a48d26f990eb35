"""ctypes wrapper of the plain-C oracle (oracle/c/oracle.c -> oracle/_build/liboracle.so).  TEST / BASELINE
INFRASTRUCTURE: used by tests/ and by bench.py's cpu_baseline / --impl reference legs only."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "liboracle.so")


class OrcModel(C.Structure):
    _fields_ = [("R", C.c_double * 9 * 6), ("p", C.c_double * 3 * 6), ("I", C.c_double * 36 * 6), ("g", C.c_double * 3)]


class OrcCfg(C.Structure):
    _fields_ = [("N", C.c_int), ("dt", C.c_double), ("dQ", C.c_double), ("R", C.c_double), ("QN", C.c_double),
                ("eps", C.c_double), ("regularize", C.c_int), ("backend", C.c_int), ("mu", C.c_double),
                ("step_tol", C.c_double), ("max_qp", C.c_int), ("limits", C.c_int), ("qmin", C.c_double * 6),
                ("qmax", C.c_double * 6), ("vmax", C.c_double * 6), ("umax", C.c_double * 6), ("rho_x", C.c_double),
                ("rho_u", C.c_double), ("max_iter", C.c_int)]


_lib = None
_dp = C.POINTER(C.c_double)


def _ptr(a):
    return None if a is None else a.ctypes.data_as(_dp)


def load(build=True):
    global _lib
    if _lib is None:
        if build and not os.path.exists(LIB):
            subprocess.run(["make", "-s", "-C", os.path.join(HERE, "c")], check=True)
        lib = C.CDLL(LIB)
        lib.orc_work_new.restype = C.c_void_p
        lib.orc_work_new.argtypes = [C.c_int]
        lib.orc_work_free.argtypes = [C.c_void_p]
        lib.orc_work_reset.argtypes = [C.c_void_p]
        lib.orc_work_state.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp]
        lib.orc_sqp.restype = C.c_int
        lib.orc_sqp.argtypes = [C.POINTER(OrcModel), C.POINTER(OrcCfg), C.c_void_p, _dp, _dp, _dp, _dp, _dp, _dp]
        lib.orc_solve_qp.restype = C.c_int
        lib.orc_solve_qp.argtypes = [C.POINTER(OrcModel), C.POINTER(OrcCfg), C.c_void_p, _dp, _dp, _dp, _dp, _dp]
        lib.orc_sqp_batch.restype = C.c_int
        lib.orc_sqp_batch.argtypes = [C.POINTER(OrcModel), C.POINTER(OrcCfg), C.POINTER(C.c_void_p), C.c_int, _dp, _dp, _dp, _dp, _dp, _dp, C.c_int]
        lib.orc_dynamics.argtypes = [C.POINTER(OrcModel)] + [_dp] * 8
        lib.orc_eepos.argtypes = [C.POINTER(OrcModel), _dp, _dp, _dp]
        lib.orc_rk4.argtypes = [C.POINTER(OrcModel), _dp, _dp, _dp, C.c_double, _dp, _dp, _dp]
        lib.orc_num_threads.restype = C.c_int
        _lib = lib
    return _lib


class COracle:
    """Batch-capable C oracle with per-problem ADMM warm-start state (the reference keeps one OSQP object per solver)."""

    def __init__(self, model, N=32, dt=0.01, dQ_cost=0.01, R_cost=1e-5, QN_cost=100.0, regularize=True, eps=1.0,
                 backend="admm", batch=1, mu=10.0, step_tol=1e-3, max_qp=2, limits=None, rho_x=0.1, rho_u=1e-3, max_iter=4000):
        """limits: None, or (qmin, qmax, vmax, umax) — four 6-vectors (description/indy7.urdf:203-238): box rows on every
        q_k, v_k (k >= 1) and u_k of the QP; needs the ADMM back-end."""
        self.lib = load()
        self.N, self.batch, self.nz = N, batch, 18 * N - 6
        m = OrcModel()
        for i in range(6):
            for a in range(9):
                m.R[i][a] = float(model.R[i].reshape(-1)[a])
            for a in range(3):
                m.p[i][a] = float(model.p[i][a])
            for a in range(36):
                m.I[i][a] = float(model.I[i].reshape(-1)[a])
        for a in range(3):
            m.g[a] = float(model.gravity[a])
        self.m = m
        self.cfg = OrcCfg(N, dt, dQ_cost, R_cost, QN_cost, eps, int(regularize), {"kkt": 0, "admm": 1}[backend], mu, step_tol, max_qp)
        self.cfg.rho_x, self.cfg.rho_u, self.cfg.max_iter = rho_x, rho_u, max_iter
        if limits is not None:
            if backend != "admm":
                raise ValueError("limits need the ADMM back-end")
            self.cfg.limits = 1
            for name, vals in zip(("qmin", "qmax", "vmax", "umax"), limits):
                arr = getattr(self.cfg, name)
                for i in range(6):
                    arr[i] = float(vals[i])
        self.works = (C.c_void_p * batch)(*[self.lib.orc_work_new(N) for _ in range(batch)])

    def __del__(self):
        try:
            for w in self.works:
                self.lib.orc_work_free(w)
        except Exception:
            pass

    def reset(self):
        for w in self.works:
            self.lib.orc_work_reset(w)

    def threads(self):
        return int(self.lib.orc_num_threads())

    def sqp(self, xs, goals, xu, f_ext=None, nthreads=0):
        xs = np.ascontiguousarray(np.asarray(xs, float).reshape(-1, 12))
        B = xs.shape[0]
        goals = np.ascontiguousarray(np.asarray(goals, float).reshape(B, 3 * self.N))
        xu = np.ascontiguousarray(np.asarray(xu, float).reshape(B, self.nz))
        f = None if f_ext is None else np.ascontiguousarray(np.asarray(f_ext, float).reshape(B, 6))
        out = np.empty_like(xu); stats = np.empty((B, 12))
        its = self.lib.orc_sqp_batch(C.byref(self.m), C.byref(self.cfg), self.works, B, _ptr(xs), _ptr(goals), _ptr(xu), _ptr(f),
                                     _ptr(out), _ptr(stats), nthreads)
        return out, stats, its

    def admm_state(self, b=0):
        """(x, y, yl, [r_prim, r_dual]) of problem b's warm-start state after its latest QP solve."""
        x = np.empty(self.nz); y = np.empty(12 * self.N); yl = np.empty(self.nz - 12); res = np.empty(2)
        self.lib.orc_work_state(self.works[b], _ptr(x), _ptr(y), _ptr(yl), _ptr(res))
        return x, y, yl, res

    def solve_qp(self, xu, xs, goals, f_ext=None):
        xu = np.ascontiguousarray(np.asarray(xu, float).reshape(-1)); xs = np.ascontiguousarray(np.asarray(xs, float).reshape(-1))
        goals = np.ascontiguousarray(np.asarray(goals, float).reshape(-1))
        f = None if f_ext is None else np.ascontiguousarray(np.asarray(f_ext, float).reshape(6))
        z = np.empty(self.nz)
        its = self.lib.orc_solve_qp(C.byref(self.m), C.byref(self.cfg), self.works[0], _ptr(xu), _ptr(xs), _ptr(goals), _ptr(f), _ptr(z))
        return z, its

    def dynamics(self, q, v, tau, f_ext=None):
        q, v, tau = (np.ascontiguousarray(np.asarray(a, float).reshape(6)) for a in (q, v, tau))
        f = None if f_ext is None else np.ascontiguousarray(np.asarray(f_ext, float).reshape(6))
        qdd = np.empty(6); dq = np.empty((6, 6)); dv = np.empty((6, 6)); mi = np.empty((6, 6))
        self.lib.orc_dynamics(C.byref(self.m), _ptr(q), _ptr(v), _ptr(tau), _ptr(f), _ptr(qdd), _ptr(dq), _ptr(dv), _ptr(mi))
        return qdd, dq, dv, mi

    def eepos(self, q):
        q = np.ascontiguousarray(np.asarray(q, float).reshape(6))
        p = np.empty(3); J = np.empty((3, 6))
        self.lib.orc_eepos(C.byref(self.m), _ptr(q), _ptr(p), _ptr(J))
        return p, J

    def rk4(self, q, v, u, dt, f_ext=None):
        q, v, u = (np.ascontiguousarray(np.asarray(a, float).reshape(6)) for a in (q, v, u))
        f = None if f_ext is None else np.ascontiguousarray(np.asarray(f_ext, float).reshape(6))
        qn = np.empty(6); vn = np.empty(6)
        self.lib.orc_rk4(C.byref(self.m), _ptr(q), _ptr(v), _ptr(u), dt, _ptr(f), _ptr(qn), _ptr(vn))
        return qn, vn
