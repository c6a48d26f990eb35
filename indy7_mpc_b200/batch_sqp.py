"""Drop-in for the external pybind11 module ``bindings.batch_sqp`` used by gato_controller.py:8,53-68,
src/gato_mpc_batch.py:5,14-55, src/gato_mpc_batch_sample.py and src/mpc/mpc/batch_solver.py (binary absent from
the reference; surface reconstructed from the call sites).

Classes ``SQPSolverfloat_<B>`` for B in {1,2,4,...,256}: zero-argument constructors, ``solve(XU_batch, dt,
xcur_batch, eepos_goals_batch) -> dict``, ``reset()``, ``resetRho()``, ``resetLambda()``,
``set_external_wrench_batch(f_ext)``, ``sim_forward(x, u, dt) -> (B, 12)``.  The horizon N is taken from the shape
of the first ``XU_batch`` (it is compiled into the reference binding).  Outputs are float32 like the reference
binding; the arithmetic type is ``COMPUTE_DTYPE`` ('fp32' mirrors SQPSolverfloat; 'fp64' is the accurate mode).

Environment (read at import): ``INDY7_B200_BATCH_DTYPE=fp64|fp32``; ``INDY7_B200_QP_MODE=exact|osqp`` — 'osqp' solves
every QP by the warm-started ADMM of the reference's own solver (then ``reset`` / ``resetRho`` / ``resetLambda`` act on
its state, as they do on the GATO solver's rho and lambda, gato_controller.py:134-138) and fills ``pcg_stats`` with its
iteration counts; ``INDY7_B200_LIMITS=1`` adds the URDF joint limits as inequality rows (osqp mode, fp64).
"""
import os

import numpy as np

from .engine import Engine
from .model import build_model_from_urdf

COMPUTE_DTYPE = os.environ.get("INDY7_B200_BATCH_DTYPE", "fp64")
QP_MODE = os.environ.get("INDY7_B200_QP_MODE", "exact")
LIMITS = os.environ.get("INDY7_B200_LIMITS", "0") not in ("0", "", "false", "False")
SUPPORTED_BATCH = (1, 2, 4, 8, 16, 32, 64, 128, 256)


class _SQPSolverBase:
    BATCH = 1

    def __init__(self):
        self.batch_size = self.BATCH
        self._model = build_model_from_urdf()
        self._engine = None
        self._f_ext = np.zeros((self.BATCH, 6))
        self._N = None
        self._dt = None

    def _ensure(self, N, dt):
        if self._engine is None or N != self._N or abs(dt - self._dt) > 0:
            if self._engine is not None:
                self._engine.close()
            self._engine = Engine(self._model, N=N, dt=dt, batch=self.BATCH, dtype=COMPUTE_DTYPE, qp_mode=QP_MODE, limits=LIMITS)
            self._N, self._dt = N, dt
            self._engine.set_external_wrench(self._f_ext)
        return self._engine

    def solve(self, XU_batch, dt, xcur_batch, eepos_goals_batch):
        XU = np.asarray(XU_batch, float)
        if XU.ndim != 2 or XU.shape[0] != self.BATCH or (XU.shape[1] + 6) % 18:
            raise ValueError(f"XU_batch must be ({self.BATCH}, 18N-6), got {XU.shape}")
        N = (XU.shape[1] + 6) // 18
        x0 = np.asarray(xcur_batch, float).reshape(self.BATCH, 12)
        goals = np.asarray(eepos_goals_batch, float).reshape(self.BATCH, -1)
        if goals.shape[1] not in (3 * N, 6 * N):
            raise ValueError(f"eepos_goals_batch must be ({self.BATCH}, {6 * N}) [x y z 0 0 0 per knot]")
        eng = self._ensure(N, float(dt))
        out, st = eng.solve(XU, x0, goals)
        st = np.asarray(st, float)
        nqp = int(st[:, 0].max())
        return {
            "xu_trajectory": np.asarray(out, np.float32),
            "solve_time_us": float(eng.last_solve_time_us()),
            "sqp_iterations": st[:, 0].astype(np.int32),
            # one entry per QP solve: exact mode — solved directly, zero iterations; osqp mode — the ADMM iterations
            "pcg_stats": [{"pcg_iterations": st[:, 10 + i].astype(np.int32)} for i in range(nqp)],
            "line_search_stats": [{"step_size": [float(a) for a in st[:, 4 + i]]} for i in range(nqp)],
        }

    def reset(self):
        if self._engine is not None:
            self._engine.reset()

    def resetRho(self):
        if self._engine is not None:
            self._engine.reset_rho()

    def resetLambda(self):
        if self._engine is not None:
            self._engine.reset_lambda()

    def set_external_wrench_batch(self, f_ext_batch):
        f = np.asarray(f_ext_batch, float)
        if f.shape != (self.BATCH, 6):
            raise ValueError(f"f_ext_batch must be ({self.BATCH}, 6)")
        self._f_ext = f.copy()
        if self._engine is not None:
            self._engine.set_external_wrench(self._f_ext)

    def sim_forward(self, x, u, dt):
        eng = self._engine or self._ensure(self._N or 32, self._dt or 0.01)
        x = np.asarray(x, float).reshape(-1)
        u = np.asarray(u, float).reshape(-1)
        return np.asarray(eng.sim_forward(x[:12], u[:6], float(dt), batch=self.BATCH), float)

    def find_best_idx(self, x_last, u_last, xs, dt, tie_last=True):
        """Extension of the reference binding: ``BatchGATO.find_best_idx`` (gato_controller.py:109-118) — sim_forward
        of all lanes, error against the measured state, `<=` argmin — as ONE launch that returns only the index."""
        eng = self._engine or self._ensure(self._N or 32, self._dt or 0.01)
        f = lambda a, n: np.asarray(a, float).reshape(-1)[:n]
        return eng.find_best_idx(f(x_last, 12), f(u_last, 6), f(xs, 12), float(dt), batch=self.BATCH,
                                 tie_last=tie_last)


    def resample_f_ext_batch(self, best_idx, noise, keep_lane=-1, keep_torque=False, decay=1.0):
        """Extension of the reference binding: ``BatchGATO.resample_f_ext_batch`` (gato_controller.py:120-129) on the
        device (i7_resample_wrench): the batch set by set_external_wrench_batch is re-centred on lane `best_idx` with the
        caller's normal draws `noise` (B x 6), becomes the solver's wrench batch and is returned."""
        eng = self._engine or self._ensure(self._N or 32, self._dt or 0.01)
        self._f_ext = eng.resample_wrench(self._f_ext, noise, int(best_idx), keep_lane=keep_lane, keep_torque=keep_torque,
                                          decay=decay, apply=True)
        return self._f_ext.copy()


def _make(B):
    return type(f"SQPSolverfloat_{B}", (_SQPSolverBase,), {"BATCH": B, "__doc__": f"batch of {B} SQP problems"})


for _B in SUPPORTED_BATCH:
    globals()[f"SQPSolverfloat_{_B}"] = _make(_B)
del _B
