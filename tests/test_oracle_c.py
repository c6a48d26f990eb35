"""The plain-C oracle port (timed CPU baseline) must agree with the numpy oracle."""
import numpy as np

from oracle import rbd
from oracle.cport import COracle
from oracle.osqp_solver import OSQPSolver
from oracle.osqp_sqp import SQP_OSQP
from tests.cases import make_cases


def rel(a, b):
    return np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(np.asarray(b)).max(), 1e-300)


def test_c_dynamics_and_plant(oracle_model):
    rng = np.random.default_rng(0)
    c = COracle(oracle_model, backend="kkt")
    for _ in range(5):
        q, v, u, f = rng.uniform(-3, 3, 6), rng.uniform(-2, 2, 6), rng.uniform(-60, 60, 6), rng.uniform(-10, 10, 6)
        qdd, dq, dv, mi = c.dynamics(q, v, u, f)
        rq, rv, rm, ra = rbd.aba_derivatives(oracle_model, q, v, u, f)
        assert rel(qdd, ra) < 1e-11 and rel(dq, rq) < 1e-10 and rel(dv, rv) < 1e-10 and rel(mi, rm) < 1e-11
        qn, vn = c.rk4(q, v, u, 0.01, f)
        rqn, rvn = rbd.rk4(oracle_model, q, v, u, 0.01, f)
        assert rel(np.concatenate([qn, vn]), np.concatenate([rqn, rvn])) < 1e-13
        p, J = c.eepos(q)
        rp, rJ = rbd.d_eepos(oracle_model, q)
        assert rel(p, rp) < 1e-14 and np.abs(J - rJ).max() < 1e-14


def test_c_sqp_matches_numpy_oracle(oracle_model):
    xu, x0, goals = make_cases(oracle_model, 4, seed=6)
    ck = COracle(oracle_model, backend="kkt", batch=4)
    out, st, _ = ck.sqp(x0, goals, xu)
    for b in range(4):
        sq = SQP_OSQP(OSQPSolver(oracle_model, qp_backend="kkt"))
        ref = sq.sqp(x0[b], goals[b], xu[b])
        assert rel(out[b], ref) < 1e-7
        assert list(st[b, 4:6]) == sq.stats["linesearch_alphas"]["values"]
        assert int(st[b, 0]) == sq.stats["qp_iters"]["values"][0]


def test_c_admm_meets_osqp_tolerances(oracle_model):
    xu, x0, goals = make_cases(oracle_model, 3, seed=3)
    ca = COracle(oracle_model, backend="admm")
    for b in range(3):
        ca.reset()
        z, it = ca.solve_qp(xu[b], x0[b], goals[b])
        ref = OSQPSolver(oracle_model, qp_backend="kkt").setup_and_solve_qp(xu[b], x0[b], goals[b]).x
        assert it % 25 == 0 and it <= 200
        assert rel(z, ref) < 1e-3
