"""The plain-C oracle port (timed CPU baseline) must agree with the numpy oracle."""
import numpy as np
import pytest

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


URDF_LIMITS = (np.array([-3.0543261909900767] * 5 + [-3.7524578917878086]), np.array([3.0543261909900767] * 5 + [3.7524578917878086]), np.array([2.6179938779914944] * 3 + [3.141592653589793] * 3),
               np.array([431.97, 431.97, 197.23, 79.79, 79.79, 79.79]))        # description/indy7.urdf:203-238
TIGHT_LIMITS = (np.full(6, -2.0), np.full(6, 2.0), np.full(6, 0.8), np.array([60.0, 60.0, 30.0, 8.0, 8.0, 8.0]))


@pytest.mark.parametrize("limits", [None, URDF_LIMITS, TIGHT_LIMITS])
def test_c_admm_equals_numpy_admm_unscaled(oracle_model, limits):
    """The two restatements of the ADMM iteration (C: banded LDL' with the limit rows eliminated by hand; numpy: sparse LU
    of the full quasi-definite KKT with the limit rows in A) run the same algorithm: same iteration count, same iterate —
    without limits (the reference's l == u QP) and with the box rows of SURVEY.md section 8 f-4."""
    xu, x0, goals = make_cases(oracle_model, 4, seed=11)
    for b in (0, 1, 3):                                       # problem 2 of this seed is infeasible under the limits
        ca = COracle(oracle_model, backend="admm", limits=limits)
        ns = OSQPSolver(oracle_model, qp_backend="admm", admm_scaling=0, limits=limits)
        for rep in range(2):                                  # second call: warm-started
            z, it = ca.solve_qp(xu[b], x0[b], goals[b])
            r = ns.setup_and_solve_qp(xu[b], x0[b], goals[b])
            assert it == r.info.iter and r.info.status == "solved"
            assert rel(z, r.x) < 1e-8
            x, y, yl, res = ca.admm_state()
            assert rel(y, r.y[:12 * 32]) < 1e-6
            if limits is not None:
                assert rel(yl, r.y[12 * 32:]) < 1e-6 or np.abs(yl - r.y[12 * 32:]).max() < 1e-9
                lo = np.tile(np.concatenate([limits[0], -limits[2], -limits[3]]), 32)[:z.size][12:]
                hi = np.tile(np.concatenate([limits[1], limits[2], limits[3]]), 32)[:z.size][12:]
                viol = np.maximum(lo - z[12:], z[12:] - hi).max()
                assert viol <= 1e-3 * (1.0 + np.abs(z).max())        # OSQP's eps_prim = eps_abs + eps_rel max(|Ax|, |z|)


def test_c_admm_limits_bind(oracle_model):
    """With tight limits the solution differs from the equality-only one and the bounds are active."""
    xu, x0, goals = make_cases(oracle_model, 1, seed=11)
    free, _ = COracle(oracle_model, backend="admm").solve_qp(xu[0], x0[0], goals[0])
    lim, _ = COracle(oracle_model, backend="admm", limits=TIGHT_LIMITS).solve_qp(xu[0], x0[0], goals[0])
    u_free = np.concatenate([free, np.zeros(6)]).reshape(32, 18)[:, 12:]
    u_lim = np.concatenate([lim, np.zeros(6)]).reshape(32, 18)[:, 12:]
    assert np.abs(u_free).max(axis=0)[3] > 8.0 * 1.05          # the free solution violates the tight effort limit ...
    assert np.abs(u_lim).max() <= 60.0 * (1 + 1e-3) and np.abs(u_lim[:, 3:]).max() <= 8.0 * (1 + 2e-3)
    assert np.abs(u_lim[:, 3:]).max() > 8.0 * 0.99             # ... the limited one sits on it
