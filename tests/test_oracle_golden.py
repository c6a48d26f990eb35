"""Pin the CPU oracle to the reference's own golden vectors (stored notebook outputs)."""
import numpy as np
import pytest

from oracle import rbd
from oracle.osqp_mpc import MPC_OSQP
from oracle.osqp_solver import OSQPSolver
from oracle.osqp_sqp import SQP_OSQP


def test_fk_known_answers(golden, oracle_model):
    # 16 printed end-effector positions from three notebooks; tolerance = print precision
    for e in golden["fk_known_answers"]:
        p = rbd.eepos(oracle_model, np.array(e["q"]))
        assert np.abs(p - np.array(e["eepos"])).max() <= 6e-9, e


def test_dimensions_and_patterns(oracle_model):
    # sizes quoted in SURVEY.md §8 (verified there by running the reference's initialize_A/P under scipy)
    s = OSQPSolver(oracle_model)
    assert (s.nq, s.nv, s.nx, s.nu, s.nxu, s.traj_len) == (6, 6, 12, 6, 18, 570)
    assert s.A.shape == (384, 570) and s.A.nnz == 11304
    assert s.P.shape == (570, 570) and s.P.nnz == 1050
    s = OSQPSolver(oracle_model, N=128)
    assert s.traj_len == 2298 and s.A.shape == (1536, 2298) and s.A.nnz == 45864 and s.P.nnz == 4218


@pytest.mark.parametrize("backend,steps", [("kkt", 21), ("admm", 12)])
def test_closed_loop_series_prefix(golden, oracle_model, backend, steps):
    """First values of the 500-step goal-distance series of notebooks/pin_mpc_indy7.ipynb.
    Value 0 is pure FK; later values pin ABA, dABA, assembly, QP, line search, rk4 and the shift jointly.
    Beyond ~25 steps line-search branch flips make the series chaotic."""
    g = golden["closed_loop"]
    s = OSQPSolver(oracle_model, qp_backend=backend)
    sqp = SQP_OSQP(s)
    mpc = MPC_OSQP(oracle_model, sqp, s)
    endpoints = np.array([s.eepos(np.array(q)) for q in g["endpoint_qs"]])
    mpc.run_mpc(np.array(g["xstart"]), endpoints, num_steps=steps)
    got = np.array(mpc.goal_distances)
    ref = np.array(g["goal_distances"][:steps])
    rel = np.abs(got - ref) / ref
    assert rel[0] < 1e-15
    assert rel.max() < 5e-6, rel
    assert all(v == 2 for v in sqp.stats["qp_iters"]["values"])


def test_admm_meets_reference_tolerances(oracle_model):
    """The OSQP-style ADMM stops at eps_abs=eps_rel=1e-3 and lands near the exact KKT solution."""
    rng = np.random.default_rng(3)
    x0 = np.concatenate([rng.uniform(-1, 1, 6), rng.uniform(-0.5, 0.5, 6)])
    goals = np.tile(rbd.eepos(oracle_model, 0.3 * np.ones(6)), 32)
    XU = np.zeros(570); XU[:12] = x0
    exact = OSQPSolver(oracle_model, qp_backend="kkt").setup_and_solve_qp(XU, x0, goals)
    admm = OSQPSolver(oracle_model, qp_backend="admm").setup_and_solve_qp(XU, x0, goals)
    assert admm.info.status == "solved" and admm.info.iter % 25 == 0
    assert np.abs(admm.x - exact.x).max() / np.abs(exact.x).max() < 1e-3
