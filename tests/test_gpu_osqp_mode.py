"""qp_mode 'osqp' (the reference's own QP regime: warm-started ADMM at eps 1e-3, src/osqp_solver.py:39-41,137-143) and the
joint-limit inequality rows against the C oracle's ADMM back-end (oracle/c/oracle.c), which
runs the same iteration through a banded LDL' of the quasi-definite KKT matrix.  All calls go through the C ABI."""
import numpy as np
import pytest

from oracle.cport import COracle
from tests.cases import make_cases
from tests.test_oracle_c import TIGHT_LIMITS, URDF_LIMITS

pytestmark = pytest.mark.gpu


def rel(a, b):
    return np.abs(np.asarray(a) - np.asarray(b)).max() / max(np.abs(np.asarray(b)).max(), 1e-300)


def _model_with(limits):
    import indy7_mpc_b200 as pkg
    model, _, _ = pkg.load_robot_model()
    if limits is not None:
        model.lowerPositionLimit, model.upperPositionLimit = np.array(limits[0]), np.array(limits[1])
        model.velocityLimit, model.effortLimit = np.array(limits[2]), np.array(limits[3])
    return model


def test_urdf_limits_parsed():
    m = _model_with(None)
    assert np.allclose(m.lowerPositionLimit, URDF_LIMITS[0]) and np.allclose(m.upperPositionLimit, URDF_LIMITS[1])
    assert np.allclose(m.velocityLimit, URDF_LIMITS[2]) and np.allclose(m.effortLimit, URDF_LIMITS[3])


@pytest.mark.parametrize("limits", [None, URDF_LIMITS, TIGHT_LIMITS])
def test_osqp_mode_qp_matches_oracle_admm(oracle_model, limits):
    """setup_and_solve_qp in 'osqp' mode: same ADMM iteration count as the oracle, iterate (x, y) within 1e-6, cold and
    warm-started, without limits (the reference's l == u QP) and with box rows."""
    from indy7_mpc_b200 import Engine
    xu, x0, goals = make_cases(oracle_model, 4, seed=11)
    eng = Engine(_model_with(limits), N=32, batch=4, qp_mode="osqp", limits=limits is not None, admm=dict(max_iter=500))
    orc = [COracle(oracle_model, backend="admm", limits=limits, max_iter=500) for _ in range(4)]
    for rep in range(2):                                          # second call: warm-started from the first
        z, y = eng.solve_qp(xu, x0, goals, want_duals=True)
        x_s, y_s, yl_s, info = eng.qp_state(4)
        for b in range(4):
            zr, it = orc[b].solve_qp(xu[b], x0[b], goals[b])
            xo, yo, ylo, res = orc[b].admm_state()
            assert int(info[b, 0]) == it, (rep, b, info[b], it)
            assert rel(z[b], zr) < 1e-6 and rel(x_s[b], xo) < 1e-6
            assert rel(y[b], yo) < 1e-6
            if limits is not None:
                assert np.abs(yl_s[b] - ylo).max() < 1e-6 * max(1.0, np.abs(ylo).max())
                if it < 500:                                  # converged lanes respect the bounds to OSQP's eps_prim
                    lo = np.tile(np.concatenate([limits[0], -limits[2], -limits[3]]), 32)[:zr.size][12:]
                    hi = np.tile(np.concatenate([limits[1], limits[2], limits[3]]), 32)[:zr.size][12:]
                    assert np.maximum(lo - z[b, 12:], z[b, 12:] - hi).max() <= 1e-3 * (1 + np.abs(z[b]).max())
            if it < 500:
                assert abs(info[b, 1] - res[0]) < 1e-6 * max(1.0, res[0]) + 1e-9
                assert info[b, 3] == 0
            else:
                assert info[b, 3] == 1


@pytest.mark.parametrize("limits", [None, URDF_LIMITS])
def test_osqp_mode_sqp_matches_oracle(oracle_model, limits):
    """Full sqp() in 'osqp' mode over two warm MPC-style calls: ADMM iterations per pass, alphas, qp_iters exact,
    trajectories within 1e-6 of the oracle (same algorithm end to end, incl. the alpha = 0 `continue` that re-solves the
    QP from the warm start, src/osqp_sqp.py:81-82)."""
    from indy7_mpc_b200 import Engine
    B = 12
    xu, x0, goals = make_cases(oracle_model, B, seed=4)
    eng = Engine(_model_with(limits), N=32, batch=B, qp_mode="osqp", limits=limits is not None, admm=dict(max_iter=300))
    orc = COracle(oracle_model, backend="admm", batch=B, limits=limits, max_iter=300)
    cur = xu
    for tick in range(2):
        out, st = eng.solve(cur, x0, goals)
        ref, rst, _ = orc.sqp(x0, goals, cur)
        for b in range(B):
            assert [int(st[b, 10]), int(st[b, 11])] == [int(rst[b, 8]), int(rst[b, 9])], (tick, b, st[b], rst[b])
            assert list(st[b, 4:6]) == list(rst[b, 4:6]), (tick, b, st[b], rst[b])
            assert list(st[b, 0:4]) == list(rst[b, 0:4]), (tick, b, st[b], rst[b])
            assert rel(out[b], ref[b]) < 1e-6
            assert np.allclose(st[b, 6:8], rst[b, 6:8], rtol=1e-6, atol=1e-9)
        cur = np.asarray(out, float)


def test_osqp_mode_reset_semantics(oracle_model):
    """i7_reset clears the warm start (the next solve repeats the cold one bit for bit); i7_reset_lambda clears the duals
    only (different iterate than either); exact mode ignores all three."""
    from indy7_mpc_b200 import Engine
    xu, x0, goals = make_cases(oracle_model, 2, seed=2)
    eng = Engine(N=32, batch=2, qp_mode="osqp")
    z0 = np.array(eng.solve_qp(xu, x0, goals))
    z1 = np.array(eng.solve_qp(xu, x0, goals))              # warm
    eng.reset()
    z2 = np.array(eng.solve_qp(xu, x0, goals))
    assert np.array_equal(z0, z2) and not np.array_equal(z0, z1)
    eng.reset_lambda()
    z3 = np.array(eng.solve_qp(xu, x0, goals))
    assert not np.array_equal(z3, z0)
    orc = COracle(oracle_model, backend="admm")
    zr, _ = orc.solve_qp(xu[0], x0[0], goals[0])
    assert rel(z0[0], zr) < 1e-6
    eng.reset_rho()


def test_osqp_mode_rejections():
    from indy7_mpc_b200 import Engine
    with pytest.raises(ValueError):
        Engine(N=32, batch=1, qp_mode="exact", limits=True)       # the exact solve is equality-only
    with pytest.raises(ValueError):
        Engine(N=32, batch=1, qp_mode="osqp", dtype="fp32")
    with pytest.raises(ValueError):
        Engine(N=32, batch=1, qp_mode="admm")


def test_osqp_mode_long_horizon_and_classes(oracle_model):
    """N = 128 (records of the linearisation in the global workspace) and the reference-shaped classes in 'osqp' mode."""
    import indy7_mpc_b200 as pkg
    from indy7_mpc_b200 import Engine
    xu, x0, goals = make_cases(oracle_model, 2, N=128, seed=8)
    eng = Engine(N=128, batch=2, qp_mode="osqp", admm=dict(max_iter=300))
    orc = COracle(oracle_model, N=128, backend="admm", batch=2, max_iter=300)
    out, st = eng.solve(xu, x0, goals)
    ref, rst, _ = orc.sqp(x0, goals, xu)
    for b in range(2):
        assert [int(st[b, 10]), int(st[b, 11])] == [int(rst[b, 8]), int(rst[b, 9])]
        assert list(st[b, 4:6]) == list(rst[b, 4:6]) and rel(out[b], ref[b]) < 1e-6
    model, _, _ = pkg.load_robot_model()
    solver = pkg.OSQPSolver(model, N=32, qp_mode="osqp")
    xu, x0, goals = make_cases(oracle_model, 1, seed=5)
    r = solver.setup_and_solve_qp(xu[0], x0[0], goals[0])
    o = COracle(oracle_model, backend="admm")
    zr, it = o.solve_qp(xu[0], x0[0], goals[0])
    assert r.info.iter == it and r.info.status == "solved" and rel(r.x, zr) < 1e-6
    sq = pkg.SQP_OSQP(pkg.OSQPSolver(model, N=32, qp_mode="osqp"))
    o2 = COracle(oracle_model, backend="admm")
    got = sq.sqp(x0[0], goals[0], xu[0])
    want, rst, _ = o2.sqp(x0[0], goals[0], xu[0])
    assert rel(got, want[0]) < 1e-6 and sq.stats["linesearch_alphas"]["values"] == list(rst[0, 4:4 + int(rst[0, 1])])


def test_osqp_mode_closed_loop_golden_prefix(golden):
    """The reference's stored closed-loop series (notebooks/pin_mpc_indy7.ipynb:98-597) was produced BY OSQP at eps 1e-3;
    MPC_OSQP.run_mpc over the 'osqp' QP mode (same regime: warm-started ADMM, same tolerances; no Ruiz scaling / adaptive
    rho) reproduces its first 12 values to 5e-6, as the CPU restatement does (tests/test_oracle_golden.py)."""
    import indy7_mpc_b200 as pkg
    g = golden["closed_loop"]
    model, _, _ = pkg.load_robot_model()
    solver = pkg.OSQPSolver(model, qp_mode="osqp")
    sqp = pkg.SQP_OSQP(solver)
    mpc = pkg.MPC_OSQP(model, sqp, solver)
    endpoints = np.array([solver.eepos(np.array(q)) for q in g["endpoint_qs"]])
    mpc.run_mpc(np.array(g["xstart"]), endpoints, num_steps=12)
    got, ref = np.array(mpc.goal_distances), np.array(g["goal_distances"][:12])
    assert (np.abs(got - ref) / ref).max() < 5e-6
    assert all(v == 2 for v in sqp.stats["qp_iters"]["values"])


def test_osqp_mode_step_tick(oracle_model):
    """Indy7SQP.step (bound buffers, x_0 of the warm start replaced inside the solve) in the osqp mode equals the generic
    path tick after tick."""
    import torch
    import bench
    import indy7_mpc_b200 as pkg
    B, N = 6, 32
    wl = bench.make_workload(B, N)
    fast = pkg.Indy7SQP(N=N, batch=B, qp_mode="osqp", admm=dict(max_iter=200))
    slow = pkg.Indy7SQP(N=N, batch=B, qp_mode="osqp", admm=dict(max_iter=200))
    p_x0 = torch.empty(B, 12, dtype=torch.float64).pin_memory()
    p_g = torch.empty(B, 3 * N, dtype=torch.float64).pin_memory()
    x0 = wl["x0"].copy()
    for t in range(3):
        goals = bench.goals_at(wl["fig"], wl["offsets"], t, N)
        p_x0.copy_(torch.from_numpy(x0)); p_g.copy_(torch.from_numpy(goals))
        a = fast.step(p_x0, p_g)
        b = slow.step(x0, goals)
        assert np.array_equal(a["costs"], b["costs"]) and a["best_index"] == b["best_index"]
        assert np.array_equal(a["x_next"], b["x_next"]) and np.array_equal(a["x_traj"], b["x_traj"])
        x0 = a["x_next"].copy()
