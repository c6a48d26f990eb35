"""Parity of the CUDA path (called through the C ABI) against the CPU oracle on identical seeded inputs.
Tolerances: fp64 trajectories 1e-6 relative (BASELINE.json north_star); kernel-level quantities much tighter."""
import numpy as np
import pytest

from oracle import rbd
from oracle.osqp_solver import OSQPSolver as OracleSolver
from oracle.osqp_sqp import SQP_OSQP as OracleSQP, ALPHAS
from tests.cases import make_cases

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.fixture(scope="module")
def pkg():
    import indy7_mpc_b200 as m
    return m


@pytest.fixture(scope="module")
def engine(pkg):
    return pkg.Engine(N=32, batch=64, dtype="fp64")


def test_fk_known_answers_gpu(golden, engine):
    q = np.array([e["q"] for e in golden["fk_known_answers"]])
    p = engine.eepos(q)
    ref = np.array([e["eepos"] for e in golden["fk_known_answers"]])
    assert np.abs(p - ref).max() <= 6e-9


def test_eepos_and_jacobian(engine, oracle_model):
    rng = np.random.default_rng(0)
    q = rng.uniform(-3, 3, (200, 6))
    p, J = engine.eepos(q, jacobian=True)
    pr, Jr = rbd.d_eepos(oracle_model, q)
    assert np.abs(p - pr).max() < 1e-14 and np.abs(J - Jr).max() < 1e-14


@pytest.mark.parametrize("with_fext", [False, True])
def test_dynamics_and_derivatives(engine, oracle_model, with_fext):
    rng = np.random.default_rng(1)
    n = 300
    q, v, u = rng.uniform(-3, 3, (n, 6)), rng.uniform(-2.6, 2.6, (n, 6)), rng.uniform(-100, 100, (n, 6))
    f = rng.uniform(-30, 30, (n, 6)) if with_fext else None
    qdd, dq, dv, mi = engine.dynamics(q, v, u, f)
    rq, rv, rm, ra = rbd.aba_derivatives(oracle_model, q, v, u, f)
    assert rel(qdd, ra) < 1e-11
    assert rel(mi, rm) < 1e-11
    assert rel(dq, rq) < 1e-10
    assert rel(dv, rv) < 1e-10


def test_qp_data_reference_order(engine, oracle_model):
    xu, x0, goals = make_cases(oracle_model, 6, seed=2)
    A, Pd, g, l = engine.dump_qp(xu, x0, goals)
    for b in range(6):
        o = OracleSolver(oracle_model)
        o.update_constraint_matrix(xu[b], x0[b])
        o.update_cost_matrix(xu[b], goals[b])
        assert rel(A[b], o.Adata) < 1e-10
        assert rel(l[b], o.l) < 1e-10
        assert rel(Pd[b], o.Pdata) < 1e-11
        assert rel(g[b], o.g) < 1e-11


def test_qp_solution_matches_exact_kkt(engine, oracle_model):
    xu, x0, goals = make_cases(oracle_model, 8, seed=3)
    z = engine.solve_qp(xu, x0, goals)
    for b in range(8):
        ref = OracleSolver(oracle_model, qp_backend="kkt").setup_and_solve_qp(xu[b], x0[b], goals[b]).x
        assert rel(z[b], ref) < 1e-6, (b, rel(z[b], ref))
        # residuals of the equality constraints and stationarity far below the reference eps_abs = 1e-3
        o = OracleSolver(oracle_model)
        o.update_constraint_matrix(xu[b], x0[b]); o.update_cost_matrix(xu[b], goals[b])
        P, A = o.matrices()
        assert np.abs(A @ z[b] - o.l).max() < 1e-8


def test_merit_terms(engine, oracle_model):
    xu, x0, goals = make_cases(oracle_model, 6, seed=4)
    m = engine.merit(xu, xu, goals)
    for b in range(6):
        sq = OracleSQP(OracleSolver(oracle_model))
        qc, vc, uc = sq.eepos_cost(goals[b], xu[b])
        ie = sq.integrator_err(xu[b])
        assert rel(m[b, :4], [qc, vc, uc, ie]) < 1e-12
        assert abs(m[b, 5] - sq.merit(xu[b], xu[b], goals[b])) < 1e-10 * abs(m[b, 5])


def test_linesearch(engine, oracle_model):
    xu, x0, goals = make_cases(oracle_model, 8, seed=5)
    for b in range(8):
        o = OracleSolver(oracle_model, qp_backend="kkt")
        z = o.setup_and_solve_qp(xu[b], x0[b], goals[b]).x
        sq = OracleSQP(o)
        a_ref = sq.linesearch(xu[b], z, goals[b])
        alpha, merits = engine.linesearch(xu[b], z, goals[b])
        base, tried = sq.last_merits
        assert abs(merits[0, 0] - base) < 1e-11 * abs(base)
        for i, m in enumerate(tried):
            assert abs(merits[0, 1 + i] - m) < 1e-11 * abs(m)
        assert alpha[0] == a_ref


def _oracle_sqp(model, xu, x0, goals, **kw):
    sq = OracleSQP(OracleSolver(model, qp_backend="kkt", **kw))
    out = sq.sqp(x0, goals, xu)
    return out, sq.stats


def test_sqp_solve_matches_oracle(engine, oracle_model):
    xu, x0, goals = make_cases(oracle_model, 12, seed=6)
    out, st = engine.solve(xu, x0, goals)
    for b in range(12):
        ref, stats = _oracle_sqp(oracle_model, xu[b], x0[b], goals[b])
        assert int(st[b, 0]) == stats["qp_iters"]["values"][0]
        n_ls = int(st[b, 1])
        assert list(st[b, 4:4 + n_ls]) == stats["linesearch_alphas"]["values"], b
        assert rel(out[b], ref) < 1e-6, (b, rel(out[b], ref))
        n_st = int(st[b, 2])
        assert n_st == len(stats["sqp_stepsizes"]["values"])
        assert rel(st[b, 6:6 + n_st], stats["sqp_stepsizes"]["values"]) < 1e-6 if n_st else True


def test_reference_shaped_classes(pkg, oracle_model):
    """OSQPSolver / SQP_OSQP used exactly as notebooks/pin_mpc_indy7.ipynb cell 0-1 does."""
    model, _, _ = pkg.load_robot_model(pkg.DEFAULT_URDF, None)
    solver = pkg.OSQPSolver(model)
    sqp = pkg.SQP_OSQP(solver)
    assert (solver.nq, solver.nv, solver.nx, solver.nu, solver.traj_len) == (6, 6, 12, 6, 570)
    xs = np.ones(12)
    goal = np.tile(solver.eepos(0.3 * np.ones(6)), 32)
    XU0 = np.zeros(570)
    XU = sqp.sqp(xs, goal, XU0)
    assert np.all(XU0 == 0)                         # inputs are never modified
    ref, stats = _oracle_sqp(oracle_model, XU0, xs, goal)
    assert rel(XU, ref) < 1e-6
    assert sqp.get_stats()["qp_iters"]["values"] == stats["qp_iters"]["values"]
    assert sqp.get_stats()["linesearch_alphas"]["values"] == stats["linesearch_alphas"]["values"]
    sol = solver.setup_and_solve_qp(XU0, xs, goal)
    assert sol.x.shape == (570,)
    with pytest.raises(ValueError):
        sqp.sqp(xs, goal[:-3], XU0)


def test_closed_loop_golden_prefix(pkg, golden):
    """The reference's stored 500-step series (notebooks/pin_mpc_indy7.ipynb cell 2): first 21 values."""
    g = golden["closed_loop"]
    model, _, _ = pkg.load_robot_model()
    solver = pkg.OSQPSolver(model)
    mpc = pkg.MPC_OSQP(model, pkg.SQP_OSQP(solver), solver)
    endpoints = np.array([solver.eepos(np.array(q)) for q in g["endpoint_qs"]])
    mpc.run_mpc(np.array(g["xstart"]), endpoints, num_steps=21)
    got, ref = np.array(mpc.goal_distances), np.array(g["goal_distances"][:21])
    assert np.abs(got - ref).max() / 1.0 < 5e-6 and (np.abs(got - ref) / ref).max() < 5e-6


def test_batch_determinism_and_independence(engine, oracle_model):
    """Identical lanes -> identical outputs (src/gato_mpc_batch.py:124-134), and a lane's result does not depend
    on what else is in the batch."""
    xu, x0, goals = make_cases(oracle_model, 5, seed=7)
    rep = lambda a: np.repeat(a[:1], 64, axis=0)
    out, st = engine.solve(rep(xu), rep(x0), rep(goals))
    assert np.all(out == out[0]) and np.all(st == st[0])
    mixed, _ = engine.solve(xu, x0, goals)
    for b in range(5):
        single, _ = engine.solve(xu[b], x0[b], goals[b])
        assert np.array_equal(single[0], mixed[b])


def test_external_wrench(pkg, oracle_model):
    xu, x0, goals = make_cases(oracle_model, 3, seed=8)
    f = np.array([[5.0, 0, 15, 0, 0, 0], [-6, 2, -4, 0.5, 0, -0.3], [0, 0, 0, 0, 0, 0]])
    eng = pkg.Engine(N=32, batch=3)
    eng.set_external_wrench(f)
    out, st = eng.solve(xu, x0, goals)
    for b in range(3):
        ref, _ = _oracle_sqp(oracle_model, xu[b], x0[b], goals[b], f_ext=f[b])
        assert rel(out[b], ref) < 1e-6
    # plant step under the same wrenches (src/utils.py:3-18)
    xn = eng.sim_forward(x0, xu[:, 12:18], 0.01)
    for b in range(3):
        qn, vn = rbd.rk4(oracle_model, x0[b, :6], x0[b, 6:], xu[b, 12:18], 0.01, f[b])
        assert rel(xn[b], np.concatenate([qn, vn])) < 1e-12


def test_goal_stride_six_and_facade(pkg, oracle_model):
    xu, x0, goals = make_cases(oracle_model, 4, seed=9)
    g6 = np.zeros((4, 32, 6)); g6[:, :, :3] = goals.reshape(4, 32, 3)
    s = pkg.Indy7SQP(N=32, batch=4)
    x_traj, u_traj, stats = s.solve(x0, g6.reshape(4, -1), xu_init=xu)
    assert x_traj.shape == (4, 32, 12) and u_traj.shape == (4, 31, 6)
    eng = pkg.Engine(N=32, batch=4)
    ref, _ = eng.solve(xu, x0, goals)
    assert np.array_equal(stats["xu"], ref)
    x1, u1, st1 = s.solve(x0[0], goals[0].reshape(32, 3), xu_init=xu[0])
    assert x1.shape == (32, 12) and u1.shape == (31, 6) and np.array_equal(st1["xu"], ref[0])


def test_batch_sqp_shim(oracle_model):
    import bindings.batch_sqp as batch_sqp
    solver = batch_sqp.SQPSolverfloat_4()
    xu, x0, goals = make_cases(oracle_model, 4, seed=10)
    g6 = np.zeros((4, 32, 6)); g6[:, :, :3] = goals.reshape(4, 32, 3)
    f = np.zeros((4, 6)); f[1:, :3] = np.random.default_rng(0).normal(0, 5, (3, 3))
    solver.set_external_wrench_batch(f)
    solver.reset(); solver.resetRho(); solver.resetLambda()
    res = solver.solve(xu, 0.01, x0, g6.reshape(4, -1))
    assert res["xu_trajectory"].shape == (4, 570) and res["xu_trajectory"].dtype == np.float32
    assert res["sqp_iterations"].shape == (4,) and res["solve_time_us"] > 0
    assert len(res["line_search_stats"]) == len(res["pcg_stats"]) and len(res["line_search_stats"][0]["step_size"]) == 4
    for b in range(4):
        ref, _ = _oracle_sqp(oracle_model, xu[b], x0[b], goals[b], f_ext=f[b])
        assert rel(res["xu_trajectory"][b], ref) < 1e-5      # float32 output of the fp64 computation
    xn = solver.sim_forward(x0[0], xu[0, 12:18], 0.01)
    assert xn.shape == (4, 12)
    with pytest.raises(ValueError):
        solver.solve(xu[:2], 0.01, x0[:2], g6[:2].reshape(2, -1))


def test_fp32_mode(pkg, oracle_model):
    """fp32 arithmetic mode: trajectories within 1e-4 relative of the fp64 oracle (north_star tolerance)."""
    xu, x0, goals = make_cases(oracle_model, 8, seed=11)
    eng = pkg.Engine(N=32, batch=8, dtype="fp32")
    z = eng.solve_qp(xu, x0, goals)
    errs = []
    for b in range(8):
        ref = OracleSolver(oracle_model, qp_backend="kkt").setup_and_solve_qp(xu[b], x0[b], goals[b]).x
        errs.append(rel(z[b], ref))
    assert max(errs) < 1e-4, errs


def test_torch_cuda_tensors(pkg, oracle_model):
    import torch
    xu, x0, goals = make_cases(oracle_model, 4, seed=12)
    eng = pkg.Engine(N=32, batch=4)
    ref, _ = eng.solve(xu, x0, goals)
    dev = torch.device("cuda:0")
    out, st = eng.solve(torch.tensor(xu, device=dev), torch.tensor(x0, device=dev), torch.tensor(goals, device=dev))
    assert out.is_cuda and np.array_equal(out.cpu().numpy(), ref)


def test_long_horizon(pkg, oracle_model):
    """N = 128 (the horizon of the reference's C++ node, trajopt_node.cu:178); fp64 spills gains to global memory."""
    xu, x0, goals = make_cases(oracle_model, 2, N=128, seed=13)
    eng = pkg.Engine(N=128, batch=2)
    z = eng.solve_qp(xu, x0, goals)
    for b in range(2):
        ref = OracleSolver(oracle_model, N=128, qp_backend="kkt").setup_and_solve_qp(xu[b], x0[b], goals[b]).x
        assert rel(z[b], ref) < 1e-6


def test_error_behaviour(pkg):
    with pytest.raises(ValueError):
        pkg.Engine(N=1)
    eng = pkg.Engine(N=32, batch=2)
    with pytest.raises(ValueError):
        eng.solve(np.zeros((3, 570)), np.zeros((3, 12)), np.zeros((3, 96)))
    with pytest.raises(ValueError):
        eng.solve(np.zeros((1, 570)), np.zeros((1, 12)), np.zeros((1, 100)))


def test_best_selection_kernels(engine, oracle_model):
    """i7_pack_best / i7_reduce_best against numpy argmin (costs = stats[:, 8])."""
    xu, x0, goals = make_cases(oracle_model, 16, seed=14)
    out, st = engine.solve(xu, x0, goals)
    packed = engine.pack_best(st, out, index_offset=100)
    bi = int(np.argmin(st[:, 8]))
    assert np.array_equal(packed[:16], st[:, 8])
    assert packed[16] == st[bi, 8] and packed[17] == 100 + bi and np.array_equal(packed[18:], out[bi])
    # the cost is the merit the last line search computed for the returned trajectory
    for b in range(4):
        sq = OracleSQP(OracleSolver(oracle_model, qp_backend="kkt"))
        sq.sqp(x0[b], goals[b], xu[b])
        base, tried = sq.last_merits
        want = tried[-1] if st[b, 5] > 0 or (st[b, 1] == 1 and st[b, 4] > 0) else base
        assert abs(st[b, 8] - want) < 1e-9 * abs(want)
    two = np.stack([packed, packed.copy()])
    two[1, 16] = packed[16] - 1.0; two[1, 17] = 7.0
    red = engine.reduce_best(two, 2)
    assert red[0] == packed[16] - 1.0 and red[1] == 7.0



# ---- the cluster-per-problem (latency) form of i7_solve: 8 CTAs per problem --------------------------------------
def test_cluster_kernel_matches_oracle(pkg, oracle_model):
    xu, x0, goals = make_cases(oracle_model, 6, seed=21)
    f = np.zeros((6, 6)); f[1] = [5.0, 0, 15, 0, 0, 0]; f[4] = [-6, 2, -4, 0.5, 0, -0.3]
    eng = pkg.Engine(N=32, batch=6, cluster=8)
    eng.set_external_wrench(f)
    out, st = eng.solve(xu, x0, goals)
    for b in range(6):
        ref, stats = _oracle_sqp(oracle_model, xu[b], x0[b], goals[b], f_ext=f[b])
        assert int(st[b, 0]) == stats["qp_iters"]["values"][0]
        n_ls = int(st[b, 1])
        assert list(st[b, 4:4 + n_ls]) == stats["linesearch_alphas"]["values"], b
        assert rel(out[b], ref) < 1e-6, (b, rel(out[b], ref))
        n_st = int(st[b, 2])
        assert rel(st[b, 6:6 + n_st], stats["sqp_stepsizes"]["values"]) < 1e-6 if n_st else True


def test_cluster_kernel_equals_cta_kernel(pkg, oracle_model):
    """Same device functions per knot / per stage: the two kernel forms agree to rounding, with identical
    line-search decisions; repeated cluster solves are bit-identical (fixed reduction order)."""
    xu, x0, goals = make_cases(oracle_model, 16, seed=22)
    a = pkg.Engine(N=32, batch=16, cluster=8)
    c = pkg.Engine(N=32, batch=16, cluster=1, threads=384)
    oa, sa = a.solve(xu, x0, goals)
    oc, sc = c.solve(xu, x0, goals)
    assert np.array_equal(sa[:, :6], sc[:, :6])                  # iteration counts, status, alphas
    assert rel(oa, oc) < 1e-11
    oa2, sa2 = a.solve(xu, x0, goals)
    assert np.array_equal(oa, oa2) and np.array_equal(sa, sa2)
    # the auto choice at batch 1 is the cluster form and returns the same trajectory
    one = pkg.Engine(N=32, batch=1)
    o1, _ = one.solve(xu[3], x0[3], goals[3])
    assert np.array_equal(o1[0], oa[3])


@pytest.mark.parametrize("N,dtype,tol", [(8, "fp64", 1e-6), (5, "fp64", 1e-6), (64, "fp64", 1e-6), (128, "fp64", 1e-6),
                                         (32, "fp32", 1e-4)])
def test_cluster_kernel_horizons_and_fp32(pkg, oracle_model, N, dtype, tol):
    xu, x0, goals = make_cases(oracle_model, 2, N=N, seed=23)
    eng = pkg.Engine(N=N, batch=2, dtype=dtype, cluster=8)
    out, st = eng.solve(xu, x0, goals)
    cta = pkg.Engine(N=N, batch=2, dtype="fp64", cluster=1)
    ref, sr = cta.solve(xu, x0, goals)
    if dtype == "fp64":
        assert np.array_equal(np.asarray(st)[:, :6], np.asarray(sr)[:, :6])
        assert rel(out, ref) < 1e-10
    if N <= 32:
        for b in range(2):
            want, _ = _oracle_sqp(oracle_model, xu[b], x0[b], goals[b], N=N)
            assert rel(out[b], want) < tol, (b, rel(out[b], want))


@pytest.mark.parametrize("cl,N,dtype", [(4, 32, "fp64"), (2, 32, "fp64"), (4, 13, "fp64"), (2, 64, "fp64"),
                                        (4, 128, "fp64"), (4, 32, "fp32"), (2, 32, "fp32")])
def test_smaller_clusters_equal_cta_kernel(pkg, oracle_model, cl, N, dtype):
    """Clusters of 4 and 2 CTAs per problem (mid-size batches): same decisions and trajectories as the CTA form."""
    xu, x0, goals = make_cases(oracle_model, 6, N=N, seed=31)
    fext = np.zeros((6, 6)); fext[1:, :3] = np.random.default_rng(4).normal(0, 3, (5, 3))
    eng = pkg.Engine(N=N, batch=6, dtype=dtype, cluster=cl)
    cta = pkg.Engine(N=N, batch=6, dtype=dtype, cluster=1)
    for e in (eng, cta):
        e.set_external_wrench(fext)
    out, st = eng.solve(xu, x0, goals)
    ref, sr = cta.solve(xu, x0, goals)
    if dtype == "fp64":
        assert np.array_equal(np.asarray(st)[:, :6], np.asarray(sr)[:, :6])
        assert rel(out, ref) < 1e-10
    else:
        assert rel(out, ref) < 2e-4
    out2, st2 = eng.solve(xu, x0, goals)
    assert np.array_equal(np.asarray(out), np.asarray(out2))


def test_batch_sample_mpc_driver_matches_oracle(pkg, oracle_model):
    """Closed loop of the batched sampled-wrench driver (src/gato_mpc_batch_sample.py:106-300): same seeds, same
    hypothesis selected at every step, same plant path as the CPU restatement (float32 hand-over included)."""
    from indy7_mpc_b200.gato_mpc_batch_sample import MPC_GATO_Batch_Sample
    from oracle.gato_mpc_batch_sample import MPC_GATO_Batch_Sample as OracleDriver
    from tests.cases import HOME
    x0 = np.concatenate([HOME, np.zeros(6)])
    endpoints = [np.array([0.3, 0.4, 0.5]), np.array([0.1, 0.5, 0.3])]
    kw = dict(N=32, dt=0.01, batch_size=8, constant_f_ext=[4.0, -6.0, 3.0], resample_f_ext=True, f_ext_std=4.0)
    np.random.seed(7)
    ref = OracleDriver(oracle_model, **kw)
    xp_ref, st_ref = ref.run_mpc(x0, endpoints, sim_dt=0.002, sim_time=0.05)
    np.random.seed(7)
    model, _, _ = pkg.load_robot_model()
    drv = MPC_GATO_Batch_Sample(model, **kw)
    best = []
    orig = drv.find_best_trajectory_id
    drv.find_best_trajectory_id = lambda *a: best.append(orig(*a)) or best[-1]
    xp, st = drv.run_mpc(x0, endpoints, sim_dt=0.002, sim_time=0.05)
    assert len(xp) == len(xp_ref) and len(xp) >= 25
    assert best == ref.best_ids
    # the selection itself (in the closed loop above the zero-wrench lane wins, as with the reference's own maths: a
    # world-origin wrench has a huge moment about the wrist): measured state produced by hypothesis 5 -> lane 5
    drv.resample_f_ext = ref.resample_f_ext = False
    rng = np.random.default_rng(5)
    xl = np.concatenate([HOME + rng.normal(0, 0.2, 6), rng.normal(0, 0.3, 6)])
    ul = rng.normal(0, 5, 6)
    p6, R6 = (a[-1] for a in rbd.forward_kinematics(oracle_model, xl[:6]))
    f5 = ref.f_ext_batch[5]
    assert np.array_equal(f5, drv.solver.f_ext_batch[5])
    loc = np.concatenate([R6.T @ f5[:3], R6.T @ (f5[3:] - np.cross(p6, f5[:3]))])
    xc = np.hstack(rbd.rk4(oracle_model, xl[:6], xl[6:], ul, 0.002, loc))
    assert ref.find_best_trajectory_id(xl, ul, xc, 0.002) == 5 and orig(xl, ul, xc, 0.002) == 5
    assert rel(np.array(xp), np.array(xp_ref)) < 1e-6
    assert np.allclose(st['goal_distances'], st_ref['goal_distances'], atol=2e-5)
    s = drv.solver.get_stats()
    assert len(s['solve_time']['values']) == len(best) + 1 and all(np.all(v >= 1) for v in s['sqp_iters']['values'])


def test_sharded_batch_is_bitwise_identical(pkg, oracle_model):
    """the batch solved as two contiguous shards (two handles, as two ranks would) + pack/reduce gives
    bitwise the same per-problem solutions and the same selected trajectory as the whole batch on one handle."""
    xu, x0, goals = make_cases(oracle_model, 24, N=32, seed=11)
    whole = pkg.Engine(N=32, batch=24, dtype="fp64", cluster=1)
    out, st = whole.solve(xu, x0, goals)
    best = np.asarray(whole.reduce_best(np.asarray(whole.pack_best(st, out))[None, :], 1))
    parts = []
    for r in range(2):
        sl = slice(12 * r, 12 * r + 12)
        e = pkg.Engine(N=32, batch=12, dtype="fp64", cluster=1)
        o, s = e.solve(xu[sl], x0[sl], goals[sl])
        assert np.array_equal(np.asarray(o), np.asarray(out)[sl])
        parts.append(np.asarray(e.pack_best(s, o, index_offset=12 * r)))
    gathered = np.stack(parts)
    best2 = np.asarray(whole.reduce_best(gathered, 2))
    assert np.array_equal(best, best2)


def test_host_pipeline_is_bitwise_identical(pkg, oracle_model, monkeypatch):
    """Host-buffer solves of big batches: pageable buffers run as a chunked copy/compute pipeline (i7_api.cu:
    run_solve_pipelined; a ragged last chunk, per-lane wrenches), page-locked buffers are read and written by the
    kernel itself in one launch (zero-copy) or, with I7_ZERO_COPY=0, through the pipeline — all bitwise equal to the
    single-launch device-resident solve."""
    import torch
    xu1, x01, g1 = make_cases(oracle_model, 40, N=16, seed=21)
    reps = 23
    xu, x0, goals = (np.tile(a, (reps, 1))[:900] for a in (xu1, x01, g1))
    rng = np.random.default_rng(2)
    x0 = x0 + rng.normal(0, 1e-3, x0.shape)
    fext = np.zeros((900, 6)); fext[:, :3] = rng.normal(0, 3, (900, 3))
    monkeypatch.setenv("I7_PIPE_CHUNK", "0")
    plain = pkg.Engine(N=16, batch=900, dtype="fp64")
    monkeypatch.setenv("I7_PIPE_CHUNK", "128")            # 8 chunks, the last one of 4 problems
    piped = pkg.Engine(N=16, batch=900, dtype="fp64")
    for e in (plain, piped):
        e.set_external_wrench(fext)
    dev = lambda a: torch.tensor(a, device="cuda")
    ref, ref_st = plain.solve(dev(xu), dev(x0), dev(goals))
    ref, ref_st = ref.cpu().numpy(), ref_st.cpu().numpy()
    n0 = piped.launch_count()
    out, st = piped.solve(xu, x0, goals)                               # pageable host buffers
    assert piped.launch_count() - n0 == 8
    assert np.array_equal(np.asarray(out), ref) and np.array_equal(np.asarray(st), ref_st)
    pin = lambda a: torch.tensor(a).pin_memory()
    n0 = piped.launch_count()
    out2, st2 = piped.solve(pin(xu), pin(x0), pin(goals))              # page-locked host buffers: zero-copy, ONE launch
    assert piped.launch_count() - n0 == 1
    assert np.array_equal(out2.numpy(), ref) and np.array_equal(st2.numpy(), ref_st)
    monkeypatch.setenv("I7_ZERO_COPY", "0")                            # page-locked buffers through the chunked pipeline
    piped2 = pkg.Engine(N=16, batch=900, dtype="fp64")
    piped2.set_external_wrench(fext)
    n0 = piped2.launch_count()
    out4, st4 = piped2.solve(pin(xu), pin(x0), pin(goals))
    assert piped2.launch_count() - n0 == 8
    assert np.array_equal(out4.numpy(), ref) and np.array_equal(st4.numpy(), ref_st)
    monkeypatch.setenv("I7_PIPE_MODE", "1")                            # the one-stream-per-chunk form of the pipeline
    piped3 = pkg.Engine(N=16, batch=900, dtype="fp64")
    piped3.set_external_wrench(fext)
    out5, st5 = piped3.solve(xu, x0, goals)
    assert np.array_equal(np.asarray(out5), ref) and np.array_equal(np.asarray(st5), ref_st)
    out3, _ = plain.solve(xu, x0, goals)                               # pipeline off: one launch
    assert np.array_equal(np.asarray(out3), ref)


def test_ragged_batch_nonfinite_and_failed_linesearch(pkg, oracle_model):
    """Edge cases of sqp(): fewer problems than the handle was created for; a non-finite lane (status 2) does not
    disturb its neighbours; an iterate the line search cannot improve returns unchanged with alpha = 0 and
    status 1, as src/osqp_sqp.py:72,81-82 (not an error)."""
    xu, x0, goals = make_cases(oracle_model, 5, N=32, seed=41)
    eng = pkg.Engine(N=32, batch=8, dtype="fp64", cluster=1)
    full, st_full = (np.asarray(a) for a in eng.solve(xu, x0, goals))
    part, st_part = (np.asarray(a) for a in eng.solve(xu[:3], x0[:3], goals[:3]))        # ragged: 3 of 8
    assert part.shape == (3, 570) and np.array_equal(part, full[:3]) and np.array_equal(st_part, st_full[:3])
    bad = xu.copy(); bad[2, 40] = np.nan
    out, st = (np.asarray(a) for a in eng.solve(bad, x0, goals))
    assert st[2, 3] == 2 and np.array_equal(out[[0, 1, 3, 4]], full[[0, 1, 3, 4]])
    # a converged iterate: solve until the accepted step vanishes, then once more
    z, g, x = xu[1:2], goals[1:2], x0[1:2]
    for _ in range(60):
        z, s = (np.asarray(a) for a in eng.solve(z, x, g))
        if s[0, 3] == 1:
            break
    if s[0, 3] == 1:                                     # line search failed: alpha = 0 recorded, iterate unchanged
        again, s2 = (np.asarray(a) for a in eng.solve(z, x, g))
        assert s2[0, 3] == 1 and s2[0, 4] == 0 and np.array_equal(again, z)
    from oracle.osqp_solver import OSQPSolver as OS
    from oracle.osqp_sqp import SQP_OSQP as OQ
    ref = OQ(OS(oracle_model, N=32, qp_backend="kkt")).sqp(x[0], g[0], z[0])
    assert rel(np.asarray(eng.solve(z, x, g)[0])[0], ref) < 1e-6


def test_mapped_host_memory_path_is_bitwise_identical(pkg, oracle_model, monkeypatch):
    """Host-buffer solves of up to 8 problems read / write mapped page-locked memory from the kernel (no copy
    operations in the stream): same bits as the copy path, for pageable and page-locked callers."""
    import torch
    xu, x0, goals = make_cases(oracle_model, 6, N=32, seed=51)
    monkeypatch.setenv("I7_DIRECT_MAX_BATCH", "0")
    copy = pkg.Engine(N=32, batch=6, dtype="fp64")
    monkeypatch.setenv("I7_DIRECT_MAX_BATCH", "8")
    direct = pkg.Engine(N=32, batch=6, dtype="fp64")
    ref, ref_st = (np.asarray(a) for a in copy.solve(xu, x0, goals))
    out, st = (np.asarray(a) for a in direct.solve(xu, x0, goals))
    assert np.array_equal(out, ref) and np.array_equal(st, ref_st)
    pin = lambda a: torch.tensor(a).pin_memory()
    out2, st2 = direct.solve(pin(xu), pin(x0), pin(goals))
    assert np.array_equal(out2.numpy(), ref) and np.array_equal(st2.numpy(), ref_st)
    one, _ = direct.solve(xu[4], x0[4], goals[4])
    assert np.array_equal(np.asarray(one)[0], ref[4])


def test_full_size_batch_properties(pkg, oracle_model):
    """BASELINE.json configs[3] at full size (4096 problems, N=32, the bench workload) through size-independent
    properties: bitwise repeatability, invariance under a permutation of the problems, the linearised dynamics of
    the QP solution satisfied far below the reference's eps_abs = eps_rel = 1e-3, no non-finite lane — plus 24 lanes
    spread over the batch against the CPU oracle (exact-KKT back-end, C port) on two consecutive MPC ticks."""
    import torch
    import bench
    from oracle.cport import COracle
    B, N = 4096, 32
    wl = bench.make_workload(B, N)
    eng = pkg.Engine(N=N, batch=B, dtype="fp64")
    eng.set_external_wrench(wl["f_ext"])
    dev = lambda a: torch.tensor(a, device="cuda")
    goals = bench.goals_at(wl["fig"], wl["offsets"], 0, N)
    out, st = (a.cpu().numpy() for a in eng.solve(dev(wl["xu"]), dev(wl["x0"]), dev(goals)))
    out2, st2 = (a.cpu().numpy() for a in eng.solve(dev(wl["xu"]), dev(wl["x0"]), dev(goals)))
    assert np.array_equal(out, out2) and np.array_equal(st, st2)
    assert np.all(np.isfinite(out)) and np.all(st[:, 3] <= 1)
    perm = np.random.default_rng(0).permutation(B)
    eng.set_external_wrench(wl["f_ext"][perm])
    outp, _ = (a.cpu().numpy() for a in eng.solve(dev(wl["xu"][perm]), dev(wl["x0"][perm]), dev(goals[perm])))
    assert np.array_equal(outp, out[perm])
    eng.set_external_wrench(wl["f_ext"])
    # equality constraints of the QP at the first iterate: z satisfies x_{k+1} = A_k x_k + B_k u_k + c_k, x_0 = xs
    lanes = np.arange(0, B, 171)[:24]
    small = pkg.Engine(N=N, batch=len(lanes), dtype="fp64")
    small.set_external_wrench(wl["f_ext"][lanes])
    z = np.asarray(small.solve_qp(wl["xu"][lanes], wl["x0"][lanes], goals[lanes]))
    Ad, Pd, g, l = (np.asarray(a) for a in small.dump_qp(wl["xu"][lanes], wl["x0"][lanes], goals[lanes]))
    from oracle.osqp_solver import OSQPSolver as OS
    tmpl = OS(oracle_model, N=N, qp_backend="kkt")
    for i in range(len(lanes)):
        A = tmpl.A.copy(); A.data[:] = Ad[i]
        res = A @ z[i] - l[i]
        assert np.abs(res).max() <= 1e-9 * max(1.0, np.abs(l[i]).max())
    # two ticks against the oracle
    orc = COracle(oracle_model, N=N, backend="kkt", batch=len(lanes))
    xu_t, x0_t = wl["xu"][lanes], wl["x0"][lanes]
    for t in range(2):
        g_t = bench.goals_at(wl["fig"], wl["offsets"][lanes], t, N)
        want, _, _ = orc.sqp(x0_t, g_t, xu_t, f_ext=wl["f_ext"][lanes])
        got = out[lanes] if t == 0 else np.asarray(small.solve(xu_t, x0_t, g_t)[0])
        assert rel(got, want) < 1e-6
        xu_t, x0_t = bench.shift_host(want, N)
        xu_t[:, :12] = x0_t


def test_long_horizon_full_size_fp32_tolerance(pkg):
    """BASELINE.json configs[4] at full size (N=128, 512 problems): fp32 mode within the stated 1e-4 relative of the
    fp64 mode on identical inputs; both repeatable."""
    import torch
    import bench
    B, N = 512, 128
    wl = bench.make_workload(B, N)
    goals = bench.goals_at(wl["fig"], wl["offsets"], 0, N)
    outs = {}
    for dt, td in (("fp64", torch.float64), ("fp32", torch.float32)):
        eng = pkg.Engine(N=N, batch=B, dtype=dt)
        eng.set_external_wrench(wl["f_ext"])
        a = [torch.tensor(x, dtype=td, device="cuda") for x in (wl["xu"], wl["x0"], goals)]
        o, s = eng.solve(*a)
        o2, _ = eng.solve(*a)
        assert torch.equal(o, o2) and bool(torch.isfinite(o).all())
        outs[dt] = o.double().cpu().numpy()
    assert rel(outs["fp32"], outs["fp64"]) < 1e-4


def test_fig8_batch_sample_driver_matches_oracle(pkg, oracle_model):
    """notebooks/gato_mpc_indy7_fig8.ipynb cell 2 (run_mpc_fig8): figure-8 tracking with 8 wrench hypotheses,
    selection through the solver's sim_forward, resampling around the best lane: same lane picked at every step and
    the same plant path as the CPU restatement."""
    from indy7_mpc_b200.gato_mpc_batch_sample import MPC_GATO_Batch_Sample
    from indy7_mpc_b200.reference_traj import figure8
    from oracle.gato_mpc_batch_sample import MPC_GATO_Batch_Sample as OracleDriver
    fig = figure8(0.4, 0.4, (0.0, 0.5, 0.75), 5.0, 0.01, cycles=5)
    kw = dict(N=32, dt=0.01, batch_size=8, constant_f_ext=[10.0, 15.0, -6.0], resample_f_ext=True, f_ext_std=5.0,
              f_ext_resample_std=3.0)
    np.random.seed(42)
    ref = OracleDriver(oracle_model, **kw)
    xp_ref, st_ref = ref.run_mpc_fig8(np.zeros(12), fig, sim_dt=0.02, sim_time=0.4)
    np.random.seed(42)
    model, _, _ = pkg.load_robot_model()
    drv = MPC_GATO_Batch_Sample(model, **kw)
    best = []
    orig = drv.find_best_trajectory_id_fig8
    drv.find_best_trajectory_id_fig8 = lambda *a: best.append(orig(*a)) or best[-1]
    xp, st = drv.run_mpc_fig8(np.zeros(12), fig, sim_dt=0.02, sim_time=0.4)
    assert len(xp) == len(xp_ref) == 40
    assert best == ref.best_ids and len(set(best)) > 2
    assert rel(np.array(xp), np.array(xp_ref)) < 1e-6
    assert np.allclose(st['goal_distances'], st_ref['goal_distances'], atol=2e-5)


def test_peer_exchange_single_rank_matches_pack_reduce(pkg, oracle_model):
    """i7_p2p_exchange with a world of one rank (the rank's own buffer is its only peer): the fused
    pack + store + selection kernel returns bitwise what i7_pack_best + i7_reduce_best return, over several ticks
    (the double-buffered sequence flags must not stall or go stale)."""
    import torch
    xu, x0, goals = make_cases(oracle_model, 24, N=32, seed=13)
    e = pkg.Engine(N=32, batch=24, dtype="fp64", cluster=1)
    h = e.p2p_create(1, 0)
    assert len(h) == 64
    e.p2p_connect([h])
    dev = lambda a: torch.tensor(a, device="cuda")
    xu_d, x0_d, g_d = dev(xu), dev(x0), dev(goals)
    for tick in range(5):
        out, st = e.solve(xu_d, x0_d, g_d)
        ref = e.reduce_best(e.pack_best(st, out, index_offset=7).view(1, -1), 1)
        got = e.p2p_exchange(st, out, index_offset=7)
        assert torch.equal(got, ref)
        # ... and as the epilogue of the solve kernel itself (i7_solve_select): one launch, same bits
        out2, st2, got2 = e.solve_select(xu_d, x0_d, g_d, index_offset=7)
        assert torch.equal(out2, out) and torch.equal(st2, st) and torch.equal(got2, ref)
        xu_d = out
    e.p2p_status(got)
    assert e.exchange_wait_ns() >= 0
    big = pkg.Engine(N=32, batch=600, dtype="fp64")          # several waves of CTAs: the last ticket still sees every result
    big.p2p_connect([big.p2p_create(1, 0)])
    xb, x0b, gb = (dev(np.tile(a, (25, 1))) for a in (xu, x0, goals))
    o, s_, b_ = big.solve_select(xb, x0b, gb)
    assert torch.equal(b_, big.reduce_best(big.pack_best(s_, o).view(1, -1), 1))
    o1, s1 = big.solve(xb, x0b, gb)
    assert torch.equal(o, o1) and torch.equal(s_, s1)


def _peer_rank(rank, world, port, q):
    import os
    import torch
    import torch.distributed as dist
    import indy7_mpc_b200 as pkg
    from indy7_mpc_b200 import dist as i7d
    from oracle.model import load_model
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        model = load_model()
        xu, x0, goals = make_cases(model, 16 * world, N=32, seed=17)
        lo, hi = i7d.shard_range(16 * world, rank, world)
        e = pkg.Engine(N=32, batch=16, dtype="fp64", cluster=1, device=rank)
        assert i7d.connect_peers(e, rank, world)
        dev = lambda a: torch.tensor(a[lo:hi], device=f"cuda:{rank}")
        xu_d, x0_d, g_d = dev(xu), dev(x0), dev(goals)
        res = []
        for tick in range(4):
            out, st = e.solve(xu_d, x0_d, g_d)
            packed = e.pack_best(st, out, index_offset=lo)
            best = e.p2p_exchange(st, out, index_offset=lo)
            parts = [torch.empty_like(packed).cpu() for _ in range(world)]
            dist.all_gather(parts, packed.cpu())
            ref = e.reduce_best(torch.stack(parts).to(best.device), world)
            res.append(bool(torch.equal(best, ref)))
            # the same tick with selection + exchange as the epilogue of the solve kernel (ONE launch)
            out2, st2, best2 = e.solve_select(xu_d, x0_d, g_d, index_offset=lo)
            res.append(bool(torch.equal(out2, out) and torch.equal(st2, st) and torch.equal(best2, ref)))
            xu_d = out
        e.p2p_status(best)
        q.put((rank, res))
    finally:
        dist.destroy_process_group()


def test_peer_exchange_two_gpus(pkg):
    """Two ranks, one per GPU: the peer-memory exchange kernel (stores over NVLink into cudaIpc-mapped buffers) selects
    bitwise the same record as pack_best -> all-gather (gloo here) -> reduce_best on every rank, tick after tick."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_peer_rank, args=(r, 2, 29611, q)) for r in range(2)]
    for p in ps:
        p.start()
    got = [q.get(timeout=240) for _ in ps]
    for p in ps:
        p.join(60)
    assert sorted(r for r, _ in got) == [0, 1]
    assert all(all(ok) for _, ok in got)


@pytest.mark.parametrize("B", [1, 7, 64, 300])
def test_find_best_idx_fused_selection(pkg, oracle_model, B):
    """i7_find_best_idx (plant step of every lane + error + argmin in one launch) against the Python loops of
    gato_controller.py:109-118 (`<=`: last of equal minima) and gato_mpc_batch_sample.py:279-285 (`<`: first):
    errors to 1e-9, indices exact, duplicated hypotheses as exact ties, a NaN hypothesis never wins; device and
    host buffers."""
    import torch
    rng = np.random.default_rng(40 + B)
    e = pkg.Engine(N=8, batch=B, dtype="fp64")
    x_last = np.concatenate([rng.uniform(-1, 1, 6), rng.uniform(-0.5, 0.5, 6)])
    u_last = rng.uniform(-5, 5, 6)
    f = np.zeros((B, 6)); f[:, :3] = rng.normal(0, 5, (B, 3))
    k = B // 2
    if B >= 7:
        f[5] = f[2]                       # exact tie
        f[3, 0] = np.nan                  # a poisoned lane
        k = 2
    step = lambda w: np.hstack(rbd.rk4(oracle_model, x_last[:6], x_last[6:], u_last, 0.002, w))
    xs = step(f[k])
    e.set_external_wrench(f)
    ref_err = np.array([np.linalg.norm(step(f[i]) - xs) for i in range(B)])

    def loop(le):
        best, bi = np.inf, None
        for i in range(B):
            if (ref_err[i] <= best) if le else (ref_err[i] < best):
                best, bi = ref_err[i], i
        return bi
    for tie_last in (True, False):
        idx, errs = e.find_best_idx(x_last, u_last, xs, 0.002, batch=B, tie_last=tie_last, return_errors=True)
        ok = ~np.isnan(ref_err)
        assert np.array_equal(np.isnan(errs), ~ok)
        assert np.abs(errs[ok] - ref_err[ok]).max() <= 1e-9
        assert idx == loop(tie_last)
        if B >= 7:
            assert idx == (5 if tie_last else 2)
        d = lambda a: torch.tensor(a, device="cuda")
        assert e.find_best_idx(d(x_last), d(u_last), d(xs), 0.002, batch=B, tie_last=tie_last) == idx
    # xs away from every prediction: a generic argmin; every lane NaN: no winner
    xs2 = xs + rng.normal(0, 1e-3, 12)
    ref_err = np.array([np.linalg.norm(step(f[i]) - xs2) for i in range(B)])
    assert e.find_best_idx(x_last, u_last, xs2, 0.002, batch=B) == loop(True)
    assert e.find_best_idx(x_last, u_last, np.full(12, np.nan), 0.002, batch=B) is None
    e32 = pkg.Engine(N=8, batch=B, dtype="fp32")
    e32.set_external_wrench(f)
    assert e32.find_best_idx(x_last, u_last, xs2, 0.002, batch=B) == loop(True)
