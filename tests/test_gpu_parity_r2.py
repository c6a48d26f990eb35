"""Round-2 parity cases (all through the C ABI, against the CPU oracle on identical inputs):
the configurations VERDICT r01 found covered only GPU-against-GPU — full sqp() at N = 128 (fp64 and fp32, the
long-horizon kernel form), the fp32 throughput kernel at N = 32, the duals of the QP, the distance to the
reference's OSQP-tolerance regime — plus the MPC tick plumbing kernels and the host-buffer hazards of ADVICE r01."""
import numpy as np
import pytest

from scipy.sparse import triu

import bench
from oracle.cport import COracle
from oracle.osqp_solver import OSQPSolver as OracleSolver
from oracle.osqp_sqp import SQP_OSQP as OracleSQP
from tests.cases import make_cases

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.fixture(scope="module")
def pkg():
    import indy7_mpc_b200 as m
    return m


def _compare_with_oracle(out, st, ref, rst, tol, what, min_same=1.0):
    """Iteration counts, alpha lists and step counts exact, trajectories and step sizes within tol — on the lanes whose
    line-search decisions agree (in fp32 a merit comparison can legitimately fall the other way); at least
    `min_same` of the lanes must agree."""
    out, st = np.asarray(out, float), np.asarray(st, float)
    same = np.all(st[:, :6] == rst[:, :6], axis=1)
    frac = float(same.mean())
    worst = max((rel(out[b], ref[b]) for b in np.nonzero(same)[0]), default=0.0)
    print(f"{what}: {int(same.sum())}/{len(same)} lanes with identical decisions, worst trajectory deviation {worst:.2e}")
    assert frac >= min_same, (what, frac, st[~same, :6], rst[~same, :6])
    for b in np.nonzero(same)[0]:
        assert rel(out[b], ref[b]) < tol, (what, b, rel(out[b], ref[b]))
        n_st = int(st[b, 2])
        if n_st:
            assert rel(st[b, 6:6 + n_st], rst[b, 6:6 + n_st]) < max(tol, 1e-6), (what, b)


@pytest.mark.parametrize("dtype,tol,min_same", [("fp64", 1e-6, 1.0), ("fp32", 1e-4, 1.0)])
def test_long_horizon_full_sqp_matches_oracle(pkg, oracle_model, dtype, tol, min_same):
    """BASELINE.json configs[4] (N = 128, the horizon of trajopt_node.cu:178): full sqp() — setup, sweep, line search,
    update — of the one-CTA-per-problem form the 512-problem batch runs, on lanes of that very workload (per-lane
    wrench hypotheses included), two consecutive MPC ticks, against the exact-KKT C oracle (src/osqp_sqp.py:76-93)."""
    N, B = 128, 16
    wl = bench.make_workload(512, N)
    lanes = np.arange(B) * 31 % 512            # spread over the batch, lane 0 (zero wrench) included
    eng = pkg.Engine(N=N, batch=B, dtype=dtype, cluster=1)
    assert eng.kernel_form()[1] == 1
    eng.set_external_wrench(wl["f_ext"][lanes])
    orc = COracle(oracle_model, N=N, backend="kkt", batch=B)
    xu, x0 = wl["xu"][lanes].copy(), wl["x0"][lanes].copy()
    for tick in range(2):
        goals = bench.goals_at(wl["fig"], wl["offsets"][lanes], tick, N)
        ref, rst, _ = orc.sqp(x0, goals, xu, f_ext=wl["f_ext"][lanes])
        out, st = eng.solve(xu, x0, goals)
        _compare_with_oracle(out, st, ref, rst, tol, f"N=128 {dtype} tick {tick}", min_same)
        xu, x0 = bench.shift_host(ref, N)          # both sides continue from the ORACLE's solution: identical inputs
        xu[:, :12] = x0


def test_fp32_throughput_kernel_full_sqp_matches_oracle(pkg, oracle_model):
    """The fp32 one-CTA-per-problem kernel (the form of the 4096-problem fp32 figure) at N = 32, full sqp(), on the
    12 seeded cases of test_sqp_solve_matches_oracle: decisions and 1e-4 trajectories against the fp64 oracle."""
    xu, x0, goals = make_cases(oracle_model, 12, seed=6)
    for threads in (128, 384):
        eng = pkg.Engine(N=32, batch=12, dtype="fp32", cluster=1, threads=threads)
        out, st = eng.solve(xu, x0, goals)
        orc = COracle(oracle_model, N=32, backend="kkt", batch=12)
        ref, rst, _ = orc.sqp(x0, goals, xu)
        _compare_with_oracle(out, st, ref, rst, 1e-4, f"N=32 fp32 CTA form ({threads} threads)", 1.0)


@pytest.mark.parametrize("N", [32, 128])
def test_qp_duals_and_kkt_residuals(pkg, oracle_model, N):
    """north_star: 'primal/dual residuals below the reference eps_abs / eps_rel' (1e-3, src/osqp_solver.py:39-41).
    i7_solve_qp returns z and the duals y of the 12N equality rows; with the oracle's P, g, A, l:
    |Az - l| and |Pz + g + A'y| are at rounding level, and y equals the exact-KKT oracle's multipliers."""
    xu, x0, goals = make_cases(oracle_model, 4, N=N, seed=41)
    eng = pkg.Engine(N=N, batch=4, cluster=1)
    z, y = eng.solve_qp(xu, x0, goals, want_duals=True)
    for b in range(4):
        o = OracleSolver(oracle_model, N=N, qp_backend="kkt")
        sol = o.setup_and_solve_qp(xu[b], x0[b], goals[b])
        Pu, A = o.matrices()
        P = Pu + triu(Pu, 1).T                     # the reference passes the upper triangle (src/osqp_solver.py:52)
        r_prim = np.abs(A @ z[b] - o.l).max()
        r_dual = np.abs(P @ z[b] + o.g + A.T @ y[b]).max()
        scale = max(np.abs(P @ z[b]).max(), np.abs(A.T @ y[b]).max(), np.abs(o.g).max())
        assert r_prim < 1e-8, (b, r_prim)
        assert r_dual < 1e-9 * max(scale, 1.0), (b, r_dual, scale)
        assert rel(y[b], sol.y) < 1e-6, (b, rel(y[b], sol.y))
    # the reference-shaped class fills .y
    if N == 32:
        s = pkg.OSQPSolver(pkg.load_robot_model()[0])
        res = s.setup_and_solve_qp(xu[0], x0[0], goals[0])
        assert res.y is not None and rel(res.y, y[0]) < 1e-12


def test_exact_qp_against_osqp_tolerance_regime(pkg, oracle_model):
    """'The same SQP iteration count where the QP tolerances match': the reference stops OSQP at eps 1e-3 with a warm
    start (src/osqp_solver.py:39-41,140-143); the product solves each QP exactly.  Over 64 seeded problems: the
    fraction of lanes on which the two regimes take the same line-search decisions (reported), and the trajectory
    distance on those lanes (the effect of the inexact QP solution alone)."""
    xu, x0, goals = make_cases(oracle_model, 64, seed=77)
    eng = pkg.Engine(N=32, batch=64)
    out, st = eng.solve(xu, x0, goals)
    out, st = np.asarray(out), np.asarray(st)
    admm = COracle(oracle_model, N=32, backend="admm", batch=64)
    ref, rst, its = admm.sqp(x0, goals, xu)
    same_iters = st[:, 0] == rst[:, 0]
    same_alpha = np.all(st[:, 4:6] == rst[:, 4:6], axis=1)
    same = same_iters & same_alpha
    dev = [rel(out[b], ref[b]) for b in np.nonzero(same)[0]]
    print(f"exact vs OSQP-tolerance regime: same qp_iters {same_iters.mean():.3f}, same alpha sequence {same_alpha.mean():.3f}, "
          f"trajectory deviation on agreeing lanes: median {np.median(dev):.2e}, max {np.max(dev):.2e}; "
          f"ADMM iterations per solve {its / 64:.1f}")
    # floors from the two CPU oracles against each other (exact KKT vs ADMM back-end: 64/64 lanes agree, median
    # deviation 2.4e-5, max 9.4e-4)
    assert same_iters.mean() >= 0.95
    assert same_alpha.mean() >= 0.9
    assert np.median(dev) < 1e-3


# ---- MPC tick plumbing -----------------------------------------------------------------------------------------
def _shift_ref(xu_new, xu_prev, x_cur, steps, tail, N):
    nz = 18 * N - 6
    out = np.empty_like(xu_new)
    s = 18 * steps
    out[:, :nz - s] = xu_new[:, s:]
    for b in range(xu_new.shape[0]):
        if tail == "osqp_mpc":
            out[b, nz - s:] = xu_prev[b, nz - s:]
        else:
            for i in range(steps):
                o = nz - s + 18 * i
                out[b, o:o + 6] = 0.0 if tail == "hold" else xu_new[b, nz - 18:nz - 12]
                out[b, o + 6:o + 18] = xu_new[b, nz - 12:]
    x0 = x_cur if x_cur is not None else xu_new[:, s:s + 12]
    out[:, :12] = x0
    if tail == "osqp_mpc":
        out[:, -12:] = np.concatenate([np.ones(6), np.zeros(6)])
    return out, np.array(x0)


@pytest.mark.parametrize("tail", ["osqp_mpc", "hold", "repeat"])
@pytest.mark.parametrize("steps", [0, 1, 3])
def test_shift_kernel(pkg, tail, steps):
    """i7_shift against the reference's trajectory shift (src/osqp_mpc.py:65-70 incl. the last-knot quirk), the
    device precedent just_shift (trajopt_solver.cuh:153-172) and the batch drivers' repeat-tail form; bit-exact."""
    import torch
    N, B = 32, 5
    rng = np.random.default_rng(5)
    eng = pkg.Engine(N=N, batch=B)
    new, prev, cur = rng.normal(size=(B, 570)), rng.normal(size=(B, 570)), rng.normal(size=(B, 12))
    for xc in (cur, None):
        want, want0 = _shift_ref(new, prev, xc, steps, tail, N)
        got, got0 = eng.shift(new, steps=steps, tail=tail, xu_prev=prev, x_cur=xc)
        assert np.array_equal(got, want) and np.array_equal(got0, want0)
        d = lambda a: None if a is None else torch.tensor(a, device="cuda")
        gd, gd0 = eng.shift(d(new), steps=steps, tail=tail, xu_prev=d(prev), x_cur=d(xc))
        assert gd.is_cuda and np.array_equal(gd.cpu().numpy(), want) and np.array_equal(gd0.cpu().numpy(), want0)
    # the reference loop itself (osqp_mpc.py): XU[:-18] = xu_new[18:]; XU[:12] = xcur; XU[-12:] = [1, 0]
    if tail == "osqp_mpc" and steps == 1:
        XU = prev[0].copy()
        XU[:-18] = new[0][18:]; XU[:12] = cur[0]; XU[-12:] = np.hstack([np.ones(6), np.zeros(6)])
        assert np.array_equal(eng.shift(new[:1], 1, "osqp_mpc", prev[:1], cur[:1])[0][0], XU)


def test_goal_window_kernel(pkg):
    import torch
    N, B = 32, 7
    eng = pkg.Engine(N=N, batch=B)
    fig = bench.make_workload(8, N)["fig"]                      # 3000 x 3
    off = (37 * np.arange(B)) % 1000
    for tick in (0, 5, 2990):
        want = bench.goals_at(fig, off, tick, N)
        assert np.array_equal(eng.goal_window(fig, off, tick=tick, wrap=True), want)
        got6 = eng.goal_window(fig, off, tick=tick, wrap=True, out_stride=6).reshape(B, N, 6)
        assert np.array_equal(got6[:, :, :3].reshape(B, -1), want) and not got6[:, :, 3:].any()
        idx = np.minimum(off[:, None] + tick + np.arange(N)[None, :], len(fig) - 1)     # clamped: padded tail
        assert np.array_equal(eng.goal_window(fig, off, tick=tick, wrap=False), fig[idx].reshape(B, -1))
    g = eng.goal_window(torch.tensor(fig, device="cuda"), torch.tensor(off, device="cuda"), tick=3)
    assert g.is_cuda and np.array_equal(g.cpu().numpy(), bench.goals_at(fig, off, 3, N))
    # GATO-format source (T x 6, gato_controller.py:316-323)
    fig6 = np.concatenate([fig, np.zeros_like(fig)], axis=1)
    assert np.array_equal(eng.goal_window(fig6, off, tick=1), bench.goals_at(fig, off, 1, N))


# ---- host-buffer hazards -----------------------------------------------------------------------------
def test_pageable_inputs_with_device_outputs_back_to_back(pkg, oracle_model):
    """Two asynchronous device_out solves with DIFFERENT pageable inputs in a row: the second call must not overwrite
    the handle's page-locked input mirror while the first launch still reads it."""
    import torch
    for B in (4, 48):               # mapped-memory path (<= 8 problems) and the staged-copy path
        xa, x0a, ga = make_cases(oracle_model, B, seed=51)
        xb, x0b, gb = make_cases(oracle_model, B, seed=52)
        eng = pkg.Engine(N=32, batch=B, cluster=1)
        dev = lambda a: torch.tensor(a, device="cuda")
        ra, _ = eng.solve(dev(xa), dev(x0a), dev(ga))
        rb, _ = eng.solve(dev(xb), dev(x0b), dev(gb))
        torch.cuda.synchronize()
        for _ in range(3):
            oa, _ = eng.solve(xa, x0a, ga, device_out=True)
            ob, _ = eng.solve(xb, x0b, gb, device_out=True)
            torch.cuda.synchronize()
            assert torch.equal(oa, ra) and torch.equal(ob, rb)


def test_pinned_results_survive_the_next_solve(pkg, oracle_model):
    """Page-locked inputs give engine-owned page-locked outputs (a ring of two): the result of call i is intact after
    call i+1 (it is that call's warm start), and Indy7SQP.solve never hands a buffer to the kernel as input and output."""
    import torch
    xu, x0, goals = make_cases(oracle_model, 3, seed=53)
    pin = lambda a: torch.tensor(a).pin_memory()
    s = pkg.Indy7SQP(N=32, batch=3, cluster=1)
    x1, u1, st1 = s.solve(pin(x0), pin(goals), xu_init=pin(xu))
    keep = x1.clone()
    x2, u2, st2 = s.solve(pin(x0), pin(goals))                    # warm start = the previous result
    assert torch.equal(x1, keep)
    assert st2["xu"].data_ptr() != st1["xu"].data_ptr()
    ref = pkg.Engine(N=32, batch=3, cluster=1)
    o1, _ = ref.solve(xu, x0, goals)
    o2, _ = ref.solve(o1, x0, goals)
    assert np.array_equal(st1["xu"].numpy(), o1) and np.array_equal(st2["xu"].numpy(), o2)


def test_nonfinite_input_returns_trajectory_unchanged(pkg):
    """A non-finite merit makes every line-search comparison false: alpha = 0, XU comes back as it went in, like the
    reference loop (src/osqp_sqp.py:66-72,81-82); status 2 in the record, no exception."""
    model, _, _ = pkg.load_robot_model()
    solver = pkg.OSQPSolver(model)
    sqp = pkg.SQP_OSQP(solver)
    xu = np.zeros(570); xu[3] = np.nan
    out = sqp.sqp(np.zeros(12), np.zeros(96), xu)
    assert np.array_equal(np.isnan(out), np.isnan(xu)) and np.array_equal(out[~np.isnan(xu)], xu[~np.isnan(xu)])
    assert int(sqp.last_record[3]) == 2 and sqp.stats['linesearch_alphas']['values'] == [0.0, 0.0]


def test_device_tick_equals_generic_calls(pkg):
    """DeviceTick (buffers bound once, one C-ABI call per stage) runs the same closed loop as the generic Engine calls —
    goal window, solve, selection, shift — bit for bit, in the local and in the fused-epilogue mode."""
    import torch
    import bench
    B, N = 48, 32
    wl = bench.make_workload(B, N)
    dev = torch.device("cuda", 0)
    for mode in ("local", "fused"):
        e = pkg.Engine(N=N, batch=B, cluster=1)
        e.set_external_wrench(wl["f_ext"])
        if mode == "fused":
            e.p2p_connect([e.p2p_create(1, 0)])
        tk = pkg.DeviceTick(e, wl["fig"], wl["offsets"], wl["xu"], wl["x0"], mode=mode, index_offset=5)
        fig_d, off_d = torch.tensor(wl["fig"], device=dev), torch.tensor(wl["offsets"], dtype=torch.int32, device=dev)
        xu, x0 = torch.tensor(wl["xu"], device=dev), torch.tensor(wl["x0"], device=dev)
        for t in range(4):
            tk.run(t)
            goals = e.goal_window(fig_d, off_d, tick=t, wrap=True, batch=B)
            out, st = e.solve(xu, x0, goals)
            pk = e.pack_best(st, out, index_offset=5)
            xu, x0 = e.shift(out, steps=1, tail="repeat")
            assert torch.equal(out, tk.out) and torch.equal(st, tk.stats) and torch.equal(pk[B:], tk.best)
            assert torch.equal(xu, tk.xu) and torch.equal(x0, tk.x0)


def test_indy7sqp_step_bound_path_equals_generic(pkg):
    """Indy7SQP.step — the closed-loop tick of the end-to-end arm — on buffers bound once (page-locked torch inputs)
    returns what the generic path (numpy inputs) returns, tick after tick, and both equal the explicit sequence
    shift -> solve -> pack_best of Engine calls."""
    import torch
    import bench
    B, N = 40, 32
    wl = bench.make_workload(B, N)
    fast = pkg.Indy7SQP(N=N, batch=B, cluster=1)
    slow = pkg.Indy7SQP(N=N, batch=B, cluster=1)
    eng = pkg.Engine(N=N, batch=B, cluster=1)
    for s in (fast, slow):
        s.set_external_wrench(wl["f_ext"])
    eng.set_external_wrench(wl["f_ext"])
    p_x0 = torch.empty(B, 12, dtype=torch.float64).pin_memory()
    p_g = torch.empty(B, 3 * N, dtype=torch.float64).pin_memory()
    x0 = wl["x0"].copy()
    prev = None
    for t in range(4):
        goals = bench.goals_at(wl["fig"], wl["offsets"], t, N)
        p_x0.copy_(torch.from_numpy(x0)); p_g.copy_(torch.from_numpy(goals))
        a = fast.step(p_x0, p_g, index_offset=3)
        b = slow.step(x0, goals, index_offset=3)
        if prev is None:
            xu = np.zeros((B, 18 * N - 6)); xu[:, :12] = x0
        else:
            xu, _ = eng.shift(prev, steps=1, tail="repeat", x_cur=x0, want_x0=False)
        out, st = eng.solve(np.asarray(xu), x0, goals)
        pk = np.asarray(eng.pack_best(st, out, index_offset=3))
        prev = np.asarray(out)
        for r in (a, b):
            assert np.array_equal(r["costs"], pk[:B]) and r["best_cost"] == pk[B] and r["best_index"] == int(pk[B + 1])
            assert np.array_equal(r["x_next"], prev[:, 18:30])
            assert np.array_equal(np.concatenate([r["x_traj"], np.vstack([r["u_traj"], np.zeros((1, 6))])], axis=1).reshape(-1)[:18 * N - 6], pk[B + 2:])
        x0 = a["x_next"].copy()
    # alternating the two paths on ONE solver keeps the closed loop intact (the bound path takes over the generic path's
    # last solution and vice versa)
    mixed = pkg.Indy7SQP(N=N, batch=B, cluster=1); mixed.set_external_wrench(wl["f_ext"])
    pure = pkg.Indy7SQP(N=N, batch=B, cluster=1); pure.set_external_wrench(wl["f_ext"])
    x0m = wl["x0"].copy()
    for t in range(4):
        goals = bench.goals_at(wl["fig"], wl["offsets"], t, N)
        p_x0.copy_(torch.from_numpy(x0m)); p_g.copy_(torch.from_numpy(goals))
        rm = mixed.step(p_x0, p_g) if t % 2 == 0 else mixed.step(x0m, goals)
        rp = pure.step(x0m, goals)
        assert np.array_equal(rm["costs"], rp["costs"]) and np.array_equal(rm["x_next"], rp["x_next"]), t
        x0m = rp["x_next"].copy()
    fast.reset()
    p_x0.copy_(torch.from_numpy(wl["x0"])); p_g.copy_(torch.from_numpy(bench.goals_at(wl["fig"], wl["offsets"], 0, N)))
    again = fast.step(p_x0, p_g, index_offset=3)           # after reset: the cold first tick again
    cold = pkg.Indy7SQP(N=N, batch=B, cluster=1); cold.set_external_wrench(wl["f_ext"])
    ref = cold.step(wl["x0"], bench.goals_at(wl["fig"], wl["offsets"], 0, N), index_offset=3)
    assert np.array_equal(again["costs"], ref["costs"])


@pytest.mark.parametrize("dtype", ["fp64", "fp32"])
def test_resample_wrench_equals_numpy_statements(pkg, dtype):
    """i7_resample_wrench against the reference's numpy statements, bit for bit: gato_controller.py:120-129 (kept lane =
    best, x 0.97), the notebook variant (lane 1 keeps the best wrench, restored after the torque part is cleared) and
    src/gato_mpc_batch_sample.py:286-298 (no kept lane); host and device buffers, in place, chained to the device result
    of the selection; the handle's wrench batch follows (checked through sim_forward)."""
    import torch
    from indy7_mpc_b200 import _lib
    B = 37
    rng = np.random.default_rng(99)
    e = pkg.Engine(N=8, batch=B, dtype=dtype)
    f0 = rng.normal(0, 5, (B, 6))

    def controller(f, best, noise, std_decay=0.97):            # gato_controller.py:120-129, statement by statement
        f = f.copy()
        fb = f[best, :].copy()
        f[:, :] = fb
        f += noise
        f[best, :] = fb
        f[:, 3:] = 0.0
        f[0] = np.array([0.0, 0.0, 0.0, 0.0, 0.0, 0.0])
        f *= std_decay
        return f

    def notebook(f, best, noise):
        fb = f[best].copy()
        g = np.tile(fb, (B, 1)) + noise
        g[:, 3:] = 0.0
        g[1] = fb
        g[0] = 0.0
        return g

    def batch_sample(f, best, noise):                          # src/gato_mpc_batch_sample.py:286-298
        g = f[best].copy() + noise
        g[:, 3:] = 0.0
        g[0] = 0.0
        return g
    f = f0.copy()
    for it, best in enumerate([5, 0, B - 1, 1]):
        noise = rng.normal(0, 2.0, (B, 6))
        ref = controller(f, best, noise)
        out = e.resample_wrench(f, noise, best, keep_lane=_lib.I7_KEEP_BEST, decay=0.97)
        assert isinstance(out, np.ndarray) and np.array_equal(out, ref), it
        assert np.array_equal(e.resample_wrench(f, noise, best, keep_lane=1, keep_torque=True, apply=False), notebook(f, best, noise))
        assert np.array_equal(e.resample_wrench(f, noise, best, keep_lane=None, apply=False), batch_sample(f, best, noise))
        f = ref                                                # the recursion of the controller, tick after tick
    # the handle's wrench batch is the applied one: same plant step as after set_external_wrench
    x, u = rng.uniform(-1, 1, 12), rng.uniform(-5, 5, 6)
    xn_applied = np.asarray(e.sim_forward(x, u, 0.002, batch=B))
    e.set_external_wrench(f)
    assert np.array_equal(xn_applied, np.asarray(e.sim_forward(x, u, 0.002, batch=B)))
    # device buffers, in place, lane taken from the device result of the selection (no host round trip)
    d = lambda a, dt=torch.float64: torch.tensor(a, device="cuda", dtype=dt)
    xs = xn_applied[7].astype(float)
    res = e.find_best_idx_async(d(x, e._torch_dtype()), d(u, e._torch_dtype()), d(xs, e._torch_dtype()), 0.002, batch=B)
    assert res.is_cuda
    noise = rng.normal(0, 2.0, (B, 6))
    # the same chain with host buffers (result of the selection in a numpy array)
    res_h = e.find_best_idx_async(x, u, xs, 0.002, batch=B)
    assert isinstance(res_h, np.ndarray) and int(res_h[0]) == 7
    assert np.array_equal(e.resample_wrench(f, noise, res_h, decay=0.97, apply=False), controller(f, 7, noise))
    fd = d(f)
    out = e.resample_wrench(fd, d(noise), res, decay=0.97, inplace=True)
    assert out.data_ptr() == fd.data_ptr() and int(res[0]) == 7
    assert np.array_equal(fd.cpu().numpy(), controller(f, 7, noise))
    # no lane qualified (-1): the batch stays as it is
    res_nan = e.find_best_idx_async(d(x, e._torch_dtype()), d(u, e._torch_dtype()), d(np.full(12, np.nan), e._torch_dtype()), 0.002, batch=B)
    before = fd.clone()
    e.resample_wrench(fd, d(noise), res_nan, inplace=True)
    assert int(res_nan[0]) == -1 and torch.equal(fd, before)
    with pytest.raises(ValueError):
        e.resample_wrench(f, noise, B)


@pytest.mark.parametrize("qp_mode,dtype", [("exact", "fp64"), ("osqp", "fp64"), ("exact", "fp32")])
def test_controller_tick_equals_joint_callback_statements(pkg, oracle_model, qp_mode, dtype):
    """ControllerTick (five launches per tick, buffers bound once, the selected lane never leaves the device between
    selection, resampling and hand-over) against GATO_Controller.joint_callback (gato_controller.py:201-256) restated
    statement by statement over host arrays and the generic calls (each of which has its own test against the CPU
    restatement): same np.random stream -> same lane, same control, same wrench batch at every tick of a closed loop."""
    from oracle import rbd
    from tests.cases import HOME
    from indy7_mpc_b200 import ControllerTick
    from indy7_mpc_b200.reference_traj import figure8
    B, N, dt = 16, 32, 0.01
    ref_traj = figure8(A_x=0.5, A_z=0.55, offset=[0.0, 0.4, 0.45], period=10, dt=dt, cycles=2)
    ref_traj = np.concatenate([np.tile(ref_traj[0:6], 20), ref_traj])            # gato_controller.py:322-323 (shorter padding)
    f_actual = np.array([-6.0, 2.0, -4.0, 0.0, 0.0, 0.0])
    x_start = np.concatenate([HOME, np.zeros(6)])
    elapsed = [0.01, 0.012, 0.009, 0.02, 0.01, 0.011, 0.01, 0.013]
    plant = lambda x, u, h: np.hstack(rbd.rk4(oracle_model, x[:6], x[6:], u, h, f_actual))

    def statements():
        np.random.seed(11)
        e = pkg.Engine(N=N, dt=dt, batch=B, dtype=dtype, qp_mode=qp_mode)
        f_ext_batch = np.random.normal(0, 20.0, (B, 6))                        # :77-80
        f_ext_batch[:, 3:] = 0.0
        f_ext_batch[0] = 0.0
        e.set_external_wrench(f_ext_batch)
        XU_batch, xs_batch, goals = np.zeros((B, 18 * N - 6)), np.zeros((B, 12)), np.zeros((B, 6 * N))
        x_last = u_last = None
        off, x_curr, log = 0.0, x_start.copy(), []
        for el in elapsed:
            if x_last is None:
                x_last, u_last = x_curr, np.zeros(6)
            off += el / dt
            offset = int(off)
            goals[:, :] = ref_traj[6 * offset:6 * (offset + N)]
            xs_batch[:, :] = x_curr
            XU_batch[:, 0:12] = x_curr
            e.reset(); e.reset_rho(); e.reset_lambda()
            XU_new, _ = e.solve(XU_batch, xs_batch, goals)
            XU_new = np.asarray(XU_new, float)
            best = e.find_best_idx(x_last, u_last, x_curr, el, batch=B, tie_last=True)
            fb = f_ext_batch[best, :].copy()                                       # :120-129
            f_ext_batch[:, :] = fb
            f_ext_batch += np.random.normal(0, 1.0, f_ext_batch.shape)
            f_ext_batch[best, :] = fb
            f_ext_batch[:, 3:] = 0.0
            f_ext_batch[0] = 0.0
            f_ext_batch *= 0.97
            e.set_external_wrench(f_ext_batch)
            XU_best = XU_new[best]
            u_curr = XU_best[12:18].copy()
            XU_batch[:, :] = XU_best
            log.append((best, u_curr, f_ext_batch.copy(), XU_best.copy()))
            x_last, u_last = x_curr, u_curr
            x_curr = plant(x_curr, u_curr, el)
        return log

    ref = statements()
    np.random.seed(11)
    e = pkg.Engine(N=N, dt=dt, batch=B, dtype=dtype, qp_mode=qp_mode)
    ct = ControllerTick(e, ref_traj, B, f_ext_std=20.0, f_ext_resample_std=1.0)
    x_curr = x_start.copy()
    picks = []
    for t, el in enumerate(elapsed):
        u, best = ct.tick(x_curr, el)
        b_ref, u_ref, f_ref, xu_ref = ref[t]
        assert best == b_ref, (t, best, b_ref)
        assert np.array_equal(u, u_ref), (t, np.abs(u - u_ref).max())
        assert np.array_equal(ct.f_ext_batch, f_ref)
        assert np.array_equal(ct.XU_best, xu_ref)
        assert np.array_equal(ct.xu.cpu().numpy(), np.tile(xu_ref, (B, 1)))      # XU_batch[:, :] = XU_best
        picks.append(best)
        x_curr = plant(x_curr, u, el)
    assert len(set(picks)) > 1                                                    # the selection is not a constant


def test_device_tick_under_cuda_graph(pkg):
    """The tick's solve + selection + shift captured into a CUDA graph (DeviceTick.capture) and replayed: bit-identical
    to the eager tick over a closed loop, for the CTA form and a latency (cluster) form."""
    import torch
    for B in (16, 300):
        wl = bench.make_workload(B, 32)
        ticks = []
        for _ in range(2):
            e = pkg.Engine(N=32, batch=B, dtype="fp64")
            e.set_external_wrench(wl["f_ext"])
            ticks.append(pkg.DeviceTick(e, wl["fig"], wl["offsets"], wl["xu"], wl["x0"], mode="local"))
        eager, graphed = ticks
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            for t in range(2):
                eager.run(t)
                graphed.run(t)
            replay = graphed.capture()
            for t in range(2, 8):
                eager.run(t)
                replay(t)
                s.synchronize()
                assert torch.equal(eager.out, graphed.out) and torch.equal(eager.stats, graphed.stats), (B, t)
                assert torch.equal(eager.packed, graphed.packed) and torch.equal(eager.xu, graphed.xu), (B, t)
        assert torch.isfinite(graphed.out).all()
