"""CPU-side checks of the host layer: the C ABI library loads and exports every declared symbol, the URDF loader
agrees with the oracle's independent loader, and the product refuses to run without a GPU (no CPU fallback)."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build():
    from indy7_mpc_b200 import build
    return build.build()


def test_library_exports_every_declared_symbol():
    _build()
    from indy7_mpc_b200 import _lib
    lib = _lib.load()
    header = open(os.path.join(ROOT, "include", "indy7_b200.h")).read()
    declared = set(re.findall(r"\b(i7_[a-z0-9_]+)\s*\(", header)) - {"i7_status"}
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    for name in declared:
        assert hasattr(lib, name)
    assert b"sm_100a" in lib.i7_version()


def test_config_struct_layout_matches_header():
    import ctypes as C
    from indy7_mpc_b200 import _lib
    lib = _lib.load()
    cfg = _lib.I7Config()
    lib.i7_default_config(C.byref(cfg))
    assert (cfg.N, cfg.batch, cfg.dtype, cfg.max_qp, cfg.regularize) == (32, 1, 0, 2, 1)
    assert (cfg.dt, cfg.dQ_cost, cfg.R_cost, cfg.QN_cost, cfg.eps, cfg.mu, cfg.step_tol) == (0.01, 0.01, 1e-5, 100.0, 1.0, 10.0, 1e-3)
    assert cfg.model.gravity[2] == -9.81
    # round-2 fields: QP mode and OSQP's default settings (src/osqp_solver.py:39-41 relies on them), limits off and unbounded
    assert (cfg.qp_mode, cfg.limits, cfg.admm_max_iter, cfg.admm_check) == (_lib.I7_QP_EXACT, 0, 4000, 25)
    assert (cfg.admm_rho, cfg.admm_rho_u, cfg.admm_sigma, cfg.admm_alpha, cfg.admm_eps_abs, cfg.admm_eps_rel) == (0.1, 1e-3, 1e-6, 1.6, 1e-3, 1e-3)
    assert cfg.model.q_min[0] == -1e30 and cfg.model.u_max[5] == 1e30
    # the model fills the limits of description/indy7.urdf:203-238 in
    from indy7_mpc_b200 import build_model_from_urdf
    m = build_model_from_urdf().to_c()
    assert abs(m.q_max[0] - 3.0543261909900767) < 1e-15 and abs(m.q_min[5] + 3.7524578917878086) < 1e-15
    assert abs(m.v_max[3] - 3.141592653589793) < 1e-15 and m.u_max[2] == 197.23


def test_product_urdf_loader_agrees_with_oracle_loader(oracle_model):
    from indy7_mpc_b200.model import build_model_from_urdf
    m = build_model_from_urdf()
    assert (m.nq, m.nv, m.njoints, len(m.joints)) == (6, 6, 7, 7)
    assert np.abs(m.R - oracle_model.R).max() < 1e-15
    assert np.array_equal(m.p, oracle_model.p) and np.array_equal(m.mass, oracle_model.mass)
    assert np.array_equal(m.com, oracle_model.com) and np.array_equal(m.Ic, oracle_model.Ic)
    c = m.to_c()
    assert c.mass[5] == m.mass[5] and c.gravity[2] == -9.81


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import indy7_mpc_b200 as pkg
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pkg.Engine()
    with pytest.raises(ValueError):
        pkg.Indy7SQP(device="cpu")


def test_product_never_imports_oracle():
    pkg_dir = os.path.join(ROOT, "indy7_mpc_b200")
    for dp, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f
                assert "oracle/" not in src, f


def test_batch_sqp_surface():
    import bindings.batch_sqp as b
    for B in (1, 2, 4, 8, 16, 32, 64, 128, 256):
        cls = getattr(b, f"SQPSolverfloat_{B}")
        for meth in ("solve", "reset", "resetRho", "resetLambda", "set_external_wrench_batch", "sim_forward"):
            assert callable(getattr(cls, meth))
    assert not hasattr(b, "SQPSolverfloat_3")


def test_figure8_generator():
    from indy7_mpc_b200.reference_traj import figure8
    ref = np.load(os.path.join(ROOT, "tests", "golden", "figure8_first_cycle.npy"))
    got = figure8(0.5, 0.55, (0.0, 0.4, 0.45), 10.0, 0.01, cycles=2)
    assert got.shape == (12000,) and np.abs(got[:6000] - ref).max() < 1e-15


def test_world_wrench_to_last_joint_matches_oracle_fk():
    """Frame bookkeeping of the batch-sample driver (oMi[6].actInv, src/gato_mpc_batch_sample.py:268-277)."""
    import numpy as np
    from indy7_mpc_b200.model import build_model_from_urdf
    from oracle.model import load_model
    from oracle import rbd
    m, om = build_model_from_urdf(), load_model()
    rng = np.random.default_rng(3)
    for _ in range(5):
        q = rng.uniform(-3, 3, 6)
        w = rng.normal(0, 2, (4, 6))
        p, Rw = rbd.forward_kinematics(om, q)
        ref = np.stack([np.concatenate([Rw[-1].T @ x[:3], Rw[-1].T @ (x[3:] - np.cross(p[-1], x[:3]))]) for x in w])
        assert np.abs(m.world_wrench_to_last_joint(q, w) - ref).max() < 1e-13


def test_oracle_batch_sample_driver_runs():
    import numpy as np
    from oracle.model import load_model
    from oracle.gato_mpc_batch_sample import MPC_GATO_Batch_Sample
    np.random.seed(1)
    d = MPC_GATO_Batch_Sample(load_model(), N=8, batch_size=2, constant_f_ext=[1.0, 0.0, 0.0], resample_f_ext=True)
    x0 = np.concatenate([0.4 * np.ones(6), np.zeros(6)])
    xp, st = d.run_mpc(x0, [np.array([0.3, 0.4, 0.5])], sim_dt=0.005, sim_time=0.03)
    assert len(xp) >= 6 and np.all(np.isfinite(np.array(xp))) and len(d.best_ids) == 6


def test_trajectory_csv_and_stats_formats(tmp_path):
    """On-disk formats of the reference drivers: the (T x 6) '%.18e' CSV of the C++ driver (byte-exact round trip of
    the head of trajectories/figure8_traj_eePos_meters.csv) and the six .npy series of gato_controller.py:281-293."""
    import os
    import numpy as np
    from indy7_mpc_b200.reference_traj import (StatsLog, figure8, figure_8, load_trajectory_csv, save_trajectory_csv)
    src = os.path.join(os.path.dirname(__file__), "golden", "figure8_traj_eePos_meters_head.csv")
    rows = load_trajectory_csv(src)
    assert rows.shape == (4, 6) and rows[0, 0] == 4.333333333333333481e-01 and np.all(rows[:, 3:] == 0)
    out = tmp_path / "t.csv"
    save_trajectory_csv(out, rows)
    assert open(out).read() == open(src).read()
    save_trajectory_csv(out, figure8(cycles=1))
    assert np.array_equal(load_trajectory_csv(out).reshape(-1), figure8(cycles=1))
    assert np.allclose(figure_8(0.5, 0.55, (0.0, 0.4, 0.45), 0.01, 10.0, 2), figure8(cycles=2), atol=1e-15)
    log = StatsLog()
    for i in range(3):
        log.record(0.01, 0.1 * i, np.ones(3) * i, np.zeros(6), np.arange(12.0), 1e-4)
    log.save(str(tmp_path / "120000"))
    names = sorted(os.listdir(tmp_path))
    assert [n for n in names if n.endswith(".npy")] == sorted(f"120000_{f}.npy" for f in StatsLog.FIELDS)
    back = StatsLog.load(str(tmp_path / "120000"))
    assert np.array(back.joint_positions).shape == (3, 6) and np.array(back.ee_ref_positions).shape == (3, 3)
