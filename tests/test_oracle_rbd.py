"""Self-validation of the oracle's rigid-body dynamics (nothing in the reference pins these individually)."""
import numpy as np

from oracle import rbd


def _rand(rng, n=8):
    return rng.uniform(-3, 3, (n, 6)), rng.uniform(-2.5, 2.5, (n, 6)), rng.uniform(-80, 80, (n, 6))


def test_aba_inverts_rnea(oracle_model):
    rng = np.random.default_rng(0)
    q, v, u = _rand(rng)
    for f in (None, rng.uniform(-20, 20, (8, 6))):
        qdd = rbd.aba(oracle_model, q, v, u, f)
        assert np.abs(rbd.rnea(oracle_model, q, v, qdd, f) - u).max() < 1e-11


def test_mass_matrix(oracle_model):
    rng = np.random.default_rng(1)
    q, v, u = _rand(rng)
    M = rbd.crba(oracle_model, q)
    assert np.abs(M - np.swapaxes(M, -1, -2)).max() == 0
    assert np.linalg.eigvalsh(M).min() > 0
    # M qdd + bias = tau
    qdd = rng.uniform(-5, 5, (8, 6))
    tau = rbd.rnea(oracle_model, q, v, qdd)
    bias = rbd.rnea(oracle_model, q, v, np.zeros_like(qdd))
    assert np.abs(np.einsum('bij,bj->bi', M, qdd) + bias - tau).max() < 1e-11


def test_analytic_derivatives_vs_complex_step(oracle_model):
    rng = np.random.default_rng(2)
    q, v, u = _rand(rng)
    for f in (None, rng.uniform(-20, 20, (8, 6))):
        an = rbd.aba_derivatives(oracle_model, q, v, u, f)
        cs = rbd.aba_derivatives_complex_step(oracle_model, q, v, u, f)
        for a, c in zip(an, cs):
            assert np.abs(a - c).max() <= 1e-12 * max(1.0, np.abs(c).max())


def test_jacobian_vs_complex_step(oracle_model):
    rng = np.random.default_rng(4)
    q = rng.uniform(-3, 3, (8, 6))
    p, J = rbd.d_eepos(oracle_model, q)
    for j in range(6):
        e = np.zeros(6, complex); e[j] = 1e-30j
        assert np.abs(J[..., :, j] - rbd.eepos(oracle_model, q + e).imag / 1e-30).max() < 1e-14
    assert np.abs(J[..., :, 5]).max() == 0        # last column: the point lies on the last joint axis


def test_rk4_energy_sanity(oracle_model):
    # zero torque, zero gravity: kinetic energy is conserved by the continuous dynamics; rk4 error is O(dt^5)
    m = oracle_model
    g = m.gravity.copy()
    try:
        m.gravity = np.zeros(3)
        q = np.array([0.3, -0.5, 0.8, 0.1, -0.4, 0.2]); v = np.array([0.5, -0.3, 0.4, 0.2, 0.1, -0.6])
        ke = lambda q, v: 0.5 * v @ rbd.crba(m, q) @ v
        e0 = ke(q, v)
        for _ in range(20):
            q, v = rbd.rk4(m, q, v, np.zeros(6), 0.01)
        assert abs(ke(q, v) - e0) / e0 < 1e-8
    finally:
        m.gravity = g
