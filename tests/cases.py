"""Seeded problem instances shared by the parity tests (sizes the numpy oracle finishes in seconds)."""
import numpy as np

from oracle import rbd

HOME = np.array([1.57994932, 0.06313131, -1.18071498, 1.09272369, -0.6255293, -0.01895007])  # sim_node.cpp:196


def figure8(N_total, dt=0.01, A_x=0.5, A_z=0.55, offset=(0.0, 0.4, 0.45), period=10.0):
    """End-effector figure-8 reference (positions only), same shape as gato_controller.py:21-46."""
    t = np.arange(N_total) * dt
    x = offset[0] + A_x * np.sin(2 * np.pi * t / period)
    y = offset[1] + 0 * t
    z = offset[2] + A_z * np.sin(4 * np.pi * t / period) / 2
    return np.stack([x, y, z], axis=1)


def make_cases(model, n, N=32, seed=0, warm=True):
    """n problems: (xu (n,18N-6), x0 (n,12), goals (n,3N)).  Mix of cold starts (XU = 0 except x0) and
    dynamically plausible warm starts (roll-out of a random smooth torque profile)."""
    rng = np.random.default_rng(seed)
    nz = 18 * N - 6
    xu = np.zeros((n, nz)); x0 = np.zeros((n, 12)); goals = np.zeros((n, 3 * N))
    fig = figure8(4000)
    for b in range(n):
        q0 = HOME + rng.normal(0, 0.3, 6) if b % 3 else rng.uniform(-1.5, 1.5, 6)
        v0 = rng.normal(0, 0.3, 6)
        x0[b] = np.concatenate([q0, v0])
        if b % 2 == 0:
            goals[b] = np.tile(rbd.eepos(model, rng.uniform(-1, 1, 6)), N)
        else:
            o = int(rng.integers(0, 3000))
            goals[b] = fig[o:o + N].reshape(-1)
        if warm and b % 4 != 0:
            q, v = q0.copy(), v0.copy()
            base = rbd.rnea(model, q, 0 * v, 0 * v)      # gravity compensation + smooth perturbation
            amp = rng.normal(0, 3, 6); ph = rng.uniform(0, 6, 6)
            traj = []
            for k in range(N):
                u = base + amp * np.sin(0.2 * k + ph)
                traj.append(np.concatenate([q, v, u]))
                a = rbd.aba(model, q, v, u)
                q, v = q + v * 0.01, v + a * 0.01
                base = rbd.rnea(model, q, 0 * v, 0 * v)
            full = np.concatenate(traj)[:nz]
            xu[b] = full + rng.normal(0, 1e-3, nz)          # slightly infeasible
            xu[b, :12] = x0[b] + rng.normal(0, 1e-3, 12)
        else:
            xu[b, :12] = x0[b]
    return xu, x0, goals
