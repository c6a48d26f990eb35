"""Rigid-body kinematics and dynamics of a serial revolute-z chain, batched numpy.  TEST INFRASTRUCTURE.

Restates the published algorithms behind the Pinocchio calls on the reference hot path:
  pin.forwardKinematics / data.oMi[6].translation     src/osqp_solver.py:146-148
  pin.computeJointJacobians + getJointJacobian(6, LOCAL_WORLD_ALIGNED)[:3]   src/osqp_solver.py:150-155
  pin.aba                                             src/osqp_sqp.py:40, src/utils.py:5-13
  pin.computeABADerivatives (ddq_dq, ddq_dv, Minv)    src/osqp_solver.py:71
Conventions: Featherstone spatial vectors [angular; linear]; link frame i = joint frame i after the joint
rotation; every joint has motion subspace S = e_z(angular).  All functions broadcast over leading
dimensions of their arguments and are dtype-generic so that complex-step differentiation can validate the
analytic derivatives to machine precision (``aba_derivatives_complex_step``).
"""
import numpy as np


def _cross(a, b):
    return np.stack([a[..., 1] * b[..., 2] - a[..., 2] * b[..., 1],
                     a[..., 2] * b[..., 0] - a[..., 0] * b[..., 2],
                     a[..., 0] * b[..., 1] - a[..., 1] * b[..., 0]], axis=-1)


def _mv(M, x):
    return np.einsum('...ij,...j->...i', M, x)


def _mtv(M, x):
    return np.einsum('...ji,...j->...i', M, x)


def joint_rotations(model, q):
    """E[..., i, :, :] = (R_i Rz(q_i))^T : rotation taking parent-frame coordinates to link-i coordinates."""
    q = np.asarray(q)
    c, s = np.cos(q), np.sin(q)
    R = model.R  # (n,3,3)
    col0 = R[:, :, 0] * c[..., None] + R[:, :, 1] * s[..., None]
    col1 = -R[:, :, 0] * s[..., None] + R[:, :, 1] * c[..., None]
    col2 = R[:, :, 2] + 0 * c[..., None]
    return np.stack([col0, col1, col2], axis=-2)  # rows of E are the columns of R Rz


def xmotion(E, r, m):
    """child <- parent motion transform: [E w ; E (v - r x w)]."""
    w, v = m[..., :3], m[..., 3:]
    return np.concatenate([_mv(E, w), _mv(E, v - _cross(r, w))], axis=-1)


def xforce_t(E, r, f):
    """parent <- child force transform X^T f: [E^T n + r x (E^T f) ; E^T f]."""
    n, l = _mtv(E, f[..., :3]), _mtv(E, f[..., 3:])
    return np.concatenate([n + _cross(r, l), l], axis=-1)


def crm(v, m):
    """spatial motion cross product v x m."""
    return np.concatenate([_cross(v[..., :3], m[..., :3]),
                           _cross(v[..., :3], m[..., 3:]) + _cross(v[..., 3:], m[..., :3])], axis=-1)


def crf(v, f):
    """spatial force cross product v x* f."""
    return np.concatenate([_cross(v[..., :3], f[..., :3]) + _cross(v[..., 3:], f[..., 3:]),
                           _cross(v[..., :3], f[..., 3:])], axis=-1)


def _S(dtype):
    S = np.zeros(6, dtype=dtype)
    S[2] = 1
    return S


def _ext(model, f_ext, dtype, shape):
    """Pinocchio-ordered local wrench on the LAST link [linear; angular] -> Featherstone [angular; linear]."""
    if f_ext is None:
        return None
    f_ext = np.asarray(f_ext)
    return np.concatenate([f_ext[..., 3:], f_ext[..., :3]], axis=-1)


def forward_kinematics(model, q):
    """World placement of every joint frame: returns (p (...,n,3), Rw (...,n,3,3)).
    eepos = p[..., n-1, :] is ``data.oMi[6].translation`` (no tcp offset; src/osqp_solver.py:146-148)."""
    q = np.asarray(q)
    E = joint_rotations(model, q)
    ps, Rs = [], []
    p = np.zeros(q.shape[:-1] + (3,), dtype=E.dtype)
    Rw = np.broadcast_to(np.eye(3, dtype=E.dtype), q.shape[:-1] + (3, 3))
    for i in range(model.n):
        p = p + _mv(Rw, model.p[i])
        Rw = Rw @ np.swapaxes(E[..., i, :, :], -1, -2)
        ps.append(p); Rs.append(Rw)
    return np.stack(ps, axis=-2), np.stack(Rs, axis=-3)


def eepos(model, q):
    return forward_kinematics(model, q)[0][..., -1, :]


def d_eepos(model, q):
    """(eepos (...,3), J (...,3,n)) with J[:, i] = z_i x (p_ee - p_i): the translational rows of the
    LOCAL_WORLD_ALIGNED Jacobian of the last joint (src/osqp_solver.py:150-155)."""
    p, Rw = forward_kinematics(model, q)
    pe = p[..., -1, :]
    cols = [_cross(Rw[..., i, :, 2], pe - p[..., i, :]) for i in range(model.n)]
    return pe, np.stack(cols, axis=-1)


def _forward_vel(model, E, qd, dtype):
    S = _S(dtype)
    vel = []
    v = None
    for i in range(model.n):
        vj = S * qd[..., i, None]
        v = vj if v is None else xmotion(E[..., i, :, :], model.p[i], v) + vj
        vel.append(v)
    return vel


def rnea(model, q, qd, qdd, f_ext=None, gravity=True, return_internal=False):
    """Inverse dynamics tau = ID(q, qd, qdd) (Featherstone RBDA Table 5.1), optional local wrench on last link."""
    q, qd, qdd = np.asarray(q), np.asarray(qd), np.asarray(qdd)
    dtype = np.result_type(q, qd, qdd, np.float64)
    n = model.n
    E = joint_rotations(model, q).astype(dtype)
    S = _S(dtype)
    a0 = np.zeros(q.shape[:-1] + (6,), dtype=dtype)
    if gravity:
        a0[..., 3:] = -model.gravity
    fx = _ext(model, f_ext, dtype, q.shape)
    vel, acc, f = [], [], []
    v, a = None, a0
    for i in range(n):
        vj = S * qd[..., i, None]
        v = vj if v is None else xmotion(E[..., i, :, :], model.p[i], v) + vj
        a = xmotion(E[..., i, :, :], model.p[i], a) + S * qdd[..., i, None] + crm(v, vj)
        fi = _mv(model.I[i], a) + crf(v, _mv(model.I[i], v))
        if fx is not None and i == n - 1:
            fi = fi - fx
        vel.append(v); acc.append(a); f.append(fi)
    tau = [None] * n
    F = [None] * n
    acc_f = None
    for i in range(n - 1, -1, -1):
        Fi = f[i] if acc_f is None else f[i] + acc_f
        F[i] = Fi
        tau[i] = Fi[..., 2]
        acc_f = xforce_t(E[..., i, :, :], model.p[i], Fi)
    tau = np.stack(tau, axis=-1)
    if return_internal:
        return tau, dict(E=E, vel=vel, acc=acc, f=f, F=F)
    return tau


def aba(model, q, qd, tau, f_ext=None):
    """Forward dynamics by the articulated-body algorithm (RBDA Table 7.1; recipe in SURVEY.md App. E)."""
    q, qd, tau = np.asarray(q), np.asarray(qd), np.asarray(tau)
    dtype = np.result_type(q, qd, tau, np.float64)
    n = model.n
    E = joint_rotations(model, q).astype(dtype)
    S = _S(dtype)
    bshape = np.broadcast_shapes(q.shape[:-1], qd.shape[:-1], tau.shape[:-1])
    fx = _ext(model, f_ext, dtype, q.shape)
    vel, c, IA, pA = [], [], [], []
    v = None
    for i in range(n):
        vj = S * qd[..., i, None]
        v = vj if v is None else xmotion(E[..., i, :, :], model.p[i], v) + vj
        vel.append(v)
        c.append(crm(v, vj))
        IA.append(np.broadcast_to(model.I[i].astype(dtype), bshape + (6, 6)).copy())
        p = crf(v, _mv(model.I[i], v))
        if fx is not None and i == n - 1:
            p = p - fx
        pA.append(np.broadcast_to(p, bshape + (6,)).copy())
    U, D, u = [None] * n, [None] * n, [None] * n
    for i in range(n - 1, -1, -1):
        U[i] = IA[i][..., :, 2]
        D[i] = U[i][..., 2]
        u[i] = tau[..., i] - pA[i][..., 2]
        if i > 0:
            Ia = IA[i] - U[i][..., :, None] * U[i][..., None, :] / D[i][..., None, None]
            pa = pA[i] + _mv(Ia, c[i]) + U[i] * (u[i] / D[i])[..., None]
            # X^T Ia X, applied column-wise through the two transforms
            Ei, ri = E[..., i, :, :], model.p[i]
            tmp = np.stack([xforce_t(Ei, ri, Ia[..., :, k]) for k in range(6)], axis=-1)      # X^T Ia
            IA[i - 1] = IA[i - 1] + np.stack([xforce_t(Ei, ri, tmp[..., k, :]) for k in range(6)], axis=-2)
            pA[i - 1] = pA[i - 1] + xforce_t(Ei, ri, pa)
    a = np.zeros(bshape + (6,), dtype=dtype)
    a[..., 3:] = -model.gravity
    qdd = []
    for i in range(n):
        ap = xmotion(E[..., i, :, :], model.p[i], a) + c[i]
        qi = (u[i] - np.einsum('...k,...k->...', U[i], ap)) / D[i]
        a = ap + S * qi[..., None]
        qdd.append(qi)
    return np.stack(qdd, axis=-1)


def crba(model, q):
    """Joint-space inertia matrix by the composite-rigid-body algorithm (RBDA Table 6.2)."""
    q = np.asarray(q)
    dtype = np.result_type(q, np.float64)
    n = model.n
    E = joint_rotations(model, q).astype(dtype)
    bshape = q.shape[:-1]
    IC = [np.broadcast_to(model.I[i].astype(dtype), bshape + (6, 6)).copy() for i in range(n)]
    for i in range(n - 1, 0, -1):
        Ei, ri = E[..., i, :, :], model.p[i]
        tmp = np.stack([xforce_t(Ei, ri, IC[i][..., :, k]) for k in range(6)], axis=-1)
        IC[i - 1] = IC[i - 1] + np.stack([xforce_t(Ei, ri, tmp[..., k, :]) for k in range(6)], axis=-2)
    M = np.zeros(bshape + (n, n), dtype=dtype)
    for i in range(n):
        F = IC[i][..., :, 2]
        M[..., i, i] = F[..., 2]
        j = i
        while j > 0:
            F = xforce_t(E[..., j, :, :], model.p[j], F)
            j -= 1
            M[..., i, j] = F[..., 2]
            M[..., j, i] = F[..., 2]
    return M


def minv(model, q):
    return np.linalg.inv(crba(model, q))


def rnea_derivatives(model, q, qd, qdd, f_ext=None):
    """Analytic d tau/dq, d tau/dqd of RNEA by forward-mode chain rule in link coordinates
    (Carpentier & Mansard 2018 in local frames; the same recursion the CUDA kernel uses)."""
    q, qd, qdd = np.asarray(q), np.asarray(qd), np.asarray(qdd)
    n = model.n
    tau, it = rnea(model, q, qd, qdd, f_ext=f_ext, return_internal=True)
    dtype = tau.dtype
    E, vel, acc, F = it['E'], it['vel'], it['acc'], it['F']
    S = _S(dtype)
    bshape = tau.shape[:-1]
    dq = np.zeros(bshape + (n, n), dtype=dtype)
    dv = np.zeros(bshape + (n, n), dtype=dtype)
    for j in range(n):
        vjS = crm(vel[j], S)                                           # v_j x S
        Xa = acc[j] - S * qdd[..., j, None] - vjS * qd[..., j, None]    # X_j a_parent
        cols = {
            'q': (vjS, crm(Xa, S) + crm(vjS, S) * qd[..., j, None]),
            'v': (np.broadcast_to(S, vjS.shape), vjS),
        }
        for name, (dvel, dacc) in cols.items():
            df = [None] * n
            for i in range(j, n):
                if i > j:
                    dvel = xmotion(E[..., i, :, :], model.p[i], dvel)
                    dacc = xmotion(E[..., i, :, :], model.p[i], dacc) + crm(dvel, S) * qd[..., i, None]
                Iv = _mv(model.I[i], vel[i])
                df[i] = _mv(model.I[i], dacc) + crf(dvel, Iv) + crf(vel[i], _mv(model.I[i], dvel))
            dF = np.zeros(bshape + (6,), dtype=dtype)
            out = dq if name == 'q' else dv
            for i in range(n - 1, -1, -1):
                if i >= j:
                    dF = dF + df[i]
                out[..., i, j] = dF[..., 2]
                if i == j and name == 'q':
                    dF = dF + crf(S, F[i])
                if i > 0:
                    dF = xforce_t(E[..., i, :, :], model.p[i], dF)
    return dq, dv


def aba_derivatives(model, q, qd, tau, f_ext=None):
    """(ddq_dq, ddq_dv, Minv, ddq) as ``pin.computeABADerivatives`` returns them plus ``data.ddq``
    (src/osqp_solver.py:71,75): dFD/dq = -Minv dID/dq |_{qdd=FD}, dFD/dv = -Minv dID/dv, dFD/dtau = Minv."""
    qdd = aba(model, q, qd, tau, f_ext=f_ext)
    Mi = minv(model, q)
    dq, dv = rnea_derivatives(model, q, qd, qdd, f_ext=f_ext)
    return -Mi @ dq, -Mi @ dv, Mi, qdd


def aba_derivatives_complex_step(model, q, qd, tau, f_ext=None, h=1e-30):
    """Same quantities by complex-step differentiation of ``aba`` (exact to rounding; validation only)."""
    q, qd, tau = (np.asarray(x, float) for x in (q, qd, tau))
    n = model.n
    dq = np.zeros(q.shape[:-1] + (n, n)); dv = np.zeros_like(dq); du = np.zeros_like(dq)
    for j in range(n):
        e = np.zeros(n, complex); e[j] = 1j * h
        dq[..., :, j] = aba(model, q + e, qd, tau, f_ext=f_ext).imag / h
        dv[..., :, j] = aba(model, q, qd + e, tau, f_ext=f_ext).imag / h
        du[..., :, j] = aba(model, q, qd, tau + e, f_ext=f_ext).imag / h
    return dq, dv, du, aba(model, q, qd, tau, f_ext=f_ext)


def rk4(model, q, v, u, dt, f_ext=None):
    """Plant integrator, literal restatement of src/utils.py:3-18 (pin.integrate == q + dq for revolute joints)."""
    k1q = v
    k1v = aba(model, q, v, u, f_ext)
    q2 = q + k1q * dt / 2
    k2q = v + k1v * dt / 2
    k2v = aba(model, q2, k2q, u, f_ext)
    q3 = q + k2q * dt / 2
    k3q = v + k2v * dt / 2
    k3v = aba(model, q3, k3q, u, f_ext)
    q4 = q + k3q * dt
    k4q = v + k3v * dt
    k4v = aba(model, q4, k4q, u, f_ext)
    v_next = v + (dt / 6) * (k1v + 2 * k2v + 2 * k3v + k4v)
    avg_v = (k1q + 2 * k2q + 2 * k3q + k4q) / 6
    q_next = q + avg_v * dt
    return q_next, v_next
