"""i7_find_best_idx_static — the hypothesis selection with its scratch memory in a static device array (no allocation
inside the call; no counterpart in the reference) — against the existing
entry point i7_find_best_idx on the same inputs, through the C ABI, on two calls in a row.  Collected last on
purpose."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-9          # the tolerance of test_find_best_idx_fused_selection


def _case(rng, B):
    x_last = np.concatenate([rng.uniform(-1, 1, 6), rng.uniform(-0.5, 0.5, 6)])
    u_last = rng.uniform(-5, 5, 6)
    f = np.zeros((B, 6))
    f[:, :3] = rng.normal(0, 5, (B, 3))
    if B >= 7:
        f[5] = f[2]                 # exact tie
        f[3, 0] = np.nan            # a poisoned lane: NaN error, never wins
    return x_last, u_last, f


def _raw(e, fn, B, x_last, u_last, xs, dt, tie_last):
    """One call straight through ctypes with host buffers: (result[2], errors[B], x_next[B, 12])."""
    dt_np = np.float64 if e.dtype == "fp64" else np.float32
    a = lambda v: np.ascontiguousarray(v, dtype=dt_np)
    x_last, u_last, xs = a(x_last), a(u_last), a(xs)
    res, errs, xn = np.full(2, -7, dt_np), np.full(B, -7, dt_np), np.full((B, 12), -7, dt_np)
    p = lambda v: v.ctypes.data_as(ctypes.c_void_p)
    rc = fn(e._h, B, p(x_last), p(u_last), p(xs), float(dt), int(tie_last), p(xn), p(errs), p(res), None)
    assert rc == 0, e.lib.i7_last_error(e._h).decode()
    return res, errs, xn


@pytest.mark.parametrize("dtype", ["fp64", "fp32"])
def test_static_scratch_selection_equals_allocating_entry_point(dtype):
    import torch
    import indy7_mpc_b200 as pkg
    rng = np.random.default_rng(2024)
    B1, B2 = 300, 64
    e = pkg.Engine(N=8, batch=B1, dtype=dtype)
    # two calls in a row on the shared array: the second with fewer lanes, another measured state and the other tie
    # rule, so that anything the first call left behind (ticket counter, errors of lanes 64..299) would show
    for call, (B, tie_last) in enumerate([(B1, True), (B2, False)]):
        x_last, u_last, f = _case(rng, B)
        fw = np.zeros((B1, 6)); fw[:B] = f
        e.set_external_wrench(fw)
        # the measured state: the prediction of lane 2 (from the existing entry point's own plant step)
        _, _, xn0 = _raw(e, e.lib.i7_find_best_idx, B, x_last, u_last, np.zeros(12), 0.002, tie_last)
        xs = xn0[2].astype(float)
        res_a, err_a, xn_a = _raw(e, e.lib.i7_find_best_idx, B, x_last, u_last, xs, 0.002, tie_last)
        res_s, err_s, xn_s = _raw(e, e.lib.i7_find_best_idx_static, B, x_last, u_last, xs, 0.002, tie_last)
        assert res_s[0] == res_a[0] == (5 if tie_last else 2), (call, res_s, res_a)
        assert abs(res_s[1] - res_a[1]) <= TOL
        ok = ~np.isnan(err_a)
        assert ok.sum() == B - 1 and np.array_equal(np.isnan(err_s), ~ok)
        assert np.abs(err_s[ok] - err_a[ok]).max() <= TOL
        assert np.abs(xn_s[ok] - xn_a[ok]).max() <= TOL
        # Engine wrapper, host and device buffers, errors not requested (they stay in the static array)
        assert e.find_best_idx_static(x_last, u_last, xs, 0.002, batch=B, tie_last=tie_last) == int(res_a[0])
        d = lambda v: torch.tensor(v, device="cuda", dtype=e._torch_dtype())
        idx, errs_d = e.find_best_idx_static(d(x_last), d(u_last), d(xs), 0.002, batch=B, tie_last=tie_last, return_errors=True)
        assert idx == int(res_a[0]) and errs_d.is_cuda
        assert np.abs(errs_d.cpu().numpy()[ok] - err_a[ok]).max() <= TOL
    # no lane qualifies: -1, as the existing entry point
    assert e.find_best_idx_static(x_last, u_last, np.full(12, np.nan), 0.002, batch=B2) is None
    assert e.find_best_idx(x_last, u_last, np.full(12, np.nan), 0.002, batch=B2) is None
    # ... and the call after that is not disturbed by it
    assert e.find_best_idx_static(x_last, u_last, xs, 0.002, batch=B2, tie_last=False) == 2


def test_static_scratch_limit_is_enforced():
    import indy7_mpc_b200 as pkg
    from indy7_mpc_b200 import _lib
    assert _lib.I7_STATIC_MAX_BATCH == 4096
    B = _lib.I7_STATIC_MAX_BATCH + 1
    e = pkg.Engine(N=8, batch=B, dtype="fp64")
    rng = np.random.default_rng(7)
    x_last, u_last, f = _case(rng, B)
    e.set_external_wrench(f)
    xs = np.zeros(12)
    with pytest.raises(ValueError, match="I7_STATIC_MAX_BATCH"):
        e.find_best_idx_static(x_last, u_last, xs, 0.002, batch=B)
    # the maximum itself is served, and agrees with the allocating entry point
    full = _lib.I7_STATIC_MAX_BATCH
    i_s, e_s = e.find_best_idx_static(x_last, u_last, xs, 0.002, batch=full, return_errors=True)
    i_a, e_a = e.find_best_idx(x_last, u_last, xs, 0.002, batch=full, return_errors=True)
    ok = ~np.isnan(e_a)
    assert i_s == i_a and np.array_equal(np.isnan(e_s), ~ok) and np.abs(e_s[ok] - e_a[ok]).max() <= TOL
