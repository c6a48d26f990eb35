"""ctypes binding of libindy7_b200.so (the C ABI declared in include/indy7_b200.h).

There is NO CPU fallback: if the CUDA library is missing or cannot be loaded the import of any solver class
raises, and `i7_create` fails when no CUDA device is present.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libindy7_b200.so")

I7_NJ, I7_STATS_LEN = 6, 12
I7_KEEP_BEST, I7_KEEP_NONE, I7_RESAMPLE_KEEP_TORQUE, I7_RESAMPLE_APPLY = -1, -2, 1, 2
I7_STATIC_MAX_BATCH = 4096     # lanes of the static scratch of i7_find_best_idx_static
I7_F64, I7_F32 = 0, 1
I7_QP_EXACT, I7_QP_OSQP = 0, 1
I7_OPT_X0_INTO_WARM_START = 1


class I7Model(C.Structure):
    _fields_ = [("R", C.c_double * 9 * I7_NJ), ("p", C.c_double * 3 * I7_NJ), ("mass", C.c_double * I7_NJ),
                ("com", C.c_double * 3 * I7_NJ), ("Ic", C.c_double * 9 * I7_NJ), ("gravity", C.c_double * 3),
                ("q_min", C.c_double * I7_NJ), ("q_max", C.c_double * I7_NJ), ("v_max", C.c_double * I7_NJ),
                ("u_max", C.c_double * I7_NJ)]


class I7Config(C.Structure):
    _fields_ = [("N", C.c_int32), ("batch", C.c_int32), ("dtype", C.c_int32), ("device", C.c_int32),
                ("dt", C.c_double), ("dQ_cost", C.c_double), ("R_cost", C.c_double), ("QN_cost", C.c_double),
                ("eps", C.c_double), ("regularize", C.c_int32), ("max_qp", C.c_int32), ("mu", C.c_double),
                ("step_tol", C.c_double), ("threads", C.c_int32), ("cluster", C.c_int32), ("qp_mode", C.c_int32),
                ("limits", C.c_int32), ("admm_rho", C.c_double), ("admm_rho_u", C.c_double), ("admm_sigma", C.c_double),
                ("admm_alpha", C.c_double), ("admm_eps_abs", C.c_double), ("admm_eps_rel", C.c_double),
                ("admm_max_iter", C.c_int32), ("admm_check", C.c_int32), ("model", I7Model)]


# every symbol include/indy7_b200.h declares: name -> (restype, argtypes)
_vp, _i32, _dbl = C.c_void_p, C.c_int32, C.c_double
SYMBOLS = {
    "i7_version": (C.c_char_p, []),
    "i7_default_config": (None, [C.POINTER(I7Config)]),
    "i7_create": (_i32, [C.POINTER(I7Config), C.POINTER(_vp)]),
    "i7_destroy": (None, [_vp]),
    "i7_last_error": (C.c_char_p, [_vp]),
    "i7_set_weights": (_i32, [_vp, _dbl, _dbl, _dbl, _i32, _dbl]),
    "i7_set_option": (_i32, [_vp, _i32, _i32]),
    "i7_set_external_wrench": (_i32, [_vp, C.POINTER(_dbl), _i32]),
    "i7_reset": (_i32, [_vp]),
    "i7_reset_rho": (_i32, [_vp]),
    "i7_reset_lambda": (_i32, [_vp]),
    "i7_solve": (_i32, [_vp, _i32, _vp, _vp, _vp, _i32, _vp, _vp, _vp]),
    "i7_last_solve_time_us": (_i32, [_vp, C.POINTER(_dbl)]),
    "i7_launch_count": (C.c_int64, [_vp]),
    "i7_kernel_form": (C.c_int, [_vp, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "i7_pack_best": (_i32, [_vp, _i32, _vp, _vp, _i32, _vp, _vp]),
    "i7_reduce_best": (_i32, [_vp, _i32, _i32, _vp, _vp, _vp]),
    "i7_p2p_create": (_i32, [_vp, _i32, _i32, C.c_char_p]),
    "i7_p2p_connect": (_i32, [_vp, C.c_char_p]),
    "i7_p2p_exchange": (_i32, [_vp, _i32, _vp, _vp, _i32, _vp, _vp]),
    "i7_p2p_gathered": (_i32, [_vp, C.POINTER(_vp)]),
    "i7_solve_select": (_i32, [_vp, _i32, _vp, _vp, _vp, _i32, _vp, _vp, _i32, _vp, _vp]),
    "i7_exchange_wait_ns": (_i32, [_vp, C.POINTER(C.c_int64), _i32]),
    "i7_p2p_status": (_i32, [_vp, _vp]),
    "i7_p2p_reset": (_i32, [_vp]),
    "i7_measure_fma_peak": (_i32, [_vp, C.POINTER(_dbl)]),
    "i7_profile_phases": (_i32, [_vp, _i32, C.POINTER(C.c_int64), _i32]),
    "i7_sim_forward": (_i32, [_vp, _i32, _vp, _vp, _dbl, _i32, _vp, _vp]),
    "i7_shift": (_i32, [_vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _vp]),
    "i7_goal_window": (_i32, [_vp, _i32, _vp, _i32, _i32, _vp, _i32, _i32, _i32, _vp, _vp]),
    "i7_find_best_idx": (_i32, [_vp, _i32, _vp, _vp, _vp, _dbl, _i32, _vp, _vp, _vp, _vp]),
    "i7_find_best_idx_static": (_i32, [_vp, _i32, _vp, _vp, _vp, _dbl, _i32, _vp, _vp, _vp, _vp]),
    "i7_resample_wrench": (_i32, [_vp, _i32, _i32, _vp, _vp, _vp, _i32, _i32, _dbl, _vp, _vp]),
    "i7_broadcast_best": (_i32, [_vp, _i32, _vp, _i32, _vp, _vp, _vp, _vp]),
    "i7_qp_state": (_i32, [_vp, _i32, _vp, _vp, _vp, _vp, _vp]),
    "i7_solve_qp": (_i32, [_vp, _i32, _vp, _vp, _vp, _i32, _vp, _vp, _vp]),
    "i7_dump_qp": (_i32, [_vp, _i32, _vp, _vp, _vp, _i32, _vp, _vp, _vp, _vp, _vp]),
    "i7_merit": (_i32, [_vp, _i32, _vp, _vp, _vp, _i32, _vp, _vp]),
    "i7_linesearch": (_i32, [_vp, _i32, _vp, _vp, _vp, _i32, _vp, _vp, _vp]),
    "i7_dynamics": (_i32, [_vp, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "i7_eepos": (_i32, [_vp, _i32, _vp, _vp, _vp, _vp]),
}

_lib = None


def load():
    """Load the CUDA library; raises RuntimeError (never falls back) when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -m indy7_mpc_b200.build` (nvcc, sm_100a). "
            "indy7_mpc_b200 has no CPU fallback.")
    try:        # a library older than the sources next to it is a silent source of wrong conclusions: say so
        from . import build as _build
        if _build.needs_build():
            import warnings
            warnings.warn(f"{LIB_PATH} is older than the sources under csrc/ (build stamp mismatch): "
                          "rebuild with `python -m indy7_mpc_b200.build`", RuntimeWarning, stacklevel=2)
    except OSError:
        pass
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError if a declared symbol is not exported
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib
