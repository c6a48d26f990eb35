"""Host helpers mirroring the reference's src/utils.py (load_robot_model :20-21, rk4 :3-18, print_stats :23-38)."""
import numpy as np

from .model import RobotModel, build_model_from_urdf

_plant_engines = {}


def load_robot_model(urdf_path=None, mesh_dir=None):
    """Same return shape as the reference (model, visual_model, collision_model); geometry models are not part of
    the SQP-MPC path and are returned as None."""
    model = build_model_from_urdf(urdf_path) if urdf_path else build_model_from_urdf()
    return model, None, None


def _plant(model):
    from .engine import Engine
    key = id(model)
    if key not in _plant_engines:
        _plant_engines[key] = Engine(model, N=2, batch=1, dtype="fp64")
    return _plant_engines[key]


def rk4(model, data, q, v, u, dt, f_ext=None):
    """One rk4 plant step on the GPU (i7_sim_forward).  ``f_ext``: None, or a 6-vector [fx fy fz tx ty tz] in the
    local frame of the last joint (the reference passes a StdVec_Force whose last entry is that wrench)."""
    eng = _plant(model)
    if f_ext is not None and not isinstance(f_ext, (list, tuple, np.ndarray)):
        raise TypeError("f_ext must be None or a 6-vector for the last joint")
    if f_ext is not None and np.ndim(f_ext) == 2:      # list of per-joint wrenches: only the last may be non-zero
        arr = np.asarray(f_ext, float)
        if np.abs(arr[:-1]).max() > 0:
            raise NotImplementedError("external wrenches are supported on the last joint only")
        f_ext = arr[-1]
    eng.set_external_wrench(None if f_ext is None else np.asarray(f_ext, float).reshape(1, 6))
    x = np.concatenate([np.asarray(q, float).reshape(-1), np.asarray(v, float).reshape(-1)])
    xn = eng.sim_forward(x, np.asarray(u, float).reshape(-1), dt)[0]
    return xn[:6].copy(), xn[6:].copy()


def print_stats(stats):
    for task, stat in stats.items():
        values = stat['values']
        if not values:
            continue
        mult, unit = stat['multiplier'], stat['unit']
        print(f"{task}:")
        print(f"  avg: {mult * sum(values) / len(values):.2f} {unit}")
        print(f"  min: {mult * min(values):.2f} {unit}")
        print(f"  max: {mult * max(values):.2f} {unit}")
        print()
