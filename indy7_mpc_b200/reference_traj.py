"""End-effector reference generators used by the reference drivers (gato_controller.py:21-46 ``figure8``)."""
import numpy as np


def figure8(A_x=0.5, A_z=0.55, offset=(0.0, 0.4, 0.45), period=10.0, dt=0.01, cycles=10, stride=6):
    """Figure-8 in the x-z plane rotated 45 degrees about z, sampled at ``int(period/dt)`` points per cycle over
    [0, 2 pi] (end point included, like the reference), tiled ``cycles`` times.  Returns a flat array of
    ``stride`` numbers per sample: [x y z 0 0 0] for stride 6 (GATO goal format) or [x y z] for stride 3."""
    t = np.linspace(0.0, 2.0 * np.pi, int(period / dt))
    px = offset[0] + A_x * np.sin(t)
    py = np.full_like(t, offset[1])
    pz = offset[2] + A_z * np.sin(2.0 * t) / 2.0 + A_z / 2.0
    c = s = np.sqrt(0.5)
    pts = np.zeros((t.size, stride))
    pts[:, 0] = c * px - s * py
    pts[:, 1] = s * px + c * py
    pts[:, 2] = pz
    return np.tile(pts.reshape(-1), cycles)
