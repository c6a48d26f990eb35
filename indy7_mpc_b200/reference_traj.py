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


def figure_8(x_amplitude, z_amplitude, offset, timestep, period, num_periods, angle_offset=np.pi / 4):
    """Same curve with the signature of src/mpc/mpc/reference_traj.py:4-28 (rotation angle as a parameter);
    returns the flat [x y z 0 0 0] * samples * num_periods array."""
    t = np.linspace(0.0, 2.0 * np.pi, int(period / timestep))
    px = offset[0] + x_amplitude * np.sin(t)
    py = np.full_like(t, offset[1])
    pz = offset[2] + z_amplitude * np.sin(2.0 * t) / 2.0 + z_amplitude / 2.0
    c, s = np.cos(angle_offset), np.sin(angle_offset)
    pts = np.zeros((t.size, 6))
    pts[:, 0] = c * px - s * py
    pts[:, 1] = s * px + c * py
    pts[:, 2] = pz
    return np.tile(pts.reshape(-1), num_periods)


# ---- on-disk formats of the reference drivers -----------------------------------------------------------------
def load_trajectory_csv(path):
    """End-effector trajectory file of the C++ driver (indy7-sdk/src/trajopt/include/csv_utils.h:34-52 reads it as one
    flat vector; trajectories/figure8_traj_eePos_meters.csv): T rows of ``x,y,z,0,0,0``.  Returns (T, 6) float64."""
    rows = np.loadtxt(path, delimiter=",", ndmin=2)
    if rows.shape[1] != 6:
        raise ValueError(f"{path}: expected 6 comma-separated values per row, found {rows.shape[1]}")
    return rows


def save_trajectory_csv(path, traj):
    """Writes a flat [x y z 0 0 0]* array, a (T, 6) or a (T, 3) array in the format of that file ('%.18e', commas)."""
    a = np.asarray(traj, float)
    if a.ndim == 1:
        a = a.reshape(-1, 6)
    if a.shape[1] == 3:
        a = np.hstack([a, np.zeros((a.shape[0], 3))])
    if a.shape[1] != 6:
        raise ValueError("trajectory must have 3 or 6 columns")
    np.savetxt(path, a, delimiter=",", fmt="%.18e")


class StatsLog:
    """The six per-tick series gato_controller.py records (:188-193,:243-247) and saves as
    ``<base>_{dts,tracking_errors,ee_positions,ee_ref_positions,joint_positions,solve_times}.npy`` (:281-293), the
    files notebooks/stat_analysis.ipynb reads."""
    FIELDS = ("dts", "tracking_errors", "ee_positions", "ee_ref_positions", "joint_positions", "solve_times")

    def __init__(self):
        for f in self.FIELDS:
            setattr(self, f, [])

    def record(self, dt, tracking_error, ee_pos, ee_ref_pos, joint_pos, solve_time):
        self.dts.append(float(dt))
        self.tracking_errors.append(float(tracking_error))
        self.ee_positions.append(np.asarray(ee_pos, float)[:3].copy())
        self.ee_ref_positions.append(np.asarray(ee_ref_pos, float)[:3].copy())
        self.joint_positions.append(np.asarray(joint_pos, float)[:6].copy())
        self.solve_times.append(float(solve_time))

    def save(self, base_filename):
        for f in self.FIELDS:
            np.save(f"{base_filename}_{f}.npy", np.array(getattr(self, f)))

    @classmethod
    def load(cls, base_filename):
        log = cls()
        for f in cls.FIELDS:
            setattr(log, f, list(np.load(f"{base_filename}_{f}.npy")))
        return log
