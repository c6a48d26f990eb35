"""The C ABI from a C++ host: examples/trajopt_node.cpp (the tick of the reference's C++ driver, TrajoptSolver::
shiftTrajectory + runTrajoptIteration, indy7-sdk/src/trajopt/include/trajopt_solver.cuh:97-175) runs as its own process —
no Python, no torch — and must produce, number for number, what the same sequence of calls gives through the Python
binding."""
import os
import re
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _python_ticks(pkg, goal, N, ticks):
    e = pkg.Engine(N=N, batch=1, dtype="fp64")
    nz, dt = 18 * N - 6, 0.01
    goal6 = goal.reshape(-1, 6)
    glen = goal6.shape[0]
    window = lambda off: np.asarray(e.goal_window(goal6, None, tick=off, wrap=False, out_stride=6, batch=1), float)
    xu, ee = np.zeros((1, nz)), window(0)
    state = np.array([1.57994932, 0.06313131, -1.18071498, 1.09272369, -0.6255293, -0.01895007, 0, 0, 0, 0, 0, 0.0])
    offset, last, rows = 0, 0.0, []
    for k in range(ticks):
        if offset >= glen:
            break
        now = 0.005 + 0.013 * k
        if last == 0:                                    # trajopt_solver.cuh:103-106: the first message only starts the clock
            last = now
        else:
            since = now - last
            n = int(since / dt)
            since_shift = since - n * dt
            n = min(n, glen - offset)
            if n > 0:
                ee = window(offset)
                offset += n
                xu, _ = e.shift(xu, steps=min(n, N - 1), tail="hold", want_x0=False)
                xu = np.asarray(xu, float)
                last = now - since_shift
            xu[0, :12] = state
        out, st = e.solve(xu, xu[:, :12].copy(), ee)
        xu = np.asarray(out, float).copy()
        st = np.asarray(st, float).reshape(-1)
        rows.append((k, offset, int(st[0]), "Success" if st[3] == 0 else "Failure", xu[0, 12:30].copy()))
        state = xu[0, 18:30].copy()
    return rows


@pytest.mark.parametrize("N", [32, 128])
def test_cpp_host_matches_python_binding(tmp_path, N):
    import indy7_mpc_b200 as pkg
    from indy7_mpc_b200.reference_traj import figure8
    exe = os.path.join(ROOT, "examples", "_build", "trajopt_node")
    if not os.path.exists(exe):
        import __graft_entry__ as g
        exe = g.build_cpp_example()
    model = pkg.build_model_from_urdf()
    mpath, gpath = str(tmp_path / "model.bin"), str(tmp_path / "goal.bin")
    open(mpath, "wb").write(bytes(model.to_c()))
    goal = figure8(A_x=0.5, A_z=0.55, offset=[0.0, 0.4, 0.45], period=10, dt=0.01, cycles=1)[:6 * 60]   # 60 points: the run ends
    goal.astype(np.float64).tofile(gpath)                                                                 # in the clamped tail
    ticks = 60
    r = subprocess.run([exe, mpath, gpath, str(N), str(ticks)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    lines = r.stdout.strip().split("\n")
    assert lines[0].startswith("indy7_b200")
    ref = _python_ticks(pkg, goal, N, ticks)
    assert len(lines) - 1 == len(ref) and 30 < len(ref) < ticks        # the C++ loop stopped at isTrajectoryComplete()
    shifts = set()
    for line, (k, offset, its, status, nums) in zip(lines[1:], ref):
        m = re.match(r"tick (\d+) offset (\d+) \| SQP iterations: (\d+), Exit status: (\w+) \|(.*)", line)
        assert m, line
        assert (int(m.group(1)), int(m.group(2)), int(m.group(3)), m.group(4)) == (k, offset, its, status), line
        got = np.array([float(x) for x in m.group(5).split()])
        assert np.array_equal(got, nums), (k, np.abs(got - nums).max())
        shifts.add(offset)
    assert len(shifts) > 20
