"""URDF -> robot constants, standing in for ``pin.buildModelsFromUrdf`` on the SQP-MPC path
(reference: src/utils.py:20-21; description/indy7.urdf:64-245).

``RobotModel`` exposes the attributes the reference callers read from a Pinocchio model (``nq``, ``nv``,
``joints``, ``njoints``, ``gravity.linear``, ``createData()``; src/osqp_solver.py:10,17-20, src/osqp_mpc.py:8-9)
and fills the ``i7_model`` struct of the C ABI.  Supported: serial chains of exactly six revolute joints about
their local z axis, root welded to the world through fixed joints.
"""
import math
import os
import xml.etree.ElementTree as ET

import numpy as np

from ._lib import I7Model

DEFAULT_URDF = os.path.join(os.path.dirname(os.path.abspath(__file__)), "description", "indy7.urdf")


class _Gravity:
    def __init__(self):
        self.linear = np.array([0.0, 0.0, -9.81])
        self.angular = np.zeros(3)


def _rpy(r, p, y):
    cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    return np.array([[cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
                     [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
                     [-sp, cp * sr, cp * cr]])


def _floats(s):
    return [float(x) for x in s.split()]


class RobotModel:
    def __init__(self, placements_R, placements_p, mass, com, Ic, limits, name="indy7"):
        self.name = name
        self.R = np.asarray(placements_R, float)
        self.p = np.asarray(placements_p, float)
        self.mass = np.asarray(mass, float)
        self.com = np.asarray(com, float)
        self.Ic = np.asarray(Ic, float)
        n = len(self.mass)
        if n != 6:
            raise ValueError(f"the sm_100a kernels are specialised for 6 joints, URDF has {n}")
        self.nq = self.nv = n
        self.njoints = n + 1                       # Pinocchio counts the universe joint
        self.joints = [None] * self.njoints        # len(model.joints) - 1 == nu (src/osqp_solver.py:20)
        self.lowerPositionLimit = np.array([l[0] for l in limits])
        self.upperPositionLimit = np.array([l[1] for l in limits])
        self.effortLimit = np.array([l[2] for l in limits])
        self.velocityLimit = np.array([l[3] for l in limits])
        self.gravity = _Gravity()

    def createData(self):
        return None

    def last_joint_placement(self, q):
        """(R, p) of the last joint frame in the world (Pinocchio ``data.oMi[6]``): frame bookkeeping for the drivers
        that re-express a world wrench in that frame (src/gato_mpc_batch_sample.py:268-277); not on the solve path."""
        Rw, pw = np.eye(3), np.zeros(3)
        for i in range(6):
            c, s_ = math.cos(float(q[i])), math.sin(float(q[i]))
            pw = pw + Rw @ self.p[i]
            Rw = Rw @ self.R[i].reshape(3, 3) @ np.array([[c, -s_, 0.0], [s_, c, 0.0], [0.0, 0.0, 1.0]])
        return Rw, pw

    def world_wrench_to_last_joint(self, q, wrench):
        """``oMi[6].actInv(pin.Force(linear, angular))`` for a wrench [fx fy fz tx ty tz] given at the world origin."""
        Rw, pw = self.last_joint_placement(q)
        w = np.asarray(wrench, float)
        f, n = w[..., :3], w[..., 3:]
        return np.concatenate([f @ Rw, (n - np.cross(pw, f)) @ Rw], axis=-1)

    def to_c(self):
        m = I7Model()
        g = np.asarray(getattr(self.gravity, "linear", self.gravity), float).reshape(-1)
        for i in range(6):
            for a in range(9):
                m.R[i][a] = float(self.R[i].reshape(-1)[a])
                m.Ic[i][a] = float(self.Ic[i].reshape(-1)[a])
            for a in range(3):
                m.p[i][a] = float(self.p[i][a])
                m.com[i][a] = float(self.com[i][a])
            m.mass[i] = float(self.mass[i])
        for a in range(3):
            m.gravity[a] = float(g[a])
        big = 1e30                                   # "no limit" (URDF joints without a <limit> element)
        for i in range(6):
            m.q_min[i] = float(max(self.lowerPositionLimit[i], -big))
            m.q_max[i] = float(min(self.upperPositionLimit[i], big))
            m.v_max[i] = float(min(self.velocityLimit[i], big))
            m.u_max[i] = float(min(self.effortLimit[i], big))
        return m


def build_model_from_urdf(urdf_path=DEFAULT_URDF):
    root = ET.parse(urdf_path).getroot()
    links = {l.get("name"): l for l in root.findall("link")}
    by_parent = {}
    child_names = set()
    for j in root.findall("joint"):
        by_parent.setdefault(j.find("parent").get("link"), []).append(j)
        child_names.add(j.find("child").get("link"))
    roots = [n for n in links if n not in child_names]
    if len(roots) != 1:
        raise ValueError(f"expected one root link, found {roots}")
    R, p, mass, com, Ic, lim = [], [], [], [], [], []
    cur = roots[0]
    while cur in by_parent:
        if len(by_parent[cur]) != 1:
            raise ValueError("only serial chains are supported")
        j = by_parent[cur][0]
        child = j.find("child").get("link")
        inertial = links[child].find("inertial")
        if j.get("type") == "fixed":
            if mass and inertial is not None:
                raise ValueError("a fixed link with inertia below a moving joint is not supported")
            cur = child
            continue
        if j.get("type") not in ("revolute", "continuous") or _floats(j.find("axis").get("xyz")) != [0.0, 0.0, 1.0]:
            raise ValueError("only revolute joints about the local z axis are supported")
        o = j.find("origin")
        R.append(_rpy(*_floats(o.get("rpy"))))
        p.append(_floats(o.get("xyz")))
        io = inertial.find("origin")
        if io is not None and any(_floats(io.get("rpy", "0 0 0"))):
            raise ValueError("rotated inertial frames are not supported")
        com.append(_floats(io.get("xyz")) if io is not None else [0.0, 0.0, 0.0])
        mass.append(float(inertial.find("mass").get("value")))
        I = inertial.find("inertia")
        v = {k: float(I.get(k)) for k in ("ixx", "ixy", "ixz", "iyy", "iyz", "izz")}
        Ic.append([[v["ixx"], v["ixy"], v["ixz"]], [v["ixy"], v["iyy"], v["iyz"]], [v["ixz"], v["iyz"], v["izz"]]])
        l = j.find("limit")
        lim.append((float(l.get("lower", "-inf")), float(l.get("upper", "inf")), float(l.get("effort", "inf")),
                    float(l.get("velocity", "inf"))) if l is not None else (-math.inf, math.inf, math.inf, math.inf))
        cur = child
    return RobotModel(R, p, mass, com, Ic, lim, name=root.get("name", "robot"))
