"""Robot model constants from a URDF, restating what ``pin.buildModelsFromUrdf`` does for the Indy7 file
(reference: src/utils.py:20-21 ``load_robot_model``; description/indy7.urdf:64-245).  TEST INFRASTRUCTURE.

Pinocchio semantics restated:
  * the URDF root link ``world`` and the fixed joint ``global`` weld ``link0`` to the universe, so joints
    ``joint0..joint5`` become model joints 1..6 and ``link0``'s inertia never enters the dynamics;
  * a joint ``<origin xyz rpy>`` is the placement parent-frame -> joint-frame with
    ``R = Rz(yaw) Ry(pitch) Rx(roll)`` evaluated on the literal rpy numbers of the file;
  * every joint is revolute about its local z axis; fixed child links with no inertia (``tcp``) add nothing;
  * gravity defaults to (0, 0, -9.81) (also forced by src/osqp_mpc.py:8-9).
"""
import os
import xml.etree.ElementTree as ET

import numpy as np

DEFAULT_URDF = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "indy7_mpc_b200", "description",
                            "indy7.urdf")


def rpy_matrix(roll, pitch, yaw):
    cr, sr = np.cos(roll), np.sin(roll)
    cp, sp = np.cos(pitch), np.sin(pitch)
    cy, sy = np.cos(yaw), np.sin(yaw)
    Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
    Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def skew(c):
    return np.array([[0, -c[2], c[1]], [c[2], 0, -c[0]], [-c[1], c[0], 0]])


class Model:
    """Minimal stand-in for ``pinocchio.Model`` restricted to serial chains of revolute-z joints."""

    def __init__(self, R, p, mass, com, Ic, lower, upper, effort, velocity):
        self.n = len(mass)
        self.nq = self.nv = self.n
        self.njoints = self.n + 1
        self.joints = [None] * self.njoints          # reference uses len(model.joints) - 1 (osqp_solver.py:20)
        self.R = np.asarray(R, float)                # (n,3,3) joint placement rotation (parent <- joint frame)
        self.p = np.asarray(p, float)                # (n,3)   joint placement translation in the parent frame
        self.mass = np.asarray(mass, float)
        self.com = np.asarray(com, float)
        self.Ic = np.asarray(Ic, float)              # (n,3,3) rotational inertia about the COM, link frame
        self.lower, self.upper = np.asarray(lower, float), np.asarray(upper, float)
        self.effort, self.velocity = np.asarray(effort, float), np.asarray(velocity, float)
        self.gravity = np.array([0.0, 0.0, -9.81])
        # 6x6 spatial inertia about the link-frame origin, Featherstone [angular; linear] ordering
        self.I = np.zeros((self.n, 6, 6))
        for i in range(self.n):
            C = skew(self.com[i])
            m = self.mass[i]
            self.I[i, :3, :3] = self.Ic[i] + m * C @ C.T
            self.I[i, :3, 3:] = m * C
            self.I[i, 3:, :3] = m * C.T
            self.I[i, 3:, 3:] = m * np.eye(3)

    def createData(self):
        return None


def load_model(urdf_path=DEFAULT_URDF):
    root = ET.parse(urdf_path).getroot()
    links = {l.get("name"): l for l in root.findall("link")}
    children = {}
    for j in root.findall("joint"):
        children.setdefault(j.find("parent").get("link"), []).append(j)
    all_children = {j.find("child").get("link") for j in root.findall("joint")}
    base = [n for n in links if n not in all_children]
    assert len(base) == 1, base
    cur = base[0]
    # follow fixed joints from the root: those links are welded to the universe
    R, p, mass, com, Ic, lo, up, eff, vel = [], [], [], [], [], [], [], [], []
    while cur in children:
        js = children[cur]
        assert len(js) == 1, "serial chain expected"
        j = js[0]
        child = j.find("child").get("link")
        if j.get("type") == "fixed":
            inert = links[child].find("inertial")
            if mass and inert is not None:
                raise NotImplementedError("fixed child with inertia after a moving joint")
            cur = child
            continue
        assert j.get("type") in ("revolute", "continuous")
        ax = [float(x) for x in j.find("axis").get("xyz").split()]
        assert ax == [0.0, 0.0, 1.0], "only local-z revolute joints are supported"
        o = j.find("origin")
        xyz = [float(x) for x in o.get("xyz").split()]
        rpy = [float(x) for x in o.get("rpy").split()]
        R.append(rpy_matrix(*rpy)); p.append(xyz)
        inert = links[child].find("inertial")
        io = inert.find("origin")
        assert [float(x) for x in io.get("rpy").split()] == [0.0, 0.0, 0.0]
        com.append([float(x) for x in io.get("xyz").split()])
        mass.append(float(inert.find("mass").get("value")))
        I = inert.find("inertia")
        ixx, ixy, ixz, iyy, iyz, izz = (float(I.get(k)) for k in ("ixx", "ixy", "ixz", "iyy", "iyz", "izz"))
        Ic.append([[ixx, ixy, ixz], [ixy, iyy, iyz], [ixz, iyz, izz]])
        lim = j.find("limit")
        lo.append(float(lim.get("lower"))); up.append(float(lim.get("upper")))
        eff.append(float(lim.get("effort"))); vel.append(float(lim.get("velocity")))
        cur = child
    return Model(R, p, mass, com, Ic, lo, up, eff, vel)
