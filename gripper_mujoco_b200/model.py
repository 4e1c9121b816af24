"""Gripper model description -> flat `GmjModel` blob.

The reference generates an MJCF file from xacro + `config/gripper.yaml` and lets MuJoCo compile
it. Neither xacro nor MuJoCo exist here, and the MJCF is git-ignored in the reference, so this
module re-derives the same kinematic tree, inertias, geoms and joint parameters directly from the
xacro recipes (closed-form) -- SURVEY.md rows M1-M10:

* base + motor joints            description/xacro/gripper.main.xacro:12-31,98-257
* task frame (height h, flip)    description/mujoco/xacro/gripper_task.urdf.xacro:35-43,140-152
* segmented fingers              description/xacro/gripper.segmented_fingers.xacro:46-128,193-273
* attachment point x_fix=0.129   description/xacro/gripper.xacro:9-11
* defaults                       description/config/gripper.yaml:1-28
* joint stiffness (runtime)      src/myfunctions.cpp:988-1048
* PD gains                       src/myfunctions.cpp:273-296
* keyframe                       description/mujoco/xml_script.py:97-99,216-223
* friction of finger geoms       description/mujoco/xml_script.py:122-127
* MuJoCo URDF import conventions: `fusestatic` merges the fixed hook link into the last segment,
  `balanceinertia=true` (description/mujoco/xacro/mujoco.xacro:6-10).
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass, field, asdict

import numpy as np

GMJ_MAXB = 88
GMJ_MAXG = 96

GEOM_ELLIPSOID, GEOM_BOX, GEOM_CYLINDER = 0, 1, 2
CLS_NONE, CLS_F1, CLS_F2, CLS_F3, CLS_PALM, CLS_GROUND, CLS_OBJECT = range(7)


class GmjModel(C.Structure):
    """ctypes mirror of `GmjModel` in include/gmj_types.h (field order must match)."""
    _fields_ = [
        ("N", C.c_int32), ("nb", C.c_int32), ("B", C.c_int32), ("ngeom", C.c_int32),
        ("parent", C.c_int32 * GMJ_MAXB), ("jtype", C.c_int32 * GMJ_MAXB),
        ("group", C.c_int32 * GMJ_MAXB),
        ("pos", (C.c_double * 3) * GMJ_MAXB), ("quat", (C.c_double * 4) * GMJ_MAXB),
        ("axis", (C.c_double * 3) * GMJ_MAXB),
        ("mass", C.c_double * GMJ_MAXB), ("ipos", (C.c_double * 3) * GMJ_MAXB),
        ("inertia", (C.c_double * 6) * GMJ_MAXB),
        ("stiffness", C.c_double * GMJ_MAXB), ("damping", C.c_double * GMJ_MAXB),
        ("range_lo", C.c_double * GMJ_MAXB), ("range_hi", C.c_double * GMJ_MAXB),
        ("qpos0", C.c_double * GMJ_MAXB),
        ("g_body", C.c_int32 * GMJ_MAXG), ("g_type", C.c_int32 * GMJ_MAXG),
        ("g_class", C.c_int32 * GMJ_MAXG),
        ("g_pos", (C.c_double * 3) * GMJ_MAXG), ("g_quat", (C.c_double * 4) * GMJ_MAXG),
        ("g_size", (C.c_double * 3) * GMJ_MAXG), ("g_friction", C.c_double * GMJ_MAXG),
        ("ground_friction", C.c_double),
        ("gravity", C.c_double), ("solref", C.c_double * 2), ("solimp", C.c_double * 5),
        ("finger_length", C.c_double), ("finger_width", C.c_double),
        ("finger_thickness", C.c_double), ("finger_E", C.c_double),
        ("hook_length", C.c_double), ("hook_angle_deg", C.c_double),
        ("fingertip_clearance", C.c_double), ("segment_length", C.c_double),
        ("EI", C.c_double),
        ("kp", C.c_double * 3), ("kd", C.c_double * 3), ("base_kp", C.c_double),
        ("base_kd", C.c_double),
        ("tip_body", C.c_int32 * 3), ("finger_body", C.c_int32 * 3),
        ("palm_body", C.c_int32), ("pad_", C.c_int32),
        ("tip_mat", (C.c_double * 9) * 3),
    ]


class GmjObject(C.Structure):
    """ctypes mirror of `GmjObject` in include/gmj_types.h."""
    _fields_ = [
        ("type", C.c_int32), ("pad_", C.c_int32),
        ("size", C.c_double * 3), ("round", C.c_double), ("quat", C.c_double * 4),
        ("mass", C.c_double), ("inertia", C.c_double * 3), ("friction", C.c_double),
        ("z_rest", C.c_double), ("bbox", C.c_double * 3),
    ]


@dataclass
class GripperParams:
    """description/config/gripper.yaml (values at the reference commit) + runtime finger settings."""
    num_segments: int = 8
    fixed_first_segment: bool = False
    fixed_hook_segment: bool = True
    xy_base_joint: bool = False
    z_base_rotation: bool = False
    angle_max_abs: float = 0.7
    finger_E: float = 193e9
    finger_length: float = 0.235
    finger_mass: float = 0.075
    finger_thickness: float = 0.9e-3
    finger_width: float = 0.028
    fingertip_clearance: float = 0.01
    hook_angle_degrees: float = 75.0
    hook_length: float = 0.035
    segment_inertia_scaling: float = 50.0
    xy_home: float = 0.13
    xy_max: float = 0.134
    xy_min: float = 0.049
    z_home: float = 0.00488
    z_max: float = 0.165
    z_min: float = 0.0


# ---------------------------------------------------------------- small rotation helpers
def rpy_to_mat(r, p, y):
    """URDF fixed-axis rpy: R = Rz(y) Ry(p) Rx(r)."""
    cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
    Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def mat_to_quat(R):
    """Rotation matrix -> unit quaternion (w, x, y, z), w >= 0 where possible."""
    R = np.asarray(R, dtype=np.float64)
    t = np.trace(R)
    if t > 0:
        s = math.sqrt(t + 1.0) * 2
        q = np.array([0.25 * s, (R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s])
    elif R[0, 0] > R[1, 1] and R[0, 0] > R[2, 2]:
        s = math.sqrt(1.0 + R[0, 0] - R[1, 1] - R[2, 2]) * 2
        q = np.array([(R[2, 1] - R[1, 2]) / s, 0.25 * s, (R[0, 1] + R[1, 0]) / s, (R[0, 2] + R[2, 0]) / s])
    elif R[1, 1] > R[2, 2]:
        s = math.sqrt(1.0 + R[1, 1] - R[0, 0] - R[2, 2]) * 2
        q = np.array([(R[0, 2] - R[2, 0]) / s, (R[0, 1] + R[1, 0]) / s, 0.25 * s, (R[1, 2] + R[2, 1]) / s])
    else:
        s = math.sqrt(1.0 + R[2, 2] - R[0, 0] - R[1, 1]) * 2
        q = np.array([(R[1, 0] - R[0, 1]) / s, (R[0, 2] + R[2, 0]) / s, (R[1, 2] + R[2, 1]) / s, 0.25 * s])
    return q / np.linalg.norm(q)


def quat_to_mat(q):
    w, x, y, z = q
    return np.array([
        [1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
        [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
        [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])


def _sym6(I):
    return [I[0, 0], I[1, 1], I[2, 2], I[0, 1], I[0, 2], I[1, 2]]


def _balance_inertia(I):
    """MuJoCo compiler `balanceinertia=true`: if the principal moments violate A+B>=C, replace
    all three with their mean (description/mujoco/xacro/mujoco.xacro:8)."""
    w, V = np.linalg.eigh(I)
    a, b, c = sorted(w)
    if a + b < c:
        m = (a + b + c) / 3.0
        return np.eye(3) * m
    return I


@dataclass
class Body:
    name: str
    parent: int
    jtype: int
    group: int
    pos: np.ndarray
    R: np.ndarray
    axis: np.ndarray
    mass: float
    ipos: np.ndarray
    inertia: np.ndarray          # 3x3 about COM in body frame
    stiffness: float = 0.0
    damping: float = 0.0
    lo: float = -1.0
    hi: float = 1.0
    qpos0: float = 0.0


@dataclass
class Geom:
    body: int
    gtype: int
    cls: int
    pos: np.ndarray
    R: np.ndarray
    size: np.ndarray
    friction: float


@dataclass
class Model:
    params: GripperParams
    bodies: list = field(default_factory=list)
    geoms: list = field(default_factory=list)
    extras: dict = field(default_factory=dict)

    @property
    def N(self):
        return self.params.num_segments

    @property
    def B(self):
        return len(self.bodies)

    @property
    def nv(self):
        return self.B + 6

    @property
    def nq(self):
        return self.B + 7

    def to_struct(self) -> GmjModel:
        m = GmjModel()
        p = self.params
        m.N, m.nb, m.B, m.ngeom = p.num_segments, 1, self.B, len(self.geoms)
        assert self.B <= GMJ_MAXB and len(self.geoms) <= GMJ_MAXG
        for i, b in enumerate(self.bodies):
            m.parent[i], m.jtype[i], m.group[i] = b.parent, b.jtype, b.group
            q = mat_to_quat(b.R)
            for k in range(3):
                m.pos[i][k] = b.pos[k]
                m.axis[i][k] = b.axis[k]
                m.ipos[i][k] = b.ipos[k]
            for k in range(4):
                m.quat[i][k] = q[k]
            m.mass[i] = b.mass
            for k, v in enumerate(_sym6(b.inertia)):
                m.inertia[i][k] = v
            m.stiffness[i], m.damping[i] = b.stiffness, b.damping
            m.range_lo[i], m.range_hi[i], m.qpos0[i] = b.lo, b.hi, b.qpos0
        for i, g in enumerate(self.geoms):
            m.g_body[i], m.g_type[i], m.g_class[i] = g.body, g.gtype, g.cls
            q = mat_to_quat(g.R)
            for k in range(3):
                m.g_pos[i][k] = g.pos[k]
                m.g_size[i][k] = g.size[k]
            for k in range(4):
                m.g_quat[i][k] = q[k]
            m.g_friction[i] = g.friction
        e = self.extras
        m.ground_friction = 1.0
        m.gravity = 9.81
        m.solref[0], m.solref[1] = 0.02, 1.0
        for k, v in enumerate((0.9, 0.95, 0.001, 0.5, 2.0)):
            m.solimp[k] = v
        m.finger_length, m.finger_width = p.finger_length, p.finger_width
        m.finger_thickness, m.finger_E = p.finger_thickness, p.finger_E
        m.hook_length, m.hook_angle_deg = p.hook_length, p.hook_angle_degrees
        m.fingertip_clearance = p.fingertip_clearance
        m.segment_length = e["segment_length"]
        m.EI = e["EI"]
        for k in range(3):
            m.kp[k], m.kd[k] = e["kp"][k], e["kd"][k]
            m.tip_body[k] = e["tip_body"][k]
            m.finger_body[k] = e["finger_body"][k]
            for j in range(9):
                m.tip_mat[k][j] = e["tip_mat"][k].reshape(-1)[j]
        m.base_kp, m.base_kd = e["base_kp"], e["base_kd"]
        m.palm_body = e["palm_body"]
        return m


def finger_joint_stiffness(p: GripperParams):
    """src/myfunctions.cpp:988-1048 -- note the reference computes `c` in single precision."""
    N = p.num_segments
    I = (p.finger_width * p.finger_thickness ** 3) / 12.0
    EI = p.finger_E * I
    out = []
    for n in range(1, N + 1):
        if n == 1:
            c = ((2 * EI) / p.finger_length) * ((N * N) / (N - (1.0 / 3.0)))
        else:
            c = (N * EI) / p.finger_length
        out.append(float(np.float32(c)))
    return out, EI


def build_model(p: GripperParams | None = None) -> Model:
    """Assemble the gripper task model for a z-only base (nb = 1)."""
    p = p or GripperParams()
    if p.xy_base_joint or p.z_base_rotation:
        raise NotImplementedError("only the world_to_base (z only) base is built in this round")
    if p.fixed_first_segment or not p.fixed_hook_segment:
        raise NotImplementedError("fixed_first_segment / free hook joint are not built in this round")
    N = p.num_segments
    if N < 2:
        raise NotImplementedError("num_segments < 2")
    mdl = Model(params=p)
    pi = math.pi
    hook_angle = p.hook_angle_degrees * (3.1415926535897 / 180.0)   # segmented_fingers.xacro:19
    hook_angle_task = p.hook_angle_degrees * (3.141592 / 180.0)      # gripper_task.urdf.xacro:16

    # ---- base (gripper_task.urdf.xacro:35-43,140-152 ; gripper.main.xacro:36-43,98-111)
    hook_extra = 0.0 if p.hook_angle_degrees > 90 else p.hook_length * math.cos(hook_angle_task)
    h = 0.1407 + p.finger_length + p.fingertip_clearance + hook_extra
    Rb_in = rpy_to_mat(pi, 0, 0)
    I_base = np.array([[9.95e-3, 2.73e-6, -14.32e-6],
                       [2.73e-6, 10.08e-3, 121.38e-6],
                       [-14.32e-6, 121.38e-6, 5.13e-3]])
    I_base = _balance_inertia(Rb_in @ I_base @ Rb_in.T)
    mdl.bodies.append(Body("gripper_base_link", -1, 0, 0, np.array([0, 0, h]), rpy_to_mat(pi, 0, 0),
                           np.array([0, 0, 1.0]), 1.5, np.array([0, 0, 0.025]), I_base,
                           0.0, 1.0, -1.0, 1.0, 0.0))

    # ---- fingers
    I_nut = _balance_inertia(np.array([[55.59e-6, -47.23e-6, 1.16e-6],
                                       [-47.23e-6, 146.36e-6, 1.00e-6],
                                       [1.16e-6, 1.00e-6, 154.18e-6]]))
    L = p.finger_length
    seg_len = L / float(N)
    w, t = p.finger_width, p.finger_thickness
    seg_mass = (p.finger_mass * seg_len) / (L + p.hook_length)
    sc = p.segment_inertia_scaling
    seg_I = np.diag([
        (1.0 / 12.0) * seg_mass * (t ** 2 + w ** 2),
        (1.0 / 12.0) * seg_mass * (seg_len ** 2 + t ** 2) + seg_mass * (seg_len / 2.0) ** 2,
        (1.0 / 12.0) * seg_mass * (seg_len ** 2 + w ** 2) + seg_mass * (seg_len / 2.0) ** 2]) * sc
    hook_mass = (p.finger_mass * p.hook_length) / (L + p.hook_length)
    hl = p.hook_length
    hook_I = np.diag([
        (1.0 / 12.0) * hook_mass * (t ** 2 + w ** 2),
        (1.0 / 12.0) * hook_mass * (hl ** 2 + t ** 2) + hook_mass * (hl / 2.0) ** 2,
        (1.0 / 12.0) * hook_mass * (hl ** 2 + w ** 2) + hook_mass * (hl / 2.0) ** 2]) * sc
    stiff, EI = finger_joint_stiffness(p)
    seg_fric = 0.5   # xml_script.py:122-127 (0.5 x MuJoCo default sliding friction)

    tip_body, finger_body, tip_mat = [], [], []
    for f in range(3):
        number = f + 1
        Ri = rpy_to_mat(3 * pi / 2, 3 * pi / 2, (pi / 2) - (number - 1) * (2 * pi / 3))
        pris = len(mdl.bodies)
        mdl.bodies.append(Body(f"finger_{number}_intermediate", 0, 0, number, np.zeros(3), Ri,
                               np.array([0, 1.0, 0]), 0.1, np.zeros(3), I_nut,
                               0.0, 1.0, p.xy_min, p.xy_max, p.xy_home))
        rev = len(mdl.bodies)
        mdl.bodies.append(Body(f"finger_{number}", pris, 1, number, np.zeros(3), np.eye(3),
                               np.array([0, 0, 1.0]), 0.1, np.zeros(3), I_nut,
                               0.0, 1.0, -p.angle_max_abs, p.angle_max_abs, 0.0))
        finger_body.append(rev)
        par = rev
        for k in range(1, N + 1):
            if k == 1:
                bpos, bR = np.array([0.129, 0, 0]), rpy_to_mat(pi / 2, 0, 0)
            else:
                bpos, bR = np.array([seg_len, 0, 0]), np.eye(3)
            mass, ipos, inertia = seg_mass, np.zeros(3), seg_I.copy()
            if k == N:
                # fixed hook link fused into the last segment (MuJoCo URDF import `fusestatic`)
                Rh = rpy_to_mat(0, -hook_angle, 0)
                ph = np.array([seg_len, 0, 0])
                mtot = seg_mass + hook_mass
                com = (seg_mass * np.zeros(3) + hook_mass * ph) / mtot
                Ih = Rh @ hook_I @ Rh.T

                def shift(I, m, d):
                    return I + m * (np.dot(d, d) * np.eye(3) - np.outer(d, d))
                inertia = shift(seg_I, seg_mass, -com) + shift(Ih, hook_mass, ph - com)
                mass, ipos = mtot, com
            inertia = _balance_inertia(inertia)
            idx = len(mdl.bodies)
            mdl.bodies.append(Body(f"finger_{number}_segment_link_{k}", par, 1, number, bpos, bR,
                                   np.array([0, 1.0, 0]), mass, ipos, inertia,
                                   stiff[k - 1], 0.01, -1.0, 1.0, 0.0))
            # segment collision box (segmented_fingers.xacro:117-123)
            mdl.geoms.append(Geom(idx, GEOM_BOX, number, np.array([seg_len / 2, 0, 0]), np.eye(3),
                                  np.array([seg_len / 2, w / 2, t / 2]), seg_fric))
            if k == N:
                # hook plate: mesh "fingertip 28mm collision.STL" (plate 28 x 35 x 1 mm) mounted
                # rpy=(0,0,-pi/2), scaled (w/28e-3, hook/35e-3, 1)  (segmented_fingers.xacro:253-259)
                Rh = rpy_to_mat(0, -hook_angle, 0)
                ph = np.array([seg_len, 0, 0])
                half = np.array([0.0175 * (hl / 35e-3), 0.014 * (w / 28e-3), 0.0005])
                centre_in_hook = np.array([half[0], 0.0, 0.0005])
                mdl.geoms.append(Geom(idx, GEOM_BOX, number, ph + Rh @ centre_in_hook, Rh, half, seg_fric))
                tip_body.append(idx)
            par = idx
    # ---- palm (gripper.main.xacro:229-257); collision hull of "palm sensor visual.STL" = cylinder
    palm = len(mdl.bodies)
    mdl.bodies.append(Body("palm", 0, 0, 4, np.zeros(3), rpy_to_mat(0, -pi / 2, 0),
                           np.array([1.0, 0, 0]), 0.05, np.zeros(3), np.eye(3) * 10.0e-6,
                           0.0, 1.0, p.z_min, p.z_max, p.z_home))
    x_lo, x_hi, r_palm = -0.08575, 0.1905, 0.0225
    mdl.geoms.append(Geom(palm, GEOM_CYLINDER, CLS_PALM, np.array([(x_lo + x_hi) / 2, 0, 0]),
                          rpy_to_mat(0, pi / 2, 0), np.array([r_palm, r_palm, (x_hi - x_lo) / 2]), seg_fric))

    # ---- control gains (src/myfunctions.cpp:273-296)
    kp = [EI * 541.3 + 49.65, EI * 44.92 - 0.846, 1000.0]
    kd = [1.0, 1.0, 1.0]
    mdl.extras = dict(segment_length=L / float(N), EI=EI, kp=kp, kd=kd, base_kp=2000.0, base_kd=100.0,
                      tip_body=tip_body, finger_body=finger_body, palm_body=palm, h=h,
                      seg_mass=seg_mass, hook_mass=hook_mass)
    # initial world orientation of the last segment of each finger (used for the calibration load)
    X = forward_kinematics(mdl, np.array([b.qpos0 for b in mdl.bodies]))
    mdl.extras["tip_mat"] = [X[1][tb] for tb in tip_body]
    return mdl


def forward_kinematics(mdl: Model, q):
    """Plain numpy FK used by host-side checks (not the hot path). Returns (xpos, xmat)."""
    nb = len(mdl.bodies)
    xpos = [None] * nb
    xmat = [None] * nb
    for i, b in enumerate(mdl.bodies):
        if b.parent < 0:
            Pp, Rp = np.zeros(3), np.eye(3)
        else:
            Pp, Rp = xpos[b.parent], xmat[b.parent]
        R = Rp @ b.R
        P = Pp + Rp @ b.pos
        if b.jtype == 0:
            P = P + R @ b.axis * q[i]
        else:
            a = b.axis / np.linalg.norm(b.axis)
            c, s = math.cos(q[i]), math.sin(q[i])
            K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
            R = R @ (np.eye(3) + s * K + (1 - c) * (K @ K))
        xpos[i], xmat[i] = P, R
    return xpos, xmat
