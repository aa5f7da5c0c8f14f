"""Flat scene arrays (``gz_scene`` / ``gz_params`` of include/gzcuda_scene.h) built from Python.

Host-side mirror of what gazebo::physics::ODEPhysics builds while loading SDF
(reference: gazebo/physics/ode/ODEPhysics.cc:175-243, ODELink.cc:63-134,
ODEModel.cc:23-27, ODEJoint.cc:83-137): models own a collision space, links become
bodies at their centre of mass, collisions become geoms with an offset, joints are
given in world coordinates and converted to the body-frame anchors/axes/qrel that
dJointSetHingeAnchor / dJointSetHingeAxis / dJointSetFixed store
(deps/opende/src/joints/joint.cpp:270-330, hinge.cpp:186-247, fixed.cpp:113-142).
"""
import ctypes as C
import math
import sys

import numpy as np

GEOM_SPHERE, GEOM_BOX, GEOM_CAPSULE, GEOM_CYLINDER, GEOM_PLANE = range(5)
JOINT_HINGE, JOINT_FIXED, JOINT_BALL, JOINT_SLIDER, JOINT_UNIVERSAL, JOINT_HINGE2, JOINT_GEARBOX, JOINT_SCREW = 1, 2, 3, 4, 5, 6, 7, 8
BODY_NO_GRAVITY, BODY_AUTO_DISABLE, BODY_DISABLED, BODY_NO_GYRO = 1, 2, 4, 8
JP_LOSTOP, JP_HISTOP, JP_VEL, JP_FMAX, JP_FUDGE, JP_BOUNCE, JP_NORMAL_CFM, \
    JP_STOP_ERP, JP_STOP_CFM, JP_ERP, JP_CFM = range(11)
JP_VEL2, JP_FMAX2, JP_SUSP_ERP, JP_SUSP_CFM, JP_V1, JP_V2, JP_C0, JP_S0 = 12, 13, 14, 15, 16, 19, 22, 23
JP_LOSTOP2, JP_HISTOP2, JP_FUDGE2, JP_BOUNCE2, JP_NORMAL_CFM2, JP_STOP_ERP2, JP_STOP_CFM2, JP_QREL2 = 24, 25, 26, 27, 28, 29, 30, 31
JP_RATIO, JP_REFBODY1, JP_REFBODY2 = 35, 14, 15
JP_THREAD_PITCH, JP_OFFSET = 35, 16
JP_COUNT = 36
SF_MU, SF_MU2, SF_KP, SF_KD, SF_SLIP1, SF_SLIP2, SF_BOUNCE, SF_BOUNCE_VEL = range(8)
SF_PATCH_RADIUS, SF_ELASTIC_MODULUS, SF_POISSON, SF_SURFACE_RADIUS, SF_USE_PATCH_RADIUS, SF_MU3, SF_SLIP3 = 8, 9, 10, 11, 12, 13, 14
SF_COUNT = 15
GZ_ALL_COLLIDE = 0x0FFFFFFF
GZ_FIXED_COLLIDE = 0x00000001


class GzParams(C.Structure):
    _fields_ = [
        ("gravity", C.c_double * 3), ("step_size", C.c_double),
        ("erp", C.c_double), ("cfm", C.c_double), ("sor_w", C.c_double),
        ("iters", C.c_int), ("friction_iters", C.c_int),
        ("contact_max_correcting_vel", C.c_double), ("contact_surface_layer", C.c_double),
        ("max_contacts", C.c_int),
        ("warm_start", C.c_double), ("smooth_contacts", C.c_double),
        ("contact_sor_scale", C.c_double),
        ("adis_idle_steps", C.c_int), ("adis_idle_time", C.c_double),
        ("adis_linear_threshold_sq", C.c_double), ("adis_angular_threshold_sq", C.c_double),
        ("damping_linear_threshold_sq", C.c_double), ("damping_angular_threshold_sq", C.c_double),
        ("friction_model", C.c_int),
        ("dynamic_inertia", C.c_int),
    ]


def default_params(**kw):
    """SDF / ODEPhysics::Load defaults (thirdparty/sdformat/sdf/1.6/physics.sdf:17-233,
    ODEPhysics.cc:197-237, ode.cpp:1650-1682)."""
    p = GzParams()
    p.gravity[:] = (0.0, 0.0, -9.8)
    p.step_size = 0.001
    p.erp, p.cfm, p.sor_w = 0.2, 0.0, 1.3
    p.iters, p.friction_iters = 50, 10
    p.contact_max_correcting_vel, p.contact_surface_layer = 100.0, 0.001
    p.max_contacts = 20
    p.warm_start, p.smooth_contacts, p.contact_sor_scale = 0.5, 0.01, 0.25
    p.adis_idle_steps, p.adis_idle_time = 5, 1.0
    p.adis_linear_threshold_sq = 0.1 * 0.1
    p.adis_angular_threshold_sq = 0.1 * 0.1
    p.damping_linear_threshold_sq = 0.01 * 0.01
    p.damping_angular_threshold_sq = 0.01 * 0.01
    for k, v in kw.items():
        if k == "gravity":
            p.gravity[:] = v
        else:
            setattr(p, k, v)
    return p


_DP = C.POINTER(C.c_double)
_IP = C.POINTER(C.c_int)
_UP = C.POINTER(C.c_uint)


class GzScene(C.Structure):
    _fields_ = [
        ("n_worlds", C.c_int),
        ("body_off", _IP), ("geom_off", _IP), ("joint_off", _IP),
        ("body_pos", _DP), ("body_quat", _DP), ("body_lvel", _DP), ("body_avel", _DP),
        ("body_mass", _DP), ("body_inertia", _DP), ("body_flags", _IP),
        ("body_damping", _DP), ("body_contactp", _DP),
        ("geom_type", _IP), ("geom_body", _IP), ("geom_space", _IP),
        ("geom_dims", _DP), ("geom_pos", _DP), ("geom_quat", _DP),
        ("geom_category_bits", _UP), ("geom_collide_bits", _UP), ("geom_collide_bitmask", _UP),
        ("geom_max_contacts", _IP), ("geom_surf", _DP), ("geom_fdir1", _DP),
        ("joint_type", _IP), ("joint_body1", _IP), ("joint_body2", _IP),
        ("joint_anchor1", _DP), ("joint_anchor2", _DP), ("joint_axis1", _DP),
        ("joint_axis2", _DP), ("joint_qrel", _DP), ("joint_param", _DP),
        ("geom_rot", _DP),      # optional, NULL here: the builders describe rotations by quaternions
    ]


# ---------------------------------------------------------------- quaternion helpers
def q_mul(a, b):
    return np.array([
        a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3],
        a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
        a[0] * b[2] + a[2] * b[0] + a[3] * b[1] - a[1] * b[3],
        a[0] * b[3] + a[3] * b[0] + a[1] * b[2] - a[2] * b[1]])


def q_conj(a):
    return np.array([a[0], -a[1], -a[2], -a[3]])


def q_to_R(q):
    """dQtoR (deps/opende/src/rotation.cpp)."""
    qq1, qq2, qq3 = 2 * q[1] * q[1], 2 * q[2] * q[2], 2 * q[3] * q[3]
    return np.array([
        [1 - qq2 - qq3, 2 * (q[1] * q[2] - q[0] * q[3]), 2 * (q[1] * q[3] + q[0] * q[2])],
        [2 * (q[1] * q[2] + q[0] * q[3]), 1 - qq1 - qq3, 2 * (q[2] * q[3] - q[0] * q[1])],
        [2 * (q[1] * q[3] - q[0] * q[2]), 2 * (q[2] * q[3] + q[0] * q[1]), 1 - qq1 - qq2]])


def q_from_axis_angle(axis, angle):
    a = np.asarray(axis, float)
    n = np.linalg.norm(a)
    if n == 0:
        return np.array([1.0, 0, 0, 0])
    s = math.sin(angle * 0.5) / n
    return np.array([math.cos(angle * 0.5), a[0] * s, a[1] * s, a[2] * s])


def q_from_rpy(r, p, y):
    return q_mul(q_from_axis_angle([0, 0, 1], y),
                 q_mul(q_from_axis_angle([0, 1, 0], p), q_from_axis_angle([1, 0, 0], r)))


def q_random(rng):
    q = rng.normal(size=4)
    return q / np.linalg.norm(q)


def box_inertia(mass, lx, ly, lz):
    return [mass * (ly * ly + lz * lz) / 12.0, mass * (lx * lx + lz * lz) / 12.0,
            mass * (lx * lx + ly * ly) / 12.0, 0.0, 0.0, 0.0]


def sphere_inertia(mass, r):
    i = 0.4 * mass * r * r
    return [i, i, i, 0.0, 0.0, 0.0]


def cylinder_inertia(mass, r, l):
    """axis along z (dMassSetCylinderTotal direction 3)."""
    it = mass * (3 * r * r + l * l) / 12.0
    return [it, it, 0.5 * mass * r * r, 0.0, 0.0, 0.0]


DEFAULT_SURFACE = dict(mu=1.0, mu2=1.0, kp=1e12, kd=1.0, slip1=0.0, slip2=0.0,
                       bounce=0.0, bounce_vel=100000.0, patch_radius=0.0, elastic_modulus=0.0,
                       poissons_ratio=0.3, surface_radius=sys.float_info.max, use_patch_radius=1.0, mu3=0.0, slip3=0.0)


class SceneBuilder:
    """Accumulates worlds; ``finish()`` returns a :class:`Scene` of contiguous numpy arrays."""

    def __init__(self):
        self.body_off, self.geom_off, self.joint_off = [0], [0], [0]
        self.b = {k: [] for k in ("pos", "quat", "lvel", "avel", "mass", "inertia", "flags",
                                  "damping", "contactp")}
        self.g = {k: [] for k in ("type", "body", "space", "dims", "pos", "quat", "cat", "col",
                                  "bitmask", "maxc", "surf", "fdir1")}
        self.j = {k: [] for k in ("type", "b1", "b2", "anchor1", "anchor2", "axis1", "axis2",
                                  "qrel", "param")}
        self._nb = self._ng = self._nj = 0
        self._w_b0 = self._w_s = 0
        self._open = False

    # -- world / model scoping
    def begin_world(self):
        assert not self._open
        self._open = True
        self._w_b0, self._w_g0, self._w_j0 = self._nb, self._ng, self._nj
        self._w_s = 0

    def end_world(self):
        self._open = False
        self.body_off.append(self._nb)
        self.geom_off.append(self._ng)
        self.joint_off.append(self._nj)

    def new_space(self):
        """One dSimpleSpace per model (ODEModel.cc:26) or per self-colliding link (ODELink.cc:231-240)."""
        s = self._w_s
        self._w_s += 1
        return s

    # -- bodies
    def add_body(self, pos, quat=(1, 0, 0, 0), mass=1.0, inertia=(1, 1, 1, 0, 0, 0),
                 lvel=(0, 0, 0), avel=(0, 0, 0), flags=0, lin_damping=0.0, ang_damping=0.0,
                 max_vel=0.01, min_depth=0.0):
        self.b["pos"].append(list(map(float, pos)))
        self.b["quat"].append(list(map(float, quat)))
        self.b["lvel"].append(list(map(float, lvel)))
        self.b["avel"].append(list(map(float, avel)))
        self.b["mass"].append(float(mass))
        self.b["inertia"].append(list(map(float, inertia)))
        self.b["flags"].append(int(flags))
        self.b["damping"].append([float(lin_damping), float(ang_damping)])
        self.b["contactp"].append([float(max_vel), float(min_depth)])
        self._nb += 1
        return self._nb - 1 - self._w_b0

    # -- geoms
    def add_geom(self, gtype, dims, space, body=-1, pos=(0, 0, 0), quat=(1, 0, 0, 0),
                 category_bits=None, collide_bits=None, collide_bitmask=0xFFFF,
                 max_contacts=20, fdir1=(0, 0, 0), **surface):
        d = list(map(float, dims)) + [0.0] * (4 - len(dims))
        s = dict(DEFAULT_SURFACE)
        s.update(surface)
        static = body < 0
        if category_bits is None:  # Collision::Load / ODECollision::Load (ODECollision.cc:70-74)
            category_bits = GZ_FIXED_COLLIDE if static else GZ_ALL_COLLIDE
        if collide_bits is None:
            collide_bits = (~GZ_FIXED_COLLIDE) & 0xFFFFFFFF if static else GZ_ALL_COLLIDE
        self.g["type"].append(int(gtype))
        self.g["body"].append(int(body))
        self.g["space"].append(int(space))
        self.g["dims"].append(d)
        self.g["pos"].append(list(map(float, pos)))
        self.g["quat"].append(list(map(float, quat)))
        self.g["cat"].append(category_bits & 0xFFFFFFFF)
        self.g["col"].append(collide_bits & 0xFFFFFFFF)
        self.g["bitmask"].append(collide_bitmask & 0xFFFFFFFF)
        self.g["maxc"].append(int(max_contacts))
        self.g["surf"].append([s["mu"], s["mu2"], s["kp"], s["kd"], s["slip1"], s["slip2"],
                               s["bounce"], s["bounce_vel"], s["patch_radius"], s["elastic_modulus"],
                               s["poissons_ratio"], s["surface_radius"], s["use_patch_radius"], s["mu3"], s["slip3"]])
        self.g["fdir1"].append(list(map(float, fdir1)))
        self._ng += 1
        return self._ng - 1 - self._w_g0

    def add_ground_plane(self, mu=100.0, mu2=50.0):
        """models/ground_plane/model.sdf: static plane z=0, mu 100 / mu2 50."""
        return self.add_geom(GEOM_PLANE, (0, 0, 1, 0), self.new_space(), mu=mu, mu2=mu2)

    # -- joints
    def _body_pose(self, lb):
        i = self._w_b0 + lb
        return np.array(self.b["pos"][i]), np.array(self.b["quat"][i])

    @staticmethod
    def joint_params(lostop=-math.inf, histop=math.inf, vel=0.0, fmax=0.0, fudge=1.0,
                     bounce=0.0, normal_cfm=0.0, stop_erp=0.2, stop_cfm=0.0, erp=-1.0, cfm=-1.0):
        return [lostop, histop, vel, fmax, fudge, bounce, normal_cfm, stop_erp, stop_cfm, erp, cfm, 0.0] + [0.0] * (JP_COUNT - 12)

    def add_hinge(self, b1, b2, anchor, axis, **params):
        """Anchor and axis in world coordinates at the current body poses."""
        p1, q1 = self._body_pose(b1)
        R1 = q_to_R(q1)
        anchor = np.asarray(anchor, float)
        axis = np.asarray(axis, float)
        axis = axis / np.linalg.norm(axis)
        a1 = R1.T @ (anchor - p1)
        ax1 = R1.T @ axis
        if b2 >= 0:
            p2, q2 = self._body_pose(b2)
            R2 = q_to_R(q2)
            a2 = R2.T @ (anchor - p2)
            ax2 = R2.T @ axis
            qrel = q_mul(q_conj(q1), q2)
        else:
            a2, ax2, qrel = anchor, axis, q_conj(q1)
        self._push_joint(JOINT_HINGE, b1, b2, a1, a2, ax1, ax2, qrel, self.joint_params(**params))
        return self._nj - 1 - self._w_j0

    def add_hinge2(self, b1, b2, anchor, axis1, axis2, vel2=0.0, fmax2=0.0, susp_erp=0.2, susp_cfm=0.0, **params):
        """dxJointHinge2 (SDF revolute2: steering / suspension axis 1 on body 1, wheel axis 2 on body 2).  Anchor and axes
        in world coordinates at the current body poses.  The members ODE derives in dJointSetHinge2Anchor / Axis1 / Axis2
        (v1, v2 to measure the angle of axis 1, c0 / s0 = cos / sin of the angle between the axes) are computed here
        and travel in joint_param: both sides of a parity test read the same numbers."""
        assert b2 >= 0, "hinge2 needs two bodies"
        p1, q1 = self._body_pose(b1)
        p2, q2 = self._body_pose(b2)
        R1, R2 = q_to_R(q1), q_to_R(q2)
        anchor = np.asarray(anchor, float)
        ax1w = np.asarray(axis1, float); ax1w = ax1w / np.linalg.norm(ax1w)
        ax2w = np.asarray(axis2, float); ax2w = ax2w / np.linalg.norm(ax2w)
        a1, a2 = R1.T @ (anchor - p1), R2.T @ (anchor - p2)
        ax1, ax2 = R1.T @ ax1w, R2.T @ ax2w
        g1, g2 = R1 @ ax1, R2 @ ax2                       # the axes back in world coordinates, as getAxisInfo sees them
        cr = np.cross(g1, g2)
        s0, c0 = math.sqrt(cr[0] * cr[0] + cr[1] * cr[1] + cr[2] * cr[2]), float(g1 @ g2)
        m2 = g2 - float(g1 @ g2) * g1                      # axis 2 made perpendicular to axis 1
        m2 = m2 / np.linalg.norm(m2)
        v1, v2 = R1.T @ m2, R1.T @ np.cross(g1, m2)
        P = self.joint_params(**params)
        P[JP_VEL2], P[JP_FMAX2], P[JP_SUSP_ERP], P[JP_SUSP_CFM] = vel2, fmax2, susp_erp, susp_cfm
        P[JP_V1:JP_V1 + 3] = list(v1); P[JP_V2:JP_V2 + 3] = list(v2); P[JP_C0], P[JP_S0] = c0, s0
        self._push_joint(JOINT_HINGE2, b1, b2, a1, a2, ax1, ax2, (1, 0, 0, 0), P)
        return self._nj - 1 - self._w_j0

    def add_fixed(self, b1, b2, **params):
        p1, q1 = self._body_pose(b1)
        R1 = q_to_R(q1)
        if b2 >= 0:
            p2, q2 = self._body_pose(b2)
            ofs = R1.T @ (p1 - p2)
            qrel = q_mul(q_conj(q1), q2)
        else:
            ofs, qrel = p1, q_conj(q1)
        z = np.zeros(3)
        self._push_joint(JOINT_FIXED, b1, b2, ofs, z, z, z, qrel, self.joint_params(**params))
        return self._nj - 1 - self._w_j0

    def add_ball(self, b1, b2, anchor, **params):
        """Ball-and-socket (dxJointBall); anchor in world coordinates at the current body poses."""
        p1, q1 = self._body_pose(b1)
        anchor = np.asarray(anchor, float)
        a1 = q_to_R(q1).T @ (anchor - p1)
        if b2 >= 0:
            p2, q2 = self._body_pose(b2)
            a2 = q_to_R(q2).T @ (anchor - p2)
        else:
            a2 = anchor
        z = np.zeros(3)
        self._push_joint(JOINT_BALL, b1, b2, a1, a2, z, z, (1, 0, 0, 0), self.joint_params(**params))
        return self._nj - 1 - self._w_j0

    def add_universal(self, b1, b2, anchor, axis1, axis2, lostop2=-math.inf, histop2=math.inf, vel2=0.0, fmax2=0.0,
                      fudge2=1.0, bounce2=0.0, normal_cfm2=0.0, stop_erp2=0.2, stop_cfm2=0.0, **params):
        """Universal joint (dxJointUniversal): axis1 is fixed in body 1, axis2 in body 2 (or the world); both in world
        coordinates at the current poses.  params = the limit motor of axis 1, the *2 arguments that of axis 2.  qrel1 /
        qrel2 (the joint's cross against the two bodies at attach time, computeInitialRelativeRotations,
        universal.cpp:716-758) are computed here and travel in the scene: both sides of a parity test read the same."""
        p1, q1 = self._body_pose(b1)
        R1 = q_to_R(q1)
        anchor = np.asarray(anchor, float)
        axis1 = np.asarray(axis1, float) / np.linalg.norm(axis1)
        axis2 = np.asarray(axis2, float) / np.linalg.norm(axis2)
        a1 = R1.T @ (anchor - p1)
        ax1 = R1.T @ axis1
        if b2 >= 0:
            p2, q2 = self._body_pose(b2)
            R2 = q_to_R(q2)
            a2 = R2.T @ (anchor - p2)
            ax2 = R2.T @ axis2
        else:
            q2 = np.array([1.0, 0, 0, 0])
            a2, ax2 = anchor, axis2

        def cross_quat(a, b):   # dRFrom2Axes + dRtoQ: the frame whose x is a and whose y is b made perpendicular to a
            a = a / np.linalg.norm(a)
            b = b - float(a @ b) * a
            b = b / np.linalg.norm(b)
            R = np.stack([a, b, np.cross(a, b)], axis=1)
            tr = R[0, 0] + R[1, 1] + R[2, 2]
            if tr >= 0:
                sq = math.sqrt(tr + 1); w = 0.5 * sq; sq = 0.5 / sq
                return np.array([w, (R[2, 1] - R[1, 2]) * sq, (R[0, 2] - R[2, 0]) * sq, (R[1, 0] - R[0, 1]) * sq])
            k = int(np.argmax([R[0, 0], R[1, 1], R[2, 2]]))
            i1, i2 = (k + 1) % 3, (k + 2) % 3
            sq = math.sqrt((R[k, k] - (R[i1, i1] + R[i2, i2])) + 1)
            q = np.zeros(4); q[1 + k] = 0.5 * sq; sq = 0.5 / sq
            q[1 + i1] = (R[k, i1] + R[i1, k]) * sq; q[1 + i2] = (R[i2, k] + R[k, i2]) * sq; q[0] = (R[i2, i1] - R[i1, i2]) * sq
            return q

        g1 = R1 @ ax1
        g2 = (q_to_R(q2) @ ax2) if b2 >= 0 else ax2
        qrel1 = q_mul(q_conj(q1), cross_quat(g1, g2))
        qrel2 = q_mul(q_conj(q2), cross_quat(g2, g1))
        P = self.joint_params(**params)
        P[JP_VEL2], P[JP_FMAX2] = vel2, fmax2
        P[JP_LOSTOP2], P[JP_HISTOP2], P[JP_FUDGE2], P[JP_BOUNCE2] = lostop2, histop2, fudge2, bounce2
        P[JP_NORMAL_CFM2], P[JP_STOP_ERP2], P[JP_STOP_CFM2] = normal_cfm2, stop_erp2, stop_cfm2
        P[JP_QREL2:JP_QREL2 + 4] = list(qrel2)
        self._push_joint(JOINT_UNIVERSAL, b1, b2, a1, a2, ax1, ax2, tuple(qrel1), P)
        return self._nj - 1 - self._w_j0

    def add_gearbox(self, b1, b2, axis1, axis2, ratio, ref1=-1, ref2=-1):
        """dxJointGearbox (SDF gearbox): angle of b1 about axis1 (against reference body ref1) + ratio x angle of b2 about
        axis2 (against ref2) is held at its initial value.  Axes in world coordinates at the current poses; ref = a body
        of the world or -1 (the world).  qrel1 / qrel2 as dJointSetGearboxReferenceBody1 / 2 compute them."""
        assert b2 >= 0, "a gearbox joint needs two bodies"
        p1, q1 = self._body_pose(b1)
        p2, q2 = self._body_pose(b2)
        ax1 = q_to_R(q1).T @ (np.asarray(axis1, float) / np.linalg.norm(axis1))
        ax2 = q_to_R(q2).T @ (np.asarray(axis2, float) / np.linalg.norm(axis2))
        qrel1 = q_mul(q_conj(q1), self._body_pose(ref1)[1]) if ref1 >= 0 else q_conj(q1)
        qrel2 = q_mul(q_conj(q2), self._body_pose(ref2)[1]) if ref2 >= 0 else q_conj(q2)
        P = self.joint_params()
        P[JP_RATIO], P[JP_REFBODY1], P[JP_REFBODY2] = float(ratio), float(ref1), float(ref2)
        P[JP_QREL2:JP_QREL2 + 4] = list(qrel2)
        z = np.zeros(3)
        self._push_joint(JOINT_GEARBOX, b1, b2, z, z, ax1, ax2, tuple(qrel1), P)
        return self._nj - 1 - self._w_j0

    def add_screw(self, b1, b2, anchor, axis, thread_pitch, **params):
        """dxJointScrew (SDF screw): rotation about the axis and translation along it coupled by thread_pitch (rad / m).
        Anchor and axis in world coordinates at the current poses; params = the stops of the rotation (a motor is refused:
        the reference writes its row behind the joint's rows)."""
        p1, q1 = self._body_pose(b1)
        R1 = q_to_R(q1)
        anchor = np.asarray(anchor, float)
        axis = np.asarray(axis, float) / np.linalg.norm(axis)
        a1, ax1 = R1.T @ (anchor - p1), R1.T @ axis
        if b2 >= 0:
            p2, q2 = self._body_pose(b2)
            R2 = q_to_R(q2)
            a2, ax2 = R2.T @ (anchor - p2), R2.T @ axis
            qrel = q_mul(q_conj(q1), q2)
            ofs = R2.T @ (p1 - p2)                 # computeOffset (screw.cpp:795-820)
        else:
            a2, ax2, qrel, ofs = anchor, axis, q_conj(q1), p1
        P = self.joint_params(**params)
        P[JP_THREAD_PITCH] = float(thread_pitch)
        P[JP_OFFSET:JP_OFFSET + 3] = list(ofs)
        self._push_joint(JOINT_SCREW, b1, b2, a1, a2, ax1, ax2, qrel, P)
        return self._nj - 1 - self._w_j0

    def add_slider(self, b1, b2, axis, **params):
        """Slider / prismatic joint (dxJointSlider): axis in world coordinates at the current poses; anchor1
        holds dxJointSlider::offset (slider.cpp computeOffset), qrel the initial relative rotation."""
        p1, q1 = self._body_pose(b1)
        R1 = q_to_R(q1)
        axis = np.asarray(axis, float)
        axis = axis / np.linalg.norm(axis)
        ax1 = R1.T @ axis
        if b2 >= 0:
            p2, q2 = self._body_pose(b2)
            ofs = q_to_R(q2).T @ (p1 - p2)
            qrel = q_mul(q_conj(q1), q2)
        else:
            ofs, qrel = p1, q_conj(q1)
        z = np.zeros(3)
        self._push_joint(JOINT_SLIDER, b1, b2, ofs, z, ax1, z, qrel, self.joint_params(**params))
        return self._nj - 1 - self._w_j0

    def _push_joint(self, t, b1, b2, a1, a2, ax1, ax2, qrel, param):
        self.j["type"].append(t)
        self.j["b1"].append(int(b1))
        self.j["b2"].append(int(b2))
        self.j["anchor1"].append(list(a1))
        self.j["anchor2"].append(list(a2))
        self.j["axis1"].append(list(ax1))
        self.j["axis2"].append(list(ax2))
        self.j["qrel"].append(list(qrel))
        self.j["param"].append(list(param))
        self._nj += 1

    def finish(self):
        assert not self._open
        return Scene(self)


def _arr(x, dtype, shape):
    a = np.ascontiguousarray(np.array(x, dtype=dtype).reshape(shape))
    return a


class Scene:
    """Contiguous numpy arrays + the ctypes ``gz_scene`` view (``.c``)."""

    def __init__(self, sb):
        nb, ng, nj = sb._nb, sb._ng, sb._nj
        self.n_worlds = len(sb.body_off) - 1
        self.nb, self.ng, self.nj = nb, ng, nj
        A = {}
        A["body_off"] = _arr(sb.body_off, np.int32, (-1,))
        A["geom_off"] = _arr(sb.geom_off, np.int32, (-1,))
        A["joint_off"] = _arr(sb.joint_off, np.int32, (-1,))
        A["body_pos"] = _arr(sb.b["pos"], np.float64, (nb, 3))
        A["body_quat"] = _arr(sb.b["quat"], np.float64, (nb, 4))
        A["body_lvel"] = _arr(sb.b["lvel"], np.float64, (nb, 3))
        A["body_avel"] = _arr(sb.b["avel"], np.float64, (nb, 3))
        A["body_mass"] = _arr(sb.b["mass"], np.float64, (nb,))
        A["body_inertia"] = _arr(sb.b["inertia"], np.float64, (nb, 6))
        A["body_flags"] = _arr(sb.b["flags"], np.int32, (nb,))
        A["body_damping"] = _arr(sb.b["damping"], np.float64, (nb, 2))
        A["body_contactp"] = _arr(sb.b["contactp"], np.float64, (nb, 2))
        A["geom_type"] = _arr(sb.g["type"], np.int32, (ng,))
        A["geom_body"] = _arr(sb.g["body"], np.int32, (ng,))
        A["geom_space"] = _arr(sb.g["space"], np.int32, (ng,))
        A["geom_dims"] = _arr(sb.g["dims"], np.float64, (ng, 4))
        A["geom_pos"] = _arr(sb.g["pos"], np.float64, (ng, 3))
        A["geom_quat"] = _arr(sb.g["quat"], np.float64, (ng, 4))
        A["geom_category_bits"] = _arr(sb.g["cat"], np.uint32, (ng,))
        A["geom_collide_bits"] = _arr(sb.g["col"], np.uint32, (ng,))
        A["geom_collide_bitmask"] = _arr(sb.g["bitmask"], np.uint32, (ng,))
        A["geom_max_contacts"] = _arr(sb.g["maxc"], np.int32, (ng,))
        A["geom_surf"] = _arr(sb.g["surf"], np.float64, (ng, SF_COUNT))
        A["geom_fdir1"] = _arr(sb.g["fdir1"], np.float64, (ng, 3))
        A["joint_type"] = _arr(sb.j["type"], np.int32, (nj,))
        A["joint_body1"] = _arr(sb.j["b1"], np.int32, (nj,))
        A["joint_body2"] = _arr(sb.j["b2"], np.int32, (nj,))
        A["joint_anchor1"] = _arr(sb.j["anchor1"], np.float64, (nj, 3))
        A["joint_anchor2"] = _arr(sb.j["anchor2"], np.float64, (nj, 3))
        A["joint_axis1"] = _arr(sb.j["axis1"], np.float64, (nj, 3))
        A["joint_axis2"] = _arr(sb.j["axis2"], np.float64, (nj, 3))
        A["joint_qrel"] = _arr(sb.j["qrel"], np.float64, (nj, 4))
        A["joint_param"] = _arr(sb.j["param"], np.float64, (nj, JP_COUNT))
        self.arrays = A
        c = GzScene()
        c.n_worlds = self.n_worlds
        for name, ctype in GzScene._fields_[1:]:
            if name in A:      # (optional fields such as geom_rot stay NULL)
                setattr(c, name, A[name].ctypes.data_as(ctype))
        self.c = c

    def __getattr__(self, k):
        A = self.__dict__.get("arrays", {})
        if k in A:
            return A[k]
        raise AttributeError(k)
