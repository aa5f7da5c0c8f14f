"""ctypes wrapper over oracle/_ref/libgzref*.so — the reference's own compiled opende
driven by oracle/ref_harness.cpp.  ORACLE / TEST INFRASTRUCTURE: only tests/,
__graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module; the
product (gazebo_b200/, libgzcuda.so) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_DIR = os.path.join(_HERE, "_ref")
_libs = {}


def build(flavour="parity", jobs=8):
    """Compile the reference in place (needs /root/reference); no-op when the .so exists."""
    target = "all" if flavour == "parity" else "fast"
    subprocess.check_call(["make", "-C", _HERE, "-j%d" % jobs, target],
                          stdout=subprocess.DEVNULL)


def available(flavour="parity"):
    return os.path.exists(_lib_path(flavour))


def _lib_path(flavour):
    return os.path.join(_REF_DIR, "libgzref.so" if flavour == "parity" else "libgzref_fast.so")


def _load(flavour):
    if flavour in _libs:
        return _libs[flavour]
    path = _lib_path(flavour)
    if not os.path.exists(path):
        raise RuntimeError("oracle library %s missing: run `make -C oracle`" % path)
    # RTLD_LOCAL + distinct sonames keep the two flavours' ODE symbols apart
    lib = C.CDLL(path, mode=os.RTLD_LOCAL | os.RTLD_NOW)
    lib.gzref_create.restype = C.c_void_p
    lib.gzref_create.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    lib.gzref_destroy.argtypes = [C.c_void_p]
    lib.gzref_set_hinge_pd.argtypes = [C.c_void_p, C.c_void_p]
    lib.gzref_step.argtypes = [C.c_void_p, C.c_int]
    lib.gzref_add_body_wrench.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.gzref_ray_query.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.gzref_step_mt.argtypes = [C.c_void_p, C.c_int, C.c_int]
    lib.gzref_collide.argtypes = [C.c_void_p, C.c_int]
    lib.gzref_get_pairs.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    lib.gzref_get_contacts.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
    lib.gzref_get_body_state.argtypes = [C.c_void_p] * 5
    lib.gzref_get_body_flags.argtypes = [C.c_void_p, C.c_void_p]
    lib.gzref_get_stats.argtypes = [C.c_void_p] * 4
    lib.gzref_get_joint_state.argtypes = [C.c_void_p] * 4
    lib.gzref_enable_feedback.argtypes = [C.c_void_p]
    lib.gzref_set_joint_velocity_cmd.argtypes = [C.c_void_p, C.c_void_p]
    lib.gzref_get_joint_feedback.argtypes = [C.c_void_p, C.c_void_p]
    lib.gzref_get_contact_feedback.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    lib.gzref_world_handles.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    lib.gzref_plugin_update.argtypes = [C.c_void_p, C.c_int]
    _libs[flavour] = lib
    return lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class RefBatch:
    """mode 1 = canonical (deterministic row order, parity oracle); mode 0 = natural gzserver order."""

    def __init__(self, scene, params, mode=1, flavour="parity"):
        self.lib = _load(flavour)
        self.scene, self.params = scene, params
        self.h = self.lib.gzref_create(C.byref(scene.c), C.byref(params), mode)
        if not self.h:
            raise RuntimeError("gzref_create failed (unsupported geom/joint type?)")

    def close(self):
        if self.h:
            self.lib.gzref_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_hinge_pd(self, pd):
        pd = np.ascontiguousarray(pd, dtype=np.float64).reshape(self.scene.nj, 6)
        self.lib.gzref_set_hinge_pd(self.h, _p(pd))

    def add_body_wrench(self, force=None, torque=None):
        """Link::AddForce / AddTorque (world frame) for the next step; [nb][3] each."""
        f = None if force is None else np.ascontiguousarray(force, dtype=np.float64).reshape(self.scene.nb, 3)
        t = None if torque is None else np.ascontiguousarray(torque, dtype=np.float64).reshape(self.scene.nb, 3)
        self.lib.gzref_add_body_wrench(self.h, None if f is None else _p(f), None if t is None else _p(t))

    def step(self, n=1, threads=0):
        if threads and threads > 1:
            self.lib.gzref_step_mt(self.h, n, threads)
        else:
            self.lib.gzref_step(self.h, n)

    def body_state(self):
        nb = self.scene.nb
        pos, quat = np.empty((nb, 3)), np.empty((nb, 4))
        lvel, avel = np.empty((nb, 3)), np.empty((nb, 3))
        self.lib.gzref_get_body_state(self.h, _p(pos), _p(quat), _p(lvel), _p(avel))
        return pos, quat, lvel, avel

    def body_flags(self):
        f = np.empty(self.scene.nb, dtype=np.int32)
        self.lib.gzref_get_body_flags(self.h, _p(f))
        return f

    def collide(self, world):
        return self.lib.gzref_collide(self.h, world)

    def pairs(self, world, cap=1 << 20):
        out = np.empty((cap, 2), dtype=np.int32)
        n = self.lib.gzref_get_pairs(self.h, world, _p(out), cap)
        return out[:n].copy()

    def contacts(self, world, cap=1 << 16):
        g = np.empty((cap, 2), dtype=np.int32)
        d = np.empty((cap, 7))
        n = self.lib.gzref_get_contacts(self.h, world, _p(g), _p(d), cap)
        return g[:n].copy(), d[:n].copy()

    def stats(self):
        nw = self.scene.n_worlds
        rms = np.empty((nw, 4))
        nc = np.empty(nw, dtype=np.int32)
        ncj = np.empty(nw, dtype=np.int32)
        self.lib.gzref_get_stats(self.h, _p(rms), _p(nc), _p(ncj))
        return rms, nc, ncj

    def joint_state(self):
        nj = self.scene.nj
        lam, lam_erp, cum = np.empty((nj, 6)), np.empty((nj, 6)), np.empty(nj)
        self.lib.gzref_get_joint_state(self.h, _p(lam), _p(lam_erp), _p(cum))
        return lam, lam_erp, cum

    def enable_feedback(self):
        """dJointSetFeedback on every persistent joint and, from the next step on, every contact joint."""
        self.lib.gzref_enable_feedback(self.h)

    def joint_feedback(self):
        fb = np.zeros((self.scene.nj, 12))
        self.lib.gzref_get_joint_feedback(self.h, _p(fb))
        return fb

    def contact_feedback(self, world, cap=1 << 16):
        fb = np.zeros((cap, 12))
        n = self.lib.gzref_get_contact_feedback(self.h, world, _p(fb), cap)
        return fb[:n].copy()

    def world_handles(self, world):
        """(dxWorld*, top dxSpace*, user pointer of gzref_surface_cb) of one world: the live ODE objects, for the
        tests of the host adapter (libgzcuda_ode.so)."""
        out = (C.c_void_p * 3)()
        self.lib.gzref_world_handles(self.h, world, out)
        return out[0], out[1], out[2]

    def surface_cb_address(self):
        return C.cast(self.lib.gzref_surface_cb, C.c_void_p).value

    def plugin_update(self, world):
        """The plugin part of a step (PD torques, joint velocity commands) applied to the live ODE objects."""
        self.lib.gzref_plugin_update(self.h, world)

    def ray_query(self, world, start, direction, length):
        """The CPU laser (ODEMultiRayShape::UpdateRays restated over the reference's ray colliders): rays of ONE world.
        Returns (range [n], world-local geom index of the hit or -1 [n])."""
        start = np.ascontiguousarray(start, dtype=np.float64).reshape(-1, 3)
        direction = np.ascontiguousarray(direction, dtype=np.float64).reshape(-1, 3)
        length = np.ascontiguousarray(length, dtype=np.float64).reshape(-1)
        n = len(length)
        rng, geom = np.empty(n), np.empty(n, dtype=np.int32)
        rc = self.lib.gzref_ray_query(C.c_void_p(self.h), int(world), n, _p(start), _p(direction), _p(length), _p(rng), _p(geom))
        if rc != 0:
            raise RuntimeError("gzref_ray_query failed")
        return rng, geom

    def phase_seconds(self, reset=False):
        """Thread-seconds in UpdateCollision / UpdatePhysics since the last reset (all batches of this flavour)."""
        out = (C.c_longlong * 2)()
        self.lib.gzref_phase_ns(out, int(reset))
        return out[0] * 1e-9, out[1] * 1e-9

    def set_joint_velocity_cmd(self, vel):
        vel = np.ascontiguousarray(vel, dtype=np.float64).reshape(self.scene.nj)
        self.lib.gzref_set_joint_velocity_cmd(self.h, _p(vel))
