"""ctypes binding of libgzcuda.so (include/gzcuda.h) — the host-side mirror of the
reference's per-step physics interface for a batch of worlds.

``Batch.update_collision()`` / ``Batch.update_physics()`` / ``Batch.step()`` correspond to
``ODEPhysics::UpdateCollision`` / ``UpdatePhysics`` (gazebo/physics/ode/ODEPhysics.cc:366-453);
errors surface as ``GzCudaError`` like ``gzerr``/``gzthrow`` do on the reference side.
There is no CPU fallback: constructing a Batch without the compiled library or without
a CUDA device raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgzcuda.so")
_lib = None


class GzCudaError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise GzCudaError("libgzcuda.so is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, ci, dp = C.c_void_p, C.c_int, C.c_void_p
    L.gzcuda_version.restype = ci
    L.gzcuda_last_error.restype = C.c_char_p
    L.gzcuda_device_count.restype = ci
    L.gzcuda_batch_create.restype = vp
    L.gzcuda_batch_create.argtypes = [vp, vp, ci]
    L.gzcuda_batch_destroy.argtypes = [vp]
    L.gzcuda_batch_step.argtypes = [vp, ci]
    L.gzcuda_batch_step_io.argtypes = [vp, ci, dp, dp]
    L.gzcuda_batch_step_log.argtypes = [vp, ci, ci, dp, ci]
    L.gzcuda_batch_set_profiling.argtypes = [vp, ci]
    L.gzcuda_batch_pgs_ms.argtypes = [vp, dp, dp]
    L.gzcuda_batch_get_counters.argtypes = [vp, dp, ci]
    L.gzcuda_batch_get_pgs_debug.argtypes = [vp, dp]
    L.gzcuda_batch_get_pgs_profile.argtypes = [vp, dp]
    L.gzcuda_batch_get_pgs_timeline.argtypes = [vp, dp]
    L.gzcuda_batch_update_collision.argtypes = [vp]
    L.gzcuda_batch_update_physics.argtypes = [vp]
    L.gzcuda_world_step.argtypes = [vp, C.c_double]
    L.gzcuda_batch_sync.argtypes = [vp]
    L.gzcuda_batch_get_body_state.argtypes = [vp, dp, dp, dp, dp]
    L.gzcuda_batch_set_body_state.argtypes = [vp, ci, ci, dp, dp, dp, dp]
    L.gzcuda_batch_set_body_state_raw.argtypes = [vp, ci, ci, dp, dp, dp, dp]
    L.gzcuda_batch_set_body_enabled.argtypes = [vp, ci, ci, dp]
    L.gzcuda_batch_add_body_wrench.argtypes = [vp, dp, dp]
    L.gzcuda_batch_get_body_flags.argtypes = [vp, dp]
    L.gzcuda_batch_set_joint_param.argtypes = [vp, ci, ci, C.c_double]
    L.gzcuda_batch_get_joint_state.argtypes = [vp, dp, dp, dp]
    L.gzcuda_batch_set_hinge_pd.argtypes = [vp, dp]
    L.gzcuda_batch_set_joint_velocity_cmd.argtypes = [vp, dp]
    L.gzcuda_batch_get_stats.argtypes = [vp, dp, dp, dp]
    L.gzcuda_batch_set_param.argtypes = [vp, C.c_char_p, C.c_double]
    L.gzcuda_batch_get_pairs.argtypes = [vp, ci, dp, ci]
    L.gzcuda_batch_get_contacts.argtypes = [vp, ci, dp, dp, ci]
    L.gzcuda_batch_ray_query.argtypes = [vp, ci, dp, dp, dp, dp, dp, dp]
    L.gzcuda_batch_get_joint_feedback.argtypes = [vp, dp]
    L.gzcuda_batch_get_contact_feedback.argtypes = [vp, ci, dp, ci]
    L.gzcuda_batch_get_joint_feedback_noalloc.argtypes = [vp, dp]
    L.gzcuda_batch_last_step_ms.argtypes = [vp, dp, dp]
    L.gzcuda_batch_stage_ms.argtypes = [vp, dp]
    _lib = L
    return L


EXPORTS = [
    "gzcuda_version", "gzcuda_last_error", "gzcuda_device_count", "gzcuda_batch_create",
    "gzcuda_batch_destroy", "gzcuda_batch_step", "gzcuda_batch_update_collision",
    "gzcuda_batch_update_physics", "gzcuda_world_step", "gzcuda_batch_sync",
    "gzcuda_batch_get_body_state", "gzcuda_batch_set_body_state", "gzcuda_batch_add_body_wrench",
    "gzcuda_batch_get_body_flags", "gzcuda_batch_set_joint_param", "gzcuda_batch_get_joint_state",
    "gzcuda_batch_set_hinge_pd", "gzcuda_batch_get_stats", "gzcuda_batch_set_param",
    "gzcuda_batch_get_pairs", "gzcuda_batch_get_contacts", "gzcuda_batch_last_step_ms",
    "gzcuda_batch_stage_ms", "gzcuda_batch_step_io", "gzcuda_batch_set_profiling",
    "gzcuda_batch_pgs_ms", "gzcuda_batch_get_counters", "gzcuda_batch_get_pgs_debug",
    "gzcuda_batch_get_pgs_profile", "gzcuda_batch_get_pgs_timeline",
    "gzcuda_batch_get_joint_feedback", "gzcuda_batch_get_contact_feedback", "gzcuda_batch_step_log",
    "gzcuda_batch_set_joint_velocity_cmd", "gzcuda_batch_get_joint_feedback_noalloc",
    "gzcuda_batch_set_body_state_raw", "gzcuda_batch_set_body_enabled", "gzcuda_batch_ray_query",
]


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Batch:
    """A batch of independent worlds resident on one GPU."""

    def __init__(self, scene, params, device=0):
        self.L = lib()
        self.scene, self.params = scene, params
        self.h = self.L.gzcuda_batch_create(C.byref(scene.c), C.byref(params), device)
        if not self.h:
            raise GzCudaError(self.L.gzcuda_last_error().decode())

    def _ck(self, rc):
        if rc < 0:
            raise GzCudaError(self.L.gzcuda_last_error().decode())
        return rc

    def close(self):
        if getattr(self, "h", None):
            self.L.gzcuda_batch_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- stepping
    def step(self, n=1):
        self._ck(self.L.gzcuda_batch_step(self.h, n))

    def step_io(self, n, wrench_ptr, state_ptr):
        """Host-buffer step: raw pointers (ints) to [nb][6] wrench in and [nb][13] state out."""
        self._ck(self.L.gzcuda_batch_step_io(self.h, n, wrench_ptr, state_ptr))

    def set_profiling(self, on):
        self._ck(self.L.gzcuda_batch_set_profiling(self.h, int(on)))

    def pgs_ms(self):
        ms = C.c_float(0)
        n = C.c_int(0)
        self._ck(self.L.gzcuda_batch_pgs_ms(self.h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def counters(self, reset=False):
        out = (C.c_ulonglong * 4)()
        self._ck(self.L.gzcuda_batch_get_counters(self.h, out, int(reset)))
        return list(out)

    def pgs_debug(self):
        out = np.empty((self.scene.n_worlds, 3), dtype=np.int32)
        self._ck(self.L.gzcuda_batch_get_pgs_debug(self.h, _p(out)))
        return out

    def pgs_profile(self):
        out = (C.c_ulonglong * 16)()
        self._ck(self.L.gzcuda_batch_get_pgs_profile(self.h, out))
        return list(out)

    def pgs_timeline(self):
        out = np.zeros((self.scene.n_worlds, 2), dtype=np.uint64)
        self._ck(self.L.gzcuda_batch_get_pgs_timeline(self.h, _p(out)))
        return out

    def update_collision(self):
        self._ck(self.L.gzcuda_batch_update_collision(self.h))

    def update_physics(self):
        self._ck(self.L.gzcuda_batch_update_physics(self.h))

    def sync(self):
        self._ck(self.L.gzcuda_batch_sync(self.h))

    # ---- state
    def body_state(self):
        nb = self.scene.nb
        pos, quat = np.empty((nb, 3)), np.empty((nb, 4))
        lvel, avel = np.empty((nb, 3)), np.empty((nb, 3))
        self._ck(self.L.gzcuda_batch_get_body_state(self.h, _p(pos), _p(quat), _p(lvel), _p(avel)))
        return pos, quat, lvel, avel

    def set_body_state(self, first, pos=None, quat=None, lvel=None, avel=None):
        arrs = [None if a is None else np.ascontiguousarray(a, dtype=np.float64) for a in (pos, quat, lvel, avel)]
        count = next(len(a) for a in arrs if a is not None)
        self._ck(self.L.gzcuda_batch_set_body_state(self.h, first, count, *[_p(a) for a in arrs]))

    def add_body_wrench(self, force=None, torque=None):
        f = None if force is None else np.ascontiguousarray(force, dtype=np.float64)
        t = None if torque is None else np.ascontiguousarray(torque, dtype=np.float64)
        self._ck(self.L.gzcuda_batch_add_body_wrench(self.h, _p(f), _p(t)))

    def body_flags(self):
        f = np.empty(self.scene.nb, dtype=np.int32)
        self._ck(self.L.gzcuda_batch_get_body_flags(self.h, _p(f)))
        return f

    def set_joint_param(self, joint, slot, value):
        self._ck(self.L.gzcuda_batch_set_joint_param(self.h, joint, slot, float(value)))

    def joint_state(self):
        nj = self.scene.nj
        lam, lam_erp, cum = np.empty((nj, 6)), np.empty((nj, 6)), np.empty(nj)
        self._ck(self.L.gzcuda_batch_get_joint_state(self.h, _p(lam), _p(lam_erp), _p(cum)))
        return lam, lam_erp, cum

    def set_hinge_pd(self, pd):
        pd = np.ascontiguousarray(pd, dtype=np.float64).reshape(self.scene.nj, 6)
        self._ck(self.L.gzcuda_batch_set_hinge_pd(self.h, _p(pd)))

    def set_joint_velocity_cmd(self, vel):
        """vel[nj] rad/s applied before every step as Joint::SetVelocity does (NaN = no command)."""
        vel = np.ascontiguousarray(vel, dtype=np.float64).reshape(self.scene.nj)
        self._ck(self.L.gzcuda_batch_set_joint_velocity_cmd(self.h, _p(vel)))

    def stats(self):
        nw = self.scene.n_worlds
        rms = np.empty((nw, 4))
        nc = np.empty(nw, dtype=np.int32)
        m = np.empty(nw, dtype=np.int32)
        self._ck(self.L.gzcuda_batch_get_stats(self.h, _p(rms), _p(nc), _p(m)))
        return rms, nc, m

    def set_param(self, key, value):
        self._ck(self.L.gzcuda_batch_set_param(self.h, key.encode(), float(value)))

    def pairs(self, world, cap=1 << 20):
        out = np.empty((cap, 2), dtype=np.int32)
        n = self._ck(self.L.gzcuda_batch_get_pairs(self.h, world, _p(out), cap))
        return out[:n].copy()

    def ray_query(self, world, start, direction, length):
        """The CPU laser on the device (gzcuda_batch_ray_query): world = one index for all rays or one per ray.
        Returns (range [n], world-local geom index of the hit or -1 [n])."""
        start = np.ascontiguousarray(start, dtype=np.float64).reshape(-1, 3)
        direction = np.ascontiguousarray(direction, dtype=np.float64).reshape(-1, 3)
        length = np.ascontiguousarray(length, dtype=np.float64).reshape(-1)
        n = len(length)
        worlds = np.ascontiguousarray(np.broadcast_to(np.asarray(world, dtype=np.int32), (n,)))
        rng, geom = np.empty(n), np.empty(n, dtype=np.int32)
        self._ck(self.L.gzcuda_batch_ray_query(self.h, n, _p(worlds), _p(start), _p(direction), _p(length), _p(rng), _p(geom)))
        return rng, geom

    def contacts(self, world, cap=1 << 16):
        g = np.empty((cap, 2), dtype=np.int32)
        d = np.empty((cap, 7))
        n = self._ck(self.L.gzcuda_batch_get_contacts(self.h, world, _p(g), _p(d), cap))
        return g[:n].copy(), d[:n].copy()

    def step_log(self, n_steps, every, frames=None):
        """n_steps steps; returns frames[n_steps // every][nb][13] = pos, quat, lvel, avel after every `every`-th step."""
        nf = n_steps // every
        if frames is None:
            frames = np.empty((max(nf, 1), self.scene.nb, 13))
        n = self._ck(self.L.gzcuda_batch_step_log(self.h, n_steps, every, _p(frames), nf))
        return frames[:n]

    def joint_feedback(self):
        """dJointFeedback of every persistent joint after the last step: [nj][12] = f1, t1, f2, t2."""
        fb = np.zeros((self.scene.nj, 12))
        self._ck(self.L.gzcuda_batch_get_joint_feedback(self.h, _p(fb)))
        return fb

    def joint_feedback_noalloc(self):
        """joint_feedback() through the entry point that allocates nothing (static device scratch inside the
        library, at most GZCUDA_NOALLOC_MAX_JOINTS joints; no counterpart in the reference)."""
        fb = np.zeros((self.scene.nj, 12))
        self._ck(self.L.gzcuda_batch_get_joint_feedback_noalloc(self.h, _p(fb)))
        return fb

    def contact_feedback(self, world, cap=1 << 16):
        fb = np.zeros((cap, 12))
        n = self._ck(self.L.gzcuda_batch_get_contact_feedback(self.h, world, _p(fb), cap))
        return fb[:n].copy()

    def last_step_ms(self):
        ms = C.c_float(0)
        n = C.c_long(0)
        self._ck(self.L.gzcuda_batch_last_step_ms(self.h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def stage_ms(self):
        ms = (C.c_float * 4)()
        self._ck(self.L.gzcuda_batch_stage_ms(self.h, ms))
        return list(ms)
