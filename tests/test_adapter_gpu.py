"""The C++ host adapter (host/gzcuda_ode.cpp -> gazebo_b200/libgzcuda_ode.so): libgzcuda behind the reference's own
step interface, `int (*physicsStepFunc)(dxWorld*, dReal)` (ODEPhysicsPrivate.hh:65), on LIVE ODE objects.

Two copies of a world are built through the public ODE C API (by the oracle harness): one is stepped by the
reference itself (canonical order), the other by gzcuda_dspace_collide + gzcuda_dworld_step on its dxWorld /
dxSpace.  What the adapter leaves in the dxBody objects (read back through the ODE objects, not through libgzcuda)
must be the reference's state; a plugin-style loop that calls dJointAddHingeTorque or writes body velocities between
steps must see them honoured; sleep flags and moved geoms must follow."""
import ctypes as C
import os

import numpy as np
import pytest

from tests import scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TIGHT = 1e-9


@pytest.fixture(scope="module")
def adapter(oracle):
    path = os.path.join(ROOT, "gazebo_b200", "libgzcuda_ode.so")
    if not os.path.exists(path):
        pytest.fail("libgzcuda_ode.so is not built (make -C host; needs the reference's headers)")
    # the adapter's ODE calls (dGeomMoved) resolve against the ODE the process carries: the oracle's liboderef.so,
    # made global here (the same instance the harness built the worlds with)
    C.CDLL(os.path.join(ROOT, "oracle", "_ref", "liboderef.so"), mode=os.RTLD_GLOBAL | os.RTLD_NOW)
    L = C.CDLL(path)
    vp = C.c_void_p
    L.gzcuda_ode_set_surface_callback.argtypes = [vp, vp, vp]
    L.gzcuda_ode_set_max_contacts.argtypes = [vp, C.c_int]
    L.gzcuda_dspace_collide.argtypes = [vp, vp]
    L.gzcuda_dworld_step.argtypes = [vp, C.c_double]
    L.gzcuda_dworld_release.argtypes = [vp]
    L.gzcuda_topology_changed.argtypes = [vp]
    L.gzcuda_dworld_ray_query.argtypes = [vp, C.c_int, vp, vp, vp, vp, vp]
    L.gzcuda_ode_last_error.restype = C.c_char_p
    return L


def _run(adapter, oracle, sc, p, pd=None, vel=None, steps=300, every=50, tol=TIGHT):
    ref = oracle.RefBatch(sc, p, mode=1)
    live = oracle.RefBatch(sc, p, mode=1)          # only built, never stepped by the harness
    for r in (ref, live):
        if pd is not None:
            r.set_hinge_pd(pd)
        if vel is not None:
            r.set_joint_velocity_cmd(vel)
    handles = [live.world_handles(w) for w in range(sc.n_worlds)]
    for world, top, user in handles:
        adapter.gzcuda_ode_set_surface_callback(world, live.surface_cb_address(), user)
        adapter.gzcuda_ode_set_max_contacts(world, p.max_contacts)
    try:
        for k in range(steps):
            for w, (world, top, user) in enumerate(handles):
                live.plugin_update(w)                                   # plugins: torques / velocities into the dxBody objects
                assert adapter.gzcuda_dspace_collide(top, world) == 0, adapter.gzcuda_ode_last_error()
                assert adapter.gzcuda_dworld_step(world, p.step_size) == 1, adapter.gzcuda_ode_last_error()
            ref.step(1)
            if (k + 1) % every == 0:
                gs, rs = live.body_state(), ref.body_state()            # read from the ODE objects on both sides
                err = max(np.abs(a - b).max() for a, b in zip(gs, rs))
                assert err <= tol, (k, err)
                assert np.array_equal(live.body_flags(), ref.body_flags())
    finally:
        for world, top, user in handles:
            adapter.gzcuda_dworld_release(world)
    return live, ref


@pytest.mark.gpu
def test_adapter_drop_test_matches_dworldquickstep(adapter, oracle):
    sc, p, _ = scenes.box_drop()
    live, ref = _run(adapter, oracle, sc, p, steps=1500, every=250)
    assert abs(live.body_state()[0][0, 2] - 0.5) < 1e-3          # PhysicsTest.DropStuff: the box rests on the plane


@pytest.mark.gpu
def test_adapter_revolve_robots_with_plugin_torques(adapter, oracle):
    """dJointAddHingeTorque between the steps (what a Revolve controller plugin does): the adapter uploads the
    accumulators, returns the hinge's cumulative angle (the PD law reads it through dJointGetHingeAngle) and zeroes
    the accumulators like the reference step."""
    sc, p, pd = scenes.revolve4()
    _run(adapter, oracle, sc, p, pd=pd, steps=400, every=100)


@pytest.mark.gpu
def test_adapter_honours_velocities_written_by_a_plugin(adapter, oracle):
    """Joint::SetVelocity -> dBodySetLinearVel / dBodySetAngularVel on the child link every update (DiffDrivePlugin)."""
    from gazebo_b200 import scene as S, worlds as W
    sc, cmd = W.pioneer_worlds(3)
    vel = np.asarray(cmd, dtype=np.float64).reshape(-1)      # two wheel hinges per world, in joint order
    live, ref = _run(adapter, oracle, sc, S.default_params(), vel=vel, steps=400, every=100)
    assert np.linalg.norm(live.body_state()[0][0, :2]) > 0.02   # the robots drive off


@pytest.mark.gpu
def test_adapter_sleep_flags_and_wake_up(adapter, oracle):
    sc, p, _ = scenes.sleepers()
    live, ref = _run(adapter, oracle, sc, p, steps=2400, every=200)
    assert live.body_flags()[1:4].all()


@pytest.mark.gpu
def test_adapter_carts_on_hinge2_joints(adapter, oracle):
    """dxJointHinge2 objects (steering limits and motor, driven axles, suspension) mirrored from the live ODE world:
    anchors, axes, v1 / v2, c0 / s0 and both limit motors read from the dxJointHinge2 the reference built."""
    sc, p, _ = scenes.carts()
    _run(adapter, oracle, sc, p, steps=300, every=50)


@pytest.mark.gpu
def test_adapter_universal_joints_with_stops_and_motors(adapter, oracle):
    """dxJointUniversal objects mirrored from the live world: qrel1 / qrel2 and both limit motors."""
    sc, p, _ = scenes.gimbals()
    _run(adapter, oracle, sc, p, steps=400, every=100)


@pytest.mark.gpu
def test_adapter_screw_joints(adapter, oracle):
    """dxJointScrew objects mirrored from the live world: thread pitch, offset, the stops of the rotation."""
    sc, p, _ = scenes.screws()
    _run(adapter, oracle, sc, p, steps=400, every=100)


@pytest.mark.gpu
def test_adapter_gearbox_joints(adapter, oracle):
    """dxJointGearbox objects mirrored from the live world: axes, qrel1 / qrel2, ratio, the reference bodies by index."""
    sc, p, _ = scenes.gears()
    _run(adapter, oracle, sc, p, steps=400, every=100)


@pytest.mark.gpu
def test_adapter_hertzian_contacts(adapter, oracle):
    """Elastic modulus, Poisson's ratio, patch / surface radius travel through the surface callback (gzcuda_surface)."""
    sc, p, _ = scenes.hertz()
    _run(adapter, oracle, sc, p, steps=400, every=100)


@pytest.mark.gpu
def test_adapter_torsional_friction(adapter, oracle):
    """MuTorsion, slipTorsion and the radii travel through the surface callback; four-row contacts over live ODE objects."""
    sc, p, _ = scenes.spinners()
    _run(adapter, oracle, sc, p, steps=400, every=100)


@pytest.mark.gpu
def test_adapter_laser_over_live_ode_objects(adapter, oracle):
    """gzcuda_dworld_ray_query: the laser against a LIVE ODE world (what ODEMultiRayShape::UpdateRays does with
    dSpaceCollide2) - ranges bit-exact against the reference's own ray colliders on the same objects, and the dxGeom
    handed back is the one the reference hits."""
    sc, p, _ = scenes.ray_targets(n_worlds=1)
    live = oracle.RefBatch(sc, p, mode=1)
    world, top, user = live.world_handles(0)
    adapter.gzcuda_ode_set_surface_callback(world, live.surface_cb_address(), user)
    adapter.gzcuda_ode_set_max_contacts(world, p.max_contacts)
    try:
        for k in range(20):   # let the shapes move a little: the adapter mirrors the ODE objects' poses
            assert adapter.gzcuda_dspace_collide(top, world) == 0, adapter.gzcuda_ode_last_error()
            assert adapter.gzcuda_dworld_step(world, p.step_size) == 1, adapter.gzcuda_ode_last_error()
        assert adapter.gzcuda_dspace_collide(top, world) == 0
        start, d, length = scenes.laser_rays(2000)
        start, d, length = np.ascontiguousarray(start), np.ascontiguousarray(d), np.ascontiguousarray(length)
        rng = np.empty(len(length)); hit = (C.c_void_p * len(length))()
        vp = lambda a: a.ctypes.data_as(C.c_void_p)
        assert adapter.gzcuda_dworld_ray_query(world, len(length), vp(start), vp(d), vp(length), vp(rng), hit) == 0, adapter.gzcuda_ode_last_error()
        rr, gr = live.ray_query(0, start, d, length)     # the reference's colliders on the same live objects
        assert np.array_equal(rng, rr)
        assert sum(1 for k in range(len(length)) if (hit[k] is not None) != (gr[k] >= 0)) == 0
        assert (gr >= 0).sum() > 300
    finally:
        adapter.gzcuda_dworld_release(world)
