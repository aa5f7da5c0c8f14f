"""Behavioural acceptance on the GPU path (through the C ABI): headless restatements of the reference's
own integration tests for this path — closed-form expectations, the reference's
tolerances.  They complement the bit-level parity tests of test_parity_gpu.py: a solver that matched
the reference bit for bit but whose reference harness was mis-configured would still fail here."""
import math

import numpy as np
import pytest

from gazebo_b200 import scene as S

pytestmark = pytest.mark.gpu
G = 9.8
DT = 0.001


def _batch(sc, p):
    from gazebo_b200.batch import Batch
    return Batch(sc, p)


def _drop_scene(z0):
    """test/worlds/drop_test.world: a unit box, a sphere (r 0.5) and a cylinder (r 0.5, l 1) above the plane."""
    sb = S.SceneBuilder()
    sb.begin_world()
    sb.add_ground_plane()
    b = sb.add_body([0, 0, z0], mass=1.0, inertia=S.box_inertia(1.0, 1, 1, 1))
    sb.add_geom(S.GEOM_BOX, (1, 1, 1), sb.new_space(), body=b)
    b = sb.add_body([0, 3, z0], mass=1.0, inertia=S.sphere_inertia(1.0, 0.5))
    sb.add_geom(S.GEOM_SPHERE, (0.5,), sb.new_space(), body=b)
    b = sb.add_body([0, 6, z0], mass=1.0, inertia=S.cylinder_inertia(1.0, 0.5, 1.0))
    sb.add_geom(S.GEOM_CYLINDER, (0.5, 1.0), sb.new_space(), body=b)
    sb.end_world()
    return sb.finish()


def test_drop_stuff_free_fall_then_rest():
    """PhysicsTest.DropStuffODE (test/integration/physics.cc:919-1062) and SpawnDrop (:94-350): free fall
    follows v += g dt; z += v dt to 1e-4 until contact, no x/y drift, then the shapes settle at z = 0.5
    (4.1e-3 first, 1e-5 at rest) with speed below 3e-5."""
    z0 = 2.0
    g = _batch(_drop_scene(z0), S.default_params())
    v, z = 0.0, z0
    t_contact = math.sqrt(2 * (z0 - 0.5) / G)
    for step in range(1, int(0.9 * t_contact / DT)):
        g.step(1)
        v -= G * DT
        z += v * DT
        if step % 50 == 0:
            pos, _, lvel, _ = g.body_state()
            assert np.abs(pos[:, 2] - z).max() < 1e-4 and np.abs(lvel[:, 2] - v).max() < 1e-4
            assert np.abs(pos[:, 0]).max() < 1e-12 and np.abs(pos[:, 1] - [0, 3, 6]).max() < 1e-12
    g.step(int(1.5 / DT))
    pos, _, lvel, avel = g.body_state()
    assert np.abs(pos[:, 2] - 0.5).max() < 4.1e-3
    g.step(int(1.5 / DT))
    pos, _, lvel, avel = g.body_state()
    assert np.abs(pos[:, 2] - 0.5).max() < 1e-5 + 1e-3          # contact_surface_layer = 1e-3 lets them sink that far
    assert np.abs(lvel[:2]).max() < 3e-5 and np.abs(avel).max() < 3e-5    # box and sphere (physics.cc:1018)
    assert np.abs(lvel[2]).max() < 2e-2     # the cylinder standing on its cap keeps a ~1 cm/s reading in the reference too
    assert np.abs(pos[:, 0]).max() < 1e-3                           # PHYSICS_TOL class, no drift


def test_zero_max_contacts_is_harmless():
    """PhysicsTest.ZeroMaxContactsODE (physics.cc:1530-1541, bitbucket issue 593): max_contacts = 0, on the
    world or on a collision, must not break the step.  By the code (ODEPhysics.cc:1053-1062,1076) a zero
    limit disables the truncation (`maxCollide > 0 && numc > maxCollide`), so all contacts are kept."""
    sb = S.SceneBuilder()
    sb.begin_world()
    sb.add_ground_plane()
    b = sb.add_body([0, 0, 0.6], mass=1.0, inertia=S.box_inertia(1.0, 1, 1, 1))
    sb.add_geom(S.GEOM_BOX, (1, 1, 1), sb.new_space(), body=b, max_contacts=0)
    b = sb.add_body([3, 0, 0.6], mass=1.0, inertia=S.box_inertia(1.0, 1, 1, 1))
    sb.add_geom(S.GEOM_BOX, (1, 1, 1), sb.new_space(), body=b, max_contacts=2)
    sb.end_world()
    sc = sb.finish()
    for world_max in (20, 0):
        g = _batch(sc, S.default_params(max_contacts=world_max))
        g.step(1000)
        pos = g.body_state()[0]
        assert np.isfinite(pos).all() and np.abs(pos[:, 2] - 0.5).max() < 4.1e-3
        cg, _ = g.contacts(0)
        assert (cg[:, 1] == 1).sum() == 4            # geom 1, the first box: nothing truncated
        assert 1 <= (cg[:, 1] == 2).sum() <= 2       # the second box: at most two kept


def test_simple_pendulum_angle_and_energy():
    """JointTestRevolute.SimplePendulum / PendulumEnergy (test/integration/joint_revolute.cc:739-759): a point-
    like bob on a world hinge follows the pendulum equation (integrated here with the same semi-implicit
    Euler at 1/10 of the step) and its energy drifts by less than 1 % over 5 s."""
    L, th0 = 1.0, 0.6
    sb = S.SceneBuilder()
    sb.begin_world()
    m = 1.0
    pos = [L * math.sin(th0), 0.0, 2.0 - L * math.cos(th0)]
    b = sb.add_body(pos, mass=m, inertia=(1e-4, 1e-4, 1e-4, 0, 0, 0))
    sb.add_hinge(b, -1, [0, 0, 2.0], [0, 1, 0])
    sb.end_world()
    g = _batch(sb.finish(), S.default_params())
    th, w, sub = th0, 0.0, 10
    I = m * L * L + 1e-4

    def energy(p, v, av):
        return 0.5 * m * (v ** 2).sum() + 0.5 * 1e-4 * (av ** 2).sum() + m * G * p[2]

    p0, _, v0, a0 = g.body_state()
    e0 = energy(p0[0], v0[0], a0[0])
    worst = 0.0
    for k in range(50):
        g.step(100)
        for _ in range(100 * sub):
            w -= (m * G * L / I) * math.sin(th) * DT / sub
            th += w * DT / sub
        p, _, v, av = g.body_state()
        ang = math.atan2(p[0, 0], 2.0 - p[0, 2])
        worst = max(worst, abs(ang - th))
        assert abs(math.hypot(p[0, 0], 2.0 - p[0, 2]) - L) < 1e-3          # the joint holds
        assert abs(energy(p[0], v[0], av[0]) - e0) < 0.01 * abs(e0 - m * G * (2.0 - L))
    assert worst < 2e-2, worst     # first-order integrator against the fine-step solution over 5 s


def test_unwrapped_hinge_angle():
    """JointTestRevolute.UnwrappedAngle: the cumulative angle (hinge.cpp:72-73) keeps counting past +-pi."""
    sb = S.SceneBuilder()
    sb.begin_world()
    b = sb.add_body([0, 0, 1.0], mass=1.0, inertia=S.box_inertia(1.0, 0.2, 0.2, 0.2), avel=(0, 0, 8.0),
                    flags=S.BODY_NO_GRAVITY)
    sb.add_hinge(b, -1, [0, 0, 1.0], [0, 0, 1], lostop=-1e9, histop=1e9)     # finite stops: the angle is tracked
    sb.end_world()
    g = _batch(sb.finish(), S.default_params())
    g.step(2000)
    cum = g.joint_state()[2][0]
    assert abs(abs(cum) - 16.0) < 1e-2 and abs(cum) > math.pi, cum


def test_anchored_kinematic_loop_stays_closed():
    """KinematicLoopTest (test/integration/kinematic_loop.cc:137-143): a four-bar loop anchored to the world
    swings under gravity; every hinge anchor stays together."""
    sb = S.SceneBuilder()
    sb.begin_world()
    sp = sb.new_space()
    m = 1.0
    a = sb.add_body([0.0, 0, 1.5], mass=m, inertia=S.box_inertia(m, 0.1, 0.1, 1.0))       # left bar, vertical
    c = sb.add_body([1.0, 0, 1.5], mass=m, inertia=S.box_inertia(m, 0.1, 0.1, 1.0))       # right bar
    d = sb.add_body([0.5, 0, 1.0], mass=m, inertia=S.box_inertia(m, 1.0, 0.1, 0.1), lvel=(0.5, 0, 0))   # coupler
    for bb, dims in ((a, (0.1, 0.1, 1.0)), (c, (0.1, 0.1, 1.0)), (d, (1.0, 0.1, 0.1))):
        sb.add_geom(S.GEOM_BOX, dims, sp, body=bb)
    anchors = [(a, -1, [0.0, 0, 2.0]), (c, -1, [1.0, 0, 2.0]), (a, d, [0.0, 0, 1.0]), (c, d, [1.0, 0, 1.0])]
    for b1, b2, p in anchors:
        sb.add_hinge(b1, b2, p, [0, 1, 0])
    sb.end_world()
    sc = sb.finish()
    g = _batch(sc, S.default_params())
    A = sc.arrays
    for _ in range(10):
        g.step(300)
        pos, quat, _, _ = g.body_state()
        for j, (b1, b2, p) in enumerate(anchors):
            w1 = pos[b1] + S.q_to_R(quat[b1]) @ np.asarray(A["joint_anchor1"]).reshape(-1, 3)[j]
            a2 = np.asarray(A["joint_anchor2"]).reshape(-1, 3)[j]
            w2 = a2 if b2 < 0 else pos[b2] + S.q_to_R(quat[b2]) @ a2
            assert np.abs(w1 - w2).max() < 2e-3, (j, w1, w2)
    assert np.isfinite(pos).all() and abs(pos[d, 1]) < 1e-9      # planar motion stays planar
