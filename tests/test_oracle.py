"""CPU tests of the oracle (the reference's own opende compiled in place, driven by
oracle/ref_harness.cpp): it must reproduce the committed golden vectors, and it must show
the behaviour the reference's own integration tests assert (test/integration/physics.cc)."""
import os

import numpy as np
import pytest

from tests import scenes

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _pairs_agree(pa, pb, asleep):
    """Pair lists as order-free sets (natural-order stepping reports pairs as the callback saw them).  The two
    sides differ by 1e-5 m in state (row order), so a box that rocks on the plane or falls asleep may touch the
    plane's AABB, or be dropped as a sleeper (ODEPhysics.cc:983-990), a step apart, and after the hit the late box bounces a little differently: at most two pairs
    may differ."""
    sa = set(tuple(sorted(map(int, x))) for x in pa)
    sb = set(tuple(sorted(map(int, x))) for x in pb)
    return len(sa ^ sb) <= 2


@pytest.mark.parametrize("name", sorted(scenes.GOLDEN))
def test_oracle_reproduces_golden(oracle, name):
    builder, steps = scenes.GOLDEN[name]
    sc, p, pd = builder()
    r = oracle.RefBatch(sc, p, mode=1)
    if pd is not None:
        r.set_hinge_pd(pd)
    r.step(steps)
    g = np.load(os.path.join(GOLD, name + ".npz"))
    pos, quat, lvel, avel = r.body_state()
    # same compiler flags, same machine class: bit-identical
    assert np.array_equal(pos, g["pos"]) and np.array_equal(quat, g["quat"])
    assert np.array_equal(lvel, g["lvel"]) and np.array_equal(avel, g["avel"])
    assert np.array_equal(r.pairs(0), g["pairs0"])
    cg, cd = r.contacts(0)
    assert np.array_equal(cg, g["contact_geoms0"]) and np.array_equal(cd, g["contact_data0"])
    assert np.array_equal(r.body_flags(), g["flags"])


def test_gearbox_test(oracle):
    """ODEGearboxJoint_TEST.GearboxTest (test/integration/joint_gearbox.cc:36-112) on the oracle: pins the harness'
    gearbox joint (created by field injection: the reference's C API has no public constructor arguments for it)."""
    sc, p, ids = scenes.gearbox_world()
    scenes.check_gearbox(oracle.RefBatch(sc, p, mode=1), sc, ids)


def test_sphere_slip(oracle):
    """PhysicsFrictionTest.SphereSlip (test/integration/physics_friction.cc:529-666): slip1 / slip2 as a viscous creep -
    a link pushed sideways by gravity on non-sliding contacts drifts at g_y x mass x slip / shapes (1.5 %)."""
    sc, p, expect = scenes.friction_spheres()
    scenes.check_friction_spheres(oracle.RefBatch(sc, p, mode=1), expect)


def test_friction_demo(oracle):
    """PhysicsFrictionTest.FrictionDemo (test/integration/physics_friction.cc:183-286): the pyramid friction's sliding
    threshold to a part in a thousand - boxes with mu just below 1 under gravity (0, -1, -1) accelerate at g_y + mu, those
    with mu >= 1 stay."""
    sc, p, mus = scenes.friction_demo()
    scenes.check_friction_demo(oracle.RefBatch(sc, p, mode=1), mus)


def test_inertia_ratio_pendulum(oracle):
    """PhysicsTest.InertiaRatioPendulum (test/integration/physics_inertia_ratio.cc:39-84): with use_dynamic_moi_rescaling
    the out-of-plane angles of both links stay below g_angle_y_tol = 0.21 / g_angle_z_tol = 0.23 (:30-31) for 3 000 steps;
    without it the light link is thrown out of the plane (which is what the reference's test guards against)."""
    sc, p, _ = scenes.inertia_ratio_pendulum(1)
    my, mz = scenes.max_out_of_plane_angles(oracle.RefBatch(sc, p, mode=1))
    assert my < 0.21 and mz < 0.23, (my, mz)
    sc, p, _ = scenes.inertia_ratio_pendulum(0)
    my0, mz0 = scenes.max_out_of_plane_angles(oracle.RefBatch(sc, p, mode=1))
    assert my0 > 0.21 and my0 > 1.5 * my, (my0, mz0)


def test_elastic_modulus_contact_golden_numbers(oracle):
    """PhysicsTest.ElasticModulusContact (test/integration/elastic_modulus.cc:60-225): the converged depth and Hertzian
    stiffnesses the reference's own test holds (k = 71454.617, k_lin = 11491.286 to float precision) come out of the
    oracle - the one golden vector the reference has for this path pins the harness' surface combination, the
    elastic-modulus normal row, min_depth and the solver."""
    sc, p, _ = scenes.elastic_modulus_contact()
    scenes.check_elastic_modulus_contact(oracle.RefBatch(sc, p, mode=1))


def test_torsional_friction_coefficient_test(oracle):
    """PhysicsTorsionalFrictionTest.CoefficientTest (test/integration/physics_torsional_friction.cc:283-381) on the
    oracle: spheres settle for 1 000 steps, then 1 000 N m about the vertical every step for 10 steps; the angular
    acceleration is (torque - friction torque) / izz with friction torque = m g x mu3 x 3 pi / 16 x patch radius, or
    zero when the friction holds; tolerance 1e-3 as in the reference's test (:29)."""
    sc, p, spec = scenes.torsion_spheres()
    r = oracle.RefBatch(sc, p, mode=1)
    _coefficient_test(r, sc, spec)


def test_torsional_friction_depth_test(oracle):
    """PhysicsTorsionalFrictionTest.DepthTest (physics_torsional_friction.cc:147-281): equilibrium depth m g / kp (1 %)."""
    sc, p, masses = scenes.depth_spheres()
    scenes.check_depth_spheres(oracle.RefBatch(sc, p, mode=1), masses)


def test_torsional_friction_radius_test(oracle):
    """PhysicsTorsionalFrictionTest.RadiusTest (physics_torsional_friction.cc:383-479): the patch radius from the surface
    radius and the depth at equilibrium, a = sqrt(R m g / kp); angular acceleration within 1 % (:475)."""
    sc, p, radii = scenes.torsion_radius_spheres()
    _radius_test(oracle.RefBatch(sc, p, mode=1), sc, radii)


def _radius_test(r, sc, radii):
    mass, izz, applied, kp = 11.309733552923259, 0.045238934211693047, 1000.0, 500000.0
    r.step(1000)
    torque = np.zeros((sc.nb, 3)); torque[:, 2] = applied
    for _ in range(10):
        w0 = r.body_state()[3].copy()
        r.add_body_wrench(torque=torque)
        r.step(1)
        acc = (r.body_state()[3] - w0) / 0.001
        assert np.abs(acc[:, :2]).max() < 1e-3
        for k, rad in enumerate(radii):
            patch = np.sqrt(rad * mass * 9.8 / kp)
            friction = mass * 9.8 * 0.005 * 3 * np.pi * patch / 16
            assert applied > friction
            expected = (applied - friction) / izz
            assert abs(acc[k, 2] - expected) / expected < 0.01
            assert acc[k, 2] < applied / izz   # (and the friction does brake: the reference's 1 % alone would not see it)


def _coefficient_test(r, sc, spec):
    mass, izz, applied = 11.309733552923259, 0.045238934211693047, 1000.0
    r.step(1000)
    torque = np.zeros((sc.nb, 3)); torque[:, 2] = applied
    err = np.zeros(sc.nb)
    moved = 0
    for _ in range(10):
        w0 = r.body_state()[3].copy()
        r.add_body_wrench(torque=torque)
        r.step(1)
        acc = (r.body_state()[3] - w0) / 0.001
        assert np.abs(acc[:, :2]).max() < 1e-3
        for k, (c, a) in enumerate(spec):
            friction = mass * 9.8 * c * 3 * np.pi / 16 * a
            if applied <= friction:
                assert abs(acc[k, 2]) < 1e-3
            else:
                err[k] = max(err[k], abs(acc[k, 2] - (applied - friction) / izz)); moved += 1
    assert moved == 100 and err.max() < 1e-3, err


def test_free_fall_then_rest(oracle):
    """PhysicsTest.DropStuffODE (test/integration/physics.cc:919-1062): free fall follows
    v += g dt; z += v dt exactly, the box settles at z = 0.5, and stays there."""
    sc, p, _ = scenes.box_drop()
    r = oracle.RefBatch(sc, p, mode=1)
    z, v = 2.0, 0.0
    for _ in range(100):
        r.step(1)
        v += -9.8 * 0.001
        z += v * 0.001
        pos, _, lvel, _ = r.body_state()
        assert abs(pos[0, 2] - z) < 1e-4 and abs(lvel[0, 2] - v) < 1e-4   # physics.cc:970-971
    r.step(1900)
    pos, _, lvel, _ = r.body_state()
    assert abs(pos[0, 2] - 0.5) < 1e-5          # physics.cc:982-986
    assert np.linalg.norm(lvel[0]) < 3e-5       # physics.cc:1018
    assert abs(pos[0, 0]) < 1e-9 and abs(pos[0, 1]) < 1e-9   # no x/y drift


def test_natural_and_canonical_order_agree_physically(oracle):
    """Canonical (deterministic row order) and natural gzserver order are two valid PGS
    orderings of the same LCP: a settled box ends at the same height in both."""
    sc, p, _ = scenes.tilted_box()
    a = oracle.RefBatch(sc, p, mode=0)
    b = oracle.RefBatch(sc, p, mode=1)
    a.step(3000)
    b.step(3000)
    za, zb = a.body_state()[0][0, 2], b.body_state()[0][0, 2]
    assert abs(za - 0.5) < 1e-3 and abs(zb - 0.5) < 1e-3


def test_canonical_stepping_matches_dworldquickstep_on_islands(oracle):
    """The parity oracle (mode 1) restates auto-disable, island building and the per-island dxQuickStepper calls
    in a canonical order; mode 0 is the reference's own dWorldQuickStep / dxProcessIslands.  On a multi-island
    world with bodies that fall asleep one by one and one that is woken again by a late hit, both must agree at
    every checkpoint on the pair set and every body's sleep flag (contact counts within one box); the states agree
    physically (rows of an island come in a different order in the two modes: 1e-3 m before the hit, 5e-3 after).
    An error in the restated island / re-enable logic would show in the flags and counts."""
    sc, p, _ = scenes.sleepers()
    a = oracle.RefBatch(sc, p, mode=0)
    b = oracle.RefBatch(sc, p, mode=1)
    slept0 = woke0 = False
    for k in range(24):
        a.step(100)
        b.step(100)
        fa, fb = a.body_flags(), b.body_flags()
        assert np.array_equal(fa, fb), (k, fa, fb)
        assert _pairs_agree(a.pairs(0), b.pairs(0), fa)
        assert np.abs(a.stats()[2] - b.stats()[2]).max() <= 4   # (one box may touch down or lift off a step apart)
        slept0 = slept0 or bool(fa[0])
        woke0 = woke0 or (slept0 and not fa[0])
        sa, sb = a.body_state(), b.body_state()
        assert max(np.abs(x - y).max() for x, y in zip(sa[:2], sb[:2])) <= (5e-3 if woke0 else 1e-3), k
    assert slept0 and woke0 and fa[1:4].all()     # 0 slept and was woken by the hit; 1..3 sleep on


def test_auto_disable_puts_stack_to_sleep(oracle):
    """ODEPhysics.cc:208-216 + util.cpp:116-252: a settled stack of jointless bodies
    goes to sleep after idle_time = 1 s; sleeping bodies have exactly zero velocity."""
    sc, p, _ = scenes.stack()
    r = oracle.RefBatch(sc, p, mode=1)
    r.step(4000)
    flags = r.body_flags()
    _, _, lvel, avel = r.body_state()
    assert flags.all()
    assert not lvel.any() and not avel.any()


def test_collide_bitmask_filters_pairs(oracle):
    """ODEPhysics::Collide (ODEPhysics.cc:1016-1018): disjoint collide bitmasks give no contacts
    although the pair is still a broadphase candidate."""
    from gazebo_b200 import scene as S
    sb = S.SceneBuilder()
    sb.begin_world()
    sb.add_geom(S.GEOM_PLANE, (0, 0, 1, 0), sb.new_space(), collide_bitmask=0x1)
    b = sb.add_body([0, 0, 0.4], mass=1.0, inertia=S.box_inertia(1, 1, 1, 1))
    sb.add_geom(S.GEOM_BOX, (1, 1, 1), sb.new_space(), body=b, collide_bitmask=0x2)
    sb.end_world()
    sc = sb.finish()
    r = oracle.RefBatch(sc, S.default_params(), mode=1)
    r.step(1)
    assert len(r.pairs(0)) == 1
    assert len(r.contacts(0)[0]) == 0


@pytest.mark.parametrize("case", ["test_drop_stuff_free_fall_then_rest", "test_zero_max_contacts_is_harmless",
                                  "test_simple_pendulum_angle_and_energy", "test_unwrapped_hinge_angle",
                                  "test_anchored_kinematic_loop_stays_closed"])
def test_oracle_passes_the_acceptance_cases(oracle, case, monkeypatch):
    """The behavioural cases of tests/test_acceptance_gpu.py (restated from the reference's integration
    tests) run against the oracle: the harness is configured like Gazebo configures ODE."""
    from tests import test_acceptance_gpu as acc
    monkeypatch.setattr(acc, "_batch", lambda sc, p: oracle.RefBatch(sc, p, mode=1))
    getattr(acc, case)()


def test_feedback_of_a_resting_box_is_its_weight(oracle):
    """dJointFeedback of contact joints (quickstep_update_bodies.cpp:79-119): a box at rest on the
    plane is held up by its four contacts, whose feedback forces sum to m g on the box (ForceTorque-
    style acceptance: the harness hook the GPU feedback getters are compared with is itself right)."""
    sc, p, _ = scenes.box_drop()
    r = oracle.RefBatch(sc, p, mode=1)
    r.enable_feedback()
    r.step(2500)
    fb = r.contact_feedback(0)
    assert fb.shape[0] >= 3
    mass = float(np.asarray(sc.body_mass)[0]) if hasattr(sc, "body_mass") else None
    total = fb[:, 0:3].sum(axis=0)
    assert abs(total[0]) < 1e-6 and abs(total[1]) < 1e-6 and total[2] > 0
    if mass is not None:
        # 0.99 m g: the lambda solve is low-pass filtered (SMOOTH_LAMBDA, smooth_contacts = 0.01,
        # quickstep_pgs_lcp.cpp:628-640), and the feedback reports that filtered lambda
        assert abs(total[2] - 9.8 * mass) < 0.02 * 9.8 * mass


def test_feedback_of_a_hanging_chain(oracle):
    """Joint feedback of a chain hanging from the world: once it swings, the top hinge carries more
    than any other and every joint's force pair is equal and opposite (f1 = -f2)."""
    sc, p, _ = scenes.chain()
    r = oracle.RefBatch(sc, p, mode=1)
    r.enable_feedback()
    r.step(300)
    fb = r.joint_feedback()
    assert np.isfinite(fb).all() and np.abs(fb[:, 0:3]).max() > 0
    two = np.linalg.norm(fb[:, 6:9], axis=1) > 0           # joints between two links (the others hang from the world)
    assert two.sum() >= fb.shape[0] - 4
    assert np.abs(fb[two, 0:3] + fb[two, 6:9]).max() < 1e-9 * max(1.0, np.abs(fb).max())


def test_laser_oracle_known_answers(oracle):
    """gzref_ray_query (ODEMultiRayShape::UpdateRays restated over the reference's own ray colliders): closed-form cases."""
    from gazebo_b200 import scene as S
    sb = S.SceneBuilder()
    sb.begin_world()
    sb.add_ground_plane()
    b = sb.add_body((0, 0, 1.0), (1, 0, 0, 0), 1.0, S.sphere_inertia(1.0, 0.5))
    sb.add_geom(S.GEOM_SPHERE, (0.5,), sb.new_space(), body=b)
    b2 = sb.add_body((3, 0, 0.5), (1, 0, 0, 0), 1.0, S.box_inertia(1.0, 1, 1, 1))
    sb.add_geom(S.GEOM_BOX, (1, 1, 1), sb.new_space(), body=b2)
    sb.end_world()
    sc = sb.finish()
    r = oracle.RefBatch(sc, S.default_params(), mode=1)
    start = np.array([[-5, 0, 1.0], [-5, 0, 0.25], [0, 0, 5], [0, 5, 5], [0, 5, 5], [3, 0, 0.5]])
    d = np.array([[1, 0, 0], [2, 0, 0], [0, 0, -1], [0, 0, -2.0], [0, 0, 1.0], [0, 1, 0]])
    L = np.array([20.0, 20, 20, 20, 7, 20])
    rng, geom = r.ray_query(0, start, d, L)
    assert np.allclose(rng, [4.5, 7.5, 3.5, 5.0, 7.0, 0.5], atol=1e-12)   # sphere front, box front, sphere top, plane, nothing, box from inside
    assert geom.tolist() == [1, 2, 1, 0, -1, 2]
