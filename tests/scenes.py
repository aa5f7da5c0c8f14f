"""Seeded scenes shared by the golden-vector generator and the parity tests."""
import math

import numpy as np

from gazebo_b200 import scene as S, worlds as W


def box_drop():
    """test/integration/physics.cc DropStuff-style: unit box dropped on the ground plane."""
    sb = S.SceneBuilder()
    sb.begin_world()
    sb.add_ground_plane()
    b = sb.add_body([0, 0, 2.0], mass=1.0, inertia=S.box_inertia(1, 1, 1, 1))
    sb.add_geom(S.GEOM_BOX, (1, 1, 1), sb.new_space(), body=b)
    sb.end_world()
    return sb.finish(), S.default_params(), None


def tilted_box():
    sb = S.SceneBuilder()
    sb.begin_world()
    sb.add_ground_plane()
    b = sb.add_body([0, 0, 0.6], S.q_from_rpy(0.3, 0.2, 0.1), mass=1.0, inertia=S.box_inertia(1, 1, 1, 1))
    sb.add_geom(S.GEOM_BOX, (1, 1, 1), sb.new_space(), body=b)
    sb.end_world()
    return sb.finish(), S.default_params(), None


def shapes():
    """sphere / capsule / cylinder / box falling on the plane and on each other's neighbours,
    plus a sphere resting on a static box and two spheres colliding."""
    sb = S.SceneBuilder()
    rng = np.random.Generator(np.random.PCG64(77))
    sb.begin_world()
    sb.add_ground_plane()
    sb.add_geom(S.GEOM_BOX, (1.0, 1.0, 0.5), sb.new_space(), pos=(3, 0, 0.25))          # static box
    b = sb.add_body([3.1, 0.05, 1.2], mass=0.5, inertia=S.sphere_inertia(0.5, 0.2))
    sb.add_geom(S.GEOM_SPHERE, (0.2,), sb.new_space(), body=b)
    for i in range(3):
        b = sb.add_body([0.0, i * 0.05, 0.3 + 0.45 * i], mass=0.5, inertia=S.sphere_inertia(0.5, 0.2))
        sb.add_geom(S.GEOM_SPHERE, (0.2,), sb.new_space(), body=b)
    for i in range(2):
        b = sb.add_body([-2.0 + i, 1.0, 0.5], S.q_random(rng), mass=1.0, inertia=S.cylinder_inertia(1.0, 0.15, 0.5))
        sb.add_geom(S.GEOM_CAPSULE, (0.15, 0.5), sb.new_space(), body=b)
    for i in range(2):
        b = sb.add_body([-2.0 + i, -1.0, 0.6], S.q_random(rng), mass=1.0, inertia=S.cylinder_inertia(1.0, 0.2, 0.3))
        sb.add_geom(S.GEOM_CYLINDER, (0.2, 0.3), sb.new_space(), body=b)
    b = sb.add_body([1.5, 1.5, 0.25], S.q_from_axis_angle([1, 0, 0], math.pi / 2), mass=1.0,
                    inertia=S.cylinder_inertia(1.0, 0.2, 0.1))
    sb.add_geom(S.GEOM_CYLINDER, (0.2, 0.1), sb.new_space(), body=b)                     # wheel-like
    sb.end_world()
    return sb.finish(), S.default_params(), None


def chain():
    return W.chain_worlds(2, 10), S.default_params(), None


def revolve4():
    sc, pd = W.revolve_worlds(4)
    return sc, S.default_params(), pd


def boxes30():
    return W.box_drop_worlds(1, 30), S.default_params(), None


def stack():
    """box-box stacking: 4 boxes on top of each other, then a 5th falling on the edge."""
    sb = S.SceneBuilder()
    sb.begin_world()
    sb.add_ground_plane()
    for i in range(4):
        b = sb.add_body([0.01 * i, -0.01 * i, 0.25 + 0.5 * i + 0.002 * (i + 1)], mass=2.0,
                        inertia=S.box_inertia(2.0, 0.5, 0.5, 0.5), flags=S.BODY_AUTO_DISABLE)
        sb.add_geom(S.GEOM_BOX, (0.5, 0.5, 0.5), sb.new_space(), body=b)
    b = sb.add_body([0.3, 0.2, 2.6], S.q_from_rpy(0.4, 0.5, 0.6), mass=1.0,
                    inertia=S.box_inertia(1.0, 0.3, 0.4, 0.2), flags=S.BODY_AUTO_DISABLE)
    sb.add_geom(S.GEOM_BOX, (0.3, 0.4, 0.2), sb.new_space(), body=b)
    sb.end_world()
    return sb.finish(), S.default_params(), None


def capsules():
    """capsule-{sphere, box, capsule, plane}: a tight heap of capsules, spheres and boxes dropped into
    a static four-box pen, two PARALLEL capsules lying on each other (the two-contact branch of
    capsule.cpp:232-283) and a capsule falling across a static box."""
    sb = S.SceneBuilder()
    rng = np.random.Generator(np.random.PCG64(2024))
    sb.begin_world()
    sb.add_ground_plane()
    pen = sb.new_space()
    for dx, dy, sx, sy in ((0.75, 0, 0.1, 1.6), (-0.75, 0, 0.1, 1.6), (0, 0.75, 1.4, 0.1), (0, -0.75, 1.4, 0.1)):
        sb.add_geom(S.GEOM_BOX, (sx, sy, 0.8), pen, pos=(dx, dy, 0.4))
    kinds = [S.GEOM_CAPSULE, S.GEOM_SPHERE, S.GEOM_CAPSULE, S.GEOM_BOX]
    n = 0
    for iz in range(3):
        for iy in range(3):
            for ix in range(3):
                kind = kinds[n % 4]
                n += 1
                pos = [0.4 * (ix - 1) + rng.uniform(-0.03, 0.03), 0.4 * (iy - 1) + rng.uniform(-0.03, 0.03), 0.35 + 0.5 * iz]
                if kind == S.GEOM_CAPSULE:
                    b = sb.add_body(pos, S.q_random(rng), mass=0.8, inertia=S.cylinder_inertia(0.8, 0.08, 0.3))
                    sb.add_geom(kind, (0.08, 0.3), sb.new_space(), body=b)
                elif kind == S.GEOM_SPHERE:
                    b = sb.add_body(pos, mass=0.5, inertia=S.sphere_inertia(0.5, 0.15))
                    sb.add_geom(kind, (0.15,), sb.new_space(), body=b)
                else:
                    b = sb.add_body(pos, S.q_random(rng), mass=1.0, inertia=S.box_inertia(1.0, 0.25, 0.2, 0.15))
                    sb.add_geom(kind, (0.25, 0.2, 0.15), sb.new_space(), body=b)
    lie = S.q_from_axis_angle([0, 1, 0], math.pi / 2)          # axis along x
    for i, (y, z) in enumerate(((3.0, 0.1), (3.0, 0.32), (3.02, 0.56))):
        b = sb.add_body([2.0 + 0.1 * i, y, z], lie, mass=1.0, inertia=S.cylinder_inertia(1.0, 0.1, 0.6))
        sb.add_geom(S.GEOM_CAPSULE, (0.1, 0.6), sb.new_space(), body=b)
    sb.add_geom(S.GEOM_BOX, (0.4, 0.4, 0.4), sb.new_space(), pos=(-3, 0, 0.2))          # static box
    b = sb.add_body([-3.0, 0.1, 0.7], S.q_from_rpy(0.2, 1.3, 0.4), mass=1.0, inertia=S.cylinder_inertia(1.0, 0.1, 0.8))
    sb.add_geom(S.GEOM_CAPSULE, (0.1, 0.8), sb.new_space(), body=b)
    sb.end_world()
    return sb.finish(), S.default_params(), None


def joints():
    """ball / universal / slider joints: a ball-jointed pendulum chain hanging from the world, a
    universal-jointed two-link arm, a slider on the world with limits and bounce, a powered slider between
    two free bodies (linear torque decoupling), all above the plane so that links also hit the ground."""
    sb = S.SceneBuilder()
    rng = np.random.Generator(np.random.PCG64(99))
    sb.begin_world()
    sb.add_ground_plane()
    sp = sb.new_space()
    # ball chain hanging from the world
    prev = -1
    for i in range(4):
        b = sb.add_body([0.3 * i + 0.15, 0.0, 1.5], mass=0.5, inertia=S.box_inertia(0.5, 0.28, 0.06, 0.06))
        sb.add_geom(S.GEOM_BOX, (0.28, 0.06, 0.06), sp, body=b)
        sb.add_ball(b, prev, [0.3 * i, 0.0, 1.5])
        prev = b
    # universal arm: base link on a world universal joint, second link on a body-body universal joint
    sp2 = sb.new_space()
    u1 = sb.add_body([0.0, 2.0, 1.0], mass=1.0, inertia=S.box_inertia(1.0, 0.1, 0.1, 0.4))
    sb.add_geom(S.GEOM_BOX, (0.1, 0.1, 0.4), sp2, body=u1)
    sb.add_universal(u1, -1, [0.0, 2.0, 1.2], [1, 0, 0], [0, 1, 0])
    u2 = sb.add_body([0.0, 2.0, 0.6], S.q_from_rpy(0.1, 0.2, 0.0), mass=0.7, inertia=S.box_inertia(0.7, 0.1, 0.1, 0.4),
                     lvel=(0.4, 0.3, 0.0))
    sb.add_geom(S.GEOM_BOX, (0.1, 0.1, 0.4), sp2, body=u2)
    sb.add_universal(u1, u2, [0.0, 2.0, 0.8], [0, 1, 0], [1, 0, 0], erp=0.3, cfm=1e-6)
    # slider on the world along a tilted axis, limits and bounce
    s1 = sb.add_body([2.0, 0.0, 1.0], S.q_random(rng), mass=1.0, inertia=S.box_inertia(1.0, 0.2, 0.2, 0.2))
    sb.add_geom(S.GEOM_BOX, (0.2, 0.2, 0.2), sb.new_space(), body=s1)
    sb.add_slider(s1, -1, [0.3, 0.0, 1.0], lostop=-0.3, histop=0.25, bounce=0.4, stop_erp=0.3, stop_cfm=1e-5)
    # powered slider between two free bodies falling onto the plane
    sp3 = sb.new_space()
    p1 = sb.add_body([2.0, 2.0, 0.8], mass=1.0, inertia=S.box_inertia(1.0, 0.3, 0.2, 0.2))
    sb.add_geom(S.GEOM_BOX, (0.3, 0.2, 0.2), sp3, body=p1)
    p2 = sb.add_body([2.5, 2.05, 0.85], S.q_from_rpy(0.0, 0.0, 0.3), mass=0.6, inertia=S.sphere_inertia(0.6, 0.12))
    sb.add_geom(S.GEOM_SPHERE, (0.12,), sp3, body=p2)
    sb.add_slider(p1, p2, [1.0, 0.2, 0.1], lostop=-0.2, histop=0.2, vel=0.5, fmax=3.0, fudge=0.7, normal_cfm=1e-6)
    sb.end_world()
    return sb.finish(), S.default_params(), None


def cylinders():
    """cylinder-sphere (barrel, cap and rim cases) and cylinder-plane: spheres dropped onto static
    standing / lying / tilted cylinders and onto free cylinders (kept apart: cylinder-cylinder is not built)."""
    sb = S.SceneBuilder()
    rng = np.random.Generator(np.random.PCG64(5))
    sb.begin_world()
    sb.add_ground_plane()
    lie = S.q_from_axis_angle([0, 1, 0], math.pi / 2)
    poses = [((0, 0, 0.3), (1, 0, 0, 0)), ((2, 0, 0.25), lie), ((4, 0, 0.4), S.q_from_rpy(0.5, 0.3, 0.0))]
    for (pos, q) in poses:                                   # static cylinders r 0.25, l 0.6
        sb.add_geom(S.GEOM_CYLINDER, (0.25, 0.6), sb.new_space(), pos=pos, quat=q)
        for k in range(4):                                   # spheres over cap centre, rim, barrel side, far side
            off = [(0.02, 0.01), (0.24, 0.03), (0.33, -0.02), (-0.1, 0.28)][k]
            b = sb.add_body([pos[0] + off[0], pos[1] + off[1], 1.0 + 0.35 * k], mass=0.3, inertia=S.sphere_inertia(0.3, 0.1))
            sb.add_geom(S.GEOM_SPHERE, (0.1,), sb.new_space(), body=b)
    for i in range(3):                                       # free cylinders with a sphere landing on each
        q = [S.q_from_rpy(0.0, 0.0, 0.0), lie, S.q_random(rng)][i]
        b = sb.add_body([2.0 * i, 3.0, 0.5], q, mass=1.0, inertia=S.cylinder_inertia(1.0, 0.2, 0.4))
        sb.add_geom(S.GEOM_CYLINDER, (0.2, 0.4), sb.new_space(), body=b)
        b = sb.add_body([2.0 * i + 0.05, 3.02, 1.2], mass=0.4, inertia=S.sphere_inertia(0.4, 0.12))
        sb.add_geom(S.GEOM_SPHERE, (0.12,), sb.new_space(), body=b)
    sb.end_world()
    return sb.finish(), S.default_params(), None


def cylboxes():
    """cylinder-box (collision_cylinder_box.cpp): cylinders standing, lying and tilted dropped onto
    static boxes (cap-on-face = the polygon clip branch, barrel-on-face and barrel-on-edge = the
    generator-line branch), boxes dropped onto static cylinders, and a free box landing on a free
    lying cylinder.  Groups are spaced so that no two cylinders ever meet (cylinder-cylinder is MPR)."""
    sb = S.SceneBuilder()
    rng = np.random.Generator(np.random.PCG64(77))
    sb.begin_world()
    sb.add_ground_plane()
    lie = S.q_from_axis_angle([0, 1, 0], math.pi / 2)
    drops = [((1, 0, 0, 0), (0.0, 0.0)), (lie, (0.02, 0.01)), (S.q_from_rpy(0.4, 0.2, 0.1), (0.05, -0.03)),
             (S.q_from_rpy(0.0, 1.2, 0.7), (0.3, 0.28)), (S.q_random(rng), (-0.1, 0.05)), ((1, 0, 0, 0), (0.33, 0.0))]
    for i, (q, off) in enumerate(drops):                      # free cylinders onto static boxes 0.8 x 0.8 x 0.4
        x = 3.0 * i
        sb.add_geom(S.GEOM_BOX, (0.8, 0.8, 0.4), sb.new_space(), pos=(x, 0, 0.2), quat=S.q_from_rpy(0, 0, 0.3 * i))
        b = sb.add_body([x + off[0], off[1], 0.85], q, mass=1.0, inertia=S.cylinder_inertia(1.0, 0.15, 0.4))
        sb.add_geom(S.GEOM_CYLINDER, (0.15, 0.4), sb.new_space(), body=b)
    poses = [((0, 5, 0.3), (1, 0, 0, 0)), ((3, 5, 0.25), lie), ((6, 5, 0.4), S.q_from_rpy(0.5, 0.3, 0.0))]
    for (pos, q) in poses:                                    # free boxes onto static cylinders r 0.25, l 0.6
        sb.add_geom(S.GEOM_CYLINDER, (0.25, 0.6), sb.new_space(), pos=pos, quat=q)
        for k in range(2):
            side = rng.uniform(0.12, 0.3, size=3)
            off = [(0.03, 0.02), (0.2, -0.15)][k]
            b = sb.add_body([pos[0] + off[0], pos[1] + off[1], 1.0 + 0.5 * k], S.q_random(rng), mass=0.6,
                            inertia=S.box_inertia(0.6, *side))
            sb.add_geom(S.GEOM_BOX, side, sb.new_space(), body=b)
    b = sb.add_body([0.0, 10.0, 0.2], lie, mass=2.0, inertia=S.cylinder_inertia(2.0, 0.2, 0.8))   # free on free
    sb.add_geom(S.GEOM_CYLINDER, (0.2, 0.8), sb.new_space(), body=b)
    b = sb.add_body([0.05, 10.02, 0.7], S.q_from_rpy(0.1, 0.0, 0.4), mass=0.5, inertia=S.box_inertia(0.5, 0.3, 0.2, 0.1))
    sb.add_geom(S.GEOM_BOX, (0.3, 0.2, 0.1), sb.new_space(), body=b)
    sb.end_world()
    return sb.finish(), S.default_params(), None


def cylcyl():
    """cylinder-cylinder and capsule-cylinder (libccd MPR, collision_libccd.cpp:305-344): cylinders
    dropped onto static cylinders (cap on cap, barrel across barrel, tilted), capsules dropped onto
    static cylinders, and a free cylinder landing on a free lying cylinder."""
    sb = S.SceneBuilder()
    rng = np.random.Generator(np.random.PCG64(91))
    sb.begin_world()
    sb.add_ground_plane()
    lie = S.q_from_axis_angle([0, 1, 0], math.pi / 2)
    liey = S.q_from_axis_angle([1, 0, 0], math.pi / 2)
    statics = [((0, 0, 0.3), (1, 0, 0, 0)), ((3, 0, 0.25), lie), ((6, 0, 0.4), S.q_from_rpy(0.5, 0.3, 0.0))]
    drops = [((1, 0, 0, 0), (0.03, 0.02)), (liey, (0.0, 0.05)), (S.q_from_rpy(0.3, 0.9, 0.2), (0.05, -0.02))]
    for (pos, q), (dq, off) in zip(statics, drops):          # free cylinders onto static cylinders r 0.25, l 0.6
        sb.add_geom(S.GEOM_CYLINDER, (0.25, 0.6), sb.new_space(), pos=pos, quat=q)
        b = sb.add_body([pos[0] + off[0], pos[1] + off[1], 1.1], dq, mass=0.8, inertia=S.cylinder_inertia(0.8, 0.12, 0.3))
        sb.add_geom(S.GEOM_CYLINDER, (0.12, 0.3), sb.new_space(), body=b)
    for i, (pos, q) in enumerate(statics):                   # capsules onto static cylinders
        p2 = (pos[0], pos[1] + 4.0, pos[2])
        sb.add_geom(S.GEOM_CYLINDER, (0.25, 0.6), sb.new_space(), pos=p2, quat=q)
        b = sb.add_body([p2[0] + 0.04, p2[1] - 0.03, 1.2], S.q_random(rng), mass=0.5, inertia=S.cylinder_inertia(0.5, 0.08, 0.3))
        sb.add_geom(S.GEOM_CAPSULE, (0.08, 0.3), sb.new_space(), body=b)
    b = sb.add_body([0.0, 9.0, 0.2], lie, mass=2.0, inertia=S.cylinder_inertia(2.0, 0.2, 0.8))      # free on free
    sb.add_geom(S.GEOM_CYLINDER, (0.2, 0.8), sb.new_space(), body=b)
    b = sb.add_body([0.03, 9.02, 0.75], liey, mass=0.5, inertia=S.cylinder_inertia(0.5, 0.1, 0.5))
    sb.add_geom(S.GEOM_CYLINDER, (0.1, 0.5), sb.new_space(), body=b)
    sb.end_world()
    return sb.finish(), S.default_params(), None


def sleepers():
    """Multi-island world for the island / auto-disable / re-enable logic (util.cpp:116-252, 408-558): six
    separated auto-disable boxes land at different times and go to sleep one by one; a late box then falls onto
    sleeper 0 and wakes it (a two-body island), the others stay asleep.  Until that hit every island is one box
    on the plane, whose rows come in the same order whatever the island order: natural (dWorldQuickStep) and
    canonical stepping agree to rounding there."""
    sb = S.SceneBuilder()
    sb.begin_world()
    sb.add_ground_plane()
    for i in range(6):
        b = sb.add_body([1.5 * i, 0, 0.3 + 0.25 * i], S.q_from_rpy(0.02 * i, 0.01, 0.3 * i), mass=1.0,
                        inertia=S.box_inertia(1, 0.4, 0.4, 0.4), flags=S.BODY_AUTO_DISABLE)
        sb.add_geom(S.GEOM_BOX, (0.4, 0.4, 0.4), sb.new_space(), body=b)
    b = sb.add_body([0.05, 0.02, 14.0], mass=0.5, inertia=S.box_inertia(0.5, 0.3, 0.3, 0.3), flags=S.BODY_AUTO_DISABLE)
    sb.add_geom(S.GEOM_BOX, (0.3, 0.3, 0.3), sb.new_space(), body=b)
    sb.end_world()
    return sb.finish(), S.default_params(), None


def dyn_inertia():
    """DYNAMIC_INERTIA (quickstep_util.cpp:170-524): a heavy base carrying thin, light arms on hinges - inertia ratios
    far beyond 200 along the hinges' constrained axes - one arm carrying a second one (a body shared by two hinges is
    adjusted twice, in joint order), plus a pair of similar bodies that must be left alone.  Standing on the plane."""
    sb = S.SceneBuilder()
    p = S.default_params(dynamic_inertia=1)
    for w in range(2):
        sb.begin_world()
        sb.add_ground_plane()
        sp = sb.new_space()
        base = sb.add_body([0, 0, 0.25], S.q_from_rpy(0, 0, 0.3 * w), mass=20.0, inertia=S.box_inertia(20.0, 0.6, 0.6, 0.5))
        sb.add_geom(S.GEOM_BOX, (0.6, 0.6, 0.5), sp, body=base)
        a1 = sb.add_body([0.5, 0, 0.55], S.q_from_rpy(0, 0.4, 0), mass=0.01, inertia=S.box_inertia(0.01, 0.4, 0.02, 0.02))
        sb.add_geom(S.GEOM_BOX, (0.4, 0.02, 0.02), sp, body=a1)
        sb.add_hinge(a1, base, [0.3, 0, 0.5], [0, 1, 0])
        a2 = sb.add_body([0.9, 0, 0.55], S.q_from_rpy(0.2, 0, 0), mass=0.004, inertia=S.box_inertia(0.004, 0.3, 0.01, 0.01))
        sb.add_geom(S.GEOM_BOX, (0.3, 0.01, 0.01), sp, body=a2)
        sb.add_hinge(a2, a1, [0.72, 0, 0.55], [0, 0, 1])
        c1 = sb.add_body([-0.6, 0.5 * w, 0.7], mass=1.0, inertia=S.box_inertia(1.0, 0.2, 0.2, 0.2))
        sb.add_geom(S.GEOM_BOX, (0.2, 0.2, 0.2), sp, body=c1)
        sb.add_hinge(c1, base, [-0.3, 0, 0.6], [1, 0, 0])
        sb.end_world()
    return sb.finish(), p, None


def fdir1():
    """ODEPhysics.cc:1120-1161 as written: only collision 1's friction direction reaches the contact.  Anisotropic
    friction (mu != mu2) and boxes sliding diagonally, with fdir1 on geom 1 only (the plane; a rotated static
    platform), on geom 2 only (no effect), and on both with differing mu (geom 1's wins whatever the mu)."""
    sb = S.SceneBuilder()
    sb.begin_world()
    c, s_ = math.cos(0.4), math.sin(0.4)
    sb.add_geom(S.GEOM_PLANE, (0, 0, 1, 0), sb.new_space(), mu=1.0, mu2=0.15, fdir1=(c, s_, 0))   # geom 0: collision 1 of every plane pair
    def slider(x, y, z, **kw):
        b = sb.add_body([x, y, z], S.q_from_rpy(0, 0, 0.2), mass=1.0, inertia=S.box_inertia(1, 0.3, 0.3, 0.2), lvel=(0.7, 0.5, 0.0))
        sb.add_geom(S.GEOM_BOX, (0.3, 0.3, 0.2), sb.new_space(), body=b, **kw)
        return b
    slider(0.0, 0.0, 0.1, mu=0.9, mu2=0.9)                                   # on the plane: the plane's fdir1
    # platform without fdir1 (geom 1), slider with fdir1 (geom 2): no effect
    sb.add_geom(S.GEOM_BOX, (3, 3, 0.5), sb.new_space(), pos=(5, 0, 0.25), mu=0.8, mu2=0.8)
    slider(5.0, 0.0, 0.6, mu=0.6, mu2=0.1, fdir1=(0, 1, 0))
    # rotated platform with fdir1 (geom 1), slider without
    sb.add_geom(S.GEOM_BOX, (3, 3, 0.5), sb.new_space(), pos=(10, 0, 0.25), quat=S.q_from_rpy(0, 0, 0.5), mu=0.7, mu2=0.1, fdir1=(1, 0, 0))
    slider(10.0, 0.0, 0.6, mu=0.9, mu2=0.9)
    # both, geom 2 with the SMALLER mu: geom 1's direction is used all the same
    sb.add_geom(S.GEOM_BOX, (3, 3, 0.5), sb.new_space(), pos=(15, 0, 0.25), quat=S.q_from_rpy(0, 0, -0.3), mu=0.8, mu2=0.2, fdir1=(0, 1, 0))
    slider(15.0, 0.0, 0.6, mu=0.3, mu2=0.05, fdir1=(1, 0, 0))
    sb.end_world()
    return sb.finish(), S.default_params(), None


def carts(n_worlds=3, seed=31):
    """Hinge2 joints (dxJointHinge2, SDF revolute2) as on a car: a chassis box on four sphere wheels; axis 1 = vertical
    (steering, limited to +-0.3..0.5 rad on the front wheels, locked by equal stops on the rear ones, one front wheel with
    a steering motor), axis 2 = the axle (driven by a motor on the rear wheels, free on the front ones), the ball row
    along axis 1 soft (suspension erp / cfm).  Carts start at different yaw, one tilted, and drive over the plane."""
    sb = S.SceneBuilder()
    rng = np.random.Generator(np.random.PCG64(seed))
    for w in range(n_worlds):
        sb.begin_world()
        sb.add_ground_plane()
        yaw = rng.uniform(-1.0, 1.0)
        q = S.q_from_rpy(0.05 * w, -0.03 * w, yaw)
        R = S.q_to_R(q)
        c = np.array([0.0, 0.0, 0.45])
        chassis = sb.add_body(c, q, 8.0, S.box_inertia(8.0, 1.2, 0.6, 0.2))
        sb.add_geom(S.GEOM_BOX, (1.2, 0.6, 0.2), sb.new_space(), body=chassis)
        space = sb.new_space()
        for k, (dx, dy) in enumerate(((0.5, 0.4), (0.5, -0.4), (-0.5, 0.4), (-0.5, -0.4))):
            pw = c + R @ np.array([dx, dy, -0.15])
            wheel = sb.add_body(pw, q, 1.0, S.sphere_inertia(1.0, 0.2))
            sb.add_geom(S.GEOM_SPHERE, (0.2,), sb.new_space(), body=wheel)
            front = dx > 0
            kw = dict(susp_erp=0.4, susp_cfm=0.02 + 0.01 * k)
            if front:
                kw.update(lostop=-0.3 - 0.05 * k, histop=0.5, stop_erp=0.3, stop_cfm=1e-4)
                if k == 0:
                    kw.update(vel=0.6, fmax=2.0, fudge=0.5)       # steers into its stop
            else:
                kw.update(lostop=0.0, histop=0.0, vel2=4.0 + w, fmax2=3.0)   # locked steering, driven axle
            sb.add_hinge2(chassis, wheel, pw, R @ np.array([0, 0, 1.0]), R @ np.array([0, 1.0, 0]), **kw)
        sb.end_world()
    return sb.finish(), S.default_params(), None


def screws(n_worlds=2, seed=71):
    """Screw joints (dxJointScrew, SDF screw): a nut on a vertical rod fixed in the world (thread pitch 30 rad / m: it
    spins down under gravity until it lands on the plane), a nut on a tilted rod with a fine pitch (0.8 rad / m, the other
    branch of screw.cpp:170-177) and stops on the rotation with bounce, and two free bodies screwed together that tumble
    onto the plane."""
    sb = S.SceneBuilder()
    for w in range(n_worlds):
        rng = np.random.Generator(np.random.PCG64(seed + w))
        sb.begin_world()
        sb.add_ground_plane()
        a = sb.add_body([0.0, 0.0, 1.2], mass=1.0, inertia=S.box_inertia(1.0, 0.3, 0.3, 0.1), avel=(0, 0, 1.0 + w))
        sb.add_geom(S.GEOM_BOX, (0.3, 0.3, 0.1), sb.new_space(), body=a)
        sb.add_screw(a, -1, [0.0, 0.0, 1.2], [0, 0, 1], 30.0)
        b = sb.add_body([1.5, 0.0, 1.0], S.q_from_rpy(0.3, 0.2, 0.0), mass=0.8, inertia=S.box_inertia(0.8, 0.2, 0.2, 0.2), avel=tuple(rng.uniform(-2, 2, 3)))
        sb.add_geom(S.GEOM_BOX, (0.2, 0.2, 0.2), sb.new_space(), body=b)
        sb.add_screw(b, -1, [1.5, 0.0, 1.0], [0.3, 0.1, 1.0], 0.8, lostop=-0.6, histop=0.5, bounce=0.3, stop_erp=0.3, stop_cfm=1e-5)
        sp = sb.new_space()
        c = sb.add_body([0.0, 2.0, 0.9], S.q_random(rng), mass=1.2, inertia=S.box_inertia(1.2, 0.4, 0.15, 0.15), lvel=(0.3, 0, 0))
        sb.add_geom(S.GEOM_BOX, (0.4, 0.15, 0.15), sp, body=c)
        d = sb.add_body([0.3, 2.0, 0.9], S.q_random(rng), mass=0.7, inertia=S.sphere_inertia(0.7, 0.1), avel=(3.0, 0, 0))
        sb.add_geom(S.GEOM_SPHERE, (0.1,), sp, body=d)
        sb.add_screw(c, d, [0.15, 2.0, 0.9], [1, 0, 0], 12.0, erp=0.3, cfm=1e-6)
        sb.end_world()
    return sb.finish(), S.default_params(), None


def gears(n_worlds=2, seed=67):
    """Gearbox joints (dxJointGearbox, SDF gearbox): world 0 - two wheels hinged to the world on parallel axes, coupled 2 : 1,
    one driven by its hinge motor; world 1 - the same pair on a free carrier body (the gearbox's reference body) that
    falls onto the plane and tumbles, ratio -1.5, axes not parallel."""
    sb = S.SceneBuilder()
    for w in range(n_worlds):
        rng = np.random.Generator(np.random.PCG64(seed + w))
        sb.begin_world()
        sb.add_ground_plane()
        sp = sb.new_space()
        if w % 2 == 0:
            a = sb.add_body([0.0, 0.0, 1.0], S.q_from_rpy(1.5707963267948966, 0, 0), mass=1.0, inertia=S.cylinder_inertia(1.0, 0.2, 0.1))
            sb.add_geom(S.GEOM_CYLINDER, (0.2, 0.1), sp, body=a)
            b = sb.add_body([0.5, 0.0, 1.0], S.q_from_rpy(1.5707963267948966, 0, 0), mass=0.5, inertia=S.cylinder_inertia(0.5, 0.1, 0.1),
                            avel=(0, 0.3, 0))
            sb.add_geom(S.GEOM_CYLINDER, (0.1, 0.1), sp, body=b)
            sb.add_hinge(a, -1, [0.0, 0.0, 1.0], [0, 1, 0], vel=2.0, fmax=0.4)
            sb.add_hinge(b, -1, [0.5, 0.0, 1.0], [0, 1, 0])
            sb.add_gearbox(a, b, [0, 1, 0], [0, 1, 0], 2.0)
        else:
            c = sb.add_body([0.0, 0.0, 0.8], S.q_from_rpy(0.2, -0.1, 0.3), mass=2.0, inertia=S.box_inertia(2.0, 0.8, 0.2, 0.1),
                            avel=tuple(rng.uniform(-1, 1, 3)))
            sb.add_geom(S.GEOM_BOX, (0.8, 0.2, 0.1), sp, body=c)
            R = S.q_to_R(S.q_from_rpy(0.2, -0.1, 0.3))
            pa, pb = np.array([0, 0, 0.8]) + R @ np.array([-0.25, 0, 0.15]), np.array([0, 0, 0.8]) + R @ np.array([0.25, 0, 0.15])
            a = sb.add_body(pa, S.q_from_rpy(0.2, -0.1, 0.3), mass=0.6, inertia=S.sphere_inertia(0.6, 0.12), avel=(0, 0, 4.0))
            sb.add_geom(S.GEOM_SPHERE, (0.12,), sp, body=a)
            b = sb.add_body(pb, S.q_from_rpy(0.2, -0.1, 0.3), mass=0.4, inertia=S.sphere_inertia(0.4, 0.1))
            sb.add_geom(S.GEOM_SPHERE, (0.1,), sp, body=b)
            sb.add_hinge(c, a, pa, R @ np.array([0, 0, 1.0]))
            sb.add_hinge(c, b, pb, R @ np.array([0, 0.6, 0.8]))
            sb.add_gearbox(a, b, R @ np.array([0, 0, 1.0]), R @ np.array([0, 0.6, 0.8]), -1.5, ref1=c, ref2=c)
        sb.end_world()
    return sb.finish(), S.default_params(), None


def gimbals(n_worlds=2, seed=61):
    """Universal joints with stops and motors on both axes (dxJointUniversal limot1 / limot2, universal.cpp:268-389):
    an arm on the world swinging into the stops of both axes (bounce on one), a two-link arm whose second joint has a
    motor on axis 1 driving into a stop (fudge factor) and stops on axis 2, a link whose axis 2 is locked by equal stops
    and powered on axis 1; links fall onto the plane as well."""
    sb = S.SceneBuilder()
    for w in range(n_worlds):
        rng = np.random.Generator(np.random.PCG64(seed + w))
        sb.begin_world()
        sb.add_ground_plane()
        sp = sb.new_space()
        a = sb.add_body([0.0, 0.0, 1.0], mass=1.0, inertia=S.box_inertia(1.0, 0.1, 0.1, 0.5), avel=tuple(rng.uniform(-3, 3, 3)))
        sb.add_geom(S.GEOM_BOX, (0.1, 0.1, 0.5), sp, body=a)
        sb.add_universal(a, -1, [0.0, 0.0, 1.25], [1, 0, 0], [0, 1, 0], lostop=-0.4, histop=0.3, stop_erp=0.3, stop_cfm=1e-5,
                         lostop2=-0.25, histop2=0.35, bounce2=0.5, stop_erp2=0.25)
        b = sb.add_body([0.0, 0.0, 0.5], S.q_from_rpy(0.05, -0.1, 0.0), mass=0.6, inertia=S.box_inertia(0.6, 0.1, 0.1, 0.4),
                        lvel=(0.5, -0.4, 0.0))
        sb.add_geom(S.GEOM_BOX, (0.1, 0.1, 0.4), sp, body=b)
        sb.add_universal(a, b, [0.0, 0.0, 0.75], [0, 1, 0], [1, 0, 0], vel=1.5, fmax=0.8, fudge=0.4, lostop=-0.5, histop=0.2,
                         lostop2=-0.3, histop2=0.3, erp=0.3, cfm=1e-6)
        sp2 = sb.new_space()
        c = sb.add_body([1.5, 0.0, 0.9], mass=0.8, inertia=S.box_inertia(0.8, 0.4, 0.1, 0.1))
        sb.add_geom(S.GEOM_BOX, (0.4, 0.1, 0.1), sp2, body=c)
        d = sb.add_body([1.9, 0.0, 0.9], S.q_from_rpy(0.0, 0.0, 0.2 * (w + 1)), mass=0.5, inertia=S.box_inertia(0.5, 0.4, 0.1, 0.1))
        sb.add_geom(S.GEOM_BOX, (0.4, 0.1, 0.1), sp2, body=d)
        sb.add_universal(c, d, [1.7, 0.0, 0.9], [0, 0, 1], [0, 1, 0], vel=-2.0, fmax=0.5, lostop2=0.0, histop2=0.0, vel2=0.7, fmax2=0.3,
                         normal_cfm=1e-6, normal_cfm2=1e-6)
        sb.end_world()
    return sb.finish(), S.default_params(), None


def hertz(n_worlds=2, seed=53):
    """Hertzian contact (FrictionPyramid elastic modulus, ODEPhysics.cc:1237-1251 / contact.cpp:163-216): soft spheres,
    boxes and capsules (E = 1e5..1e7 Pa, Poisson's ratio 0.2..0.45) dropped on a soft plane and on each other; half of
    them use the patch radius (0.02..0.1 m), the others the surface radius (their own radius of curvature, combined as
    1/R = 1/R1 + 1/R2), a few have no modulus at all (the pair then keeps the kp / kd contact)."""
    sb = S.SceneBuilder()
    for w in range(n_worlds):
        rng = np.random.Generator(np.random.PCG64(seed + w))
        sb.begin_world()
        # the plane: soft, flat (surface radius = infinity: Gazebo's default DBL_MAX), patch radius 0.05
        sb.add_geom(S.GEOM_PLANE, (0, 0, 1, 0), sb.new_space(), mu=100.0, mu2=50.0, elastic_modulus=5e6, poissons_ratio=0.3,
                    patch_radius=0.05, use_patch_radius=float(w % 2))
        for k in range(14):
            pos = np.array([rng.uniform(-0.6, 0.6), rng.uniform(-0.6, 0.6), 0.4 + 0.45 * k])
            q = S.q_random(rng)
            em = 0.0 if k % 5 == 4 else 10 ** rng.uniform(5, 7)
            use_patch = float((k + w) % 2)
            kw = dict(elastic_modulus=em, poissons_ratio=rng.uniform(0.2, 0.45), patch_radius=rng.uniform(0.02, 0.1),
                      use_patch_radius=use_patch, kp=1e6, kd=10.0)
            t = k % 3
            if t == 0:
                r = rng.uniform(0.1, 0.25); b = sb.add_body(pos, q, 1.0, S.sphere_inertia(1.0, r))
                sb.add_geom(S.GEOM_SPHERE, (r,), sb.new_space(), body=b, surface_radius=r, **kw)
            elif t == 1:
                d = tuple(rng.uniform(0.15, 0.4, 3)); b = sb.add_body(pos, q, 1.0, S.box_inertia(1.0, *d))
                sb.add_geom(S.GEOM_BOX, d, sb.new_space(), body=b, **kw)      # flat faces: the default surface radius
            else:
                r, l = rng.uniform(0.08, 0.15), rng.uniform(0.2, 0.5); b = sb.add_body(pos, q, 1.0, S.cylinder_inertia(1.0, r, l))
                sb.add_geom(S.GEOM_CAPSULE, (r, l), sb.new_space(), body=b, surface_radius=r, **kw)
        sb.end_world()
    return sb.finish(), S.default_params(), None


def spinners(n_worlds=3, seed=91):
    """Torsional friction (dContactMu3: a fourth contact row, contact.cpp:362-437; switched on per pair as in
    ODEPhysics.cc:1168-1232): spheres, upright cylinders, capsules and boxes set down spinning about the vertical on a
    plane with mu3 = 0.8.  Two thirds use the patch radius (0.02..0.12 m), the others sqrt(surface radius x depth);
    every fifth body has mu3 = 0 (its contacts keep three rows among four-row contacts), every fourth a torsional slip.
    A few bodies are stacked so that torsional rows join two dynamic bodies."""
    sb = S.SceneBuilder()
    for w in range(n_worlds):
        rng = np.random.Generator(np.random.PCG64(seed + w))
        sb.begin_world()
        sb.add_geom(S.GEOM_PLANE, (0, 0, 1, 0), sb.new_space(), mu=1.0, mu2=1.0, mu3=0.8, patch_radius=0.05,
                    use_patch_radius=1.0)
        for k in range(12):
            col, lvl = k % 6, k // 6
            pos = np.array([1.2 * (col - 2.5), 0.0, 0.0])
            spin = rng.uniform(4.0, 15.0) * (1 if k % 2 else -1)
            kw = dict(mu3=0.0 if k % 5 == 4 else rng.uniform(0.2, 1.5), patch_radius=rng.uniform(0.02, 0.12),
                      use_patch_radius=0.0 if (k + w) % 3 == 2 else 1.0, slip3=0.05 if k % 4 == 3 else 0.0, kp=1e6, kd=100.0)
            t = k % 4
            if t == 0:
                r = rng.uniform(0.15, 0.3); pos[2] = r + 0.9 * lvl + 0.002
                b = sb.add_body(pos, (1, 0, 0, 0), 1.0, S.sphere_inertia(1.0, r), avel=(0, 0, spin))
                sb.add_geom(S.GEOM_SPHERE, (r,), sb.new_space(), body=b, surface_radius=r, **kw)
            elif t == 1:
                r, l = rng.uniform(0.15, 0.3), rng.uniform(0.2, 0.4); pos[2] = 0.5 * l + 0.9 * lvl + 0.002
                b = sb.add_body(pos, (1, 0, 0, 0), 1.0, S.cylinder_inertia(1.0, r, l), avel=(0, 0, spin))
                sb.add_geom(S.GEOM_CYLINDER, (r, l), sb.new_space(), body=b, surface_radius=5.0, **kw)
            elif t == 2:
                d = tuple(rng.uniform(0.3, 0.5, 3)); pos[2] = 0.5 * d[2] + 0.9 * lvl + 0.002
                b = sb.add_body(pos, (1, 0, 0, 0), 1.0, S.box_inertia(1.0, *d), avel=(0, 0, spin))
                sb.add_geom(S.GEOM_BOX, d, sb.new_space(), body=b, surface_radius=2.0, **kw)
            else:
                r, l = rng.uniform(0.12, 0.2), rng.uniform(0.2, 0.4); pos[2] = 0.5 * l + r + 0.9 * lvl + 0.002
                b = sb.add_body(pos, (1, 0, 0, 0), 1.0, S.cylinder_inertia(1.0, r, l), avel=(0.3, 0, spin))
                sb.add_geom(S.GEOM_CAPSULE, (r, l), sb.new_space(), body=b, surface_radius=r, **kw)
        sb.end_world()
    return sb.finish(), S.default_params(), None


def torsion_spheres():
    """test/worlds/torsional_friction_test.world restated (the reference's PhysicsTorsionalFrictionTest.CoefficientTest,
    test/integration/physics_torsional_friction.cc:283-381): ten 0.1 m spheres of 11.31 kg (izz 0.04524) with kp = 5e5 on the
    stock ground plane (MuTorsion 1, no patch radius of its own): five with patch radius 2 m and coefficients 0 .. 0.1,
    five with coefficient 0.005 and patch radii 0 .. 1 m.  Returns the scene, the parameters and (coefficient, patch)."""
    sb = S.SceneBuilder()
    sb.begin_world()
    sb.add_geom(S.GEOM_PLANE, (0, 0, 1, 0), sb.new_space(), mu=100.0, mu2=50.0, mu3=1.0)
    spec = [(c, 2.0) for c in (0.0, 0.001, 0.005, 0.01, 0.1)] + [(0.005, a) for a in (0.0, 0.01, 0.1, 0.5, 1.0)]
    mass, izz = 11.309733552923259, 0.045238934211693047
    for k, (c, a) in enumerate(spec):
        b = sb.add_body([0.4 * (k % 5 + 1), 1.0 + k // 5, 0.1], mass=mass, inertia=(izz, izz, izz, 0, 0, 0))   # (allow_auto_disable 0)
        sb.add_geom(S.GEOM_SPHERE, (0.1,), sb.new_space(), body=b, mu3=c, patch_radius=a, kp=5e5)
    sb.end_world()
    return sb.finish(), S.default_params(), spec


def friction_demo():
    """worlds/friction_demo.world restated (PhysicsFrictionTest.FrictionDemo, test/integration/physics_friction.cc:183-286):
    gravity (0, -1, -1); six 1 kg unit boxes (inertia 10, centre of mass on the bottom face so that they cannot tip) with
    mu = 0.998 .. 1.01 on a plane with mu = 50.  Returns the scene, the parameters and the friction coefficients."""
    sb = S.SceneBuilder()
    sb.begin_world()
    sb.add_geom(S.GEOM_PLANE, (0, 0, 1, 0), sb.new_space(), mu=50.0, mu2=50.0, kp=1e9, kd=1.0)
    mus = (0.998, 0.9985, 0.999, 0.9995, 1.0, 1.01)
    for k, mu in enumerate(mus):
        b = sb.add_body([2.0 * (k + 1), 8.0, 0.0], mass=1.0, inertia=(10, 10, 10, 0, 0, 0), max_vel=100.0, min_depth=0.001)
        sb.add_geom(S.GEOM_BOX, (1, 1, 1), sb.new_space(), body=b, pos=(0, 0, 0.5), mu=mu, mu2=mu, kp=1e5, kd=1.0)
    sb.end_world()
    return sb.finish(), S.default_params(gravity=(0.0, -1.0, -1.0)), mus


def check_friction_demo(r, mus):
    """physics_friction.cc:239-281: every 500 steps for 10 s, v_x = v_z = 0 and v_y = (g_y + mu) t below mu = 1, 0 from
    there on, all within g_friction_tolerance = 1e-3."""
    for k in range(20):
        r.step(500)
        t = 0.5 * (k + 1)
        v = r.body_state()[2]
        assert np.abs(v[:, 0]).max() < 1e-3 and np.abs(v[:, 2]).max() < 1e-3
        for i, mu in enumerate(mus):
            assert abs(v[i, 1] - (0.0 if mu >= 1.0 else (-1.0 + mu) * t)) < 1e-3, (t, mu, v[i, 1])


def friction_spheres():
    """test/worlds/friction_spheres.world.erb restated (PhysicsFrictionTest.SphereSlip,
    test/integration/physics_friction.cc:529-666): gravity (0, 2, -9.81); links of one ("lowball": centre of mass hanging
    0.3 m below the ball so that it slides), two and three collision spheres of radius 0.2, and boxes of edge 0.4, each
    in six mass / slip combinations (mass x slip = 0.2, 0.4, 0.6 twice), mu = 0.5, on the stock ground plane.
    Returns the scene, the parameters and per body (mass x slip, number of collision shapes sharing the load)."""
    sb = S.SceneBuilder()
    sb.begin_world()
    sb.add_ground_plane()
    rad, alt = 0.2, 0.5
    side = alt / (np.sqrt(3) / 2.0)
    kinds = [("lowball", [(0, 0, 0)], (0, 0, -0.3)),
             ("twoball", [(0, side / 2, 0), (0, -side / 2, 0)], (0, 0, 0)),
             ("triball", [(2 / 3 * alt, 0, 0), (-1 / 3 * alt, side / 2, 0), (-1 / 3 * alt, -side / 2, 0)], (0, 0, 0))]
    mass_slip = [(1.0, 0.2), (2.0, 0.1), (2.0, 0.2), (4.0, 0.1), (3.0, 0.2), (6.0, 0.1)]
    expect = []
    for i, (_, balls, off) in enumerate(kinds):
        n = len(balls)
        for ii, (mass, slip) in enumerate(mass_slip):
            bm = mass / n
            I = np.diag([n * 0.4 * bm * rad ** 2] * 3)
            for (dx, dy, dz) in balls:   # (the balls' centre is the origin for all three kinds)
                I += bm * np.array([[dy * dy + dz * dz, -dx * dy, -dx * dz], [-dx * dy, dz * dz + dx * dx, -dy * dz],
                                    [-dx * dz, -dy * dz, dx * dx + dy * dy]])
            b = sb.add_body([2.0 * ii + off[0], -2.0 * i + off[1], rad + off[2]], mass=mass,
                            inertia=(I[0, 0], I[1, 1], I[2, 2], I[0, 1], I[0, 2], I[1, 2]))
            sp = sb.new_space()
            for (x, y, z) in balls:
                sb.add_geom(S.GEOM_SPHERE, (rad,), sp, body=b, pos=(x - off[0], y - off[1], z - off[2]), mu=0.5, mu2=0.5,
                            slip1=slip, slip2=slip)
            expect.append((mass * slip, float(n)))
    for ii, (mass, slip) in enumerate(mass_slip):
        ixx = 1.0 / 12.0 * mass * 2 * (2 * rad) ** 2
        b = sb.add_body([2.0 * (ii + 0.5), 0.0, rad], mass=mass, inertia=(ixx, ixx, ixx, 0, 0, 0))
        sb.add_geom(S.GEOM_BOX, (2 * rad,) * 3, sb.new_space(), body=b, mu=0.5, mu2=0.5, slip1=slip, slip2=slip)
        expect.append((mass * slip, 1.0))   # (the slip of a pair is multiplied by its number of contacts, ODEPhysics.cc:1175-1185)
    sb.end_world()
    return sb.finish(), S.default_params(gravity=(0.0, 2.0, -9.81)), expect


def check_friction_spheres(r, expect):
    """physics_friction.cc:604-665: after 5 000 steps v_x = v_z = 0 (1e-3) and v_y = g_y x mass x slip / shapes within 1.5 %."""
    r.step(5000)
    v = r.body_state()[2]
    assert np.abs(v[:, 0]).max() < 1e-3 and np.abs(v[:, 2]).max() < 1e-3
    for k, (ms, n) in enumerate(expect):
        ve = 2.0 * ms / n
        assert abs(v[k, 1] - ve) < 0.015 * ve, (k, v[k, 1], ve)


def gearbox_world():
    """test/worlds/gearbox.world restated (ODEGearboxJoint_TEST.GearboxTest, test/integration/joint_gearbox.cc:36-112): no
    gravity, 200 iterations, sor 1, erp 1; link_2 on a hinge to the world, link_1 and link_3 on hinges (about x) to link_2, a
    gearbox of ratio -1.5 between link_1 (parent) and link_3 (child: ODE's body 1) with link_2 as reference body.
    Returns the scene, the parameters and (link_1, link_2, link_3, joint_12, joint_23)."""
    sb = S.SceneBuilder()
    sb.begin_world()
    q = S.q_from_rpy(-1.57079, 0, 0)
    R, T = S.q_to_R(q), np.array([0, 0, 2.1])
    sp = sb.new_space()
    I = (1, 1, 1, 0, 0, 0)
    l1 = sb.add_body(T + R @ np.array([0, 0, -0.5]), q, mass=10.0, inertia=I)          # (bodies sit at the centres of mass)
    sb.add_geom(S.GEOM_CYLINDER, (0.1, 1.0), sp, body=l1)
    l2 = sb.add_body(T + R @ np.array([0.5, 0, -0.5]), q, mass=10.0, inertia=I)
    sb.add_geom(S.GEOM_BOX, (0.2, 0.2, 0.2), sp, body=l2, pos=(0, 0, 0.5))
    l3 = sb.add_body(T + R @ np.array([1, 1, 0.5]), q, mass=10.0, inertia=I)
    sb.add_geom(S.GEOM_CYLINDER, (0.1, 1.0), sp, body=l3)
    x = np.array([1.0, 0, 0])
    sb.add_hinge(l2, -1, T + R @ np.array([0.5, 0, 0]), R @ np.array([0, 0, 1.0]))
    j12 = sb.add_hinge(l1, l2, T, x)
    j23 = sb.add_hinge(l3, l2, T + R @ np.array([1, 1, 1.0]), x)
    sb.add_gearbox(l3, l1, x, x, -1.5, ref1=l2, ref2=l2)
    sb.end_world()
    p = S.default_params(gravity=(0, 0, 0), iters=200, sor_w=1.0, cfm=0.0, erp=1.0)
    return sb.finish(), p, (l1, l2, l3, j12, j23)


def check_gearbox(r, sc, ids, steps=2000):
    """joint_gearbox.cc:52-110 for the joint_02 angle 0 (2 000 of its 10 000 steps per phase): hinge torques in the
    gearbox's own proportion (force1 = ratio x force3) hold everything still to 1e-6; with -1 / +1 both joints turn forwards
    and velocity1 x ratio = -velocity3, position1 x ratio = -position3 to 1e-6."""
    l1, l2, l3, j12, j23 = ids
    x, ratio = np.array([1.0, 0, 0]), -1.5

    def drive(f1, f3):   # Joint::SetForce = dJointAddHingeTorque: +f on the child, -f on the parent, about the axis
        for _ in range(steps):
            tq = np.zeros((sc.nb, 3))
            tq[l1] += f1 * x; tq[l2] -= f1 * x
            tq[l3] += f3 * x; tq[l2] -= f3 * x
            r.add_body_wrench(torque=tq)
            r.step(1)
        av, cum = r.body_state()[3], r.joint_state()[2]
        return (av[l1] - av[l2]) @ x, (av[l3] - av[l2]) @ x, cum[j12], cum[j23]
    v1, v3, p1, p3 = drive(1.0 * ratio, 1.0)
    assert max(abs(v1), abs(v3), abs(p1), abs(p3)) < 1e-6
    v1, v3, p1, p3 = drive(-1.0, 1.0)
    assert min(v1, v3, p1, p3) > 0
    assert abs(v1 * ratio + v3) < 1e-6 and abs(p1 * ratio + p3) < 1e-6
    return v1, v3, p1, p3


def depth_spheres():
    """The sphere_mass_1..5 models of torsional_friction_test.world (PhysicsTorsionalFrictionTest.DepthTest,
    physics_torsional_friction.cc:147-281): 0.1 m spheres of 0.42 .. 41.9 kg, kp = 5e5, max_vel 10 000.  Returns the
    scene, the parameters and the masses."""
    sb = S.SceneBuilder()
    sb.begin_world()
    sb.add_geom(S.GEOM_PLANE, (0, 0, 1, 0), sb.new_space(), mu=100.0, mu2=50.0, mu3=1.0)
    masses = (0.41887902047863923, 4.188790204786391, 11.309733552923259, 20.94395102393196, 41.88790204786392)
    izzs = (0.0016755160819145573, 0.01675516081914557, 0.045238934211693047, 0.08377580409572787, 0.16755160819145573)
    for k, (m, izz) in enumerate(zip(masses, izzs)):
        b = sb.add_body([0.4 * (k + 1), 0.0, 0.1], mass=m, inertia=(izz, izz, izz, 0, 0, 0), max_vel=10000.0)
        sb.add_geom(S.GEOM_SPHERE, (0.1,), sb.new_space(), body=b, mu3=0.005, kp=5e5)
    sb.end_world()
    return sb.finish(), S.default_params(), masses


def check_depth_spheres(r, masses):
    """physics_torsional_friction.cc:226-273: after 1 000 steps every sphere's contact has the normal +z exactly and the
    depth m g / kp within 1 %."""
    r.step(1000)
    cg, cd = r.contacts(0)
    assert len(cg) == len(masses)
    for (g1, g2), c in zip(cg, cd):
        m = masses[max(g1, g2) - 1]
        assert c[3] == 0.0 and c[4] == 0.0 and abs(c[5]) == 1.0   # (exactly along z; its sign follows the pair's geom order)
        expected = m * 9.8 / 500000.0
        assert abs(c[6] - expected) / expected < 0.01, (m, c[6], expected)


def torsion_radius_spheres():
    """The sphere_radius_1..5 models of torsional_friction_test.world (PhysicsTorsionalFrictionTest.RadiusTest,
    physics_torsional_friction.cc:383-479): spheres of radius 0.01 .. 10 m, coefficient 0.005, use_patch_radius false and
    their own radius as surface radius: the patch is sqrt(R x depth), the plane's curvature being zero.
    Returns the scene, the parameters and the radii."""
    sb = S.SceneBuilder()
    sb.begin_world()
    sb.add_geom(S.GEOM_PLANE, (0, 0, 1, 0), sb.new_space(), mu=100.0, mu2=50.0, mu3=1.0)
    radii = (0.01, 0.1, 0.8, 2.0, 10.0)
    mass, izz = 11.309733552923259, 0.045238934211693047
    for k, rad in enumerate(radii):
        b = sb.add_body([(0.04, 0.8, 9.6, 32.0, 200.0)[k], 3.0, rad], mass=mass, inertia=(izz, izz, izz, 0, 0, 0))
        sb.add_geom(S.GEOM_SPHERE, (rad,), sb.new_space(), body=b, mu3=0.005, surface_radius=rad, use_patch_radius=0.0, kp=5e5)
    sb.end_world()
    return sb.finish(), S.default_params(), radii


def inertia_ratio_pendulum(dynamic_inertia=1):
    """test/worlds/inertia_ratio_pendulum.world restated (PhysicsTest.InertiaRatioPendulum,
    test/integration/physics_inertia_ratio.cc:39-84): a 27 g upper link on a hinge to the world carrying a 3.375 kg lower
    link on a second hinge (inertia ratio ~3 000 along the constrained axes), both about x, released at 45 / 90 degrees
    with gravity (0.1, 0, -9.81): 250 iterations, sor 1.4, use_dynamic_moi_rescaling."""
    sb = S.SceneBuilder()
    sb.begin_world()
    sp = sb.new_space()
    a = np.pi / 4
    up = sb.add_body([0, 0.05 * np.sin(a), 0.9 - 0.05 * np.cos(a)], S.q_from_rpy(a, 0, 0), mass=0.02700000000000001,
                     inertia=(2.272500000000001e-05, 2.272500000000001e-05, 4.5000000000000035e-07, 0, 0, 0))
    sb.add_geom(S.GEOM_BOX, (0.010000000000000002, 0.010000000000000002, 0.1), sp, body=up)
    y, z = 0.17677669529663687, -0.2767766952966369
    lo = sb.add_body([0, y * np.cos(a) - z * np.sin(a), 0.9 + y * np.sin(a) + z * np.cos(a)], S.q_from_rpy(2 * a, 0, 0),
                     mass=3.375, inertia=(0.071015625, 0.071015625, 0.0014062500000000004, 0, 0, 0))
    sb.add_geom(S.GEOM_BOX, (0.05, 0.05, 0.5), sp, body=lo)
    sb.add_hinge(up, -1, [0, 0, 0.9], [1, 0, 0])
    sb.add_hinge(lo, up, [0, 0.1 * np.sin(a), 0.9 - 0.1 * np.cos(a)], [1, 0, 0])
    sb.end_world()
    p = S.default_params(gravity=(0.1, 0, -9.81), iters=250, sor_w=1.4, cfm=0.0, erp=0.2, contact_max_correcting_vel=0.0,
                         contact_surface_layer=0.0, dynamic_inertia=dynamic_inertia)
    return sb.finish(), p, None


def max_out_of_plane_angles(r, steps=3000, every=10):
    """max |Euler Y|, max |Euler Z| of all links over the run (physics_inertia_ratio.cc:66-78)."""
    my = mz = 0.0
    for _ in range(steps // every):
        r.step(every)
        w, x, y, z = r.body_state()[1].T
        my = max(my, float(np.abs(np.arcsin(np.clip(2 * (w * y - z * x), -1, 1))).max()))
        mz = max(mz, float(np.abs(np.arctan2(2 * (w * z + x * y), 1 - 2 * (y * y + z * z))).max()))
    return my, mz


def elastic_modulus_contact():
    """test/worlds/elastic_modulus_contact_test.world restated (PhysicsTest.ElasticModulusContact,
    test/integration/elastic_modulus.cc:60-225): gravity -100 along x, 300 iterations, sor 1; a static 1 m box at the origin
    (E = 5e5, nu = 0.3, surface radius 0.08) and a 3 kg sphere of radius 0.5 (E = 6e5, nu = 0.4, surface radius 0.06, min
    depth 0.001) that falls onto the box's +x face; both surfaces: kp 1, kd 100, torsional coefficient 1 on the surface
    radius.  The reference's test holds the converged numbers (:115-117): the only golden vector in its tests for this path."""
    sb = S.SceneBuilder()
    sb.begin_world()
    common = dict(mu=1.0, mu2=1.0, kp=1.0, kd=100.0, mu3=1.0, use_patch_radius=0.0)
    sb.add_geom(S.GEOM_BOX, (1, 1, 1), sb.new_space(), pos=(0, 0, 0), poissons_ratio=0.3, elastic_modulus=5e5,
                surface_radius=0.08, **common)
    b = sb.add_body([2, 0.3, 0.2], mass=3.0, inertia=(1, 1, 1, 0, 0, 0), max_vel=100.0, min_depth=0.001)
    sb.add_geom(S.GEOM_SPHERE, (0.5,), sb.new_space(), body=b, poissons_ratio=0.4, elastic_modulus=6e5,
                surface_radius=0.06, **common)
    sb.end_world()
    p = S.default_params(gravity=(-100.0, 0, 0), iters=300, sor_w=1.0, cfm=0.0, erp=0.5,
                         contact_max_correcting_vel=1000.0, contact_surface_layer=0.0)
    return sb.finish(), p, None


def check_elastic_modulus_contact(r):
    """The assertions of elastic_modulus.cc:108-190 on a stepper with .step / .contacts / .body_state."""
    d_converged, k_converged, k_lin_converged = 0.0258629, 71454.617, 11491.286   # :115-117
    nu1, e1, m1, r1, nu2, e2, r2 = 0.4, 600000.0, 3.0, 0.06, 0.3, 500000.0, 0.08
    e_star = 1.0 / ((1.0 - nu1 * nu1) / e1 + (1.0 - nu2 * nu2) / e2)
    r_star = 1.0 / (1.0 / r1 + 1.0 / r2)
    f1 = 100.0 * m1
    float_eq = lambda a, b: abs(np.float32(a) - np.float32(b)) <= 4 * np.spacing(np.float32(b))   # EXPECT_FLOAT_EQ: 4 ULPs
    r.step(3000)
    for _ in range(10):
        r.step(1)
        cg, cd = r.contacts(0)
        assert len(cg) == 1
        min_depth = min(0.0015, 0.001)
        d1 = cd[0, 6] - min_depth
        assert d1 < d_converged and abs((d1 - d_converged) / d_converged) < 1e-2
        d2 = 1.0 - r.body_state()[0][0, 0] - min_depth
        assert float_eq(d1, d2)
        d3 = (f1 / (4.0 / 3.0 * e_star * np.sqrt(np.sqrt(r_star)))) ** (1.0 / 1.75)
        assert abs(d1 - d3) / d3 < 1e-2
        patch_radius = np.sqrt(r_star * d1)
        assert float_eq(4.0 / 3.0 * e_star * np.sqrt(patch_radius), k_converged)
        assert float_eq(4.0 / 3.0 * e_star * np.sqrt(patch_radius * d1), k_lin_converged)
    return d1


def ray_targets(n_worlds=2, per_type=6, seed=77):
    """Worlds for the laser: a ground plane and, floating at rest (no gravity flag needed: they are never stepped far), boxes,
    spheres, capsules and cylinders at random poses inside a 6 m cube, one sensor-category geom that rays must not see."""
    sb = S.SceneBuilder()
    for w in range(n_worlds):
        rng = np.random.Generator(np.random.PCG64(seed + w))
        sb.begin_world()
        sb.add_ground_plane()
        for k in range(4 * per_type):
            pos = rng.uniform(-3, 3, 3) + np.array([0, 0, 3.5])
            q = S.q_random(rng)
            t = k % 4
            if t == 0:
                dims = tuple(rng.uniform(0.2, 1.0, 3)); b = sb.add_body(pos, q, 1.0, S.box_inertia(1.0, *dims)); sb.add_geom(S.GEOM_BOX, dims, sb.new_space(), body=b)
            elif t == 1:
                r = rng.uniform(0.1, 0.6); b = sb.add_body(pos, q, 1.0, S.sphere_inertia(1.0, r)); sb.add_geom(S.GEOM_SPHERE, (r,), sb.new_space(), body=b)
            elif t == 2:
                r, l = rng.uniform(0.1, 0.4), rng.uniform(0.2, 1.2); b = sb.add_body(pos, q, 1.0, S.cylinder_inertia(1.0, r, l)); sb.add_geom(S.GEOM_CAPSULE, (r, l), sb.new_space(), body=b)
            else:
                r, l = rng.uniform(0.1, 0.5), rng.uniform(0.2, 1.2); b = sb.add_body(pos, q, 1.0, S.cylinder_inertia(1.0, r, l)); sb.add_geom(S.GEOM_CYLINDER, (r, l), sb.new_space(), body=b)
        # a big box of the sensor category right in the middle (GZ_SENSOR_COLLIDE = 2, collides with nothing of its kind): invisible to the rays
        sb.add_geom(S.GEOM_BOX, (2.0, 2.0, 2.0), sb.new_space(), pos=(0, 0, 3.5), category_bits=0x2, collide_bits=0xFFFFFFFD)
        sb.end_world()
    return sb.finish(), S.default_params(), None


def laser_rays(n, seed=5):
    """n rays: origins inside and around the 6 m cube of ray_targets (some inside shapes), random directions of random
    length (dGeomRaySet normalises), lengths 0.5..12 m; a few axis-parallel ones (the colliders' parallel branches)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    start = rng.uniform(-4, 4, (n, 3)) + np.array([0, 0, 3.5])
    d = rng.normal(size=(n, 3)) * rng.uniform(0.1, 5.0, (n, 1))
    for k in range(0, n, 17):
        d[k] = 0.0; d[k, (k // 17) % 3] = 1.0 if (k // 51) % 2 == 0 else -1.0
    length = rng.uniform(0.5, 12.0, n)
    return start, d, length


GOLDEN = {
    "box_drop": (box_drop, 1500),
    "tilted_box": (tilted_box, 600),
    "shapes": (shapes, 600),
    "chain": (chain, 500),
    "revolve4": (revolve4, 400),
    "boxes30": (boxes30, 500),
    "stack": (stack, 1500),
    "capsules": (capsules, 500),
    "joints": (joints, 600),
    "cylinders": (cylinders, 500),
    "cylboxes": (cylboxes, 500),
    "cylcyl": (cylcyl, 400),
    "fdir1": (fdir1, 300),
    "dyn_inertia": (dyn_inertia, 400),
    "carts": (carts, 600),
    "hertz": (hertz, 700),
    "gimbals": (gimbals, 800),
    "gears": (gears, 700),
    "screws": (screws, 900),
    "spinners": (spinners, 800),
}
