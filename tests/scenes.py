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


GOLDEN = {
    "box_drop": (box_drop, 1500),
    "tilted_box": (tilted_box, 600),
    "shapes": (shapes, 600),
    "chain": (chain, 500),
    "revolve4": (revolve4, 400),
    "boxes30": (boxes30, 500),
    "stack": (stack, 1500),
}
