"""Synthetic worlds for BASELINE.json's configs.

The reference ships no Revolve robot and no pioneer2dx model (models/ holds only
ground_plane and sun), so every config is generated here from a seed.  RNG:
numpy PCG64 seeded with ``1000 * config + world_index``; both the GPU path and the
CPU oracle consume the arrays this module emits, so they see identical inputs.
"""
import math

import numpy as np

from . import scene as S

# Revolve-style module dimensions / masses
CORE = dict(size=(0.089, 0.089, 0.089), mass=0.25)
HINGE_A = dict(size=(0.052, 0.053, 0.035), mass=0.011)   # frame half, rigidly on the parent link
HINGE_B = dict(size=(0.052, 0.053, 0.035), mass=0.058)   # servo half, start of the child link
BRICK = dict(size=(0.041, 0.041, 0.035), mass=0.03)
HINGE_LIMIT = math.pi / 3
HINGE_EFFORT = 1.8


def _compound(parts):
    """parts: list of (mass, centre(3), size(3)) axis-aligned boxes in one link frame.
    Returns total mass, CoM, inertia(6) about the CoM (parallel-axis theorem)."""
    m = sum(p[0] for p in parts)
    com = sum(p[0] * np.asarray(p[1], float) for p in parts) / m
    I = np.zeros((3, 3))
    for pm, pc, ps in parts:
        d = np.asarray(pc, float) - com
        lx, ly, lz = ps
        I += np.diag([pm * (ly * ly + lz * lz) / 12, pm * (lx * lx + lz * lz) / 12,
                      pm * (lx * lx + ly * ly) / 12])
        I += pm * (np.dot(d, d) * np.eye(3) - np.outer(d, d))
    return m, com, [I[0, 0], I[1, 1], I[2, 2], I[0, 1], I[0, 2], I[1, 2]]


def add_revolve_robot(sb, rng, clearance=0.002):
    """8 links / 7 revolute joints, spider (4 arms) or snake, self_collide false.

    Link 0 = core (+ the frame halves of the hinges mounted on it); every further
    link = servo half + brick (+ frame half of the next hinge on that arm).
    Returns the PD table rows for its 7 joints."""
    space = sb.new_space()
    spider = rng.random() < 0.5
    arms = [2, 2, 2, 1] if spider else [7]
    dirs = [np.array(d, float) for d in ((1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0))]
    hl = HINGE_A["size"][0]
    bl = BRICK["size"][0]

    # Build link part lists in the WORLD frame first (all links axis-aligned at start,
    # arms along +-x / +-y; a yaw of pi/2 swaps the x/y extents of the boxes).
    zc = CORE["size"][2] / 2 + clearance     # every module starts just above the plane
    z0 = HINGE_A["size"][2] / 2 + clearance
    links = [[(CORE["mass"], np.array([0, 0, zc]), CORE["size"])]]
    joints = []  # (parent link, child link, anchor, axis)
    for a, n in enumerate(arms):
        d = dirs[a]
        swap = abs(d[1]) > 0.5

        def sz(s):
            return (s[1], s[0], s[2]) if swap else s
        parent = 0
        x = CORE["size"][0] / 2
        for _ in range(n):
            # frame half on the parent, servo half + brick on the child
            ca = d * (x + hl / 2) + np.array([0, 0, z0])
            links[parent].append((HINGE_A["mass"], ca, sz(HINGE_A["size"])))
            anchor = d * (x + hl) + np.array([0, 0, z0])
            cb = d * (x + hl + hl / 2) + np.array([0, 0, z0])
            cbrick = d * (x + 2 * hl + bl / 2) + np.array([0, 0, z0])
            child = len(links)
            links.append([(HINGE_B["mass"], cb, sz(HINGE_B["size"])),
                          (BRICK["mass"], cbrick, sz(BRICK["size"]))])
            # hinge axis horizontal, perpendicular to the arm (alternating with vertical)
            axis = np.cross(d, [0, 0, 1]) if (len(joints) % 2 == 0) else np.array([0, 0, 1.0])
            joints.append((parent, child, anchor, axis))
            parent = child
            x = x + 2 * hl + bl
    assert len(links) == 8 and len(joints) == 7

    yaw = rng.uniform(-math.pi, math.pi)
    q = S.q_from_axis_angle([0, 0, 1], yaw)
    R = S.q_to_R(q)
    bidx = []
    for parts in links:
        m, com, I = _compound(parts)
        # rotate the whole robot about the world z axis through the origin
        b = sb.add_body(R @ com, q, m, I)
        bidx.append(b)
        for pm, pc, ps in parts:
            sb.add_geom(S.GEOM_BOX, ps, space, body=b, pos=np.asarray(pc) - com)
    pd = []
    for (p, c, anchor, axis) in joints:
        sb.add_hinge(bidx[p], bidx[c], R @ anchor, R @ axis,
                     lostop=-HINGE_LIMIT, histop=HINGE_LIMIT)
        pd.append([rng.uniform(0.2, 1.0), rng.uniform(0.5, 2.0), rng.uniform(0, 2 * math.pi),
                   1.0, 0.05, HINGE_EFFORT])
    return pd


def revolve_worlds(n_worlds, config=2, first_world=0):
    """C2: plane + one 8-link robot per world.  Returns (Scene, pd[nj][6])."""
    sb = S.SceneBuilder()
    pd = []
    for w in range(n_worlds):
        rng = np.random.Generator(np.random.PCG64(1000 * config + first_world + w))
        sb.begin_world()
        sb.add_ground_plane()
        pd += add_revolve_robot(sb, rng)
        sb.end_world()
    return sb.finish(), np.array(pd, dtype=np.float64).reshape(-1, 6)


def box_drop_worlds(n_worlds=1, n_boxes=500, config=1, auto_disable=True, first_world=0):
    """C1: plane + n_boxes single-box models on a jittered grid, random orientation."""
    sb = S.SceneBuilder()
    for w in range(n_worlds):
        rng = np.random.Generator(np.random.PCG64(1000 * config + first_world + w))
        sb.begin_world()
        sb.add_ground_plane()
        nx = ny = 10
        nz = max(1, (n_boxes + nx * ny - 1) // (nx * ny))
        k = 0
        for iz in range(nz):
            for iy in range(ny):
                for ix in range(nx):
                    if k >= n_boxes:
                        break
                    side = rng.uniform(0.1, 0.4, size=3)
                    mass = 500.0 * side[0] * side[1] * side[2]
                    pos = [(-2 + 4 * (ix + 0.5) / nx) + rng.uniform(-0.05, 0.05),
                           (-2 + 4 * (iy + 0.5) / ny) + rng.uniform(-0.05, 0.05),
                           0.5 + (5.5 * iz / max(1, nz - 1) if nz > 1 else 0.0) + rng.uniform(0, 0.1)]
                    b = sb.add_body(pos, S.q_random(rng), mass, S.box_inertia(mass, *side),
                                    flags=S.BODY_AUTO_DISABLE if auto_disable else 0)
                    sb.add_geom(S.GEOM_BOX, side, sb.new_space(), body=b)
                    k += 1
        sb.end_world()
    return sb.finish()


def chain_worlds(n_worlds=256, n_links=50, config=5, first_world=0):
    """C5: hanging chain of n_links boxes joined by hinges about y, first joint to the world,
    every joint bent by 0.3 rad, no ground, no limits, no motor."""
    sb = S.SceneBuilder()
    L = 0.2
    for w in range(n_worlds):
        sb.begin_world()
        top = np.array([0.0, 0.0, 0.0])
        ang = 0.0
        prev = -1
        space = sb.new_space()
        for i in range(n_links):
            ang += 0.3 if i < 3 else 0.0   # bend only the first joints so the chain stays compact
            q = S.q_from_axis_angle([0, 1, 0], ang)
            R = S.q_to_R(q)
            com = top + R @ np.array([0, 0, -L / 2])
            b = sb.add_body(com, q, 0.2, S.box_inertia(0.2, 0.05, 0.05, L))
            sb.add_geom(S.GEOM_BOX, (0.05, 0.05, L), space, body=b)
            if prev < 0:
                sb.add_hinge(b, -1, top, [0, 1, 0])
            else:
                sb.add_hinge(prev, b, top, [0, 1, 0])
            prev = b
            top = top + R @ np.array([0, 0, -L])
        sb.end_world()
    return sb.finish()


def pioneer_worlds(n_worlds=1024, config=3, first_world=0, drive=False, torque=5.0):
    """C3: pioneer2dx restated (asset not vendored in the reference): chassis box + castor
    sphere on one link, two cylinder wheels on hinge joints about y.  Returns
    (Scene, wheel_cmd[n_worlds][2]) = target wheel joint velocities (rad/s).
    drive=True: the wheel hinges run their velocity motor (dParamVel = wheelSpeed / wheelRadius,
    dParamFMax = torque: Joint::SetParam("vel" / "fmax"), the way diff-drive plugins command wheels
    through the solver).  This fork's DiffDrivePlugin::OnUpdate (plugins/DiffDrivePlugin.cc:100-122)
    instead calls Joint::SetVelocity -> SetVelocityMaximal (ODEHingeJoint.cc:153-156), a host write
    of the wheel link velocities every update: that is gzcuda_batch_set_body_state on this side."""
    sb = S.SceneBuilder()
    cmds = []
    for w in range(n_worlds):
        rng = np.random.Generator(np.random.PCG64(1000 * config + first_world + w))
        sb.begin_world()
        sb.add_ground_plane()
        space = sb.new_space()
        chassis = sb.add_body([0, 0, 0.16], (1, 0, 0, 0), 5.67, [0.07, 0.08, 0.10, 0, 0, 0])
        sb.add_geom(S.GEOM_BOX, (0.445, 0.277, 0.17), space, body=chassis, pos=(0, 0, 0.0))
        sb.add_geom(S.GEOM_SPHERE, (0.04,), space, body=chassis, pos=(-0.2, 0, -0.12),
                    mu=0.0, mu2=0.0, slip1=1.0, slip2=1.0)
        qw = S.q_from_axis_angle([1, 0, 0], math.pi / 2)
        v, yaw = 0.2, rng.uniform(-0.5, 0.5)
        sep, r = 0.34, 0.11
        cmd = [(v + yaw * sep / 2) / r, (v - yaw * sep / 2) / r]
        wheels = []
        for k, sy in enumerate((+1, -1)):
            wb = sb.add_body([0.1, sy * 0.17, 0.11], qw, 1.5, [0.0051, 0.0051, 0.0090, 0, 0, 0])
            sb.add_geom(S.GEOM_CYLINDER, (0.11, 0.05), space, body=wb)
            if drive:
                sb.add_hinge(wb, chassis, [0.1, sy * 0.17, 0.11], [0, 1, 0], vel=cmd[k], fmax=torque)
            else:
                sb.add_hinge(wb, chassis, [0.1, sy * 0.17, 0.11], [0, 1, 0])
            wheels.append(wb)
        cmds.append(cmd)
        sb.end_world()
    return sb.finish(), np.array(cmds)


def rubble_world(n_bodies=20000, config=4, pen=None, auto_disable=False):
    """C4: plane + 4 static wall boxes + n_bodies single-body models: 60 % boxes (half "2x4" planks 0.6-2.4 x 0.1016 x 0.0508 m,
    half cinder blocks 0.2032 x 0.2032 x 0.4064 m as plugins/RubblePlugin.cc:71-91, scaled by 0.5), 25 % spheres r U(0.05, 0.15),
    15 % capsules r U(0.04, 0.08), l U(0.1, 0.4); mass U(0.1, 1.0) kg; dropped in layers into a square pen (24 x 24 m at
    20 000 bodies: 7 layers, the top one falls 3.7 m; the side scaled with sqrt(n) below that so the pile is as deep).
    A 10 m pen (40 layers, the top one falls 22 m) ends in the reference as it does here: bodies shot out of the pile."""
    rng = np.random.Generator(np.random.PCG64(1000 * config))
    if pen is None:
        pen = 24.0 * math.sqrt(n_bodies / 20000.0)
    sb = S.SceneBuilder()
    sb.begin_world()
    sb.add_ground_plane()
    wall_h, wall_t = 2.0, 0.2
    for sx, sy in ((1, 0), (-1, 0), (0, 1), (0, -1)):
        size = (wall_t, pen + 2 * wall_t, wall_h) if sx else (pen + 2 * wall_t, wall_t, wall_h)
        sb.add_geom(S.GEOM_BOX, size, sb.new_space(), pos=(sx * (pen + wall_t) / 2, sy * (pen + wall_t) / 2, wall_h / 2))
    per_layer = max(1, int((pen / 0.45) ** 2))
    side = int(math.ceil(math.sqrt(per_layer)))
    flags = S.BODY_AUTO_DISABLE if auto_disable else 0
    for k in range(n_bodies):
        layer, cell = divmod(k, per_layer)
        ix, iy = divmod(cell, side)
        pos = [-pen / 2 + (ix + 0.5) * pen / side + rng.uniform(-0.05, 0.05),
               -pen / 2 + (iy + 0.5) * pen / side + rng.uniform(-0.05, 0.05),
               0.4 + 0.55 * layer + rng.uniform(0, 0.1)]
        pos[0] = min(max(pos[0], -pen / 2 + 0.3), pen / 2 - 0.3)
        pos[1] = min(max(pos[1], -pen / 2 + 0.3), pen / 2 - 0.3)
        mass = rng.uniform(0.1, 1.0)
        u = rng.uniform()
        q = S.q_random(rng)
        if u < 0.30:
            dims = (0.5 * rng.uniform(0.6, 2.4) * 0.35, 0.5 * 0.1016, 0.5 * 0.0508)   # plank (length cut so it fits its cell)
            b = sb.add_body(pos, q, mass, S.box_inertia(mass, *dims), flags=flags)
            sb.add_geom(S.GEOM_BOX, dims, sb.new_space(), body=b)
        elif u < 0.60:
            dims = (0.5 * 0.2032, 0.5 * 0.2032, 0.5 * 0.4064)
            b = sb.add_body(pos, q, mass, S.box_inertia(mass, *dims), flags=flags)
            sb.add_geom(S.GEOM_BOX, dims, sb.new_space(), body=b)
        elif u < 0.85:
            r = rng.uniform(0.05, 0.15)
            b = sb.add_body(pos, q, mass, S.sphere_inertia(mass, r), flags=flags)
            sb.add_geom(S.GEOM_SPHERE, (r,), sb.new_space(), body=b)
        else:
            r, l = rng.uniform(0.04, 0.08), rng.uniform(0.1, 0.4)
            b = sb.add_body(pos, q, mass, S.cylinder_inertia(mass, r, l), flags=flags)
            sb.add_geom(S.GEOM_CAPSULE, (r, l), sb.new_space(), body=b)
    sb.end_world()
    return sb.finish()
