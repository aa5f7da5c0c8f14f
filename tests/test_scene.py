"""Host-side scene construction: joint frames, CSR offsets, defaults."""
import numpy as np

from gazebo_b200 import scene as S, worlds as W


def test_hinge_frames_reconstruct_world_anchor():
    sc, pd = W.revolve_worlds(3)
    for w in range(sc.n_worlds):
        b0, j0, j1 = sc.body_off[w], sc.joint_off[w], sc.joint_off[w + 1]
        for j in range(j0, j1):
            b1, b2 = b0 + sc.joint_body1[j], b0 + sc.joint_body2[j]
            R1, R2 = S.q_to_R(sc.body_quat[b1]), S.q_to_R(sc.body_quat[b2])
            a1 = R1 @ sc.joint_anchor1[j] + sc.body_pos[b1]
            a2 = R2 @ sc.joint_anchor2[j] + sc.body_pos[b2]
            assert np.allclose(a1, a2, atol=1e-12)
            assert np.allclose(R1 @ sc.joint_axis1[j], R2 @ sc.joint_axis2[j], atol=1e-12)
            assert abs(np.linalg.norm(sc.joint_qrel[j]) - 1) < 1e-12


def test_revolve_world_shape():
    sc, pd = W.revolve_worlds(5)
    assert sc.n_worlds == 5 and sc.nb == 40 and sc.nj == 35 and pd.shape == (35, 6)
    assert (np.diff(sc.body_off) == 8).all() and (np.diff(sc.joint_off) == 7).all()
    assert (sc.geom_type[sc.geom_off[:-1]] == S.GEOM_PLANE).all()
    assert (sc.body_mass > 0).all()


def test_seeding_is_per_world():
    a, _ = W.revolve_worlds(4)
    b, _ = W.revolve_worlds(2, first_world=2)
    assert np.array_equal(a.body_pos[16:], b.body_pos)


def test_static_geom_bits_follow_gazebo():
    sc, _ = W.revolve_worlds(1)
    assert sc.geom_category_bits[0] == S.GZ_FIXED_COLLIDE
    assert sc.geom_collide_bits[0] == (~S.GZ_FIXED_COLLIDE) & 0xFFFFFFFF
    assert sc.geom_category_bits[1] == S.GZ_ALL_COLLIDE


def test_default_params_match_sdf_defaults():
    p = S.default_params()
    assert p.iters == 50 and p.friction_iters == 10 and p.sor_w == 1.3
    assert p.step_size == 0.001 and p.max_contacts == 20 and p.cfm == 0.0 and p.erp == 0.2
    assert p.contact_surface_layer == 0.001 and p.contact_max_correcting_vel == 100.0
