"""gzcuda_batch_get_joint_feedback_noalloc (include/gzcuda.h): the joint-feedback entry point that allocates nothing
(static device scratch inside the library, no counterpart in the
reference).  Called through the C ABI and checked against the existing gzcuda_batch_get_joint_feedback on two calls
in a row - the second after more steps, so a stale scratch would show - and beyond its stated maximum.  Collected
last (file name) so that a failure here hides no other test under `pytest -x`."""
import numpy as np
import pytest

TIGHT = 1e-9   # the tolerance of tests/test_parity_gpu.py for states and wrenches


@pytest.mark.gpu
def test_noalloc_joint_feedback_matches_the_allocating_entry_point():
    from gazebo_b200 import scene as S, worlds as W
    from gazebo_b200.batch import Batch
    sc, pd = W.revolve_worlds(24)
    g = Batch(sc, S.default_params())
    g.set_hinge_pd(pd)
    for steps in (60, 45):                      # two calls in a row through the same static scratch
        g.step(steps)
        ref = g.joint_feedback()                # existing entry point (cudaMalloc per call)
        got = g.joint_feedback_noalloc()        # new entry point
        assert got.shape == ref.shape == (sc.nj, 12)
        assert np.abs(ref).max() > 0            # robots standing on the plane: the joints carry load
        assert np.abs(got - ref).max() <= TIGHT * max(1.0, np.abs(ref).max())
    # a smaller batch afterwards: the scratch still holds the larger batch's wrenches beyond its own range
    sc2 = W.chain_worlds(2, 5)
    g2 = Batch(sc2, S.default_params())
    g2.step(30)
    ref2 = g2.joint_feedback()
    got2 = g2.joint_feedback_noalloc()
    assert np.abs(got2 - ref2).max() <= TIGHT * max(1.0, np.abs(ref2).max())
    # and the first batch again
    assert np.abs(g.joint_feedback_noalloc() - g.joint_feedback()).max() == 0.0


@pytest.mark.gpu
def test_noalloc_rejects_batches_beyond_its_maximum():
    from gazebo_b200 import scene as S, worlds as W
    from gazebo_b200.batch import Batch, GzCudaError
    n_worlds = 32768 // 7 + 1                   # 7 hinges per robot: one robot more than GZCUDA_NOALLOC_MAX_JOINTS allows
    sc, pd = W.revolve_worlds(n_worlds)
    assert sc.nj > 32768
    g = Batch(sc, S.default_params())
    g.step(1)
    with pytest.raises(GzCudaError, match="GZCUDA_NOALLOC_MAX_JOINTS"):
        g.joint_feedback_noalloc()
    assert g.joint_feedback().shape == (sc.nj, 12)   # the allocating entry point still serves it
