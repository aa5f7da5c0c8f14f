"""Developer check (GPU box): cycles of k_schedule's phases per world (GZCUDA_PGS_PROF=1)."""
import os, sys; sys.path.insert(0, "/root/repo")
os.environ["GZCUDA_PGS_PROF"] = "1"
import numpy as np
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
sc, pd = W.revolve_worlds(4096)
b = Batch(sc, S.default_params()); b.set_hinge_pd(pd)
b.step(100); b.sync()
b.pgs_profile()   # the getter resets the sums
b.step(20); b.sync(); p1 = b.pgs_profile(); p0 = [0] * 16
n = 20 * 4096
print("k_schedule cycles per world: order/bodies %.0f | items, k, probe %.0f | serial placement %.0f" % (
    (p1[6] - p0[6]) / n, (p1[7] - p0[7]) / n, (p1[15] - p0[15]) / n))
