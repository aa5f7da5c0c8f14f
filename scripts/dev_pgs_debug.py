"""Developer check (GPU box): distribution of the PGS schedule over the headline batch."""
import sys; sys.path.insert(0, "/root/repo")
import numpy as np
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
sc, pd = W.revolve_worlds(4096)
b = Batch(sc, S.default_params()); b.set_hinge_pd(pd)
for tot in (5, 60, 150):
    b.step(tot)
    d = b.pgs_debug(); rms, nc, m = b.stats()
    ms = b.stage_ms()
    R, k = d[:, 1] >> 8, d[:, 1] & 255
    print("step", b.counters()[3] // 4096, "stage ms", ms)
    for name, v in (("rows", m), ("warp rounds", d[:, 0]), ("R", R), ("k", k), ("fill", m / np.maximum(R * k, 1)),
                    ("cycles", d[:, 2]), ("cyc/round", d[:, 2] / np.maximum(d[:, 0], 1))):
        print("  %-12s mean %10.2f  p50 %10.2f  p90 %10.2f  p99 %10.2f  max %10.2f" % (name, v.mean(), np.percentile(v, 50), np.percentile(v, 90), np.percentile(v, 99), v.max()))
    print("  k histogram", {int(x): int((k == x).sum()) for x in np.unique(k)})
