"""Developer check (GPU box): tail of the per-world PGS period / cycle distribution of the headline batch."""
import os, sys; sys.path.insert(0, "/root/repo")
import numpy as np
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
sc, pd = W.revolve_worlds(4096)
b = Batch(sc, S.default_params()); b.set_hinge_pd(pd)
b.step(120)
d = b.pgs_debug(); rms, nc, m = b.stats()
R, k, cyc = d[:, 1] >> 8, d[:, 1] & 255, d[:, 2]
print("R percentiles", [int(np.percentile(R, p)) for p in (50, 75, 90, 95, 98, 99, 99.5, 100)])
print("cycles percentiles", [int(np.percentile(cyc, p)) for p in (50, 75, 90, 95, 98, 99, 99.5, 100)])
for t in (30, 35, 40, 50, 60, 70, 80):
    print("R >= %d: %d worlds, rows mean %.0f" % (t, (R >= t).sum(), m[R >= t].mean() if (R >= t).any() else 0))
for t in (1.2e6, 1.5e6, 1.8e6, 2.1e6, 2.4e6):
    print("cycles >= %.1fM: %d worlds" % (t / 1e6, (cyc >= t).sum()))
top = np.argsort(-cyc)[:12]
for w in top:
    print("world %4d cycles %8d rounds %5d R %3d k %2d rows %4d contacts %3d" % (w, cyc[w], d[w, 0], R[w], k[w], m[w], nc[w]))
b.set_profiling(True); b.step(20); pgs_ms, n = b.pgs_ms(); print("pgs ms", pgs_ms / n, "stage", b.stage_ms())
