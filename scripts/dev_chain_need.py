"""Developer check (GPU box): shared memory the chain-mode worlds of the headline batch need (compact records)."""
import os, sys; sys.path.insert(0, "/root/repo")
import numpy as np
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
sc, pd = W.revolve_worlds(4096)
b = Batch(sc, S.default_params()); b.set_hinge_pd(pd)
b.step(int(os.environ.get("STEPS", "120")))
d = b.pgs_debug(); rms, nc, m = b.stats()
R, k = d[:, 1] >> 8, d[:, 1] & 255
ch = k == 16
m1 = 3 * nc[ch]; m2 = m[ch] - m1
need = (12 * 9 + 322 + 2 * m[ch] + 20 * m1 + 32 * m2 + 32 + 32 * R[ch]) * 8 / 1024.0
old = (12 * 9 + 322 + 2 * m[ch] + 34 * (m[ch] + 1) + 32 * R[ch]) * 8 / 1024.0
print("chain worlds %d; rows mean %.0f max %d; R mean %.1f max %d" % (ch.sum(), m[ch].mean(), m[ch].max(), R[ch].mean(), R[ch].max()))
print("need KB (compact): p10 %.1f p25 %.1f p50 %.1f p75 %.1f p90 %.1f p99 %.1f max %.1f | sum %.1f MB" % (*np.percentile(need, [10, 25, 50, 75, 90, 99]), need.max(), need.sum() / 1024))
print("need KB (packed) : p10 %.1f p25 %.1f p50 %.1f p75 %.1f p90 %.1f p99 %.1f max %.1f | sum %.1f MB" % (*np.percentile(old, [10, 25, 50, 75, 90, 99]), old.max(), old.sum() / 1024))
print("cycles of the sweeps by need quartile:", [int(d[ch, 2][(need >= a) & (need < c)].mean()) for a, c in zip(np.percentile(need, [0, 25, 50, 75]), list(np.percentile(need, [25, 50, 75])) + [1e9])])
print("regular worlds: R hist", {int(x): int((R[~ch] == x).sum()) for x in np.unique(R[~ch])})
