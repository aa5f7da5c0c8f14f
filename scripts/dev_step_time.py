"""Developer check (GPU box): wall-clock ms per step of the headline batch (device-resident), current env knobs."""
import os, sys, time; sys.path.insert(0, "/root/repo")
import numpy as np
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
nw = int(os.environ.get("NW", "4096"))
sc, pd = W.revolve_worlds(nw)
b = Batch(sc, S.default_params()); b.set_hinge_pd(pd)
b.step(100); b.sync()
ts = []
for rep in range(5):
    t = time.perf_counter(); b.step(100); b.sync(); ts.append((time.perf_counter() - t) * 10)
b.set_profiling(True)
per = []
for i in range(12):
    b.step(1); per.append(b.pgs_ms()[0])
tag = " ".join("%s=%s" % (e[7:], os.environ[e]) for e in sorted(os.environ) if e.startswith("GZCUDA_"))
print("%-36s ms/step %s | pgs per step %s" % (tag, ["%.3f" % x for x in ts], ["%.2f" % x for x in per]), flush=True)
for i in range(30):
    b.step(1); t = b.pgs_ms()[0]
    if t > 1.8:
        d = b.pgs_debug(); rms, nc, m = b.stats()
        for w in np.argsort(-d[:, 2])[:3]:
            print("  slow step %d (%.2f ms): world %d cycles %d rounds %d R %d k %d rows %d contacts %d" % (i, t, w, d[w, 2], d[w, 0], d[w, 1] >> 8, d[w, 1] & 255, m[w], nc[w]))
