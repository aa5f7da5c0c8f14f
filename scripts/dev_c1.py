"""Developer check (GPU box): BASELINE.json configs[0] (500 boxes on a plane, one world) — GPU vs CPU reference."""
import sys, time; sys.path.insert(0, "/root/repo")
import numpy as np
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
from oracle.refharness import RefBatch
n = int(sys.argv[1]) if len(sys.argv) > 1 else 500
sc = W.box_drop_worlds(1, n)
p = S.default_params()
g = Batch(sc, p)
r = RefBatch(sc, p, mode=1)
for k in range(6):
    t0 = time.perf_counter(); g.step(250); g.sync(); tg = time.perf_counter() - t0
    t0 = time.perf_counter(); r.step(250); tr = time.perf_counter() - t0
    gs, rs = g.body_state(), r.body_state()
    d = [np.abs(a - b).max() for a, b in zip(gs, rs)]
    rms, nc, m = g.stats()
    print("step %4d rows %5d contacts %5d | gpu %.2f ms/step, cpu(canonical, 1 thread) %.2f ms/step | dpos %.2e dq %.2e | stage ms %s" % (
        250 * (k + 1), m[0], nc[0], 1e3 * tg / 250, 1e3 * tr / 250, d[0], d[1], ["%.3f" % x for x in g.stage_ms()]), flush=True)
