"""Developer run (GPU box): C1 (500 boxes dropped on the plane, one world) time per step for each PGS implementation
(GZCUDA_PGS_IMPL in the environment), with parity against the reference's opende when REF=1."""
import os, sys, time; sys.path.insert(0, "/root/repo")
import numpy as np
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
from oracle import refharness
nbox = int(os.environ.get("N", "500")); steps = int(os.environ.get("STEPS", "1500")); chk = int(os.environ.get("CHK", "500"))
sc = W.box_drop_worlds(1, nbox, auto_disable=os.environ.get("AUTO_DISABLE", "1") == "1")
p = S.default_params()
g = Batch(sc, p)
r = refharness.RefBatch(sc, p, mode=1) if os.environ.get("REF", "0") == "1" else None
for k in range(0, steps, chk):
    t = time.perf_counter(); g.step(chk); g.sync(); tg = (time.perf_counter() - t) / chk
    rms, nc, m = g.stats()
    line = "impl %s step %4d: gpu %.3f ms/step stages %s rows %d contacts %d" % (os.environ.get("GZCUDA_PGS_IMPL", "-"), k + chk, tg * 1e3, ["%.2f" % x for x in g.stage_ms()], m[0], nc[0])
    if r is not None:
        t = time.perf_counter(); r.step(chk); tr = (time.perf_counter() - t) / chk
        gs, rs = g.body_state(), r.body_state()
        line += " | cpu %.2f ms/step | max |dpos| %.2e |dq| %.2e pairs equal %s" % (tr * 1e3, np.abs(gs[0] - rs[0]).max(), np.abs(gs[1] - rs[1]).max(), np.array_equal(g.pairs(0), r.pairs(0)))
    line += " | pgs debug %s" % (g.pgs_debug()[0].tolist(),)
    print(line, flush=True)
