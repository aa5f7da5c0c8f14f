"""Developer run (GPU box): the large-world path (k_pgsl, cooperative grid) on a rubble pile of N bodies against the
reference's own opende (oracle/_ref, canonical order): parity along the way, then time per step on both sides."""
import os, sys, time; sys.path.insert(0, "/root/repo")
import numpy as np
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
from oracle import refharness
n = int(os.environ.get("N", "1000")); steps = int(os.environ.get("STEPS", "300")); chk = int(os.environ.get("CHK", "100"))
sc = W.rubble_world(n, pen=float(os.environ["PEN"]) if "PEN" in os.environ else None)
p = S.default_params()
t = time.perf_counter(); g = Batch(sc, p); print("bodies %d geoms %d; batch created in %.1f s" % (sc.nb, sc.ng, time.perf_counter() - t), flush=True)
r = refharness.RefBatch(sc, p, mode=1) if os.environ.get("REF", "1") == "1" else None
for k in range(0, steps, chk):
    t = time.perf_counter(); g.step(chk); g.sync(); tg = (time.perf_counter() - t) / chk
    rms, nc, m = g.stats()
    line = "step %4d: gpu %.2f ms/step stages %s rows %d contacts %d" % (k + chk, tg * 1e3, ["%.2f" % x for x in g.stage_ms()], m[0], nc[0])
    if r is not None:
        t = time.perf_counter(); r.step(chk); tr = (time.perf_counter() - t) / chk
        gs, rs = g.body_state(), r.body_state()
        line += " | cpu %.2f ms/step | max |dpos| %.2e |dq| %.2e pairs equal %s" % (tr * 1e3, np.abs(gs[0] - rs[0]).max(), np.abs(gs[1] - rs[1]).max(), np.array_equal(g.pairs(0), r.pairs(0)))
    line += " | nan bodies %d" % int((~np.isfinite(g.body_state()[0]).all(axis=1)).sum())
    line += " | pgs debug %s" % (g.pgs_debug()[0].tolist(),)
    print(line, flush=True)
