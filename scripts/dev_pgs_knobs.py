"""Developer check (GPU box): PGS time of the headline batch under the current env knobs."""
import os, sys; sys.path.insert(0, "/root/repo")
import numpy as np
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
nw = int(os.environ.get("NW", "4096"))
sc, pd = W.revolve_worlds(nw)
b = Batch(sc, S.default_params()); b.set_hinge_pd(pd)
b.step(100)
b.pgs_profile()
b.set_profiling(True); b.step(20); pgs_ms, n = b.pgs_ms()
pf = b.pgs_profile()
d = b.pgs_debug(); rms, nc, m = b.stats(); ms = b.stage_ms()
R, k = d[:, 1] >> 8, d[:, 1] & 255
tag = " ".join("%s=%s" % (e[7:], os.environ[e]) for e in sorted(os.environ) if e.startswith("GZCUDA_"))
print("%-40s pgs %.3f ms | rounds mean %.0f max %d | cyc/round mean %.0f | cycles mean %.0f max %d | k %s | stage %s" % (
    tag, pgs_ms / n, d[:, 0].mean(), d[:, 0].max(), (d[:, 2] / np.maximum(d[:, 0], 1)).mean(), d[:, 2].mean(), d[:, 2].max(),
    {int(x): int((k == x).sum()) for x in np.unique(k)}, ["%.3f" % x for x in ms]), flush=True)
if os.environ.get("GZCUDA_PGS_PROF"):
    r = max(pf[5], 1)
    if pf[14]:
        print("   chain: %d worlds, run rounds %.0f/world at %.0f cycles (longest run %.1f rows), other rounds %.0f/world at %.0f cycles, sweep cycles %.0f/world" % (
            pf[14] // 20, pf[9] / pf[14], pf[8] / max(pf[9], 1), pf[10] / max(pf[9], 1), pf[12] / pf[14], pf[11] / max(pf[12], 1), pf[13] / pf[14]))
    print("   per round: fetch %.0f wait %.0f load %.0f row %.0f cycles, active lanes %.1f" % (pf[0] / r, pf[1] / r, pf[2] / r, pf[3] / r, pf[4] / r))
