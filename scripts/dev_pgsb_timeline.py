"""(pad[0] = passes of the world's warp) Developer check (GPU box): timeline of k_pgsb over one step of the headline batch: when every world's CTA
started and ended, its rounds, and the cycles per round that follow."""
import os, sys; sys.path.insert(0, "/root/repo")
import numpy as np
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
nw = int(os.environ.get("NW", "4096"))
sc, pd = W.revolve_worlds(nw)
b = Batch(sc, S.default_params()); b.set_hinge_pd(pd)
b.step(120)
tl = b.pgs_timeline().astype(np.int64); d = b.pgs_debug()
rounds = d[:, 0]; R = d[:, 1] >> 8; staged = d[:, 1] & 1
t0 = tl[:, 0].min(); st = (tl[:, 0] - t0) / 1e3; en = ((tl[:, 1] & ~511) - t0) / 1e3
dur = en - st
print("PGS phase %.0f us; worlds %d staged %d; stage_ms %s" % (en.max(), nw, staged.sum(), ["%.3f" % x for x in b.stage_ms()]))
print("start p0 %.0f p50 %.0f p90 %.0f p100 %.0f | end p50 %.0f p90 %.0f p100 %.0f | duration mean %.0f max %.0f us" % (
    st.min(), np.percentile(st, 50), np.percentile(st, 90), st.max(), np.percentile(en, 50), np.percentile(en, 90), en.max(), dur.mean(), dur.max()))
cyc = dur * 1965.0 / np.maximum(rounds, 1)
for name, sel in (("staged", staged == 1), ("pool", staged == 0)):
    if sel.any():
        print("  %-7s worlds %5d rounds mean %.0f max %d | cycles/round p10 %.0f p50 %.0f p90 %.0f | duration p50 %.0f p90 %.0f max %.0f" % (
            name, sel.sum(), rounds[sel].mean(), rounds[sel].max(), np.percentile(cyc[sel], 10), np.percentile(cyc[sel], 50),
            np.percentile(cyc[sel], 90), np.percentile(dur[sel], 50), np.percentile(dur[sel], 90), dur[sel].max()))
for w in np.argsort(-en)[:6]:
    print("  last to end: world %d start %.0f end %.0f rounds %d R %d staged %d cycles/round %.0f" % (w, st[w], en[w], rounds[w], R[w], staged[w], cyc[w]))
edges = np.linspace(0, en.max(), 11)
for a, c in zip(edges[:-1], edges[1:]):
    mid = (a + c) / 2
    print("  t=%5.0f us: worlds in flight %4d" % (mid, ((st <= mid) & (en > mid)).sum()))
