"""Developer check (GPU box): how many worlds' sweeps are in flight over the PGS phase of one step."""
import os, sys; sys.path.insert(0, "/root/repo")
import numpy as np
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
sc, pd = W.revolve_worlds(4096)
b = Batch(sc, S.default_params()); b.set_hinge_pd(pd)
b.step(120)
tl = b.pgs_timeline().astype(np.int64); d = b.pgs_debug()
k = d[:, 1] & 255
ok = tl[:, 0] > 0
t0 = tl[ok, 0].min(); st = (tl[:, 0] - t0) / 1e3; en = (tl[:, 1] - t0) / 1e3
print("PGS phase %.0f us; worlds %d (chain %d)" % (en[ok].max(), ok.sum(), (k[ok] == 16).sum()))
for name, sel in (("chain", ok & (k == 16)), ("regular", ok & (k != 16))):
    if sel.any():
        print("  %-8s start p0 %.0f p50 %.0f p90 %.0f p100 %.0f | end p50 %.0f p90 %.0f p100 %.0f | duration mean %.0f max %.0f us" % (
            name, st[sel].min(), np.percentile(st[sel], 50), np.percentile(st[sel], 90), st[sel].max(),
            np.percentile(en[sel], 50), np.percentile(en[sel], 90), en[sel].max(), (en - st)[sel].mean(), (en - st)[sel].max()))
edges = np.linspace(0, en[ok].max(), 13)
for a, c in zip(edges[:-1], edges[1:]):
    mid = (a + c) / 2
    fl = ok & (st <= mid) & (en > mid)
    print("  t=%5.0f us: in flight chain %4d regular worlds %4d" % (mid, (fl & (k == 16)).sum(), (fl & (k != 16)).sum()))
sm = tl[:, 1] & 511
for tq in (100.0, 300.0):
    fl = ok & (st <= tq) & (en > tq)
    cc = np.bincount(sm[fl & (k == 16)], minlength=160)[:148]; rr = np.bincount(sm[fl & (k != 16)], minlength=160)[:148]
    print("  t=%.0f us, per SM: chain warps min %d mean %.1f max %d | SMs by chain warps %s | regular worlds per SM by chain warps %s" % (
        tq, cc.min(), cc.mean(), cc.max(), {int(x): int((cc == x).sum()) for x in np.unique(cc)},
        {int(x): round(float(rr[cc == x].mean()), 1) for x in np.unique(cc)}))
