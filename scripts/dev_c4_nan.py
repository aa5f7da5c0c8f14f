"""Developer run (GPU box): steps the pile to the first NaN in the body states and prints what touches those bodies."""
import os, sys; sys.path.insert(0, "/root/repo")
import numpy as np
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
n = int(os.environ.get("N", "20000")); start = int(os.environ.get("START", "2080")); end = int(os.environ.get("END", "2110"))
sc = W.rubble_world(n, pen=float(os.environ["PEN"]) if "PEN" in os.environ else None)
a = Batch(sc, S.default_params())
a.step(start); a.sync()
prev = a.body_state()
for k in range(start, end):
    a.update_collision()
    pa = a.pairs(0, 1 << 22); ca = a.contacts(0, 1 << 18)
    a.update_physics()
    st = a.body_state()
    bad = np.nonzero(~np.isfinite(st[0]).all(axis=1) | ~np.isfinite(st[2]).all(axis=1))[0]
    rms, nc, m = a.stats()
    if len(bad) or k % 100 == 0: print("step %d rows %d contacts %d pairs %d nan bodies %d rms %s dbg %s" % (k, m[0], nc[0], len(pa), len(bad), rms[0].tolist() if hasattr(rms[0], "tolist") else rms[0], a.pgs_debug()[0].tolist()), flush=True)
    if len(bad):
        print("nan bodies:", bad[:20].tolist())
        g = ca[0]
        print(' velocities before:', [(int(bb), np.abs(prev[2][bb]).max(), np.abs(prev[3][bb]).max()) for bb in bad[:8]])
        fast = np.argsort(-np.abs(prev[2]).max(axis=1))[:5]
        print(' fastest bodies before:', [(int(f), float(np.abs(prev[2][f]).max())) for f in fast])
        for bb in bad[:3].tolist():
            gi = bb + 5
            sel = np.nonzero((g[:, 0] == gi) | (g[:, 1] == gi))[0]
            print(" body", bb, "prev pos", prev[0][bb].tolist(), "prev lvel", prev[2][bb].tolist(), "prev avel", prev[3][bb].tolist(), "prev q", prev[1][bb].tolist(), "geom type/dims", sc.geom_type[gi] if hasattr(sc, "geom_type") else None, "contacts:", [(g[i].tolist(), ca[1][i].tolist()) for i in sel[:6]])
        # duplicates in the pair list?
        pp = pa[:, 0].astype(np.int64) * (1 << 20) + pa[:, 1]
        print("pairs sorted:", bool(np.all(np.diff(pp) > 0)), "dups:", int((np.diff(np.sort(pp)) == 0).sum()))
        break
    prev = st
