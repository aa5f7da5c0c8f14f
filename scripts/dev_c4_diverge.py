"""Developer run (GPU box): two batches of the same 20 000-body pile stepped in lockstep; reports the first step at which
pair list, contacts or body states differ (a fingerprint of a race: both must stay identical forever)."""
import os, sys; sys.path.insert(0, "/root/repo")
import numpy as np
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
n = int(os.environ.get("N", "20000")); start = int(os.environ.get("START", "1950")); end = int(os.environ.get("END", "2300"))
sc = W.rubble_world(n)
envb = dict(x.split("=") for x in os.environ.get("ENVB", "").split(",") if x)
a = Batch(sc, S.default_params())
os.environ.update(envb)
b = Batch(sc, S.default_params())
a.step(start); b.step(start); a.sync(); b.sync()
for k in range(start, end):
    a.update_collision(); b.update_collision()
    pa, pb = a.pairs(0, 1 << 22), b.pairs(0, 1 << 22)
    ca, cb = a.contacts(0, 1 << 18), b.contacts(0, 1 << 18)
    same_p = np.array_equal(pa, pb)
    same_c = np.array_equal(ca[0], cb[0]) and np.array_equal(ca[1], cb[1])
    a.update_physics(); b.update_physics()
    sa, sb = a.body_state(), b.body_state()
    same_s = all(np.array_equal(x, y) for x, y in zip(sa, sb))
    ma, mb = a.stats()[2][0], b.stats()[2][0]
    if not (same_p and same_c and same_s):
        print("step %d: pairs equal %s (%d vs %d) contacts equal %s (%d vs %d) states equal %s rows %d vs %d" % (k, same_p, len(pa), len(pb), same_c, len(ca[0]), len(cb[0]), same_s, ma, mb), flush=True)
        if not same_p:
            sa_, sb_ = set(map(tuple, pa.tolist())), set(map(tuple, pb.tolist()))
            print("  only in A:", sorted(sa_ - sb_)[:10], " only in B:", sorted(sb_ - sa_)[:10], flush=True)
        if not same_s:
            d = np.abs(sa[0] - sb[0]).max(axis=1)
            print("  bodies differing:", int((d > 0).sum()), "first", np.nonzero(d > 0)[0][:10].tolist(), "max", d.max(), flush=True)
        break
else:
    print("identical through step", end, flush=True)
