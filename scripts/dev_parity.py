"""Developer check (GPU box): libgzcuda vs the compiled reference on the same scenes."""
import sys, time
import numpy as np
sys.path.insert(0, "/root/repo")
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
from oracle.refharness import RefBatch

def cmp(name, sc, p, pd=None, steps=(1, 1, 8, 90, 900)):
    g = Batch(sc, p); r = RefBatch(sc, p, mode=1)
    if pd is not None:
        g.set_hinge_pd(pd); r.set_hinge_pd(pd)
    tot = 0
    for n in steps:
        g.step(n); r.step(n); tot += n
        gs = g.body_state(); rs = r.body_state()
        d = [np.abs(a - b).max() for a, b in zip(gs, rs)]
        rg, ncg, mg = g.stats(); rr, ncr, ncjr = r.stats()
        pe = all(np.array_equal(g.pairs(w), r.pairs(w)) for w in range(min(sc.n_worlds, 4)))
        ce = True
        for w in range(min(sc.n_worlds, 4)):
            gg, gd = g.contacts(w); rgm, rd = r.contacts(w)
            ce = ce and np.array_equal(gg, rgm) and (gd.shape == rd.shape) and (np.abs(gd - rd).max() if gd.size else 0) < 1e-12
        print("%-10s step %5d  dpos %.3e dq %.3e dv %.3e dw %.3e | nc gpu/ref %s/%s pairs_eq %s contacts_eq %s | res gpu %.3e ref %.3e" % (
            name, tot, d[0], d[1], d[2], d[3], ncg[:4], ncjr[:4], pe, ce, rg[0, 3], rr[0, 3]), flush=True)

p = S.default_params()
sb = S.SceneBuilder(); sb.begin_world(); sb.add_ground_plane()
b = sb.add_body([0, 0, 0.6], S.q_from_rpy(0.3, 0.2, 0.1), mass=1.0, inertia=S.box_inertia(1, 1, 1, 1))
sb.add_geom(S.GEOM_BOX, (1, 1, 1), sb.new_space(), body=b); sb.end_world()
cmp("boxdrop", sb.finish(), p)
cmp("chain", W.chain_worlds(2, 10), p)
sc, pd = W.revolve_worlds(8)
cmp("revolve0", sc, p)          # no actuation
cmp("revolve", sc, p, pd)
cmp("boxes", W.box_drop_worlds(1, 40), p, steps=(1, 1, 8, 90, 400))
