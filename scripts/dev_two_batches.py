"""Developer experiment (GPU box): one batch of 4096 worlds against two batches of 2048 stepped side by side
(their kernels overlap across streams): does the assemble stage of one hide under the PGS sweeps of the other?"""
import sys, time; sys.path.insert(0, "/root/repo")
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
def mk(n, first):
    sc, pd = W.revolve_worlds(n, first_world=first)
    b = Batch(sc, S.default_params()); b.set_hinge_pd(pd); return b
for parts in (1, 2, 4):
    n = 4096 // parts
    bs = [mk(n, i * n) for i in range(parts)]
    for _ in range(120):
        for b in bs: b.step(1)
    for b in bs: b.sync()
    best = 1e9
    for rep in range(3):
        t = time.perf_counter()
        for _ in range(200):
            for b in bs: b.step(1)
        for b in bs: b.sync()
        best = min(best, (time.perf_counter() - t) / 200 * 1e3)
    print("%d batch(es) of %d worlds: %.3f ms per step of all 4096 worlds" % (parts, n, best), flush=True)
    del bs
