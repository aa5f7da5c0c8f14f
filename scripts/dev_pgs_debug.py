import sys; sys.path.insert(0, "/root/repo")
import numpy as np
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
sc, pd = W.revolve_worlds(4096)
b = Batch(sc, S.default_params()); b.set_hinge_pd(pd)
for tot in (5, 60, 150):
    b.step(tot)
    d = b.pgs_debug(); rms, nc, m = b.stats()
    ms = b.stage_ms()
    print("step", b.counters()[3] // 4096, "pgs ms", ms[2])
    for name, v in (("rows", m), ("loops", d[:, 0]), ("execs", d[:, 1]), ("cycles", d[:, 2]), ("cyc/exec", d[:, 2] / np.maximum(d[:, 1], 1)), ("loops/exec", d[:,0]/np.maximum(d[:,1],1))):
        print("  %-10s mean %10.1f  p50 %10.1f  p90 %10.1f  p99 %10.1f  max %10.1f" % (name, v.mean(), np.percentile(v, 50), np.percentile(v, 90), np.percentile(v, 99), v.max()))
