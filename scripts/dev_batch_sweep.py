"""Developer run (GPU box): world-steps/s of the headline workload against the batch size per GPU."""
import sys, time; sys.path.insert(0, "/root/repo")
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
for n in (256, 512, 1024, 2048, 4096, 8192, 16384):
    sc, pd = W.revolve_worlds(n)
    b = Batch(sc, S.default_params()); b.set_hinge_pd(pd)
    b.step(120); b.sync()
    b.step(100); b.sync()
    ms, _ = b.last_step_ms()
    print("worlds %6d: %.3f ms/step, %10.0f world-steps/s, stages %s" % (n, ms / 100, n * 100 / (ms * 1e-3), ["%.3f" % x for x in b.stage_ms()]), flush=True)
    b.close()
