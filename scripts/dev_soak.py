"""Developer run (GPU box): soak of the large-world kernels - C1 (k_pgsm: sleep / wake transitions) and C4 (k_pgsd) for many
thousands of steps: no error, no hang, the number of bodies that left the number line stays what it was."""
import os, sys, time; sys.path.insert(0, "/root/repo")
import numpy as np
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
for name, sc, steps in (("C1 500 boxes", W.box_drop_worlds(1, 500), int(os.environ.get("C1_STEPS", "20000"))),
                        ("C4 20000-body pile", W.rubble_world(20000), int(os.environ.get("C4_STEPS", "10000")))):
    b = Batch(sc, S.default_params())
    t = time.perf_counter(); done = 0
    while done < steps:
        n = min(2500, steps - done); b.step(n); b.sync(); done += n
        pos = b.body_state()[0]
        print("%s: step %6d, %.2f ms/step so far, rows %d, bodies not finite %d, asleep %d" % (
            name, done, (time.perf_counter() - t) / done * 1e3, b.stats()[2][0], int((~np.isfinite(pos).all(axis=1)).sum()), int((b.body_flags() != 0).sum())), flush=True)
    b.close()
