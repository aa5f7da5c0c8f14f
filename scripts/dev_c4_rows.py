"""Developer run (GPU box): row / contact counts of the 20 000-body pile at a few steps (a fingerprint of the trajectory:
every variant of the large-world kernels must print the same numbers)."""
import os, sys; sys.path.insert(0, "/root/repo")
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
sc = W.rubble_world(int(os.environ.get("N", "20000")))
b = Batch(sc, S.default_params())
out = []
done = 0
for k in [int(x) for x in os.environ.get("AT", "2000,2200,2400").split(",")]:
    b.step(k - done); b.sync(); done = k
    rms, nc, m = b.stats()
    out.append("%d: rows %d contacts %d" % (k, m[0], nc[0]))
print(os.environ.get("TAG", ""), " | ".join(out), flush=True)
