import sys; sys.path.insert(0, "/root/repo")
import numpy as np
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
from tests import scenes
for name, builder in (("box_drop", scenes.box_drop), ("tilted", scenes.tilted_box), ("chain2x10", scenes.chain), ("stack", scenes.stack)):
    sc, p, pd = builder()
    b = Batch(sc, p)
    b.step(1400 if name != "chain2x10" else 50)
    d = b.pgs_debug(); rms, nc, m = b.stats()
    print(name, "rows", m, "loops", d[:, 0], "execs", d[:, 1], "cycles", d[:, 2], "cyc/exec", d[:, 2] / np.maximum(d[:, 1], 1), "stage ms", b.stage_ms())
