"""Target for ncu (GPU box): C1 (500 boxes, one world) stepped until the pile has settled."""
import os, sys; sys.path.insert(0, "/root/repo")
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
sc = W.box_drop_worlds(1, 500)
b = Batch(sc, S.default_params())
b.step(int(os.environ.get("STEPS", "1502"))); b.sync()
