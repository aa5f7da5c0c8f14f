"""Target for ncu (GPU box): the 20 000-body rubble pile stepped until the pile has formed."""
import os, sys; sys.path.insert(0, "/root/repo")
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
sc = W.rubble_world(int(os.environ.get("N", "20000")))
b = Batch(sc, S.default_params())
b.step(int(os.environ.get("STEPS", "2002"))); b.sync()
