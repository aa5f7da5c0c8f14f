"""Target for `ncu --set full` (GPU box): the headline batch, stepped past its transient so that the profiled
k_pgs / k_pgs_chain launches see the steady-state contact load of bench.py's timed region."""
import os, sys; sys.path.insert(0, "/root/repo")
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
sc, pd = W.revolve_worlds(4096)
b = Batch(sc, S.default_params()); b.set_hinge_pd(pd)
b.step(int(os.environ.get("STEPS", "104"))); b.sync()
