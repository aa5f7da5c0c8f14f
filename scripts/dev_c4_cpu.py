"""Developer run (CPU only): the reference's own opende on the C4 rubble pile (oracle/_ref fast flavour, natural order,
dWorldQuickStep, one thread: one world is one thread's work), ms per step as the pile forms.  Needs `ulimit -s unlimited`
for dxHashSpace::collide's n x n / 8-byte `tested` matrix on the stack (collision_space.cpp:486-488)."""
import os, sys, time; sys.path.insert(0, "/root/repo")
from gazebo_b200 import scene as S, worlds as W
from oracle import refharness
n = int(os.environ.get("N", "20000")); steps = int(os.environ.get("STEPS", "2400")); chk = int(os.environ.get("CHK", "400"))
sc = W.rubble_world(n)
r = refharness.RefBatch(sc, S.default_params(), mode=0, flavour="fast")
for k in range(0, steps, chk):
    t = time.perf_counter(); r.step(chk); dt = (time.perf_counter() - t) / chk
    rms, nc, ncj = r.stats()
    print("step %4d: cpu %.1f ms/step, contacts %d" % (k + chk, dt * 1e3, ncj[0]), flush=True)
