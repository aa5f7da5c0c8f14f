"""Developer run (GPU box): C4 - one world, a 20 000-body rubble pile - time per step on the GPU once the pile has
formed, and the reference's own opende (oracle/_ref fast flavour, natural order, dWorldQuickStep, one thread: one world is
one thread's work) ON THE SAME PILE: the reference would need ~20 minutes to get there by itself, so it is created from
the GPU's body states at that step (contacts are new joints every step: nothing else carries over) and timed over a few
steps.  Bodies that left the number line on the way (a handful, shot out of the pile: NaN) are parked far away first."""
import os, sys, time, resource; sys.path.insert(0, "/root/repo")
import numpy as np
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
from oracle import refharness
try:   # dxHashSpace::collide keeps an n x n / 8-byte matrix on the stack (collision_space.cpp:486-488): 50 MB at C4
    resource.setrlimit(resource.RLIMIT_STACK, (resource.RLIM_INFINITY, resource.RLIM_INFINITY))
except Exception as e:
    print("stack limit not raised:", e)
n = int(os.environ.get("N", "20000")); warm = int(os.environ.get("WARM", "3000")); steps = int(os.environ.get("STEPS", "200")); cpu_steps = int(os.environ.get("CPU_STEPS", "5"))
sc = W.rubble_world(n)
p = S.default_params()
g = Batch(sc, p)
g.step(warm); g.sync()
t0 = time.perf_counter(); g.step(steps); g.sync(); tg = (time.perf_counter() - t0) / steps
ms, _ = g.last_step_ms()
rms, nc, m = g.stats()
print("C4 %d bodies after %d steps: rows %d contacts %d | gpu %.3f ms/step (device %.3f), stages %s" % (n, warm + steps, m[0], nc[0], tg * 1e3, ms / steps, ["%.2f" % x for x in g.stage_ms()]), flush=True)
if refharness.available("fast") and cpu_steps > 0:
    pos, quat, lvel, avel = g.body_state()
    bad = ~(np.isfinite(pos).all(axis=1) & np.isfinite(quat).all(axis=1) & np.isfinite(lvel).all(axis=1) & np.isfinite(avel).all(axis=1))
    far = np.abs(pos).max(axis=1) > 1e3
    park = np.nonzero(bad | far)[0]
    for k, b in enumerate(park):
        pos[b] = (1000.0 + 2.0 * k, 1000.0, 1000.0); quat[b] = (1, 0, 0, 0); lvel[b] = 0; avel[b] = 0
    A = sc.arrays
    A["body_pos"][:] = pos; A["body_quat"][:] = quat; A["body_lvel"][:] = lvel; A["body_avel"][:] = avel
    r = refharness.RefBatch(sc, p, mode=0, flavour="fast")
    r.step(1)
    t0 = time.perf_counter(); r.step(cpu_steps); tc = (time.perf_counter() - t0) / cpu_steps
    print("   reference opende on the same pile (%d bodies parked): contacts %d, %.1f ms/step on 1 thread -> gpu %.0fx" % (len(park), r.stats()[2][0], tc * 1e3, tc / tg), flush=True)
