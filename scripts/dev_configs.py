"""Developer run (GPU box): every BASELINE.json config that fits the current build, GPU (libgzcuda, canonical
order) beside the reference's opende on the host cores (oracle/_ref fast flavour, natural order, all threads)."""
import os, sys, time; sys.path.insert(0, "/root/repo")
import numpy as np
from gazebo_b200 import scene as S, worlds as W
from gazebo_b200.batch import Batch
from oracle import refharness

def run(name, sc, pd, warm, steps, cpu_worlds_scene=None, cpu_pd=None, cpu_steps=None):
    p = S.default_params()
    g = Batch(sc, p)
    if pd is not None: g.set_hinge_pd(pd)
    g.step(warm); g.sync()
    t0 = time.perf_counter(); g.step(steps); g.sync(); tg = time.perf_counter() - t0
    ms, _ = g.last_step_ms()
    rms, nc, m = g.stats()
    line = "%-34s worlds %5d rows/world %7.1f | gpu %9.0f world-steps/s (%.3f ms/step, stages %s)" % (
        name, sc.n_worlds, m.mean(), sc.n_worlds * steps / (ms * 1e-3), ms / steps, ["%.3f" % x for x in g.stage_ms()])
    if refharness.available("fast"):
        csc = cpu_worlds_scene if cpu_worlds_scene is not None else sc
        r = refharness.RefBatch(csc, p, mode=0, flavour="fast")
        if cpu_pd is not None: r.set_hinge_pd(cpu_pd)
        th = os.cpu_count()
        r.step(warm, threads=th)
        n = cpu_steps or steps
        t0 = time.perf_counter(); r.step(n, threads=th); tc = time.perf_counter() - t0
        line += " | cpu %9.0f world-steps/s on %d threads (%d worlds sampled)" % (csc.n_worlds * n / tc, th, csc.n_worlds)
    print(line, flush=True)

sc = W.box_drop_worlds(1, 500); run("C1 500 boxes, 1 world", sc, None, 1500, 200)   # (one world = one CPU thread)
sc, pd = W.revolve_worlds(4096); csc, cpd = W.revolve_worlds(256)
run("C2 4096 8-link robots", sc, pd, 100, 100, csc, cpd, 100)
sc, _ = W.pioneer_worlds(1024, drive=True); csc, _ = W.pioneer_worlds(256, drive=True)
run("C3 1024 pioneer2dx (wheel motors)", sc, None, 100, 200, csc, None, 200)
sc = W.chain_worlds(256, 50); csc = W.chain_worlds(64, 50)
run("C5 256 x 50-link hinge chains", sc, None, 50, 100, csc, None, 50)
