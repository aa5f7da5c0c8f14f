#!/usr/bin/env python
"""bench.py — batched world-steps/s of the ODE quickstep hot path (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            our arm (libgzcuda on B200)
  python bench.py --impl reference --gpus N --steps K ...  the reference's own CPU code

Workload: BASELINE.json configs[1] — 4096 independent Revolve-style 8-link robot worlds on a
plane PER GPU (weak scaling: worlds never couple, so each rank owns its own 4096 worlds and
there is no data-path collective).  A "step" is one world step (UpdateCollision +
UpdatePhysics) of every world of the batch.  One JSON line is printed by rank 0.

  value  world-steps/s with all state resident in HBM (CUDA events, max over ranks)
  e2e    the same through gzcuda_batch_step_io with HOST buffers: every step uploads a
         per-body wrench array and downloads every body's pose + velocity
  roofline      k_pgsb: algorithmic bytes (288 B per row per sweep) / launch time
  strong        (N > 1) the same 4096 worlds split over the N ranks
  --config c1   a line for BASELINE.json configs[0] (ONE world: 500 boxes), same keys
  --config c4   a line for BASELINE.json configs[3] (ONE world: a 20 000-body rubble pile), same keys
  cpu_baseline  the reference's opende (oracle/_ref, -O3 -msse2 -DODE_SSE, natural gzserver
                order, dWorldQuickStep) on this box's host cores, bounded sample
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORLDS_PER_GPU = 4096
METRIC = "batched world-steps/sec (8-link robots)"
UNIT = "world-steps/s"
CPU_PHASES = None


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """Samples nvidia-smi clocks and throttle reasons during the timed region."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.check_output(
                    ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                    timeout=5).decode().strip()
                self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].startswith("Active") for r in self.rows)]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


def cpu_reference_rates(n_worlds, steps, warmup, threads, reps=3, flavour="fast"):
    """World-steps/s of the reference's CPU path (natural order, dWorldQuickStep): all worlds dealt to `threads`
    persistent host threads (every thread calls dAllocateODEDataForThread once), `reps` timed windows of
    `steps` steps each after `warmup` steps.  Returns (rates per window, seconds per window, flavour)."""
    from gazebo_b200 import scene as S, worlds as W
    from oracle import refharness
    if not refharness.available(flavour):
        flavour = "parity"
    sc, pd = W.revolve_worlds(n_worlds)
    r = refharness.RefBatch(sc, S.default_params(), mode=0, flavour=flavour)
    r.set_hinge_pd(pd)
    r.step(warmup, threads=threads)
    r.phase_seconds(reset=True)
    rates, secs = [], []
    for _ in range(reps):
        t0 = time.perf_counter()
        r.step(steps, threads=threads)
        dt = time.perf_counter() - t0
        rates.append(n_worlds * steps / dt)
        secs.append(dt)
    global CPU_PHASES
    col, phy = r.phase_seconds()
    CPU_PHASES = {"UpdateCollision": col / max(col + phy, 1e-30), "UpdatePhysics": phy / max(col + phy, 1e-30),
                  "note": "share of the host threads' time inside the two calls of the path (host clock around them in the harness)"}
    return rates, secs, flavour


def run_reference(args, rank, world):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    n_worlds = args.worlds   # the SAME batch as one GPU of our arm steps (config[1]: 4096 worlds)
    # each "step" of this arm = one world step of all worlds of the batch on all host threads; three windows of
    # K steps, value = the median window
    rates, secs, flavour = cpu_reference_rates(n_worlds, args.steps, args.warmup, cores)
    order = sorted(range(len(rates)), key=lambda i: rates[i])
    mid = order[len(order) // 2]
    rate, dt = rates[mid], secs[mid]
    sample = "%d worlds x %d steps on %d persistent threads, median of %d windows (best %.0f) (oracle/_ref %s flavour, natural order, dWorldQuickStep)" % (
        n_worlds, args.steps, cores, len(rates), max(rates), flavour)
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "configs[1]: %d Revolve-style 8-link robot worlds on a plane, ODE quickstep 50+10 iters, dt=1ms "
                               "(the batch ONE GPU of the gzcuda arm steps; the host has one set of cores whatever --gpus says)" % n_worlds,
                   "worlds": n_worlds, "host_threads": cores, "windows": rates},
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample, "phases": CPU_PHASES},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def run_gpu(args, rank, world, local_rank):
    import torch
    from gazebo_b200 import scene as S, worlds as W
    from gazebo_b200.batch import Batch

    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist = dist_mod
        torch.cuda.set_device(local_rank)
        dist.init_process_group("gloo")   # start barrier + one scalar max only: worlds never couple, NCCL is not used

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    from gazebo_b200 import sharding

    def max_over_ranks(x):
        return sharding.max_over_ranks(x, dist, "cpu")

    n_worlds = args.worlds
    first_world, _ = sharding.world_block(rank, n_worlds)
    sc, pd = W.revolve_worlds(n_worlds, first_world=first_world)
    p = S.default_params()
    b = Batch(sc, p, device=local_rank)
    b.set_hinge_pd(pd)
    K, Wm = args.steps, args.warmup

    # ---- kernel-resident number: K steps in one call, state never leaves HBM
    b.step(Wm)
    b.sync()
    b.counters(reset=True)
    b.set_profiling(True)
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    b.step(K)
    b.sync()
    barrier()
    ms, launches = b.last_step_ms()
    pgs_ms, pgs_launches = b.pgs_ms()
    cnt = b.counters()
    stage = b.stage_ms()
    b.set_profiling(False)
    ms = max_over_ranks(ms)
    value = sharding.job_throughput(n_worlds, world, K, ms * 1e-3)

    # ---- end to end through the host-buffer ABI call: pinned wrench in, pinned state out, every step
    nb = sc.nb
    wrench = torch.zeros((nb, 6), dtype=torch.float64).pin_memory()
    state = torch.empty((nb, 13), dtype=torch.float64).pin_memory()
    for _ in range(Wm):
        b.step_io(1, wrench.data_ptr(), state.data_ptr())
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        b.step_io(1, wrench.data_ptr(), state.data_ptr())
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    sampler.stop_flag = True
    sampler.join(timeout=2)
    e2e_value = sharding.job_throughput(n_worlds, world, K, e2e_s)
    finite = bool(torch.isfinite(state).all())

    # ---- the stated multi-GPU configuration read literally (BASELINE.json configs[1]: "4096 ... worlds, batched
    # across 8 x B200"): the same 4096 worlds SPLIT over the ranks (strong scaling), device-resident, K steps
    strong = None
    if world > 1:
        s_first, s_n = sharding.strong_block(rank, world, WORLDS_PER_GPU)
        sc2, pd2 = W.revolve_worlds(s_n, first_world=s_first)
        b2 = Batch(sc2, p, device=local_rank)
        b2.set_hinge_pd(pd2)
        b2.step(Wm)
        b2.sync()
        barrier()
        b2.step(K)
        b2.sync()
        barrier()
        ms2 = max_over_ranks(b2.last_step_ms()[0])
        strong = {"value": WORLDS_PER_GPU * K / (ms2 * 1e-3), "unit": UNIT, "ms_per_step": ms2 / K,
                  "worlds_total": WORLDS_PER_GPU, "worlds_per_gpu": s_n}
        b2.close()

    if rank != 0:
        return
    peak, peak_kind = load_peaks()
    rows, fric, contacts, wsteps = cnt
    iters, fiters = p.iters, p.friction_iters
    bytes_pgs = 288.0 * ((rows - fric) * iters + fric * (iters + fiters))   # rows sweep 50x, friction rows 60x
    bytes_per_launch = bytes_pgs / max(pgs_launches, 1)
    pgs_avg_ms = pgs_ms / max(pgs_launches, 1)
    achieved = bytes_per_launch / (pgs_avg_ms * 1e-3) / 1e9 if pgs_avg_ms > 0 else 0.0
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "pgs_traffic.json")) as f:
            traffic = json.load(f).get("dram_bytes_per_launch")
    except Exception:
        pass
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {
            "workload": "configs[1]: %d Revolve-style 8-link robot worlds on a plane per GPU (value and e2e are WEAK scaling: every rank "
                        "owns its own %d worlds; the same 4096 worlds split over the ranks are under `strong` when N > 1), ODE quickstep "
                        "50+10 iters, dt=1ms, PD-driven sinusoid gaits, deterministic row-order mode" % (n_worlds, n_worlds),
            "worlds_per_gpu": n_worlds, "bodies_per_gpu": sc.nb, "geoms_per_gpu": sc.ng, "joints_per_gpu": sc.nj,
            "rows_per_world_step": rows / max(wsteps, 1), "contacts_per_world_step": contacts / max(wsteps, 1),
            "l2": "inputs larger than L2, no flush: a step touches bodies+geoms+joints (%.0f MB), the row array and the "
                  "PGS entry pool (~%.0f MB at %.0f rows/world), more than the 126 MB L2, so a step does not find its "
                  "inputs in L2 from the previous one" % ((sc.nb * 704 + sc.ng * 448 + sc.nj * 432) / 1e6,
                                                          (192 + 1.8 * 184) * rows / max(wsteps, 1) * n_worlds / 1e6, rows / max(wsteps, 1)),
            "stage_ms_last_step": {"collide": stage[0], "assemble": stage[1], "pgs": stage[2], "integrate": stage[3]},
            "finite": finite,
        },
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(nb * 6 * 8), "d2h_bytes_per_step": int(nb * 13 * 8),
                "api": "gzcuda_batch_step_io (pinned host buffers, 1 step per call)"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": "k_pgsb (the PGS sweeps: one launch per shared-memory class, side by side, timed together "
                                                  "as one launch)", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak if peak else None, "traffic": traffic, "peak_kind": peak_kind,
                     "bytes_per_launch": bytes_per_launch, "avg_launch_ms": pgs_avg_ms,
                     "share_of_step": pgs_ms / ms if ms > 0 else None,
                     "dram_achieved": (traffic / (pgs_avg_ms * 1e-3) / 1e9) if (traffic and pgs_avg_ms > 0) else None,
                     "limiter": "instruction issue and the exact-order dependent FP64 chain per body (ncu: profiles/), not memory",
                     "note": "achieved = ALGORITHMIC bytes (288 B per row per sweep, SURVEY 8d) / CUDA-event time of the PGS "
                             "sweeps; the entries are staged in shared memory once per step and stay there for the 60 sweeps, "
                             "so measured DRAM traffic (traffic, per launch, from profiles/; dram_achieved = traffic / time) "
                             "is far lower: this kernel is not HBM-bound and frac is not an HBM utilisation (DESIGN.md 4)"},
        "clocks": sampler.summary(),
    }
    # CPU baseline on this box's host cores (rank 0, N=1 only), bounded sample
    if world == 1 and not args.no_cpu_baseline:
        try:
            cores = os.cpu_count() or 1
            nw = n_worlds     # the same batch
            steps = 300       # x 3 windows: ~10-30 s of CPU work on all cores
            rates, secs, flavour = cpu_reference_rates(nw, steps, 100, cores)
            rate = sorted(rates)[len(rates) // 2]
            line["cpu_baseline"] = {
                "value": rate, "unit": UNIT, "cores": cores, "kind": "reference",
                "sample": "%d worlds x %d steps, median of %d windows (%.1f s each) on %d persistent threads (oracle/_ref %s flavour: "
                          "the reference's own opende, natural order, dWorldQuickStep)" % (nw, steps, len(rates), sorted(secs)[len(secs) // 2], cores, flavour),
                "phases": CPU_PHASES}
        except Exception as e:   # the oracle library did not travel: say so instead of inventing a number
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "reference",
                                    "sample": "unavailable: %s" % e}
    if strong is not None:
        line["strong"] = strong
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def run_c4(args, rank, world, local_rank):
    """--config c1: BASELINE.json configs[0], ONE world - 500 boxes dropped on the plane, auto-disable on, timed once the
    pile has settled (warm-up >= 1500 steps); --config c4: BASELINE.json configs[3], ONE world - a 20 000-body rubble pile in a walled pen - on one GPU (under
    torchrun every rank steps its own replica: a single world does not shard).  The timed region starts once the pile
    has formed (warm-up >= 2500 steps).  cpu_baseline: the reference's opende created from the GPU's body states at that
    moment (it would need ~20 minutes to get there itself) and timed over a few steps on one thread."""
    import resource
    import torch
    from gazebo_b200 import scene as S, worlds as W
    from gazebo_b200.batch import Batch
    try:   # dxHashSpace::collide's n x n / 8-byte matrix on the stack (collision_space.cpp:486-488): 50 MB at C4
        resource.setrlimit(resource.RLIMIT_STACK, (resource.RLIM_INFINITY, resource.RLIM_INFINITY))
    except Exception:
        pass
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist = dist_mod
        torch.cuda.set_device(local_rank)
        dist.init_process_group("gloo")
    from gazebo_b200 import sharding

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    c1 = args.config == "c1"
    n_bodies = 500 if c1 else args.bodies
    sc = W.box_drop_worlds(1, n_bodies) if c1 else W.rubble_world(n_bodies)
    p = S.default_params()
    b = Batch(sc, p, device=local_rank)
    K, Wm = args.steps, max(args.warmup, 1500 if c1 else 2500)
    b.step(Wm)
    b.sync()
    b.counters(reset=True)
    b.set_profiling(True)
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    b.step(K)
    b.sync()
    barrier()
    ms, launches = b.last_step_ms()
    pgs_ms, pgs_launches = b.pgs_ms()
    cnt = b.counters()
    stage = b.stage_ms()
    b.set_profiling(False)
    ms = sharding.max_over_ranks(ms, dist, "cpu")
    value = world * K / (ms * 1e-3)
    nb = sc.nb
    wrench = torch.zeros((nb, 6), dtype=torch.float64).pin_memory()
    state = torch.empty((nb, 13), dtype=torch.float64).pin_memory()
    for _ in range(3):
        b.step_io(1, wrench.data_ptr(), state.data_ptr())
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        b.step_io(1, wrench.data_ptr(), state.data_ptr())
    torch.cuda.synchronize()
    e2e_s = sharding.max_over_ranks(time.perf_counter() - t0, dist, "cpu")
    barrier()
    sampler.stop_flag = True
    sampler.join(timeout=2)
    if rank != 0:
        return
    peak, peak_kind = load_peaks()
    rows, fric, contacts, wsteps = cnt
    bytes_pgs = 288.0 * ((rows - fric) * p.iters + fric * (p.iters + p.friction_iters))
    bytes_per_launch = bytes_pgs / max(pgs_launches, 1)
    pgs_avg_ms = pgs_ms / max(pgs_launches, 1)
    achieved = bytes_per_launch / (pgs_avg_ms * 1e-3) / 1e9 if pgs_avg_ms > 0 else 0.0
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "pgs_traffic_c1.json" if c1 else "pgs_traffic_c4.json")) as f:
            traffic = json.load(f).get("dram_bytes_per_launch")
    except Exception:
        pass
    unit = "world-steps/s"
    line = {
        "metric": "world-steps/sec (500 boxes, one world)" if c1 else "world-steps/sec (20k-body rubble pile, one world)", "value": value, "unit": unit, "n_gpus": world, "steps": K,
        "warmup": Wm, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "replicas only", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": ("configs[0]: ONE world, %d boxes dropped on a plane, auto-disable on, timed settled (one island of ~340 bodies "
                                "among ~120 awake ones); ODE quickstep 50+10 iters, dt=1ms, deterministic row-order mode; N > 1 = N replicas" % n_bodies) if c1 else
                               ("configs[3]: ONE world, %d-body mixed box / sphere / capsule rubble pile in a walled pen on a plane, one "
                                "island once formed; ODE quickstep 50+10 iters, dt=1ms, deterministic row-order mode; a single world does "
                                "not shard: N > 1 = N replicas" % n_bodies),
                   "bodies": sc.nb, "geoms": sc.ng, "rows_per_step": rows / max(wsteps, 1), "contacts_per_step": contacts / max(wsteps, 1),
                   "l2": "no flush: the step's working set (entries 160 B x ~170 k, rows, contacts, bodies, geoms: ~90 MB) is below "
                         "the 126 MB L2 and is rewritten every step; DRAM traffic is in roofline.traffic",
                   "stage_ms_last_step": {"collide": stage[0], "assemble": stage[1], "pgs": stage[2], "integrate": stage[3]},
                   "finite_bodies": int(torch.isfinite(state).all(dim=1).sum())},
        "e2e": {"value": world * K / e2e_s, "unit": unit, "h2d_bytes_per_step": int(nb * 6 * 8), "d2h_bytes_per_step": int(nb * 13 * 8),
                "api": "gzcuda_batch_step_io (pinned host buffers, 1 step per call)"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": "k_pgsm (the PGS sweeps on one thread-block cluster, entries in shared memory)" if c1 else
                     "k_pgsd (the PGS sweeps of the pile as dataflow over the whole GPU)", "achieved": achieved,
                     "peak": peak, "unit": "GB/s", "frac": achieved / peak if peak else None, "traffic": traffic, "peak_kind": peak_kind,
                     "bytes_per_launch": bytes_per_launch, "avg_launch_ms": pgs_avg_ms, "share_of_step": pgs_ms / ms if ms > 0 else None,
                     "dram_achieved": (traffic / (pgs_avg_ms * 1e-3) / 1e9) if (traffic and pgs_avg_ms > 0) else None,
                     "limiter": ("the exact-order dependency chain through the pile: ~3 100 hops of ~590 cycles (push into the partner CTA's "
                                 "shared memory -> poll -> row update) per step, not bytes (DESIGN.md 4)") if c1 else
                                ("the exact-order dependency chain through the pile: ~2 500 hops of ~1 500 cycles (publish -> L2 -> poll -> "
                                 "row update) per step, not bytes (DESIGN.md 4)"),
                     "note": "achieved = ALGORITHMIC bytes (288 B per row per sweep, SURVEY 8d) / CUDA-event time of the sweeps"},
        "clocks": sampler.summary(),
    }
    if world == 1 and not args.no_cpu_baseline:
        try:
            from oracle import refharness
            flavour = "fast" if refharness.available("fast") else "parity"
            if c1:   # the reference gets there by itself in a few seconds
                r = refharness.RefBatch(sc, p, mode=0, flavour=flavour)
                r.step(Wm)
                n_cpu = 1000
                t0 = time.perf_counter(); r.step(n_cpu); dt = time.perf_counter() - t0
                line["cpu_baseline"] = {"value": n_cpu / dt, "unit": unit, "cores": 1, "kind": "reference",
                                        "sample": "%d steps (%.1f s) after %d of the reference's own opende (oracle/_ref %s flavour, natural order, "
                                                  "dWorldQuickStep) on ONE thread (one world is one thread's work)" % (n_cpu, dt, Wm, flavour)}
                raise StopIteration
            pos, quat, lvel, avel = b.body_state()
            bad = ~(np.isfinite(pos).all(axis=1) & np.isfinite(quat).all(axis=1) & np.isfinite(lvel).all(axis=1) & np.isfinite(avel).all(axis=1))
            park = np.nonzero(bad | (np.abs(np.nan_to_num(pos)).max(axis=1) > 1e3))[0]
            for k, bi in enumerate(park):   # bodies shot out of the pile on the way: parked far away
                pos[bi] = (1000.0 + 2.0 * k, 1000.0, 1000.0); quat[bi] = (1, 0, 0, 0); lvel[bi] = 0; avel[bi] = 0
            A = sc.arrays
            A["body_pos"][:] = pos; A["body_quat"][:] = quat; A["body_lvel"][:] = lvel; A["body_avel"][:] = avel
            r = refharness.RefBatch(sc, p, mode=0, flavour=flavour)
            r.step(1)
            n_cpu = 20
            t0 = time.perf_counter(); r.step(n_cpu); dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": n_cpu / dt, "unit": unit, "cores": 1, "kind": "reference",
                                    "sample": "%d steps (%.1f s) of the reference's own opende (oracle/_ref %s flavour, natural order, dWorldQuickStep) "
                                              "on ONE thread (one world is one thread's work), created from the GPU's body states after %d steps "
                                              "(%d bodies parked)" % (n_cpu, dt, flavour, Wm + 2 * K + 3, len(park))}
        except StopIteration:
            pass
        except Exception as e:
            line["cpu_baseline"] = {"value": None, "unit": unit, "cores": 1, "kind": "reference", "sample": "unavailable: %s" % e}
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="gzcuda", choices=["gzcuda", "reference"])
    ap.add_argument("--worlds", type=int, default=WORLDS_PER_GPU, help="worlds per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--config", default="c2", choices=["c2", "c1", "c4"], help="c2: BASELINE.json configs[1] (the headline); c1: configs[0], 500 boxes in one world; c4: configs[3], one 20k-body pile")
    ap.add_argument("--bodies", type=int, default=20000, help="--config c4: bodies of the pile")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under it so that one process drives each GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    if args.config in ("c1", "c4"):
        run_c4(args, rank, world, local_rank)
        return
    run_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
