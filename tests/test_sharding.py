"""N > 1 host logic on CPU: two gloo ranks partition the worlds, agree on the slowest rank's time
and the shards reproduce exactly the worlds a single rank would have built."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gazebo_b200 import sharding, worlds as W

PER_GPU = 3
FIELDS = ["body_pos", "body_quat", "body_mass", "body_inertia", "geom_dims", "geom_pos", "joint_anchor1", "joint_axis1"]


def _worker(rank, world_size, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    first, n = sharding.world_block(rank, PER_GPU)
    sc, pd = W.revolve_worlds(n, first_world=first)
    # every rank reports a different device time; all must agree on the maximum
    slowest = sharding.max_over_ranks(1.0 + rank, dist)
    value = sharding.job_throughput(PER_GPU, world_size, 10, slowest)
    gathered = [None] * world_size
    dist.all_gather_object(gathered, {f: np.array(getattr(sc, f)) for f in FIELDS} | {"pd": pd, "first": first})
    if rank == 0:
        out.put((slowest, value, gathered))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_partition_worlds_and_aggregate():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29500 + os.getpid() % 400
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    slowest, value, gathered = out.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert slowest == 2.0
    assert value == PER_GPU * 2 * 10 / 2.0
    assert [g["first"] for g in gathered] == [0, PER_GPU]
    whole, pd_whole = W.revolve_worlds(2 * PER_GPU)
    for f in FIELDS:
        both = np.concatenate([g[f] for g in gathered])
        assert np.array_equal(both, np.array(getattr(whole, f))), f
    assert np.array_equal(np.concatenate([g["pd"] for g in gathered]), pd_whole)


def test_strong_block_splits_the_stated_batch():
    """BASELINE.json configs[1] read literally: 4096 worlds split over the ranks."""
    for n in (1, 2, 3, 4, 8):
        blocks = [sharding.strong_block(r, n, 4096) for r in range(n)]
        assert blocks[0][0] == 0 and sum(c for _, c in blocks) == 4096
        for (f0, c0), (f1, _) in zip(blocks, blocks[1:]):
            assert f1 == f0 + c0
        assert max(c for _, c in blocks) - min(c for _, c in blocks) <= 1


def test_single_rank_is_identity():
    assert sharding.max_over_ranks(3.5) == 3.5
    assert sharding.world_block(5, 4096) == (5 * 4096, 4096)
