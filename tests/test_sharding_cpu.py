"""N>1 host logic on CPU: world_size-2 gloo run of the video sharding, the order-independent index
plan and the stats gather (the only collective of the job)."""
import os
import random

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from lvc_b200 import sharding
from oracle import samplers as osm


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n_frames = [48, 300, 5, 33, 1000, 64, 31]
    mine = sharding.shard_indices(len(n_frames), rank, world)
    table = sharding.plan_epoch(n_frames, "random_window", 32)
    checksum = int(table[mine].astype(np.int64).sum())
    stats = sharding.gather_stats(len(mine), 0.5 + rank, checksum)
    q.put((rank, mine, table[mine].tolist(), stats))
    dist.destroy_process_group()


def test_world_size_2_gloo():
    world, port = 2, 29500 + random.randint(0, 2000)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    n_frames = [48, 300, 5, 33, 1000, 64, 31]
    # shards partition the videos
    assert sorted(res[0][1] + res[1][1]) == list(range(len(n_frames)))
    # the sharded index rows are those of ONE sequential reference stream (seed 42), whatever the world size
    rng = random.Random(42)
    want = [osm.dataset_indices("random_window", n, 32, rng) for n in n_frames]
    for rank, mine, rows, stats in res:
        assert rows == [want[v] for v in mine]
        assert len(stats) == world and [s[0] for s in stats] == [4.0, 3.0]
        assert stats == res[0][3]  # every rank holds the same gathered report
    assert sum(s[2] for s in res[0][3]) == float(sum(sum(w) for w in want))


def test_single_process_gather_and_reseed_plan():
    assert sharding.gather_stats(3, 1.5, 7) == [(3.0, 1.5, 7.0)]
    t = sharding.plan_epoch([100, 100], "random", 8, reseed_each_call=True)
    assert t[0].tolist() == t[1].tolist() == osm.dataset_indices("random", 100, 8, random.Random(42))
