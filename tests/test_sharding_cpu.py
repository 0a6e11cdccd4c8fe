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
    # shards for the next pass sized by what every rank measured (seconds per clip), identical on every rank
    counts = sharding.weighted_counts(128, sharding.ingest_weights([sec / n for n, sec, _ in stats]))
    q.put((rank, mine, table[mine].tolist(), stats, counts))
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
    for rank, mine, rows, stats, counts in res:
        assert rows == [want[v] for v in mine]
        assert len(stats) == world and [s[0] for s in stats] == [4.0, 3.0]
        assert stats == res[0][3]  # every rank holds the same gathered report
        # rank 0: 0.5 s / 4 clips, rank 1: 1.5 s / 3 clips -> rates 8 : 2 clips/s, clamped to 2 x the median (5): 8 : 2.5
        assert counts == res[0][4] == [97, 31] and sum(counts) == 128
    assert sum(s[2] for s in res[0][3]) == float(sum(sum(w) for w in want))


def test_single_process_gather_and_reseed_plan():
    assert sharding.gather_stats(3, 1.5, 7) == [(3.0, 1.5, 7.0)]
    t = sharding.plan_epoch([100, 100], "random", 8, reseed_each_call=True)
    assert t[0].tolist() == t[1].tolist() == osm.dataset_indices("random", 100, 8, random.Random(42))


def test_weighted_counts_and_shards():
    # the measured eight-GPU box: four links at 23 GB/s, four at 35 GB/s, 512 clips per step
    w = sharding.ingest_weights([1 / 23.2] * 4 + [1 / 35.4] * 4)
    c = sharding.weighted_counts(512, w)
    assert sum(c) == 512
    assert max(c[:4]) - min(c[:4]) <= 1 and max(c[4:]) - min(c[4:]) <= 1 and 76 <= c[4] <= 78 and 50 <= c[0] <= 52
    # every step then ends together: time = clips / rate is equal to within one clip
    t = [c[i] / (23.2 if i < 4 else 35.4) for i in range(8)]
    assert max(t) / min(t) < 1.03
    assert sharding.weighted_counts(10, [1, 1, 1]) == [4, 3, 3]                 # ties: lowest rank first
    assert sharding.weighted_counts(3, [100, 1, 1]) == [1, 1, 1]                # floor: nobody starves
    assert sharding.weighted_counts(64, [1.0]) == [64]
    sh = sharding.weighted_shards(7, [2, 1])
    assert sh == [[0, 1, 2, 3, 4], [5, 6]] and sorted(sum(sh, [])) == list(range(7))
    assert sharding.ingest_weights([1.0, 1.0, 100.0]) == [1.0, 1.0, 0.5]        # a stalled measurement is clamped
    import pytest
    with pytest.raises(ValueError):
        sharding.weighted_counts(2, [1, 1, 1])
    with pytest.raises(ValueError):
        sharding.weighted_counts(8, [1, 0])
