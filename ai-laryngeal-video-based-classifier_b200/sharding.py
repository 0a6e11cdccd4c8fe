"""Multi-GPU partitioning of the clip path: one process per GPU, videos sharded by index, no
collective on the data path (a clip depends on one video only -- SURVEY.md 8e).  The only
communication is one ``all_gather`` of per-rank stats (clips, seconds, checksum) for the report;
backend nccl on GPUs, gloo in the CPU tests.
"""
import numpy as np

from . import sampling


def shard_indices(n_videos, rank, world_size):
    """Video v belongs to rank v mod world_size."""
    return list(range(rank, n_videos, world_size))


def plan_epoch(total_frames, method, num_frames, seed=42, reseed_each_call=False):
    """Index table (n_videos, T) for a whole pass IN DATASET ORDER on one stream seeded like the
    reference's ``VideoDataset.__init__`` (vivit .../dataset.py:39-40): exactly what a sequential,
    unshuffled single-process pass of the reference draws, so the result does not depend on how the
    videos are later sharded over ranks.  Short videos may yield fewer distinct frames but always T
    entries (the dataset variant never clamps T)."""
    rng = sampling.SamplerRng(seed, "py")
    table = np.empty((len(total_frames), num_frames), np.int32)
    for v, n in enumerate(total_frames):
        if reseed_each_call:
            rng.seed(seed)
        table[v] = sampling.get_sampling_indices(method, int(n), num_frames, rng)
    return table


def gather_stats(clips, seconds, checksum, device=None):
    """all_gather of (clips, seconds, checksum mod 2^40) over the default process group; returns a list
    of per-rank tuples on every rank (a single entry when torch.distributed is not initialised)."""
    import torch
    import torch.distributed as dist

    mine = torch.tensor([float(clips), float(seconds), float(int(checksum) % (1 << 40))], dtype=torch.float64,
                        device=device)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [tuple(mine.tolist())]
    out = [torch.empty_like(mine) for _ in range(dist.get_world_size())]
    dist.all_gather(out, mine)
    return [tuple(t.tolist()) for t in out]
