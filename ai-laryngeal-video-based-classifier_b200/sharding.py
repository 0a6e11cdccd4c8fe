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


def weighted_counts(n_items, weights, floor=1):
    """Split ``n_items`` over ranks in proportion to ``weights`` (largest-remainder rounding, every rank at least
    ``floor``).  Used to shard by MEASURED ingest rate: on a box whose GPUs do not all reach the same host-to-device
    rate (shared PCIe uplinks: 23 vs 35 GB/s per GPU were measured with eight GPUs active, profiles/r02_pcie_probe_n8.json)
    equal shards make every step wait for the slowest link; proportional shards finish together."""
    w = np.asarray(weights, dtype=np.float64)
    if w.ndim != 1 or len(w) == 0 or not np.all(np.isfinite(w)) or np.any(w <= 0):
        raise ValueError("weights must be positive finite numbers, one per rank")
    n_items, floor = int(n_items), int(floor)
    if n_items < floor * len(w):
        raise ValueError("fewer items than ranks x floor")
    ideal = w / w.sum() * (n_items - floor * len(w))
    counts = np.floor(ideal).astype(np.int64)
    rest = int(n_items - floor * len(w) - counts.sum())
    order = np.argsort(-(ideal - counts), kind="stable")          # ties: lowest rank first
    counts[order[:rest]] += 1
    return (counts + floor).tolist()


def weighted_shards(n_videos, weights):
    """Contiguous blocks of videos per rank, sized by ``weighted_counts``: ``[range(a0, a1), range(a1, a2), ...]``."""
    counts = weighted_counts(n_videos, weights, floor=0)
    out, a = [], 0
    for c in counts:
        out.append(list(range(a, a + c)))
        a += c
    return out


def ingest_weights(seconds_per_clip, clamp=2.0):
    """Per-rank weights from measured seconds per clip (all ranks ingesting concurrently): rate = 1 / seconds, clamped
    to within ``clamp`` x of the median so that one noisy measurement cannot starve a rank."""
    r = 1.0 / np.asarray(seconds_per_clip, dtype=np.float64)
    med = float(np.median(r))
    return np.clip(r, med / clamp, med * clamp).tolist()


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
