"""Timed CPU baseline: the reference's CPU recipe for the clip path, run on host cores.

Test/bench infrastructure (only bench.py's cpu_baseline / --impl reference legs call this).
The per-clip function calls the SAME third-party routines the reference calls (cv2.resize,
numpy/torch float ops), in the reference's order -- resnet50-2d-lstm/inference.py:158-178 for the
`resnet_infer` recipe (BASELINE configs 1 and 5) -- so the timing is the reference's arithmetic,
not the slower integer restatement in oracle/resize.py (that one is the bit-exact checker).
"""
import os
import time

import numpy as np


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def uniform_indices(N, T):
    if T == 1:
        return [N // 2]
    step = (N - 1) / (T - 1)
    return [min(int(i * step), N - 1) for i in range(T)]


def clip_resnet_infer(video, T):
    """One clip, reference order: sample -> per-frame cv2.resize -> stack -> /255 -> -mean -> /std
    -> permute(3,0,1,2) -> unsqueeze(0)."""
    import cv2
    import torch

    frames = [cv2.resize(video[i], (224, 224)) for i in uniform_indices(video.shape[0], T)]
    frames = np.stack(frames, axis=0)
    frames = frames.astype(np.float32) / 255.0
    frames = torch.FloatTensor(frames)
    frames = frames - torch.tensor([0.45, 0.45, 0.45]).view(1, 1, 1, 3)
    frames = frames / torch.tensor([0.225, 0.225, 0.225]).view(1, 1, 1, 3)
    frames = frames.permute(3, 0, 1, 2).unsqueeze(0).contiguous()
    return frames


_VIDEO = None


def _init():
    import cv2
    import torch

    cv2.setNumThreads(1)
    torch.set_num_threads(1)


def _work(args):
    T, clips = args
    t0 = time.perf_counter()
    s = 0.0
    for _ in range(clips):
        s += float(clip_resnet_infer(_VIDEO, T)[0, 0, 0, 0, 0])
    return time.perf_counter() - t0, s


class Pool:
    """`workers` single-threaded processes (the reference parallelises with DataLoader workers,
    vivit_transformer/main.py:49) sharing one resident synthetic video (copy-on-write after fork)."""

    def __init__(self, height, width, n_frames, T, workers):
        global _VIDEO
        import multiprocessing as mp

        _VIDEO = np.random.default_rng(1000).integers(0, 256, (n_frames, height, width, 3), dtype=np.uint8)
        self.T, self.workers = T, workers
        self.pool = mp.get_context("fork").Pool(workers, initializer=_init)

    def step(self, clips_per_worker=1):
        """All workers build `clips_per_worker` clips; returns (clips built, seconds)."""
        t0 = time.perf_counter()
        self.pool.map(_work, [(self.T, clips_per_worker)] * self.workers, chunksize=1)
        return self.workers * clips_per_worker, time.perf_counter() - t0

    def close(self):
        self.pool.terminate()
        self.pool.join()


def run(height, width, n_frames, T, workers, clips_per_worker, repeats=3):
    """clips/s of the pool: median over `repeats` steps after one warm-up step."""
    pool = Pool(height, width, n_frames, T, workers)
    try:
        pool.step(1)
        vals = []
        for _ in range(repeats):
            n, dt = pool.step(clips_per_worker)
            vals.append(n / dt)
    finally:
        pool.close()
    return float(np.median(vals))
