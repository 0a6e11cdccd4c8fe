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


def clip_vivit(video, T):
    """ViViT recipe, reference order (vivit .../dataset.py:271-277 + trainers/trainer.py:62-95): per-frame cv2.resize,
    then the HF VivitImageProcessor on the list of frames."""
    import cv2
    from transformers import VivitImageProcessor

    global _PROC
    if _PROC is None:
        _PROC = VivitImageProcessor()
    frames = [cv2.resize(video[i], (224, 224)) for i in uniform_indices(video.shape[0], T)]
    return _PROC(frames, return_tensors="pt")["pixel_values"]


def clip_timesformer(video, T):
    """TimeSformer recipe (timesformer .../trainers/trainer.py:91-98): cv2.resize, x/255, (x-mean)/std, CHW."""
    import cv2
    import torch

    frames = np.stack([cv2.resize(video[i], (224, 224)) for i in uniform_indices(video.shape[0], T)])
    x = torch.from_numpy(frames.astype(np.float64) * (1 / 255)).float()
    x = (x - 0.45) / 0.225
    return x.permute(0, 3, 1, 2).unsqueeze(0).contiguous()


def clip_swin_roi(video, T):
    """cfg 4: per-frame ROI crop (numpy slice of a W/3 x H/3 box) + cv2.resize + manual normalise, (1,C,T,H,W)."""
    import cv2
    import torch

    H, W = video.shape[1:3]
    frames = np.stack([cv2.resize(video[i][H // 3:2 * H // 3, W // 3:2 * W // 3], (224, 224))
                       for i in uniform_indices(video.shape[0], T)])
    x = torch.FloatTensor(frames.astype(np.float32) / 255.0)
    x = (x - torch.tensor([0.45] * 3).view(1, 1, 1, 3)) / torch.tensor([0.225] * 3).view(1, 1, 1, 3)
    return x.permute(3, 0, 1, 2).unsqueeze(0).contiguous()


RECIPES = {"resnet_infer": clip_resnet_infer, "vivit": clip_vivit, "timesformer": clip_timesformer, "swin_roi": clip_swin_roi}
_PROC = None
_RECIPE = clip_resnet_infer


def embed_cpu_ms(kind, batch, T, threads, repeats=2):
    """fp32 patch embedding of the reference modules on `threads` CPU threads: ms per batch (median)."""
    import torch
    import torch.nn.functional as F

    torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(0)
    if kind == "vivit":
        x, w = torch.randn(batch, 3, T, 224, 224, generator=g), torch.randn(768, 3, 2, 16, 16, generator=g) * 0.02
        fn = lambda: F.conv3d(x, w, None, stride=(2, 16, 16)).flatten(2).transpose(1, 2)  # noqa: E731
    elif kind == "timesformer":
        x, w = torch.randn(batch * T, 3, 224, 224, generator=g), torch.randn(768, 3, 16, 16, generator=g) * 0.02
        fn = lambda: F.conv2d(x, w, None, stride=16).flatten(2).transpose(1, 2)  # noqa: E731
    else:
        x, w = torch.randn(batch, 3, T, 224, 224, generator=g), torch.randn(128, 3, 2, 4, 4, generator=g) * 0.02
        ln = (torch.ones(128), torch.zeros(128))
        fn = lambda: F.layer_norm(F.conv3d(x, w, None, stride=(2, 4, 4)).permute(0, 2, 3, 4, 1), (128,), *ln, 1e-5)  # noqa: E731
    with torch.no_grad():
        fn()
        ts = []
        for _ in range(repeats):
            t0 = time.perf_counter()
            fn()
            ts.append(time.perf_counter() - t0)
    return 1e3 * float(np.median(ts))


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
        s += float(_RECIPE(_VIDEO, T).reshape(-1)[0])
    return time.perf_counter() - t0, s


class Pool:
    """`workers` single-threaded processes (the reference parallelises with DataLoader workers,
    vivit_transformer/main.py:49) sharing one resident synthetic video (copy-on-write after fork)."""

    def __init__(self, height, width, n_frames, T, workers, recipe="resnet_infer"):
        global _VIDEO, _RECIPE
        import multiprocessing as mp

        _RECIPE = RECIPES[recipe]   # inherited by the forked workers

        _VIDEO = np.random.default_rng(1000).integers(0, 256, (n_frames, height, width, 3), dtype=np.uint8)
        self.T, self.workers = T, workers
        self.pool = mp.get_context("fork").Pool(workers, initializer=_init)

    def step(self, clips_per_worker=1):
        """The pool builds `workers * clips_per_worker` clips, handed out one clip at a time from a shared queue (a
        worker that finishes early takes the next clip: no straggler wait beyond one clip); returns (clips, seconds)."""
        total = self.workers * clips_per_worker
        t0 = time.perf_counter()
        self.pool.map(_work, [(self.T, 1)] * total, chunksize=1)
        return total, time.perf_counter() - t0

    def close(self):
        self.pool.terminate()
        self.pool.join()


def run(height, width, n_frames, T, workers, clips_per_worker, repeats=3, recipe="resnet_infer"):
    """clips/s of the pool: median over `repeats` steps after one warm-up step."""
    pool = Pool(height, width, n_frames, T, workers, recipe)
    try:
        pool.step(1)
        vals = []
        for _ in range(repeats):
            n, dt = pool.step(clips_per_worker)
            vals.append(n / dt)
    finally:
        pool.close()
    return float(np.median(vals))
