"""Oracle: frame-index samplers.  Test infrastructure.

Every function is a restatement of a reference function and cites it.  The RNG is
passed in explicitly (``rng``: anything with CPython ``random`` semantics --
the ``random`` module itself or a ``random.Random``; ``nprng``: a numpy legacy
``RandomState`` or the ``numpy.random`` module) because the reference consumes
process-global streams.
"""
import math
import random as _random

import numpy as np

METHODS = ("uniform", "random_window", "random")


def dataset_indices(method, total_frames, num_frames, rng=_random):
    """a1 -- VideoDataset.get_sampling_indices.

    Follows vivit_transformer/vivit_classifier/data_config/dataset.py:129-193
    (identical arithmetic in timesformer/.../dataset.py:137-207,
    videoswintransformer/.../dataset.py:64-149, resnet50-3d-video/.../dataset.py:79-169,
    resnet50-2d-lstm/src/data_config/dataset.py:85-163 -- the last one re-seeds
    ``random.seed(42)`` on every call, :97; callers model that by passing a fresh
    ``random.Random(42)``).
    """
    N, T = int(total_frames), int(num_frames)
    if N >= T:
        if method == "random":
            return sorted(rng.sample(range(N), T))
        if method == "random_window":
            w = N / T
            out = []
            for i in range(T):
                start = int(i * w)
                end = min(int((i + 1) * w), N)
                end = max(end, start + 1)
                out.append(rng.randint(start, end - 1))
            return out
        if T == 1:
            return [N // 2]
        step = (N - 1) / (T - 1)
        return [min(int(i * step), N - 1) for i in range(T)]
    if method == "random":
        return sorted(rng.choices(range(N), k=T))
    if method == "random_window":
        w = N / T
        out = []
        for i in range(T):
            vs = i * w
            ve = (i + 1) * w
            out.append(min(int(math.floor(vs + (ve - vs) * rng.random())), N - 1))
        return out
    if T == 1:
        return [N // 2]
    step = N / T
    return [min(int(i * step), N - 1) for i in range(T)]


def dynamic_fps(total_frames, num_frames, original_fps):
    """Second return value of the Swin copy, videoswintransformer/.../dataset.py:104-121."""
    if total_frames >= num_frames:
        return None
    if original_fps <= 0:
        original_fps = 30.0
    return num_frames / (total_frames / original_fps)


def sample_frame_indices(clip_len, frame_sample_rate, seg_len, nprng=np.random):
    """a2 -- data_handling.py:41-56 (numpy legacy global RNG)."""
    converted_len = int(clip_len * frame_sample_rate)
    end_idx = nprng.randint(converted_len, seg_len)
    start_idx = end_idx - converted_len
    indices = np.linspace(start_idx, end_idx, num=clip_len)
    return np.clip(indices, start_idx, end_idx - 1).astype(np.int64)


def inference_indices_np(method, total_frames, num_frames, seed=42):
    """a3 -- numpy-RNG inference samplers, resnet50-2d-lstm/inference.py:74-130.

    ``np.random.seed(42)`` precedes every random draw there (:77, :114); T is clamped
    to N first (:80, :97, :117).
    """
    N = int(total_frames)
    T = min(int(num_frames), N)
    if method == "random":
        st = np.random.RandomState(seed)
        return sorted(int(v) for v in st.choice(N, T, replace=False))
    if method == "random_window":
        st = np.random.RandomState(seed)
        w = N / T
        out = []
        for i in range(T):
            start = int(i * w)
            end = min(int((i + 1) * w), N)
            end = max(end, start + 1)
            out.append(int(st.randint(start, end)))
        return out
    if T == 1:
        return [N // 2]
    step = (N - 1) / (T - 1)
    return [min(int(i * step), N - 1) for i in range(T)]


def uniform_temporal_subsample(t, num_samples):
    """a4 -- pytorchvideo ``uniform_temporal_subsample`` indices (package absent, restated
    from its published definition: ``torch.linspace(0, t-1, n).clamp(0, t-1).long()``).
    Call sites: videoswintransformer/.../dataset.py:157,168; resnet50-2d-lstm/.../dataset.py:67,78.
    """
    import torch

    idx = torch.linspace(0, t - 1, num_samples)
    return torch.clamp(idx, 0, t - 1).long().tolist()


def gather_positions(indices, decoded_frames):
    """a5 -- clip-relative frame positions, vivit .../dataset.py:246-255:
    ``pos = clamp(idx - min(idx), 0, decoded-1)``."""
    start = min(indices)
    return [max(0, min(i - start, decoded_frames - 1)) for i in indices]


def pure_indices(method, total_frames, num_frames, rng=_random):
    """visualize_sampling.py:105-178 -- the stand-alone copies: T is clamped to N first
    (:116, :134, :158), then the N >= T formulas of a1 on the CPython stream."""
    N = int(total_frames)
    return dataset_indices(method, N, min(int(num_frames), N), rng)
