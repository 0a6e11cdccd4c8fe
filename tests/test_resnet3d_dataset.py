"""The resnet50-3d-video Dataset copy's sampler contract (resnet50-3d-video/video_classifier/data_config/dataset.py:79-169):
per-video seed, index cache, re-seed to 42 -- against indices produced by the reference class itself
(tests/golden/samplers3d.npz, `hash` pinned to crc32 in the reference module's namespace)."""
import csv
import os

import numpy as np
import pytest

from lvc_b200 import sampling

METHODS = ["uniform", "random_window", "random"]


def _bare(method, T):
    from lvc_b200.datasets import Resnet3dVideoDataset, crc32_video_seed

    ds = object.__new__(Resnet3dVideoDataset)   # no ClipBuilder (needs a GPU): the sampler is host code
    ds.sampling_method, ds.num_frames, ds.seed_fn, ds.cached_indices = method, T, crc32_video_seed, {}
    ds.rng = sampling.SamplerRng(42, "py")
    return ds


def test_golden_indices_seed_cache_reseed(golden):
    g = golden("samplers3d.npz")
    names = [str(n) for n in g["names"]]
    rows, want = g["rows"], g["indices"]
    ds, key = None, None
    for (mi, N, T, call), w in zip(rows.tolist(), want):
        if (mi, N, T) != key:
            key, ds = (mi, N, T), _bare(METHODS[mi], T)
        if call == 3:   # stray draws on the dataset's stream must not matter: every video seeds its own
            for _ in range(3):
                ds.rng.random()
        got = ds.get_sampling_indices("/data/train/referral/" + names[call], N)
        assert got == [int(v) for v in w[w >= 0]], (METHODS[mi], N, T, call)
        # after every call the stream is random.seed(42) again (dataset.py:163-164)
        assert ds.rng.random() == sampling.SamplerRng(42, "py").random()
        ds.rng.seed(42)


def test_cache_returns_the_same_list_and_custom_seed_fn():
    ds = _bare("random", 8)
    a = ds.get_sampling_indices("x/v1.mp4", 100)
    assert ds.get_sampling_indices("x/v1.mp4", 100) is a          # cached object, as in the reference
    assert ds.get_sampling_indices("x/v1.mp4", 5) is a            # the cache ignores total_frames (reference: :91-93)
    ds2 = _bare("random", 8)
    ds2.seed_fn = lambda name: 7
    want = sampling.get_sampling_indices("random", 100, 8, sampling.SamplerRng(7, "py"))
    assert ds2.get_sampling_indices("y/other.mp4", 100) == want


@pytest.mark.gpu
def test_dataset_items_and_csv(tmp_path):
    import torch

    from lvc_b200.datasets import Resnet3dVideoDataset, SyntheticFrameSource, window_subsample
    from oracle import processors as op

    src = SyntheticFrameSource(3, 20, 120, 160)
    ds = Resnet3dVideoDataset(frame_source=src, mode="val", sampling_method="random_window", num_frames=8,
                              log_dir=str(tmp_path))
    clips, labels = ds[1]
    assert tuple(clips.shape) == (1, 3, 8, 224, 224) and labels.dtype == torch.long
    idx = ds.cached_indices[ds.video_paths[1]]
    host = src.frames(1).cpu().numpy()
    want = op.ptv_eval_chain(host[window_subsample(idx, 8)])
    assert np.allclose(clips[0].cpu().numpy(), want, atol=1.1e-2, rtol=1e-5)
    path = ds.save_sampled_indices()
    with open(path) as f:
        rows = list(csv.reader(f))
    assert rows[0] == ["video_filename", "total_frames", "sampled_frames"] and len(rows) == 4
    assert rows[2] == [os.path.basename(ds.video_paths[1]), "20", ",".join(map(str, idx))]
