"""Row f1: real files.  A tiny mp4 dataset tree is written with cv2.VideoWriter, read back through
the drop-in VideoDataset (host decode with the reference's own cv2 calls, frames resident in HBM, fused clip
kernel) and compared with the reference's OpenCV fallback recipe (vivit .../dataset.py:317-350: seek, read,
BGR->RGB, cv2.resize) run on the CPU."""
import random

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _write_tree(root, n_frames=24, h=120, w=160):
    import cv2

    rng = np.random.default_rng(3)
    paths = []
    for cls, count in (("referral", 2), ("non_referral", 1)):
        d = root / "val" / cls
        d.mkdir(parents=True)
        for k in range(count):
            p = d / f"v{k}.mp4"
            wr = cv2.VideoWriter(str(p), cv2.VideoWriter_fourcc(*"mp4v"), 25.0, (w, h))
            assert wr.isOpened()
            base = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
            for i in range(n_frames):
                wr.write(np.roll(base, 3 * i, axis=1))
            wr.release()
            paths.append(str(p))
    return paths


def test_dataset_on_real_mp4_files(tmp_path):
    import cv2

    from lvc_b200.datasets import VideoDataset
    from oracle import samplers as osm

    _write_tree(tmp_path)
    ds = VideoDataset(str(tmp_path), mode="val", sampling_method="random_window", num_frames=8)
    assert len(ds) == 3 and sorted(ds.labels) == [0, 1, 1]
    rng = random.Random(42)
    for i in range(len(ds)):
        s = ds[i]
        cap = cv2.VideoCapture(ds.video_paths[i])
        total = int(cap.get(cv2.CAP_PROP_FRAME_COUNT))
        want_idx = osm.dataset_indices("random_window", total, 8, rng)
        assert s["frame_indices"] == want_idx
        frames = []
        for fi in want_idx:  # the reference's fallback: seek + read + BGR2RGB + resize
            cap.set(cv2.CAP_PROP_POS_FRAMES, fi)
            ok, fr = cap.read()
            assert ok
            frames.append(cv2.resize(cv2.cvtColor(fr, cv2.COLOR_BGR2RGB), (224, 224)))
        cap.release()
        assert np.array_equal(s["pixel_values"].cpu().numpy(), np.stack(frames))
        assert ds.source.fps(i) == 25.0
