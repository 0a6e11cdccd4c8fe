"""Host-side dataset logic that needs no GPU."""
import os

import pytest

from lvc_b200 import sampling
from oracle import samplers as osm


def test_window_subsample_matches_pytorchvideo_rule():
    from lvc_b200.datasets import window_subsample

    for raw in ([3, 9, 40, 41], [0, 0, 0, 0], [5, 6, 7, 8, 9, 10, 11, 12], list(range(0, 300, 10))):
        T = len(raw)
        lo, hi = min(raw), max(raw)
        assert window_subsample(raw, T) == [lo + k for k in osm.uniform_temporal_subsample(hi - lo + 1, T)]


def test_scan_split_layout(tmp_path):
    from lvc_b200.datasets import scan_split

    for cls, names in {"referral": ["b.mp4", "a.mp4"], "non_referral": ["c.mp4", "note.txt"]}.items():
        d = tmp_path / "val" / cls
        d.mkdir(parents=True)
        for n in names:
            (d / n).write_bytes(b"")
    paths, labels = scan_split(tmp_path, "val")
    assert [os.path.basename(p) for p in paths] == ["c.mp4", "a.mp4", "b.mp4"] and labels == [0, 1, 1]
    with pytest.raises(ValueError):
        scan_split(tmp_path, "train")


def test_collate_functions():
    import torch

    from lvc_b200.dataloader import clip_collate_fn, collate_fn, video_collate_fn

    s = [{"pixel_values": torch.zeros(4, 224, 224, 3, dtype=torch.uint8), "labels": torch.tensor(i)} for i in range(3)]
    b = video_collate_fn(s)
    assert len(b["pixel_values"]) == 3 and b["labels"].tolist() == [0, 1, 2]
    c, l = clip_collate_fn([(torch.zeros(2, 3, 4, 8, 8), torch.zeros(2, dtype=torch.long)),
                            (torch.zeros(1, 3, 4, 8, 8), torch.zeros(1, dtype=torch.long))])
    assert tuple(c.shape) == (2, 1, 3, 4, 8, 8) and tuple(l.shape) == (2, 1)
    assert tuple(collate_fn(s)["pixel_values"].shape) == (3, 4, 224, 224, 3)
    assert sampling.uniform_sampling(10, 4) == [0, 3, 6, 9]
