"""Shared set-up of the profiling scripts: one secondary config of bench.py as a prepared clip batch."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from lvc_b200 import ClipBuilder, sampling  # noqa: E402
from lvc_b200.synth import synth_boxes  # noqa: E402


def setup_clip(name, dev=None):
    dev = dev or torch.device("cuda", 0)
    c = bench.SECONDARY[name]
    g = torch.Generator(device=dev).manual_seed(7)
    vids = [torch.randint(0, 256, (c["frames"], c["h"], c["w"], 3), dtype=torch.uint8, device=dev, generator=g)
            for _ in range(c["videos"])]
    rng = sampling.SamplerRng(42)
    idx = np.array([sampling.get_sampling_indices(c["sampling"], c["frames"], c["T"], rng) for _ in vids], np.int32)
    boxes = None
    if c.get("boxes"):
        boxes = np.stack([synth_boxes(2000 + v, c["frames"], c["h"], c["w"])[idx[v]] for v in range(len(vids))])
    cb = ClipBuilder(c["preset"], num_frames=c["T"], layout="TOKENS", patch=c["patch"])
    batch = cb.prepare(vids, idx, boxes)
    tok = torch.empty(batch.shape, dtype=torch.bfloat16, device=dev)
    return cb, batch, tok, vids
