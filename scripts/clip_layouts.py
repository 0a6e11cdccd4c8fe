"""Clip stage of one secondary config under different output layouts / dtypes (where do the stores cost?):
python scripts/clip_layouts.py [cfg]   -- CUDA events, L2 flushed, median of 15."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from lvc_b200 import ClipBuilder, sampling  # noqa: E402
from lvc_b200.synth import synth_boxes  # noqa: E402

dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
name = sys.argv[1] if len(sys.argv) > 1 else "swin_cfg4"
c = bench.SECONDARY[name]
g = torch.Generator(device=dev).manual_seed(7)
vids = [torch.randint(0, 256, (c["frames"], c["h"], c["w"], 3), dtype=torch.uint8, device=dev, generator=g) for _ in range(c["videos"])]
rng = sampling.SamplerRng(42)
idx = np.array([sampling.get_sampling_indices(c["sampling"], c["frames"], c["T"], rng) for _ in vids], np.int32)
boxes = None
if c.get("boxes"):
    boxes = np.stack([synth_boxes(2000 + v, c["frames"], c["h"], c["w"])[idx[v]] for v in range(len(vids))])
VARIANTS = [("TOKENS", c["patch"], None, True), ("TOKENS", (2, 16, 16), None, True), ("BCTHW", None, torch.bfloat16, True),
            ("BCTHW", None, torch.float32, True), ("BTHWC", None, torch.uint8, True), ("TOKENS", c["patch"], None, False)]
for layout, patch, dt, use_boxes in VARIANTS:
    preset = "dataset_u8" if dt == torch.uint8 else c["preset"]
    cb = ClipBuilder(preset, num_frames=c["T"], layout=layout, patch=patch, dtype=dt)
    batch = cb.prepare(vids, idx, boxes if use_boxes else None)
    out = torch.empty(batch.shape, dtype=cb.dtype, device=dev)
    ts = []
    for _ in range(18):
        flush.zero_(); flush.view(torch.int64).sum()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        cb.run(batch, out)
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    print(f"{name} {layout:7s} patch={patch} dtype={dt} boxes={use_boxes and boxes is not None}: {np.median(ts[3:]):7.1f} us", flush=True)
