"""Times the letterbox kernel (row f4) with CUDA events: python scripts/perf_letterbox.py"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lvc_b200 import yolo  # noqa: E402

flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for (n, h, w) in ((64, 1080, 1920), (64, 720, 1280), (64, 480, 640)):
    fr = torch.randint(0, 256, (n, h, w, 3), dtype=torch.uint8, device="cuda")
    out, g = yolo.letterbox_batch(fr)
    ts = []
    for _ in range(13):
        flush.zero_(); flush.view(torch.int64).sum()
        torch.cuda._sleep(400000)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        yolo.letterbox_batch(fr, out=out)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts[3:])[5]
    rows = min(2 * g.new_h, h)
    phys = n * (rows * w * 3 + out[0].numel() * 4)
    print(json.dumps({"frames": n, "src": [h, w], "out": [g.out_h, g.out_w], "ms": round(ms, 4), "frames_per_s": round(n / ms * 1e3),
                      "sector_floor_GBs": round(phys / ms / 1e6, 1)}))
