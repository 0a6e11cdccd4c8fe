"""Randomised parity sweep of the raw-frame HF processor kernel: python tests/fuzz_hfraw.py [cases] [seed]
Random frame sizes (tiny, odd, portrait, up to 4K), mixed sizes per batch, both presets, edges, layouts, BGR."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lvc_b200 import ClipBuilder  # noqa: E402
from oracle import processors as op  # noqa: E402

cases = int(sys.argv[1]) if len(sys.argv) > 1 else 30
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
bad = 0
for k in range(cases):
    B = int(rng.integers(1, 4))
    T = int(rng.choice([1, 2, 3, 8]))
    edge = int(rng.choice([224, 256, 256, 300]))
    vids = []
    for _ in range(B):
        if rng.random() < 0.4:
            h, w = [(240, 320), (480, 640), (720, 1280), (1080, 1920), (1920, 1080), (224, 224), (2160, 3840)][int(rng.integers(0, 7))]
        else:
            h, w = int(rng.integers(1, 1500)), int(rng.integers(1, 2200))
        if min(h, w) > 11 * edge:
            h, w = h // 2, w // 2
        n = int(rng.integers(1, 4))
        vids.append(rng.integers(0, 256, (n, h, w, 3), dtype=np.uint8))
    idx = np.stack([rng.integers(-1, v.shape[0] + 1, T) for v in vids])   # out-of-range indices clamp
    bgr = bool(rng.random() < 0.3)
    mode = int(rng.integers(0, 3))
    try:
        dv = [torch.from_numpy(v[..., ::-1].copy() if bgr else v).cuda() for v in vids]
        kw = dict(num_frames=T, shortest_edge=edge, src_order="BGR" if bgr else "RGB")
        u8 = np.stack([op.hf_shortest_edge_crop_u8(v[np.clip(idx[b], 0, v.shape[0] - 1)], edge) for b, v in enumerate(vids)])
        if mode == 0:
            got = ClipBuilder("vivit_raw", layout="BTHWC", dtype=torch.uint8, **kw)(dv, idx).cpu().numpy()
            ok = np.array_equal(got, u8)
        elif mode == 1:
            got = ClipBuilder("timesformer_raw", **kw)(dv, idx).cpu().numpy()
            ok = np.array_equal(got, np.stack([op.videomae_processor(u) for u in u8]))
        else:
            got = ClipBuilder("timesformer_raw", layout="BCTHW", **kw)(dv, idx).cpu().numpy()
            ok = np.array_equal(got, np.stack([op.videomae_processor(u) for u in u8]).transpose(0, 2, 1, 3, 4))
    except Exception as e:  # noqa: BLE001
        ok = False
        print("ERROR", repr(e))
    if not ok:
        bad += 1
        print("MISMATCH", k, [v.shape for v in vids], edge, T, bgr, mode, flush=True)
print(f"fuzz_hfraw: {cases - bad}/{cases} cases match", flush=True)
sys.exit(1 if bad else 0)
