"""Randomised parity sweep of the detector pre / post-processing kernels: python tests/fuzz_yolo.py [cases] [seed]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lvc_b200 import yolo  # noqa: E402
from oracle import yolo as oy  # noqa: E402

cases = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
bad = 0
for k in range(cases):
    h = int(rng.choice([1, 2, 37, 240, 360, 480, 640, 720, 1080])) if rng.random() < 0.5 else int(rng.integers(1, 1300))
    w = int(rng.choice([1, 3, 150, 320, 640, 1280, 1920])) if rng.random() < 0.5 else int(rng.integers(1, 2000))
    shape = int(rng.choice([320, 640, 416]))
    stride = int(rng.choice([32, 64]))
    n = int(rng.integers(1, 4))
    frames = rng.integers(0, 256, (n, h, w, 3), dtype=np.uint8)
    gh = yolo_geom_or_none = None
    try:
        gh = yolo.letterbox_geometry(h, w, (shape, shape), stride)
    except ValueError:
        # the frame collapses to zero rows / columns: the reference's cv2.resize raises as well
        try:
            oy.preprocess(frames[0], (shape, shape), stride)
            ok = False
        except Exception:  # noqa: BLE001
            ok = True
        if not ok:
            bad += 1
            print("MISMATCH (degenerate accepted by the reference)", k, (h, w), flush=True)
        continue
    try:
        im, g = yolo.letterbox_batch(torch.from_numpy(frames).cuda(), new_shape=(shape, shape), stride=stride)
        want = torch.cat([oy.preprocess(f, (shape, shape), stride) for f in frames])
        ok = tuple(im.shape) == tuple(want.shape) and torch.equal(im.cpu(), want)
        nd = int(rng.integers(1, 30))
        det = np.concatenate([rng.uniform(-30, shape + 30, (nd, 4)), rng.uniform(0, 1, (nd, 2))], 1).astype(np.float32)
        got = yolo.scale_boxes(torch.from_numpy(det).cuda(), g).cpu().numpy()
        wb = np.array(oy.detections_to_int_boxes(torch.from_numpy(det), (g.out_h, g.out_w), (h, w, 3)))
        ok = ok and np.array_equal(got, wb)
    except Exception as e:  # noqa: BLE001
        ok = False
        print("ERROR", repr(e))
    if not ok:
        bad += 1
        print("MISMATCH", k, (h, w), shape, stride, flush=True)
print(f"fuzz_yolo: {cases - bad}/{cases} cases match", flush=True)
sys.exit(1 if bad else 0)
