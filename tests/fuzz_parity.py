"""Randomised parity sweep of the fused clip kernel against the oracle (GPU box): random frame sizes, ROI boxes,
index tables, presets, layouts and source orders.  python tests/fuzz_parity.py [cases] [seed]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lvc_b200 import ClipBuilder, sampling  # noqa: E402
from oracle import processors as op  # noqa: E402
from oracle.resize import cv2_resize_linear_f32_trunc  # noqa: E402


def main():
    cases = int(sys.argv[1]) if len(sys.argv) > 1 else 60
    rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
    bad = 0
    for k in range(cases):
        B = int(rng.integers(1, 4))
        T = int(rng.choice([1, 2, 3, 4]))
        preset = str(rng.choice(["dataset_u8", "resnet_infer", "timesformer", "vivit", "ptv_eval", "ptv_train", "float_u8"]))
        vids, host = [], []
        for _ in range(B):
            h = int(rng.choice([1, 2, 7, 100, 223, 224, 225, 300, 448, 449, 480, 720])) if rng.random() < 0.7 else int(rng.integers(1, 900))
            w = int(rng.choice([1, 3, 150, 224, 300, 447, 448, 640, 1000, 1280])) if rng.random() < 0.7 else int(rng.integers(1, 1500))
            if preset.startswith("ptv"):
                h, w = max(h, 8), max(w, 8)
            n = int(rng.integers(1, 6))
            v = rng.integers(0, 256, (n, h, w, 3), dtype=np.uint8)
            host.append(v)
            vids.append(torch.from_numpy(v).cuda())
        idx = [[int(rng.integers(-2, v.shape[0] + 2)) for _ in range(T)] for v in host]   # out-of-range indices clamp
        cidx = [[min(max(i, 0), v.shape[0] - 1) for i in row] for row, v in zip(idx, host)]
        boxes = None
        if preset in ("dataset_u8", "resnet_infer", "vivit") and rng.random() < 0.5:
            boxes = np.zeros((B, T, 4), np.int32)
            for b, v in enumerate(host):
                for t in range(T):
                    H, W = v.shape[1:3]
                    x1, y1 = int(rng.integers(-5, W + 3)), int(rng.integers(-5, H + 3))
                    boxes[b, t] = (x1, y1, x1 + int(rng.integers(-2, W + 5)), y1 + int(rng.integers(-2, H + 5)))
                    if rng.random() < 0.2:
                        boxes[b, t] = -1
        bgr = bool(rng.random() < 0.3)
        src = [torch.from_numpy(np.ascontiguousarray(v[..., ::-1])).cuda() for v in host] if bgr else vids
        kw = dict(src_order="BGR") if bgr else {}
        try:
            if preset == "float_u8":
                out = ClipBuilder("dataset_u8", num_frames=T, float_frames=True, **kw)(src, idx).cpu().numpy()
                want = np.stack([np.stack([cv2_resize_linear_f32_trunc(v[i]) for i in row]) for row, v in zip(cidx, host)])
            elif preset == "ptv_eval":
                out = ClipBuilder("ptv_eval", num_frames=T, **kw)(src, idx).cpu().numpy()
                want = np.stack([op.ptv_eval_chain(v[row]) for row, v in zip(cidx, host)])
            elif preset == "ptv_train":
                r = sampling.TorchRng(k)
                aug = [sampling.train_aug_params(v.shape[1], v.shape[2], r) for v in host]
                out = ClipBuilder("ptv_eval", num_frames=T, **kw)(src, idx, aug=aug).cpu().numpy()
                torch.manual_seed(k)
                want = np.stack([op.ptv_train_chain(v[row], seed=None)[0] for row, v in zip(cidx, host)])
            else:
                out = ClipBuilder(preset, num_frames=T, **kw)(src, idx, boxes=boxes).cpu().numpy()
                res = [op.resize_clip_cv2(v[row], None if boxes is None else boxes[b]) for b, (row, v) in enumerate(zip(cidx, host))]
                if preset == "dataset_u8":
                    want = np.stack(res)
                elif preset == "resnet_infer":
                    want = np.stack([op.manual_normalize(r_) for r_ in res])
                elif preset == "timesformer":
                    want = np.stack([op.videomae_processor(r_) for r_ in res])
                else:
                    want = np.stack([op.vivit_processor(r_) for r_ in res])
            tol = 1.1e-2 if preset.startswith("ptv") else 0.0
            err = float(np.abs(out.astype(np.float64) - want.astype(np.float64)).max())
            ok = out.shape == want.shape and err <= tol
        except Exception as e:  # noqa: BLE001
            ok, err = False, repr(e)
        if not ok:
            bad += 1
            print("MISMATCH", k, preset, [v.shape for v in host], "T", T, "bgr", bgr, "boxes", None if boxes is None else boxes.tolist(), "err", err, flush=True)
    print(f"fuzz: {cases - bad}/{cases} cases match", flush=True)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
