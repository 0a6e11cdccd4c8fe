"""Seeded synthetic decoded frames and ROI boxes: the same bytes feed the CUDA
path, the oracle and the golden generator (tests/golden/make_golden.py has the same code)."""
import numpy as np


def synth_video(seed, n, h, w, mode="noise"):
    """uint8 RGB frames (n, h, w, 3).  ``noise``: iid 0..255 (worst case for rounding parity);
    ``ramp``: (x + 2y + 3t + 85c) mod 256 (catches index / stride / channel-order bugs)."""
    if mode == "noise":
        return np.random.default_rng(seed).integers(0, 256, (n, h, w, 3), dtype=np.uint8)
    t, y, x, c = np.meshgrid(np.arange(n), np.arange(h), np.arange(w), np.arange(3), indexing="ij")
    return ((x + 2 * y + 3 * t + c * 85) % 256).astype(np.uint8)


def synth_boxes(seed, n, h, w, p_none=0.1):
    """Per-frame int32 xyxy boxes in the reference's record format (src/yolo_detector.py:76-85):
    width in [w/4, w/2], height in [h/4, h/2], inside the frame; ``p_none`` of the frames have no
    detection, encoded as (-1,-1,-1,-1) = use the full frame."""
    r = np.random.default_rng(seed)
    bw = r.integers(max(w // 4, 1), max(w // 2, 2), n)
    bh = r.integers(max(h // 4, 1), max(h // 2, 2), n)
    x1 = (r.random(n) * (w - bw)).astype(np.int64)
    y1 = (r.random(n) * (h - bh)).astype(np.int64)
    b = np.stack([x1, y1, x1 + bw, y1 + bh], 1).astype(np.int32)
    b[r.random(n) < p_none] = -1
    return b
