"""Decoded-frame ingest from real files (the step BEFORE the accelerated path).

Decode itself stays on the host with the reference's own call -- ``cv2.VideoCapture.read()`` (vivit
.../dataset.py:317-350, resnet50-2d-lstm/inference.py:55-72).  The reference's per-frame
``cv2.cvtColor(frame, COLOR_BGR2RGB)`` pass is NOT run on the host: frames are uploaded as decoded (BGR),
kept resident in HBM, and the fused clip kernel swaps the channels in its gather (``src_order="BGR"``)."""
import numpy as np
import torch

from .datasets import FrameSource


class OpenCVFrameSource(FrameSource):
    def __init__(self, paths, labels=None, device="cuda", cache=True, bgr=True):
        # bgr=True: keep cv2's channel order and let the kernel swap; False: cvtColor on the host
        self.channel_order = "BGR" if bgr else "RGB"
        self.video_paths = [str(p) for p in paths]
        self.labels = list(labels) if labels is not None else [0] * len(self.video_paths)
        self.device, self._cache, self._do_cache, self._fps = device, {}, cache, {}

    def _decode(self, i):
        import cv2

        cap = cv2.VideoCapture(self.video_paths[i])
        if not cap.isOpened():
            raise ValueError(f"Could not open video: {self.video_paths[i]}")
        self._fps[i] = float(cap.get(cv2.CAP_PROP_FPS))
        frames = []
        while True:
            ok, fr = cap.read()
            if not ok:
                break
            frames.append(fr if self.channel_order == "BGR" else cv2.cvtColor(fr, cv2.COLOR_BGR2RGB))
        cap.release()
        if not frames:
            raise ValueError(f"Invalid video: {self.video_paths[i]} - no frames")
        return torch.from_numpy(np.stack(frames)).to(self.device)

    def frames(self, i):
        if i in self._cache:
            return self._cache[i]
        v = self._decode(i)
        if self._do_cache:
            self._cache[i] = v
        return v

    def fps(self, i):
        if i not in self._fps:
            self.frames(i)
        return self._fps[i]
