"""Decoded-frame ingest from real files (the step BEFORE the accelerated path).

Decode itself stays on the host with the reference's own calls, and -- like the reference -- touches ONLY the frames
the sampler picked: ``cap.set(CAP_PROP_POS_FRAMES, idx)`` + ``cap.read()`` per index, the sequential ``grab()``
fallback and the grey placeholder of vivit .../dataset.py:316-341 (frame count and fps come from the container,
:195-214, not from decoding the file).  Decoded frames land in a bounded ring of PINNED host memory that the fused clip
kernel reads in place over PCIe (``ClipBuilder.build_from_host``): no whole-video decode, no whole-video residency in
HBM (a 3-minute 1080p file is 33 GB decoded; a 32-frame clip of it is 199 MB), no staging copy.  The reference's
per-frame ``cv2.cvtColor(frame, COLOR_BGR2RGB)`` pass is NOT run on the host: frames stay BGR as decoded and the
kernel swaps the channels in its gather (``src_order="BGR"``).

On-device decode (NVDEC) would keep the surfaces in HBM and lift the end-to-end rate off the PCIe floor;
``nvdec_probe()`` reports whether the driver's decoder library is loadable on the box.
"""
import collections

import numpy as np
import torch

from .datasets import FrameSource


class PinnedRing:
    """Bounded arena of page-locked host memory handed out in FIFO order.

    ``alloc(nbytes)`` returns a uint8 view; ``fence(event)`` ties every view handed out since the previous fence to a
    CUDA event (recorded behind the kernel that reads them); space is reused only after its event completed --
    ``alloc`` waits for the oldest fences when the ring is full.  ``capacity`` is the residency cap of the ingest."""

    def __init__(self, capacity, write_combined=False):
        from . import hostmem

        self.capacity = int(capacity)
        self.buf = hostmem.pinned_empty((self.capacity,), write_combined=write_combined)
        self.head = 0
        self.live = collections.deque()      # (start, end, event or None) in allocation order
        self.high_water = 0

    def _free_oldest(self):
        start, end, ev = self.live.popleft()
        if ev is None:
            raise RuntimeError("PinnedRing is full of frames no kernel has been fenced behind: call fence(event) after "
                               "enqueuing the consumer, or raise ring_bytes")
        ev.synchronize()

    def _overlaps(self, a, b):
        return any(s < b and a < e for s, e, _ in self.live)

    def alloc(self, nbytes):
        nbytes = (int(nbytes) + 255) & ~255
        if nbytes > self.capacity:
            raise ValueError(f"one clip needs {nbytes} bytes of decoded frames; the pinned ring holds {self.capacity} "
                             "(raise ring_bytes)")
        start = self.head
        if start + nbytes > self.capacity:
            start = 0
        while self._overlaps(start, start + nbytes):
            self._free_oldest()
        self.head = start + nbytes
        self.live.append((start, start + nbytes, None))
        self.high_water = max(self.high_water, sum(e - s for s, e, _ in self.live))
        return self.buf[start:start + nbytes]

    def release_last(self):
        """Give back the most recent allocation (nothing will ever be fenced behind it: the decode that asked for it failed)."""
        if self.live and self.live[-1][2] is None:
            start, _, _ = self.live.pop()
            self.head = start

    def fence(self, event):
        self.live = collections.deque((s, e, ev if ev is not None else event) for s, e, ev in self.live)
        while self.live and self.live[0][2] is not None and self.live[0][2].query():
            self.live.popleft()


class OpenCVFrameSource(FrameSource):
    """``cv2.VideoCapture`` ingest.  ``read_frames(i, indices)`` decodes the requested frames only (index-directed,
    bounded); ``frames(i)`` -- the whole video resident in HBM -- remains for short files and is refused beyond
    ``max_resident_bytes``."""

    index_directed = True

    def __init__(self, paths, labels=None, device="cuda", cache=True, bgr=True, ring_bytes=1 << 30,
                 max_resident_bytes=4 << 30, write_combined=False):
        # bgr=True: keep cv2's channel order and let the kernel swap; False: cvtColor on the host
        self.channel_order = "BGR" if bgr else "RGB"
        self.video_paths = [str(p) for p in paths]
        self.labels = list(labels) if labels is not None else [0] * len(self.video_paths)
        self.device, self._cache, self._do_cache, self._props = device, {}, cache, {}
        self.ring_bytes, self.max_resident_bytes, self._wc = int(ring_bytes), int(max_resident_bytes), write_combined
        self._ring = None
        # what the ingest actually did (tests assert decoded == requested)
        self.decoded_frames = self.grabbed_frames = self.seeks = self.placeholders = 0

    # -- container metadata: vivit .../dataset.py:195-214 (get_video_properties) -------------------------------
    def properties(self, i):
        """(total_frames, fps, width, height) from the container header; fps 0 -> 30.0 as the reference does."""
        if i not in self._props:
            import cv2

            cap = cv2.VideoCapture(self.video_paths[i])
            if not cap.isOpened():
                raise ValueError(f"Could not open video: {self.video_paths[i]}")
            total = int(cap.get(cv2.CAP_PROP_FRAME_COUNT))
            fps = float(cap.get(cv2.CAP_PROP_FPS))
            w, h = int(cap.get(cv2.CAP_PROP_FRAME_WIDTH)), int(cap.get(cv2.CAP_PROP_FRAME_HEIGHT))
            cap.release()
            if fps <= 0:
                fps = 30.0
            self._props[i] = (total, fps, w, h)
        return self._props[i]

    def num_frames(self, i):
        return self.properties(i)[0]

    def fps(self, i):
        return self.properties(i)[1]

    @property
    def ring(self):
        if self._ring is None:
            self._ring = PinnedRing(self.ring_bytes, self._wc)
        return self._ring

    # -- index-directed decode: vivit .../dataset.py:316-341 ---------------------------------------------------
    def read_frames(self, i, indices):
        """Decode the DISTINCT frames ``indices`` name into the pinned ring.  Returns ``(frames, remap)``: a pinned uint8
        ``(K, H, W, 3)`` tensor (channel order as decoded) and ``remap`` with ``frames[remap[t]]`` = frame
        ``indices[t]``.  Seek + read per index (no seek when the decoder already stands on the frame), sequential
        ``grab()`` from the start when the seek fails, a grey 127 frame when that fails too."""
        import cv2

        total, _, w, h = self.properties(i)
        if total <= 0 or w <= 0 or h <= 0:
            raise ValueError(f"Invalid video: {self.video_paths[i]} - no frames")
        want = np.clip(np.asarray(indices, dtype=np.int64), 0, total - 1)
        uniq, remap = np.unique(want, return_inverse=True)
        frames = self.ring.alloc(len(uniq) * h * w * 3).narrow(0, 0, len(uniq) * h * w * 3).view(len(uniq), h, w, 3)
        dst = frames.numpy()
        cap = cv2.VideoCapture(self.video_paths[i])
        if not cap.isOpened():
            raise ValueError(f"Could not open video: {self.video_paths[i]}")
        pos = -1   # index of the frame the next read() returns, when known
        for k, fi in enumerate(uniq.tolist()):
            if pos != fi:
                cap.set(cv2.CAP_PROP_POS_FRAMES, fi)
                self.seeks += 1
            ok, fr = cap.read()
            if ok:
                self.decoded_frames += 1
            else:
                cap.set(cv2.CAP_PROP_POS_FRAMES, 0)
                for _ in range(fi):
                    cap.grab()
                    self.grabbed_frames += 1
                ok, fr = cap.read()
                self.decoded_frames += int(ok)
            pos = fi + 1 if ok else -1
            if ok and fr.shape == (h, w, 3):
                dst[k] = fr if self.channel_order == "BGR" else cv2.cvtColor(fr, cv2.COLOR_BGR2RGB)
            else:
                dst[k] = 127
                self.placeholders += 1
        cap.release()
        return frames, remap.astype(np.int32)

    # -- sequential decode: data_handling.py:12-38 / vivit_transformer/inference.py:81-103 (read_video_pyav) -----------
    def read_frames_sequential(self, i, indices):
        """The frame SELECTION of the reference's PyAV reader: rewind, decode from frame 0 in order, keep frame ``k`` when
        ``indices[0] <= k <= indices[-1]`` and ``k in indices`` -- each frame once however often it is listed, so a list
        with repeats yields fewer frames than entries (the reference's ``np.stack`` has that shape too) -- and stop after
        ``indices[-1]``.  No seek (robust on streams whose seek lands on key frames only); frames between the samples
        are ``grab()``-bed (demuxed + decoded by the codec, never converted or copied), the kept ones land in the pinned
        ring as decoded.  Returns ``(frames, kept)``: pinned uint8 ``(K, H, W, 3)`` and the K frame numbers, ascending.
        The reader's ``frame.reformat(width=224, height=224)`` (libswscale) is NOT applied here: resizing is the clip
        kernel's job with the cv2 arithmetic of the Dataset path; raises ValueError("No frames were extracted!") like
        the reference when nothing is kept."""
        import cv2

        total, _, w, h = self.properties(i)
        idx = [int(v) for v in indices]
        if not idx:
            raise ValueError("No frames were extracted!")
        start, end = idx[0], idx[-1]
        wanted = sorted({k for k in idx if start <= k <= end and 0 <= k < max(total, end + 1)})
        frames = self.ring.alloc(max(len(wanted), 1) * h * w * 3)
        dst = frames.narrow(0, 0, len(wanted) * h * w * 3).view(len(wanted), h, w, 3).numpy() if wanted else None
        cap = cv2.VideoCapture(self.video_paths[i])
        if not cap.isOpened():
            self.ring.release_last()
            raise ValueError(f"Could not open video: {self.video_paths[i]}")
        kept, want_set = [], set(wanted)
        for k in range(end + 1):
            if not cap.grab():           # the stream ended before indices[-1]: the reference's loop ends there too
                break
            self.grabbed_frames += 1
            if k in want_set:
                ok, fr = cap.retrieve()
                if not ok or fr.shape != (h, w, 3):
                    break
                dst[len(kept)] = fr if self.channel_order == "BGR" else cv2.cvtColor(fr, cv2.COLOR_BGR2RGB)
                kept.append(k)
                self.decoded_frames += 1
        cap.release()
        if not kept:
            self.ring.release_last()
            raise ValueError("No frames were extracted!")
        return frames.narrow(0, 0, len(kept) * h * w * 3).view(len(kept), h, w, 3), np.asarray(kept, dtype=np.int64)

    def fence(self, event):
        """Tell the ring which event the frames handed out so far are read behind."""
        if self._ring is not None:
            self._ring.fence(event)

    # -- whole-video residency (short files only) --------------------------------------------------------------
    def _decode(self, i):
        import cv2

        total, _, w, h = self.properties(i)
        if total * h * w * 3 > self.max_resident_bytes:
            raise ValueError(f"{self.video_paths[i]}: {total} frames of {h}x{w} are {total * h * w * 3 / 2**30:.1f} GiB decoded, "
                             f"over max_resident_bytes; use the index-directed path (read_frames / Dataset.get_batch)")
        cap = cv2.VideoCapture(self.video_paths[i])
        if not cap.isOpened():
            raise ValueError(f"Could not open video: {self.video_paths[i]}")
        frames = []
        while True:
            ok, fr = cap.read()
            if not ok:
                break
            self.decoded_frames += 1
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


def nvdec_probe():
    """Is the driver's hardware video decoder library loadable here?  (No decode is attempted.)  NVDEC output surfaces
    live in device memory, so an NVDEC FrameSource would hand the clip kernel HBM-resident frames."""
    import ctypes

    out = {}
    for name in ("libnvcuvid.so.1", "libnvcuvid.so"):
        try:
            lib = ctypes.CDLL(name)
            out[name] = {"loaded": True, "cuvidCreateDecoder": hasattr(lib, "cuvidCreateDecoder"),
                         "cuvidCreateVideoParser": hasattr(lib, "cuvidCreateVideoParser")}
        except OSError as e:
            out[name] = {"loaded": False, "error": str(e)[:120]}
    return out
