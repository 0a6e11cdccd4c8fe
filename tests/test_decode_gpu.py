"""Row f1: real files.  A tiny mp4 dataset tree is written with cv2.VideoWriter, read back through
the drop-in VideoDataset (host decode with the reference's own cv2 calls, frames resident in HBM, fused clip
kernel) and compared with the reference's OpenCV fallback recipe (vivit .../dataset.py:317-350: seek, read,
BGR->RGB, cv2.resize) run on the CPU."""
import random

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _write_tree(root, n_frames=24, h=120, w=160):
    import cv2

    rng = np.random.default_rng(3)
    paths = []
    for cls, count in (("referral", 2), ("non_referral", 1)):
        d = root / "val" / cls
        d.mkdir(parents=True)
        for k in range(count):
            p = d / f"v{k}.mp4"
            wr = cv2.VideoWriter(str(p), cv2.VideoWriter_fourcc(*"mp4v"), 25.0, (w, h))
            assert wr.isOpened()
            base = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
            for i in range(n_frames):
                wr.write(np.roll(base, 3 * i, axis=1))
            wr.release()
            paths.append(str(p))
    return paths


def test_dataset_on_real_mp4_files(tmp_path):
    import cv2

    from lvc_b200.datasets import VideoDataset
    from oracle import samplers as osm

    _write_tree(tmp_path)
    ds = VideoDataset(str(tmp_path), mode="val", sampling_method="random_window", num_frames=8)
    assert len(ds) == 3 and sorted(ds.labels) == [0, 1, 1]
    rng = random.Random(42)
    for i in range(len(ds)):
        s = ds[i]
        cap = cv2.VideoCapture(ds.video_paths[i])
        total = int(cap.get(cv2.CAP_PROP_FRAME_COUNT))
        want_idx = osm.dataset_indices("random_window", total, 8, rng)
        assert s["frame_indices"] == want_idx
        frames = []
        for fi in want_idx:  # the reference's fallback: seek + read + BGR2RGB + resize
            cap.set(cv2.CAP_PROP_POS_FRAMES, fi)
            ok, fr = cap.read()
            assert ok
            frames.append(cv2.resize(cv2.cvtColor(fr, cv2.COLOR_BGR2RGB), (224, 224)))
        cap.release()
        assert np.array_equal(s["pixel_values"].cpu().numpy(), np.stack(frames))
        assert ds.source.fps(i) == 25.0


def test_index_directed_ingest_decodes_only_the_sampled_frames(tmp_path):
    """The ingest decodes the frames the sampler picked and nothing else (the reference seeks + reads per index,
    vivit .../dataset.py:316-341), into a bounded pinned ring, and never holds a whole video."""
    import cv2

    from lvc_b200.datasets import VideoDataset

    _write_tree(tmp_path, n_frames=60)
    ds = VideoDataset(str(tmp_path), mode="val", sampling_method="uniform", num_frames=8)
    src = ds.source
    assert src.index_directed and src.num_frames(0) == 60           # from the container header, nothing decoded
    assert src.decoded_frames == 0
    b = ds.get_batch([0, 1, 2])
    torch.cuda.synchronize()
    requested = sum(len(set(fi)) for fi in b["frame_indices"])
    assert src.decoded_frames == requested == 24                    # decoded == indices requested (distinct)
    assert src.grabbed_frames == 0 and src.placeholders == 0
    assert not src._cache                                           # no whole-video residency
    assert src.ring.high_water <= 3 * 8 * 120 * 160 * 3 + 3 * 256   # three clips' frames, not three videos
    for k, i in enumerate([0, 1, 2]):
        cap = cv2.VideoCapture(ds.video_paths[i])
        frames = []
        for fi in b["frame_indices"][k]:
            cap.set(cv2.CAP_PROP_POS_FRAMES, fi)
            ok, fr = cap.read()
            assert ok
            frames.append(cv2.resize(cv2.cvtColor(fr, cv2.COLOR_BGR2RGB), (224, 224)))
        cap.release()
        assert np.array_equal(b["pixel_values"][k].cpu().numpy(), np.stack(frames))


def test_pinned_ring_is_a_residency_cap(tmp_path):
    from lvc_b200.datasets import VideoDataset
    from lvc_b200.decode import OpenCVFrameSource, PinnedRing

    paths = _write_tree(tmp_path, n_frames=40)
    clip_bytes = 8 * 120 * 160 * 3
    # ring of 2.5 clips: the third batch must reuse the first one's space, after its kernel is done
    src = OpenCVFrameSource(paths, [1, 1, 0], ring_bytes=int(2.5 * clip_bytes))
    ds = VideoDataset(frame_source=src, mode="val", sampling_method="random", num_frames=8)
    ref = VideoDataset(frame_source=OpenCVFrameSource(paths, [1, 1, 0]), mode="val", sampling_method="random", num_frames=8)
    for rep in range(4):
        for i in range(3):
            a, b = ds[i], ref[i]
            assert a["frame_indices"] == b["frame_indices"]
            assert torch.equal(a["pixel_values"], b["pixel_values"]), (rep, i)
    assert src.ring.high_water <= src.ring.capacity
    with pytest.raises(ValueError, match="ring"):
        OpenCVFrameSource(paths, ring_bytes=1000).read_frames(0, [0, 1])
    r = PinnedRing(4096)
    r.alloc(3000)
    with pytest.raises(RuntimeError, match="fence"):
        r.alloc(3000)     # would overwrite frames nobody has been fenced behind
    # whole-video residency is refused beyond the cap
    with pytest.raises(ValueError, match="max_resident_bytes"):
        OpenCVFrameSource(paths, max_resident_bytes=1000).frames(0)


def test_seek_failure_falls_back_to_grab_then_placeholder(tmp_path, monkeypatch):
    """A container whose seek fails: sequential grab() from the start; when that fails too a grey 127 frame
    (vivit .../dataset.py:328-341)."""
    import cv2

    from lvc_b200.decode import OpenCVFrameSource

    paths = _write_tree(tmp_path, n_frames=12)
    real = cv2.VideoCapture

    class NoSeek:
        def __init__(self, path):
            self.c, self.broken = real(path), False

        def isOpened(self):
            return self.c.isOpened()

        def get(self, p):
            return self.c.get(p)

        def set(self, p, v):
            if p == cv2.CAP_PROP_POS_FRAMES and v != 0:
                self.broken = True      # a failed seek leaves the decoder unable to read
                return False
            self.broken = False
            return self.c.set(p, v)

        def read(self):
            return (False, None) if self.broken else self.c.read()

        def grab(self):
            return self.c.grab()

        def release(self):
            self.c.release()

    monkeypatch.setattr(cv2, "VideoCapture", NoSeek)
    src = OpenCVFrameSource(paths[:1])
    fr, remap = src.read_frames(0, [5, 5, 9])
    assert remap.tolist() == [0, 0, 1] and src.grabbed_frames == 5 + 9 and src.placeholders == 0
    monkeypatch.setattr(cv2, "VideoCapture", real)
    cap = real(paths[0])
    seq = [cap.read()[1] for _ in range(10)]
    cap.release()
    assert np.array_equal(fr[0].numpy(), seq[5]) and np.array_equal(fr[1].numpy(), seq[9])
    # reads that never succeed -> placeholder
    class Dead(NoSeek):
        def read(self):
            return False, None

    monkeypatch.setattr(cv2, "VideoCapture", Dead)
    src = OpenCVFrameSource(paths[:1])
    fr, _ = src.read_frames(0, [3])
    assert src.placeholders == 1 and bool((fr == 127).all())


def test_sequential_reader_keeps_the_frames_the_pyav_reader_keeps(tmp_path):
    """read_frames_sequential mirrors the frame selection of read_video_pyav (data_handling.py:12-38): decode in order
    from frame 0, keep k when indices[0] <= k <= indices[-1] and k in indices (once each), stop after indices[-1]."""
    import cv2

    from lvc_b200 import ClipBuilder
    from lvc_b200.decode import OpenCVFrameSource

    paths = _write_tree(tmp_path, n_frames=30)
    src = OpenCVFrameSource(paths, [1, 1, 0])
    indices = [4, 4, 9, 10, 17, 17, 21]                      # repeats: the reference's np.stack holds 5 frames, not 7
    frames, kept = src.read_frames_sequential(0, indices)
    assert kept.tolist() == [4, 9, 10, 17, 21] and tuple(frames.shape) == (5, 120, 160, 3) and frames.is_pinned()
    assert src.decoded_frames == 5 and src.grabbed_frames == 22 and src.seeks == 0     # frames 0..21 demuxed, 5 converted
    cap = cv2.VideoCapture(paths[0])                          # the reference's loop, with cv2 standing in for PyAV
    want = []
    for k in range(indices[-1] + 1):
        ok, fr = cap.read()
        assert ok
        if k >= indices[0] and k in indices:
            want.append(fr)
    cap.release()
    assert np.array_equal(frames.numpy(), np.stack(want))
    # ... and they feed the clip kernel in place (BGR as decoded, swapped in the gather)
    clip = ClipBuilder("dataset_u8", num_frames=5, src_order="BGR").build_from_host([frames], [list(range(5))])
    torch.cuda.synchronize()
    ref = np.stack([cv2.resize(cv2.cvtColor(f, cv2.COLOR_BGR2RGB), (224, 224)) for f in want])
    assert np.array_equal(clip[0].cpu().numpy(), ref)
    # a list that starts past the end of the stream keeps nothing: the reference raises
    with pytest.raises(ValueError, match="No frames were extracted"):
        src.read_frames_sequential(0, [40, 45])
    # indices[-1] bounds the scan: out-of-order lists drop what lies outside [indices[0], indices[-1]]
    _, kept = src.read_frames_sequential(1, [10, 3, 25, 12])
    assert kept.tolist() == [10, 12]
