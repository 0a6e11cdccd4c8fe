#!/usr/bin/env python
"""Generate tests/golden/*.npz by EXECUTING the reference's own functions and the
third-party calls it makes, in the build container (needs /root/reference; cannot run on
the GPU box).  Run:  python tests/golden/make_golden.py

What each fixture pins:
  samplers.npz    a1  VideoDataset.get_sampling_indices of the vivit, swin and 2d-lstm copies,
                  a2  data_handling.sample_frame_indices, a3 resnet50-2d-lstm/inference.py samplers,
                      visualize_sampling.py pure functions
  samplers3d.npz  a1  the resnet50-3d-video copy: per-video seed (hash pinned to crc32), index cache, re-seed to 42
  resize.npz      a6  cv2.resize(.,(224,224)) and Pillow BILINEAR on seeded images
  pixels.npz      a7  VivitImageProcessor, a8 VideoMAEImageProcessorPil, a9 the real
                      EnhancedVideoInference.predict_video of resnet50-2d-lstm/inference.py run on a fake
                      cv2.VideoCapture (config 1 of BASELINE.json end to end), a10 torch chain
  embed.npz       a12-a14 the HF / torchvision patch-embed modules with seeded weights
Missing third-party packages (pytorchvideo, av, matplotlib, seaborn) and the reference's model
classes are stubbed in sys.modules: none of them takes part in the arithmetic pinned here.
"""
import hashlib
import importlib
import os
import random
import sys
import types

import numpy as np

REF = os.environ.get("LVC_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))


class _Anything(types.ModuleType):
    """Module stub: any attribute is a dummy class / callable; works as a package."""

    __path__ = []

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return type(name, (), {"__init__": lambda self, *a, **k: None})


def stub(*names):
    for n in names:
        parts = n.split(".")
        for i in range(1, len(parts) + 1):
            sys.modules.setdefault(".".join(parts[:i]), _Anything(".".join(parts[:i])))


def load(path, name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, path))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def sha(a):
    a = np.ascontiguousarray(a)
    return hashlib.sha256(a.tobytes()).hexdigest()


class _Log:
    def info(self, *a, **k):
        pass

    warning = error = debug = info


N_GRID = [1, 5, 20, 31, 32, 33, 64, 277, 278, 300, 1000]
T_GRID = [1, 8, 16, 32]
SEEDS = [0, 42, 1234]
METHODS = ["uniform", "random_window", "random"]


def make_samplers():
    stub("pytorchvideo.data.encoded_video", "pytorchvideo.transforms", "av", "matplotlib.pyplot",
         "seaborn")
    vivit = load("vivit_transformer/vivit_classifier/data_config/dataset.py", "ref_vivit_ds")
    swin = load("videoswintransformer/swin_video_classifier/data_config/dataset.py", "ref_swin_ds")
    lstm = load("resnet50-2d-lstm/src/data_config/dataset.py", "ref_lstm_ds")
    dh = load("data_handling.py", "ref_data_handling")
    vis = load("visualize_sampling.py", "ref_vis")
    stub("src.models.model", "src.utils.logger", "src.utils.visualization", "src.config.config")
    inf = load("resnet50-2d-lstm/inference.py", "ref_lstm_inf")

    def mk(cls, **attrs):
        o = object.__new__(cls)
        o.logger = _Log()
        for k, v in attrs.items():
            setattr(o, k, v)
        return o

    rows, outs = [], []  # rows: (copy, method, N, T, seed, call#) ; outs: padded int64[32]

    def put(copy, m, N, T, seed, call, idx):
        rows.append((copy, METHODS.index(m), N, T, seed, call))
        a = np.full(32, -1, np.int64)
        a[: len(idx)] = idx
        outs.append(a)

    for m in METHODS:
        for N in N_GRID:
            for T in T_GRID:
                for seed in SEEDS:
                    # copy 0: vivit dataset -- seeded once, stream continues across 2 calls
                    ds = mk(vivit.VideoDataset, sampling_method=m, num_frames=T)
                    random.seed(seed)
                    for call in range(2):
                        put(0, m, N, T, seed, call, ds.get_sampling_indices("x.mp4", N))
                    # copy 1: swin dataset (returns (indices, dynamic_fps)); cv2 reports fps 0 for a
                    # missing file -> the reference substitutes 30.0
                    ds = mk(swin.VideoDataset, sampling_method=m, num_frames=T)
                    random.seed(seed)
                    idx, dyn = ds.get_sampling_indices("x.mp4", N)
                    put(1, m, N, T, seed, 0, idx)
                    assert (dyn is None) == (N >= T)
                    if dyn is not None:
                        assert dyn == T / (N / 30.0)
                # copy 2: 2d-lstm dataset re-seeds 42 inside every call (N >= T only: the short
                # branch divides by the fps of a real file)
                if N >= T:
                    ds = mk(lstm.VideoDataset, sampling_method=m, sequence_length=T)
                    random.seed(7)
                    for call in range(2):
                        put(2, m, N, T, 42, call, ds.get_sampling_indices("x.mp4", N))
                # copy 3: resnet50-2d-lstm/inference.py numpy-RNG samplers (seed 42 inside)
                o = mk(inf.EnhancedVideoInference)
                f = {"uniform": o.uniform_sampling, "random_window": o.random_window_sampling,
                     "random": o.random_sampling}[m]
                put(3, m, N, T, 42, 0, [int(v) for v in f(N, T)])
                # copy 4: visualize_sampling.py pure functions (global CPython stream)
                for seed in SEEDS:
                    random.seed(seed)
                    f = {"uniform": vis.uniform_sampling, "random_window": vis.random_window_sampling,
                         "random": vis.random_sampling}[m]
                    put(4, m, N, T, seed, 0, f(N, T))
    # a2: data_handling.sample_frame_indices
    a2_rows, a2_out = [], []
    for clip_len, rate, seg in [(32, 1, 300), (32, 2, 100), (8, 4, 64), (16, 1, 17), (32, 8, 1000)]:
        for seed in SEEDS:
            np.random.seed(seed)
            for call in range(2):
                idx = dh.sample_frame_indices(clip_len, rate, seg)
                a2_rows.append((clip_len, rate, seg, seed, call))
                a = np.full(32, -1, np.int64)
                a[:clip_len] = idx
                a2_out.append(a)
    np.savez_compressed(os.path.join(HERE, "samplers.npz"),
                        rows=np.array(rows, np.int64), indices=np.array(outs),
                        a2_rows=np.array(a2_rows, np.int64), a2_indices=np.array(a2_out))
    print("samplers.npz", len(rows), len(a2_rows))


def synth_video(seed, n, h, w, mode="noise"):
    """Seeded synthetic decoded RGB frames.  Same generator as lvc_b200.synth."""
    if mode == "noise":
        return np.random.default_rng(seed).integers(0, 256, (n, h, w, 3), dtype=np.uint8)
    t, y, x, c = np.meshgrid(np.arange(n), np.arange(h), np.arange(w), np.arange(3), indexing="ij")
    return ((x + 2 * y + 3 * t + c * 85) % 256).astype(np.uint8)


RESIZE_CASES = [  # (seed, H, W, mode)
    (11, 720, 1280, "noise"), (12, 1080, 1920, "noise"), (13, 480, 640, "ramp"), (14, 448, 448, "noise"),
    (15, 224, 224, "noise"), (16, 100, 150, "noise"), (17, 223, 225, "noise"), (18, 37, 1000, "noise"),
    (19, 1, 1, "noise"), (20, 2, 3, "noise"), (21, 300, 200, "ramp"), (22, 449, 447, "noise"),
]


def make_resize():
    import cv2
    from PIL import Image

    cv2.setNumThreads(1)
    out = {"cases": np.array([(s, h, w, m == "ramp") for s, h, w, m in RESIZE_CASES], np.int64)}
    for i, (s, h, w, m) in enumerate(RESIZE_CASES):
        img = synth_video(s, 1, h, w, m)[0]
        r = cv2.resize(img, (224, 224))
        out[f"cv2_sha_{i}"] = np.array(sha(r))
        if i in (0, 5, 7):
            out[f"cv2_full_{i}"] = r
        else:
            out[f"cv2_crop_{i}"] = r[:16, :16].copy()
    # ROI crops (numpy slice + cv2.resize; the semantics this repo defines)
    img = synth_video(31, 1, 720, 1280)[0]
    boxes = np.array([[100, 50, 740, 410], [0, 0, 1280, 720], [3, 5, 227, 229], [1000, 600, 1280, 720],
                      [17, 33, 465, 481], [5, 5, 60, 40]], np.int64)
    out["roi_boxes"] = boxes
    for i, (x1, y1, x2, y2) in enumerate(boxes):
        out[f"roi_sha_{i}"] = np.array(sha(cv2.resize(np.ascontiguousarray(img[y1:y2, x1:x2]), (224, 224))))
    # Pillow bilinear 224 -> 256 (the ViViT processor's resize)
    img = synth_video(41, 1, 224, 224)[0]
    r = np.asarray(Image.fromarray(img).resize((256, 256), Image.BILINEAR))
    out["pil_sha_224_256"] = np.array(sha(r))
    out["pil_crop_224_256"] = r[:16, :16].copy()
    np.savez_compressed(os.path.join(HERE, "resize.npz"), **out)
    print("resize.npz", len(out))


class FakeCapture:
    """Stands in for cv2.VideoCapture: serves synthetic decoded frames (as BGR, like OpenCV)."""

    videos = {}

    def __init__(self, path):
        self.v = self.videos.get(path)
        self.pos = 0

    def isOpened(self):
        return self.v is not None

    def get(self, prop):
        import cv2

        n, h, w = self.v.shape[:3]
        return {cv2.CAP_PROP_FRAME_COUNT: n, cv2.CAP_PROP_FPS: 30.0, cv2.CAP_PROP_FRAME_WIDTH: w,
                cv2.CAP_PROP_FRAME_HEIGHT: h}[prop]

    def set(self, prop, val):
        self.pos = int(val)

    def read(self):
        if self.pos >= len(self.v):
            return False, None
        f = np.ascontiguousarray(self.v[self.pos][..., ::-1])
        self.pos += 1
        return True, f

    def release(self):
        pass


def make_pixels():
    import cv2
    import torch
    from transformers import VivitImageProcessor
    from transformers.models.videomae.image_processing_pil_videomae import VideoMAEImageProcessorPil

    cv2.setNumThreads(1)
    out = {}
    # a9 / config 1: the reference's real predict_video on a fake capture and a capturing model
    stub("src.models.model", "src.utils.logger", "src.utils.visualization", "src.config.config")
    inf = load("resnet50-2d-lstm/inference.py", "ref_lstm_inf2")
    real_cap = cv2.VideoCapture
    cv2.VideoCapture = FakeCapture
    try:
        for k, (seed, n, h, w, method, T) in enumerate([(101, 40, 360, 640, "uniform", 32),
                                                        (102, 12, 120, 160, "random_window", 8),
                                                        (103, 20, 96, 128, "random", 16)]):
            FakeCapture.videos = {"v.mp4": synth_video(seed, n, h, w)}
            o = object.__new__(inf.EnhancedVideoInference)
            o.device = torch.device("cpu")
            o.sequence_length, o.sampling_method, o.fps = T, method, 30
            o.mean, o.std = [0.45, 0.45, 0.45], [0.225, 0.225, 0.225]
            seen = {}

            def model(x, seen=seen):
                seen["x"] = x.clone()
                return torch.zeros(1, 1)

            o.model = model
            res = o.predict_video("v.mp4")
            assert res["status"] == "success", res
            x = seen["x"].numpy()  # (1, 3, T, 224, 224)
            out[f"infer_case_{k}"] = np.array([seed, n, h, w, METHODS.index(method), T], np.int64)
            out[f"infer_idx_{k}"] = np.array(res["frame_indices"], np.int64)
            out[f"infer_sha_{k}"] = np.array(sha(x))
            out[f"infer_crop_{k}"] = x[0, :, :2, :8, :8].copy()
    finally:
        cv2.VideoCapture = real_cap
    # a7: VivitImageProcessor exactly as constructed at vivit .../trainers/trainer.py:22-26
    fr = synth_video(111, 4, 224, 224)
    p = VivitImageProcessor(num_frames=32, image_size=224, patch_size=16)
    x = p(list(fr), return_tensors="pt")["pixel_values"].numpy()
    out["vivit_sha"], out["vivit_crop"], out["vivit_shape"] = np.array(sha(x)), x[0, :2, :, :8, :8].copy(), np.array(x.shape)
    # a8: VideoMAE-class processor forced to 224 (timesformer .../trainers/trainer.py:91-98), both mean/std guesses
    fr = synth_video(112, 4, 224, 224)
    for tag, (m, s) in {"045": (0.45, 0.225), "05": (0.5, 0.5)}.items():
        p = VideoMAEImageProcessorPil(image_mean=[m] * 3, image_std=[s] * 3)
        p.size = {"height": 224, "width": 224}
        p.crop_size = {"height": 224, "width": 224}
        x = p(images=list(fr), return_tensors="pt", do_resize=True, size={"height": 224, "width": 224},
              do_center_crop=True, crop_size={"height": 224, "width": 224})["pixel_values"].numpy()
        out[f"videomae_sha_{tag}"], out[f"videomae_crop_{tag}"] = np.array(sha(x)), x[0, :2, :, :8, :8].copy()
        out[f"videomae_shape_{tag}"] = np.array(x.shape)
    np.savez_compressed(os.path.join(HERE, "pixels.npz"), **out)
    print("pixels.npz", len(out))


def seeded(shape, seed, scale=1.0):
    return (np.random.default_rng(seed).standard_normal(shape) * scale).astype(np.float32)


def make_embed():
    import torch
    from functools import partial
    from transformers import TimesformerConfig, VivitConfig
    from transformers.models.timesformer.modeling_timesformer import TimesformerPatchEmbeddings
    from transformers.models.vivit.modeling_vivit import VivitTubeletEmbeddings
    from torchvision.models.video.swin_transformer import PatchEmbed3d

    torch.set_num_threads(1)
    out = {}
    with torch.no_grad():
        # a12 ViViT: default tubelet [2,16,16], hidden 768; small image to keep the fixture small
        m = VivitTubeletEmbeddings(VivitConfig(image_size=64, num_frames=4))
        m.projection.weight.copy_(torch.from_numpy(seeded(m.projection.weight.shape, 1, 0.02)))
        m.projection.bias.copy_(torch.from_numpy(seeded(m.projection.bias.shape, 2, 0.02)))
        x = torch.from_numpy(seeded((2, 4, 3, 64, 64), 3))
        out["vivit_out"] = m(x).numpy()
        # a13 TimeSformer: patch 16, hidden 768
        m = TimesformerPatchEmbeddings(TimesformerConfig(image_size=64, num_frames=3))
        m.projection.weight.copy_(torch.from_numpy(seeded(m.projection.weight.shape, 4, 0.02)))
        m.projection.bias.copy_(torch.from_numpy(seeded(m.projection.bias.shape, 5, 0.02)))
        x = torch.from_numpy(seeded((2, 3, 3, 64, 64), 6))
        emb, nf, pw = m(x)
        out["tsf_out"], out["tsf_meta"] = emb.numpy(), np.array([nf, pw], np.int64)
        # a14 Swin-B: PatchEmbed3d([2,4,4], 3, 128, LayerNorm eps 1e-5) as built by swin3d_b
        m = PatchEmbed3d([2, 4, 4], 3, 128, partial(torch.nn.LayerNorm, eps=1e-5))
        m.proj.weight.copy_(torch.from_numpy(seeded(m.proj.weight.shape, 7, 0.02)))
        m.proj.bias.copy_(torch.from_numpy(seeded(m.proj.bias.shape, 8, 0.02)))
        m.norm.weight.copy_(torch.from_numpy(1 + seeded(m.norm.weight.shape, 9, 0.1)))
        m.norm.bias.copy_(torch.from_numpy(seeded(m.norm.bias.shape, 10, 0.1)))
        x = torch.from_numpy(seeded((2, 3, 4, 32, 32), 11))
        out["swin_out"] = m(x).numpy()
        # odd sizes exercise PatchEmbed3d's padding (swin_transformer.py:353-358)
        x = torch.from_numpy(seeded((1, 3, 3, 30, 29), 12))
        out["swin_out_pad"] = m(x).numpy()
    np.savez_compressed(os.path.join(HERE, "embed.npz"), **out)
    print("embed.npz", {k: v.shape for k, v in out.items()})


def make_samplers3d():
    """a1, resnet50-3d-video copy: per-video seed hash(basename) % 1e7 (with `hash` replaced by a FIXED function in
    the reference module's namespace -- CPython salts str hashes per process), index cache, re-seed to 42.  For every
    (method, N, T): a sequence of calls [a.mp4, b.mp4, a.mp4 (cached), c.mp4 after 3 stray draws on the global stream]
    -- the stray draws must not matter (per-video seed) and the cached call must return the first answer."""
    import zlib

    stub("pytorchvideo.data.encoded_video", "pytorchvideo.data.clip_sampling", "pytorchvideo.transforms", "av")
    ref = load("resnet50-3d-video/video_classifier/data_config/dataset.py", "ref_r3d_ds")
    ref.hash = lambda name: zlib.crc32(name.encode("utf-8"))   # the module-level name shadows the builtin

    class _Cap:   # the short-video branch reads the file's fps for a log line only (:119-123)
        def __init__(self, path):
            pass

        def get(self, prop):
            return 30.0

        def release(self):
            pass

    ref.cv2 = types.SimpleNamespace(VideoCapture=_Cap, CAP_PROP_FPS=5)
    names = ["a.mp4", "b.mp4", "a.mp4", "c.mp4"]
    rows, outs = [], []
    for mi, m in enumerate(METHODS):
        for N in N_GRID:
            for T in T_GRID:
                ds = object.__new__(ref.VideoDataset)
                ds.logger, ds.sampling_method, ds.num_frames, ds.fps, ds.cached_indices = _Log(), m, T, 30, {}
                random.seed(42)
                for call, name in enumerate(names):
                    if call == 3:
                        for _ in range(3):
                            random.random()
                    idx = ds.get_sampling_indices("/data/train/referral/" + name, N)
                    rows.append((mi, N, T, call))
                    a = np.full(32, -1, np.int64)
                    a[: len(idx)] = idx
                    outs.append(a)
                # after any call the global stream is random.seed(42) again
                assert random.random() == random.Random(42).random()
    np.savez_compressed(os.path.join(HERE, "samplers3d.npz"), rows=np.array(rows, np.int64), indices=np.array(outs),
                        names=np.array(names))
    print("samplers3d.npz", len(rows))


if __name__ == "__main__":
    if sys.argv[1:] == ["samplers3d"]:
        make_samplers3d()
        sys.exit(0)
    make_samplers()
    make_samplers3d()
    make_resize()
    make_pixels()
    make_embed()
