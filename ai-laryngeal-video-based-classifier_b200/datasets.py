"""Drop-in Dataset layer: the reference's ``VideoDataset`` classes with the clip construction
moved onto the GPU.

Constructor signatures, attributes (``video_paths``, ``labels``, ``class_labels``), method names and
return structures follow the reference copies:

  VivitVideoDataset / TimesformerVideoDataset
      vivit_transformer/vivit_classifier/data_config/dataset.py:10-381 (timesformer copy :25-394)
      __getitem__ -> {'pixel_values': uint8 (T,224,224,3), 'labels', 'video_path', 'frame_indices'}
  LstmVideoDataset   resnet50-2d-lstm/src/data_config/dataset.py:19-219
      __getitem__ -> (float (C,T,224,224), float label); re-seeds random.seed(42) in every sampler call
  SwinVideoDataset   videoswintransformer/swin_video_classifier/data_config/dataset.py:18-235
      (resnet50-3d-video copy :22-289)   __getitem__ -> ((1,C,T,224,224), LongTensor (1,))

Differences that are the point of this package: decoded frames come from a ``FrameSource`` (decode
itself is outside the accelerated path); tensors stay on the device;
``get_batch(indices)`` builds a whole batch with ONE fused kernel launch and is what the loaders in
``dataloader.py`` use (the GPU path runs in the main process, ``num_workers=0``).
"""
import os

import numpy as np
import torch

from . import _lib, sampling
from .clips import KINETICS, ClipBuilder

CLASS_LABELS = ["non_referral", "referral"]


class FrameSource:
    """Decoded RGB frames of a set of videos.  ``frames(i)`` returns a CUDA uint8 (N,H,W,3) tensor."""

    video_paths = ()
    labels = ()
    channel_order = "RGB"  # "BGR": frames as cv2 decodes them; the clip kernel swaps channels in its gather

    def frames(self, i):
        raise NotImplementedError

    def fps(self, i):
        return 30.0

    def num_frames(self, i):
        return int(self.frames(i).shape[0])


class TensorFrameSource(FrameSource):
    """Videos handed over as tensors / arrays (uploaded once, kept resident in HBM)."""

    def __init__(self, videos, labels=None, paths=None, fps=30.0, device="cuda"):
        self._v = [(torch.from_numpy(v) if isinstance(v, np.ndarray) else v).to(device).contiguous() for v in videos]
        self.labels = list(labels) if labels is not None else [0] * len(self._v)
        self.video_paths = list(paths) if paths is not None else [f"video_{i:04d}.mp4" for i in range(len(self._v))]
        self._fps = fps

    def frames(self, i):
        return self._v[i]

    def fps(self, i):
        return self._fps[i] if isinstance(self._fps, (list, tuple)) else self._fps


class SyntheticFrameSource(TensorFrameSource):
    """Seeded synthetic videos: video v = ``synth_video(1000 + v, ...)``."""

    def __init__(self, n_videos, n_frames, height, width, mode="noise", device="cuda", first=0):
        from .synth import synth_video

        vids = [synth_video(1000 + first + i, n_frames, height, width, mode) for i in range(n_videos)]
        super().__init__(vids, labels=[(first + i) % 2 for i in range(n_videos)], device=device)


def scan_split(root_dir, split):
    """Paths and labels the reference's ``_load_dataset`` collects: ``root/split/<class>/*.mp4``
    (vivit .../dataset.py:84-100), referral = 1, non_referral = 0, sorted for determinism."""
    d = os.path.join(str(root_dir), split)
    if not os.path.isdir(d):
        raise ValueError(f"Directory not found: {d}")
    paths, labels = [], []
    for li, cls in enumerate(CLASS_LABELS):
        cd = os.path.join(d, cls)
        if os.path.isdir(cd):
            for f in sorted(os.listdir(cd)):
                if f.lower().endswith(".mp4"):
                    paths.append(os.path.join(cd, f))
                    labels.append(li)
    if not paths:
        raise ValueError(f"No videos found in {d}")
    return paths, labels


class _BaseDataset(torch.utils.data.Dataset):
    reseed_each_call = False

    def _setup(self, root_dir, split, sampling_method, num_frames, logger, frame_source, preset, **builder_kw):
        self.root_dir, self.sampling_method, self.num_frames, self.logger = root_dir, sampling_method, int(num_frames), logger
        self.class_labels = list(CLASS_LABELS)
        if frame_source is None:
            from .decode import OpenCVFrameSource  # host decode of real files (row f1)

            paths, labels = scan_split(root_dir, split)
            frame_source = OpenCVFrameSource(paths, labels)
        self.source = frame_source
        self.video_paths = list(frame_source.video_paths)
        self.labels = list(frame_source.labels)
        # the reference seeds the global streams once in __init__ (vivit .../dataset.py:39-40)
        self.rng = sampling.SamplerRng(42, "py")
        builder_kw.setdefault("src_order", getattr(frame_source, "channel_order", "RGB"))
        self.builder = ClipBuilder(preset, num_frames=self.num_frames, **builder_kw)

    def __len__(self):
        return len(self.video_paths)

    def _indices(self, total_frames):
        if self.reseed_each_call:  # resnet50-2d-lstm/src/data_config/dataset.py:97
            self.rng.seed(42)
        return sampling.get_sampling_indices(self.sampling_method, total_frames, self.num_frames, self.rng)

    def get_sampling_indices(self, video_path, total_frames):
        return self._indices(total_frames)

    # -- frames of a batch: resident videos, or (index-directed sources) only the sampled frames, decoded into pinned
    #    host memory that the kernel reads in place -- the reference decodes just the sampled indices too
    #    (vivit .../dataset.py:316-341)
    def _build(self, indices, pick, **kw):
        """``pick(i, total_frames)`` -> (frame indices the clip uses, what to report as frame_indices).  Returns
        (clips, reported indices, list of per-video frame arrays' shapes)."""
        src = self.source
        if getattr(src, "index_directed", False):
            frames, remaps, reported, shapes = [], [], [], []
            for i in indices:
                use, rep = pick(i, src.num_frames(i))
                fr, rm = src.read_frames(i, use)
                frames.append(fr)
                remaps.append(rm)
                reported.append(rep)
                shapes.append(tuple(fr.shape[1:3]))
            clips = self.builder.build_from_host(frames, np.stack(remaps), **kw)
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.builder.device))
            src.fence(ev)   # the pinned ring reuses these frames' space only after the kernel has read them
            return clips, reported, shapes
        vids = [src.frames(i) for i in indices]
        picks = [pick(i, v.shape[0]) for i, v in zip(indices, vids)]
        clips = self.builder(vids, [u for u, _ in picks], **kw)
        return clips, [r for _, r in picks], [tuple(v.shape[1:3]) for v in vids]


class VivitVideoDataset(_BaseDataset):
    """``VideoDataset(root_dir, mode='train', sampling_method='uniform', num_frames=32, logger=None)``."""

    def __init__(self, root_dir=None, mode="train", sampling_method="uniform", num_frames=32, logger=None,
                 frame_source=None, float_frames=False):
        # float_frames=True reproduces the reference's PRIMARY branch (pytorchvideo decode -> float32 frames ->
        # cv2.resize in float -> astype(uint8), vivit .../dataset.py:244-284); False its OpenCV fallback
        # (uint8 frames, :317-350), which is also what the cv2-based ingest of this package decodes
        self.mode = mode
        self._setup(root_dir, mode, sampling_method, num_frames, logger, frame_source, "dataset_u8",
                    float_frames=float_frames)

    def get_batch(self, indices):
        def pick(i, total):
            fi = self.get_sampling_indices(self.video_paths[i], total)
            return fi, fi

        clips, fidx, _ = self._build(indices, pick)  # (B, T, 224, 224, 3) uint8, one launch
        return {"pixel_values": list(clips), "labels": torch.tensor([self.labels[i] for i in indices], dtype=torch.long),
                "video_paths": [str(self.video_paths[i]) for i in indices], "frame_indices": fidx}

    def __getitem__(self, idx):
        b = self.get_batch([idx])
        return {"pixel_values": b["pixel_values"][0], "labels": torch.tensor(self.labels[idx]),
                "video_path": b["video_paths"][0], "frame_indices": b["frame_indices"][0]}


class TimesformerVideoDataset(VivitVideoDataset):
    """Same as the ViViT copy; accepts ``mode='inference'`` (timesformer .../dataset.py:25-28)."""


def window_subsample(indices, num_frames):
    """Frames the pytorchvideo datasets really use: the custom indices only pick the decoded window
    [idx[0], idx[-1]] (videoswintransformer/.../dataset.py:213-218) and UniformTemporalSubsample(T)
    then picks inside it (:157) -- SURVEY.md 3.3."""
    lo, hi = min(indices), max(indices)
    return [lo + k for k in sampling.uniform_temporal_subsample(hi - lo + 1, num_frames)]


class SwinVideoDataset(_BaseDataset):
    """``VideoDataset(root_dir, mode='train', sampling_method='uniform', num_frames=32, fps=30, stride=0.5,
    logger=None, log_dir=None)`` -> ``((1, C, T, 224, 224) float, LongTensor (1,))`` with the eval
    transform chain (ShortSideScale 256, CenterCrop 224, Normalize 0.45/0.225 on 0..255 values)."""

    def __init__(self, root_dir=None, mode="train", sampling_method="uniform", num_frames=32, fps=30, stride=0.5,
                 logger=None, log_dir=None, frame_source=None, augment=None, aug_seed=0):
        self.mode, self.fps, self.stride = mode, fps, stride
        self._setup(root_dir, mode, sampling_method, num_frames, logger, frame_source, "ptv_eval")
        # mode == 'train' uses RandomShortSideScale(256,320) + RandomCrop(224) + RandomHorizontalFlip(0.5)
        # (videoswintransformer/.../dataset.py:153-162); the draws come from an explicit torch-compatible
        # stream (sampling.TorchRng(aug_seed) == torch.manual_seed(aug_seed)), one draw per clip in item order
        self.augment = (mode == "train") if augment is None else bool(augment)
        self.aug_rng = sampling.TorchRng(aug_seed)
        self.last_aug = None

    def get_sampling_indices(self, video_path, total_frames):
        i = self.video_paths.index(video_path) if video_path in self.video_paths else 0
        return self._indices(total_frames), sampling.dynamic_fps(total_frames, self.num_frames, self.source.fps(i))

    def _frame_size(self, i):
        src = self.source
        if getattr(src, "index_directed", False):
            _, _, w, h = src.properties(i)
            return h, w
        return tuple(src.frames(i).shape[1:3])

    def _ptv_batch(self, indices, sample):
        """Shared by the pytorchvideo-style copies: ``sample(i, total)`` -> raw sampler indices; the clip uses
        UniformTemporalSubsample inside the decoded window (``window_subsample``)."""
        aug = None
        if self.augment:   # drawn per clip in item order, before the frames are touched (needs only the frame size)
            aug = [sampling.train_aug_params(*self._frame_size(i), self.aug_rng) for i in indices]
        self.last_aug = aug

        def pick(i, total):
            use = window_subsample(sample(i, total), self.num_frames)
            return use, use

        clips, _, _ = self._build(indices, pick, **({"aug": aug} if aug is not None else {}))
        return clips

    def get_batch(self, indices):
        clips = self._ptv_batch(indices, lambda i, total: self._indices(total))  # (B, C, T, 224, 224)
        return clips.unsqueeze(1), torch.tensor([[self.labels[i]] for i in indices], dtype=torch.long)

    def __getitem__(self, idx):
        clips, labels = self.get_batch([idx])
        return clips[0], labels[0]


def crc32_video_seed(basename):
    """Stable stand-in for the reference's ``hash(os.path.basename(video_path)) % 10000000``: CPython salts ``hash``
    of str per process (PYTHONHASHSEED), so the reference's per-video seed is only reproducible inside one process.
    Pass ``seed_fn=lambda n: hash(n) % 10000000`` to get the reference's own values under a fixed PYTHONHASHSEED."""
    import zlib

    return zlib.crc32(basename.encode("utf-8")) % 10000000


class Resnet3dVideoDataset(SwinVideoDataset):
    """The ``resnet50-3d-video`` copy (resnet50-3d-video/video_classifier/data_config/dataset.py:22-289): same
    constructor and transforms as the Swin copy, but its sampler (:79-169) returns a plain list and
      * seeds the stream PER VIDEO with ``hash(basename) % 10000000`` (:95-98) -> ``seed_fn(basename)`` here,
      * caches the indices per video path (``cached_indices``, :91-93, :166-167): a second epoch reuses them,
      * re-seeds the global stream to 42 after every draw (:163-164),
      * can dump the cache as CSV (``save_sampled_indices``, :241-289)."""

    def __init__(self, root_dir=None, mode="train", sampling_method="uniform", num_frames=32, fps=30, stride=0.5,
                 logger=None, log_dir=None, frame_source=None, augment=None, aug_seed=0, seed_fn=None):
        super().__init__(root_dir, mode, sampling_method, num_frames, fps, stride, logger, log_dir, frame_source,
                         augment, aug_seed)
        self.log_dir = log_dir
        self.seed_fn = seed_fn or crc32_video_seed
        self.cached_indices = {}

    def set_random_seed(self, seed):
        self.rng.seed(seed)

    def get_sampling_indices(self, video_path, total_frames):
        video_path = str(video_path)
        if video_path in self.cached_indices:
            return self.cached_indices[video_path]
        self.rng.seed(int(self.seed_fn(os.path.basename(video_path))))
        indices = sampling.get_sampling_indices(self.sampling_method, total_frames, self.num_frames, self.rng)
        self.set_random_seed(42)
        self.cached_indices[video_path] = indices
        return indices

    def get_batch(self, indices):
        clips = self._ptv_batch(indices, lambda i, total: self.get_sampling_indices(self.video_paths[i], total))
        return clips.unsqueeze(1), torch.tensor([[self.labels[i]] for i in indices], dtype=torch.long)

    def save_sampled_indices(self):
        """CSV ``video_filename,total_frames,sampled_frames`` of every video's cached indices (:241-289); videos not
        sampled yet are sampled now.  Returns the path, or None without a ``log_dir`` (the reference only warns)."""
        import csv

        if not self.log_dir:
            return None
        totals = {}
        for i, vp in enumerate(self.video_paths):
            totals[str(vp)] = self.source.num_frames(i)
            self.get_sampling_indices(vp, totals[str(vp)])
        path = os.path.join(str(self.log_dir), f"sampled_frames_{self.mode}_{self.sampling_method}.csv")
        with open(path, "w", newline="") as f:
            w = csv.writer(f)
            w.writerow(["video_filename", "total_frames", "sampled_frames"])
            for vp in sorted(str(v) for v in self.video_paths):
                w.writerow([os.path.basename(vp), totals[vp], ",".join(map(str, self.cached_indices[vp]))])
        return path


class LstmVideoDataset(_BaseDataset):
    """``VideoDataset(root_dir, split='train', sampling_method='uniform', sequence_length=32, transform=None,
    logger=None)`` -> ``((C, T, 224, 224) float, float label)``; the sampler re-seeds 42 on every call."""

    reseed_each_call = True

    def __init__(self, root_dir=None, split="train", sampling_method="uniform", sequence_length=32, transform=None,
                 logger=None, frame_source=None, augment=None, aug_seed=0):
        if transform is not None:
            raise NotImplementedError("custom transforms are not supported; the eval chain is fused in the kernel")
        self.split, self.sequence_length = split, sequence_length
        self._setup(root_dir, split, sampling_method, sequence_length, logger, frame_source, "ptv_eval")
        # split == 'train': the same random chain as the Swin dataset (resnet50-2d-lstm/.../dataset.py:63-72)
        self.augment = (split == "train") if augment is None else bool(augment)
        self.aug_rng = sampling.TorchRng(aug_seed)
        self.last_aug = None

    _frame_size = SwinVideoDataset._frame_size
    _ptv_batch = SwinVideoDataset._ptv_batch

    def get_batch(self, indices):
        clips = self._ptv_batch(indices, lambda i, total: self._indices(total))
        return clips, torch.tensor([self.labels[i] for i in indices], dtype=torch.float32)

    def __getitem__(self, idx):
        clips, labels = self.get_batch([idx])
        return clips[0], labels[0]


VideoDataset = VivitVideoDataset  # the name every reference package uses
__all__ = ["FrameSource", "TensorFrameSource", "SyntheticFrameSource", "VideoDataset", "VivitVideoDataset",
           "TimesformerVideoDataset", "SwinVideoDataset", "LstmVideoDataset", "Resnet3dVideoDataset", "crc32_video_seed", "window_subsample", "scan_split",
           "KINETICS", "_lib"]
