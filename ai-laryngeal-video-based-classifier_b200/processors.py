"""Drop-in image processors / transforms: the callables the reference's trainers invoke in
``_process_batch`` (on the CPU, per frame), re-backed by the fused CUDA clip kernel.

  VivitImageProcessor            vivit .../trainers/trainer.py:22-26,62-95; evaluator.py:25,64-67;
                                 vivit_transformer/inference.py:181-187; data_preprocessing.py:7,16
  TimesformerImageProcessor      timesformer .../trainers/trainer.py:22-24,91-98 (a VideoMAE-class
                                 processor forced to size = crop_size = 224)
  PtvEvalTransform               the pytorchvideo eval chain applied as ``transform({"video": ...})``
                                 (resnet50-2d-lstm/src/data_config/dataset.py:63-83,204-208)

Call signature and result follow the HF processors: ``proc(images, return_tensors="pt", **kw)`` returns
a mapping with ``"pixel_values"``: float32 ``(B, T, 3, 224, 224)`` (B = 1 for one video), supporting
``.keys()``, ``[k]``, ``.items()`` -- but the tensor is already on the GPU, so the caller's ``.to(device)``
is a no-op.  Frames must be uint8 RGB.  224x224 frames (what the reference's Datasets hand over) take the
fused clip kernel's presets; frames of any other size -- the stand-alone scripts data_preprocessing.py:16,
data_preprocessing_timesformer.py:16 pass decoded frames as they are -- take the processor's own geometry:
Pillow BILINEAR resize of the shortest edge (256 ViViT / 224 VideoMAE-class), centre crop 224
(``lvc_hf_clip_forward``), bit-exact against transformers.
"""
import numpy as np
import torch

from .clips import KINETICS, ClipBuilder


class BatchFeature(dict):
    """Minimal stand-in for transformers.BatchFeature: a dict with ``.to()``."""

    def to(self, *a, **k):
        return BatchFeature({key: v.to(*a, **k) for key, v in self.items()})


def _as_videos(images, device):
    """Normalise the accepted inputs to a list of CUDA uint8 (T,H,W,3) tensors."""
    def one(v):
        t = torch.from_numpy(np.ascontiguousarray(v)) if isinstance(v, np.ndarray) else v
        if t.dtype != torch.uint8:
            raise ValueError("frames must be uint8 RGB (the reference casts to uint8 before the processor)")
        return t.to(device, non_blocking=True).contiguous()

    if isinstance(images, (np.ndarray, torch.Tensor)):
        if images.ndim == 3:
            return [one(images)[None]]
        if images.ndim == 4:
            return [one(images)]
        if images.ndim == 5:
            return [one(v) for v in images]
        raise ValueError(f"unsupported input shape {tuple(images.shape)}")
    images = list(images)
    if not images:
        raise ValueError("no frames given")
    if images[0].ndim == 3:  # list of HWC frames = one video
        if isinstance(images[0], np.ndarray):
            return [one(np.stack(images))]
        return [one(torch.stack(list(images)))]
    return [one(v) for v in images]  # list of (T,H,W,3) videos


class _ClipProcessor:
    preset = None
    raw_preset = None   # used when the frames are not already 224x224
    shortest_edge = None

    def __init__(self, image_mean=None, image_std=None, device=None, **ignored):
        # the reference passes extra kwargs (num_frames, image_size, patch_size) the HF class ignores
        self.image_mean, self.image_std = image_mean, image_std
        # settable like the HF attributes (timesformer .../trainers/trainer.py:23-24 assigns height/width = 224)
        self.size = {"shortest_edge": self.shortest_edge}
        self.crop_size = {"height": 224, "width": 224}
        self.device = device
        self._builders = {}

    def _builder(self, T, raw=False, exact=False):
        key = (T, raw, exact)
        if key not in self._builders:
            self._builders[key] = ClipBuilder(self.raw_preset if raw else self.preset, num_frames=T,
                                              mean=self.image_mean, std=self.image_std, device=self.device,
                                              shortest_edge=(0 if exact else self.shortest_edge) if raw else None)
        return self._builders[key]

    def _check(self, kw):
        for key in ("size", "crop_size"):
            v = kw.get(key) or getattr(self, key)
            if v.get("height", 224) != 224 or v.get("width", 224) != 224 or \
                    v.get("shortest_edge", self.shortest_edge) != self.shortest_edge:
                raise NotImplementedError(f"{key}={v}: only 224x224 or shortest_edge={self.shortest_edge} is supported")
        for key in ("do_resize", "do_center_crop", "do_rescale", "do_normalize"):
            if kw.get(key) is False:
                raise NotImplementedError(f"{key}=False: the fused kernel always runs the whole chain")

    def __call__(self, images=None, return_tensors="pt", videos=None, **kw):
        if images is None:
            images = videos
        if return_tensors not in ("pt", None):
            raise ValueError("only return_tensors='pt' is supported")
        self._check(kw)
        dev = torch.device(self.device if self.device is not None else f"cuda:{torch.cuda.current_device()}")
        vids = _as_videos(images, dev)
        T = vids[0].shape[0]
        if any(v.shape[0] != T for v in vids):
            raise ValueError("all videos of a batch must have the same number of frames")
        idx = np.tile(np.arange(T, dtype=np.int32), (len(vids), 1))
        raw = any(tuple(v.shape[1:3]) != (224, 224) for v in vids)
        # size={"height": 224, "width": 224} (given per call, or set on the instance as timesformer
        # .../trainers/trainer.py:23-24 does) on other sizes: Pillow resizes both axes to 224, aspect not kept
        exact = raw and "height" in (kw.get("size") or self.size)
        return BatchFeature(pixel_values=self._builder(T, raw, exact)(vids, idx))

    preprocess = __call__


class VivitImageProcessor(_ClipProcessor):
    """VivitImageProcessor defaults: shortest_edge 256 (Pillow bilinear), centre crop 224,
    ``x/127.5 - 1``, mean/std 0.5 -> values in [-3, 1]."""

    preset = "vivit"
    raw_preset = "vivit_raw"
    shortest_edge = 256



class TimesformerImageProcessor(_ClipProcessor):
    """VideoMAE-class processor at size = crop_size = 224: ``x/255`` then ``(x - mean)/std``.  The
    checkpoint's mean/std are not available offline: pass them, default 0.45/0.225."""

    preset = "timesformer"
    raw_preset = "timesformer_raw"   # the checkpoint's preprocessor_config: shortest_edge 224, crop 224
    shortest_edge = 224

    def __init__(self, image_mean=None, image_std=None, **kw):
        super().__init__(image_mean if image_mean is not None else KINETICS[0],
                         image_std if image_std is not None else KINETICS[1], **kw)


class PtvEvalTransform:
    """``transform({"video": float (C,T,H,W) 0..255, ...})`` -> same dict with "video" replaced by the
    (C,T,224,224) result of UniformTemporalSubsample(T) -> ShortSideScale(256) -> CenterCrop(224) ->
    Normalize(mean, std).  A uint8 (T,H,W,3) tensor under "video" skips the float round trip."""

    def __init__(self, num_frames, mean=KINETICS[0], std=KINETICS[1], device=None):
        self.num_frames = num_frames
        self.builder = ClipBuilder("ptv_eval", num_frames=num_frames, mean=mean, std=std, device=device)

    def __call__(self, sample):
        from .sampling import uniform_temporal_subsample

        v = sample["video"]
        if v.dtype != torch.uint8:  # (C,T,H,W) float holding decoded integers
            v = v.permute(1, 2, 3, 0).round().clamp(0, 255).to(torch.uint8)
        v = v.to(self.builder.device).contiguous()
        idx = uniform_temporal_subsample(v.shape[0], self.num_frames)
        out = dict(sample)
        out["video"] = self.builder([v], [idx])[0]
        return out


class PtvTrainTransform(PtvEvalTransform):
    """Train-mode chain of the pytorchvideo datasets (videoswintransformer/.../dataset.py:153-162):
    UniformTemporalSubsample(T) -> RandomShortSideScale(256, 320) -> RandomCrop(224) ->
    RandomHorizontalFlip(0.5) -> Normalize, one draw per clip.  The draws come from ``rng`` (a
    ``sampling.TorchRng``: the stream ``torch.manual_seed(seed)`` gives the reference), so a seeded run is
    reproducible and comparable with the reference; geometry, flip and normalisation run in the fused kernel."""

    def __init__(self, num_frames, mean=KINETICS[0], std=KINETICS[1], device=None, rng=None, min_size=256,
                 max_size=320, flip_p=0.5):
        super().__init__(num_frames, mean, std, device)
        from .sampling import TorchRng

        self.rng = rng if rng is not None else TorchRng(0)
        self.min_size, self.max_size, self.flip_p = min_size, max_size, flip_p
        self.last_params = None

    def __call__(self, sample):
        from .sampling import train_aug_params, uniform_temporal_subsample

        v = sample["video"]
        if v.dtype != torch.uint8:
            v = v.permute(1, 2, 3, 0).round().clamp(0, 255).to(torch.uint8)
        v = v.to(self.builder.device).contiguous()
        idx = uniform_temporal_subsample(v.shape[0], self.num_frames)
        self.last_params = train_aug_params(v.shape[1], v.shape[2], self.rng, self.min_size, self.max_size, self.flip_p)
        out = dict(sample)
        out["video"] = self.builder([v], [idx], aug=[self.last_params])[0]
        return out
