"""Frame-index samplers: Python face of the C++ samplers in csrc/lvc_sampler.cpp.

Mirrors the reference's sampler entry points (same names, argument meaning, return types):
  VideoDataset.get_sampling_indices   vivit_transformer/vivit_classifier/data_config/dataset.py:129-193
  sample_frame_indices                data_handling.py:41-56
  random_sampling / uniform_sampling / random_window_sampling
                                      visualize_sampling.py:105-178, resnet50-2d-lstm/inference.py:74-130
The reference draws from process-global RNG streams; here the stream is an explicit ``SamplerRng``
whose sequence is bit-identical to ``random.seed(s)`` / ``np.random.seed(s)``.
"""
import ctypes as C
import math

from . import _lib

METHODS = ("uniform", "random_window", "random")


class SamplerRng:
    """MT19937 stream seeded like CPython ``random.seed(int)`` (``kind='py'``) or numpy's legacy
    ``np.random.seed(int)`` (``kind='np'``)."""

    def __init__(self, seed=42, kind="py"):
        self._st = _lib.Rng()
        self.kind = kind
        self.seed(seed)

    def seed(self, seed):
        seed = int(seed)
        if self.kind == "py":
            _lib.check(_lib.lib().lvc_rng_seed_py(C.byref(self._st), abs(seed)), "SamplerRng.seed")
        elif self.kind == "np":
            if not 0 <= seed < 2 ** 32:
                raise ValueError("Seed must be between 0 and 2**32 - 1")
            _lib.check(_lib.lib().lvc_rng_seed_np(C.byref(self._st), seed), "SamplerRng.seed")
        else:
            raise ValueError(f"unknown RNG kind {self.kind!r}")

    def random(self):
        return _lib.lib().lvc_rng_random(C.byref(self._st))

    def u32(self):
        return _lib.lib().lvc_rng_u32(C.byref(self._st))

    @property
    def ref(self):
        return C.byref(self._st)


def _method_id(method):
    # the reference treats every unknown name as uniform ("else: # Default to uniform sampling")
    return _lib.SAMPLE.get(method, 0)


def get_sampling_indices(method, total_frames, num_frames, rng=None, variant=_lib.VARIANT_DATASET):
    """Indices exactly as ``VideoDataset.get_sampling_indices`` returns them (a Python list)."""
    if total_frames <= 0:
        raise ValueError(f"total_frames must be positive, got {total_frames}")
    out = (C.c_int64 * int(num_frames))()
    n = C.c_int32(0)
    st = rng.ref if rng is not None else None
    _lib.check(_lib.lib().lvc_sample_indices(_method_id(method), variant, int(total_frames), int(num_frames),
                                             st, out, C.byref(n)), "get_sampling_indices")
    return list(out[: n.value])


def dynamic_fps(total_frames, num_frames, original_fps):
    """Second return value of the Swin copy (videoswintransformer/.../dataset.py:104-121)."""
    v = _lib.lib().lvc_dynamic_fps(int(total_frames), int(num_frames), float(original_fps))
    return None if math.isnan(v) else v


def sample_frame_indices(clip_len, frame_sample_rate, seg_len, rng):
    """data_handling.py:41-56 on an explicit numpy-legacy stream; returns a list of ints."""
    out = (C.c_int64 * int(clip_len))()
    _lib.check(_lib.lib().lvc_sample_frame_indices(int(clip_len), float(frame_sample_rate), int(seg_len),
                                                   rng.ref, out), "sample_frame_indices")
    return list(out)


def uniform_temporal_subsample(t, num_samples):
    """pytorchvideo UniformTemporalSubsample index rule."""
    out = (C.c_int64 * int(num_samples))()
    _lib.check(_lib.lib().lvc_uniform_temporal_subsample(int(t), int(num_samples), out),
               "uniform_temporal_subsample")
    return list(out)


# --- the stand-alone function names of visualize_sampling.py / resnet50-2d-lstm/inference.py -----
def uniform_sampling(total_frames, num_frames):
    return get_sampling_indices("uniform", total_frames, num_frames, None, _lib.VARIANT_CLAMPED)


def random_sampling(total_frames, num_frames, rng=None, numpy_stream=False):
    """``numpy_stream=True`` = the inference.py copy (``np.random.seed(42)`` then ``choice``)."""
    if numpy_stream:
        return get_sampling_indices("random", total_frames, num_frames, SamplerRng(42, "np"),
                                    _lib.VARIANT_NUMPY_INFER)
    return get_sampling_indices("random", total_frames, num_frames, rng, _lib.VARIANT_CLAMPED)


def random_window_sampling(total_frames, num_frames, rng=None, numpy_stream=False):
    if numpy_stream:
        return get_sampling_indices("random_window", total_frames, num_frames, SamplerRng(42, "np"),
                                    _lib.VARIANT_NUMPY_INFER)
    return get_sampling_indices("random_window", total_frames, num_frames, rng, _lib.VARIANT_CLAMPED)


class TorchRng:
    """The torch CPU generator stream after ``torch.manual_seed(seed)`` (at::mt19937), emulated in the C
    library: ``randint(lo, hi)`` == ``torch.randint(lo, hi, (1,)).item()``, ``rand()`` == ``torch.rand(1).item()``.
    The explicit, reproducible RNG contract of the train-time augmentation chain."""

    def __init__(self, seed=0):
        self.state = _lib.Rng()
        self.seed(seed)

    def seed(self, seed):
        _lib.check(_lib.lib().lvc_rng_seed_np(C.byref(self.state), int(seed) & 0xFFFFFFFF), "rng_seed")

    def randint(self, low, high):
        o = C.c_int64()
        _lib.check(_lib.lib().lvc_torch_randint(C.byref(self.state), int(low), int(high), C.byref(o)), "torch_randint")
        return int(o.value)

    def rand(self):
        return float(_lib.lib().lvc_torch_rand(C.byref(self.state)))


def train_aug_params(height, width, rng, min_size=256, max_size=320, flip_p=0.5):
    """One clip's draws of RandomShortSideScale(min,max) -> RandomCrop(224) -> RandomHorizontalFlip(p)
    in the reference's order (videoswintransformer/.../dataset.py:156-161) -> [size, top, left, flip]."""
    out = (C.c_int32 * 4)()
    _lib.check(_lib.lib().lvc_train_aug_draw(C.byref(rng.state), int(height), int(width), int(min_size), int(max_size),
                                             float(flip_p), out), "train_aug_draw")
    return [int(v) for v in out]
