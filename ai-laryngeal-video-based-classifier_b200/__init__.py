"""lvc_b200 -- B200-native clip construction and patch embedding.

Importing the package is cheap and works without a GPU (the samplers are host code); anything
that touches the device raises if the CUDA library or a CUDA device is missing -- there is no
CPU fallback.
"""
__version__ = "0.2.0"

from . import _lib, sampling, synth  # noqa: F401
from .sampling import SamplerRng, get_sampling_indices  # noqa: F401


def __getattr__(name):  # torch-dependent parts load lazily
    if name in ("ClipBuilder", "ClipBatch", "build_lut", "PRESETS"):
        from . import clips

        return getattr(clips, name)
    if name in ("VivitImageProcessor", "TimesformerImageProcessor", "PtvEvalTransform", "PtvTrainTransform"):
        from . import processors

        return getattr(processors, name)
    if name in ("clips", "datasets", "dataloader", "processors", "patch_embed", "decode", "sharding", "hostmem", "pipeline", "yolo"):
        import importlib

        return importlib.import_module(f"{__name__}.{name}")
    raise AttributeError(name)
