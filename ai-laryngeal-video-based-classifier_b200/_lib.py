"""ctypes binding of liblvc_b200.so (the C ABI declared in include/lvc_b200.h).

The library is built in-tree by ``__graft_entry__.build()`` / ``make -C csrc``.  There is no
fallback: a missing library raises ImportError with the build command.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# LVC_B200_LIB: an alternative build of the same library (A/B experiments of compile-time variants; profiling only)
LIB_PATH = os.environ.get("LVC_B200_LIB") or os.path.join(_HERE, "liblvc_b200.so")

OK, E_INVAL, E_SHAPE, E_CUDA, E_ALIGN, E_NOTSUP = 0, -1, -2, -3, -4, -5
SAMPLE = {"uniform": 0, "random_window": 1, "random": 2}
VARIANT_DATASET, VARIANT_CLAMPED, VARIANT_NUMPY_INFER = 0, 1, 2
LUT_DIV, LUT_MUL64 = 0, 1
HF_SRC_HOST = 1
HOST_WRITE_COMBINED = 1
DEVERR_FRAME_TOO_WIDE, DEVERR_FRAME_TOO_LARGE, DEVERR_INTERNAL = 1, 2, 4
B200_SMS, B200_MAX_SMEM = 148, 232448
RESAMPLE_CV2, RESAMPLE_CV2_PIL256, RESAMPLE_PTV_EVAL, RESAMPLE_PIL_SHORT = 0, 1, 2, 3
LAYOUT_BTCHW, LAYOUT_BCTHW, LAYOUT_BTHWC, LAYOUT_TOKENS = 0, 1, 2, 3
DT_F32, DT_BF16, DT_U8 = 0, 1, 2
SRC_RGB, SRC_BGR = 0, 1
EMBED_VIVIT, EMBED_TIMESFORMER, EMBED_SWIN = 0, 1, 2


class Rng(C.Structure):
    _fields_ = [("mt", C.c_uint32 * 624), ("pos", C.c_int32), ("reserved", C.c_int32)]


class Video(C.Structure):
    _fields_ = [("base", C.c_uint64), ("nbytes", C.c_int64), ("n_frames", C.c_int32),
                ("height", C.c_int32), ("width", C.c_int32), ("reserved", C.c_int32)]


class ClipDesc(C.Structure):
    _fields_ = [("resample", C.c_int32), ("layout", C.c_int32), ("out_dtype", C.c_int32),
                ("out_size", C.c_int32), ("batch", C.c_int32), ("num_frames", C.c_int32),
                ("max_width", C.c_int32), ("tok_kt", C.c_int32), ("tok_kh", C.c_int32), ("tok_kw", C.c_int32),
                ("mean", C.c_float * 3), ("std", C.c_float * 3), ("src_order", C.c_int32), ("src_float", C.c_int32)]


class ClipPlan(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("rows_per_stage", "ring_bytes", "ring_offset", "slot_pitch", "stages_of_widest",
                                          "bands", "rows_per_band", "producers", "grid", "threads", "smem_bytes")] + \
               [("reserved", C.c_int32 * 5)]


class LetterboxGeom(C.Structure):
    _fields_ = [("src_h", C.c_int32), ("src_w", C.c_int32), ("new_h", C.c_int32), ("new_w", C.c_int32),
                ("top", C.c_int32), ("left", C.c_int32), ("out_h", C.c_int32), ("out_w", C.c_int32)]


class EmbedDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("x_dtype", C.c_int32), ("x_layout", C.c_int32),
                ("out_dtype", C.c_int32), ("batch", C.c_int32), ("frames", C.c_int32),
                ("height", C.c_int32), ("width", C.c_int32), ("kt", C.c_int32), ("kh", C.c_int32),
                ("kw", C.c_int32), ("embed_dim", C.c_int32), ("ln_eps", C.c_float),
                ("reserved", C.c_int32)]


_SIGS = {
    "lvc_last_error": (C.c_char_p, []),
    "lvc_version": (C.c_int, []),
    "lvc_rng_seed_py": (C.c_int, [C.POINTER(Rng), C.c_uint64]),
    "lvc_rng_seed_np": (C.c_int, [C.POINTER(Rng), C.c_uint32]),
    "lvc_rng_random": (C.c_double, [C.POINTER(Rng)]),
    "lvc_rng_u32": (C.c_uint32, [C.POINTER(Rng)]),
    "lvc_sample_indices": (C.c_int, [C.c_int, C.c_int, C.c_int64, C.c_int32, C.POINTER(Rng),
                                     C.POINTER(C.c_int64), C.POINTER(C.c_int32)]),
    "lvc_sample_frame_indices": (C.c_int, [C.c_int32, C.c_double, C.c_int64, C.POINTER(Rng),
                                           C.POINTER(C.c_int64)]),
    "lvc_uniform_temporal_subsample": (C.c_int, [C.c_int64, C.c_int32, C.POINTER(C.c_int64)]),
    "lvc_dynamic_fps": (C.c_double, [C.c_int64, C.c_int32, C.c_double]),
    "lvc_build_lut": (C.c_int, [C.c_int, C.c_double, C.c_int, C.POINTER(C.c_float),
                                C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "lvc_clip_forward": (C.c_int, [C.POINTER(ClipDesc), C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lvc_clip_launch_count": (C.c_int, [C.POINTER(ClipDesc)]),
    "lvc_clip_forward_aug": (C.c_int, [C.POINTER(ClipDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lvc_hf_clip_forward": (C.c_int, [C.POINTER(ClipDesc), C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p]),
    "lvc_hf_clip_forward_ex": (C.c_int, [C.POINTER(ClipDesc), C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_void_p]),
    "lvc_clip_validate": (C.c_int, [C.POINTER(ClipDesc), C.c_int32, C.c_void_p]),
    "lvc_device_status": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint32), C.c_int32]),
    "lvc_clip_plan_query": (C.c_int, [C.POINTER(ClipDesc), C.c_int32, C.c_int32, C.c_int32, C.POINTER(ClipPlan)]),
    "lvc_torch_randint": (C.c_int, [C.POINTER(Rng), C.c_int64, C.c_int64, C.POINTER(C.c_int64)]),
    "lvc_torch_rand": (C.c_float, [C.POINTER(Rng)]),
    "lvc_train_aug_draw": (C.c_int, [C.POINTER(Rng), C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_double,
                                     C.POINTER(C.c_int32)]),
}
# entry points added by later translation units; bound when the library exports them
_OPTIONAL_SIGS = {
    "lvc_device_bytes_fetched": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64), C.c_int32]),
    "lvc_host_alloc": (C.c_int, [C.c_int64, C.c_int32, C.POINTER(C.c_void_p)]),
    "lvc_host_free": (C.c_int, [C.c_void_p]),
    "lvc_patch_embed_workspace": (C.c_int64, [C.POINTER(EmbedDesc)]),
    "lvc_patch_embed_pack_weights": (C.c_int, [C.POINTER(EmbedDesc), C.c_void_p, C.c_void_p, C.c_void_p]),
    "lvc_patch_embed": (C.c_int, [C.POINTER(EmbedDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lvc_patch_embed_launch_count": (C.c_int, [C.POINTER(EmbedDesc)]),
    "lvc_patch_embed_static_capacity": (C.c_int64, []),
    "lvc_patch_embed_static": (C.c_int, [C.POINTER(EmbedDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_void_p]),
    "lvc_clip_swin_embed_workspace": (C.c_int, [C.POINTER(EmbedDesc), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "lvc_clip_swin_embed_forward": (C.c_int, [C.POINTER(ClipDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                              C.POINTER(EmbedDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                              C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lvc_letterbox_geometry": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                         C.POINTER(LetterboxGeom)]),
    "lvc_letterbox_forward": (C.c_int, [C.POINTER(LetterboxGeom), C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "lvc_scale_boxes": (C.c_int, [C.POINTER(LetterboxGeom), C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "lvc_best_boxes": (C.c_int, [C.POINTER(LetterboxGeom), C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_float,
                                 C.c_void_p, C.c_void_p]),
    "lvc_embeddings_forward": (C.c_int, [C.POINTER(EmbedDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
}

_lib = None


def lib():
    """The loaded library (loads on first use)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                f"or `make -C {os.path.join(_HERE, 'csrc')}` (there is no CPU fallback)")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in {**_SIGS, **_OPTIONAL_SIGS}.items():
            try:
                fn = getattr(l, name)
            except AttributeError:
                if name in _OPTIONAL_SIGS:
                    continue
                raise
            fn.restype, fn.argtypes = res, args
        _lib = l
    return _lib


def declared_symbols():
    """Every entry point include/lvc_b200.h declares."""
    return sorted({**_SIGS, **_OPTIONAL_SIGS})


def check(status, what=""):
    """Map a negative lvc_status to the exception the reference's Python seam would raise."""
    if status == OK:
        return
    msg = lib().lvc_last_error().decode("utf-8", "replace")
    if status in (E_INVAL, E_SHAPE, E_ALIGN):
        raise ValueError(f"{what}: {msg}")
    if status == E_NOTSUP:
        raise NotImplementedError(f"{what}: {msg}")
    raise RuntimeError(f"{what}: {msg}")
