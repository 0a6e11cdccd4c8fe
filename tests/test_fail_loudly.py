"""Contract violations are reported, not swallowed and not fatal: host-side validation of the video table
(lvc_clip_validate), the device error word (lvc_device_status), gated profiling knobs, forward-only modules under
autograd."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _desc(max_width=1280, batch=1, T=4, resample=0, layout=0, dtype=2):
    from lvc_b200 import _lib

    return _lib.ClipDesc(resample, layout, dtype, 224, batch, T, max_width, 0, 0, 0, (C.c_float * 3)(0, 0, 0),
                         (C.c_float * 3)(1, 1, 1), 0, 0)


def test_validate_rejects_bad_video_tables_on_the_host():
    from lvc_b200 import _lib

    lib = _lib.lib()
    d = _desc(max_width=640)
    vt = (_lib.Video * 1)()
    vt[0].base, vt[0].nbytes, vt[0].n_frames, vt[0].height, vt[0].width = 4096, 3 * 480 * 640 * 3, 3, 480, 640
    assert lib.lvc_clip_validate(C.byref(d), 0, vt) == _lib.OK
    vt[0].width = 641                                       # wider than desc.max_width
    vt[0].nbytes = 3 * 480 * 641 * 3
    assert lib.lvc_clip_validate(C.byref(d), 0, vt) == _lib.E_SHAPE
    assert b"max_width" in lib.lvc_last_error()
    vt[0].width, vt[0].n_frames = 640, 0                    # no frames: the reference raises ValueError (dataset.py:217)
    assert lib.lvc_clip_validate(C.byref(d), 0, vt) == _lib.E_INVAL
    vt[0].n_frames, vt[0].base = 3, 4097                    # misaligned
    assert lib.lvc_clip_validate(C.byref(d), 0, vt) == _lib.E_ALIGN
    vt[0].base, vt[0].nbytes = 4096, 100                    # buffer smaller than the frames it claims
    assert lib.lvc_clip_validate(C.byref(d), 0, vt) == _lib.E_INVAL
    vt[0].nbytes = 3 * 480 * 640 * 3
    assert lib.lvc_clip_validate(C.byref(d), 479, vt) == _lib.E_SHAPE   # lvc_hf_clip_forward's max_height
    assert lib.lvc_clip_validate(C.byref(d), 480, vt) == _lib.OK
    assert lib.lvc_clip_validate(None, 0, vt) == _lib.E_INVAL


def test_plan_vivit_large_batches_keep_a_ring():
    """from ~1024 frames per call the ViViT preset used to pick whole-frame items, whose Pillow
    intermediate left a 1080p source no room for the ring (LVC_E_SHAPE) and a 720p source two starved stages."""
    from lvc_b200 import _lib
    from lvc_b200.clips import clip_plan

    for width, frames in [(1920, 256), (1920, 1024), (1920, 4096), (1280, 1024), (1280, 8192), (3840, 2048)]:
        d = _desc(max_width=width, batch=frames // 32, T=32, resample=_lib.RESAMPLE_CV2_PIL256, dtype=_lib.DT_F32)
        plan = clip_plan(d, False, _lib.B200_SMS, _lib.B200_MAX_SMEM)
        assert plan["stages_of_widest"] >= 2, (width, frames, plan)
        assert plan["rows_per_stage"] >= 4 or width > 1920, (width, frames, plan)
        assert plan["smem_bytes"] <= _lib.B200_MAX_SMEM
    # small batches still slice frames into bands (quantisation), large 720p batches use few bands
    d = _desc(max_width=1280, batch=8, T=32, resample=_lib.RESAMPLE_CV2_PIL256, dtype=_lib.DT_F32)
    assert clip_plan(d, False, _lib.B200_SMS, _lib.B200_MAX_SMEM)["bands"] == 4


def test_plan_is_byte_granular_and_bounded():
    from lvc_b200 import _lib
    from lvc_b200.clips import clip_plan

    for width in (64, 320, 640, 1280, 1920, 3840, 7000):
        for boxes in (False, True):
            p = clip_plan(_desc(max_width=width, batch=8, T=32), boxes, _lib.B200_SMS, _lib.B200_MAX_SMEM)
            assert p["ring_offset"] + p["ring_bytes"] == p["smem_bytes"] <= _lib.B200_MAX_SMEM
            assert p["ring_bytes"] >= 2 * 2 * p["rows_per_stage"] * p["slot_pitch"], (width, p)
            assert 224 % p["rows_per_band"] == 0 and p["rows_per_band"] % p["rows_per_stage"] == 0
            assert 1 <= p["grid"] <= 2 * _lib.B200_SMS
    with pytest.raises(ValueError):
        clip_plan(_desc(max_width=40000), False, _lib.B200_SMS, _lib.B200_MAX_SMEM)


def test_plan_keeps_eight_row_stages_for_swin_token_batches():
    """The whole-sector token stores of the Video Swin layout (4 x 4 patches: one patch row per consumer group and stage)
    need 8-row stages: ROI batches get them up to 1080p sources (two stages of the widest rows must fit the ring), and
    the bands are whole patch rows."""
    from lvc_b200 import _lib
    from lvc_b200.clips import clip_plan

    for width in (320, 640, 1280, 1920):
        d = _lib.ClipDesc(_lib.RESAMPLE_CV2, _lib.LAYOUT_TOKENS, _lib.DT_BF16, 224, 8, 32, width, 2, 4, 4,
                          (C.c_float * 3)(0, 0, 0), (C.c_float * 3)(1, 1, 1), 0, 0)
        p = clip_plan(d, True, _lib.B200_SMS, _lib.B200_MAX_SMEM)
        assert p["rows_per_stage"] == 8 and p["rows_per_band"] % 4 == 0 and p["stages_of_widest"] >= 2, (width, p)
    # full-frame 720p batches (no boxes) prefer the finely sliced ring: 4-row stages (HBM-bound there, the stores do not matter)
    d = _lib.ClipDesc(_lib.RESAMPLE_CV2, _lib.LAYOUT_TOKENS, _lib.DT_BF16, 224, 8, 32, 1280, 2, 4, 4,
                      (C.c_float * 3)(0, 0, 0), (C.c_float * 3)(1, 1, 1), 0, 0)
    assert clip_plan(d, False, _lib.B200_SMS, _lib.B200_MAX_SMEM)["rows_per_stage"] == 4


_KNOB_PROBE = """
import ctypes as C, json, sys
sys.path.insert(0, {root!r})
from lvc_b200 import _lib
from lvc_b200.clips import clip_plan
d = _lib.ClipDesc(0, 0, 2, 224, 8, 32, 1280, 0, 0, 0, (C.c_float * 3)(0, 0, 0), (C.c_float * 3)(1, 1, 1), 0, 0)
print(json.dumps(clip_plan(d, False, 148, 232448)))
"""


def _plan_in_subprocess(env):
    e = {k: v for k, v in os.environ.items() if not k.startswith("LVC_")}
    e.update(env)
    out = subprocess.run([sys.executable, "-c", _KNOB_PROBE.format(root=ROOT)], env=e, check=True, capture_output=True,
                         text=True).stdout
    import json

    return json.loads(out.strip().splitlines()[-1])


def test_profiling_knobs_are_ignored_unless_gated():
    base = _plan_in_subprocess({})
    stray = _plan_in_subprocess({"LVC_CLIP_RPS": "2", "LVC_CLIP_NP": "1", "LVC_CLIP_BAND": "2", "LVC_CLIP_SMEM_KB": "100"})
    assert stray == base                                           # a stray variable changes nothing
    gated = _plan_in_subprocess({"LVC_PROFILING_KNOBS": "1", "LVC_CLIP_RPS": "2", "LVC_CLIP_NP": "1"})
    assert gated["rows_per_stage"] == 2 and gated["producers"] == 1 and gated != base
    assert _plan_in_subprocess({"LVC_PROFILING_KNOBS": "yes", "LVC_CLIP_RPS": "2"}) == base   # the gate is exactly "1"


def test_no_getenv_outside_the_gate():
    """Every environment read of the native library goes through lvc_knob_* (csrc/lvc_host.cpp)."""
    csrc = os.path.join(ROOT, "ai-laryngeal-video-based-classifier_b200", "csrc")
    for fn in os.listdir(csrc):
        if fn.endswith((".cu", ".cpp", ".h")) and fn != "lvc_host.cpp":
            assert "getenv(" not in open(os.path.join(csrc, fn)).read(), fn


@pytest.mark.gpu
def test_too_wide_video_is_a_value_error_not_a_trap():
    import torch

    from lvc_b200 import ClipBuilder, _lib
    from lvc_b200.clips import device_status
    from oracle import processors as op

    cb = ClipBuilder("dataset_u8", num_frames=2)
    g = torch.Generator(device="cuda").manual_seed(3)
    small = torch.randint(0, 256, (2, 64, 96, 3), dtype=torch.uint8, device="cuda", generator=g)
    huge = torch.randint(0, 256, (2, 8, 30000, 3), dtype=torch.uint8, device="cuda", generator=g)
    batch = cb.prepare([small, small], [[0, 1], [1, 0]])
    # a caller that lies about max_width on the raw C ABI: host validation says so ...
    vt = (_lib.Video * 2).from_buffer(batch.host.numpy())
    vt[1].base, vt[1].nbytes, vt[1].height, vt[1].width = huge.data_ptr(), huge.numel(), 8, 30000
    assert _lib.lib().lvc_clip_validate(C.byref(batch.desc), 0, vt) == _lib.E_SHAPE
    # ... and a launch that carries it anyway neither faults nor kills the context: the frames of the offending video
    # stay unwritten, the others are right, the device error word says what happened
    batch.tables.copy_(batch.host)
    out = torch.full(batch.shape, 7, dtype=torch.uint8, device="cuda")
    cb.run(batch, out)
    assert device_status() == {"frame_too_wide"}
    assert device_status() == set()                                    # cleared by the read
    assert np.array_equal(out[0].cpu().numpy(), op.resize_clip_cv2(small.cpu().numpy()[[0, 1]]))
    assert bool((out[1] == 7).all())
    # the Python face never gets that far
    with pytest.raises(ValueError):
        bad = cb.prepare([small, small], [[0, 1], [1, 0]])
        bad.desc.max_width = 32
        _lib.check(_lib.lib().lvc_clip_validate(C.byref(bad.desc), 0, bad.host.data_ptr()), "validate")
    again = cb([small], [[1, 1]])                                      # context still healthy
    assert np.array_equal(again[0].cpu().numpy(), op.resize_clip_cv2(small.cpu().numpy()[[1, 1]]))


@pytest.mark.gpu
def test_raw_preset_too_large_video_is_reported():
    import torch

    from lvc_b200 import ClipBuilder, _lib
    from lvc_b200.clips import device_status

    cb = ClipBuilder("vivit_raw", num_frames=2)
    small = torch.randint(0, 256, (2, 240, 320, 3), dtype=torch.uint8, device="cuda")
    big = torch.randint(0, 256, (2, 480, 640, 3), dtype=torch.uint8, device="cuda")
    batch = cb.prepare([small, small], [[0, 1], [1, 0]])
    want = cb.run(batch).clone()
    vt = (_lib.Video * 2).from_buffer(batch.host.numpy())
    vt[1].base, vt[1].nbytes, vt[1].height, vt[1].width = big.data_ptr(), big.numel(), 480, 640
    assert _lib.lib().lvc_clip_validate(C.byref(batch.desc), batch.max_h, vt) == _lib.E_SHAPE
    batch.tables.copy_(batch.host)
    out = torch.full(batch.shape, -5.0, dtype=torch.float32, device="cuda")
    cb.run(batch, out)
    assert device_status() == {"frame_too_large"}
    assert torch.equal(out[0], want[0])        # the other video is right; the offender's frames are undefined (but the
    torch.cuda.synchronize()                   # kernel stayed inside its buffers: no fault, context alive)
    assert torch.equal(cb.run(cb.prepare([small, small], [[0, 1], [1, 0]])), want)


@pytest.mark.gpu
def test_patch_embed_refuses_to_run_under_autograd():
    import torch

    from lvc_b200 import patch_embed as pe

    m = pe.TimesformerPatchEmbeddings(image_size=32).cuda()
    x = torch.randn(1, 2, 3, 32, 32, device="cuda")
    with torch.enable_grad():
        with pytest.raises(RuntimeError, match="forward-only"):
            m(x)                                   # parameters want gradients: the reference trains this layer
        with torch.no_grad():
            y = m(x)[0]
        m.requires_grad_(False)
        assert torch.equal(m(x)[0], y)             # frozen layer: fine while autograd records elsewhere
        with pytest.raises(RuntimeError, match="forward-only"):
            m(x.requires_grad_(True))              # a gradient would have to flow THROUGH the layer
    e = pe.VivitEmbeddings(num_frames=2, image_size=32).cuda()
    with torch.enable_grad(), pytest.raises(RuntimeError, match="forward-only"):
        e(torch.randn(1, 2, 3, 32, 32, device="cuda"))


@pytest.mark.gpu
def test_bf16_module_parameters_do_not_alias():
    """fp32 copies of bf16 parameters were temporaries freed before the launch (bias, LayerNorm weight and
    bias could all point at one recycled block)."""
    from functools import partial

    import torch

    from lvc_b200 import patch_embed as pe
    from oracle import patch_embed as ope

    torch.manual_seed(1)
    m = pe.PatchEmbed3d([2, 4, 4], 3, 128, partial(torch.nn.LayerNorm, eps=1e-5)).cuda()
    with torch.no_grad():
        m.proj.bias.normal_(0, 0.5)
        m.norm.weight.normal_(1, 0.3)
        m.norm.bias.normal_(0, 0.5)
    m = m.to(torch.bfloat16)
    x = torch.randn(1, 3, 4, 32, 32, device="cuda")
    y = m(x).float().cpu()
    f = lambda t: t.detach().float().cpu()  # noqa: E731
    ref = ope.swin_patch_embed(x.cpu(), f(m.proj.weight), f(m.proj.bias), f(m.norm.weight), f(m.norm.bias))
    e_max, e_l2 = ope.rel_err(y, ref)
    assert e_max <= 1e-2 and e_l2 <= 1e-2, (e_max, e_l2)
