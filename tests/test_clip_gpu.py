"""Parity of the fused CUDA clip kernel (through the C ABI) against the oracle and the golden
fixtures generated from the reference.  Needs a B200: run with `pytest -m gpu`.

Tolerances (BASELINE.json north_star): resized uint8 within 1 LSB -- asserted at 0 LSB;
normalised fp32 within 1e-5 -- asserted bit-exact where the LUT makes it so; bf16 = rounding of fp32.
"""
import hashlib

import numpy as np
import pytest
import torch

from oracle import processors as op
from oracle import samplers as osm

pytestmark = pytest.mark.gpu

METHODS = ["uniform", "random_window", "random"]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _videos(specs):
    from lvc_b200.synth import synth_video

    host = [synth_video(*s) for s in specs]
    return host, [torch.from_numpy(v).cuda() for v in host]


SHAPES = [  # (seed, N, H, W, mode)
    (1, 6, 720, 1280, "noise"), (2, 5, 480, 640, "ramp"), (3, 4, 100, 150, "noise"),
    (4, 4, 448, 448, "noise"), (5, 4, 449, 447, "noise"), (6, 3, 37, 1000, "noise"),
    (7, 3, 1, 1, "noise"), (8, 4, 224, 224, "noise"), (9, 4, 1080, 1920, "noise"),
    (10, 4, 223, 225, "ramp"), (11, 3, 2, 3, "noise"), (12, 4, 300, 200, "noise"),
]


def test_resized_uint8_bit_exact_all_shapes():
    from lvc_b200 import ClipBuilder

    T = 4
    host, dev = _videos(SHAPES)
    idx = [osm.dataset_indices("uniform", v.shape[0], T) for v in host]
    cb = ClipBuilder("dataset_u8", num_frames=T)
    out = cb(dev, idx).cpu().numpy()
    assert out.shape == (len(host), T, 224, 224, 3)
    for b, v in enumerate(host):
        want = op.resize_clip_cv2(v[idx[b]])
        diff = np.abs(out[b].astype(int) - want.astype(int))
        assert diff.max() == 0, (SHAPES[b], int(diff.max()), int((diff > 0).sum()))


@pytest.mark.parametrize("preset,layout", [("resnet_infer", "BCTHW"), ("timesformer", "BTCHW")])
def test_normalised_fp32_matches_reference_recipes(preset, layout):
    from lvc_b200 import ClipBuilder

    T = 8
    host, dev = _videos(SHAPES[:5])
    idx = [osm.dataset_indices("random_window", v.shape[0], T, __import__("random").Random(3)) for v in host]
    cb = ClipBuilder(preset, num_frames=T)
    assert cb.layout == layout
    out = cb(dev, idx).cpu().numpy()
    for b, v in enumerate(host):
        r = op.resize_clip_cv2(v[idx[b]])
        want = op.manual_normalize(r) if preset == "resnet_infer" else op.videomae_processor(r)
        assert out[b].shape == want.shape
        assert np.abs(out[b] - want).max() <= 1e-5
        assert (out[b] != want).sum() == 0  # LUT makes it bit-exact


def test_golden_config1_predict_video(golden):
    """The GPU path reproduces the sha256 of the tensor the reference's own predict_video fed its model."""
    from lvc_b200 import ClipBuilder, sampling, _lib
    from lvc_b200.synth import synth_video

    g = golden("pixels.npz")
    for k in range(3):
        seed, n, h, w, m, T = (int(v) for v in g[f"infer_case_{k}"])
        vid = synth_video(seed, n, h, w)
        idx = sampling.get_sampling_indices(METHODS[m], n, T, sampling.SamplerRng(42, "np"), _lib.VARIANT_NUMPY_INFER)
        assert idx == g[f"infer_idx_{k}"].tolist()
        out = ClipBuilder("resnet_infer", num_frames=T)([torch.from_numpy(vid).cuda()], [idx]).cpu().numpy()
        assert out.shape == (1, 3, T, 224, 224)
        assert sha(out) == str(g[f"infer_sha_{k}"])


def test_golden_cv2_resize(golden):
    from lvc_b200 import ClipBuilder
    from lvc_b200.synth import synth_video

    g = golden("resize.npz")
    cb = ClipBuilder("dataset_u8", num_frames=1)
    for i, (seed, h, w, ramp) in enumerate(g["cases"]):
        img = synth_video(int(seed), 1, int(h), int(w), "ramp" if ramp else "noise")
        out = cb([torch.from_numpy(img).cuda()], [[0]]).cpu().numpy()[0, 0]
        assert sha(out) == str(g[f"cv2_sha_{i}"]), (i, h, w)
    # ROI crops: numpy slice + cv2.resize
    img = synth_video(31, 1, 720, 1280)
    dev = torch.from_numpy(img).cuda()
    for i, box in enumerate(g["roi_boxes"]):
        out = cb([dev], [[0]], boxes=[[box]]).cpu().numpy()[0, 0]
        assert sha(out) == str(g[f"roi_sha_{i}"]), i


def test_roi_boxes_layouts_bf16_and_stats():
    from lvc_b200 import ClipBuilder
    from lvc_b200.synth import synth_boxes

    T = 8
    host, dev = _videos([(21, 12, 720, 1280, "noise"), (22, 9, 360, 640, "ramp"), (23, 8, 1080, 1920, "noise")])
    idx = [osm.dataset_indices("uniform", v.shape[0], T) for v in host]
    boxes = np.stack([synth_boxes(2000 + b, T, v.shape[1], v.shape[2], p_none=0.25) for b, v in enumerate(host)])
    boxes[0, 0] = [-5, -7, 100000, 100000]  # x1 < 0 = no detection
    boxes[1, 1] = [10, 20, 10, 20]          # empty box -> 1x1 crop
    boxes[2, 2] = [1900, 1000, 5000, 5000]  # clipped to the frame
    want_u8 = np.stack([op.resize_clip_cv2(v[idx[b]], boxes[b]) for b, v in enumerate(host)])
    stats = torch.zeros(len(host), dtype=torch.int64, device="cuda")
    u8 = ClipBuilder("dataset_u8", num_frames=T)(dev, idx, boxes=boxes, stats=stats).cpu().numpy()
    assert np.array_equal(u8, want_u8)
    assert stats.cpu().tolist() == [int(w.astype(np.int64).sum()) for w in want_u8]
    want = np.stack([op.manual_normalize(w) for w in want_u8])  # (B,C,T,H,W)
    f32 = ClipBuilder("resnet_infer", num_frames=T)(dev, idx, boxes=boxes)
    assert np.array_equal(f32.cpu().numpy(), want)
    btchw = ClipBuilder("resnet_infer", num_frames=T, layout="BTCHW")(dev, idx, boxes=boxes)
    assert torch.equal(btchw, f32.permute(0, 2, 1, 3, 4))
    bf = ClipBuilder("resnet_infer", num_frames=T, dtype=torch.bfloat16)(dev, idx, boxes=boxes)
    assert torch.equal(bf, f32.to(torch.bfloat16))
    planar = ClipBuilder("dataset_u8", num_frames=T, layout="BTCHW")(dev, idx, boxes=boxes)
    assert torch.equal(planar.cpu(), torch.from_numpy(want_u8).permute(0, 1, 4, 2, 3))


def test_edge_cases_empty_ragged_short_and_out_of_range():
    from lvc_b200 import ClipBuilder, sampling

    cb = ClipBuilder("dataset_u8", num_frames=16)
    assert tuple(cb([], []).shape) == (0, 16, 224, 224, 3)
    # short video (N < T): duplicate-allowing branch; ragged sizes in one batch
    host, dev = _videos([(31, 5, 120, 160, "noise"), (32, 40, 90, 300, "ramp")])
    rng = sampling.SamplerRng(42)
    idx = [sampling.get_sampling_indices("random_window", v.shape[0], 16, rng) for v in host]
    out = cb(dev, idx).cpu().numpy()
    for b, v in enumerate(host):
        assert np.array_equal(out[b], op.resize_clip_cv2(v[idx[b]]))
    # indices beyond the video repeat the last frame (vivit .../dataset.py:256-265), negatives clamp to 0
    bad = [[-3, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 99, 10 ** 6]] * 2
    out = cb(dev, bad).cpu().numpy()
    for b, v in enumerate(host):
        cl = np.clip(bad[b], 0, v.shape[0] - 1)
        assert np.array_equal(out[b], op.resize_clip_cv2(v[cl]))
    with pytest.raises(ValueError):
        cb([dev[0].float()], [idx[0]])
    with pytest.raises(ValueError):
        ClipBuilder("nope")
    with pytest.raises(ValueError):   # a video without frames (the reference raises too, vivit .../dataset.py:217)
        cb([torch.empty((0, 120, 160, 3), dtype=torch.uint8, device="cuda")], [list(range(16))])


def test_full_size_properties_1080p():
    """BASELINE config 5 frame size: size-independent checks (checksum of checksums, identity,
    constant images, determinism) plus oracle spot checks."""
    from lvc_b200 import ClipBuilder

    B, N, T = 8, 48, 32
    g = torch.Generator(device="cuda").manual_seed(5)
    vids = [torch.randint(0, 256, (N, 1080, 1920, 3), dtype=torch.uint8, device="cuda", generator=g) for _ in range(B)]
    idx = [osm.dataset_indices("uniform", N, T) for _ in range(B)]
    cb = ClipBuilder("dataset_u8", num_frames=T)
    stats = torch.zeros(B, dtype=torch.int64, device="cuda")
    u8 = cb(vids, idx, stats=stats)
    assert torch.equal(stats, u8.to(torch.int64).sum(dim=(1, 2, 3, 4)))
    assert torch.equal(u8, cb(vids, idx))  # deterministic
    for b, t in [(0, 0), (3, 17), (7, 31)]:
        want = op.resize_clip_cv2(vids[b][idx[b][t]][None].cpu().numpy())[0]
        assert np.array_equal(u8[b, t].cpu().numpy(), want)
    f32 = ClipBuilder("resnet_infer", num_frames=T, layout="BTCHW")(vids, idx)
    lut = ClipBuilder("resnet_infer").lut
    want = torch.stack([lut[c][u8[..., c].long()] for c in range(3)], dim=2)
    assert torch.equal(f32, want)
    # identity: a 224x224 source passes through unchanged; constant image stays constant
    small = torch.randint(0, 256, (T, 224, 224, 3), dtype=torch.uint8, device="cuda", generator=g)
    assert torch.equal(cb([small], [list(range(T))])[0], small)
    const = torch.full((2, 1080, 1920, 3), 137, dtype=torch.uint8, device="cuda")
    assert bool((ClipBuilder("dataset_u8", num_frames=2)([const], [[0, 1]]) == 137).all())


def test_vivit_preset_fuses_pillow_stage(golden):
    """cv2 resize -> VivitImageProcessor (Pillow 224->256, crop 224, rescale+offset, normalise) in one kernel."""
    from lvc_b200 import ClipBuilder
    from lvc_b200.synth import synth_video

    # golden: the HF processor's own output on 224x224 frames (the cv2 stage is the identity there)
    g = golden("pixels.npz")
    fr = synth_video(111, 4, 224, 224)
    out = ClipBuilder("vivit", num_frames=4)([torch.from_numpy(fr).cuda()], [[0, 1, 2, 3]]).cpu().numpy()
    assert list(out.shape) == g["vivit_shape"].tolist()
    assert sha(out) == str(g["vivit_sha"])
    # general sources, ROI boxes, ragged batch
    T = 8
    host, dev = _videos([(51, 10, 720, 1280, "noise"), (52, 9, 480, 640, "ramp"), (53, 8, 100, 150, "noise"),
                         (54, 8, 1080, 1920, "noise")])
    idx = [osm.dataset_indices("uniform", v.shape[0], T) for v in host]
    out = ClipBuilder("vivit", num_frames=T)(dev, idx).cpu().numpy()
    for b, v in enumerate(host):
        want = op.vivit_processor(op.resize_clip_cv2(v[idx[b]]))
        assert np.abs(out[b] - want).max() <= 1e-5
        assert (out[b] != want).sum() == 0


def test_ptv_eval_preset_matches_torch_chain():
    """ShortSideScale(256) + CenterCrop(224) + Normalize on 0..255 floats.  Tolerance:
    values reach 1131, so 1e-5 is read relative to the range: atol 1.1e-2, rtol 1e-5."""
    from lvc_b200 import ClipBuilder

    T = 4
    host, dev = _videos([(61, 6, 720, 1280, "noise"), (62, 5, 480, 640, "ramp"), (63, 5, 640, 480, "noise"),
                         (64, 4, 300, 300, "noise"), (65, 4, 1080, 1920, "noise")])
    idx = [osm.dataset_indices("uniform", v.shape[0], T) for v in host]
    cb = ClipBuilder("ptv_eval", num_frames=T)
    out = cb(dev, idx).cpu().numpy()
    assert out.shape == (len(host), 3, T, 224, 224)
    worst = 0.0
    for b, v in enumerate(host):
        want = op.ptv_eval_chain(v[idx[b]])
        assert np.allclose(out[b], want, atol=1.1e-2, rtol=1e-5)
        worst = max(worst, float(np.abs(out[b] - want).max()))
    print("ptv_eval max abs err", worst)
    with pytest.raises(NotImplementedError):
        cb(dev, idx, boxes=np.zeros((len(host), T, 4), np.int32))


@pytest.mark.parametrize("preset", ["resnet_infer", "vivit", "dataset_u8", "ptv_eval"])
def test_bgr_source_swaps_channels_in_the_gather(preset):
    """Row f1: frames as cv2.VideoCapture.read() returns them (BGR).  src_order="BGR" must give exactly what the
    RGB path gives on cv2.cvtColor(frame, COLOR_BGR2RGB) input (vivit .../dataset.py:336), incl. the checksum."""
    from lvc_b200 import ClipBuilder

    T = 4
    host, dev = _videos([(71, 6, 360, 640, "noise"), (72, 5, 100, 150, "ramp"), (73, 4, 448, 448, "noise")])
    bgr = [torch.from_numpy(np.ascontiguousarray(v[..., ::-1])).cuda() for v in host]
    idx = [osm.dataset_indices("uniform", v.shape[0], T) for v in host]
    a = ClipBuilder(preset, num_frames=T)
    b = ClipBuilder(preset, num_frames=T, src_order="BGR")
    sa = torch.zeros(len(host), dtype=torch.int64, device="cuda")
    sb = torch.zeros_like(sa)
    if preset == "ptv_eval":
        oa, ob = a(dev, idx), b(bgr, idx)
    else:
        oa, ob = a(dev, idx, stats=sa), b(bgr, idx, stats=sb)
    assert torch.equal(oa, ob)
    assert torch.equal(sa, sb)
    with pytest.raises(ValueError):
        ClipBuilder(preset, num_frames=T, src_order="GBR")


def test_ptv_train_chain_matches_torchvision_under_seed():
    """Row f2: RandomShortSideScale(256,320) + RandomCrop(224) + RandomHorizontalFlip(0.5) + Normalize, one draw
    per clip on the torch-compatible stream.  Oracle = torch CPU + torchvision's own RandomCrop / RandomHorizontalFlip
    under torch.manual_seed (RandomShortSideScale restated, pytorchvideo absent).  Same tolerance as the eval chain."""
    from lvc_b200 import ClipBuilder, sampling

    T = 4
    host, dev = _videos([(81, 6, 720, 1280, "noise"), (82, 5, 480, 640, "ramp"), (83, 5, 640, 480, "noise"),
                         (84, 4, 300, 300, "noise"), (85, 4, 224, 224, "noise"), (86, 4, 1080, 1920, "noise")])
    idx = [osm.dataset_indices("uniform", v.shape[0], T) for v in host]
    for seed in (0, 3):
        rng = sampling.TorchRng(seed)
        aug = [sampling.train_aug_params(v.shape[1], v.shape[2], rng) for v in host]
        out = ClipBuilder("ptv_eval", num_frames=T)(dev, idx, aug=aug).cpu().numpy()
        torch.manual_seed(seed)
        worst, flips = 0.0, 0
        for b, v in enumerate(host):
            want, params = op.ptv_train_chain(v[idx[b]], seed=None)
            assert params == aug[b]
            flips += params[3]
            assert np.allclose(out[b], want, atol=1.1e-2, rtol=1e-5)
            worst = max(worst, float(np.abs(out[b] - want).max()))
        print("ptv_train seed", seed, "max abs err", worst, "flips", flips)
    with pytest.raises(ValueError):
        ClipBuilder("resnet_infer", num_frames=T)(dev, idx, aug=aug)
    with pytest.raises(ValueError):
        ClipBuilder("ptv_eval", num_frames=T)(dev, idx, aug=[[100, 0, 0, 0]] * len(host))


def test_swin_dataset_train_mode_draws_per_clip():
    from lvc_b200.datasets import SwinVideoDataset, TensorFrameSource

    host, _ = _videos([(91, 40, 240, 320, "noise"), (92, 36, 320, 240, "ramp")])
    src = TensorFrameSource(host, labels=[0, 1])
    ds = SwinVideoDataset(mode="train", num_frames=8, frame_source=src, aug_seed=11)
    ev = SwinVideoDataset(mode="val", num_frames=8, frame_source=src)
    clips, labels = ds.get_batch([0, 1])
    assert clips.shape == (2, 1, 3, 8, 224, 224) and labels.tolist() == [[0], [1]]
    assert ds.last_aug is not None and ev.augment is False
    torch.manual_seed(11)
    for b in range(2):
        fidx = ds.get_sampling_indices(ds.video_paths[b], host[b].shape[0])[0]
        lo, hi = min(fidx), max(fidx)
        from lvc_b200.sampling import uniform_temporal_subsample
        pick = [lo + k for k in uniform_temporal_subsample(hi - lo + 1, 8)]
        want, params = op.ptv_train_chain(host[b][pick], seed=None)
        assert params == ds.last_aug[b]
        assert np.allclose(clips[b, 0].cpu().numpy(), want, atol=1.1e-2, rtol=1e-5)


def test_float_frame_branch_of_cv2_resize():
    """a6, primary Dataset branch (vivit .../dataset.py:244-284): float32 frames -> cv2.resize in float ->
    truncation.  Kernel vs the oracle restatement: bit-exact; vs the installed cv2: within 1 LSB, > 99.5 % equal."""
    import cv2

    from lvc_b200 import ClipBuilder
    from oracle.resize import cv2_resize_linear_f32_trunc

    T = 3
    host, dev = _videos([(101, 4, 720, 1280, "noise"), (102, 4, 480, 640, "ramp"), (103, 3, 100, 150, "noise"),
                         (104, 3, 448, 448, "noise"), (105, 3, 224, 224, "noise"), (106, 3, 37, 1000, "noise")])
    idx = [osm.dataset_indices("uniform", v.shape[0], T) for v in host]
    out = ClipBuilder("dataset_u8", num_frames=T, float_frames=True)(dev, idx).cpu().numpy()
    plain = ClipBuilder("dataset_u8", num_frames=T)(dev, idx).cpu().numpy()
    for b, v in enumerate(host):
        for t, fi in enumerate(idx[b]):
            want = cv2_resize_linear_f32_trunc(v[fi])
            assert np.array_equal(out[b, t], want), (b, t)
            ref = cv2.resize(v[fi].astype(np.float32), (224, 224)).astype(np.uint8)
            d = np.abs(ref.astype(np.int16) - out[b, t].astype(np.int16))
            assert d.max() <= 1 and (d == 0).mean() > 0.995
            assert np.abs(plain[b, t].astype(np.int16) - out[b, t].astype(np.int16)).max() <= 1
    # the variant feeds every downstream stage unchanged (LUT normalisation here)
    f = ClipBuilder("resnet_infer", num_frames=T, float_frames=True)(dev, idx).cpu().numpy()
    want = op.manual_normalize(np.stack([cv2_resize_linear_f32_trunc(host[0][fi]) for fi in idx[0]]))
    assert np.array_equal(f[0], want)
    with pytest.raises(ValueError):
        ClipBuilder("ptv_eval", num_frames=T, float_frames=True)


def test_very_wide_frames_small_stages():
    """Frames so wide that only 2 / 4 output rows fit a ring stage (rows per stage falls back 8 -> 4 -> 2):
    same bytes as cv2.resize."""
    from lvc_b200 import ClipBuilder

    T = 3
    specs = [(301, 3, 64, 7000, "noise"), (302, 3, 300, 4096, "ramp")]
    for spec in specs:
        host, dev = _videos([spec])
        idx = [osm.dataset_indices("uniform", host[0].shape[0], T)]
        out = ClipBuilder("dataset_u8", num_frames=T)(dev, idx).cpu().numpy()
        want = op.resize_clip_cv2(host[0][idx[0]])
        assert np.array_equal(out[0], want), spec


_KNOB_SCRIPT = """
import hashlib, random, sys
import numpy as np, torch
sys.path.insert(0, {root!r})
from lvc_b200 import ClipBuilder
from lvc_b200.synth import synth_video
from oracle import samplers as osm
specs = {specs!r}
host = [synth_video(*s) for s in specs]
dev = [torch.from_numpy(v).cuda() for v in host]
idx = [osm.dataset_indices("random", v.shape[0], 5, random.Random(11)) for v in host]
out = {{}}
for preset in ("resnet_infer", "vivit"):
    out[preset] = hashlib.sha256(ClipBuilder(preset, num_frames=5)(dev, idx).cpu().numpy().tobytes()).hexdigest()
# the Video Swin token layout (4 x 4 patches: whole-sector stores through the lane transpose), with and without a checksum
idx4 = [i[:4] for i in idx]
cb = ClipBuilder("resnet_infer", num_frames=4, layout="TOKENS", patch=(2, 4, 4), src_order="BGR")
dev4, idx4 = dev[1:], idx4[1:]   # without the 720p video: 8-row stages, which is where the whole-sector path runs
st = torch.zeros(len(dev4), dtype=torch.int64, device="cuda")
tok = cb(dev4, idx4).view(torch.int16).cpu().numpy().tobytes()
tok_s = cb(dev4, idx4, stats=st).view(torch.int16).cpu().numpy().tobytes()
assert tok == tok_s
out["tok44"] = hashlib.sha256(tok + st.cpu().numpy().tobytes()).hexdigest()
print("SHA", out["resnet_infer"], out["vivit"], out["tok44"])
"""


@pytest.mark.parametrize("knobs", [{}, {"LVC_CLIP_NP": "1"}, {"LVC_CLIP_NP": "2"}, {"LVC_CLIP_NP": "4", "LVC_CLIP_RPS": "4"},
                                   {"LVC_CLIP_MERGE": "0"}, {"LVC_CLIP_MERGE": "0", "LVC_CLIP_NP": "1", "LVC_CLIP_RPS": "8"},
                                   {"LVC_CLIP_NP": "4", "LVC_CLIP_RPS": "2", "LVC_CLIP_BAND": "2"},
                                   {"LVC_CLIP_SMEM_KB": "100", "LVC_CLIP_VBANDS": "14"},
                                   {"LVC_CLIP_NO_TOK44": "1"}, {"LVC_CLIP_NO_FMA": "1"}, {"LVC_CLIP_NO_FMA": "1", "LVC_CLIP_NO_TOK44": "1"}])
def test_producer_count_and_stage_knobs_do_not_change_bytes(knobs):
    """The launch geometry (producer warps, rows per stage, band size, ring size) is a performance choice: results are
    identical.  The knobs are honoured only under LVC_PROFILING_KNOBS=1, read once per process -> a subprocess."""
    import os
    import subprocess
    import sys

    from lvc_b200 import ClipBuilder

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    T = 5
    host, dev = _videos(SHAPES[:6])
    idx = [osm.dataset_indices("random", v.shape[0], T, __import__("random").Random(11)) for v in host]
    base = {p: sha(ClipBuilder(p, num_frames=T)(dev, idx).cpu().numpy()) for p in ("resnet_infer", "vivit")}
    # Video Swin tokens: the shipped path (lane transpose, affine bf16 form) must equal every fallback, and the planar clip
    dev4, idx4 = dev[1:], [i[:4] for i in idx][1:]   # without the 720p video: 8-row stages (the whole-sector path)
    cbt = ClipBuilder("resnet_infer", num_frames=4, layout="TOKENS", patch=(2, 4, 4), src_order="BGR")
    st = torch.zeros(len(dev4), dtype=torch.int64, device="cuda")
    tok = cbt(dev4, idx4, stats=st)
    base["tok44"] = sha(np.frombuffer(tok.view(torch.int16).cpu().numpy().tobytes() + st.cpu().numpy().tobytes(), dtype=np.uint8))
    planar = ClipBuilder("resnet_infer", num_frames=4, layout="BCTHW", dtype=torch.bfloat16, src_order="BGR")(dev4, idx4)
    B = len(dev4)   # A[m][k]: m = (b, t/2, y/4, x/4), k = (c, t%2, y%4, x%4)
    want_tok = planar.view(B, 3, 2, 2, 56, 4, 56, 4).permute(0, 2, 4, 6, 1, 3, 5, 7).reshape(B * 2 * 56 * 56, 96)
    assert torch.equal(tok.view(torch.int16), want_tok.contiguous().view(torch.int16))
    env = {k: v for k, v in os.environ.items() if not k.startswith("LVC_")}
    env.update(knobs, LVC_PROFILING_KNOBS="1")
    r = subprocess.run([sys.executable, "-c", _KNOB_SCRIPT.format(root=root, specs=SHAPES[:6])], env=env, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    got = [ln for ln in r.stdout.splitlines() if ln.startswith("SHA")][-1].split()[1:]
    assert got == [base["resnet_infer"], base["vivit"], base["tok44"]], knobs
    want = op.manual_normalize(op.resize_clip_cv2(host[0][idx[0]]))
    assert sha(want) == sha(ClipBuilder("resnet_infer", num_frames=T)(dev[:1], idx[:1]).cpu().numpy()[0])
