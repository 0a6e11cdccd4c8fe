"""Parity of the tcgen05 patch-embedding GEMM against the fp32 oracle and the golden outputs of the
HF / torchvision modules.  Tolerance (north star): bf16 embeddings within 1e-2 relative of the fp32
torch path: max|y-ref|/max|ref| <= 1e-2 and ||y-ref||/||ref|| <= 1e-2."""
from functools import partial

import numpy as np
import pytest
import torch

from oracle import patch_embed as ope

pytestmark = pytest.mark.gpu
TOL = 1e-2


def sd(shape, seed, scale=1.0):
    return torch.from_numpy((np.random.default_rng(seed).standard_normal(shape) * scale).astype(np.float32))


def check(y, ref, tol=TOL):
    e_max, e_l2 = ope.rel_err(y.float().cpu(), ref)
    assert e_max <= tol and e_l2 <= tol, (e_max, e_l2)
    return e_max, e_l2


def _vivit(image, frames, out_dtype=torch.bfloat16):
    from lvc_b200.patch_embed import VivitTubeletEmbeddings

    m = VivitTubeletEmbeddings(num_frames=frames, image_size=image, out_dtype=out_dtype)
    m.projection.weight.data.copy_(sd(m.projection.weight.shape, 1, 0.02))
    m.projection.bias.data.copy_(sd(m.projection.bias.shape, 2, 0.02))
    return m.cuda()


def _swin(dim, out_dtype=torch.bfloat16):
    from lvc_b200.patch_embed import PatchEmbed3d

    s = PatchEmbed3d([2, 4, 4], 3, dim, partial(torch.nn.LayerNorm, eps=1e-5), out_dtype=out_dtype)
    s.proj.weight.data.copy_(sd(s.proj.weight.shape, 7, 0.02))
    s.proj.bias.data.copy_(sd(s.proj.bias.shape, 8, 0.02))
    s.norm.weight.data.copy_(1 + sd(s.norm.weight.shape, 9, 0.1))
    s.norm.bias.data.copy_(sd(s.norm.bias.shape, 10, 0.1))
    return s.cuda()


def _swin_ref(s, x):
    return ope.swin_patch_embed(x, s.proj.weight.detach().cpu(), s.proj.bias.detach().cpu(),
                                s.norm.weight.detach().cpu(), s.norm.bias.detach().cpu())


def test_golden_small_modules(golden):
    from lvc_b200.patch_embed import TimesformerPatchEmbeddings

    g = golden("embed.npz")
    m = _vivit(64, 4)
    y = m(sd((2, 4, 3, 64, 64), 3).cuda())
    assert y.dtype == torch.bfloat16 and tuple(y.shape) == g["vivit_out"].shape
    check(y, torch.from_numpy(g["vivit_out"]))
    t = TimesformerPatchEmbeddings(image_size=64)
    t.projection.weight.data.copy_(sd(t.projection.weight.shape, 4, 0.02))
    t.projection.bias.data.copy_(sd(t.projection.bias.shape, 5, 0.02))
    emb, nf, pw = t.cuda()(sd((2, 3, 3, 64, 64), 6).cuda())
    assert [nf, pw] == g["tsf_meta"].tolist() and tuple(emb.shape) == g["tsf_out"].shape
    check(emb, torch.from_numpy(g["tsf_out"]))
    s = _swin(128)
    y = s(sd((2, 3, 4, 32, 32), 11).cuda())
    assert tuple(y.shape) == g["swin_out"].shape
    check(y, torch.from_numpy(g["swin_out"]))
    y = s(sd((1, 3, 3, 30, 29), 12).cuda())  # padded T/H/W (swin_transformer.py:353-358)
    assert tuple(y.shape) == g["swin_out_pad"].shape
    check(y, torch.from_numpy(g["swin_out_pad"]))


def test_vivit_config2_full_size():
    """BASELINE config 2: B=8, T=32, 224x224 -> M=25088, K=1536, N=768."""
    m = _vivit(224, 32, torch.float32)
    x = sd((8, 32, 3, 224, 224), 21)
    y = m(x.cuda())
    assert tuple(y.shape) == (8, 3136, 768) and y.dtype == torch.float32
    ref = ope.vivit_tubelet_embed(x, m.projection.weight.detach().cpu(), m.projection.bias.detach().cpu())
    print("vivit cfg2 rel err", check(y, ref))
    m.out_dtype = torch.bfloat16  # bf16 input / bf16 output path
    yb = m(x.cuda().to(torch.bfloat16))
    assert yb.dtype == torch.bfloat16
    check(yb, ref)
    with pytest.raises(ValueError):
        m(torch.zeros(1, 32, 3, 200, 224, device="cuda"))


def test_timesformer_config3_full_size():
    from lvc_b200.patch_embed import TimesformerPatchEmbeddings

    t = TimesformerPatchEmbeddings()
    t.projection.weight.data.copy_(sd(t.projection.weight.shape, 4, 0.02))
    t.projection.bias.data.copy_(sd(t.projection.bias.shape, 5, 0.02))
    t = t.cuda()
    x = sd((16, 8, 3, 224, 224), 22)
    emb, nf, pw = t(x.cuda())
    assert tuple(emb.shape) == (128, 196, 768) and (nf, pw) == (8, 14)
    ref, _, _ = ope.timesformer_patch_embed(x, t.projection.weight.detach().cpu(), t.projection.bias.detach().cpu())
    print("timesformer cfg3 rel err", check(emb, ref))


def test_swin_config4_full_size():
    s = _swin(128, torch.float32)
    x = sd((8, 3, 32, 224, 224), 23)
    y = s(x.cuda())
    assert tuple(y.shape) == (8, 16, 56, 56, 128)
    print("swin cfg4 rel err", check(y, _swin_ref(s, x)))
    s96 = _swin(96, torch.float32)  # Swin-T width, the reference's default model_size
    x = sd((1, 3, 8, 64, 64), 24)
    check(s96(x.cuda()), _swin_ref(s96, x))


def test_clip_kernel_feeds_patch_embed():
    """configs[2] end to end on the device: fused clip kernel (TimeSformer preset) -> tcgen05 embedding."""
    from lvc_b200 import ClipBuilder
    from lvc_b200.patch_embed import TimesformerPatchEmbeddings
    from lvc_b200.synth import synth_video
    from oracle import processors as op
    from oracle import samplers as osm

    host = [synth_video(40 + i, 12, 240, 320) for i in range(2)]
    idx = [osm.dataset_indices("uniform", 12, 8) for _ in host]
    clips = ClipBuilder("timesformer", num_frames=8)([torch.from_numpy(v).cuda() for v in host], idx)
    t = TimesformerPatchEmbeddings().cuda()
    emb, _, _ = t(clips)
    want_clip = np.stack([op.videomae_processor(op.resize_clip_cv2(v[i])) for v, i in zip(host, idx)])
    ref, _, _ = ope.timesformer_patch_embed(torch.from_numpy(want_clip), t.projection.weight.detach().cpu(),
                                            t.projection.bias.detach().cpu())
    check(emb, ref)


@pytest.mark.parametrize("preset,kind", [("vivit", "vivit"), ("timesformer", "timesformer"), ("resnet_infer", "swin")])
def test_fused_token_layout_skips_im2col(preset, kind):
    """The clip kernel writes the bf16 patch matrix A[M][K] itself (layout TOKENS); the GEMM then runs
    alone.  Must equal: clip kernel (bf16, planar) -> module forward (im2col + GEMM) bit for bit."""
    from lvc_b200 import ClipBuilder
    from lvc_b200.patch_embed import TimesformerPatchEmbeddings, VivitTubeletEmbeddings
    from lvc_b200.synth import synth_video
    from oracle import samplers as osm

    T = 8
    host = [synth_video(70 + i, 12, 180, 320) for i in range(3)]
    dev = [torch.from_numpy(v).cuda() for v in host]
    idx = [osm.dataset_indices("uniform", 12, T) for _ in host]
    if kind == "vivit":
        m, patch, lay = _vivit(224, T), (2, 16, 16), "BTCHW"
    elif kind == "timesformer":
        m, patch, lay = TimesformerPatchEmbeddings().cuda(), (1, 16, 16), "BTCHW"
    else:
        m, patch, lay = _swin(128), (2, 4, 4), "BCTHW"
    planar = ClipBuilder(preset, num_frames=T, layout=lay, dtype=torch.bfloat16)(dev, idx)
    want = m(planar)
    want = want[0] if isinstance(want, tuple) else want
    tokens = ClipBuilder(preset, num_frames=T, layout="TOKENS", patch=patch)(dev, idx)
    assert tokens.dtype == torch.bfloat16 and tokens.shape[1] == 3 * patch[0] * patch[1] * patch[2]
    launches = m.launches
    got = m.forward_tokens(tokens, len(host), T)
    assert m.launches - launches == 1
    assert torch.equal(got.reshape(want.shape), want)


# ------------------------------------------------------------------ row f3: fused embeddings epilogue ----
def _fill(params, seed0, scale=0.02):
    for k, t in enumerate(params):
        t.data.copy_(sd(t.shape, seed0 + k, scale))


@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float32])
def test_vivit_embeddings_match_hf_module(out_dtype):
    """VivitEmbeddings (modeling_vivit.py:141-155): cls token + position embeddings fused into the GEMM epilogue.
    Oracle = the HF module itself (fp32, CPU) with the same state dict."""
    from transformers import VivitConfig
    from transformers.models.vivit.modeling_vivit import VivitEmbeddings as HF

    from lvc_b200.patch_embed import VivitEmbeddings

    cfg = VivitConfig(num_frames=8, hidden_dropout_prob=0.0)
    ref = HF(cfg).eval()
    _fill(list(ref.parameters()), 20)
    mine = VivitEmbeddings(cfg, out_dtype=out_dtype)
    missing = mine.load_state_dict(ref.state_dict(), strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    mine = mine.cuda()
    x = sd((3, 8, 3, 224, 224), 5)
    with torch.no_grad():
        want = ref(x)
    y = mine(x.cuda())
    assert y.dtype == out_dtype and tuple(y.shape) == tuple(want.shape) == (3, 1 + 4 * 196, 768)
    check(y, want)
    # the cls row is cls_token + position_embeddings[0], identical for every batch entry
    c = (ref.cls_token[0, 0] + ref.position_embeddings[0, 0]).detach()
    assert torch.allclose(y[:, 0].float().cpu(), c.expand(3, -1), atol=2e-2 if out_dtype == torch.bfloat16 else 1e-6)


def test_timesformer_embeddings_match_hf_module():
    """TimesformerEmbeddings (modeling_timesformer.py:92-144), divided space-time: position + time embeddings and the
    (B, patch*T + t) token order are produced by the GEMM epilogue's scatter."""
    from transformers import TimesformerConfig
    from transformers.models.timesformer.modeling_timesformer import TimesformerEmbeddings as HF

    from lvc_b200.patch_embed import TimesformerEmbeddings

    cfg = TimesformerConfig(num_frames=8, hidden_dropout_prob=0.0)
    ref = HF(cfg).eval()
    _fill(list(ref.parameters()), 40)
    mine = TimesformerEmbeddings(cfg, out_dtype=torch.float32)
    mine.load_state_dict(ref.state_dict(), strict=True)
    mine = mine.cuda()
    x = sd((5, 8, 3, 224, 224), 6)   # 5*8*196 = 7840 rows: the last 256-row pair tile is ragged
    with torch.no_grad():
        want = ref(x)
    y = mine(x.cuda())
    assert tuple(y.shape) == tuple(want.shape) == (5, 1 + 196 * 8, 768)
    check(y, want)
    with pytest.raises(ValueError):
        mine(sd((1, 4, 3, 224, 224), 7).cuda())


def test_fused_clip_to_embeddings_pipeline():
    """clip kernel (token layout) -> one GEMM launch with the embeddings epilogue == module applied to the
    normalised clip tensor."""
    from lvc_b200 import ClipBuilder
    from lvc_b200.patch_embed import VivitEmbeddings
    from lvc_b200.synth import synth_video
    from oracle import samplers as osm

    T = 8
    host = [synth_video(31 + i, 12, 240, 320) for i in range(2)]
    dev = [torch.from_numpy(v).cuda() for v in host]
    idx = [osm.dataset_indices("uniform", 12, T) for _ in host]
    m = VivitEmbeddings(num_frames=T, out_dtype=torch.float32)
    _fill(list(m.parameters()), 60)
    m = m.cuda()
    clip = ClipBuilder("vivit", num_frames=T)(dev, idx)                      # (B,T,3,224,224) fp32
    tok = ClipBuilder("vivit", num_frames=T, layout="TOKENS", patch=(2, 16, 16))(dev, idx)
    a = m(clip)
    b = m(tokens=tok, batch=2)
    assert tuple(a.shape) == tuple(b.shape) == (2, 1 + 4 * 196, 768)
    check(b, a.float().cpu())


@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float32])
def test_vivit_embeddings_16_frames_ragged_last_tile(out_dtype):
    """num_frames = 16, B = 3: 4704 token rows leave a ragged last 256-row pair tile; rows of one warp stay inside
    one batch entry (1568 tokens per clip is a multiple of 32) while 128-row CTA tiles straddle entries."""
    from transformers import VivitConfig
    from transformers.models.vivit.modeling_vivit import VivitEmbeddings as HF

    from lvc_b200.patch_embed import VivitEmbeddings

    cfg = VivitConfig(num_frames=16, hidden_dropout_prob=0.0)
    ref = HF(cfg).eval()
    _fill(list(ref.parameters()), 80)
    mine = VivitEmbeddings(cfg, out_dtype=out_dtype)
    mine.load_state_dict(ref.state_dict(), strict=True)
    mine = mine.cuda()
    x = sd((3, 16, 3, 224, 224), 9)
    with torch.no_grad():
        want = ref(x)
    y = mine(x.cuda())
    assert tuple(y.shape) == tuple(want.shape) == (3, 1 + 8 * 196, 768)
    check(y, want)
