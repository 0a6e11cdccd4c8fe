"""lvc_clip_swin_embed_forward (clip kernel and PatchEmbed3d kernel co-resident, tile-level hand-over through ready
counters) against the two separate calls: same kernels' arithmetic, so the bar is byte equality of the token matrix
and of the embedding; the embedding is also checked against the fp32 oracle at the embed tolerance (1e-2 relative)."""
from functools import partial

import numpy as np
import pytest
import torch

from oracle import patch_embed as ope
from oracle import processors as op

pytestmark = pytest.mark.gpu


def _module(out_dtype=torch.bfloat16):
    from lvc_b200.patch_embed import PatchEmbed3d

    g = np.random.default_rng(5)
    m = PatchEmbed3d([2, 4, 4], 3, 128, partial(torch.nn.LayerNorm, eps=1e-5), out_dtype=out_dtype)
    m.proj.weight.data.copy_(torch.from_numpy((g.standard_normal(m.proj.weight.shape) * 0.02).astype(np.float32)))
    m.proj.bias.data.copy_(torch.from_numpy((g.standard_normal(128) * 0.02).astype(np.float32)))
    m.norm.weight.data.copy_(torch.from_numpy((1 + g.standard_normal(128) * 0.1).astype(np.float32)))
    m.norm.bias.data.copy_(torch.from_numpy((g.standard_normal(128) * 0.1).astype(np.float32)))
    return m.cuda()


def _batch(B, T, sizes, boxes, seed):
    from lvc_b200 import ClipBuilder
    from lvc_b200.synth import synth_boxes

    rng = np.random.default_rng(seed)
    host = [rng.integers(0, 256, (T + 3, h, w, 3), dtype=np.uint8) for (h, w) in (sizes * B)[:B]]
    vids = [torch.from_numpy(v).cuda() for v in host]
    idx = np.stack([np.sort(rng.integers(0, T + 3, T)) for _ in range(B)]).astype(np.int32)
    bx = None
    if boxes:
        bx = np.stack([synth_boxes(2000 + b, T + 3, host[b].shape[1], host[b].shape[2])[idx[b]] for b in range(B)])
    cb = ClipBuilder("resnet_infer", num_frames=T, layout="TOKENS", patch=(2, 4, 4))
    return cb, cb.prepare(vids, idx, bx), host, idx, bx


@pytest.mark.parametrize("B,T,sizes,boxes,out_dtype", [
    (2, 4, [(240, 320), (120, 200)], True, torch.bfloat16),
    (1, 2, [(480, 640)], False, torch.float32),
    (8, 32, [(720, 1280)], True, torch.bfloat16),          # BASELINE config 4 at full size: 3136 tiles
])
def test_fused_equals_the_two_separate_calls(B, T, sizes, boxes, out_dtype):
    m = _module(out_dtype)
    cb, batch, host, idx, bx = _batch(B, T, sizes, boxes, seed=B * 100 + T)
    tok = cb.run(batch)
    want = m.forward_tokens(tok, B, T, out_shape=(B, T // 2, 56, 56, 128))
    torch.cuda.synchronize()
    for rep in range(3):      # the ready counters are cleared by every call
        got = m.forward_clips(cb, batch)
        torch.cuda.synchronize()
        assert torch.equal(m.tokens_view(B, T).view(torch.int16), tok.view(torch.int16)), rep
        assert tuple(got.shape) == (B, T // 2, 56, 56, 128) and got.dtype == out_dtype
        assert torch.equal(got.view(torch.int32 if out_dtype == torch.float32 else torch.int16),
                           want.view(torch.int32 if out_dtype == torch.float32 else torch.int16)), rep
    if B <= 2:   # ... and the fp32 reference chain: cv2 resize + normalise + Conv3d + LayerNorm on the CPU
        clips = []
        for b in range(B):
            fr = host[b][idx[b]]
            clips.append(op.manual_normalize(op.resize_clip_cv2(fr, None if bx is None else bx[b])))   # (3, T, 224, 224)
        x = torch.from_numpy(np.stack(clips))
        ref = ope.swin_patch_embed(x, m.proj.weight.detach().cpu(), m.proj.bias.detach().cpu(),
                                   m.norm.weight.detach().cpu(), m.norm.bias.detach().cpu())
        e_max, e_l2 = ope.rel_err(got.float().cpu(), ref)
        assert e_max <= 1e-2 and e_l2 <= 1e-2, (e_max, e_l2)


def test_fused_rejects_what_it_cannot_do():
    from lvc_b200 import ClipBuilder

    m = _module()
    vid = [torch.zeros((4, 64, 64, 3), dtype=torch.uint8, device="cuda")]
    cb = ClipBuilder("resnet_infer", num_frames=4, layout="TOKENS", patch=(1, 16, 16))
    with pytest.raises(ValueError, match="patch"):
        m.forward_clips(cb, cb.prepare(vid, [[0, 1, 2, 3]]))
    cb = ClipBuilder("resnet_infer", num_frames=4, layout="BCTHW")
    with pytest.raises(ValueError, match="TOKENS"):
        m.forward_clips(cb, cb.prepare(vid, [[0, 1, 2, 3]]))
