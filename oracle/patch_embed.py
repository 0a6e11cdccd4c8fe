"""Oracle: fp32 patch / tubelet embeddings.  Test infrastructure.

Restated on torch CPU fp32 functional ops; pinned against the third-party modules the
reference instantiates (transformers VivitTubeletEmbeddings / TimesformerPatchEmbeddings,
torchvision PatchEmbed3d) by tests/golden/make_golden.py.
"""
import torch
import torch.nn.functional as F


def vivit_tubelet_embed(pixel_values, weight, bias):
    """a12 -- transformers modeling_vivit.py:59-77 (created at
    vivit_transformer/vivit_classifier/models/vivit_model.py:27-41):
    (B,T,C,H,W) -> permute(0,2,1,3,4) -> Conv3d(k=s=(2,16,16)) -> flatten(2).transpose(1,2)."""
    x = pixel_values.permute(0, 2, 1, 3, 4)
    x = F.conv3d(x, weight, bias, stride=tuple(weight.shape[2:]))
    return x.flatten(2).transpose(1, 2)


def timesformer_patch_embed(pixel_values, weight, bias):
    """a13 -- transformers modeling_timesformer.py:55-64 (created at
    timesformer/timesformer_classifier/models/timesformer_model.py:27-42):
    (B,T,C,H,W) -> (B*T,C,H,W) -> Conv2d(k=s=16) -> flatten(2).transpose(1,2);
    returns (embeddings, num_frames, patch_width)."""
    B, T, C, H, W = pixel_values.shape
    x = pixel_values.reshape(B * T, C, H, W)
    x = F.conv2d(x, weight, bias, stride=tuple(weight.shape[2:]))
    pw = x.size(-1)
    return x.flatten(2).transpose(1, 2), T, pw


def swin_patch_embed(x, weight, bias, ln_weight, ln_bias, eps=1e-5):
    """a14 -- torchvision models/video/swin_transformer.py:353-363 (created at
    videoswintransformer/swin_video_classifier/models/swin3d.py:22-44):
    (B,C,T,H,W) -> pad to multiples of (2,4,4) -> Conv3d(k=s=(2,4,4)) -> permute to
    (B,T',H',W',E) -> LayerNorm(E, eps=1e-5)."""
    kt, kh, kw = weight.shape[2:]
    _, _, t, h, w = x.shape
    pad = ((kw - w % kw) % kw, (kh - h % kh) % kh, (kt - t % kt) % kt)
    x = F.pad(x, (0, pad[0], 0, pad[1], 0, pad[2]))
    x = F.conv3d(x, weight, bias, stride=(kt, kh, kw))
    x = x.permute(0, 2, 3, 4, 1)
    if ln_weight is not None:
        x = F.layer_norm(x, (weight.shape[0],), ln_weight, ln_bias, eps)
    return x


def rel_err(y, ref):
    """The '1e-2 relative' metric of the north star: max-abs error over
    max-abs reference, and relative L2 error."""
    y = y.double()
    ref = ref.double()
    return (
        float((y - ref).abs().max() / ref.abs().max()),
        float((y - ref).norm() / ref.norm()),
    )
