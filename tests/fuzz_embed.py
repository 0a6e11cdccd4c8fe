"""Randomised parity sweep of the patch-embedding kernels (ragged token counts, both output dtypes, both input
layouts) against fp32 torch on the CPU.  python tests/fuzz_embed.py [cases] [seed]"""
import os
import sys
from functools import partial

import numpy as np
import torch

torch.set_grad_enabled(False)  # forward-only kernels
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lvc_b200 import patch_embed as pe  # noqa: E402


def rel(y, ref):
    y, ref = y.float().cpu(), ref.float()
    return float((y - ref).abs().max() / ref.abs().max()), float((y - ref).norm() / ref.norm())


def main():
    cases = int(sys.argv[1]) if len(sys.argv) > 1 else 30
    g = torch.Generator().manual_seed(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
    ri = lambda lo, hi: int(torch.randint(lo, hi, (1,), generator=g))  # noqa: E731
    bad = 0
    for k in range(cases):
        kind = ["vivit", "timesformer", "swin", "vivit_emb", "timesformer_emb"][ri(0, 5)]
        od = [torch.bfloat16, torch.float32][ri(0, 2)]
        B = ri(1, 4)
        if kind.startswith("vivit"):
            T = 2 * ri(1, 5)
            x = torch.randn(B, T, 3, 224, 224, generator=g)
            m = (pe.VivitEmbeddings if kind.endswith("emb") else pe.VivitTubeletEmbeddings)(num_frames=T, out_dtype=od)
        elif kind.startswith("timesformer"):
            T = ri(1, 9)
            x = torch.randn(B, T, 3, 224, 224, generator=g)
            m = pe.TimesformerEmbeddings(num_frames=T, out_dtype=od) if kind.endswith("emb") else pe.TimesformerPatchEmbeddings(out_dtype=od)
        else:
            T = 2 * ri(1, 4)
            x = torch.randn(B, 3, T, 224, 224, generator=g)
            m = pe.PatchEmbed3d([2, 4, 4], 3, 128, partial(torch.nn.LayerNorm, eps=1e-5), out_dtype=od)
        for p_ in m.parameters():
            p_.data.copy_(torch.randn(p_.shape, generator=g) * (0.02 if p_.dim() > 1 else 0.2))
        if kind == "swin":
            m.norm.weight.data.add_(1.0)
        mc = m.cuda()
        y = mc(x.cuda())
        y = y[0] if isinstance(y, tuple) else y
        with torch.no_grad():
            if kind.startswith("vivit"):
                pm = m.patch_embeddings if kind.endswith("emb") else m
                ref = F.conv3d(x.permute(0, 2, 1, 3, 4), pm.projection.weight.cpu(), pm.projection.bias.cpu(), stride=(2, 16, 16)).flatten(2).transpose(1, 2)
                if kind.endswith("emb"):
                    ref = torch.cat((m.cls_token.cpu().expand(B, -1, -1), ref), 1) + m.position_embeddings.cpu()
            elif kind.startswith("timesformer"):
                pm = m.patch_embeddings if kind.endswith("emb") else m
                ref = F.conv2d(x.reshape(B * T, 3, 224, 224), pm.projection.weight.cpu(), pm.projection.bias.cpu(), stride=16).flatten(2).transpose(1, 2)
                if kind.endswith("emb"):
                    e = torch.cat((m.cls_token.cpu().expand(B * T, -1, -1), ref), 1) + m.position_embeddings.cpu()
                    cls = e[:B, 0:1]
                    e = e[:, 1:].reshape(B, T, 196, -1).permute(0, 2, 1, 3).reshape(B * 196, T, -1) + m.time_embeddings.cpu()
                    ref = torch.cat((cls, e.reshape(B, 196 * T, -1)), 1)
            else:
                c = F.conv3d(x, m.proj.weight.cpu(), m.proj.bias.cpu(), stride=(2, 4, 4)).permute(0, 2, 3, 4, 1)
                ref = F.layer_norm(c, (128,), m.norm.weight.cpu(), m.norm.bias.cpu(), 1e-5)
        e1, e2 = rel(y, ref)
        ok = tuple(y.shape) == tuple(ref.shape) and e1 <= 1e-2 and e2 <= 1e-2
        if not ok:
            bad += 1
            print("MISMATCH", k, kind, od, "B", B, "T", T, tuple(y.shape), tuple(ref.shape), e1, e2, flush=True)
        del mc, y
    print(f"fuzz_embed: {cases - bad}/{cases} cases match", flush=True)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
