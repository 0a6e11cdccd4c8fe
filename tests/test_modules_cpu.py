"""Host-side checks that need no GPU: the C-ABI library loads and exports every declared symbol, the
drop-in modules keep the reference modules' parameter names and shapes, and device entry points fail
loudly (no CPU fallback)."""
import ctypes as C
import os
import re
from functools import partial

import pytest
import torch

from lvc_b200 import _lib


def test_library_exports_every_declared_symbol():
    lib = _lib.lib()
    hdr = open(os.path.join(os.path.dirname(_lib.LIB_PATH), "..", "include", "lvc_b200.h")).read()
    declared = set(re.findall(r"\b(lvc_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_lib.declared_symbols())
    for s in declared:
        assert hasattr(lib, s), s
    assert lib.lvc_version() >= 100


def test_state_dict_names_match_reference_modules():
    from torchvision.models.video.swin_transformer import PatchEmbed3d as TvPatchEmbed3d
    from transformers import TimesformerConfig, VivitConfig
    from transformers.models.timesformer.modeling_timesformer import TimesformerPatchEmbeddings as HfTsf
    from transformers.models.vivit.modeling_vivit import VivitTubeletEmbeddings as HfVivit

    from lvc_b200 import patch_embed as pe

    pairs = [(pe.VivitTubeletEmbeddings(VivitConfig()), HfVivit(VivitConfig())),
             (pe.TimesformerPatchEmbeddings(TimesformerConfig()), HfTsf(TimesformerConfig())),
             (pe.PatchEmbed3d([2, 4, 4], 3, 128, partial(torch.nn.LayerNorm, eps=1e-5)),
              TvPatchEmbed3d([2, 4, 4], 3, 128, partial(torch.nn.LayerNorm, eps=1e-5)))]
    for mine, ref in pairs:
        a, b = mine.state_dict(), ref.state_dict()
        assert {k: tuple(v.shape) for k, v in a.items()} == {k: tuple(v.shape) for k, v in b.items()}
        mine.load_state_dict(b)  # reference checkpoints load unchanged
    assert pairs[0][0].num_patches == pairs[0][1].num_patches == 3136
    assert pairs[1][0].num_patches == pairs[1][1].num_patches == 196


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    from lvc_b200 import ClipBuilder
    from lvc_b200 import patch_embed as pe

    with pytest.raises(RuntimeError):
        ClipBuilder("resnet_infer")
    with pytest.raises(RuntimeError):
        pe.VivitTubeletEmbeddings()(torch.zeros(1, 32, 3, 224, 224))
    d = _lib.ClipDesc(0, 0, 0, 224, 1, 1, 8, 0, 0, 0, (C.c_float * 3)(), (C.c_float * 3)())
    rc = _lib.lib().lvc_clip_forward(C.byref(d), 1, 1, None, 1, 1, None, None)
    assert rc == _lib.E_CUDA
    with pytest.raises(RuntimeError):
        _lib.check(rc, "lvc_clip_forward")


def test_argument_validation_without_gpu():
    lib = _lib.lib()
    d = _lib.ClipDesc(0, 0, 0, 200, 1, 1, 8, 0, 0, 0, (C.c_float * 3)(), (C.c_float * 3)())
    assert lib.lvc_clip_forward(C.byref(d), 1, 1, None, 1, 1, None, None) == _lib.E_SHAPE
    assert lib.lvc_clip_forward(None, 1, 1, None, 1, 1, None, None) == _lib.E_INVAL
    e = _lib.EmbedDesc(_lib.EMBED_VIVIT, 0, 0, 1, 8, 32, 224, 224, 2, 16, 16, 768, 0.0, 0)
    assert lib.lvc_patch_embed_workspace(C.byref(e)) == 25088 * 1536 * 2
    e.embed_dim = 100
    assert lib.lvc_patch_embed_workspace(C.byref(e)) == -1
