"""lvc_patch_embed_static (no allocation, scratch in a file-scope device array) against lvc_patch_embed (caller
workspace), both through the C ABI.  Same kernels on the same inputs, so the bar is byte equality; the oracle check
beside it uses the embed tolerance of tests/test_embed_gpu.py (1e-2 relative).

Collected last on purpose (file name).  Passes alone in a fresh process and after
the other tests (it shares no state with them except the library's one scratch array, which it rewrites)."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import patch_embed as ope

pytestmark = pytest.mark.gpu
TOL = 1e-2


def _rand(shape, seed, scale=1.0):
    return torch.from_numpy((np.random.default_rng(seed).standard_normal(shape) * scale).astype(np.float32))


def _setup(kind, B, T, H, W, k, E, x_layout, seed):
    from lvc_b200 import _lib

    lib = _lib.lib()
    d = _lib.EmbedDesc(kind, _lib.DT_F32, x_layout, _lib.DT_BF16, B, T, H, W, k[0], k[1], k[2], E, 1e-5, 0)
    xshape = (B, T, 3, H, W) if x_layout == _lib.LAYOUT_BTCHW else (B, 3, T, H, W)
    x = _rand(xshape, seed).cuda()
    wshape = (E, 3, k[1], k[2]) if kind == _lib.EMBED_TIMESFORMER else (E, 3, k[0], k[1], k[2])
    w = _rand(wshape, seed + 1, 0.02).cuda()
    bias = _rand((E,), seed + 2, 0.02).cuda()
    ln_w = (1 + _rand((E,), seed + 3, 0.1)).cuda()
    ln_b = _rand((E,), seed + 4, 0.1).cuda()
    packed = torch.empty(w.numel(), dtype=torch.bfloat16, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.lvc_patch_embed_pack_weights(C.byref(d), w.data_ptr(), packed.data_ptr(), st), "pack")
    M = B * (T // k[0]) * (H // k[1]) * (W // k[2])
    return lib, d, x, w, bias, ln_w, ln_b, packed, M, st


def _both(lib, d, x, bias, ln_w, ln_b, packed, M, E, st, swin):
    from lvc_b200 import _lib

    need = lib.lvc_patch_embed_workspace(C.byref(d))
    assert 0 < need <= lib.lvc_patch_embed_static_capacity()
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    want = torch.empty((M, E), dtype=torch.bfloat16, device="cuda")
    got = torch.full((M, E), float("nan"), dtype=torch.bfloat16, device="cuda")
    lw, lb = (ln_w.data_ptr(), ln_b.data_ptr()) if swin else (None, None)
    _lib.check(lib.lvc_patch_embed(C.byref(d), x.data_ptr(), packed.data_ptr(), bias.data_ptr(), lw, lb,
                                   want.data_ptr(), ws.data_ptr(), st), "lvc_patch_embed")
    _lib.check(lib.lvc_patch_embed_static(C.byref(d), x.data_ptr(), packed.data_ptr(), bias.data_ptr(), lw, lb,
                                          got.data_ptr(), st), "lvc_patch_embed_static")
    torch.cuda.synchronize()
    return want, got


def test_static_scratch_matches_workspace_entry_point_two_calls_in_a_row():
    from lvc_b200 import _lib

    # call 1: ViViT tubelets (CTA-pair tcgen05 kernel), one 8-frame 224x224 clip: A = 784 x 1536 bf16 = 2.4 MB
    lib, d, x, w, bias, ln_w, ln_b, packed, M, st = _setup(_lib.EMBED_VIVIT, 1, 8, 224, 224, (2, 16, 16), 768,
                                                           _lib.LAYOUT_BTCHW, 11)
    want, got = _both(lib, d, x, bias, ln_w, ln_b, packed, M, 768, st, swin=False)
    assert torch.equal(want.view(torch.int16), got.view(torch.int16))
    ref = ope.vivit_tubelet_embed(x.cpu(), w.cpu(), bias.cpu()).reshape(M, 768)
    e_max, e_l2 = ope.rel_err(got.float().cpu(), ref)
    assert e_max <= TOL and e_l2 <= TOL, (e_max, e_l2)

    # call 2, straight after, on the SAME array holding call 1's patch matrix: a different shape and kernel
    # (Video Swin short-K GEMM + LayerNorm), then call 1's shape again with new data
    lib, d2, x2, w2, b2, lw2, lb2, p2, M2, st = _setup(_lib.EMBED_SWIN, 2, 4, 96, 128, (2, 4, 4), 128,
                                                       _lib.LAYOUT_BCTHW, 23)
    want2, got2 = _both(lib, d2, x2, b2, lw2, lb2, p2, M2, 128, st, swin=True)
    assert torch.equal(want2.view(torch.int16), got2.view(torch.int16))
    ref2 = ope.swin_patch_embed(x2.cpu(), w2.cpu(), b2.cpu(), lw2.cpu(), lb2.cpu()).reshape(M2, 128)
    e_max, e_l2 = ope.rel_err(got2.float().cpu(), ref2)
    assert e_max <= TOL and e_l2 <= TOL, (e_max, e_l2)

    x.copy_(_rand(tuple(x.shape), 99).cuda())
    want3, got3 = _both(lib, d, x, bias, ln_w, ln_b, packed, M, 768, st, swin=False)
    assert torch.equal(want3.view(torch.int16), got3.view(torch.int16))
    assert not torch.equal(want3.view(torch.int16), want.view(torch.int16))


def test_static_scratch_allocates_nothing_and_rejects_oversize():
    from lvc_b200 import _lib

    lib, d, x, w, bias, ln_w, ln_b, packed, M, st = _setup(_lib.EMBED_TIMESFORMER, 2, 1, 224, 224, (1, 16, 16), 768,
                                                           _lib.LAYOUT_BTCHW, 5)
    out = torch.empty((M, 768), dtype=torch.bfloat16, device="cuda")
    args = (C.byref(d), x.data_ptr(), packed.data_ptr(), bias.data_ptr(), None, None, out.data_ptr(), st)
    _lib.check(lib.lvc_patch_embed_static(*args), "warm")    # first call on the device resolves the symbol address
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    reserved0 = torch.cuda.memory_reserved()
    for _ in range(3):
        _lib.check(lib.lvc_patch_embed_static(*args), "lvc_patch_embed_static")
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free0 and torch.cuda.memory_reserved() == reserved0

    # cfg 2 (8 clips x 32 frames): 77 MB of patch matrix > the 16 MiB array -> LVC_E_SHAPE and a message, no launch
    big = _lib.EmbedDesc(_lib.EMBED_VIVIT, _lib.DT_F32, _lib.LAYOUT_BTCHW, _lib.DT_BF16, 8, 32, 224, 224, 2, 16, 16, 768, 0.0, 0)
    assert lib.lvc_patch_embed_workspace(C.byref(big)) > lib.lvc_patch_embed_static_capacity() == 16 * 1024 * 1024
    rc = lib.lvc_patch_embed_static(C.byref(big), x.data_ptr(), packed.data_ptr(), bias.data_ptr(), None, None,
                                    out.data_ptr(), st)
    assert rc == _lib.E_SHAPE
    msg = lib.lvc_last_error().decode()
    assert "LVC_STATIC_SCRATCH_BYTES" in msg and "lvc_patch_embed" in msg
    with pytest.raises(ValueError):
        _lib.check(rc, "lvc_patch_embed_static")
    assert lib.lvc_patch_embed_static(None, None, None, None, None, None, None, None) == _lib.E_INVAL


def test_static_scratch_under_stream_capture():
    """What the entry point is for: inside a CUDA graph capture nothing may allocate.  One warm call outside the capture
    (resolves the array's address, loads the kernels), then the call is captured and the graph replayed twice on new inputs."""
    from lvc_b200 import _lib

    lib, d, x, w, bias, ln_w, ln_b, packed, M, st = _setup(_lib.EMBED_SWIN, 1, 4, 224, 224, (2, 4, 4), 128,
                                                           _lib.LAYOUT_BCTHW, 41)
    need = lib.lvc_patch_embed_workspace(C.byref(d))
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    want = torch.empty((M, 128), dtype=torch.bfloat16, device="cuda")
    got = torch.zeros((M, 128), dtype=torch.bfloat16, device="cuda")
    args = (C.byref(d), x.data_ptr(), packed.data_ptr(), bias.data_ptr(), ln_w.data_ptr(), ln_b.data_ptr(), got.data_ptr())
    _lib.check(lib.lvc_patch_embed_static(*args, st), "warm")
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            _lib.check(lib.lvc_patch_embed_static(*args, torch.cuda.current_stream().cuda_stream), "captured")
    torch.cuda.current_stream().wait_stream(side)
    for seed in (7, 8):
        x.copy_(_rand(tuple(x.shape), seed).cuda())
        got.zero_()
        g.replay()
        _lib.check(lib.lvc_patch_embed(C.byref(d), x.data_ptr(), packed.data_ptr(), bias.data_ptr(), ln_w.data_ptr(), ln_b.data_ptr(),
                                       want.data_ptr(), ws.data_ptr(), torch.cuda.current_stream().cuda_stream), "lvc_patch_embed")
        torch.cuda.synchronize()
        assert torch.equal(want.view(torch.int16), got.view(torch.int16)), seed
