"""Times the Video Swin front end (BASELINE config 4) three ways with CUDA events, L2 flushed: clip kernel alone, clip + embed as
two launches, and the co-resident pair (PatchEmbed3d.forward_clips): python scripts/perf_swin_fused.py"""
import os
import statistics
import sys
from functools import partial

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from lvc_b200 import patch_embed as pe  # noqa: E402
from scripts_common import setup_clip  # noqa: E402

torch.set_grad_enabled(False)
cb, batch, tok, vids = setup_clip("swin_cfg4")
m = pe.PatchEmbed3d([2, 4, 4], 3, 128, partial(torch.nn.LayerNorm, eps=1e-5)).cuda()
B, T = 8, 32
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
out = torch.empty((B, T // 2, 56, 56, 128), dtype=torch.bfloat16, device="cuda")


def timed(fn, n=15):
    ts = []
    for _ in range(n + 3):
        bench.l2_flush(flush)
        torch.cuda._sleep(400000)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return statistics.median(ts[3:])


print(f"clip alone            {timed(lambda: cb.run(batch, tok)):8.1f} us")
print(f"embed alone           {timed(lambda: m.forward_tokens(tok, B, T)):8.1f} us")
print(f"clip + embed (2 launches, PDL) {timed(lambda: m.forward_tokens(cb.run(batch, tok), B, T)):8.1f} us")
print(f"co-resident pair      {timed(lambda: m.forward_clips(cb, batch, out=out)):8.1f} us")
want = m.forward_tokens(cb.run(batch, tok), B, T).view(out.shape)
got = m.forward_clips(cb, batch, out=out)
torch.cuda.synchronize()
print("bit-identical:", torch.equal(want.view(torch.int16), got.view(torch.int16)))
tiles = (B * (T // 2) * 56 * 56 + 127) // 128
st = m._fused_ws[-256:].view(torch.int64)[:4].cpu().tolist()
t0 = min(st[0], st[2])
print(f"timeline (us, CTA 0 of each kernel): clip {(st[0] - t0) / 1e3:.1f} .. {(st[1] - t0) / 1e3:.1f}   embed {(st[2] - t0) / 1e3:.1f} .. {(st[3] - t0) / 1e3:.1f}")
