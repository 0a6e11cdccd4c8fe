"""Times the patch-embedding path (im2col + tcgen05 GEMM) of BASELINE configs 2-4 with CUDA events.
Usage (GPU box): python scripts/perf_embed.py [vivit|timesformer|swin ...]"""
import json
import os
import sys
from functools import partial

import torch

torch.set_grad_enabled(False)  # forward-only kernels

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lvc_b200 import patch_embed as pe  # noqa: E402


def run(name, iters=20):
    dt = torch.bfloat16 if os.environ.get("X_BF16") else torch.float32
    if name == "vivit":
        m, x = pe.VivitTubeletEmbeddings().cuda(), torch.randn(8, 32, 3, 224, 224, device="cuda", dtype=dt)
        M, K, N = 25088, 1536, 768
    elif name == "timesformer":
        m, x = pe.TimesformerPatchEmbeddings().cuda(), torch.randn(16, 8, 3, 224, 224, device="cuda", dtype=dt)
        M, K, N = 25088, 768, 768
    else:
        m = pe.PatchEmbed3d([2, 4, 4], 3, 128, partial(torch.nn.LayerNorm, eps=1e-5)).cuda()
        x = torch.randn(8, 3, 32, 224, 224, device="cuda", dtype=dt)
        M, K, N = 401408, 96, 128
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        m(x)
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_(); flush.view(torch.int64).sum()  # > L2
        torch.cuda._sleep(400000)  # host call fully enqueued before e0 fires
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        m(x)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    ms = ts[len(ts) // 2]
    print(json.dumps({"embed": name, "x_dtype": str(dt), "ms_module_median": ms, "ms_min": ts[0],
                      "gflop": 2 * M * K * N / 1e9, "tflops_module": 2 * M * K * N / ms / 1e9}))


if __name__ == "__main__":
    for n in (sys.argv[1:] or ["vivit", "timesformer", "swin"]):
        run(n)
