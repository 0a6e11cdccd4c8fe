"""Directed timing experiments on the CTA-pair GEMM (profiling aid, GPU box only):
python scripts/gemm_probe.py  -> one JSON line per (shape, knob) with CUDA-event medians."""
import json
import os

os.environ["LVC_PROFILING_KNOBS"] = "1"  # opt in to the LVC_GEMM_DIAG experiment knobs
import sys

import torch

torch.set_grad_enabled(False)  # forward-only kernels

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lvc_b200 import patch_embed as pe  # noqa: E402


def timed(fn, flush, iters=15):
    ts = []
    for _ in range(3 + iters):
        if flush is not None:
            flush.zero_(); flush.view(torch.int64).sum()
        torch.cuda._sleep(400000)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts = sorted(ts[3:])
    return ts[len(ts) // 2] * 1e3


def main():
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    shapes = {"vivit": (pe.VivitTubeletEmbeddings, 8, 32, 1536), "timesformer": (pe.TimesformerPatchEmbeddings, 16, 8, 768)}
    knobs = [{"LVC_GEMM_DIAG": v} for v in (sys.argv[1:] or ["0", "7", "8", "16", "23", "20", "18"])]
    for name, (cls, B, T, K) in shapes.items():
        m = cls().cuda()
        tok = torch.randn(25088, K, device="cuda").to(torch.bfloat16)
        for kn in knobs:
            for k in ("LVC_GEMM_DIAG", "LVC_GEMM_STAGES"):
                os.environ.pop(k, None)
            os.environ.update(kn)
            cold = timed(lambda: m.forward_tokens(tok, B, T), flush)
            warm = timed(lambda: m.forward_tokens(tok, B, T), None)
            print(json.dumps({"shape": name, "knobs": kn, "us_cold": round(cold, 2), "us_warm": round(warm, 2)}), flush=True)


if __name__ == "__main__":
    main()
