"""Directed timing experiments on the short-K GEMM + LayerNorm kernel (profiling aid, GPU box only)."""
import json
import os

os.environ["LVC_PROFILING_KNOBS"] = "1"  # opt in to the LVC_GEMM_DIAG experiment knobs
import sys
from functools import partial

import torch

torch.set_grad_enabled(False)  # forward-only kernels

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lvc_b200 import patch_embed as pe  # noqa: E402
from gemm_probe import timed  # noqa: E402

flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
m = pe.PatchEmbed3d([2, 4, 4], 3, 128, partial(torch.nn.LayerNorm, eps=1e-5)).cuda()
tok = torch.randn(401408, 96, device="cuda").to(torch.bfloat16)
for spec in (sys.argv[1:] or ["0", "1", "2", "4", "16", "6", "22", "23", "0:3", "0:4", "0:6"]):
    d, _, st = spec.partition(":")
    os.environ["LVC_GEMM_DIAG"] = d
    os.environ.pop("LVC_GEMM_STAGES", None)
    if st:
        os.environ["LVC_GEMM_STAGES"] = st
    cold = timed(lambda: m.forward_tokens(tok, 8, 32), flush)
    warm = timed(lambda: m.forward_tokens(tok, 8, 32), None)
    print(json.dumps({"diag": d, "stages": st, "us_cold": round(cold, 2), "us_warm": round(warm, 2)}), flush=True)
