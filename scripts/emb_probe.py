"""Times the fused-embeddings GEMM variants (profiling aid): python scripts/emb_probe.py"""
import json
import os

os.environ["LVC_PROFILING_KNOBS"] = "1"  # opt in to the LVC_GEMM_DIAG experiment knobs
import sys

import torch

torch.set_grad_enabled(False)  # forward-only kernels

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lvc_b200 import patch_embed as pe  # noqa: E402
from gemm_probe import timed  # noqa: E402

flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
me = pe.VivitEmbeddings().cuda()
tok = torch.randn(25088, 1536, device="cuda").to(torch.bfloat16)
for env in ({}, {"LVC_GEMM_DIAG": "1"}, {"LVC_GEMM_DIAG": "2"}, {"LVC_GEMM_DIAG": "4"}):
    for k in ("LVC_GEMM_DIAG",):
        os.environ.pop(k, None)
    os.environ.update(env)
    cold = timed(lambda: me(tokens=tok, batch=8), flush)
    plain = timed(lambda: me.patch_embeddings.forward_tokens(tok, 8, 32), flush)
    print(json.dumps({"env": env, "fused_us": round(cold, 1), "plain_us": round(plain, 1)}), flush=True)
