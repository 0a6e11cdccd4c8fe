"""Host-inclusive throughput of the clip -> embed pipeline for the small configs (python overhead included):
python scripts/host_inclusive.py"""
import json
import os
import sys
import time

import numpy as np
import torch

torch.set_grad_enabled(False)  # forward-only kernels

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from lvc_b200 import ClipBuilder, sampling  # noqa: E402
from lvc_b200 import patch_embed as pe  # noqa: E402

dev = torch.device("cuda", 0)
for name in ("vivit_cfg2", "timesformer_cfg3"):
    c = bench.SECONDARY[name]
    g = torch.Generator(device=dev).manual_seed(7)
    vids = [torch.randint(0, 256, (c["frames"], c["h"], c["w"], 3), dtype=torch.uint8, device=dev, generator=g) for _ in range(c["videos"])]
    rng = sampling.SamplerRng(42)
    cb = ClipBuilder(c["preset"], num_frames=c["T"], layout="TOKENS", patch=c["patch"])
    m = (pe.VivitTubeletEmbeddings() if c["embed"] == "vivit" else pe.TimesformerPatchEmbeddings()).to(dev)

    def step():
        idx = np.array([sampling.get_sampling_indices(c["sampling"], c["frames"], c["T"], rng) for _ in vids], np.int32)
        tok = cb(vids, idx)
        return m.forward_tokens(tok, c["videos"], c["T"])

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    n = 200
    t0 = time.perf_counter()
    for _ in range(n):
        step()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / n
    res = {"config": name, "host_inclusive_ms_per_batch": round(dt * 1e3, 4), "clips_per_s": round(c["videos"] / dt)}
    if hasattr(sys.modules.get("lvc_b200.pipeline"), "ClipEmbedPipeline") or os.path.exists(os.path.join(os.path.dirname(pe.__file__), "pipeline.py")):
        from lvc_b200.pipeline import ClipEmbedPipeline

        pipe = ClipEmbedPipeline(cb, m, vids, c["T"])
        for _ in range(5):
            pipe(np.array([sampling.get_sampling_indices(c["sampling"], c["frames"], c["T"], rng) for _ in vids], np.int32))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n):
            pipe(np.array([sampling.get_sampling_indices(c["sampling"], c["frames"], c["T"], rng) for _ in vids], np.int32))
        torch.cuda.synchronize()
        dg = (time.perf_counter() - t0) / n
        res["graph_ms_per_batch"] = round(dg * 1e3, 4)
        res["graph_clips_per_s"] = round(c["videos"] / dg)
    print(json.dumps(res), flush=True)
