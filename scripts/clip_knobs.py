"""Times the clip stage of the secondary configs under the launch knobs of lvc_clip.cu (profiling aid):
python scripts/clip_knobs.py [cfg ...]   -- CUDA events, L2 flushed before every launch, median of 15."""
import os
import sys

import numpy as np

os.environ["LVC_PROFILING_KNOBS"] = "1"  # opt in to the launch knobs (read once per process, before the library loads)
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from lvc_b200 import ClipBuilder, sampling  # noqa: E402
from lvc_b200.synth import synth_boxes  # noqa: E402

dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def setup(name):
    c = bench.SECONDARY[name]
    g = torch.Generator(device=dev).manual_seed(7)
    vids = [torch.randint(0, 256, (c["frames"], c["h"], c["w"], 3), dtype=torch.uint8, device=dev, generator=g) for _ in range(c["videos"])]
    rng = sampling.SamplerRng(42)
    idx = np.array([sampling.get_sampling_indices(c["sampling"], c["frames"], c["T"], rng) for _ in vids], np.int32)
    boxes = None
    if c.get("boxes"):
        boxes = np.stack([synth_boxes(2000 + v, c["frames"], c["h"], c["w"])[idx[v]] for v in range(len(vids))])
    cb = ClipBuilder(c["preset"], num_frames=c["T"], layout="TOKENS", patch=c["patch"])
    batch = cb.prepare(vids, idx, boxes)
    tok = torch.empty(batch.shape, dtype=torch.bfloat16, device=dev)
    return cb, batch, tok, vids


def timeit(cb, batch, tok, n=15):
    ts = []
    for _ in range(n + 3):
        flush.zero_(); flush.view(torch.int64).sum()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        cb.run(batch, tok)
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return float(np.median(ts[3:]))


KNOBS = [{}, {"LVC_CLIP_SMEM_KB": "227"}, {"LVC_CLIP_RPS": "16"}, {"LVC_CLIP_RPS": "4"}, {"LVC_CLIP_RPS": "2"}, {"LVC_CLIP_SMEM_KB": "100"}, {"LVC_CLIP_BAND": "56"},
         {"LVC_CLIP_BAND": "16"}, {"LVC_CLIP_BAND": "224"}]
if os.environ.get("CLIP_KNOBS_AB"):   # the consumer variants of round 2
    KNOBS = [{}, {"LVC_CLIP_NO_FMA": "1"}]
if os.environ.get("CLIP_KNOBS_MATRIX"):   # producers x rows-per-stage
    KNOBS = [{"LVC_CLIP_NP": str(n), **({"LVC_CLIP_RPS": r} if r else {})} for n in (1, 2, 4) for r in ("", "4")]
if os.environ.get("CLIP_KNOBS_JSON"):   # explicit list of knob dicts, e.g. '[{}, {"LVC_CLIP_SMEM_KB": "110"}]'
    import json
    KNOBS = json.loads(os.environ["CLIP_KNOBS_JSON"])
for name in (sys.argv[1:] or ["swin_cfg4", "timesformer_cfg3", "vivit_cfg2"]):
    cb, batch, tok, vids = setup(name)
    for k in KNOBS:
        for kk in ("LVC_CLIP_RPS", "LVC_CLIP_SMEM_KB", "LVC_CLIP_BAND", "LVC_CLIP_NP", "LVC_CLIP_NO_FMA", "LVC_CLIP_NO_TOK44"):
            os.environ.pop(kk, None)
        os.environ.update(k)
        try:
            print(f"{name:18s} {str(k):72s} {timeit(cb, batch, tok):8.1f} us", flush=True)
        except Exception as e:  # a knob the shape does not admit
            print(f"{name:18s} {str(k):72s} failed: {e}", flush=True)
    del vids
