"""Per-stage timeline of CTA 0 of the clip kernel (profiling aid; needs the knob gate):
python scripts/clip_trace.py swin_cfg4 [KNOB=VALUE ...]
For every ring stage: when its producer issued the copies, when consumer warp 0 started waiting for it, when it had it, when it
was done with it.  Prints how the CTA's time splits into consumer work, consumer starvation and per-item gaps."""
import ctypes as C
import os
import sys

_here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ.setdefault("LVC_B200_LIB", os.path.join(_here, "ai-laryngeal-video-based-classifier_b200", "liblvc_b200_trace.so"))  # `make -C csrc trace`
os.environ["LVC_PROFILING_KNOBS"] = "1"
os.environ["LVC_CLIP_TRACE"] = "1"
for a in sys.argv[2:]:
    k, _, v = a.partition("=")
    os.environ[k] = v

import numpy as np  # noqa: E402
import torch  # noqa: E402

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lvc_b200 import _lib  # noqa: E402
from scripts_common import setup_clip  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "swin_cfg4"
cb, batch, tok, vids = setup_clip(name)
for _ in range(3):
    cb.run(batch, tok)
torch.cuda.synchronize()
n = 4096
buf = (C.c_ulonglong * (4 * n + 2048))()
fn = _lib.lib().lvc_clip_trace_read
fn.restype, fn.argtypes = C.c_int, [C.c_void_p, C.c_int]
assert fn(buf, n) == 0
raw = np.frombuffer(buf, dtype=np.uint64).astype(np.int64)
cta = raw[4 * n:].reshape(1024, 2)
cta = cta[(cta > 0).all(axis=1)]
t = raw[:4 * n].reshape(n, 4)
t = t[(t > 0).all(axis=1)]
if len(cta):
    s0 = cta[:, 0].min()
    dur = cta[:, 1] - cta[:, 0]
    print(f"{name}: {len(cta)} CTAs; start spread {(cta[:, 0].max() - s0) / 1e3:.1f} us; CTA duration min {dur.min() / 1e3:.1f} median "
          f"{np.median(dur) / 1e3:.1f} max {dur.max() / 1e3:.1f} us; first start -> last end {(cta[:, 1].max() - s0) / 1e3:.1f} us; "
          f"CTA 0: {dur[0] / 1e3:.1f} us" if len(dur) else "")
t0 = t[:, 0].min()
t = t - t0
issue, wait0, have, done = t[:, 0], t[:, 1], t[:, 2], t[:, 3]
total = done.max() - min(issue.min(), wait0.min())
print(f"{name}: {len(t)} stages on CTA 0, {total / 1e3:.1f} us from first issue to last stage done")
print(f"  consumer warp 0: working {np.sum(done - have) / 1e3:.1f} us, waiting for data {np.sum(have - wait0) / 1e3:.1f} us, "
      f"between stages {np.sum(wait0[1:] - done[:-1]) / 1e3:.1f} us")
lead = have - issue
print(f"  issue -> consumed-by-warp-0 lead: median {np.median(lead) / 1e3:.2f} us, min {lead.min() / 1e3:.2f}, max {lead.max() / 1e3:.2f}")
print(f"  per stage: work median {np.median(done - have):.0f} ns, p90 {np.percentile(done - have, 90):.0f}; wait median "
      f"{np.median(have - wait0):.0f} ns, p90 {np.percentile(have - wait0, 90):.0f}, max {np.max(have - wait0)}")
gaps = issue[1:] - issue[:-1]
print(f"  producer issue spacing: median {np.median(gaps):.0f} ns, p90 {np.percentile(gaps, 90):.0f}, max {gaps.max()}")
w = have - wait0
big = np.argsort(-w)[:12]
print("  longest waits (stage, wait ns, issued-before-wait-start ns):", [(int(i), int(w[i]), int(wait0[i] - issue[i])) for i in sorted(big)])
if os.environ.get("CLIP_TRACE_STAGES"):
    print("  first 24 stages [issue, wait0, have, done] (ns):")
    for i in range(min(24, len(t))):
        print("   ", i, t[i].tolist())
