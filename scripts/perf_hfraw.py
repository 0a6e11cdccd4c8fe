"""Timing of the raw-frame HF processor kernel (lvc_hf_clip_forward) on the device."""
import sys
import torch
sys.path.insert(0, ".")
from lvc_b200 import ClipBuilder

def run(preset, shape, B=8, T=32, n=40, iters=20):
    vids = [torch.randint(0, 256, (n,) + shape + (3,), dtype=torch.uint8, device="cuda") for _ in range(B)]
    idx = torch.randint(0, n, (B, T)).numpy()
    cb = ClipBuilder(preset, num_frames=T)
    batch = cb.prepare(vids, idx)
    out = cb.run(batch)
    for _ in range(3):
        cb.run(batch, out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        cb.run(batch, out)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    E = cb.shortest_edge
    h, w = shape
    side = 224 / E * min(h, w)
    alg = B * T * (side * side * 3 + 224 * 224 * 3 * 4)
    print(f"{preset:16s} {h}x{w} B={B} T={T}: {ms:.3f} ms  {B / ms * 1e3:.0f} clips/s  {alg / ms / 1e6:.0f} GB/s algorithmic")

def run_host(preset, shape, B=8, T=32, n=36, iters=5):
    """frames in pinned host memory, read by the kernel over PCIe (zero-copy) vs upload-then-run"""
    vids = [torch.randint(0, 256, (n,) + shape + (3,), dtype=torch.uint8).pin_memory() for _ in range(B)]
    idx = torch.stack([torch.randperm(n)[:T].sort().values for _ in range(B)]).numpy()
    cb = ClipBuilder(preset, num_frames=T)
    out = torch.empty((B, T, 3, 224, 224), device="cuda")
    for zc in (True, False):
        for _ in range(2):
            cb.build_from_host(vids, idx, out=out, zero_copy=zc)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            cb.build_from_host(vids, idx, out=out, zero_copy=zc)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        print(f"{preset:16s} {shape[0]}x{shape[1]} host frames, zero_copy={zc}: {ms:.2f} ms  {B / ms * 1e3:.0f} clips/s  "
              f"{cb.last_h2d_bytes / ms / 1e6:.1f} GB/s over PCIe ({cb.last_h2d_bytes / 1e6:.0f} MB)")


if len(sys.argv) > 1 and sys.argv[1] == "host":
    run_host("vivit_raw", (1080, 1920))
    run_host("vivit_raw", (720, 1280))
elif len(sys.argv) > 1:   # one configuration (for ncu): preset H W
    run(sys.argv[1], (int(sys.argv[2]), int(sys.argv[3])), iters=2)
else:
    for preset in ("vivit_raw", "timesformer_raw"):
        for shape in ((1080, 1920), (720, 1280), (480, 640), (256, 340)):
            run(preset, shape)
