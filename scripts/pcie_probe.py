"""Host -> device ingest rate of the box, per GPU and aggregate, at 1 / 2 / 4 / 8 GPUs pulling CONCURRENTLY:

  memcpy   plain pinned cudaMemcpyAsync (the ceiling the runtime itself reaches)
  kernel   the fused clip kernel reading pinned frames in place (bench.py's e2e leg: tapped rows only)
each from ordinary pinned memory and from write-combined pinned memory (device reads are not snooped).

  python scripts/pcie_probe.py                                  one GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
      scripts/pcie_probe.py                                     ranks 0..c-1 pull together for c in 1, 2, 4, 8 <= N

Prints one JSON line (rank 0): per concurrency level the per-GPU GB/s and their sum.  Used by DESIGN.md section 5 to
tell a kernel-side limit from a box-side limit of the end-to-end scaling curve.
"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lvc_b200 import ClipBuilder, hostmem, sampling  # noqa: E402
from lvc_b200.clips import device_bytes_fetched  # noqa: E402


def main():
    rank, local, world = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("LOCAL_RANK", 0), ("WORLD_SIZE", 1)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = hostmem.bind_rank_to_gpu(local)
    if world > 1:
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    N, H, W, T, B = 48, 1080, 1920, 32, 16
    n = 1 << 30
    dst = torch.empty(n, dtype=torch.uint8, device=dev)
    src_dev = torch.randint(0, 256, (N, H, W, 3), dtype=torch.uint8, device=dev)
    idx = np.array([sampling.get_sampling_indices("uniform", N, T) for _ in range(B)], dtype=np.int32)
    cb = ClipBuilder("resnet_infer", num_frames=T, layout="BTCHW")
    out = torch.empty((B, T, 3, 224, 224), dtype=torch.float32, device=dev)
    res = {}
    for wc in (False, True):
        flat = hostmem.pinned_empty((n,), write_combined=wc)
        vids = [hostmem.pinned_empty((N, H, W, 3), write_combined=wc) for _ in range(4)]
        for v in vids:
            v.copy_(src_dev)
        torch.cuda.synchronize()
        host = [vids[i % 4] for i in range(B)]
        levels = [c for c in (1, 2, 4, 8) if c <= world]
        for c in levels:
            active = rank < c
            for what in ("memcpy", "kernel"):
                barrier()
                gbs = 0.0
                if active:
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    reps = 4
                    if what == "memcpy":
                        dst.copy_(flat, non_blocking=True)
                        torch.cuda.synchronize()
                    else:
                        cb.build_from_host(host, idx, out=out)
                        torch.cuda.synchronize()
                        device_bytes_fetched(clear=True)
                if world > 1:
                    dist.barrier()
                if active:
                    e0.record()
                    for _ in range(reps):
                        if what == "memcpy":
                            dst.copy_(flat, non_blocking=True)
                        else:
                            cb.build_from_host(host, idx, out=out)
                    e1.record()
                    torch.cuda.synchronize()
                    nbytes = reps * n if what == "memcpy" else device_bytes_fetched(clear=True)
                    gbs = nbytes / e0.elapsed_time(e1) / 1e6
                t = torch.tensor([gbs], dtype=torch.float64, device=dev)
                allv = [t]
                if world > 1:
                    allv = [torch.empty_like(t) for _ in range(world)]
                    dist.all_gather(allv, t)
                per = [round(float(v.item()), 2) for v in allv][:c]
                res[f"{'wc' if wc else 'pinned'}_{what}_x{c}"] = {"per_gpu_gb_s": per, "aggregate_gb_s": round(sum(per), 2)}
        del flat, vids, host
    if rank == 0:
        print(json.dumps({"probe": "pcie_h2d", "world": world, "host_cores": len(os.sched_getaffinity(0)), "numa": numa,
                          "frames": [H, W], "results": res}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
