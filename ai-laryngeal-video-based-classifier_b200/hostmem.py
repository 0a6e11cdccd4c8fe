"""Host memory for decoded frames that the kernels read in place (zero-copy ingest), and NUMA placement of a rank.

``pinned_empty`` allocates page-locked memory through the C ABI (``lvc_host_alloc``): portable, mapped, optionally
write-combined -- device reads of write-combined memory are not snooped by the CPU caches, which matters when eight
GPUs pull frames from one host at once; the CPU should only WRITE such a buffer (a decoder filling it), never read it.
``bind_rank_to_gpu`` pins the calling process to the CPUs of its GPU's NUMA node BEFORE it allocates pinned memory, so
first-touch places the frames on the memory controller next to the GPU's PCIe root complex.
"""
import ctypes as C
import os

import numpy as np
import torch

from . import _lib


class _Owner:
    def __init__(self, ptr):
        self.ptr = ptr

    def __del__(self):
        try:
            if self.ptr:
                _lib.lib().lvc_host_free(self.ptr)
        except Exception:
            pass
        self.ptr = None


def pinned_empty(shape, write_combined=False):
    """uint8 CPU tensor of ``shape`` in page-locked host memory (``is_pinned()`` is True; usable with
    ``ClipBuilder.build_from_host``).  The allocation is freed when the tensor's storage dies."""
    if not torch.cuda.is_available():
        raise RuntimeError("pinned_empty needs a CUDA device (cudaHostAlloc)")
    n = int(np.prod(shape))
    ptr = C.c_void_p()
    _lib.check(_lib.lib().lvc_host_alloc(max(n, 1), _lib.HOST_WRITE_COMBINED if write_combined else 0, C.byref(ptr)),
               "lvc_host_alloc")
    arr = (C.c_uint8 * max(n, 1)).from_address(ptr.value)
    arr._owner = _Owner(ptr.value)   # the buffer object torch keeps alive owns the allocation
    return torch.frombuffer(arr, dtype=torch.uint8, count=n).view(*shape)


def gpu_numa_info(index):
    """NUMA node and CPU list of CUDA device ``index`` as sysfs reports them (None when unknown, e.g. inside a VM that
    shows a single node)."""
    try:
        p = torch.cuda.get_device_properties(index)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
    except Exception:
        return {"pci": None, "node": None, "cpus": None}
    base = f"/sys/bus/pci/devices/{bdf}"
    node, cpus = None, None
    try:
        node = int(open(os.path.join(base, "numa_node")).read().strip())
        txt = open(os.path.join(base, "local_cpulist")).read().strip()
        cpus = set()
        for part in txt.split(","):
            if part:
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
    except Exception:
        pass
    return {"pci": bdf, "node": node, "cpus": sorted(cpus) if cpus else None}


def bind_rank_to_gpu(index):
    """Restrict this process to the CPUs local to GPU ``index`` (call before allocating pinned memory).  Returns a
    record of what was done; a box that exposes no topology (node -1 / one node for every GPU) is left alone."""
    info = gpu_numa_info(index)
    rec = {"pci": info["pci"], "numa_node": info["node"], "bound": False}
    try:
        have = os.sched_getaffinity(0)
    except AttributeError:
        rec["why"] = "no sched_getaffinity"
        return rec
    nodes = set()
    try:
        nodes = {d for d in os.listdir("/sys/devices/system/node") if d.startswith("node") and d[4:].isdigit()}
    except Exception:
        pass
    rec["host_numa_nodes"] = len(nodes) or None
    if info["node"] is None or info["node"] < 0 or not info["cpus"]:
        rec["why"] = "the box reports no NUMA node for this GPU"
        return rec
    want = have & set(info["cpus"])
    if not want or want == have:
        rec["why"] = "GPU-local CPUs are all (or none) of the CPUs this process may use: nothing to narrow"
        return rec
    os.sched_setaffinity(0, want)
    rec.update(bound=True, cpus=len(want))
    return rec
