"""Clip construction -> patch embedding as ONE replayable CUDA graph.

For small batches (BASELINE configs 2-4) the device needs 85-210 us per batch while the Python + ctypes call path of
two launches costs about as much again, so a loop over batches is host-bound.  ``ClipEmbedPipeline`` captures
[table upload -> fused clip kernel (token layout) -> tcgen05 GEMM] once; per batch the host only rewrites the small
pinned table (frame indices, optionally ROI boxes) and replays the graph.  Buffers are static: the returned tensor
is overwritten by the call after the next one (two graphs alternate, see ``ClipEmbedPipeline``).
"""
import ctypes as C

import numpy as np
import torch

from . import _lib


class _Slot:
    """One captured graph with its own pinned host table, device table and output buffer."""

    def __init__(self):
        self.batch = self.hv = self.graph = self.out = None
        self.done = None   # event recorded behind the slot's last replay


class ClipEmbedPipeline:
    """Two graphs with private pinned tables alternate: the host rewrites slot k % 2 while the replay of slot (k-1) % 2
    may still be queued, and waits (normally not at all) for the replay that last READ the table it is about to
    overwrite.  Rewriting a single table right after ``replay()`` would race with the graph's own upload node whenever
    the GPU runs behind the host (batch k built with batch k+1's indices).  The returned tensor belongs to the slot: it
    is overwritten by the call after the next one."""

    SLOTS = 2

    def __init__(self, clip_builder, embed_module, videos, num_frames, boxes=False, embeddings=None):
        """clip_builder: ``ClipBuilder(..., layout="TOKENS", patch=...)``; embed_module: a patch-embedding module of
        ``patch_embed`` (or ``VivitEmbeddings`` / ``TimesformerEmbeddings`` via ``embeddings=``); videos: the resident
        CUDA uint8 (N,H,W,3) tensors of the batch slots (replaceable later by same-shaped ones with ``set_videos``)."""
        if clip_builder.layout != "TOKENS":
            raise ValueError("ClipEmbedPipeline needs a ClipBuilder with layout='TOKENS'")
        self.cb, self.m, self.emb = clip_builder, embed_module, embeddings
        self.B, self.T = len(videos), int(num_frames)
        self._shapes = [tuple(v.shape[1:]) for v in videos]
        idx0 = np.zeros((self.B, self.T), np.int32)
        box0 = np.full((self.B, self.T, 4), -1, np.int32) if boxes else None
        dev = self.cb.device
        self.tok = None
        self.slots = []
        for _ in range(self.SLOTS):
            sl = _Slot()
            sl.batch = self.cb.prepare(videos, idx0, box0)
            sl.hv = sl.batch.host.numpy()
            self.slots.append(sl)
        b0 = self.slots[0].batch
        self._off_idx = b0.p_idx - b0.p_vid
        self._off_box = (b0.p_box - b0.p_vid) if boxes else None
        self.tok = torch.empty(b0.shape, dtype=torch.bfloat16, device=dev)   # shared: replays are stream-ordered
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side), torch.no_grad():   # warm-up outside the capture: packs weights, sets function attributes
            self._enqueue(self.slots[0])
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        for sl in self.slots:
            sl.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(sl.graph), torch.no_grad():
                sl.out = self._enqueue(sl)
        self.replays = 0

    @property
    def batch(self):
        return self.slots[0].batch

    def _enqueue(self, sl):
        sl.batch.tables.copy_(sl.batch.host, non_blocking=True)
        self.cb.run(sl.batch, self.tok)
        if self.emb is not None:
            return self.emb(tokens=self.tok, batch=self.B)
        return self.m.forward_tokens(self.tok, self.B, self.T)

    def _host_table(self, sl):
        """The slot's pinned table, safe to rewrite: the replay that last read it has finished its upload."""
        if sl.done is not None:
            sl.done.synchronize()
        return sl.hv

    def set_videos(self, videos):
        """Point the batch slots at other resident videos of the SAME shapes (no re-capture needed)."""
        if [tuple(v.shape[1:]) for v in videos] != self._shapes:
            raise ValueError("set_videos: frame sizes differ from the captured batch; build a new pipeline")
        for v in videos:
            if not (v.is_cuda and v.dtype == torch.uint8 and v.is_contiguous() and v.data_ptr() % 16 == 0):
                raise ValueError("each video must be a contiguous, 16-byte aligned CUDA uint8 tensor")
        for sl in self.slots:
            vt = (_lib.Video * self.B).from_buffer(self._host_table(sl))
            for i, v in enumerate(videos):
                vt[i].base, vt[i].nbytes, vt[i].n_frames = v.data_ptr(), v.numel(), v.shape[0]
            sl.batch.videos = list(videos)

    def __call__(self, indices, boxes=None):
        idx = np.ascontiguousarray(np.asarray(indices, dtype=np.int32).reshape(self.B, self.T))
        bx = None
        if boxes is not None:
            if self._off_box is None:
                raise ValueError("the pipeline was captured without an ROI table (boxes=True)")
            bx = np.ascontiguousarray(np.asarray(boxes, dtype=np.int32).reshape(self.B, self.T, 4))
        sl = self.slots[self.replays % self.SLOTS]
        hv = self._host_table(sl)
        hv[self._off_idx:self._off_idx + idx.nbytes] = idx.view(np.uint8).reshape(-1)
        if bx is not None:
            hv[self._off_box:self._off_box + bx.nbytes] = bx.view(np.uint8).reshape(-1)
        sl.graph.replay()
        if sl.done is None:
            sl.done = torch.cuda.Event()
        sl.done.record(torch.cuda.current_stream(self.cb.device))
        self.replays += 1
        return sl.out
