"""Drop-in patch / tubelet embedding modules backed by the tcgen05 GEMM (csrc/lvc_embed.cu).

Same class names, constructor meaning, parameter names (so reference checkpoints load unchanged:
``projection.weight/bias``, ``proj.weight/bias``, ``norm.weight/bias``), forward signatures, return
values and ValueError behaviour as the modules the reference instantiates:

  VivitTubeletEmbeddings      transformers modeling_vivit.py:36-77      (vivit_model.py:27-41)
  TimesformerPatchEmbeddings  transformers modeling_timesformer.py:38-64 (timesformer_model.py:27-42)
  PatchEmbed3d                torchvision swin_transformer.py:321-363   (swin3d.py:22-44)

Forward only (inference / frozen first layer): inputs are rounded to bf16, products accumulate in
fp32 on the tensor cores, output is bf16 by default (``out_dtype=torch.float32`` to keep fp32).
There is no fallback: forward raises if the input is not on a CUDA device.
"""
import ctypes as C
from types import SimpleNamespace

import torch
from torch import nn

from . import _lib

_DT = {torch.float32: _lib.DT_F32, torch.bfloat16: _lib.DT_BF16}


class _EmbedBase(nn.Module):
    kind = None
    x_layout = _lib.LAYOUT_BTCHW

    def _init_runtime(self, out_dtype):
        self.out_dtype = out_dtype
        self._packed = None
        self._packed_key = None
        self._ws = None
        self._f32 = {}
        self.launches = 0

    def _fp32(self, name, t):
        """fp32 contiguous view/copy of a parameter that OUTLIVES the launch (a raw pointer taken from a temporary
        would dangle for bf16 / half modules: the caching allocator hands the block to the next temporary).  Cached
        per parameter on (data_ptr, version, device, dtype)."""
        if t is None:
            return None
        key = (t.data_ptr(), t._version, t.device, t.dtype, tuple(t.shape))
        hit = self._f32.get(name)
        if hit is None or hit[0] != key:
            hit = (key, t.detach().float().contiguous().view(-1))
            self._f32[name] = hit
        return hit[1]

    def _check_no_grad(self, *extra):
        """Forward only: the reference FINE-TUNES these layers (vivit .../trainers/trainer.py:145-146), and a silent
        no_grad forward would freeze them without a word.  Raise when autograd is recording and any parameter of
        the module wants a gradient; wrap the call in torch.no_grad() / inference_mode(), or freeze the module."""
        if torch.is_grad_enabled() and any(q.requires_grad for q in list(self.parameters()) + [e for e in extra if e is not None]):
            raise RuntimeError(
                f"{type(self).__name__}: forward-only kernel (no autograd).  Call it under torch.no_grad() or freeze its "
                "parameters (requires_grad_(False)); to fine-tune this layer keep the reference's torch module.")

    def _conv(self):
        raise NotImplementedError

    def _ln(self):
        return None, None, 0.0

    def _desc(self, x, B, T, H, W):
        conv = self._conv()
        ks = conv.kernel_size
        kt, kh, kw = (1, ks[0], ks[1]) if len(ks) == 2 else ks
        return _lib.EmbedDesc(self.kind, _DT[x.dtype], self.x_layout, _DT[self.out_dtype], B, T, H, W,
                              kt, kh, kw, conv.out_channels, float(self._ln()[2]), 0)

    def forward_tokens(self, tokens, batch, frames, height=224, width=224, emb=None, out_shape=None):
        """Embedding of a clip batch the fused clip kernel already wrote as the bf16 patch matrix
        ``A[M][K]`` (``ClipBuilder(..., layout="TOKENS", patch=(kt,kh,kw))``): no im2col pass, one
        launch.  Returns the flat ``(M, N)`` result; callers reshape to the module's output shape."""
        conv = self._conv()
        M = tokens.shape[0]
        if tokens.dtype != torch.bfloat16 or tokens.dim() != 2 or not tokens.is_contiguous():
            raise ValueError("tokens must be a contiguous bf16 (M, K) matrix")
        return self._run(tokens, batch, frames, height, width, out_shape or (M, conv.out_channels),
                         layout=_lib.LAYOUT_TOKENS, emb=emb)

    def _run(self, x, B, T, H, W, out_shape, layout=None, emb=None):
        if not x.is_cuda:
            raise RuntimeError(f"{type(self).__name__}: input must be on a CUDA device (no CPU fallback)")
        self._check_no_grad(x, *(emb or ()))
        with torch.no_grad():
            return self._launch(x, B, T, H, W, out_shape, layout, emb)

    def _launch(self, x, B, T, H, W, out_shape, layout, emb):
        if x.dtype not in _DT:
            x = x.float()
        x = x.contiguous()
        conv = self._conv()
        w, bias = conv.weight, conv.bias
        if w.device != x.device:
            raise RuntimeError("module parameters and input are on different devices")
        lib = _lib.lib()
        d = self._desc(x, B, T, H, W)
        if layout is not None:
            d.x_layout = layout
        stream = torch.cuda.current_stream(x.device).cuda_stream
        key = (w.data_ptr(), w._version, tuple(w.shape))
        if self._packed is None or self._packed_key != key:
            self._packed = torch.empty(w.numel(), dtype=torch.bfloat16, device=x.device)
            _lib.check(lib.lvc_patch_embed_pack_weights(C.byref(d), self._fp32("weight", w).data_ptr(),
                                                        self._packed.data_ptr(), stream), "pack_weights")
            self._packed_key = key
            self.launches += 1
        need = lib.lvc_patch_embed_workspace(C.byref(d))
        if need < 0:
            _lib.check(_lib.E_SHAPE, type(self).__name__)
        if layout == _lib.LAYOUT_TOKENS:
            need = 0
        if self._ws is None or self._ws.numel() < need or self._ws.device != x.device:
            self._ws = torch.empty(max(need, 256), dtype=torch.uint8, device=x.device)
        out = torch.empty(out_shape, dtype=self.out_dtype, device=x.device)
        if bias is None:
            if getattr(self, "_zero_bias", None) is None or self._zero_bias.device != x.device:
                self._zero_bias = torch.zeros(conv.out_channels, device=x.device)
            bias = self._zero_bias
        lw, lb, _ = self._ln()
        bias_f, lw_f, lb_f = self._fp32("bias", bias), self._fp32("ln_w", lw), self._fp32("ln_b", lb)
        if emb is not None:  # fused cls / position / time embeddings
            cls, pos, temb = (self._fp32(n, t) for n, t in zip(("cls", "pos", "temb"), emb))
            st = lib.lvc_embeddings_forward(C.byref(d), x.data_ptr(), self._packed.data_ptr(),
                                            bias_f.data_ptr(), cls.data_ptr(), pos.data_ptr(),
                                            temb.data_ptr() if temb is not None else None, out.data_ptr(),
                                            self._ws.data_ptr(), stream)
            _lib.check(st, type(self).__name__)
            self.launches += lib.lvc_patch_embed_launch_count(C.byref(d))
            return out
        st = lib.lvc_patch_embed(C.byref(d), x.data_ptr(), self._packed.data_ptr(), bias_f.data_ptr(),
                                 lw_f.data_ptr() if lw_f is not None else None,
                                 lb_f.data_ptr() if lb_f is not None else None,
                                 out.data_ptr(), self._ws.data_ptr(), stream)
        _lib.check(st, type(self).__name__)
        self.launches += lib.lvc_patch_embed_launch_count(C.byref(d))
        return out


class VivitTubeletEmbeddings(_EmbedBase):
    """(B, T, C, H, W) -> (B, T/2 * H/16 * W/16, hidden).  ``config`` needs num_frames, image_size,
    tubelet_size, num_channels, hidden_size (a ``transformers.VivitConfig`` works)."""

    kind = _lib.EMBED_VIVIT

    def __init__(self, config=None, out_dtype=torch.bfloat16, **kw):
        super().__init__()
        if config is None:
            config = SimpleNamespace(num_frames=32, image_size=224, tubelet_size=[2, 16, 16], num_channels=3,
                                     hidden_size=768)
            for k, v in kw.items():
                setattr(config, k, v)
        self.num_frames = config.num_frames
        self.image_size = config.image_size if isinstance(config.image_size, (tuple, list)) else (config.image_size,) * 2
        self.patch_size = list(config.tubelet_size)
        self.num_patches = ((self.image_size[1] // self.patch_size[2]) * (self.image_size[0] // self.patch_size[1])
                            * (self.num_frames // self.patch_size[0]))
        self.embed_dim = config.hidden_size
        self.projection = nn.Conv3d(config.num_channels, config.hidden_size, kernel_size=tuple(config.tubelet_size),
                                    stride=tuple(config.tubelet_size))
        self._init_runtime(out_dtype)

    def _conv(self):
        return self.projection

    def forward(self, pixel_values, interpolate_pos_encoding=False):
        B, T, Cc, H, W = pixel_values.shape
        if not interpolate_pos_encoding and (H != self.image_size[0] or W != self.image_size[1]):
            raise ValueError(f"Image image size ({H}*{W}) doesn't match model ({self.image_size[0]}*{self.image_size[1]}).")
        if Cc != 3:
            raise ValueError("lvc_b200 patch embedding supports 3 input channels")
        kt, kh, kw = self.patch_size
        return self._run(pixel_values, B, T, H, W, (B, (T // kt) * (H // kh) * (W // kw), self.embed_dim))


class TimesformerPatchEmbeddings(_EmbedBase):
    """(B, T, C, H, W) -> (embeddings (B*T, H/16*W/16, hidden), T, W/16)."""

    kind = _lib.EMBED_TIMESFORMER

    def __init__(self, config=None, out_dtype=torch.bfloat16, **kw):
        super().__init__()
        if config is None:
            config = SimpleNamespace(image_size=224, patch_size=16, num_channels=3, hidden_size=768)
            for k, v in kw.items():
                setattr(config, k, v)
        image_size = config.image_size if isinstance(config.image_size, (tuple, list)) else (config.image_size,) * 2
        patch_size = config.patch_size if isinstance(config.patch_size, (tuple, list)) else (config.patch_size,) * 2
        self.image_size, self.patch_size = tuple(image_size), tuple(patch_size)
        self.num_patches = (image_size[1] // patch_size[1]) * (image_size[0] // patch_size[0])
        self.projection = nn.Conv2d(config.num_channels, config.hidden_size, kernel_size=self.patch_size,
                                    stride=self.patch_size)
        self._init_runtime(out_dtype)

    def _conv(self):
        return self.projection

    def forward(self, pixel_values):
        B, T, Cc, H, W = pixel_values.shape
        if Cc != 3:
            raise ValueError("lvc_b200 patch embedding supports 3 input channels")
        kh, kw = self.patch_size
        emb = self._run(pixel_values, B, T, H, W, (B * T, (H // kh) * (W // kw), self.projection.out_channels))
        return emb, T, W // kw


class PatchEmbed3d(_EmbedBase):
    """(B, C, T, H, W) -> (B, T/2, H/4, W/4, embed_dim), padded like torchvision's, LayerNorm fused."""

    kind = _lib.EMBED_SWIN
    x_layout = _lib.LAYOUT_BCTHW

    def __init__(self, patch_size, in_channels=3, embed_dim=96, norm_layer=None, out_dtype=torch.bfloat16):
        super().__init__()
        if norm_layer is None:
            raise NotImplementedError("lvc_b200.PatchEmbed3d fuses the LayerNorm every swin3d_* model uses; "
                                      "norm_layer=None is not supported")
        self.tuple_patch_size = (patch_size[0], patch_size[1], patch_size[2])
        self.proj = nn.Conv3d(in_channels, embed_dim, kernel_size=self.tuple_patch_size, stride=self.tuple_patch_size)
        self.norm = norm_layer(embed_dim)
        if not isinstance(self.norm, nn.LayerNorm):
            raise NotImplementedError("norm_layer must build an nn.LayerNorm")
        self._init_runtime(out_dtype)

    def _conv(self):
        return self.proj

    def _ln(self):
        return self.norm.weight, self.norm.bias, self.norm.eps

    def forward(self, x):
        B, Cc, T, H, W = x.shape
        if Cc != 3:
            raise ValueError("lvc_b200 patch embedding supports 3 input channels")
        kt, kh, kw = self.tuple_patch_size
        shape = (B, -(-T // kt), -(-H // kh), -(-W // kw), self.proj.out_channels)
        return self._run(x, B, T, H, W, shape)


    def forward_clips(self, clip_builder, batch, out=None):
        """The whole Video Swin front end -- gather, ROI crop, cv2 resize, normalise (``clip_builder``, preset of the
        cv2 chain with ``layout="TOKENS", patch=(2, 4, 4)``) and this module's Conv3d + LayerNorm -- as TWO CO-RESIDENT
        kernels: the embedding kernel is launched behind the clip kernel with programmatic stream serialization and
        consumes each 128-token tile of the patch matrix as soon as the clip kernel has completed it
        (``lvc_clip_swin_embed_forward``).  ``batch = clip_builder.prepare(videos, indices, boxes)``.
        Returns (B, T/2, 56, 56, embed_dim); bit-identical to ``self.forward_tokens(clip_builder.run(batch), ...)``."""
        self._check_no_grad()
        cb = clip_builder
        d_clip = batch.desc
        B, T = int(d_clip.batch), int(d_clip.num_frames)
        kt, kh, kw = self.tuple_patch_size
        if cb.layout != "TOKENS" or tuple(cb.patch) != (kt, kh, kw):
            raise ValueError(f"the clip builder must write layout='TOKENS' with patch={(kt, kh, kw)}")
        w, bias = self.proj.weight, self.proj.bias
        dev = cb.device
        if w.device != dev:
            raise RuntimeError("module parameters and clip builder are on different devices")
        with torch.no_grad():
            lib = _lib.lib()
            d = _lib.EmbedDesc(self.kind, _lib.DT_BF16, _lib.LAYOUT_TOKENS, _DT[self.out_dtype], B, T, 224, 224,
                               kt, kh, kw, self.proj.out_channels, float(self.norm.eps), 0)
            stream = torch.cuda.current_stream(dev).cuda_stream
            key = (w.data_ptr(), w._version, tuple(w.shape))
            if self._packed is None or self._packed_key != key:
                self._packed = torch.empty(w.numel(), dtype=torch.bfloat16, device=dev)
                _lib.check(lib.lvc_patch_embed_pack_weights(C.byref(d), self._fp32("weight", w).data_ptr(),
                                                            self._packed.data_ptr(), stream), "pack_weights")
                self._packed_key = key
                self.launches += 1
            tb, rb = C.c_int64(), C.c_int64()
            _lib.check(lib.lvc_clip_swin_embed_workspace(C.byref(d), C.byref(tb), C.byref(rb)), "clip_swin_embed_workspace")
            need = tb.value + rb.value
            ws = getattr(self, "_fused_ws", None)
            if ws is None or ws.numel() < need or ws.device != dev:
                ws = self._fused_ws = torch.empty(need, dtype=torch.uint8, device=dev)
            shape = (B, T // kt, 224 // kh, 224 // kw, self.proj.out_channels)
            if out is None:
                out = torch.empty(shape, dtype=self.out_dtype, device=dev)
            elif tuple(out.shape) != shape or out.dtype != self.out_dtype or not out.is_contiguous():
                raise ValueError("out has the wrong shape, dtype or strides")
            if bias is None:
                if getattr(self, "_zero_bias", None) is None or self._zero_bias.device != dev:
                    self._zero_bias = torch.zeros(self.proj.out_channels, device=dev)
                bias = self._zero_bias
            bias_f, lw_f, lb_f = self._fp32("bias", bias), self._fp32("ln_w", self.norm.weight), self._fp32("ln_b", self.norm.bias)
            st = lib.lvc_clip_swin_embed_forward(C.byref(d_clip), batch.p_vid, batch.p_idx, batch.p_box, cb.lut.data_ptr(),
                                                 C.byref(d), self._packed.data_ptr(), bias_f.data_ptr(), lw_f.data_ptr(),
                                                 lb_f.data_ptr(), ws.data_ptr(), ws.data_ptr() + tb.value, out.data_ptr(), stream)
            _lib.check(st, "lvc_clip_swin_embed_forward")
            self.launches += 2
        return out

    def tokens_view(self, batch_size, frames):
        """The bf16 patch matrix the last forward_clips call left in the module's workspace (tests, debugging)."""
        kt, kh, kw = self.tuple_patch_size
        M, K = batch_size * (frames // kt) * (224 // kh) * (224 // kw), 3 * kt * kh * kw
        return self._fused_ws[:M * K * 2].view(torch.bfloat16).view(M, K)


class VivitEmbeddings(nn.Module):
    """``transformers`` ``VivitEmbeddings`` (modeling_vivit.py:141-155): tubelet embedding, cls token and position
    embeddings, produced by ONE GEMM launch whose epilogue adds the tables and writes the (B, 1 + tokens, hidden)
    layout.  Parameter names match the HF module (``cls_token``, ``position_embeddings``, ``patch_embeddings.*``)."""

    def __init__(self, config=None, out_dtype=torch.bfloat16, **kw):
        super().__init__()
        self.patch_embeddings = VivitTubeletEmbeddings(config, out_dtype=out_dtype, **kw)
        E, P = self.patch_embeddings.embed_dim, self.patch_embeddings.num_patches
        self.cls_token = nn.Parameter(torch.zeros(1, 1, E))
        self.position_embeddings = nn.Parameter(torch.zeros(1, P + 1, E))

    def forward(self, pixel_values=None, interpolate_pos_encoding=False, tokens=None, batch=None):
        pe = self.patch_embeddings
        if interpolate_pos_encoding:
            raise NotImplementedError("interpolate_pos_encoding is not fused; sizes must match the config")
        emb = (self.cls_token, self.position_embeddings, None)
        if tokens is not None:  # bf16 patch matrix written by the clip kernel (layout="TOKENS")
            P = pe.num_patches
            return pe.forward_tokens(tokens, batch, pe.num_frames, emb=emb, out_shape=(batch, 1 + P, pe.embed_dim))
        B, T, Cc, H, W = pixel_values.shape
        if H != pe.image_size[0] or W != pe.image_size[1] or T != pe.num_frames:
            raise ValueError(f"Input ({T}x{H}*{W}) doesn't match model ({pe.num_frames}x{pe.image_size[0]}*{pe.image_size[1]}).")
        return pe._run(pixel_values, B, T, H, W, (B, 1 + pe.num_patches, pe.embed_dim), emb=emb)


class TimesformerEmbeddings(nn.Module):
    """``transformers`` ``TimesformerEmbeddings`` (modeling_timesformer.py:70-144) for matching sizes: patch
    embedding, cls token, position embeddings, time embeddings and the (B, patch*T + t) token order, fused into
    the GEMM epilogue.  ``attention_type='space_only'`` is not supported (its output keeps the (B*T, 197, E) shape)."""

    def __init__(self, config=None, out_dtype=torch.bfloat16, num_frames=8, attention_type="divided_space_time", **kw):
        super().__init__()
        self.patch_embeddings = TimesformerPatchEmbeddings(config, out_dtype=out_dtype, **kw)
        E = self.patch_embeddings.projection.out_channels
        self.num_patches = self.patch_embeddings.num_patches
        self.num_frames = getattr(config, "num_frames", num_frames) if config is not None else num_frames
        self.attention_type = getattr(config, "attention_type", attention_type) if config is not None else attention_type
        if self.attention_type == "space_only":
            raise NotImplementedError("space_only attention keeps per-frame cls rows; not fused")
        self.cls_token = nn.Parameter(torch.zeros(1, 1, E))
        self.position_embeddings = nn.Parameter(torch.zeros(1, self.num_patches + 1, E))
        self.time_embeddings = nn.Parameter(torch.zeros(1, self.num_frames, E))

    def forward(self, pixel_values=None, tokens=None, batch=None):
        pe = self.patch_embeddings
        E = pe.projection.out_channels
        emb = (self.cls_token, self.position_embeddings, self.time_embeddings)
        if tokens is not None:
            return pe.forward_tokens(tokens, batch, self.num_frames, emb=emb,
                                     out_shape=(batch, 1 + self.num_patches * self.num_frames, E))
        B, T, Cc, H, W = pixel_values.shape
        if (H, W) != tuple(pe.image_size) or T != self.num_frames:
            raise ValueError("input size does not match the position / time tables (resizing is not fused)")
        return pe._run(pixel_values, B, T, H, W, (B, 1 + self.num_patches * T, E), emb=emb)
