"""Detector pre- and post-processing on the GPU, so ROI boxes are produced and consumed on the
device.  The YOLOv5 network and NMS themselves stay out of scope; these are the steps around them:

  letterbox_batch   src/yolo_detector.py:52-61   letterbox(640, stride) + transpose + [::-1] + float()/255
  scale_boxes       src/yolo_detector.py:76-81   scale_boxes(im.shape[2:], det[:, :4], im0.shape).round() -> int
  best_boxes        src/yolo_detector.py:161     highest-confidence detection per frame -> rows of the clip
                                                 kernel's ROI table (x1 = -1: no detection, full frame)
"""
import ctypes as C

import torch

from . import _lib

_DT = {torch.float32: _lib.DT_F32, torch.bfloat16: _lib.DT_BF16}


def letterbox_geometry(src_h, src_w, new_shape=(640, 640), stride=32, auto=True, scaleup=True):
    """yolov5 ``letterbox`` geometry for one frame size -> ``_lib.LetterboxGeom`` (src, new, top/left, out)."""
    if isinstance(new_shape, int):
        new_shape = (new_shape, new_shape)
    g = _lib.LetterboxGeom()
    _lib.check(_lib.lib().lvc_letterbox_geometry(int(src_h), int(src_w), int(new_shape[0]), int(new_shape[1]), int(stride),
                                                 int(bool(auto)), int(bool(scaleup)), C.byref(g)), "letterbox_geometry")
    return g


def letterbox_batch(frames, new_shape=(640, 640), stride=32, dtype=torch.float32, out=None):
    """frames: CUDA uint8 (N, H, W, 3) in the detector's input order (BGR, as cv2 decodes).  Returns
    ``(im, geom)``: ``im`` = (N, 3, out_h, out_w) network input in [0, 1] -- what ``process_image`` builds frame
    by frame on the host -- from ONE kernel launch."""
    if not (frames.is_cuda and frames.dtype == torch.uint8 and frames.dim() == 4 and frames.shape[-1] == 3
            and frames.is_contiguous()):
        raise ValueError("frames must be a contiguous CUDA uint8 tensor of shape (N, H, W, 3) (there is no CPU fallback)")
    if dtype not in _DT:
        raise ValueError("dtype must be float32 or bfloat16")
    n, h, w, _ = frames.shape
    g = letterbox_geometry(h, w, new_shape, stride)
    if out is None:
        out = torch.empty((n, 3, g.out_h, g.out_w), dtype=dtype, device=frames.device)
    stream = torch.cuda.current_stream(frames.device).cuda_stream
    _lib.check(_lib.lib().lvc_letterbox_forward(C.byref(g), frames.data_ptr(), n, _DT[dtype], out.data_ptr(), stream),
               "letterbox_forward")
    return out, g


def scale_boxes(det, geom):
    """det: CUDA float32 (n, >=4) rows starting with xyxy in letterboxed coordinates (yolov5 NMS rows are
    xyxy, conf, cls).  Returns int32 (n, 4) boxes in source-frame pixels, rounded and clipped like the reference."""
    if not (det.is_cuda and det.dtype == torch.float32 and det.dim() == 2 and det.shape[1] >= 4 and det.is_contiguous()):
        raise ValueError("det must be a contiguous CUDA float32 (n, >=4) tensor")
    out = torch.empty((det.shape[0], 4), dtype=torch.int32, device=det.device)
    stream = torch.cuda.current_stream(det.device).cuda_stream
    _lib.check(_lib.lib().lvc_scale_boxes(C.byref(geom), det.data_ptr(), det.shape[0], det.shape[1], out.data_ptr(), stream),
               "scale_boxes")
    return out


def best_boxes(det, frame_of, n_frames, geom, conf_thres=0.25):
    """det: CUDA float32 (n, 6) xyxy, conf, cls rows of all frames (letterboxed coordinates); frame_of: CUDA int32 (n,)
    frame index per row.  Returns int32 (n_frames, 4): per frame the highest-confidence box >= conf_thres scaled
    back to the frame, or -1s (no detection = full frame) -- the ROI table ``ClipBuilder`` takes as ``boxes``."""
    if not (det.is_cuda and det.dtype == torch.float32 and det.dim() == 2 and det.shape[1] == 6 and det.is_contiguous()):
        raise ValueError("det must be a contiguous CUDA float32 (n, 6) tensor")
    if not (frame_of.is_cuda and frame_of.dtype == torch.int32 and frame_of.numel() == det.shape[0]):
        raise ValueError("frame_of must be a CUDA int32 (n,) tensor")
    out = torch.empty((n_frames, 4), dtype=torch.int32, device=det.device)
    stream = torch.cuda.current_stream(det.device).cuda_stream
    _lib.check(_lib.lib().lvc_best_boxes(C.byref(geom), det.data_ptr(), frame_of.data_ptr(), det.shape[0], int(n_frames),
                                         float(conf_thres), out.data_ptr(), stream), "best_boxes")
    return out
