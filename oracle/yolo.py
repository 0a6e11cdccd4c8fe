"""CPU oracle of the detector pre / post-processing (TEST INFRASTRUCTURE ONLY, like everything under oracle/).

``letterbox`` and ``scale_boxes`` belong to ultralytics/yolov5 (utils/augmentations.py, utils/general.py), which the
reference vendors as an EMPTY, unpinned git submodule (`yolov5/`): they are restated from the published
definitions -> PARITY UNPINNED for these two; the call sites that fix their arguments and the surrounding
arithmetic are the reference's own (src/yolo_detector.py:52-61, :76-81)."""
import numpy as np


def letterbox(im, new_shape=(640, 640), color=(114, 114, 114), auto=True, scale_fill=False, scaleup=True, stride=32):
    """yolov5 utils/augmentations.py::letterbox, restated.  Returns (image, ratio, (dw, dh))."""
    import cv2

    shape = im.shape[:2]
    if isinstance(new_shape, int):
        new_shape = (new_shape, new_shape)
    r = min(new_shape[0] / shape[0], new_shape[1] / shape[1])
    if not scaleup:
        r = min(r, 1.0)
    ratio = r, r
    new_unpad = int(round(shape[1] * r)), int(round(shape[0] * r))
    dw, dh = new_shape[1] - new_unpad[0], new_shape[0] - new_unpad[1]
    if auto:
        dw, dh = np.mod(dw, stride), np.mod(dh, stride)
    elif scale_fill:
        dw, dh = 0.0, 0.0
        new_unpad = (new_shape[1], new_shape[0])
        ratio = new_shape[1] / shape[1], new_shape[0] / shape[0]
    dw /= 2
    dh /= 2
    if shape[::-1] != new_unpad:
        im = cv2.resize(im, new_unpad, interpolation=cv2.INTER_LINEAR)
    top, bottom = int(round(dh - 0.1)), int(round(dh + 0.1))
    left, right = int(round(dw - 0.1)), int(round(dw + 0.1))
    im = cv2.copyMakeBorder(im, top, bottom, left, right, cv2.BORDER_CONSTANT, value=color)
    return im, ratio, (dw, dh)


def preprocess(image_array, imgsz=(640, 640), stride=32):
    """src/yolo_detector.py:52-61: letterbox -> transpose((2,0,1))[::-1] -> float() -> /= 255 -> [None]."""
    import torch

    im = letterbox(image_array, imgsz, stride=stride)[0]
    im = im.transpose((2, 0, 1))[::-1].copy()
    im = torch.from_numpy(np.ascontiguousarray(im)).float()
    im /= 255
    return im[None]


def scale_boxes(img1_shape, boxes, img0_shape):
    """yolov5 utils/general.py::scale_boxes (+ clip_boxes), restated on a float32 torch tensor (in place)."""
    gain = min(img1_shape[0] / img0_shape[0], img1_shape[1] / img0_shape[1])
    pad = (img1_shape[1] - img0_shape[1] * gain) / 2, (img1_shape[0] - img0_shape[0] * gain) / 2
    boxes[..., [0, 2]] -= pad[0]
    boxes[..., [1, 3]] -= pad[1]
    boxes[..., :4] /= gain
    boxes[..., 0].clamp_(0, img0_shape[1])
    boxes[..., 1].clamp_(0, img0_shape[0])
    boxes[..., 2].clamp_(0, img0_shape[1])
    boxes[..., 3].clamp_(0, img0_shape[0])
    return boxes


def detections_to_int_boxes(det, im_shape, im0_shape):
    """src/yolo_detector.py:76-81: det[:, :4] = scale_boxes(...).round(); bbox = [int(x) for x in xyxy]."""
    det = det.clone()
    det[:, :4] = scale_boxes(im_shape, det[:, :4], im0_shape).round()
    return [[int(x) for x in row[:4]] for row in det]
