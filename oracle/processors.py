"""Oracle: per-clip pixel pipelines of the reference (ROI crop).

Test infrastructure.  Each function restates one reference recipe on numpy / torch CPU
and cites the lines it follows.  Inputs are decoded RGB uint8 frames ``(T, H, W, 3)``.
Pinned against the reference's own third-party calls by tests/golden/make_golden.py;
``roi_crop`` and ``ptv_eval_chain`` are PARITY UNPINNED (no reference implementation /
package absent) -- see DESIGN.md.
"""
import numpy as np

from .resize import cv2_resize_linear_u8, pil_resize_bilinear_u8


def roi_crop(frame, box):
    """ROI crop -- semantics defined by this repo: the reference only
    defines the box record ``[x1, y1, x2, y2]`` int pixels (src/yolo_detector.py:76-85).
    Crop = numpy half-open slice of the box clipped to the frame, at least 1x1;
    ``box is None`` or an all-negative box (= no detection) keeps the full frame."""
    H, W = frame.shape[:2]
    if box is None or box[0] < 0:
        return frame
    x1, y1, x2, y2 = (int(v) for v in box)
    x1 = min(max(x1, 0), W - 1)
    y1 = min(max(y1, 0), H - 1)
    x2 = min(max(x2, x1 + 1), W)
    y2 = min(max(y2, y1 + 1), H)
    return frame[y1:y2, x1:x2]


def resize_clip_cv2(frames, boxes=None, size=224):
    """a6 (+ROI): per-frame ``cv2.resize(frame, (224, 224))`` -- vivit .../dataset.py:271-277,
    resnet50-2d-lstm/inference.py:158-163 -- on the bit-exact restatement."""
    out = np.empty((len(frames), size, size, 3), np.uint8)
    for t, fr in enumerate(frames):
        src = roi_crop(fr, None if boxes is None else boxes[t])
        out[t] = cv2_resize_linear_u8(np.ascontiguousarray(src), size, size)
    return out


def manual_normalize(frames224, mean=(0.45, 0.45, 0.45), std=(0.225, 0.225, 0.225)):
    """a9 -- resnet50-2d-lstm/inference.py:166-178: ``f32(x)/255.0``, ``- mean``, ``/ std``,
    ``permute(3,0,1,2)`` -> (C, T, H, W) float32 (the caller adds the batch dim)."""
    import torch

    x = torch.from_numpy(np.asarray(frames224).astype(np.float32) / 255.0)
    x = x - torch.tensor(mean).view(1, 1, 1, 3)
    x = x / torch.tensor(std).view(1, 1, 1, 3)
    return x.permute(3, 0, 1, 2).contiguous().numpy()


def vivit_processor(frames224):
    """a7 -- VivitImageProcessor defaults as called at vivit .../trainers/trainer.py:77:
    Pillow BILINEAR to shortest_edge 256 -> centre crop 224 -> ``f32(f64(x)*(1/127.5)) - 1``
    -> ``(x - 0.5) / 0.5`` -> CHW; stacked to (T, 3, 224, 224) float32
    (transformers image_processing_vivit.py:223-283, image_transforms.py:89-124,394-442)."""
    T, H, W, _ = frames224.shape
    short = min(H, W)
    # get_resize_output_image_size(shortest_edge=256, default_to_square=False)
    if H <= W:
        nh, nw = 256, int(256 * W / H)
    else:
        nh, nw = int(256 * H / W), 256
    del short
    top = (nh - 224) // 2
    left = (nw - 224) // 2
    out = np.empty((T, 3, 224, 224), np.float32)
    mean = np.array([0.5, 0.5, 0.5], np.float32)
    std = np.array([0.5, 0.5, 0.5], np.float32)
    for t in range(T):
        r = pil_resize_bilinear_u8(frames224[t], nh, nw)[top : top + 224, left : left + 224]
        x = (r.astype(np.float64) * (1 / 127.5)).astype(np.float32)
        x = x - 1
        x = (x - mean) / std
        out[t] = x.transpose(2, 0, 1)
    return out


def hf_shortest_edge_crop_u8(frames, edge, crop=224):
    """Geometry the HF video processors apply to RAW frames (data_preprocessing.py:16,
    data_preprocessing_timesformer.py:16): ``resize(size={"shortest_edge": edge}, resample=BILINEAR)`` through
    Pillow, then ``center_crop(crop)``; uint8 (T, crop, crop, 3).  transformers image_transforms.py
    get_resize_output_image_size / resize / center_crop."""
    T, H, W, _ = frames.shape
    if H <= W:
        nh, nw = edge, int(edge * W / H)
    else:
        nh, nw = int(edge * H / W), edge
    top, left = (nh - crop) // 2, (nw - crop) // 2
    return np.stack([pil_resize_bilinear_u8(frames[t], nh, nw)[top : top + crop, left : left + crop] for t in range(T)])


def videomae_processor_exact(frames, mean=(0.45, 0.45, 0.45), std=(0.225, 0.225, 0.225)):
    """The call form of timesformer .../trainers/trainer.py:91-98 on frames that are NOT 224x224:
    ``size={"height": 224, "width": 224}`` resizes both axes to 224 with Pillow BILINEAR (aspect not kept), the
    224 crop is the identity; then ``videomae_processor``'s rescale + normalise."""
    frames = np.asarray(frames)
    return videomae_processor(np.stack([pil_resize_bilinear_u8(f, 224, 224) for f in frames]), mean, std)


def videomae_processor_raw(frames, mean=(0.45, 0.45, 0.45), std=(0.225, 0.225, 0.225), edge=224):
    """VideoMAEImageProcessor defaults of the TimeSformer checkpoint on raw frames: shortest edge 224
    (Pillow BILINEAR), centre crop 224, then ``videomae_processor``'s rescale + normalise."""
    return videomae_processor(hf_shortest_edge_crop_u8(np.asarray(frames), edge), mean, std)


def videomae_processor(frames224, mean=(0.45, 0.45, 0.45), std=(0.225, 0.225, 0.225)):
    """a8 -- VideoMAE-class processor as forced to size=crop=224 at
    timesformer .../trainers/trainer.py:91-98 (resize and crop are identities on 224x224
    frames): ``f32(f64(x)*(1/255))`` then ``(x - mean)/std`` in float32, CHW.
    mean/std are parameters: the checkpoint's preprocessor_config.json is not available
    offline."""
    x = (np.asarray(frames224).astype(np.float64) * (1 / 255)).astype(np.float32)
    m = np.array(mean, np.float32)
    s = np.array(std, np.float32)
    x = (x - m) / s
    return np.ascontiguousarray(x.transpose(0, 3, 1, 2))


def ptv_eval_geometry(H, W, short=256, crop=224):
    """Sizes used by ShortSideScale(256) + CenterCrop(224) (a10):
    ``new_long = int(floor(long/short_src * 256))``; crop offsets ``int(round((n - 224)/2.0))``
    (torchvision CenterCrop).  Returns (new_h, new_w, top, left)."""
    import math

    if W < H:
        nw, nh = short, int(math.floor((float(H) / W) * short))
    else:
        nh, nw = short, int(math.floor((float(W) / H) * short))
    top = int(round((nh - crop) / 2.0))
    left = int(round((nw - crop) / 2.0))
    return nh, nw, top, left


def ptv_eval_chain(frames, mean=(0.45, 0.45, 0.45), std=(0.225, 0.225, 0.225)):
    """a10 eval chain -- videoswintransformer/.../dataset.py:164-172 and
    videoswintransformer/inference.py:50-63 restated on torch CPU (pytorchvideo absent):
    float32 (C,T,H,W) 0..255 -> F.interpolate(bilinear, align_corners=False) to short side
    256 -> CenterCrop(224) -> ``(x - mean)/std`` WITHOUT /255.  PARITY UNPINNED."""
    import torch
    import torch.nn.functional as F

    x = torch.from_numpy(np.asarray(frames)).permute(3, 0, 1, 2).float()  # (C,T,H,W)
    H, W = x.shape[2:]
    nh, nw, top, left = ptv_eval_geometry(H, W)
    x = F.interpolate(x, size=(nh, nw), mode="bilinear", align_corners=False)
    x = x[:, :, top : top + 224, left : left + 224]
    m = torch.tensor(mean, dtype=torch.float32).view(3, 1, 1, 1)
    s = torch.tensor(std, dtype=torch.float32).view(3, 1, 1, 1)
    return ((x - m) / s).contiguous().numpy()


def ptv_train_chain(frames, seed=None, mean=(0.45, 0.45, 0.45), std=(0.225, 0.225, 0.225), min_size=256,
                    max_size=320, flip_p=0.5, generator_seeded=False):
    """f2 train chain -- videoswintransformer/.../dataset.py:153-162 (= resnet50-2d-lstm/.../dataset.py:63-72,
    resnet50-3d-video/.../dataset.py:173-182): RandomShortSideScale(256, 320) -> transforms.RandomCrop(224)
    -> transforms.RandomHorizontalFlip(0.5) -> Normalize, applied to the whole (C,T,H,W) clip.
    RandomCrop / RandomHorizontalFlip are torchvision's own classes (installed); RandomShortSideScale is
    restated from pytorchvideo's published definition (``size = torch.randint(min, max + 1, (1,)).item()``
    then short_side_scale, package absent -> PARITY UNPINNED for that step).  All draws come from torch's
    global CPU generator: ``seed`` calls torch.manual_seed first (unless the caller already seeded it).
    Returns (out (C,T,224,224) float32, [size, top, left, flip])."""
    import math

    import torch
    import torch.nn.functional as F
    from torchvision import transforms

    if seed is not None and not generator_seeded:
        torch.manual_seed(seed)
    x = torch.from_numpy(np.asarray(frames)).permute(3, 0, 1, 2).float()  # (C,T,H,W)
    H, W = x.shape[2:]
    size = int(torch.randint(min_size, max_size + 1, (1,)).item())
    if W < H:
        nh, nw = int(math.floor((float(H) / W) * size)), size
    else:
        nh, nw = size, int(math.floor((float(W) / H) * size))
    x = F.interpolate(x, size=(nh, nw), mode="bilinear", align_corners=False)
    # record what the torchvision classes are about to draw, from a cloned generator state
    state = torch.get_rng_state()
    top, left = 0, 0
    if not (nh == 224 and nw == 224):
        top = int(torch.randint(0, nh - 224 + 1, size=(1,)).item())
        left = int(torch.randint(0, nw - 224 + 1, size=(1,)).item())
    flip = int(bool(torch.rand(1) < flip_p))
    torch.set_rng_state(state)
    x = transforms.RandomCrop(224)(x)
    x = transforms.RandomHorizontalFlip(p=flip_p)(x)
    m = torch.tensor(mean, dtype=torch.float32).view(3, 1, 1, 1)
    s = torch.tensor(std, dtype=torch.float32).view(3, 1, 1, 1)
    return ((x - m) / s).contiguous().numpy(), [size, top, left, flip]
