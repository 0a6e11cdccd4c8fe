"""Oracle: the three bilinear resamplers the reference calls.

Test infrastructure.  The algorithms live in third-party binaries (opencv-python pinned
4.11.0.86, Pillow pinned 10.4.0 in the reference's requirements.txt:65,70; torch 2.6.0
:103) that are not under /root/reference; each function restates the library's
published algorithm in integer / float32 numpy and is pinned against the installed
library by tests/golden/make_golden.py (cv2 4.13.0, Pillow 12.2.0, torch 2.11.0).
"""
import numpy as np

# ----------------------------------------------------------------------------------
# cv2.resize(src, (dw, dh)) -- default INTER_LINEAR, uint8
# call sites: vivit .../dataset.py:275,348; timesformer .../dataset.py:289,362;
# resnet50-2d-lstm/inference.py:162; resnet50-3d-video/inference.py:378
# ----------------------------------------------------------------------------------


def cv2_linear_taps(src, dst, vertical=False):
    """Per-axis tap table of OpenCV's INTER_LINEAR for 8-bit images.

    ``vertical=True`` gives the row table: cv::resize does NOT reset the fraction of
    out-of-range rows, it only clamps the two row indices when it fetches them, so the
    result is (clip(sy), clip(sy+1)) blended with the unclamped coefficients -- which
    differs from the horizontal rule by 1 LSB on the first/last rows of an up-scale
    because both products are truncated separately.  Returns (y0, y1, b0, b1) then.

    Returns (ofs int32[dst], a0 int32[dst], a1 int32[dst]): the first tap index and the two
    11-bit fixed-point coefficients.  Follows cv::resize's table set-up:
    ``fx = (float)((dx+0.5)*scale - 0.5)`` with ``scale = 1./((double)dst/src)``,
    ``sx = floor(fx)``, ``fx -= sx``; ``sx < 0 -> (0, 0)``; ``sx >= src-1 -> (src-1, 0)``;
    coefficients ``saturate_cast<short>((1-fx)*2048)`` / ``(fx*2048)`` (round half even).
    """
    inv = np.float64(dst) / np.float64(src)
    scale = np.float64(1.0) / inv
    d = np.arange(dst, dtype=np.float64)
    f = ((d + 0.5) * scale - 0.5).astype(np.float32)
    s = np.floor(f).astype(np.int32)
    fr = (f - s.astype(np.float32)).astype(np.float32)
    if vertical:
        b0 = np.rint((np.float32(1.0) - fr) * np.float32(2048.0)).astype(np.int32)
        b1 = np.rint(fr * np.float32(2048.0)).astype(np.int32)
        return np.clip(s, 0, src - 1), np.clip(s + 1, 0, src - 1), b0, b1
    lo = s < 0
    s[lo] = 0
    fr[lo] = 0.0
    hi = s >= src - 1
    s[hi] = src - 1
    fr[hi] = 0.0
    a0 = np.rint((np.float32(1.0) - fr) * np.float32(2048.0)).astype(np.int32)
    a1 = np.rint(fr * np.float32(2048.0)).astype(np.int32)
    return s, a0, a1


def cv2_resize_linear_u8(img, dw=224, dh=224):
    """Bit-exact restatement of ``cv2.resize(img, (dw, dh))`` for uint8 HxWxC.

    Horizontal pass in int32 (``S = a0*p[x0] + a1*p[x0+1]``), vertical pass with OpenCV's
    8-bit formula ``(((b0*(S0>>4))>>16) + ((b1*(S1>>4))>>16) + 2) >> 2``.
    An exact 2x2 down-scale is switched to INTER_AREA by cv::resize (mean of the 2x2
    block, round half up); an identity size is a copy.
    """
    img = np.asarray(img)
    assert img.dtype == np.uint8 and img.ndim == 3
    sh, sw = img.shape[:2]
    if sh == dh and sw == dw:
        return img.copy()
    if sw == 2 * dw and sh == 2 * dh:
        p = img.astype(np.int32)
        s = p[0::2, 0::2] + p[0::2, 1::2] + p[1::2, 0::2] + p[1::2, 1::2]
        return ((s + 2) >> 2).astype(np.uint8)
    x0, ax0, ax1 = cv2_linear_taps(sw, dw)
    y0, y1, ay0, ay1 = cv2_linear_taps(sh, dh, vertical=True)
    x1 = np.minimum(x0 + 1, sw - 1)
    p = img.astype(np.int32)
    r0 = p[y0]
    r1 = p[y1]
    S0 = r0[:, x0] * ax0[None, :, None] + r0[:, x1] * ax1[None, :, None]
    S1 = r1[:, x0] * ax0[None, :, None] + r1[:, x1] * ax1[None, :, None]
    b0 = ay0[:, None, None]
    b1 = ay1[:, None, None]
    out = (((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2
    return np.clip(out, 0, 255).astype(np.uint8)


# ----------------------------------------------------------------------------------
# Pillow Image.resize(size, BILINEAR) on 8-bit RGB -- what transformers'
# image_transforms.resize() executes for VivitImageProcessor
# (vivit .../trainers/trainer.py:77 -> image_processing_vivit.py resize)
# ----------------------------------------------------------------------------------

PIL_PRECISION_BITS = 22


def pil_bilinear_coeffs(in_size, out_size):
    """Pillow's precompute_coeffs + normalize_coeffs_8bpc for the triangle filter.

    Returns (xmin int32[out], count int32[out], kk int32[out, ksize]) with 22-bit
    fixed-point weights.  The support widens with the down-scale (antialiasing); for the
    224 -> 256 up-scale the support is 1 and each output has 2-3 taps.
    """
    scale = float(in_size) / float(out_size)
    filterscale = max(scale, 1.0)
    support = 1.0 * filterscale
    ksize = int(np.ceil(support)) * 2 + 1
    xmins = np.zeros(out_size, np.int32)
    counts = np.zeros(out_size, np.int32)
    kk = np.zeros((out_size, ksize), np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        xmin = int(center - support + 0.5)
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > in_size:
            xmax = in_size
        n = xmax - xmin
        w = np.zeros(n, np.float64)
        for x in range(n):
            t = (x + xmin - center + 0.5) * ss
            t = -t if t < 0 else t
            w[x] = 1.0 - t if t < 1.0 else 0.0
        ww = 0.0
        for x in range(n):  # Pillow accumulates in tap order (numpy's sum is pairwise from 8 terms up)
            ww += float(w[x])
        if ww != 0.0:
            w = w / ww
        for x in range(n):
            v = w[x] * (1 << PIL_PRECISION_BITS)
            kk[xx, x] = int(v - 0.5) if w[x] < 0 else int(v + 0.5)
        xmins[xx] = xmin
        counts[xx] = n
    return xmins, counts, kk


def _pil_pass(img, out_size, axis):
    xmins, counts, kk = pil_bilinear_coeffs(img.shape[axis], out_size)
    src = np.moveaxis(img.astype(np.int64), axis, 0)
    out = np.empty((out_size,) + src.shape[1:], np.uint8)
    for i in range(out_size):
        n = counts[i]
        acc = np.int64(1 << (PIL_PRECISION_BITS - 1)) + np.tensordot(
            kk[i, :n].astype(np.int64), src[xmins[i] : xmins[i] + n], axes=(0, 0)
        )
        out[i] = np.clip(acc >> PIL_PRECISION_BITS, 0, 255).astype(np.uint8)
    return np.moveaxis(out, 0, axis)


def pil_resize_bilinear_u8(img, out_h, out_w):
    """Bit-exact restatement of ``PIL.Image.fromarray(img).resize((out_w, out_h), BILINEAR)``:
    horizontal pass, round to uint8, vertical pass, round to uint8."""
    img = np.asarray(img)
    assert img.dtype == np.uint8 and img.ndim == 3
    out = img
    if out_w != img.shape[1]:
        out = _pil_pass(out, out_w, 1)
    if out_h != img.shape[0]:
        out = _pil_pass(out, out_h, 0)
    return out.copy() if out is img else out


def cv2_resize_linear_f32_trunc(img, out_h=224, out_w=224):
    """a6, float-frame branch -- vivit .../dataset.py:244-284: pytorchvideo hands the Dataset float32
    frames (0..255), ``cv2.resize(frame, (224, 224))`` then runs in FLOAT and ``.astype(np.uint8)``
    truncates.  Restatement: tap = floor / fraction of ``(d + 0.5) * scale - 0.5`` kept in double until the
    fraction's final cast to float32; per pass two float32 products and one add; column taps clamp with the
    fraction reset, row taps clamp without (as in the uint8 path); an exact 2x2 down-scale is INTER_AREA
    (mean of four).  PINNED LOOSELY: the installed opencv-python (4.13.0, IPP build; reference pins 4.11.0.86)
    agrees to 1 float32 ulp before truncation, i.e. the uint8 result matches on > 99.5 % of pixels and never
    differs by more than 1 LSB (tests/test_oracle_golden.py::test_cv2_float_branch)."""
    img = np.asarray(img)
    H, W = img.shape[:2]
    f32 = np.float32
    src = img.astype(f32)
    if H == 2 * out_h and W == 2 * out_w:
        m = (src[0::2, 0::2] + src[0::2, 1::2] + src[1::2, 0::2] + src[1::2, 1::2]) * f32(0.25)
        return m.astype(np.uint8)

    def taps(dst, n):
        scale = 1.0 / (np.float64(dst) / np.float64(n))
        f = (np.arange(dst, dtype=np.float64) + 0.5) * scale - 0.5
        s = np.floor(f).astype(np.int64)
        return s, (f - s).astype(f32)

    sy, fy = taps(out_h, H)
    sx, fx = taps(out_w, W)
    lo, hi = sx < 0, sx >= W - 1
    sx = np.where(lo, 0, np.where(hi, W - 1, sx))
    fx = np.where(lo | hi, f32(0), fx).astype(f32)
    x1 = np.minimum(sx + 1, W - 1)
    a0, a1 = (f32(1) - fx).astype(f32), fx
    y0, y1 = np.clip(sy, 0, H - 1), np.clip(sy + 1, 0, H - 1)
    b0, b1 = (f32(1) - fy).astype(f32), fy
    h = (src[:, sx] * a0[None, :, None]).astype(f32) + (src[:, x1] * a1[None, :, None]).astype(f32)
    out = (h[y0] * b0[:, None, None]).astype(f32) + (h[y1] * b1[:, None, None]).astype(f32)
    return out.astype(np.uint8)
