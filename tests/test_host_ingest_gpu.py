"""Parity of the path the headline end-to-end number is measured on: ``ClipBuilder.build_from_host`` with the decoded
frames in PINNED HOST memory -- the fused kernel's bulk-copy producer reads the tapped rows over PCIe (zero-copy) -- and
its staged ``zero_copy=False`` branch.  Both must equal the device-resident result byte for byte AND the oracle's
restatement of the reference recipe (resnet50-2d-lstm/inference.py:158-178; vivit .../trainers/trainer.py:62-95;
vivit .../dataset.py:271-292)."""
import numpy as np
import pytest
import torch

from oracle import processors as op
from oracle import samplers as osm

pytestmark = pytest.mark.gpu

# ragged batch: 720p, 1080p, odd sizes, a one-row video, an exact-2x video (INTER_AREA switch)
SPECS = [(21, 5, 720, 1280, "noise"), (22, 4, 1080, 1920, "noise"), (23, 6, 449, 447, "ramp"), (24, 3, 37, 1000, "noise"),
         (25, 4, 448, 448, "noise"), (26, 3, 1, 5, "noise"), (27, 5, 223, 225, "ramp")]


def _host_videos(specs, exact_tail=False):
    """Pinned host tensors.  exact_tail: each tensor's storage ends with its last row (no slack behind the buffer), so
    the producer's hand-copied tail (last bytes that are not a whole 16-byte granule) is exercised."""
    from lvc_b200.synth import synth_video

    arrs = [synth_video(*s) for s in specs]
    pinned = []
    for a in arrs:
        t = torch.empty(a.shape, dtype=torch.uint8).pin_memory()
        t.copy_(torch.from_numpy(a))
        pinned.append(t)
    return arrs, pinned


def _oracle(preset, frames):
    r = op.resize_clip_cv2(frames)
    if preset == "dataset_u8":
        return r
    if preset == "resnet_infer":
        return op.manual_normalize(r)
    return op.vivit_processor(r)


@pytest.mark.parametrize("preset", ["resnet_infer", "vivit", "dataset_u8"])
@pytest.mark.parametrize("zero_copy", [True, False])
def test_build_from_host_equals_device_resident_and_oracle(preset, zero_copy):
    from lvc_b200 import ClipBuilder

    T = 6
    arrs, pinned = _host_videos(SPECS)
    dev = [torch.from_numpy(a).cuda() for a in arrs]
    rng = __import__("random").Random(5)
    idx = [osm.dataset_indices("random", a.shape[0], T, rng) for a in arrs]
    cb = ClipBuilder(preset, num_frames=T)
    stats_d = torch.zeros(len(arrs), dtype=torch.int64, device="cuda")
    stats_h = torch.zeros_like(stats_d)
    want = cb(dev, idx, stats=stats_d)
    got = cb.build_from_host(pinned, idx, stats=stats_h, zero_copy=zero_copy)
    torch.cuda.synchronize()
    assert torch.equal(got, want), (preset, zero_copy)
    assert torch.equal(stats_d, stats_h)
    assert cb.last_h2d_bytes > 0
    got = got.cpu().numpy()
    for b, a in enumerate(arrs):
        ref = _oracle(preset, a[idx[b]])
        assert np.array_equal(got[b], ref), (preset, zero_copy, SPECS[b])
    from lvc_b200.clips import device_status

    assert device_status() == set()


@pytest.mark.parametrize("zero_copy", [True, False])
def test_build_from_host_with_roi_boxes(zero_copy):
    from lvc_b200 import ClipBuilder
    from lvc_b200.synth import synth_boxes

    T = 5
    specs = SPECS[:3]
    arrs, pinned = _host_videos(specs)
    idx = [osm.dataset_indices("uniform", a.shape[0], T) for a in arrs]
    boxes = np.stack([synth_boxes(2000 + v, a.shape[0], a.shape[1], a.shape[2])[np.asarray(idx[v])] for v, a in enumerate(arrs)])
    cb = ClipBuilder("dataset_u8", num_frames=T)
    got = cb.build_from_host(pinned, idx, boxes=boxes, zero_copy=zero_copy).cpu().numpy()
    for b, a in enumerate(arrs):
        want = op.resize_clip_cv2(a[idx[b]], boxes[b])
        assert np.array_equal(got[b], want), (zero_copy, specs[b])


def test_zero_copy_last_row_ends_the_allocation():
    """A pinned allocation whose last byte is the video's last byte, at sizes whose row length is not a multiple of 16:
    the producer copies the trailing bytes by hand instead of reading a whole 16-byte granule past the buffer."""
    from lvc_b200 import ClipBuilder

    for (n, h, w) in [(2, 33, 21), (1, 224, 225), (3, 50, 333)]:
        nbytes = n * h * w * 3
        raw = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
        g = torch.Generator().manual_seed(nbytes)
        raw.copy_(torch.randint(0, 256, (nbytes,), dtype=torch.uint8, generator=g))
        vid = raw.view(n, h, w, 3)
        idx = [[n - 1, 0, n - 1]]
        got = ClipBuilder("dataset_u8", num_frames=3).build_from_host([vid], idx, zero_copy=True).cpu().numpy()
        want = op.resize_clip_cv2(vid.numpy()[idx[0]])
        assert np.array_equal(got[0], want), (n, h, w)


def test_unpinned_host_frames_take_the_staged_branch():
    from lvc_b200 import ClipBuilder

    arrs, _ = _host_videos(SPECS[:2])
    pageable = [torch.from_numpy(a) for a in arrs]
    idx = [[0, 1, 2, 3], [3, 2, 1, 0]]
    cb = ClipBuilder("resnet_infer", num_frames=4)
    got = cb.build_from_host(pageable, idx).cpu().numpy()   # zero_copy defaults to False for pageable memory
    for b, a in enumerate(arrs):
        assert np.array_equal(got[b], op.manual_normalize(op.resize_clip_cv2(a[idx[b]])))
    # distinct frames are uploaded once each
    assert cb.last_h2d_bytes == sum(4 * a[0].size for a in arrs)
    with pytest.raises(ValueError):
        cb.build_from_host(pageable, idx, zero_copy=True)   # zero-copy needs pinned memory
