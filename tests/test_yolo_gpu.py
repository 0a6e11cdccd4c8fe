"""Row f4: detector pre / post-processing kernels against the oracle (restated yolov5 letterbox /
scale_boxes around the reference's own arithmetic, src/yolo_detector.py:52-61, :76-81).  uint8 resize bit-exact
(cv2), float32 network input bit-exact vs torch CPU ``float() / 255``, integer boxes exact."""
import numpy as np
import pytest
import torch

from oracle import yolo as oy

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("hw", [(720, 1280), (1080, 1920), (480, 640), (640, 480), (100, 150), (640, 640), (333, 517)])
def test_letterbox_batch_matches_reference_preprocessing(hw):
    from lvc_b200 import yolo

    h, w = hw
    frames = np.random.default_rng(h + w).integers(0, 256, (3, h, w, 3), dtype=np.uint8)
    im, g = yolo.letterbox_batch(torch.from_numpy(frames).cuda())
    want = torch.cat([oy.preprocess(f) for f in frames])
    assert tuple(im.shape) == tuple(want.shape)
    assert torch.equal(im.cpu(), want)
    bf, _ = yolo.letterbox_batch(torch.from_numpy(frames).cuda(), dtype=torch.bfloat16)
    assert torch.equal(bf.cpu(), want.to(torch.bfloat16))


@pytest.mark.parametrize("hw", [(64, 20000), (2160, 3840), (1080, 1921), (31, 9001)])
def test_letterbox_wide_and_odd_frames(hw):
    """Rows too wide for the staged kernel's shared-memory slots take the direct-global kernel (20000 px = 60 KB per row,
    eight slots do not fit); 4K rows get one block per SM; a width that is not a multiple of 16 bytes exercises the
    aligned-superset bulk copies and the hand-copied tail of the last frame's last row."""
    from lvc_b200 import yolo

    h, w = hw
    frames = np.random.default_rng(h * 7 + w).integers(0, 256, (2, h, w, 3), dtype=np.uint8)
    im, g = yolo.letterbox_batch(torch.from_numpy(frames).cuda())
    want = torch.cat([oy.preprocess(f) for f in frames])
    assert tuple(im.shape) == tuple(want.shape)
    assert torch.equal(im.cpu(), want)


def test_scale_boxes_and_best_boxes_feed_the_clip_kernel():
    from lvc_b200 import ClipBuilder, yolo
    from oracle import processors as op

    h, w, n_frames = 720, 1280, 6
    g = yolo.letterbox_geometry(h, w)
    rng = np.random.default_rng(9)
    n = 40
    xy = rng.uniform(-20, 660, (n, 2)).astype(np.float32)
    wh = rng.uniform(5, 300, (n, 2)).astype(np.float32)
    det = np.concatenate([xy, xy + wh, rng.uniform(0.05, 0.99, (n, 1)).astype(np.float32), np.zeros((n, 1), np.float32)], 1)
    det[:8, :4] = np.round(det[:8, :4]) + 0.5  # ties of round-half-to-even after the gain
    frame_of = rng.integers(0, n_frames - 1, n).astype(np.int32)  # the last frame has no detection
    det_d, fo_d = torch.from_numpy(det).cuda(), torch.from_numpy(frame_of).cuda()
    got = yolo.scale_boxes(det_d, g).cpu().numpy()
    want = np.array(oy.detections_to_int_boxes(torch.from_numpy(det), (g.out_h, g.out_w), (h, w, 3)))
    assert np.array_equal(got, want)

    best = yolo.best_boxes(det_d, fo_d, n_frames, g, conf_thres=0.25)
    wb = np.full((n_frames, 4), -1, np.int64)
    for f in range(n_frames):
        rows = [i for i in range(n) if frame_of[i] == f and det[i, 4] >= 0.25]
        if rows:
            wb[f] = want[max(rows, key=lambda i: (det[i, 4], -i))]
    assert np.array_equal(best.cpu().numpy(), wb)
    assert (wb[-1] == -1).all()

    # the table is consumed on the device: one clip whose frames are cropped by these boxes
    video = rng.integers(0, 256, (n_frames, h, w, 3), dtype=np.uint8)
    idx = list(range(n_frames))
    out = ClipBuilder("dataset_u8", num_frames=n_frames)([torch.from_numpy(video).cuda()], [idx], boxes=best.view(1, n_frames, 4))
    ref = op.resize_clip_cv2(video, boxes=wb)
    assert np.array_equal(out[0].cpu().numpy(), ref)
