"""Row f4, host side: letterbox geometry of the C library == the restated yolov5 letterbox (oracle/yolo.py)."""
import numpy as np
import pytest

from lvc_b200 import yolo
from oracle import yolo as oy

SIZES = [(720, 1280), (1080, 1920), (480, 640), (640, 480), (100, 150), (640, 640), (333, 517), (1, 1), (405, 811)]


@pytest.mark.parametrize("hw", SIZES)
@pytest.mark.parametrize("stride", [32, 64])
def test_letterbox_geometry_matches_restated_yolov5(hw, stride):
    h, w = hw
    g = yolo.letterbox_geometry(h, w, (640, 640), stride)
    im, ratio, (dw, dh) = oy.letterbox(np.zeros((h, w, 3), np.uint8), (640, 640), stride=stride)
    assert (g.out_h, g.out_w) == im.shape[:2]
    assert (g.new_w, g.new_h) == (int(round(w * ratio[0])), int(round(h * ratio[1])))
    assert (g.top, g.left) == (int(round(dh - 0.1)), int(round(dw - 0.1)))
    assert g.out_h % stride == 0 and g.out_w % stride == 0


def test_letterbox_geometry_argument_checks():
    with pytest.raises(ValueError):
        yolo.letterbox_geometry(0, 10)
    with pytest.raises(ValueError):
        yolo.letterbox_geometry(10, 10, stride=0)
