"""Oracle pin for the HF processors on RAW frames (any size): oracle/processors.py against transformers + Pillow
installed here (reference call sites data_preprocessing.py:16,
data_preprocessing_timesformer.py:16, vivit_transformer/inference.py:187)."""
import numpy as np
import pytest

from oracle import processors as op
from oracle.resize import pil_resize_bilinear_u8

SHAPES = [(720, 1280), (480, 640), (300, 200), (256, 256), (231, 517), (1080, 1920), (224, 224), (1100, 240)]


@pytest.mark.parametrize("shape", SHAPES)
def test_oracle_vs_hf_vivit_raw(shape):
    tr = pytest.importorskip("transformers")
    rng = np.random.default_rng(hash(shape) % 2**32)
    frames = rng.integers(0, 256, (2,) + shape + (3,), dtype=np.uint8)
    ref = tr.VivitImageProcessor()(list(frames), return_tensors="np")["pixel_values"][0]
    assert np.array_equal(op.vivit_processor(frames), ref)


@pytest.mark.parametrize("shape", SHAPES)
def test_oracle_vs_hf_videomae_raw(shape):
    tr = pytest.importorskip("transformers")
    rng = np.random.default_rng(hash(shape) % 2**32 + 1)
    frames = rng.integers(0, 256, (2,) + shape + (3,), dtype=np.uint8)
    mean, std = (0.45, 0.45, 0.45), (0.225, 0.225, 0.225)
    # The reference pins transformers 4.48.2, whose VideoMAEImageProcessor is the numpy + Pillow pipeline
    # (resize -> center_crop -> rescale -> normalize).  The transformers installed here (5.x) moved that class to a
    # torchvision backend with different rounding, but still ships the legacy pipeline as VivitImageProcessor:
    # with offset=False and these sizes it is step for step the 4.48.2 VideoMAE chain.
    proc = tr.VivitImageProcessor(size={"shortest_edge": 224}, crop_size={"height": 224, "width": 224}, offset=False,
                                  rescale_factor=1 / 255, image_mean=list(mean), image_std=list(std))
    ref = proc(list(frames), return_tensors="np")["pixel_values"][0]
    assert np.array_equal(op.videomae_processor_raw(frames, mean, std), ref)


def test_pil_many_taps_vs_pillow():
    """Down-scales with >= 8 taps per sample (where a pairwise weight sum would differ from Pillow's)."""
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(5)
    for (h, w), (oh, ow) in [((1080, 1920), (256, 455)), ((2160, 3840), (224, 398)), ((901, 1203), (224, 299))]:
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        ref = np.asarray(Image.fromarray(img).resize((ow, oh), Image.BILINEAR))
        assert np.array_equal(pil_resize_bilinear_u8(img, oh, ow), ref)


@pytest.mark.parametrize("shape", [(720, 1280), (300, 200), (231, 517)])
def test_oracle_vs_hf_exact_224(shape):
    """size={"height": 224, "width": 224} on frames of another size (aspect-changing Pillow resize)."""
    tr = pytest.importorskip("transformers")
    frames = np.random.default_rng(shape[0]).integers(0, 256, (2,) + shape + (3,), dtype=np.uint8)
    mean, std = (0.45, 0.45, 0.45), (0.225, 0.225, 0.225)
    proc = tr.VivitImageProcessor(size={"height": 224, "width": 224}, crop_size={"height": 224, "width": 224},
                                  offset=False, rescale_factor=1 / 255, image_mean=list(mean), image_std=list(std))
    ref = proc(list(frames), return_tensors="np")["pixel_values"][0]
    assert np.array_equal(op.videomae_processor_exact(frames, mean, std), ref)
