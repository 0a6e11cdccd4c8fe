"""Row f2, host side: the explicit RNG contract of the train-time augmentation chain.
The C library's torch-stream emulation must reproduce torch's CPU generator bit for bit, and the per-clip
draws must be the ones torchvision's RandomCrop / RandomHorizontalFlip make under the same seed."""
import numpy as np
import pytest
import torch

from lvc_b200 import sampling
from oracle import processors as op


@pytest.mark.parametrize("seed", [0, 42, 1234, 2**31 + 5])
def test_torch_stream_bit_exact(seed):
    rng = sampling.TorchRng(seed)
    torch.manual_seed(seed)
    for k in range(200):
        assert rng.randint(256, 321) == torch.randint(256, 321, (1,)).item()
        assert rng.randint(0, 1 + k) == torch.randint(0, 1 + k, (1,)).item()
        assert rng.rand() == torch.rand(1).item()


@pytest.mark.parametrize("hw", [(72, 128), (128, 72), (90, 90), (224, 224), (240, 320)])
def test_train_aug_draws_match_torchvision(hw):
    h, w = hw
    frames = np.random.default_rng(5).integers(0, 256, (2, h, w, 3), dtype=np.uint8)
    rng = sampling.TorchRng(7)
    torch.manual_seed(7)
    flips = 0
    for _ in range(12):  # consecutive clips continue the same stream on both sides
        _, want = op.ptv_train_chain(frames, seed=None)
        got = sampling.train_aug_params(h, w, rng)
        assert got == want
        assert 256 <= got[0] <= 320
        flips += got[3]
    assert 0 < flips < 12


def test_train_aug_argument_checks():
    rng = sampling.TorchRng(0)
    with pytest.raises(ValueError):
        sampling.train_aug_params(100, 100, rng, min_size=200)
    with pytest.raises(ValueError):
        rng.randint(5, 5)
