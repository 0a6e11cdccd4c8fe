"""C++ samplers (through the C ABI) against the reference's golden indices and the oracle.  CPU only."""
import random

import numpy as np
import pytest

from lvc_b200 import _lib, sampling
from oracle import samplers as osm

METHODS = ["uniform", "random_window", "random"]


def test_streams_match_cpython_and_numpy():
    for seed in (0, 1, 42, 1234, 2 ** 31, 2 ** 32 + 5, 2 ** 63 + 11):
        r, p = sampling.SamplerRng(seed, "py"), random.Random(seed)
        assert [r.random() for _ in range(700)] == [p.random() for _ in range(700)]
    for seed in (0, 42, 2 ** 32 - 1):
        r, n = sampling.SamplerRng(seed, "np"), np.random.RandomState(seed)
        assert [r.u32() for _ in range(700)] == [int(v) for v in n.randint(0, 2 ** 32, 700, dtype=np.uint64)]


def test_golden_indices(golden):
    g = golden("samplers.npz")
    state, bad = {}, []
    for (copy, m, N, T, seed, call), w in zip(g["rows"], g["indices"]):
        method, key = METHODS[m], (copy, m, N, T, seed)
        if copy in (0, 1):
            if call == 0:
                state[key] = sampling.SamplerRng(int(seed), "py")
            got = sampling.get_sampling_indices(method, N, T, state[key])
        elif copy == 2:
            got = sampling.get_sampling_indices(method, N, T, sampling.SamplerRng(42, "py"))
        elif copy == 3:
            got = sampling.get_sampling_indices(method, N, T, sampling.SamplerRng(42, "np"), _lib.VARIANT_NUMPY_INFER)
        else:
            got = sampling.get_sampling_indices(method, N, T, sampling.SamplerRng(int(seed), "py"), _lib.VARIANT_CLAMPED)
        if got != [int(v) for v in w if v >= 0]:
            bad.append((copy, method, N, T, seed, call))
    assert not bad, bad[:10]


def test_golden_sample_frame_indices(golden):
    g = golden("samplers.npz")
    st = None
    for (clip_len, rate, seg, seed, call), w in zip(g["a2_rows"], g["a2_indices"]):
        if call == 0:
            st = sampling.SamplerRng(int(seed), "np")
        assert sampling.sample_frame_indices(clip_len, rate, seg, st) == w[:clip_len].tolist()


@pytest.mark.parametrize("method", METHODS)
def test_dense_grid_against_oracle(method):
    # every N up to 400 (pool<->set switch of random.sample at 21 + 4**ceil(log4(3T))) and a few T
    for T in (1, 2, 5, 6, 8, 16, 21, 32, 64):
        for N in list(range(1, 400)) + [1000, 4097, 100000]:
            for seed in (42, 7):
                want = osm.dataset_indices(method, N, T, random.Random(seed))
                got = sampling.get_sampling_indices(method, N, T, sampling.SamplerRng(seed))
                assert got == want, (method, N, T, seed)


def test_numpy_variant_dense():
    for method in METHODS:
        for T in (1, 8, 32):
            for N in list(range(1, 200)) + [1000]:
                want = osm.inference_indices_np(method, N, T, 42)
                got = sampling.get_sampling_indices(method, N, T, sampling.SamplerRng(42, "np"), _lib.VARIANT_NUMPY_INFER)
                assert got == want, (method, N, T)


def test_uniform_temporal_subsample_matches_torch():
    for T in (1, 2, 8, 16, 32, 33):
        for t in list(range(1, 600)) + [1000, 5000]:
            assert sampling.uniform_temporal_subsample(t, T) == osm.uniform_temporal_subsample(t, T), (t, T)


def test_dynamic_fps_and_errors():
    assert sampling.dynamic_fps(300, 32, 30.0) is None
    assert sampling.dynamic_fps(5, 32, 25.0) == osm.dynamic_fps(5, 32, 25.0)
    assert sampling.dynamic_fps(5, 32, 0.0) == osm.dynamic_fps(5, 32, 0.0)
    with pytest.raises(ValueError):
        sampling.get_sampling_indices("uniform", 0, 8)
    with pytest.raises(ValueError):
        sampling.get_sampling_indices("random", 10, 8, None)
    with pytest.raises(ValueError):  # numpy raises ValueError for low >= high
        sampling.sample_frame_indices(32, 1, 32, sampling.SamplerRng(0, "np"))
    assert sampling.get_sampling_indices("no_such_method", 10, 4) == sampling.get_sampling_indices("uniform", 10, 4)
