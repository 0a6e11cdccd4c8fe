"""Bounded randomised parity sweep (tests/fuzz_parity.py): random frame sizes, out-of-range indices, ROI boxes that
stick out of the frame, all presets, RGB / BGR sources -- every case bit-exact against the oracle (float chain within
the tolerance of tests/test_clip_gpu.py)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu


def test_randomised_parity_sweep():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tests", "fuzz_parity.py"), "40", "7"], capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert "40/40 cases match" in r.stdout


def test_randomised_embed_sweep():
    """Ragged token counts (partial 128/256-row tiles), both dtypes, plain and fused-embeddings epilogues."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tests", "fuzz_embed.py"), "24", "11"], capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert "24/24 cases match" in r.stdout


def test_randomised_detector_pre_post_sweep():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tests", "fuzz_yolo.py"), "40", "13"], capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert "40/40 cases match" in r.stdout


def test_randomised_raw_processor_sweep():
    """Random frame sizes (1 px to 4K, portrait, odd), mixed per batch, both HF presets, edges 224/256/300."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tests", "fuzz_hfraw.py"), "30", "17"], capture_output=True,
                       text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert "30/30 cases match" in r.stdout
