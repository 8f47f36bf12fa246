"""Randomised cross-check of the TF32 tensor-core mode against the exact FP32 mode of the library itself (scripts/fuzz_modes.py):
random level counts, odd sizes, channel counts with and without padding, batch sizes, both up-sampling methods, refinement."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_tf32_mode_tracks_fp32_mode_on_random_configs():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "fuzz_modes.py"), "0", "24"], cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("case")]
    assert len(lines) == 24
    assert not [l for l in lines if "CHECK" in l], "\n".join(l for l in lines if "CHECK" in l)
