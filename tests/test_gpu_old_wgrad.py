"""The earlier weight-gradient kernel generations stay selectable (DINCAE_WGRAD9=0, DINCAE_WGRADM=0); the switches are read once
per process, so the bit-exact kernel-family test is re-run in a child process with them set."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_previous_wgrad_generations_bit_exact():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    env = dict(os.environ, DINCAE_WGRAD9="0", DINCAE_WGRADM="0")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_gpu_kernels.py"), "-m", "gpu", "-q", "-x",
                        "-k", "wgrad"], env=env, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "passed" in r.stdout
