"""The C-ABI driven from plain C (tests/c_abi_smoke.c, gcc): context -> model -> forward -> train steps -> interp/adjoint on
the reference's golden vectors, without Python, torch or C++ in the calling process."""
import os
import subprocess
import pytest


def _harness():
    import __graft_entry__ as ge
    return ge.build_c_harness(), ge.LIB


@pytest.mark.gpu
def test_c_harness_runs_the_hot_path():
    exe, lib = _harness()
    r = subprocess.run([exe, lib], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "c_abi_smoke: OK" in r.stdout


def test_c_harness_builds_and_fails_loudly_without_gpu():
    import torch
    exe, lib = _harness()
    assert os.path.exists(exe)
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    r = subprocess.run([exe, lib], capture_output=True, text=True, timeout=60)
    assert r.returncode == 3 and "no CPU path" in r.stderr          # no fallback behind the header
