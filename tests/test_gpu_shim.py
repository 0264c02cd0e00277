"""The C++ drop-in shim (sdmm_mitsuba_b200/csrc/host_shim.hpp) driven with the REFERENCE's own types:
oracle/_ref/shim_check fits identical SDMMContext objects with the reference's optimize() loop body on
the CPU and with sdmm_b200::optimize on the GPU and compares them member by member (1e-3)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "shim_check")


def _run(*args):
    env = dict(os.environ)
    return subprocess.run([BIN] + [str(a) for a in args], capture_output=True, text=True, timeout=300, env=env)


@pytest.mark.gpu
@pytest.mark.parametrize("K,C,N", [(16, 6, 700), (64, 5, 1500), (24, 3, 400)])
def test_shim_optimize_matches_reference_objects(K, C, N):
    if not os.path.exists(BIN):
        pytest.skip("oracle/_ref/shim_check not built (needs /root/reference at build time)")
    r = _run(K, C, N)
    assert r.returncode == 0 and "PASS" in r.stdout, r.stdout + r.stderr


def test_shim_fails_loudly_without_gpu():
    """No CPU fallback behind the shim: without a device it throws (exit 2), it does not compute."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    if not os.path.exists(BIN):
        pytest.skip("oracle/_ref/shim_check not built")
    r = _run(16, 2, 100)
    assert r.returncode == 2 and "CUDA runtime error" in r.stdout
