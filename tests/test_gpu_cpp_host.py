"""The C++ host binding (include/prt_rt_reference.hpp) driven by a compiled C++ program, no Python between it and the C ABI."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BINARY = os.path.join(ROOT, "tests", "cpp", "test_abi")


def _binary():
    if not os.path.exists(BINARY):
        raise RuntimeError(f"{BINARY} is missing: __graft_entry__.build() compiles it")
    return BINARY


@pytest.mark.gpu
def test_cpp_host_renders_through_the_abi():
    r = subprocess.run([_binary()], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "cpp host ok" in r.stdout


def test_cpp_host_fails_loudly_without_a_device():
    """No CPU fallback: on a machine without a CUDA device the wrapper's World constructor throws prt_last_error()."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    r = subprocess.run([_binary()], capture_output=True, text=True, timeout=60)
    assert r.returncode == 2 and "no CUDA device" in r.stderr
