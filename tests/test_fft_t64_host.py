"""The scaled-tangent radix-64 transform of the frame kernel (csrc/fft4096_t64.cuh) compiled for the HOST and checked
against a double-precision DFT: full transform, empty guard band, one tone, and the two pruned forms the frame kernel
uses (level A of the first transform, last level of the second).  No GPU needed."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("nvcc") is None and not os.path.isfile("/usr/local/cuda/bin/nvcc"), reason="needs nvcc")
def test_t64_transform_on_the_host(tmp_path):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    exe = str(tmp_path / "fft_t64_hostcheck")
    src = os.path.join(ROOT, "tests", "host", "fft_t64_hostcheck.cu")
    res = subprocess.run([nvcc, "-O1", "-std=c++17", "-o", exe, src], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    out = subprocess.run([exe, "5"], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "worst" in out.stdout
