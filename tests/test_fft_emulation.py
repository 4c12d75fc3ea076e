"""Host emulation of the warp-level FFT core (csrc/pf_fft.cuh): the same
templates the kernels use, compiled for the host, lanes executed
sequentially, compared with a naive long-double DFT for N = 2..1024."""
import os
import shutil
import subprocess

import pytest

from util import ROOT


@pytest.mark.skipif(shutil.which("nvcc") is None, reason="nvcc not available")
def test_line_fft_host_emulation(tmp_path):
    exe = str(tmp_path / "fft_emul")
    src = os.path.join(ROOT, "tests", "host", "fft_emul.cu")
    inc = os.path.join(ROOT, "dune-parafields_b200", "csrc")
    subprocess.check_call(
        ["nvcc", "-std=c++17", "-O1", "-Wno-deprecated-gpu-targets", "-I", inc, "-o", exe, src]
    )
    out = subprocess.run([exe], stdout=subprocess.PIPE, text=True)
    assert out.returncode == 0, out.stdout
    assert "OK" in out.stdout
