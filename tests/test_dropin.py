"""Drop-in proof: the UNMODIFIED reference parafields::RandomField driving
this repository's CudaMatrix plugin next to its own CPU matrices
(oracle/dropin_test.cpp, built into oracle/_ref/dropin_test where the
reference checkout exists; the binary travels to the GPU box)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "dropin_test")


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(BIN), reason="oracle/_ref/dropin_test not built (no /root/reference)")
def test_reference_randomfield_with_cuda_matrix():
    out = subprocess.run([BIN], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    print(out.stdout)
    assert out.returncode == 0, out.stdout
    assert "DROP-IN OK" in out.stdout
    assert "FAIL" not in out.stdout
