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


CLI = os.path.join(ROOT, "oracle", "_ref", "fieldgenerator_cuda")


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(CLI), reason="oracle/_ref/fieldgenerator_cuda not built (no /root/reference)")
def test_reference_cli_on_the_cuda_path(tmp_path):
    """the reference's own `fieldgenerator` (src/fieldgenerator.cc, unmodified)
    with CudaMatrix as RandomField's matrix: ini file + -key value options in,
    legacy VTK out; the values are those of pf_generate for the same seed"""
    import importlib
    import sys

    import numpy as np

    sys.path.insert(0, ROOT)
    pfb = importlib.import_module("dune-parafields_b200")
    ini = tmp_path / "field.ini"
    ini.write_text("[grid]\nextensions = 1 1\ncells = 64 32\n[stochastic]\ncovariance = exponential\nvariance = 1.\ncorrLength = 0.1\n")
    out = subprocess.run([CLI, str(ini), "-input.seed", "7", "-output.legacyVtk", str(tmp_path / "f")], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300, cwd=tmp_path)
    print(out.stdout)
    assert out.returncode == 0, out.stdout
    files = [f for f in os.listdir(tmp_path) if f.startswith("f") and "vtk" in f]
    assert files, os.listdir(tmp_path)
    vals = []
    started = False
    for line in open(tmp_path / files[0]):
        if started:
            vals += [float(x) for x in line.split()]
        elif line.startswith("LOOKUP_TABLE"):
            started = True
    vals = np.array(vals)
    assert vals.size == 64 * 32
    with pfb.Plan([64, 32], covariance="exponential", corr_length=0.1) as p:
        ref = p.generate(seed=7, realization=0).ravel()
    assert np.max(np.abs(vals - ref)) < 1e-4 * np.max(np.abs(ref))  # ASCII output precision


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(CLI), reason="oracle/_ref/fieldgenerator_cuda not built (no /root/reference)")
def test_reference_cli_merge_periodization(tmp_path):
    """the ini keys of the merge periodisation (embedding.periodization,
    embedding.sigmoid, stochastic.sigmoidCombine, matrix.hh:530-590,669-692)
    reach the CUDA path through the unmodified CLI"""
    import importlib
    import sys

    import numpy as np

    sys.path.insert(0, ROOT)
    pfb = importlib.import_module("dune-parafields_b200")
    ini = tmp_path / "field.ini"
    ini.write_text(
        "[grid]\nextensions = 1 1\ncells = 64 64\n[stochastic]\ncovariance = exponential\nvariance = 1.\ncorrLength = 0.5\n"
        "sigmoidCombine = radial\n[embedding]\nfactor = 4\nperiodization = merge\nsigmoid = smooth\n"
    )
    out = subprocess.run([CLI, str(ini), "-input.seed", "9", "-output.legacyVtk", str(tmp_path / "f")], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300, cwd=tmp_path)
    print(out.stdout)
    assert out.returncode == 0, out.stdout
    files = [f for f in os.listdir(tmp_path) if f.startswith("f") and "vtk" in f]
    vals = []
    started = False
    for line in open(tmp_path / files[0]):
        if started:
            vals += [float(x) for x in line.split()]
        elif line.startswith("LOOKUP_TABLE"):
            started = True
    vals = np.array(vals)
    with pfb.Plan([64, 64], covariance="exponential", corr_length=0.5, factor=4, periodization="merge", sigmoid="smooth", sigmoid_combine="radial") as p:
        ref = p.generate(seed=9, realization=0).ravel()
    assert np.max(np.abs(vals - ref)) < 1e-4 * np.max(np.abs(ref))  # ASCII output precision
    # the classical embedding of this configuration is refused (negative eigenvalues)
    ini.write_text(ini.read_text().replace("periodization = merge", "periodization = classical"))
    out = subprocess.run([CLI, str(ini), "-input.seed", "9", "-output.legacyVtk", str(tmp_path / "g")], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300, cwd=tmp_path)
    assert out.returncode != 0 and "negative eigenvalues" in out.stdout
