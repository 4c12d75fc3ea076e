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


BIN_HOST = os.path.join(ROOT, "oracle", "_ref", "dropin_test_host")


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(BIN_HOST), reason="oracle/_ref/dropin_test_host not built (no /root/reference)")
def test_reference_randomfield_with_the_reference_host_stochastic_part():
    """the same program with PARAFIELDS_CUDA_HOST_STOCHASTIC: CudaMatrix under the
    reference's own host StochasticPart (values staged per call)"""
    out = subprocess.run([BIN_HOST], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert out.returncode == 0, out.stdout
    assert "DROP-IN OK" in out.stdout and "FAIL" not in out.stdout


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(BIN), reason="oracle/_ref/dropin_test not built (no /root/reference)")
def test_operator_chain_keeps_fields_on_the_device():
    """timesMatrix -> timesInverseMatrix -> axpy, *=, scalar product through
    the reference's RandomField: with the StochasticPart specialisation of
    cudastochastic.hh the host mirror of the field is never filled"""
    import json

    out = subprocess.run([BIN, "chain", "64", "2"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["fields"] == "device-resident StochasticPart, host mirror never filled"
    assert line["chain_ms"] > 0


# product binary: the reference's src/fieldgenerator.cc linked against
# libparafields_cuda.so and the SIZES-ONLY FFTW stand-in (every CPU transform
# entry point aborts), dune-parafields_b200/cli/Makefile
CLI = os.path.join(ROOT, "dune-parafields_b200", "cli", "_build", "fieldgenerator_cuda")


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(CLI), reason="dune-parafields_b200/cli/_build/fieldgenerator_cuda not built (no /root/reference)")
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
    with pfb.Plan([64, 32], covariance="exponential", corr_length=0.1, rng_mode="fp64") as p:
        ref = p.generate(seed=7, realization=0).ravel()
    assert np.max(np.abs(vals - ref)) < 1e-4 * np.max(np.abs(ref))  # ASCII output precision


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(CLI), reason="dune-parafields_b200/cli/_build/fieldgenerator_cuda not built (no /root/reference)")
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
    with pfb.Plan([64, 64], covariance="exponential", corr_length=0.5, factor=4, periodization="merge", sigmoid="smooth", sigmoid_combine="radial", rng_mode="fp64") as p:
        ref = p.generate(seed=9, realization=0).ravel()
    assert np.max(np.abs(vals - ref)) < 1e-4 * np.max(np.abs(ref))  # ASCII output precision
    # the classical embedding of this configuration is refused (negative eigenvalues)
    ini.write_text(ini.read_text().replace("periodization = merge", "periodization = classical"))
    out = subprocess.run([CLI, str(ini), "-input.seed", "9", "-output.legacyVtk", str(tmp_path / "g")], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300, cwd=tmp_path)
    assert out.returncode != 0 and "negative eigenvalues" in out.stdout


def _read_legacy_vtk(path):
    import numpy as np

    vals, started = [], False
    for line in open(path):
        if started:
            vals += [float(x) for x in line.split()]
        elif line.startswith("LOOKUP_TABLE"):
            started = True
    return np.array(vals)


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(CLI), reason="dune-parafields_b200/cli/_build/fieldgenerator_cuda not built (no /root/reference)")
def test_reference_cli_list_mode(tmp_path):
    """randomField.types: the reference's RandomFieldList (randomfield.hh:
    1455-1500, src/fieldgenerator.cc:178-248) on the CUDA path.  List mode of
    the CLI cannot be seeded (its generate(seed) binds the seed to a bool,
    SURVEY Appendix C.5), so the check is structural and statistical: one
    legacy VTK file per field name, each with the configured marginal variance
    and correlation, the `fast` noise option and the fast/fp64 switch honoured."""
    import numpy as np

    (tmp_path / "list.ini").write_text("[grid]\nextensions = 1 1\ncells = 256 256\n[randomField]\ntypes = smooth rough\n")
    (tmp_path / "smooth.field").write_text("[stochastic]\ncovariance = gaussian\nvariance = 4.\ncorrLength = 0.02\n[embedding]\napproximate = 1\n")
    (tmp_path / "rough.field").write_text("[stochastic]\ncovariance = exponential\nvariance = 0.25\ncorrLength = 0.01\n[cuda]\nrng = fast\n")
    out = subprocess.run([CLI, str(tmp_path / "list.ini"), "-input.seed", "3", "-output.legacyVtk", str(tmp_path / "f")], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300, cwd=tmp_path)
    print(out.stdout)
    assert out.returncode == 0, out.stdout
    fields = {}
    for name in ("smooth", "rough"):
        files = [f for f in os.listdir(tmp_path) if f.startswith("f." + name) and "vtk" in f]
        assert files, os.listdir(tmp_path)
        fields[name] = _read_legacy_vtk(tmp_path / files[0]).reshape(256, 256)
    h = 1.0 / 256
    # ~2500 (smooth) / ~10000 (rough) correlation areas: the sample variance is
    # within 10 % of the target, the lag-1 correlation follows the model
    for name, var, corr1 in (("smooth", 4.0, np.exp(-((h / 0.02) ** 2))), ("rough", 0.25, np.exp(-h / 0.01))):
        f = fields[name]
        assert np.isfinite(f).all()
        v = f.var()
        assert abs(v / var - 1.0) < 0.1, (name, v)
        c1 = np.mean(f[:, :-1] * f[:, 1:]) / v
        assert abs(c1 - corr1) < 0.05, (name, c1, corr1)
    assert abs(np.corrcoef(fields["smooth"].ravel(), fields["rough"].ravel())[0, 1]) < 0.1
