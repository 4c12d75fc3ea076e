"""On-disk formats.  The reference's OWN I/O layer --
RandomField::writeToFile (randomfield.hh:656-663: `.stoch.h5` through
writeParallelToHDF5 io.hh:171-286, the XDMF sidecar io.hh:306-402, `.trend`
trend.hh:1517-1530, the `.field` report) and the file-reading constructor
(randomfield.hh:190-225, stochastic.hh:80-107) -- runs unmodified on the
dependency-free HDF5 stand-in of dune-parafields_b200/host/hdf5_npy/hdf5.h,
which stores the values as a NumPy .npy file in the reference's dimension order.

CPU part: the reference's CPU build (oracle/_ref) writes, numpy and the
reference's constructor read back.  GPU part (-m gpu): the reference's CLI on
the CUDA path writes `-output.dune`, numpy reads what pf_generate produced."""
import os
import subprocess

import numpy as np
import pytest

from oracle import ref_lib as rl
from util import ROOT

CASES = [
    ([24], {}),
    ([16, 12], dict(extensions=[2.0, 0.5])),
    ([8, 6, 10], {}),
]


@pytest.mark.skipif(not rl.available(), reason="oracle/_ref not built (no /root/reference)")
@pytest.mark.parametrize("cells,kw", CASES, ids=["1d", "2d", "3d"])
def test_reference_files_round_trip_on_the_npy_container(tmp_path, cells, kw):
    trend = "[mean]\nmean = 0.7\nvariance = 0.01\n"
    r = rl.Reference(cells, covariance="exponential", corr_length=0.1, extra=trend, **kw)
    f = r.generate(5)
    base = str(tmp_path / "field")
    r.write(base)
    for ext in (".stoch.h5", ".xdmf", ".field", ".trend"):
        assert os.path.exists(base + ext), ext
    # the values, in the reference's dimension order (slowest index first)
    v = np.load(base + ".stoch.h5")
    assert v.dtype == np.float64 and v.shape == tuple(reversed(cells))
    assert np.array_equal(v, f)
    # the XDMF sidecar names the file and its shape (io.hh:343-350)
    x = open(base + ".xdmf").read()
    assert os.path.basename(base) + ".stoch.h5:/stochastic" in x
    assert 'Dimensions="%s "' % " ".join(str(c) for c in reversed(cells)) in x
    # the .field report holds the configuration the constructor re-reads
    rep = open(base + ".field").read()
    assert 'cells = "%s"' % " ".join(map(str, cells)) in rep and 'covariance = "exponential"' in rep
    # read back through the reference's own constructor
    back = rl.Reference.read(base, cells)
    assert np.array_equal(back.get(), f)
    # RandomField::operator* adds the trend part's product: the .trend file
    # round-trips too
    # (the reference prints trend coefficients with 6 significant digits)
    assert abs(back.dot_self() - r.dot_self()) <= 1e-5 * r.dot_self()


def test_npy_header_is_plain_numpy(tmp_path):
    """a file written by numpy itself is accepted by the stand-in's parser --
    checked through the reference's reading constructor when it is available"""
    if not rl.available():
        pytest.skip("oracle/_ref not built (no /root/reference)")
    cells = [6, 4]
    r = rl.Reference(cells, covariance="exponential", corr_length=0.1)
    r.generate(1)
    base = str(tmp_path / "f")
    r.write(base)
    new = np.arange(24, dtype=np.float64).reshape(4, 6) / 7.0
    with open(base + ".stoch.h5", "wb") as fh:
        np.save(fh, new)
    assert np.array_equal(rl.Reference.read(base, cells).get(), new)


CLI = os.path.join(ROOT, "dune-parafields_b200", "cli", "_build", "fieldgenerator_cuda")


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(CLI), reason="dune-parafields_b200/cli/_build/fieldgenerator_cuda not built (no /root/reference)")
def test_cli_writes_field_files_from_the_cuda_path(tmp_path):
    """`-output.dune <base>` of the reference's CLI (src/fieldgenerator.cc:145)
    with CudaMatrix: the values on disk are pf_generate's, bit for bit"""
    import importlib
    import sys

    sys.path.insert(0, ROOT)
    pfb = importlib.import_module("dune-parafields_b200")
    ini = tmp_path / "field.ini"
    ini.write_text("[grid]\nextensions = 1 1 1\ncells = 32 16 8\n[stochastic]\ncovariance = exponential\nvariance = 1.\ncorrLength = 0.1\n")
    out = subprocess.run([CLI, str(ini), "-input.seed", "11", "-output.dune", str(tmp_path / "f")], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300, cwd=tmp_path)
    print(out.stdout)
    assert out.returncode == 0, out.stdout
    for ext in (".stoch.h5", ".xdmf", ".field", ".trend"):
        assert os.path.exists(str(tmp_path / "f") + ext), ext
    v = np.load(tmp_path / "f.stoch.h5")
    assert v.shape == (8, 16, 32)
    with pfb.Plan([32, 16, 8], covariance="exponential", corr_length=0.1, rng_mode="fp64") as p:
        ref = p.generate(seed=11, realization=0)
    assert np.array_equal(v, ref)
