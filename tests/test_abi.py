"""CPU-side checks of the drop-in boundary: the C-ABI library loads and
exports every symbol include/parafields_cuda.h declares; without a GPU the
compute entry points fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import pytest

from util import ROOT, pfb


def header_symbols():
    text = open(os.path.join(ROOT, "include", "parafields_cuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pf_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    m = pfb()
    L = m.lib()
    syms = header_symbols()
    assert len(syms) >= 30
    for s in syms:
        assert hasattr(L, s), "missing symbol %s" % s
    # the Python binding's list and the header agree
    assert sorted(m.ABI_SYMBOLS) == syms


def test_version_and_desc_default():
    m = pfb()
    assert b"sm_100a" in m.lib().pf_version()
    d = m.PfDesc()
    m.lib().pf_desc_default(C.byref(d))
    assert d.embedding_factor == 2 and d.threshold == 1e-14 and d.cg_iterations == 100


def test_no_cpu_fallback_without_gpu():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    m = pfb()
    with pytest.raises(m.PfError) as e:
        m.Plan([16, 16])
    assert e.value.code == m.PF_ERR_CUDA
    assert "no CPU" in str(e.value) or "CUDA" in str(e.value)


def test_argument_validation_messages():
    m = pfb()
    # validation happens before any device work, with the reference's texts
    d = m.make_desc([8, 8, 8], covariance="cubic")
    h = C.c_void_p()
    rc = m.lib().pf_plan_create(C.byref(d), C.byref(h))
    assert rc == m.PF_ERR_INVALID
    assert m.lib().pf_last_error() == b"cubic covariance only applicable in 1D or 2D"
    d = m.make_desc([8, 8], covariance="gammaExponential", cov_param=(3.0, 0.0))
    rc = m.lib().pf_plan_create(C.byref(d), C.byref(h))
    assert rc == m.PF_ERR_INVALID
    assert m.lib().pf_last_error() == b"exponent gamma has to be between 0 and 2"


def test_sass_is_sm100a():
    so = pfb().LIB_PATH
    out = subprocess.run(["cuobjdump", "-lelf", so], stdout=subprocess.PIPE, text=True).stdout
    assert "sm_100a" in out
