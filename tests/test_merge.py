"""embedding.periodization = merge (Matrix::computeMatrixEntriesWithMerge,
matrix.hh:662-787; sigmoids covariance.hh:145-275): the covariance on the
torus is the sum of 2^dim mirrored copies tapered by sigmoids.

CPU part: the oracle reproduces the fixtures recorded from the UNMODIFIED
reference (tests/golden/merge_*.npz, made by tests/golden/make_golden_merge.py)
and the live reference where oracle/_ref exists.  GPU part (-m gpu): the CUDA
path (covariance evaluation fused into the first eigenvalue FFT pass) against
the same fixtures and the oracle, 1e-10 relative to the max norm."""
import glob
import json
import os

import numpy as np
import pytest

from oracle import oracle_lib as ol
from oracle import ref_lib as rl
from util import pfb, rel

HERE = os.path.dirname(os.path.abspath(__file__))
FIXTURES = sorted(glob.glob(os.path.join(HERE, "golden", "merge_*.npz")))
TOL = 1e-10


def fid(f):
    return os.path.basename(f)[len("merge_") : -4]


def load(f):
    g = np.load(f)
    case = json.loads(str(g["case"]))
    return g, case, case.get("anisotropy") == "geometric"


def test_fixtures_present():
    assert len(FIXTURES) == 9


@pytest.mark.parametrize("fixture", FIXTURES, ids=fid)
def test_oracle_matches_reference_fixtures(fixture):
    g, case, quirk = load(fixture)
    o = ol.Oracle(periodization="merge", **case)
    o.eigenvalues(census_quirk=quirk)
    assert rel(o.lambda_stored().ravel(), g["lambda_stored"]) < 1e-13
    f = o.generate(int(g["seed"]), want_spare=True)[0] if o.dim == 1 else o.generate(int(g["seed"]))
    assert rel(f, g["field"]) < 1e-12


def test_merge_repairs_the_embedding():
    """what the periodisation is for: at factor 4 the classical embedding of a
    long-range covariance has negative eigenvalues, the merged one has none
    (at factor 2 the default taper has no width and both coincide)."""
    case = dict(cells=[64, 64], covariance="exponential", corr_length=0.5, factor=4, approximate=True)
    classical = ol.Oracle(**case)
    classical.eigenvalues()
    merged = ol.Oracle(periodization="merge", **case)
    merged.eigenvalues()
    assert classical.stats["negative"] > 0
    assert merged.stats["negative"] == 0


def test_recursion_heuristic():
    # matrix.hh:693-719
    assert ol.merge_recursions_for("exponential", 3) == 2
    assert ol.merge_recursions_for("matern32", 2) == 4
    assert ol.merge_recursions_for("matern52", 2) == 6
    assert ol.merge_recursions_for("cauchy", 1) == 3
    assert ol.merge_recursions_for("matern", 3, p0=1.3) == 4
    assert ol.merge_recursions_for("gammaExponential", 2, p0=1.5) == 2
    assert ol.merge_recursions_for("gaussian", 2, recursions=7) == 7
    with pytest.raises(RuntimeError, match="magic recursions value 99"):
        ol.merge_recursions_for("gaussian", 2)


@pytest.mark.skipif(not rl.available(), reason="oracle/_ref not built (no /root/reference)")
@pytest.mark.parametrize("cov", ["exponential", "matern32", "matern52", "cauchy"])
@pytest.mark.parametrize("sigmoid", ["smoothstep", "smooth"])
def test_oracle_matches_live_reference(cov, sigmoid):
    cells = [12, 10]
    r = rl.Reference(cells, covariance=cov, corr_length=0.15, approximate=True, factor=3, extra="[embedding]\nperiodization = merge\nsigmoid = %s\n" % sigmoid)
    o = ol.Oracle(cells, covariance=cov, corr_length=0.15, approximate=True, factor=3, periodization="merge", sigmoid=sigmoid)
    assert rel(o.lambda_stored().ravel(), r.lambda_stored()) < 1e-13
    assert rel(o.generate(5), r.generate(5)) < 1e-12


# ---------------------------------------------------------------------------
# GPU
# ---------------------------------------------------------------------------
@pytest.fixture(scope="module")
def m():
    return pfb()


@pytest.mark.gpu
@pytest.mark.parametrize("fixture", FIXTURES, ids=fid)
def test_cuda_matches_reference_fixtures(m, fixture):
    g, case, quirk = load(fixture)
    o = ol.Oracle(periodization="merge", **case)
    noise = o.reference_noise(int(g["seed"]))
    with m.Plan(periodization="merge", census_quirk=quirk, **case) as p:
        p.compute_eigenvalues()
        lam = p.eigenvalues(m.LAMBDA_REFERENCE).ravel()
        # the reference leaves some negative eigenvalues of the geometric
        # cases unclamped (census quirk); the CUDA path
        # clamps them all, so compare the clamped tables
        assert rel(np.maximum(lam, 0.0), np.maximum(g["lambda_stored"], 0.0)) < TOL
        if o.dim == 1:
            f = p.generate_from_noise(noise, spare=np.empty(case["cells"][0]))
        else:
            f = p.generate_from_noise(noise)
        if g["lambda_stored"].min() >= 0.0:
            assert rel(f, g["field"]) < TOL
        else:
            # sqrt of a negative eigenvalue is NaN in the reference's field
            o.eigenvalues()
            assert rel(f, o.generate_from_noise(noise)) < TOL


MERGE_CASES = [
    dict(cells=[64, 32], approximate=True, covariance="exponential", corr_length=0.3),
    dict(cells=[128, 16], approximate=True, covariance="matern32", corr_length=0.2, sigmoid="smooth", extensions=[2.0, 0.5]),
    dict(cells=[16, 32, 8], approximate=True, covariance="exponential", corr_length=[0.3, 0.2, 0.1], anisotropy="axiparallel"),
    dict(cells=[32, 16, 8], approximate=True, covariance="cauchy", corr_length=0.2, factor=4),
    dict(cells=[64, 64], approximate=True, covariance="exponential", corr_length=0.25, factor=4, sigmoid_combine="radial"),
    dict(cells=[256], approximate=True, covariance="matern52", corr_length=0.1),
    dict(cells=[32, 32], approximate=True, covariance="spherical", corr_length=0.6, factor=4, merge_recursions=3, sigmoid_start=0.9, sigmoid_end=2.5),
]


@pytest.mark.gpu
@pytest.mark.parametrize("case", MERGE_CASES, ids=lambda c: "x".join(map(str, c["cells"])) + "-" + c["covariance"])
def test_cuda_matches_oracle(m, case):
    o = ol.Oracle(periodization="merge", **case)
    lam_o = o.lambda_natural().ravel()
    noise = o.reference_noise(7)
    with m.Plan(periodization="merge", **case) as p:
        st = p.compute_eigenvalues()
        assert rel(p.eigenvalues(m.LAMBDA_NATURAL).ravel(), lam_o) < TOL
        assert st["negative"] == o.stats["negative"] or case.get("approximate")
        if o.dim == 1:
            f_o = o.generate_from_noise(noise, want_spare=True)[0]
            f = p.generate_from_noise(noise, spare=np.empty_like(f_o))
        else:
            f_o = o.generate_from_noise(noise)
            f = p.generate_from_noise(noise)
        assert rel(f, f_o) < TOL


@pytest.mark.gpu
def test_cuda_merge_repairs_the_embedding(m):
    case = dict(cells=[64, 64], covariance="exponential", corr_length=0.5, factor=4)
    with m.Plan(**case) as p:
        with pytest.raises(m.PfError, match="negative eigenvalues in covariance matrix"):
            p.compute_eigenvalues()
    with m.Plan(periodization="merge", **case) as p:
        st = p.compute_eigenvalues()
        assert st["negative"] == 0


def test_argument_validation(m):
    import ctypes as C

    h = C.c_void_p()
    d = m.make_desc([8, 8], covariance="gaussian", periodization="merge")
    assert m.lib().pf_plan_create(C.byref(d), C.byref(h)) == m.PF_ERR_INVALID
    assert m.lib().pf_last_error() == b"magic recursions value 99 not implemented for choice of covariance"
    d = m.make_desc([8, 8], periodization="merge")
    d.sigmoid = 5
    assert m.lib().pf_plan_create(C.byref(d), C.byref(h)) == m.PF_ERR_INVALID
    assert m.lib().pf_last_error() == b'embedding.sigmoid must be "smooth" or "smoothstep"'
    d = m.make_desc([8, 8], periodization="merge")
    d.sigmoid_combine = 3
    assert m.lib().pf_plan_create(C.byref(d), C.byref(h)) == m.PF_ERR_INVALID
    assert m.lib().pf_last_error() == b"sigmoidCombine must be tensor, radial or radialPlus"
    d = m.make_desc([8, 8])
    d.periodization = 2
    assert m.lib().pf_plan_create(C.byref(d), C.byref(h)) == m.PF_ERR_INVALID
    assert b"embedding.periodization" in m.lib().pf_last_error()
