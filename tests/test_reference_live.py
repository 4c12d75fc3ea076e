"""Oracle against the UNMODIFIED reference run live (oracle/_ref/libpf_ref.so).
Only where the reference checkout existed at build time (this container); on
the GPU box the committed fixtures of tests/golden/ stand in."""
import numpy as np
import pytest

from oracle import oracle_lib as ol
from oracle import ref_lib as rl
from util import rel

pytestmark = pytest.mark.skipif(not rl.available(), reason="oracle/_ref not built (no /root/reference)")

COVS = [
    ("exponential", {}),
    ("gaussian", {}),
    ("spherical", {}),
    ("separableExponential", {}),
    ("matern32", {}),
    ("matern52", {}),
    ("dampedOscillation", {}),
    ("cauchy", {}),
    ("cubic", {}),
    ("whiteNoise", {}),
    ("gammaExponential", {"expGamma": 1.5}),
    ("generalizedCauchy", {"cauchyAlpha": 1.2, "cauchyBeta": 0.8}),
]


@pytest.mark.parametrize("cov,params", COVS, ids=[c[0] for c in COVS])
@pytest.mark.parametrize("cells", [[32], [16, 8], [8, 4, 16]], ids=["1d", "2d", "3d"])
def test_every_covariance_functor(cov, params, cells):
    if cov == "cubic" and len(cells) == 3:
        pytest.skip("cubic covariance only applicable in 1D or 2D (covariance.hh:842-845)")
    r = rl.Reference(cells, covariance=cov, corr_length=0.07, variance=1.7, approximate=True, cov_params=params)
    p = list(params.values()) + [0.0, 0.0]
    o = ol.Oracle(cells, covariance=cov, corr_length=0.07, variance=1.7, approximate=True, p0=p[0], p1=p[1])
    lam_r = r.lambda_stored()
    assert rel(o.lambda_stored().ravel(), lam_r) < 1e-12
    f_r = r.generate(11)
    f_o = o.generate(11, want_spare=True)[0] if len(cells) == 1 else o.generate(11)
    assert rel(f_o, f_r) < 1e-10


def test_negative_eigenvalues_message():
    with pytest.raises(rl.RefError, match="negative eigenvalues in covariance matrix"):
        rl.Reference([16, 16], covariance="gaussian", corr_length=0.5).generate(1)
    o = ol.Oracle([16, 16], covariance="gaussian", corr_length=0.5)
    with pytest.raises(Exception, match="negative eigenvalues in covariance matrix"):
        o.eigenvalues()


def test_operator_identities_of_fieldtest():
    """ext/parafields-core/test/fieldtest.cc:32-130 prints these; here the
    oracle must agree with the reference on each composed operator."""
    for cells in ([32], [16, 16], [8, 8, 8]):
        case = dict(cells=cells, covariance="exponential", corr_length=0.05)
        r = rl.Reference(**case)
        o = ol.Oracle(**case)
        x = r.generate(5)
        # M^1/2 M^-1 M^1/2
        r.set(x)
        r.times("root")
        r.times("inverse")
        y_r = r.times("root")
        y_o = o.multiply(o.multiply_inverse(o.multiply(x, "root")), "root")
        assert rel(y_o, y_r) < 1e-8
        # M M^-1 and M^-1 M reach CG tolerance
        r.set(x)
        r.times("inverse")
        z = r.times("matrix")
        assert np.linalg.norm(z - x) / np.linalg.norm(x) < 1e-5


def test_cache_state_machine():
    """randomfield.hh:1131-1260: with cacheInvMatvec the inverse after a
    timesMatrix is the stored pre-image (a copy, no CG)."""
    r = rl.Reference([16, 16], covariance="exponential", corr_length=0.05, cache_inv=True, cache_inv_root=True)
    x = r.generate(3)
    r.times("matrix")
    back = r.times("inverse")
    assert np.array_equal(back, x)
