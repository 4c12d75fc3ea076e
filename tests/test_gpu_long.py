"""GPU parity for 2D fields whose extended size exceeds the 1024-point warp
FFT (pf_long.cuh, the route BASELINE.json's K4 -- 16384^2 torus -- takes):
eigenvalues and fields from the reference noise stream against the CPU oracle,
1e-10 relative to the max norm (fp64)."""
import numpy as np
import pytest

from oracle import oracle_lib as ol
from util import pfb, rel

pytestmark = pytest.mark.gpu
TOL = 1e-10

GEO = dict(covariance="cubic", corr_length=[0.08, 0.03, 0.03, 0.05], anisotropy="geometric", approximate=True)
CASES = [
    # long c2r axis only (n0 = 2048), short pass-A axis
    dict(cells=[1024, 8], covariance="exponential", corr_length=0.05),
    # long pass-A axis only (n1 = 4096)
    dict(cells=[8, 2048], covariance="exponential", corr_length=0.05),
    # both long, different splits
    dict(cells=[2048, 1024], covariance="exponential", corr_length=0.05),
    # K4's matrix backend (geometric anisotropy: half-spectrum table), scaled down
    dict(cells=[1024, 1024], **GEO),
    # the longest lines: 16384 on either axis
    dict(cells=[8192, 4], covariance="spherical", corr_length=0.1),
    dict(cells=[4, 8192], covariance="matern32", corr_length=0.02),
    dict(cells=[4096, 16], **GEO),
    # a single line (the second line of the c2r pair is absent) and embedding factor 4
    # (a quarter of the line kept: the kernels without compile-time pruning)
    dict(cells=[2048, 1], covariance="exponential", corr_length=0.05),
    dict(cells=[1024, 2], covariance="exponential", corr_length=0.05, factor=4, approximate=True),
]


def ids(c):
    return "x".join(map(str, c["cells"])) + "-" + c["covariance"] + "-" + c.get("anisotropy", "none")


@pytest.fixture(scope="module")
def m():
    return pfb()


@pytest.mark.parametrize("case", CASES, ids=ids)
def test_long_eigenvalues_and_field(m, case):
    o = ol.Oracle(**case)
    lam_o = o.lambda_natural().ravel()
    noise = o.reference_noise(20240611)
    f_o = o.generate_from_noise(noise)
    with m.Plan(**case) as p:
        st = p.compute_eigenvalues()
        lam = p.eigenvalues(m.LAMBDA_NATURAL)
        assert lam.shape == lam_o.shape
        assert rel(lam, lam_o) < TOL
        assert st["negative"] == o.stats["negative"] or case.get("approximate")
        assert rel(p.eigenvalues(m.LAMBDA_REFERENCE), o.lambda_stored()) < TOL
        assert p.noise_entries == o.noise_entries
        f = p.generate_from_noise(noise)
        assert f.shape == f_o.shape
        assert rel(f, f_o) < TOL


def test_long_device_noise_statistics(m):
    # device Philox path: finite, reproducible, unit-ish variance for a short
    # correlation length, fused value transform consistent with the raw field
    case = dict(cells=[2048, 1024], covariance="exponential", corr_length=0.01, transform="logNormal")
    with m.Plan(**case) as p:
        a = p.generate(seed=3, realization=1)
        b = p.generate(seed=3, realization=1)
        c = p.generate(seed=3, realization=2)
        t = p.generate(seed=3, realization=1, apply_transform=True)
    assert np.isfinite(a).all() and np.array_equal(a, b) and not np.array_equal(a, c)
    assert abs(a.var() - 1.0) < 0.05 and abs(a.mean()) < 0.05
    assert rel(t, np.exp(a)) < 1e-13


OP_CASES = [
    dict(cells=[1024, 8], covariance="exponential", corr_length=0.05),
    dict(cells=[8, 2048], covariance="exponential", corr_length=0.05),
    dict(cells=[2048, 1024], covariance="exponential", corr_length=0.05),
    dict(cells=[1024, 1024], **GEO),  # half-spectrum table (geometric anisotropy)
    dict(cells=[8192, 4], covariance="spherical", corr_length=0.1),
    dict(cells=[4096, 1], covariance="spherical", corr_length=0.1),
    dict(cells=[1024, 2], covariance="exponential", corr_length=0.05, factor=4, approximate=True),
]


@pytest.mark.parametrize("case", OP_CASES, ids=ids)
def test_long_operators(m, case):
    """multiplyExtended / multiplyRootExtended / multiplyInverseExtended
    (matrix.hh:1459-1662) on long lines: full complex two-step transforms"""
    o = ol.Oracle(**case)
    rng = np.random.default_rng(5)
    x = rng.standard_normal(tuple(reversed(case["cells"])))
    with m.Plan(**case) as p:
        for op in ("matrix", "root", "inverse_ext"):
            y = p.apply(op, x)
            y_o = o.multiply(x, op)
            assert y.shape == y_o.shape
            # 1 / lambda on clamped (approximate) spectra amplifies the rounding
            # of lambda itself (values ~1e11): looser there
            tol = 1e-7 if (op == "inverse_ext" and case.get("approximate")) else TOL
            assert rel(y, y_o) < tol, op


def test_long_pcg(m):
    # Matrix::multiplyInverse (innerCG, matrix.hh:1331-1448), pinned iteration count
    case = dict(cells=[1024, 16], covariance="exponential", corr_length=0.05, cg_iterations=4)
    o = ol.Oracle(**case)
    x = np.random.default_rng(6).standard_normal((16, 1024))
    with m.Plan(**case) as p:
        assert rel(p.solve(x), o.multiply_inverse(x)) < 1e-8


def test_long_3d_limits(m):
    # 3D axes of 2048 / 4096 take the generic route (tests/test_gpu_anysize.py);
    # beyond the shared-memory transform length they are refused
    with m.Plan([1024, 4, 4], covariance="exponential", corr_length=0.05) as p:
        assert p.extended_size == 2048 * 8 * 8
    with pytest.raises(m.PfError, match="unsupported"):
        m.Plan([4096, 4, 4], covariance="exponential", corr_length=0.05)


def test_long_batch_equals_single_calls(m):
    """pf_generate_batch on long lines: step 1 of pass A takes up to four
    realizations from one read of the sigma table -- same fields as one call
    per realization, bit for bit (5 = one full group + a remainder)"""
    import torch

    for case in (dict(cells=[2048, 64], covariance="exponential", corr_length=0.05),
                 dict(cells=[64, 1024], covariance="cubic", corr_length=[0.08, 0.03, 0.03, 0.05], anisotropy="geometric", approximate=True, transform="logNormal")):
        with m.Plan(**case) as p:
            D = p.domain_size
            batch = torch.empty((5, D), dtype=torch.float64, device="cuda")
            one = torch.empty(D, dtype=torch.float64, device="cuda")
            tr = "transform" in case
            p.generate_batch(9, 100, 5, out=batch, apply_transform=tr)
            p.synchronize()
            for r in range(5):
                p.generate(9, 100 + r, out=one, apply_transform=tr)
                p.synchronize()
                assert torch.equal(batch[r], one), r
