"""Size-independent properties at BASELINE.json's FULL sizes (the oracle cannot
finish these in seconds, so parity with it is checked on the scaled-down cases
of test_gpu_parity.py / test_golden.py and the full-size runs are pinned by
properties instead): trace of the eigenvalue table, census, determinism,
target variance and correlation of the sampled fields, symmetry of the
covariance operator, fused value transform."""
import numpy as np
import pytest

from util import pfb

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def m():
    return pfb()


def spherical(h):
    return np.where(h < 1.0, 1.0 - 1.5 * h + 0.5 * h**3, 0.0)


def test_k3_full_size_properties(m):
    import torch

    n = 512
    with m.Plan([n, n, n], covariance="spherical", corr_length=0.1) as p:
        st = p.compute_eigenvalues()
        assert st["negative"] == 0 and st["factor_used"] == 2
        # sum of ALL eigenvalues = c(0) = variance: compact table (513^3) with the
        # multiplicity of every mirrored index
        lam = p.eigenvalues(m.LAMBDA_REFERENCE).reshape(n + 1, n + 1, n + 1)
        w = np.full(n + 1, 2.0)
        w[0] = w[n] = 1.0
        trace = np.einsum("ijk,i,j,k->", lam, w, w, w, optimize=True)
        assert abs(trace - 1.0) < 1e-10
        del lam
        dev = torch.device("cuda")
        a = torch.empty(p.domain_size, dtype=torch.float64, device=dev)
        b = torch.empty_like(a)
        p.generate(20240611, 0, out=a)
        p.generate(20240611, 0, out=b)
        p.synchronize()
        assert torch.isfinite(a).all() and torch.equal(a, b)
        # target variance and correlation (spherical model), averaged over a few
        # realizations; about 1000 correlation volumes per field
        var, c10, c40, R = 0.0, 0.0, 0.0, 6
        for r in range(R):
            p.generate(20240611, r, out=a)
            p.synchronize()
            f = a.view(n, n, n)
            var += float((f * f).mean())
            c10 += float((f[:, :, :-10] * f[:, :, 10:]).mean())
            c40 += float((f[:, :-40, :] * f[:, 40:, :]).mean())
        var, c10, c40 = var / R, c10 / R, c40 / R
        assert abs(var - 1.0) < 0.1
        assert abs(c10 - spherical(10 / 512 / 0.1)) < 0.1
        assert abs(c40 - spherical(40 / 512 / 0.1)) < 0.1
        # covariance operator: symmetric, positive
        p.generate(1, 100, out=b)
        Ma, Mb = torch.empty_like(a), torch.empty_like(a)
        p.apply("matrix", a, out=Ma)
        p.apply("matrix", b, out=Mb)
        p.synchronize()
        ab, ba, aa = float(torch.dot(b, Ma)), float(torch.dot(a, Mb)), float(torch.dot(a, Ma))
        assert aa > 0
        assert abs(ab - ba) <= 1e-10 * max(abs(aa), abs(ab))


def test_k2_full_size_lognormal(m):
    import torch

    case = dict(cells=[256, 256, 256], covariance="gaussian", corr_length=[0.05, 0.03, 0.02], anisotropy="axiparallel", transform="logNormal", approximate=True)
    with m.Plan(**case) as p:
        p.compute_eigenvalues()
        raw = torch.empty(p.domain_size, dtype=torch.float64, device="cuda")
        tr = torch.empty_like(raw)
        p.generate(3, 7, out=raw)
        p.generate(3, 7, out=tr, apply_transform=True)
        p.synchronize()
        assert torch.isfinite(raw).all()
        assert float((tr - torch.exp(raw)).abs().max()) <= 1e-13 * float(tr.abs().max())
        assert abs(float((raw * raw).mean()) - 1.0) < 0.1


def test_k4_full_size_properties(m):
    import torch

    n = 8192
    case = dict(cells=[n, n], covariance="cubic", corr_length=[0.08, 0.03, 0.03, 0.05], anisotropy="geometric", approximate=True)
    with m.Plan(**case) as p:
        st = p.compute_eigenvalues()
        assert st["negative"] == 0
        a = torch.empty(p.domain_size, dtype=torch.float64, device="cuda")
        b = torch.empty_like(a)
        p.generate(5, 1, out=a)
        p.generate(5, 1, out=b)
        p.synchronize()
        assert torch.isfinite(a).all() and torch.equal(a, b)
        var = 0.0
        for r in range(8):  # about 250 correlation areas per field
            p.generate(5, r, out=a)
            p.synchronize()
            var += float((a * a).mean()) / 8
        assert abs(var - 1.0) < 0.15


def test_k4_full_size_operators(m):
    """timesMatrix on long lines at K4's size (16384^2 torus): symmetric and
    positive; applied to a delta it returns the first column of the circulant
    matrix, c(x) / N (the reference's timesMatrix is C_ext / N)"""
    import torch

    case = dict(cells=[8192, 8192], covariance="cubic", corr_length=[0.08, 0.03, 0.03, 0.05], anisotropy="geometric", approximate=True)
    next_cells = 16384.0 * 16384.0
    with m.Plan(**case) as p:
        p.compute_eigenvalues()
        x = torch.empty(p.domain_size, dtype=torch.float64, device="cuda")
        y = torch.empty_like(x)
        p.generate(5, 0, out=x)
        p.generate(5, 1, out=y)
        Mx, My = torch.empty_like(x), torch.empty_like(x)
        p.apply("matrix", x, out=Mx)
        p.apply("matrix", y, out=My)
        p.synchronize()
        xy, yx, xx = float(torch.dot(y, Mx)), float(torch.dot(x, My)), float(torch.dot(x, Mx))
        assert xx > 0
        assert abs(xy - yx) <= 1e-10 * max(abs(xx), abs(xy))
        d = torch.zeros(p.domain_size, dtype=torch.float64, device="cuda")
        d[0] = 1.0
        col = torch.empty_like(d)
        p.apply("matrix", d, out=col)
        p.synchronize()
        col = col.view(8192, 8192) * next_cells
        # variance at lag 0 (the clamped negative eigenvalues of the approximate
        # embedding move it slightly), decaying, zero beyond the cubic model's range
        assert abs(float(col[0, 0]) - 1.0) < 1e-2
        assert 0.0 < float(col[0, 8]) < float(col[0, 0])
        assert abs(float(col[4000, 4000])) < 1e-3
