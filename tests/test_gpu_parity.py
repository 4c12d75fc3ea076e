"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on the
same seeded inputs.  Tolerance 1e-10 relative to the max norm (fp64), as
BASELINE.json's north star states."""
import numpy as np
import pytest

from oracle import oracle_lib as ol
from util import pfb, rel

pytestmark = pytest.mark.gpu
TOL = 1e-10

CASES = [
    # reference test grids (ext/parafields-core/test/randomfield{1,2,3}d.ini)
    dict(cells=[32], covariance="exponential", corr_length=0.05),
    dict(cells=[16, 16], covariance="exponential", corr_length=0.05),
    dict(cells=[8, 8, 8], covariance="exponential", corr_length=0.05),
    dict(cells=[16, 16], covariance="gaussian", corr_length=0.05, approximate=True),
    dict(cells=[8, 8, 8], covariance="spherical", corr_length=0.05),
    # long lines (N >= 256): persistent TMA-pipelined pass kernels
    dict(cells=[128, 128, 8], covariance="exponential", corr_length=0.1),
    dict(cells=[8, 8, 128], covariance="spherical", corr_length=0.2),
    dict(cells=[256, 2, 128], covariance="exponential", corr_length=[0.1, 0.3, 0.05], anisotropy="axiparallel"),
    # ragged / anisotropic
    dict(cells=[64, 16], covariance="exponential", corr_length=0.1, extensions=[2.0, 0.5]),
    dict(cells=[16, 8, 32], covariance="gaussian", corr_length=[0.05, 0.03, 0.02], anisotropy="axiparallel", approximate=True),
    dict(cells=[32, 64], covariance="cubic", corr_length=[0.08, 0.03, 0.03, 0.05], anisotropy="geometric", approximate=True),
    dict(cells=[16, 32, 8], covariance="exponential", corr_length=[0.1, 0.02, 0.0, 0.02, 0.2, 0.01, 0.0, 0.01, 0.15], anisotropy="geometric", approximate=True),
    dict(cells=[128, 128], covariance="matern32", corr_length=0.05),
    dict(cells=[512], covariance="exponential", corr_length=0.01),
    dict(cells=[512, 4], covariance="spherical", corr_length=0.2),
    dict(cells=[32, 32, 32], covariance="exponential", corr_length=0.1),
    dict(cells=[2, 2], covariance="exponential", corr_length=0.5),
    # pass A lines of 512 / 1024: grouped persistent kernel (kp_gen3), ragged
    # tile counts so that clipped tensor stores are covered
    dict(cells=[8, 2, 256], covariance="exponential", corr_length=0.1),
    dict(cells=[4, 1, 512], covariance="spherical", corr_length=0.2),
    dict(cells=[64, 512], covariance="exponential", corr_length=0.05),
    dict(cells=[32, 256], covariance="cubic", corr_length=[0.08, 0.03, 0.03, 0.05], anisotropy="geometric", approximate=True),
]


def ids(c):
    return "x".join(map(str, c["cells"])) + "-" + c["covariance"] + "-" + c.get("anisotropy", "none")


@pytest.fixture(scope="module")
def m():
    return pfb()


@pytest.mark.parametrize("case", CASES, ids=ids)
def test_eigenvalues_match_oracle(m, case):
    o = ol.Oracle(**case)
    lam_o = o.lambda_natural().ravel()
    with m.Plan(**case) as p:
        st = p.compute_eigenvalues()
        lam = p.eigenvalues(m.LAMBDA_NATURAL)
        assert lam.shape == lam_o.shape
        assert rel(lam, lam_o) < TOL
        # census: equal except for values within 1e-12*max of a threshold
        assert st["negative"] == o.stats["negative"] or case.get("approximate")
        assert abs(st["smallest"] - o.stats["smallest"]) <= 1e-10 * lam_o.max()
        # reference storage layout
        ref = p.eigenvalues(m.LAMBDA_REFERENCE)
        assert rel(ref, o.lambda_stored()) < TOL


@pytest.mark.parametrize("case", CASES, ids=ids)
def test_generate_from_reference_noise(m, case):
    o = ol.Oracle(**case)
    noise = o.reference_noise(20240611)
    with m.Plan(**case) as p:
        assert p.noise_entries == o.noise_entries
        if o.dim == 1:
            f_o, s_o = o.generate_from_noise(noise, want_spare=True)
            spare = np.empty_like(f_o)
            f = p.generate_from_noise(noise, spare=spare)
            assert rel(spare, s_o) < TOL
        else:
            f_o = o.generate_from_noise(noise)
            f = p.generate_from_noise(noise)
        assert f.shape == f_o.shape
        assert rel(f, f_o) < TOL


@pytest.mark.parametrize("case", CASES[:13], ids=ids)
@pytest.mark.parametrize("op", ["matrix", "root", "inverse_ext"])
def test_operators_match_oracle(m, case, op):
    o = ol.Oracle(**case)
    lam = o.lambda_stored()
    if op == "inverse_ext" and lam.min() <= 0:
        pytest.skip("clamped zero eigenvalues: 1/lambda = inf in both")
    rng = np.random.default_rng(3)
    x = rng.standard_normal(o.domain_size).reshape(tuple(reversed(o.cells)))
    y_o = o.multiply(x, op)
    with m.Plan(**case) as p:
        y = p.apply(op, x)
    tol = TOL if op != "inverse_ext" else 1e-8  # 1/lambda amplifies rounding
    assert rel(y, y_o) < tol


def test_negative_eigenvalues_error(m):
    with m.Plan([16, 16], covariance="gaussian", corr_length=0.5) as p:
        with pytest.raises(m.PfError, match="negative eigenvalues in covariance matrix") as e:
            p.compute_eigenvalues()
        assert e.value.code == m.PF_ERR_NEGATIVE_EIGENVALUES
        assert p.stats["negative"] > 0


def test_embedding_factor_retry(m):
    # new behaviour (off by default): double the factor until the census is clean
    # (oracle census: l=0.3 has 120 negatives at factor 2, none at factor 4)
    with m.Plan([16, 16], covariance="gaussian", corr_length=0.3, max_factor=16) as p:
        st = p.compute_eigenvalues()
        assert st["negative"] == 0 and st["factor_used"] == 4
        o = ol.Oracle([16, 16], covariance="gaussian", corr_length=0.3, factor=4, approximate=True)
        assert rel(p.eigenvalues(m.LAMBDA_NATURAL), o.lambda_natural().ravel()) < TOL
        noise = o.reference_noise(9)
        # the Gaussian spectrum decays to ~1e-17: eigenvalues that agree to
        # 1e-16 absolute give sqrt(lambda) that agree to ~1e-8 only, so the
        # field tolerance here is sqrt(eps)-limited by construction
        assert rel(p.generate_from_noise(noise), o.generate_from_noise(noise)) < 1e-7


@pytest.mark.parametrize("kind", ["logNormal", "foldedNormal", "sign", "boxCox"])
def test_fused_value_transform(m, kind):
    case = dict(cells=[16, 16], covariance="exponential", corr_length=0.1)
    with m.Plan(transform=kind, boxcox_lambda=0.25, **case) as p:
        raw = p.generate(seed=5, realization=2)
        tr = p.generate(seed=5, realization=2, apply_transform=True)
    if kind == "boxCox":
        raw = np.maximum(raw, -3.9)  # keep the base positive in both
        tr = np.where(raw <= -3.9, ol.transform(kind, raw, 0.25), tr)
    assert rel(tr, ol.transform(kind, raw, 0.25)) < 1e-13


def test_cg_inverse_small(m):
    for cells in ([32], [16, 16], [8, 8, 8]):
        case = dict(cells=cells, covariance="exponential", corr_length=0.05)
        o = ol.Oracle(cg_iterations=100, **case)
        rng = np.random.default_rng(1)
        b = rng.standard_normal(o.domain_size).reshape(tuple(reversed(cells)))
        with m.Plan(cg_iterations=100, **case) as p:
            x = p.solve(b)
            r = p.apply("matrix", x) - b
            assert np.linalg.norm(r) / np.linalg.norm(b) < 1e-6
            assert p.cg_iterations_used <= 100
        # pinned iteration count: both sides stop on the count
        o3 = ol.Oracle(cg_iterations=3, **case)
        x_o = o3.multiply_inverse(b)
        with m.Plan(cg_iterations=3, **case) as p:
            x3 = p.solve(b)
        assert rel(x3, x_o) < 1e-8


def test_zero_rhs_returns_copy(m):
    with m.Plan([16, 16]) as p:
        x = p.solve(np.zeros((16, 16)))
        assert np.all(x == 0) and p.cg_iterations_used == 0


@pytest.mark.parametrize("case", [CASES[-4], CASES[-3], CASES[-2], dict(cells=[256, 256], covariance="exponential", corr_length=0.1), dict(cells=[16, 16, 512], covariance="spherical", corr_length=0.1)], ids=ids)
@pytest.mark.parametrize("rng", ["fast", "fp64"])
def test_grouped_pass_a_equals_plain_kernels(m, case, rng, monkeypatch):
    """device noise: the grouped persistent pass A (default) and the plain
    per-tile kernels (PF_PIPE=0) compute the same products in the same order"""
    monkeypatch.setenv("PF_GEN3", "force")  # small grids: skip the size heuristic
    with m.Plan(rng_mode=rng, **case) as p:
        a = [p.generate(seed=7, realization=r) for r in range(3)]
        noise = np.random.default_rng(2).standard_normal(2 * p.noise_entries)
        fa = p.generate_from_noise(noise)
    monkeypatch.delenv("PF_GEN3")
    monkeypatch.setenv("PF_PIPE", "0")
    with m.Plan(rng_mode=rng, **case) as p:
        b = [p.generate(seed=7, realization=r) for r in range(3)]
        fb = p.generate_from_noise(noise)
    assert np.array_equal(fa, fb)
    for x, y in zip(a, b):
        assert np.isfinite(x).all() and np.array_equal(x, y)
