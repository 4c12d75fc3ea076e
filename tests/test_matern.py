"""General-nu Matern covariance (covariance.hh:527-581, `covariance = matern`,
`stochastic.maternNu`).  The reference evaluates it with GSL's
gsl_sf_bessel_Knu (GSL is absent here, so the reference build in oracle/_ref
throws for it); pins used instead:
  * nu = 1/2, 3/2, 5/2 reduce to the reference's own closed-form functors
    (exponential with the argument scaled, matern32, matern52);
  * scipy.special.kv as an independent K_nu for other orders.
The CUDA path evaluates K_nu on the device (Temme series / Steed CF2,
pf_kernels.cuh: bessel_k_nu)."""
import os
import subprocess

import numpy as np
import pytest

from oracle import oracle_lib as ol
from util import ROOT, pfb, rel

TOL = 1e-10


def scipy_lambda_2d(cells, ell, nu, factor=2):
    """eigenvalues of the mirrored torus covariance, field-backend index set
    (matrix.hh:624-648 + r2c half spectrum), K_nu from scipy"""
    from scipy.special import gamma, kv

    n = [factor * c for c in cells]
    ax = []
    for i in range(2):
        h = 1.0 / cells[i]
        x = np.arange(n[i]) * h
        x = np.where(x > 0.5 * factor, x - factor, x)
        ax.append(x / ell)
    r = np.sqrt(ax[1][:, None] ** 2 + ax[0][None, :] ** 2)
    z = np.sqrt(2.0 * nu) * r
    with np.errstate(invalid="ignore", over="ignore"):
        c = 2.0 ** (1.0 - nu) / gamma(nu) * z**nu * kv(nu, z)
    c[r < 1e-10] = 1.0
    return np.real(np.fft.rfft2(c)) / (n[0] * n[1])


@pytest.mark.parametrize("nu,closed", [(1.5, "matern32"), (2.5, "matern52")])
def test_oracle_matern_reduces_to_closed_forms(nu, closed):
    a = ol.Oracle([16, 16], covariance="matern", p0=nu, corr_length=0.1, approximate=True)
    b = ol.Oracle([16, 16], covariance=closed, corr_length=0.1, approximate=True)
    assert rel(a.lambda_natural(), b.lambda_natural()) < 1e-12


def test_oracle_matern_half_is_exponential():
    # nu = 1/2: sqrt(2 nu) = 1, C = exp(-h)
    a = ol.Oracle([8, 8, 8], covariance="matern", p0=0.5, corr_length=0.2)
    b = ol.Oracle([8, 8, 8], covariance="exponential", corr_length=0.2)
    assert rel(a.lambda_natural(), b.lambda_natural()) < 1e-12


@pytest.mark.parametrize("nu", [0.3, 0.8, 1.0, 2.3, 3.5, 6.25])
def test_oracle_matern_against_scipy(nu):
    o = ol.Oracle([32, 16], covariance="matern", p0=nu, corr_length=0.07, approximate=True)
    lam = o.lambda_natural().ravel()
    ref = scipy_lambda_2d([32, 16], 0.07, nu).ravel()
    assert rel(lam, ref) < 1e-12


MATERN_CASES = [
    dict(cells=[32, 16], nu=0.3, ell=0.07),
    dict(cells=[32, 16], nu=0.5, ell=0.07),
    dict(cells=[32, 16], nu=0.8, ell=0.07),
    dict(cells=[32, 16], nu=1.0, ell=0.07),
    dict(cells=[32, 16], nu=1.49, ell=0.07),
    dict(cells=[32, 16], nu=1.5, ell=0.07),
    dict(cells=[32, 16], nu=2.3, ell=0.07),
    dict(cells=[32, 16], nu=6.25, ell=0.3),
    dict(cells=[16, 8, 8], nu=1.2, ell=0.15),
    dict(cells=[64], nu=2.0, ell=0.05),
    dict(cells=[256, 128], nu=0.75, ell=0.01),  # arguments up to ~170: CF2 branch, underflow region
]


@pytest.mark.gpu
@pytest.mark.parametrize("case", MATERN_CASES, ids=lambda c: "x".join(map(str, c["cells"])) + "-nu%g" % c["nu"])
def test_cuda_matern_matches_oracle(case):
    m = pfb()
    o = ol.Oracle(case["cells"], covariance="matern", p0=case["nu"], corr_length=case["ell"], approximate=True)
    lam_o = o.lambda_natural().ravel()
    noise = o.reference_noise(11)
    with m.Plan(case["cells"], covariance="matern", cov_param=(case["nu"], 0.0), corr_length=case["ell"], approximate=True) as p:
        p.compute_eigenvalues()
        lam = p.eigenvalues(m.LAMBDA_NATURAL)
        assert rel(lam, lam_o) < TOL
        f = p.generate_from_noise(noise)
        f_o = o.generate_from_noise(noise)
        if isinstance(f_o, tuple):
            f_o = f_o[0]
        assert rel(f, f_o) < TOL


@pytest.mark.gpu
def test_cuda_matern_equals_matern32_plan():
    m = pfb()
    with m.Plan([64, 64], covariance="matern", cov_param=(1.5, 0.0), corr_length=0.05) as a, m.Plan([64, 64], covariance="matern32", corr_length=0.05) as b:
        a.compute_eigenvalues()
        b.compute_eigenvalues()
        assert rel(a.eigenvalues(), b.eigenvalues()) < TOL
        assert rel(a.generate(seed=3, realization=1), b.generate(seed=3, realization=1)) < 1e-9


def test_matern_nu_must_be_positive():
    m = pfb()
    d = m.make_desc([8, 8], covariance="matern", cov_param=(0.0, 0.0))
    import ctypes as C

    plan = C.c_void_p()
    rc = m.lib().pf_plan_create(C.byref(d), C.byref(plan))
    assert rc == m.PF_ERR_INVALID
    assert b"matern nu has to be positive" in m.lib().pf_last_error()


CLI = os.path.join(ROOT, "dune-parafields_b200", "cli", "_build", "fieldgenerator_cuda")


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(CLI), reason="dune-parafields_b200/cli/_build/fieldgenerator_cuda not built (no /root/reference)")
def test_reference_cli_general_matern(tmp_path):
    """`covariance = matern` + `maternNu` through the reference's own CLI
    (src/fieldgenerator.cc, unmodified) on the CUDA path: same field as the
    matern32 functor for nu = 1.5"""
    m = pfb()
    ini = tmp_path / "field.ini"
    ini.write_text("[grid]\nextensions = 1 1\ncells = 32 32\n[stochastic]\ncovariance = matern\nmaternNu = 1.5\nvariance = 1.\ncorrLength = 0.1\n")
    out = subprocess.run([CLI, str(ini), "-input.seed", "5", "-output.legacyVtk", str(tmp_path / "f")], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300, cwd=tmp_path)
    print(out.stdout)
    assert out.returncode == 0, out.stdout
    files = [f for f in os.listdir(tmp_path) if f.startswith("f") and "vtk" in f]
    vals, started = [], False
    for line in open(tmp_path / files[0]):
        if started:
            vals += [float(x) for x in line.split()]
        elif line.startswith("LOOKUP_TABLE"):
            started = True
    vals = np.array(vals)
    with m.Plan([32, 32], covariance="matern32", corr_length=0.1, rng_mode="fp64") as p:
        ref = p.generate(seed=5, realization=0).ravel()
    assert vals.size == ref.size
    assert np.max(np.abs(vals - ref)) < 1e-4 * np.max(np.abs(ref))
