"""The product's OWN noise path (device Philox + Box-Muller fused into the first
inverse pass -- the variant bench.py times) against its CPU restatement
(oracle/pf_oracle.cpp: philox_noise -> generate_from_noise), plus the
statistical tests BASELINE.json's north star asks for: per-frequency chi-square
of the sampled spectrum against w_k lambda_k and a read-back of
the normals themselves.

Tolerances: rng_mode fp64 1e-10 relative to the max norm (the north star's
fp64 tolerance); rng_mode fast 3e-5 (fp32 MUFU Box-Muller: ~1e-6 per normal)."""
import numpy as np
import pytest

from oracle import oracle_lib as ol
from util import pfb, rel

pytestmark = pytest.mark.gpu

CASES = [
    # reference test grids and scaled-down BASELINE configs (K1..K5)
    dict(cells=[32], covariance="exponential", corr_length=0.05),
    dict(cells=[512], covariance="exponential", corr_length=0.01),
    dict(cells=[16, 16], covariance="exponential", corr_length=0.05),
    dict(cells=[64, 64], covariance="exponential", corr_length=0.05),  # K1
    dict(cells=[16, 8, 32], covariance="gaussian", corr_length=[0.05, 0.03, 0.02], anisotropy="axiparallel", approximate=True),  # K2
    dict(cells=[16, 16, 16], covariance="spherical", corr_length=0.1),  # K3
    dict(cells=[32, 64], covariance="cubic", corr_length=[0.08, 0.03, 0.03, 0.05], anisotropy="geometric", approximate=True),  # K4
    dict(cells=[8, 8, 8], covariance="exponential", corr_length=0.05),  # K5
    dict(cells=[2, 2], covariance="exponential", corr_length=0.5),
    # pass A lines of 512 / 1024 (with PF_GEN3=force: the grouped persistent
    # kernel kp_gen3 that the K3 bench runs), ragged tile counts
    dict(cells=[8, 2, 256], covariance="exponential", corr_length=0.1),
    dict(cells=[4, 1, 512], covariance="spherical", corr_length=0.2),
    dict(cells=[64, 512], covariance="exponential", corr_length=0.05),
    dict(cells=[16, 4, 512], covariance="exponential", corr_length=0.1, extensions=[1.0, 0.25, 2.0], approximate=True),
    # 2D lines above 1024 (four-step route of K4), either axis
    dict(cells=[1024, 16], covariance="exponential", corr_length=0.05),
    dict(cells=[16, 2048], covariance="spherical", corr_length=0.1),
]


def ids(c):
    return "x".join(map(str, c["cells"])) + "-" + c["covariance"]


@pytest.fixture(scope="module")
def m():
    return pfb()


def oracle_field(o, seed, realization, mode):
    noise = o.philox_noise(seed, realization, mode)
    if o.dim == 1:
        return o.generate_from_noise(noise, want_spare=True)
    return o.generate_from_noise(noise), None


@pytest.mark.parametrize("case", CASES, ids=ids)
@pytest.mark.parametrize("rng,tol", [("fp64", 1e-10), ("fast", 3e-5)])
@pytest.mark.parametrize("gen3", [False, True], ids=["default", "gen3"])
def test_device_noise_field_matches_oracle(m, case, rng, tol, gen3, monkeypatch):
    n_last = 2 * case["cells"][-1]
    if gen3:
        if len(case["cells"]) == 1 or n_last not in (512, 1024):
            pytest.skip("kp_gen3 serves pass A lines of 512 / 1024")
        monkeypatch.setenv("PF_GEN3", "force")
    o = ol.Oracle(**case)
    # a seed with high key bits and a realization index above 2^32: every
    # counter / key word of the Philox block is exercised
    with m.Plan(rng_mode=rng, **case) as p:
        for seed, r in ((20240611, 0), (0x1234567890ABCDEF, 3), (7, (1 << 40) + 5)):
            if o.dim == 1:
                # DFT backend: realizations 2t / 2t+1 are the real / imaginary
                # part of transform t (stateless C ABI)
                re_o, im_o = oracle_field(o, seed, r, rng)
                assert rel(p.generate(seed, 2 * r), re_o) < tol
                assert rel(p.generate(seed, 2 * r + 1), im_o) < tol
                continue
            f_o, _ = oracle_field(o, seed, r, rng)
            f = p.generate(seed, r)
            assert f.shape == f_o.shape
            assert rel(f, f_o) < tol


def test_batch_equals_single_calls(m):
    case = dict(cells=[16, 8, 32], covariance="exponential", corr_length=0.1)
    with m.Plan(**case) as p:
        batch = p.generate_batch(5, 2, 3)
        for i in range(3):
            assert np.array_equal(batch[i], p.generate(5, 2 + i))


@pytest.mark.parametrize("cells", [[64, 64], [512, 512], [48 * 0 + 32, 256]], ids=["64x64", "K1", "32x256"])
@pytest.mark.parametrize("rng", ["fast", "fp64"])
def test_2d_batch_in_two_launches_equals_single_calls(m, cells, rng):
    """pf_generate_batch on a device buffer, 2D: pass A takes the realization
    as its second line index, pass C sees the lines of the whole batch"""
    import torch

    case = dict(cells=cells, covariance="exponential", corr_length=0.05, transform="logNormal", rng_mode=rng)
    with m.Plan(**case) as p:
        dev = torch.empty((7, p.domain_size), dtype=torch.float64, device="cuda")
        p.generate(3, 0, out=dev[0])  # eigenvalues, tables
        l0 = p.launch_count
        p.generate_batch(3, (1 << 33) + 5, 7, out=dev, apply_transform=True)
        p.synchronize()
        assert p.launch_count - l0 == 2
        one = torch.empty(p.domain_size, dtype=torch.float64, device="cuda")
        for i in range(7):
            p.generate(3, (1 << 33) + 5 + i, out=one, apply_transform=True)
            p.synchronize()
            assert torch.equal(one, dev[i])


def test_1d_is_a_pure_function_of_seed_and_realization(m):
    """pf_generate(seed, r) must not depend on the call history"""
    case = dict(cells=[64], covariance="exponential", corr_length=0.05)
    noise = np.random.default_rng(0).standard_normal(2 * 128)
    with m.Plan(**case) as p:
        batch = p.generate_batch(9, 0, 4)
    with m.Plan(**case) as p:
        one = p.generate(9, 1)  # fresh plan, odd realization first
        assert np.array_equal(one, batch[1])
        assert np.array_equal(p.generate(9, 1), one)  # twice the same
        assert np.array_equal(p.generate(9, 0), batch[0])
        p.generate_from_noise(noise)  # overwrites the plan's spare buffer
        assert np.array_equal(p.generate(9, 1), batch[1])
        assert np.array_equal(p.generate(9, 3), batch[3])
        assert np.array_equal(p.generate(9, 2), batch[2])
        assert not np.array_equal(batch[0], batch[2])
    with m.Plan(transform="logNormal", **case) as p:
        t = p.generate_batch(9, 0, 2, apply_transform=True)
        assert rel(t, np.exp(batch[:2])) < 1e-13


def test_generate_uncorrelated_matches_restated_stream(m):
    # Matrix::generateUncorrelatedField (matrix.hh:450-462) on the device stream
    for cells in ([32], [2, 1], [4, 8, 16], [2, 2, 1]):  # 2 and 4 cells: the partial last quad
        with m.Plan(cells) as p:
            f = p.generate_uncorrelated(0xABCDEF0123456789)
            ref = ol.philox_uncorrelated(0xABCDEF0123456789, f.size).reshape(f.shape)
            assert np.max(np.abs(f - ref)) < 1e-14
            g = p.generate_uncorrelated(0xABCDEF0123456789)
            assert np.array_equal(f, g)
            assert not np.array_equal(f, p.generate_uncorrelated(1))


def test_set_variance_as_field(m):
    # Matrix::setVarianceAsField, matrix.hh:473-479: host and device outputs
    import torch

    with m.Plan([16, 8, 4], variance=2.75) as p:
        assert np.all(p.variance_field() == 2.75)
        d = torch.zeros(p.domain_size, dtype=torch.float64, device="cuda")
        p.variance_field(out=d)
        p.synchronize()
        assert bool((d == 2.75).all())


@pytest.mark.parametrize("n", [1, 1000, (1 << 20) + 3])
def test_vector_reductions(m, n):
    # StochasticPart::operator*, oneNorm, infNorm (stochastic.hh:420-467)
    import torch

    rng = np.random.default_rng(n)
    x, y = rng.standard_normal(n), rng.standard_normal(n)
    x[n // 2] = -7.5
    with m.Plan([8, 8]) as p:
        for a, b in ((x, y), (torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda())):
            assert abs(p.dot(a, b) - np.dot(x, y)) <= 1e-13 * np.sum(np.abs(x * y))
            assert abs(p.norm1(a) - np.sum(np.abs(x))) <= 1e-13 * np.sum(np.abs(x))
            assert p.norminf(a) == 7.5


CENSUS = [
    # (small, zero, negative) of the oracle: (270, 667, 0), (3498, 376, 0),
    # (26, 59, 0), (3108, 0, 0), (98, 0, 99), (121, 0, 120), (0, 0, 0) x 2
    dict(cells=[32, 32], covariance="gaussian", corr_length=0.15, approximate=True),
    dict(cells=[16, 16, 16], covariance="gaussian", corr_length=0.15, approximate=True),
    dict(cells=[64], covariance="gaussian", corr_length=0.1, approximate=True),
    dict(cells=[64, 64], covariance="cauchy", corr_length=0.1, approximate=True),
    dict(cells=[128], covariance="matern52", corr_length=0.2, approximate=True),
    dict(cells=[16, 16], covariance="gaussian", corr_length=0.3, approximate=True),
    dict(cells=[16, 16, 16], covariance="spherical", corr_length=0.1),
    dict(cells=[32, 64], covariance="cubic", corr_length=[0.08, 0.03, 0.03, 0.05], anisotropy="geometric", approximate=True),
]


@pytest.mark.parametrize("case", CENSUS, ids=ids)
def test_census_counts(m, case):
    """small / zero / negative counts of Matrix::fillTransformedMatrix
    (matrix.hh:289-306) equal the oracle's, up to the number of eigenvalues
    that sit within rounding distance of one of the three thresholds."""
    o = ol.Oracle(**case)
    lam = o.eigenvalues()  # unclamped census happens inside; stats in o.stats
    thr = 1e-14
    with m.Plan(**case) as p:
        st = p.compute_eigenvalues()
        nat = p.eigenvalues(m.LAMBDA_REFERENCE)
    assert nat.size == lam.size
    # eigenvalues near a threshold may be classified either way (they differ
    # by ~1e-16 between two FFT implementations); the stored table is clamped,
    # so the bands are taken at the two positive thresholds, and the cases are
    # chosen so that nothing sits near -thr (tiny negatives ~1e-17 are "zero",
    # the negatives of the last case are ~1e-3)
    band = 2e-15 * max(1.0, lam.max())
    near = sum(int(np.count_nonzero(np.abs(lam - t) < band)) for t in (1e-6, thr))
    for key in ("small", "zero"):
        assert abs(st[key] - o.stats[key]) <= near, (key, st, o.stats, near)
    assert st["negative"] == o.stats["negative"], (st, o.stats)
    assert abs(st["smallest"] - o.stats["smallest"]) <= 1e-12 * lam.max()
    total = lambda d: d["small"] + d["zero"] + d["negative"]
    assert total(st) == total(o.stats)
    if case["covariance"] not in ("spherical", "cubic"):
        assert total(st) > 0


# ---------------------------------------------------------------------------
# statistics of the sampled fields
# ---------------------------------------------------------------------------
def _periodic_plan(m, rng_mode, parity_mode, n=128):
    # randomField.periodic: the field IS the torus (factor 1), so its DFT gives
    # back every spectral coefficient the sampler drew
    return m.Plan([n, n, n], covariance="exponential", corr_length=0.05, periodic=True, approximate=True, rng_mode=rng_mode, parity_mode=parity_mode)


def _weights(n, parity_mode, device):
    """w_k of SURVEY A.6 on the half spectrum (k2, k1, k0 <= n/2) and the mask
    of self-conjugate entries"""
    import torch

    m0 = n // 2 + 1
    k0 = torch.arange(m0, device=device)
    kk = torch.arange(n, device=device)
    plane = (k0 == 0) | (k0 == n // 2)
    selfc = ((2 * kk) % n == 0)
    self_mask = selfc[:, None, None] & selfc[None, :, None] & plane[None, None, :]
    w = torch.ones((n, n, m0), dtype=torch.float64, device=device)
    if parity_mode == "reference":
        w[:, :, plane] = 0.5
        w[self_mask] = 1.0
    return w, self_mask, plane


@pytest.mark.parametrize("rng_mode", ["fast", "fp64"])
@pytest.mark.parametrize("parity_mode", ["reference", "corrected"])
def test_sampled_spectrum_chi_square(m, rng_mode, parity_mode):
    """Average periodogram of R = 32 realizations against w_k lambda_k, binned
    over |k| shells: every shell's chi-square statistic within 4 sigma.  A
    shell holds 10^3..10^5 coefficients, so a variance defect of a fraction of
    a percent (or the 50 % effect of the k0-plane rule of r2cfieldbackend.hh:
    305-325 if it were mis-stated) fails."""
    import torch

    n, R = 128, 32
    dev = torch.device("cuda")
    with _periodic_plan(m, rng_mode, parity_mode, n) as p:
        p.compute_eigenvalues()
        lam = torch.from_numpy(p.eigenvalues(m.LAMBDA_NATURAL).reshape(n, n, n // 2 + 1)).to(dev)
        w, self_mask, plane = _weights(n, parity_mode, dev)
        acc = torch.zeros_like(lam)
        f = torch.empty(n * n * n, dtype=torch.float64, device=dev)
        for r in range(R):
            p.generate(99, r, out=f)
            p.synchronize()
            F = torch.fft.rfftn(f.view(n, n, n)) / float(n**3)
            acc += F.real**2 + F.imag**2
    ok = lam > 1e-10 * lam.max()
    ratio = torch.where(ok, acc / (w * lam).clamp_min(1e-300), torch.zeros_like(acc))
    # integer |k| shells (frequencies folded to [-n/2, n/2])
    kf = torch.arange(n, device=dev)
    kf = torch.minimum(kf, n - kf).double()
    k0 = torch.arange(n // 2 + 1, device=dev).double()
    radius = torch.sqrt(kf[:, None, None] ** 2 + kf[None, :, None] ** 2 + k0[None, None, :] ** 2)
    shell = radius.round().long()
    interior = ok & ~plane[None, None, :].expand_as(ok)
    zs = []
    for b in range(1, int(shell.max()) + 1):
        sel = interior & (shell == b)
        cnt = int(sel.sum())
        if cnt < 50:
            continue
        # each coefficient: |F|^2 / lambda = (r1^2 + r2^2) / 2; summed over R
        # realizations and cnt coefficients: chi^2 with 2 R cnt dof, halved
        dof = 2.0 * R * cnt
        T = 2.0 * float(ratio[sel].sum())
        zs.append((T - dof) / np.sqrt(2.0 * dof))
    assert len(zs) > 60
    assert max(abs(z) for z in zs) < 4.0, zs
    # the two k0 planes: entries k and -k are conjugates of each other after
    # the c2r, so the plane sum counts every independent coefficient twice
    sel = ok & plane[None, None, :].expand_as(ok) & ~self_mask
    cnt = int(sel.sum())
    dof = 2.0 * R * (cnt / 2)
    T = 2.0 * float(ratio[sel].sum()) / 2
    z_plane = (T - dof) / np.sqrt(2.0 * dof)
    assert abs(z_plane) < 4.0, z_plane
    # power of the test: with the wrong weight on the planes it must fail
    if parity_mode == "reference":
        T_wrong = T * 0.5  # as if w_k were 1
        assert abs((T_wrong - dof) / np.sqrt(2.0 * dof)) > 50
    # self-conjugate entries: real, full weight: chi^2 with R dof each
    sel = ok & self_mask
    cnt = int(sel.sum())
    T = float(ratio[sel].sum())
    assert abs(T - R * cnt) < 5.0 * np.sqrt(2.0 * R * cnt)


@pytest.mark.parametrize("rng_mode,tol", [("fast", 1e-4), ("fp64", 1e-9)])
def test_normals_read_back_from_the_field(m, rng_mode, tol):
    """The unit normals the device drew, recovered from a periodic field
    (z_k = sqrt(2) F_k / sqrt(lambda_k)), equal the restated Philox / Box-Muller
    stream entry by entry (10^6 entries), and pass moment, kurtosis,
    Kolmogorov-Smirnov and tail-count checks."""
    import torch
    from scipy import stats

    n = 128
    m0 = n // 2 + 1
    with _periodic_plan(m, rng_mode, "reference", n) as p:
        p.compute_eigenvalues()
        lam = p.eigenvalues(m.LAMBDA_NATURAL).reshape(n, n, m0)
        f = torch.empty(n**3, dtype=torch.float64, device="cuda")
        p.generate(2024, 11, out=f)
        p.synchronize()
        F = (torch.fft.rfftn(f.view(n, n, n)) / float(n**3)).cpu().numpy()
    o = ol.Oracle([n, n, n], covariance="exponential", corr_length=0.05, periodic=True, approximate=True)
    # reference stream order s = k0 + m0 (k2 + n2 k1) -> [k1][k2][k0][2] -> natural
    z_o = o.philox_noise(2024, 11, rng_mode).reshape(n, n, m0, 2).transpose(1, 0, 2, 3)
    ok = lam > 1e-8 * lam.max()
    ok[:, :, 0] = False  # k0 planes hold sums of two draws: checked spectrally
    ok[:, :, m0 - 1] = False
    z = np.sqrt(2.0) * F[ok] / np.sqrt(lam[ok])
    zo = z_o[ok]
    assert z.size > 900000
    assert np.max(np.abs(z.real - zo[:, 0])) < tol
    assert np.max(np.abs(z.imag - zo[:, 1])) < tol
    v = np.concatenate([z.real, z.imag])
    N = v.size
    assert abs(v.mean()) < 4.0 / np.sqrt(N)
    assert abs(v.var() - 1.0) < 4.0 * np.sqrt(2.0 / N)
    assert abs(stats.skew(v)) < 4.0 * np.sqrt(6.0 / N)
    assert abs(stats.kurtosis(v)) < 4.0 * np.sqrt(24.0 / N)
    assert stats.kstest(v, "norm").pvalue > 1e-3
    for t in (2.0, 3.0, 4.0, 4.5):
        pr = 2 * stats.norm.sf(t)
        k = np.count_nonzero(np.abs(v) > t)
        assert abs(k - N * pr) < 5.0 * np.sqrt(N * pr) + 1, (t, k, N * pr)


def test_fast_and_fp64_modes_agree_in_distribution(m):
    """Two-sample Kolmogorov-Smirnov test between the point values of fields
    drawn in the two rng modes (independent streams: 7 vs 10 rounds).  Samples
    are taken 32 cells apart (five correlation lengths), so they are
    independent to better than one percent."""
    import torch
    from scipy import stats

    n = 128
    vals = {}
    for mode in ("fast", "fp64"):
        with m.Plan([n, n, n], covariance="exponential", corr_length=0.05, rng_mode=mode) as p:
            f = torch.empty(n**3, dtype=torch.float64, device="cuda")
            acc = []
            for r in range(128):
                p.generate(31, r, out=f)
                p.synchronize()
                acc.append(f.view(n, n, n)[::32, ::32, ::32].reshape(-1).cpu().numpy())
            vals[mode] = np.concatenate(acc)
    assert vals["fast"].size == 8192
    assert stats.ks_2samp(vals["fast"], vals["fp64"]).pvalue > 1e-3
    for v in vals.values():
        assert stats.kstest(v, "norm").pvalue > 1e-3
        assert abs(v.var() - 1.0) < 5.0 * np.sqrt(2.0 / v.size)


def test_fp64_box_muller_routines_match_library_forms(tmp_path):
    """neg2_log_uniform / sincos_turn_fraction / sqrt_positive (pf_rng.cuh) against
    log / sincospi / sqrt on the device, 2^27 normals + edge words"""
    import os
    import shutil
    import subprocess

    if shutil.which("nvcc") is None:
        pytest.skip("nvcc not available")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "box_muller_f64")
    subprocess.check_call(["nvcc", "-std=c++17", "-O2", "-gencode", "arch=compute_100a,code=sm_100a", "-I", os.path.join(root, "dune-parafields_b200", "csrc"), "-o", exe, os.path.join(root, "tests", "host", "box_muller_f64.cu")])
    out = subprocess.run([exe], stdout=subprocess.PIPE, text=True, check=True, timeout=300).stdout
    worst = float(out.split()[-1])
    # normals reach 6.8: a few ulp of that (the library forms carry 1-2 ulp themselves)
    assert 0.0 <= worst < 1e-14, out
