"""Self-validation of the CPU oracle (oracle/pf_oracle.cpp).

The reference ships no golden vectors, so the oracle
is pinned against first principles: naive O(N^2) DFTs, dense circulant
algebra, numpy's pocketfft as an independent FFT, and the libstdc++ RNG
known-answer values.
"""
import numpy as np
import pytest

from oracle import oracle_lib as ol

TOL = 1e-10

REF_GRIDS = {1: [32], 2: [16, 16], 3: [8, 8, 8]}  # test/randomfield{1,2,3}d.ini


def rel(a, b):
    return np.max(np.abs(np.asarray(a) - np.asarray(b))) / max(np.max(np.abs(b)), 1e-300)


# ---------------------------------------------------------------- RNG KAT
def test_reference_rng_known_answer():
    # std::default_random_engine(42) + normal_distribution
    golden = [
        -1.7141127395876619,
        0.17805682714513033,
        0.057178867692959344,
        -1.4097969926348002,
        0.75628398756772552,
        -0.58227370160874969,
    ]
    o = ol.Oracle([4])
    noise = o.reference_noise(42)
    assert np.array_equal(noise[:6], np.array(golden))


# ------------------------------------------------------------ eigenvalues
def numpy_lambda_full(o):
    c = o.torus_covariance()  # (n2,n1,n0)
    return np.real(np.fft.fftn(c)) / c.size


@pytest.mark.parametrize("dim", [1, 2, 3])
@pytest.mark.parametrize("cov", ["exponential", "gaussian", "spherical"])
def test_lambda_matches_numpy_fft(dim, cov):
    o = ol.Oracle(REF_GRIDS[dim], covariance=cov, corr_length=0.05, approximate=True)
    lam = o.lambda_natural()
    full = numpy_lambda_full(o)
    b0 = lam.shape[2]
    ref = np.maximum(full[:, :, :b0], 0.0)  # clamp, matrix.hh:304-305
    assert rel(lam, ref) < TOL


def test_lambda_naive_dft_2d():
    o = ol.Oracle([4, 3], covariance="exponential", corr_length=0.3, extensions=[1.0, 0.7])
    c = o.torus_covariance()[0]  # (n1,n0)
    n1, n0 = c.shape
    lam = np.zeros((n1, n0 // 2 + 1))
    for k1 in range(n1):
        for k0 in range(n0 // 2 + 1):
            acc = 0.0
            for j1 in range(n1):
                for j0 in range(n0):
                    acc += c[j1, j0] * np.cos(2 * np.pi * (j0 * k0 / n0 + j1 * k1 / n1))
            lam[k1, k0] = acc / (n0 * n1)
    assert rel(o.lambda_natural()[0], np.maximum(lam, 0)) < TOL


def test_matrix_backends_agree():
    # compact (DCT), half-spectrum (R2C) and the transposed storage must give
    # the same eigenvalue for the same frequency
    kw = dict(covariance="gaussian", corr_length=[0.05, 0.1, 0.07], anisotropy="axiparallel", approximate=True)
    a = ol.Oracle([8, 6, 4], matrix_backend="DCT", **kw).lambda_natural()
    b = ol.Oracle([8, 6, 4], matrix_backend="R2C", **kw).lambda_natural()
    c = ol.Oracle([8, 6, 4], matrix_backend="DCT", transposed=False, **kw).lambda_natural()
    assert rel(a, b) < TOL and rel(a, c) < TOL


def test_transposed_storage_layout():
    # dctmatrixbackend.hh:221-231: after the transform the last two dims are
    # swapped: value(k0,k1,k2) at [k0 + d0*(k2 + d2*k1)]
    o = ol.Oracle([4, 6, 8], covariance="exponential", corr_length=0.2)
    stored = o.eigenvalues()
    nat = o.lambda_natural()
    d0, d1, d2 = 5, 7, 9
    for (k0, k1, k2) in [(1, 2, 3), (4, 6, 8), (0, 5, 1)]:
        assert stored[k0 + d0 * (k2 + d2 * k1)] == nat[k2, k1, k0]


def test_geometric_anisotropy_lambda():
    G = [0.08, 0.03, 0.03, 0.05]
    o = ol.Oracle([16, 12], covariance="cubic", corr_length=G, anisotropy="geometric", approximate=True)
    lam = o.lambda_natural()[0]
    full = numpy_lambda_full(o)[0]
    assert rel(lam, np.maximum(full[:, : lam.shape[1]], 0)) < TOL
    # geometry: the torus covariance is c(A h) with A = G^-1
    A = np.linalg.inv(np.array(G).reshape(2, 2))
    c = o.torus_covariance()[0]
    h = np.array([3 / 16.0, -2 / 12.0])  # torus cell (3, 24-2)
    r = np.linalg.norm(A @ h)
    expect = 0.0 if r > 1 else 1 - 7 * r**2 + 8.75 * r**3 - 3.5 * r**5 + 0.75 * r**7
    assert abs(c[24 - 2, 3] - expect) < 1e-13


def test_census_and_negative_throws():
    # the Gaussian covariance with a long correlation length is the classic
    # non-embeddable case
    o = ol.Oracle([16, 16], covariance="gaussian", corr_length=0.5)
    with pytest.raises(RuntimeError, match="negative eigenvalues in covariance matrix"):
        o.eigenvalues()
    assert o.stats["negative"] > 0
    o2 = ol.Oracle([16, 16], covariance="gaussian", corr_length=0.5, approximate=True)
    lam = o2.eigenvalues()
    assert lam.min() == 0.0
    full = numpy_lambda_full(o2)[0][:17, :17]
    thr = 1e-14
    assert o2.stats["negative"] == int(np.sum(full <= -thr))
    assert o2.stats["smallest"] == pytest.approx(full.min(), rel=1e-6)


# --------------------------------------------------------------- generate
def numpy_generate(o, noise):
    """Independent restatement of SURVEY.md A.5 with numpy's irfftn."""
    n = o.n + [1] * (3 - o.dim)
    n0, n1, n2 = n
    m0 = n0 // 2 + 1
    lam = o.lambda_natural()  # (n2,n1,m0)
    r = noise.reshape(-1, 2)
    if o.dim == 3:  # stream s = k0 + m0*(k2 + n2*k1)
        r1 = r[:, 0].reshape(n1, n2, m0).transpose(1, 0, 2)
        r2 = r[:, 1].reshape(n1, n2, m0).transpose(1, 0, 2)
    else:
        r1 = r[:, 0].reshape(n2, n1, m0)
        r2 = r[:, 1].reshape(n2, n1, m0)
    k = np.meshgrid(np.arange(n2), np.arange(n1), np.arange(m0), indexing="ij")
    selfc = ((2 * k[2]) % n0 == 0) & ((2 * k[1]) % n1 == 0) & ((2 * k[0]) % n2 == 0)
    sig = np.sqrt(lam)
    X = np.where(selfc, sig * r1 + 0j, sig / np.sqrt(2.0) * (r1 + 1j * r2))
    axes = {2: (1, 2), 3: (0, 1, 2)}[o.dim]
    s = {2: (n1, n0), 3: (n2, n1, n0)}[o.dim]
    y = np.fft.irfftn(X, s=s, axes=axes) * (n0 * n1 * n2)
    c = o.cells + [1] * (3 - o.dim)
    return y[: c[2], : c[1], : c[0]].reshape(tuple(reversed(o.cells)))


@pytest.mark.parametrize("cells", [[16, 16], [8, 8, 8], [8, 6, 4], [12, 10]])
def test_generate_matches_numpy_irfftn(cells):
    o = ol.Oracle(cells, covariance="exponential", corr_length=0.05)
    noise = o.reference_noise(7)
    assert rel(o.generate_from_noise(noise), numpy_generate(o, noise)) < TOL


def test_generate_naive_inverse_dft_2d():
    # A.4: y = Re sum_k w.. evaluated literally (complex inverse along axis 1,
    # then half-complex->real along axis 0 ignoring Im at k0 in {0,n0/2})
    o = ol.Oracle([4, 4], covariance="exponential", corr_length=0.2)
    noise = o.reference_noise(3)
    n0 = n1 = 8
    m0 = 5
    lam = o.lambda_natural()[0]
    r = noise.reshape(n1, m0, 2)
    y = np.zeros((4, 4))
    for x1 in range(4):
        for x0 in range(4):
            acc = 0.0
            for k0 in range(m0):
                g = 0j
                for k1 in range(n1):
                    sc = (2 * k0) % n0 == 0 and (2 * k1) % n1 == 0
                    s = np.sqrt(lam[k1, k0])
                    X = s * r[k1, k0, 0] if sc else s / np.sqrt(2) * (r[k1, k0, 0] + 1j * r[k1, k0, 1])
                    g += X * np.exp(2j * np.pi * k1 * x1 / n1)
                if k0 == 0 or 2 * k0 == n0:
                    acc += g.real * np.cos(2 * np.pi * k0 * x0 / n0)
                else:
                    acc += 2 * (g * np.exp(2j * np.pi * k0 * x0 / n0)).real
            y[x1, x0] = acc
    assert rel(o.generate_from_noise(noise), y) < TOL


def test_generate_1d_dft_backend_with_spare():
    o = ol.Oracle([32], covariance="exponential", corr_length=0.05)
    noise = o.reference_noise(11)
    f, spare = o.generate_from_noise(noise, want_spare=True)
    lam = o.lambda_natural()[0, 0]
    Z = np.sqrt(lam) * (noise[0::2] + 1j * noise[1::2])
    z = np.fft.ifft(Z) * 64
    assert rel(f, z.real[:32]) < TOL and rel(spare, z.imag[:32]) < TOL


# -------------------------------------------------------------- operators
def dense_operator(o, g):
    """R (C_ext/N)^g P as a dense matrix built from the torus covariance."""
    c = o.torus_covariance()
    N = c.size
    lam = np.real(np.fft.fftn(c)) / N
    lam = np.maximum(lam, 0)
    cells = o.cells + [1] * (3 - o.dim)
    D = o.domain_size
    M = np.zeros((D, D))
    for j in range(D):
        e = np.zeros(D)
        e[j] = 1
        x = np.zeros(c.shape)
        x[: cells[2], : cells[1], : cells[0]] = e.reshape(cells[2], cells[1], cells[0])
        y = np.real(np.fft.ifftn(g(lam) * np.fft.fftn(x) / N)) * N
        M[:, j] = y[: cells[2], : cells[1], : cells[0]].ravel()
    return M


@pytest.mark.parametrize("cells", [[8], [4, 4], [4, 2, 2]])
def test_multiply_against_dense(cells):
    o = ol.Oracle(cells, covariance="exponential", corr_length=0.3)
    rng = np.random.default_rng(0)
    x = rng.standard_normal(o.domain_size)
    for op, g in [("matrix", lambda l: l), ("root", np.sqrt), ("inverse_ext", lambda l: 1 / l)]:
        M = dense_operator(o, g)
        assert rel(o.multiply(x, op).ravel(), M @ x) < 1e-9


def test_timesmatrix_is_cext_over_n():
    # SURVEY finding 1: multiplyExtended computes (C_ext/N) x
    o = ol.Oracle([4, 4], covariance="exponential", corr_length=0.3)
    c = o.torus_covariance()[0]
    e = np.zeros(16)
    e[0] = 1
    y = o.multiply(e, "matrix")
    assert rel(y, c[:4, :4] / 64.0) < 1e-12


@pytest.mark.parametrize("dim", [1, 2, 3])
def test_cg_inverse(dim):
    o = ol.Oracle(REF_GRIDS[dim], covariance="exponential", corr_length=0.05, cg_iterations=100)
    rng = np.random.default_rng(1)
    b = rng.standard_normal(o.domain_size)
    x = o.multiply_inverse(b)
    r = o.multiply(x, "matrix").ravel() - b
    assert np.linalg.norm(r) / np.linalg.norm(b) < 1e-6
    assert o.cg_iterations_used <= 100


def test_inverse_of_zero_field_is_copy():
    o = ol.Oracle([8, 8])
    x = o.multiply_inverse(np.zeros(64))
    assert np.all(x == 0)


# ------------------------------------------------------------- transforms
def test_value_transforms():
    v = np.array([-1.5, -0.0, 0.0, 0.3, 2.0])
    assert np.allclose(ol.transform("logNormal", v), np.exp(v), rtol=1e-15)
    assert np.array_equal(ol.transform("foldedNormal", v), np.abs(v))
    assert np.array_equal(ol.transform("sign", v), np.array([-1.0, 1.0, 1.0, 1.0, 1.0]))
    bc = ol.transform("boxCox", np.array([0.3, 2.0]), 0.5)
    assert np.allclose(bc, (np.array([0.3, 2.0]) * 0.5 + 1) ** -0.5, rtol=1e-15)


# ---------------------------------------------- effective spectrum (A.6)
def test_effective_variance_of_r2c_sampler():
    # finding 2: the default nD sampler under-disperses; the point variance is
    # sum_k w_k lambda_k with w=1/2 on the k0 in {0,n0/2} planes
    o = ol.Oracle([16, 16], covariance="exponential", corr_length=0.2)
    lam = o.lambda_natural()[0]  # (n1, m0)
    n0 = n1 = 32
    k1, k0 = np.meshgrid(np.arange(n1), np.arange(17), indexing="ij")
    selfc = ((2 * k0) % n0 == 0) & ((2 * k1) % n1 == 0)
    plane = (k0 == 0) | (k0 == 16)
    # half-spectrum weights: interior k0 count twice (k and -k)
    w = np.where(plane, np.where(selfc, 1.0, 0.5), 2.0)
    predicted = np.sum(w * lam)
    rng = np.random.default_rng(5)
    acc = 0.0
    R = 300
    for _ in range(R):
        f = o.generate_from_noise(rng.standard_normal(2 * o.noise_entries))
        acc += np.mean(f**2)
    emp = acc / R
    assert abs(emp - predicted) < 0.03 * predicted
    assert predicted < 0.97  # visibly below the nominal variance 1
