"""CPU check of the index algebra behind pf_long.cuh (lines longer than the
1024-point warp FFT): both two-step decompositions, the permuted half-spectrum
columns and the Hermitian row pairing of the long c2r axis, restated with
numpy and compared with numpy's own transforms.  Small sizes (N2 = 8 stands
for 1024)."""
import numpy as np
import pytest


def two_step_inverse_I(X, N1, N2):
    """decomposition (I): k = N2 k1 + k2, j = j1 + N1 j2 (pass A / eigenvalues)"""
    n = N1 * N2
    x = X.reshape(N1, N2)  # [k1][k2]
    # step 1: N1-point DFT over k1, twiddle W_n^{+ j1 k2}  (k_long1)
    y = np.fft.ifft(x, axis=0) * N1
    j1 = np.arange(N1)[:, None]
    k2 = np.arange(N2)[None, :]
    y = y * np.exp(2j * np.pi * j1 * k2 / n)
    # step 2: N2-point FFT over k2 (k_c2t), stored at j1 + N1 j2
    z = np.fft.ifft(y, axis=1) * N2  # [j1][j2]
    return z.T.reshape(n)  # index j2 * N1 + j1


@pytest.mark.parametrize("N1,N2", [(1, 8), (2, 8), (4, 16), (16, 8)])
def test_decomposition_I_matches_fft(N1, N2):
    rng = np.random.default_rng(1)
    n = N1 * N2
    X = rng.standard_normal(n) + 1j * rng.standard_normal(n)
    ref = np.fft.ifft(X) * n
    assert np.allclose(two_step_inverse_I(X, N1, N2), ref, atol=1e-11)


def long_c2r_pair(A, B, N1, N2):
    """decomposition (II) on the packed pair V = A + iB (Hermitian-extended),
    reading the PERMUTED rows [k1][k2 <= N2/2] as k_long_c2r1 does"""
    n = N1 * N2
    M = N2 // 2 + 1
    # permuted storage: column k1 * M + k2 holds k0 = k1 + N1 k2 (k0 <= n/2)
    pa = np.zeros((N1, M), complex)
    pb = np.zeros((N1, M), complex)
    for k1 in range(N1):
        for k2 in range(M):
            k0 = k1 + N1 * k2
            if k0 <= n // 2:
                pa[k1, k2], pb[k1, k2] = A[k0], B[k0]
    U = np.zeros((N1, N2), complex)
    for k1 in range(N1):
        kr = (N1 - k1) % N1
        v = np.zeros(N2, complex)
        for e in range(N2):
            direct = 2 * e < N2 or (k1 == 0 and 2 * e == N2)
            col = e if direct else (N2 - e if k1 == 0 else N2 - 1 - e)
            xa = pa[k1 if direct else kr, col]
            xb = pb[k1 if direct else kr, col]
            if k1 == 0 and (e == 0 or 2 * e == N2):
                v[e] = xa.real + 1j * xb.real
            elif direct:
                v[e] = xa + 1j * xb
            else:
                v[e] = np.conj(xa) + 1j * np.conj(xb)
        u = np.fft.ifft(v) * N2  # over k2'
        U[k1] = u * np.exp(2j * np.pi * np.arange(N2) * k1 / n)
    # step 2 (k_long_c2r2): N1-point DFT over k1 -> x[N2 j1 + j2]
    x = np.fft.ifft(U, axis=0) * N1  # [j1][j2]
    return x.reshape(n)


@pytest.mark.parametrize("N1,N2", [(2, 8), (4, 8), (16, 16)])
def test_long_c2r_matches_irfft(N1, N2):
    rng = np.random.default_rng(2)
    n = N1 * N2
    m0 = n // 2 + 1
    A = rng.standard_normal(m0) + 1j * rng.standard_normal(m0)
    B = rng.standard_normal(m0) + 1j * rng.standard_normal(m0)
    x = long_c2r_pair(A, B, N1, N2)
    # FFTW / numpy c2r semantics on non-Hermitian input: Im at k = 0, n/2 ignored
    ra = np.fft.irfft(A, n) * n
    rb = np.fft.irfft(B, n) * n
    assert np.allclose(x.real, ra, atol=1e-10)
    assert np.allclose(x.imag, rb, atol=1e-10)


def test_permuted_columns_cover_the_half_spectrum_once():
    for N1, N2 in [(1, 16), (2, 1024), (16, 1024)]:
        n = N1 * N2
        M = N2 // 2 + 1
        k0 = np.array([k1 + N1 * k2 for k1 in range(N1) for k2 in range(M)])
        valid = k0[k0 <= n // 2]
        assert sorted(valid) == list(range(n // 2 + 1))
