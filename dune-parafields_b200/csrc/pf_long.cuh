// pf_long.cuh -- lines longer than the 1024-point warp FFT (2D, n <= 16384).
//
// A line of n = N1 * N2 points (N2 = 1024, N1 = 2..16) is transformed in two
// steps through HBM (the "four-step" FFT):
//
//   (I)  k = N2 k1 + k2, j = j1 + N1 j2 :
//        step 1 (here, elementwise): N1-point DFT over k1, twiddle W_n^{j1 k2},
//                                    written as Y[line][j1][k2];
//        step 2 (k_c2t of pf_kernels.cuh): N2-point FFT over k2 of the
//                                    contiguous sub-lines, tile store.
//        Used by pass A (noise -> x1) and by the eigenvalue transforms.
//   (II) k = k1 + N1 k2, j = N2 j1 + j2 : step 1 = k_long_c2r1 (N2-point FFTs,
//        pf_kernels.cuh), step 2 (here): N1-point DFT over k1 -> contiguous
//        real output.  Used by the c2r axis.
//
// The sub-line order doubles as the transposition the next pass needs, so no
// separate transpose runs per realization.  Replaces the FFTW plans of
// backends/r2cfieldbackend.hh:250-274 / r2cmatrixbackend.hh:254-285 for the
// sizes of BASELINE.json's K4 (16384^2 torus).  N1 = 1 is accepted everywhere
// (a short axis next to a long one takes the same route).
#pragma once
#include "pf_kernels.cuh"

namespace pf {

struct LongArgs
{
  int N2;            // sub-line length (<= 1024)
  long long n;       // N1 * N2
  const cplx* tw;    // [j1][k2] = exp(-2 pi i j1 k2 / n), j1 < N1, k2 < N2
  // noise source: output column a' = k1_0 * perm_m + k2_0 stands for
  // k0 = k1_0 + perm_n1 * k2_0 (identity when perm_n1 == 1)
  int perm_n1, perm_m;
  long long k0_max;  // columns with k0 > k0_max are padding (zeros)
  long long n_half;  // n / 2
  // complex source (in_real: the array holds doubles, the imaginary part is
  // zero); entries x >= in_valid are zero padding (0: the whole line is present)
  const cplx* in;
  long long in_pitch;
  long long in_valid;
  int in_real;
  // herm_n = n: the line holds the half spectrum x <= n / 2 of a real signal,
  // entries beyond are conj(in[n - x]) (0: a full complex line)
  long long herm_n;
  cplx* out;         // Y[line][j1][k2]
  // noise sources: `batch` realizations (p.realization + r) from ONE read of the
  // sigma row, written to out + r * batch_stride (0 / 1: a single realization)
  int batch;
  long long batch_stride;
};

enum
{
  LSRC_NOISE_FAST = 0,
  LSRC_NOISE_F64 = 1,
  LSRC_NOISE_INJECTED = 2,
  LSRC_CPLX = 3,
  LSRC_COV = 4
};

__device__ __forceinline__ double
long_sigma(const PassArgs& p, const double* row, long long e, long long n)
{
  const long long em = (p.tab_mirror_e && 2 * e > n) ? n - e : e;
  return __ldg(row + em);
}

// step 1 of decomposition (I); grid = (ceil(N2 / 256), lines)
// three blocks per SM (85 registers; measured at K4: 2.10 -> 2.02 ms per
// realization; four blocks spill and lose, 2.19 ms).  BATCH: g.batch
// realizations from one read of the sigma row (more registers: two blocks)
template <int N1, int SRC, int DIR, bool BATCH = false>
__global__ void __launch_bounds__(256, BATCH ? 2 : 3)
k_long1(const LongArgs g, const PassArgs p)
{
  constexpr double SQRT2 = 1.41421356237309504880168872420969808;
  const int k2 = blockIdx.x * blockDim.x + threadIdx.x;
  const long long line = blockIdx.y;
  const int N2 = g.N2;
  cplx* out = g.out + line * g.n;
  if constexpr (SRC <= LSRC_NOISE_INJECTED) {
    constexpr int NOISE = SRC;
    const long long k0 =
      (line / g.perm_m) + (long long)g.perm_n1 * (line % g.perm_m);
    // N1 == 1: a thread owns k2 and k2 + N2/2 (the two halves of one Philox
    // draw); otherwise k2 and all k1
    constexpr int H = N1 > 1 ? N1 / 2 : 1;
    if (k2 >= (N1 > 1 ? N2 : N2 / 2))
      return;
    cplx v[N1 > 1 ? N1 : 2];
    if (k0 > g.k0_max) {
      for (int r = 0; r < (g.batch > 1 ? g.batch : 1); r++) {
        cplx* o = out + (long long)r * g.batch_stride;
        if constexpr (N1 > 1) {
#pragma unroll
          for (int j = 0; j < N1; j++)
            o[(long long)j * N2 + k2] = make_double2(0., 0.);
        } else {
          o[k2] = make_double2(0., 0.);
          o[k2 + N2 / 2] = make_double2(0., 0.);
        }
      }
      return;
    }
    if constexpr (BATCH) {
    const double* row = table_row(p, (int)k0, 0);
    const unsigned long long cbase = (unsigned long long)k0 * p.cma;
    const long long ibase = k0 * p.ima;
    double s1[H], s2[H];
#pragma unroll
    for (int k1 = 0; k1 < H; k1++) {
      const long long e = (long long)k1 * N2 + k2; // < n / 2
      s1[k1] = long_sigma(p, row, e, g.n);
      s2[k1] = long_sigma(p, row, e + g.n_half, g.n);
    }
    const int nbatch = g.batch > 1 ? g.batch : 1;
    for (int r = 0; r < nbatch; r++) {
      const unsigned long long rea = p.realization + (unsigned long long)r;
#pragma unroll
      for (int k1 = 0; k1 < H; k1++) {
        const long long e = (long long)k1 * N2 + k2;
        double2 n1, n2;
        if constexpr (NOISE == NOISE_INJECTED) {
          n1 = __ldg(p.noise + ibase + e * p.ime);
          n2 = __ldg(p.noise + ibase + (e + g.n_half) * p.ime);
        } else {
          const unsigned long long ctr = cbase + (unsigned long long)e * p.cme;
          if constexpr (NOISE == NOISE_PHILOX_FAST) {
            const uint4 w = philox4x32_r<7>(ctr, rea, p.keys);
            n1 = box_muller_fast(w.x, w.y);
            n2 = box_muller_fast(w.z, w.w);
          } else {
            const uint4 w = philox4x32_r<10>(ctr, rea, p.keys);
            n1 = box_muller_f64(w.x, w.y);
            n2 = box_muller_f64(w.z, w.w);
          }
        }
        v[k1] = make_double2(s1[k1] * n1.x, s1[k1] * n1.y);
        v[k1 + H] = make_double2(s2[k1] * n2.x, s2[k1] * n2.y);
      }
      if (p.hermitian) {
        // FieldBackend::set, r2cfieldbackend.hh:305-325 (2D: a = k0, e = k1)
        const bool a_self = (2 * k0) % p.na_self == 0;
        const bool plane = p.corrected && a_self;
        if (plane) {
#pragma unroll
          for (int j = 0; j < 2 * H; j++)
            v[j] = make_double2(v[j].x * SQRT2, v[j].y * SQRT2);
        }
        if (k2 == 0 && a_self) {
          const double f = plane ? 1.0 : SQRT2;
          v[0] = make_double2(v[0].x * f, 0.);
          v[H] = make_double2(v[H].x * f, 0.);
        }
      }
      cplx* o = out + (long long)r * g.batch_stride;
      if constexpr (N1 > 1) {
        Dft<N1, DIR, 1>::run(v);
#pragma unroll
        for (int j = 0; j < N1; j++) {
          const cplx w = __ldg(g.tw + (long long)j * N2 + k2);
          o[(long long)j * N2 + k2] = DIR < 0 ? cmul(v[j], w) : cmulc(v[j], w);
        }
      } else {
        o[k2] = v[0];
        o[k2 + N2 / 2] = v[1];
      }
    }
    } else {
    const double* row = table_row(p, (int)k0, 0);
    const unsigned long long cbase = (unsigned long long)k0 * p.cma;
    const long long ibase = k0 * p.ima;
#pragma unroll
    for (int k1 = 0; k1 < H; k1++) {
      const long long e = (long long)k1 * N2 + k2; // < n / 2
      double2 n1, n2;
      if constexpr (NOISE == NOISE_INJECTED) {
        n1 = __ldg(p.noise + ibase + e * p.ime);
        n2 = __ldg(p.noise + ibase + (e + g.n_half) * p.ime);
      } else {
        const unsigned long long ctr = cbase + (unsigned long long)e * p.cme;
        if constexpr (NOISE == NOISE_PHILOX_FAST) {
          const uint4 w = philox4x32_r<7>(ctr, p.realization, p.keys);
          n1 = box_muller_fast(w.x, w.y);
          n2 = box_muller_fast(w.z, w.w);
        } else {
          const uint4 w = philox4x32_r<10>(ctr, p.realization, p.keys);
          n1 = box_muller_f64(w.x, w.y);
          n2 = box_muller_f64(w.z, w.w);
        }
      }
      const double s1 = long_sigma(p, row, e, g.n);
      const double s2 = long_sigma(p, row, e + g.n_half, g.n);
      v[k1] = make_double2(s1 * n1.x, s1 * n1.y);
      v[k1 + H] = make_double2(s2 * n2.x, s2 * n2.y);
    }
    if (p.hermitian) {
      // FieldBackend::set, r2cfieldbackend.hh:305-325 (2D: a = k0, e = k1)
      const bool a_self = (2 * k0) % p.na_self == 0;
      const bool plane = p.corrected && a_self;
      if (plane) {
#pragma unroll
        for (int j = 0; j < 2 * H; j++)
          v[j] = make_double2(v[j].x * SQRT2, v[j].y * SQRT2);
      }
      if (k2 == 0 && a_self) {
        const double f = plane ? 1.0 : SQRT2;
        v[0] = make_double2(v[0].x * f, 0.);
        v[H] = make_double2(v[H].x * f, 0.);
      }
    }
    if constexpr (N1 > 1) {
      Dft<N1, DIR, 1>::run(v);
#pragma unroll
      for (int j = 0; j < N1; j++) {
        const cplx w = __ldg(g.tw + (long long)j * N2 + k2);
        out[(long long)j * N2 + k2] = DIR < 0 ? cmul(v[j], w) : cmulc(v[j], w);
      }
    } else {
      out[k2] = v[0];
      out[k2 + N2 / 2] = v[1];
    }
    }
  } else {
    if (k2 >= N2)
      return;
    cplx v[N1];
#pragma unroll
    for (int k1 = 0; k1 < N1; k1++) {
      const long long x = (long long)k1 * N2 + k2;
      if constexpr (SRC == LSRC_CPLX) {
        if (g.herm_n > 0 && 2 * x > g.herm_n) {
          const cplx h = __ldg(g.in + line * g.in_pitch + (g.herm_n - x));
          v[k1] = make_double2(h.x, -h.y);
        } else if (g.in_valid > 0 && x >= g.in_valid)
          v[k1] = make_double2(0., 0.);
        else if (g.in_real)
          v[k1] = make_double2(
            __ldg(reinterpret_cast<const double*>(g.in) + line * g.in_pitch + x), 0.);
        else
          v[k1] = __ldg(g.in + line * g.in_pitch + x);
      } else
        v[k1] = make_double2(cov_torus(p.cov, (unsigned)x, (unsigned)line, 0), 0.);
    }
    Dft<N1, DIR, 1>::run(v);
#pragma unroll
    for (int j = 0; j < N1; j++) {
      const cplx w = __ldg(g.tw + (long long)j * N2 + k2);
      out[(long long)j * N2 + k2] =
        j == 0 ? v[0] : (DIR < 0 ? cmul(v[j], w) : cmulc(v[j], w));
    }
  }
}

// step 2 of decomposition (II): U[pair][k1][j2] -> N1-point inverse DFT over
// k1 -> x[N2 j1 + j2]; real part = line 2*pair, imaginary part = line
// 2*pair + 1; restriction to x < keep and value transform fused
// (extendedFieldToField, r2cfieldbackend.hh:423-441).  grid = (N2/256, pairs)
template <int N1>
__global__ void __launch_bounds__(256)
k_long_c2r2(const cplx* __restrict__ U, double* __restrict__ out, int N2,
            long long nlines, long long out_pitch, int keep, double scale,
            int transform, double tparam)
{
  const int j2 = blockIdx.x * blockDim.x + threadIdx.x;
  const long long pair = blockIdx.y;
  if (j2 >= N2)
    return;
  cplx v[N1];
  const cplx* u = U + pair * (long long)N1 * N2 + j2;
#pragma unroll
  for (int k1 = 0; k1 < N1; k1++)
    v[k1] = __ldg(u + (long long)k1 * N2);
  Dft<N1, +1, 1>::run(v);
  const long long L0 = 2 * pair, L1 = L0 + 1;
  double* o0 = out + L0 * out_pitch;
  double* o1 = out + L1 * out_pitch;
#pragma unroll
  for (int j1 = 0; j1 < N1; j1++) {
    const long long x = (long long)j1 * N2 + j2;
    if (x < keep) {
      o0[x] = apply_transform(v[j1].x * scale, transform, tparam);
      if (L1 < nlines)
        o1[x] = apply_transform(v[j1].y * scale, transform, tparam);
    }
  }
}

// complex transpose out[c][r] = in[r][c], r < rows, c < cols (eigenvalue setup)
__global__ void
k_long_transpose(const cplx* __restrict__ in, long long in_pitch,
                 cplx* __restrict__ out, long long out_pitch, long long rows,
                 long long cols)
{
  __shared__ cplx tile[32][33];
  const long long c0 = (long long)blockIdx.x * 32, r0 = (long long)blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const long long r = r0 + i, c = c0 + threadIdx.x;
    if (r < rows && c < cols)
      tile[i][threadIdx.x] = in[r * in_pitch + c];
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const long long c = c0 + i, r = r0 + threadIdx.x;
    if (r < rows && c < cols)
      out[c * out_pitch + r] = tile[threadIdx.x][i];
  }
}

// eigenvalues from the transformed covariance: real part / N, census and clamp
// (Matrix::fillTransformedMatrix, matrix.hh:283-306), table[k0 * P + k1]
__global__ void __launch_bounds__(256)
k_long_lambda(const cplx* __restrict__ C, long long c_pitch, double* __restrict__ tab,
              long long rows, long long P, double scale, double threshold,
              unsigned long long* census, unsigned long long* smallest_bits)
{
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned nsmall = 0, nzero = 0, nneg = 0;
  double smallest = 1.7976931348623157e308;
  if (i < rows * P) {
    const long long r = i / P, k = i % P;
    double val = C[r * c_pitch + k].x * scale;
    smallest = val;
    if (val < 1e-6) {
      if (val < threshold) {
        if (val > -threshold)
          nzero++;
        else
          nneg++;
      } else
        nsmall++;
    }
    if (val < 0.)
      val = 0.;
    tab[i] = val;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    nsmall += __shfl_xor_sync(0xffffffffu, nsmall, o);
    nzero += __shfl_xor_sync(0xffffffffu, nzero, o);
    nneg += __shfl_xor_sync(0xffffffffu, nneg, o);
    smallest = fmin(smallest, __shfl_xor_sync(0xffffffffu, smallest, o));
  }
  if ((threadIdx.x & 31) == 0) {
    if (nsmall)
      atomicAdd(census + 0, (unsigned long long)nsmall);
    if (nzero)
      atomicAdd(census + 1, (unsigned long long)nzero);
    if (nneg)
      atomicAdd(census + 2, (unsigned long long)nneg);
    atomicMin(smallest_bits, ordered_bits(smallest));
  }
}

// covariance operators on long lines: S[k0][k1] *= g(lambda(k0, k1)) * scale,
// table[k0' * P + k1'] holds rows k0' <= n0/2 (lambda(k) = lambda(-k) gives the
// others) and, when mirror is set, entries k1' <= n1/2
__global__ void __launch_bounds__(256)
k_long_spec(cplx* __restrict__ S, long long n0, long long n1,
            const double* __restrict__ tab, long long m0, long long P, int mirror,
            double scale)
{
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n0 * n1)
    return;
  long long k0 = i / n1, k1 = i % n1;
  if (k0 >= m0) {
    k0 = n0 - k0;
    k1 = k1 ? n1 - k1 : 0;
  }
  if (mirror && 2 * k1 > n1)
    k1 = n1 - k1;
  const double g = __ldg(tab + k0 * P + k1) * scale;
  const cplx v = S[i];
  S[i] = make_double2(v.x * g, v.y * g);
}

// real part of a complex matrix, restricted: out[r][c] = Re in[r][c]
__global__ void __launch_bounds__(256)
k_long_real(const cplx* __restrict__ in, long long in_pitch, double* __restrict__ out,
            long long rows, long long cols)
{
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < rows * cols)
    out[i] = in[(i / cols) * in_pitch + i % cols].x;
}

// tw[a * N2 + b] = exp(-2 pi i a b / n), a < N1, b < N2: the inter-step
// twiddles laid out so that consecutive threads read consecutive entries
__global__ void
k_long_twiddles(cplx* __restrict__ tw, long long n, int N2)
{
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const long long a = i / N2, b = i % N2;
    double s, c;
    sincospi(-2.0 * (double)(a * b) / (double)n, &s, &c);
    tw[i] = make_double2(c, s);
  }
}

} // namespace pf
