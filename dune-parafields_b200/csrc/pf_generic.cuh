// pf_generic.cuh -- transforms of ARBITRARY length (the reference hands FFTW any
// extended size n_i = embedding.factor * cells_i, fieldtraits.hh:268; nothing
// there asks for powers of two).  Sizes the specialised pass kernels do not
// cover -- cells = 100 100, 96 96 96, 48 80 112, 3D axes of 2048 -- take this
// route: natural-order arrays in HBM, one shared-memory transform per line.
//
//  * n a power of two (<= 4096): Stockham radix-2 autosort FFT in shared memory;
//  * any other n (2n - 1 <= 4096): Bluestein's chirp-z algorithm on top of it
//    (L. Bluestein 1970): X_k = w_k sum_j (x_j w_j) conj(w)_{k-j},
//    w_j = exp(dir * pi i j^2 / n), the convolution done with two FFTs of
//    length M >= 2n - 1 against the precomputed spectrum of conj(w).
//
// Transform conventions are FFTW's (backends/fftwwrapper.hh:230-415): dir = -1
// forward, +1 backward, both unnormalised; the c2r loader rebuilds the full
// Hermitian line from its stored half and ignores Im at k = 0 and k = n/2
// (r2cfieldbackend.hh:250-274).
#pragma once
#include "pf_kernels.cuh"

namespace pf {

struct GenLineArgs
{
  int n, M;            // line length, shared-memory transform length
  int dir;             // -1 forward, +1 backward
  const cplx* twM;     // exp(-2 pi i m / M), m < M / 2
  const cplx* chirp;   // w_j = exp(dir pi i j^2 / n), j < n; nullptr: n == M
  const cplx* bhat;    // FFT_M of the wrapped conj(w) kernel
  long long L0, L1;    // lines l = l0 + L0 * l1
  int in_mode;         // 0 complex, 1 real, 2 Hermitian half line (n/2+1 entries)
  const void* in;
  long long in_s0, in_s1, in_se; // line base l0*in_s0 + l1*in_s1, element stride
  int n_valid;         // entries present (zero beyond); modes 0 and 1
  int out_mode;        // 0 complex, 1 real part, 2 real part -> out, imaginary -> out2
  void* out;
  void* out2;
  long long out_s0, out_s1, out_se;
  int keep;            // entries stored
  double scale;
  int transform;       // value transform of the real output (mutators.hh)
  double tparam;
};

// in-place-looking Stockham FFT of length M in shared memory: input in `a`,
// `b` is scratch; returns the buffer holding the result.  conj_tw: backward.
__device__ __forceinline__ cplx*
smem_fft(cplx* a, cplx* b, int M, const cplx* __restrict__ tw, bool backward)
{
  const int T = M >> 1;
  for (int Ns = 1; Ns < M; Ns <<= 1) {
    const int tstep = T / Ns; // twiddle index step: W_M^{k * M / (2 Ns)}
    for (int j = threadIdx.x; j < T; j += blockDim.x) {
      const int k = j & (Ns - 1);
      cplx w = __ldg(tw + (long long)k * tstep);
      if (backward)
        w.y = -w.y;
      const cplx u0 = a[j];
      const cplx u1 = cmul(a[j + T], w);
      const int j0 = ((j - k) << 1) + k;
      b[j0] = cadd(u0, u1);
      b[j0 + Ns] = csub(u0, u1);
    }
    __syncthreads();
    cplx* t = a;
    a = b;
    b = t;
  }
  return a;
}

__global__ void __launch_bounds__(256)
k_generic_lines(const GenLineArgs g)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* A = reinterpret_cast<cplx*>(smem_raw);
  cplx* B = A + g.M;
  const long long l = blockIdx.x;
  const long long l0 = l % g.L0, l1 = l / g.L0;
  const int n = g.n, M = g.M;
  // ---- load (with the chirp when Bluestein) ---------------------------------
  const long long ibase = l0 * g.in_s0 + l1 * g.in_s1;
  for (int j = threadIdx.x; j < M; j += blockDim.x) {
    cplx v = make_double2(0., 0.);
    if (j < n) {
      if (g.in_mode == 0) {
        if (j < g.n_valid)
          v = reinterpret_cast<const cplx*>(g.in)[ibase + (long long)j * g.in_se];
      } else if (g.in_mode == 1) {
        if (j < g.n_valid)
          v.x = reinterpret_cast<const double*>(g.in)[ibase + (long long)j * g.in_se];
      } else {
        const int h = n / 2;
        if (j <= h) {
          v = reinterpret_cast<const cplx*>(g.in)[ibase + (long long)j * g.in_se];
          if (j == 0 || (2 * j == n))
            v.y = 0.;
        } else {
          v = reinterpret_cast<const cplx*>(g.in)[ibase + (long long)(n - j) * g.in_se];
          v.y = -v.y;
        }
      }
      if (g.chirp)
        v = cmul(v, __ldg(g.chirp + j));
    }
    A[j] = v;
  }
  __syncthreads();
  cplx* R;
  if (!g.chirp) {
    R = smem_fft(A, B, M, g.twM, g.dir > 0);
  } else {
    cplx* F = smem_fft(A, B, M, g.twM, false);
    cplx* O = F == A ? B : A;
    for (int j = threadIdx.x; j < M; j += blockDim.x)
      F[j] = cmul(F[j], __ldg(g.bhat + j));
    __syncthreads();
    R = smem_fft(F, O, M, g.twM, true);
    const double invM = 1.0 / (double)M;
    for (int j = threadIdx.x; j < n; j += blockDim.x) {
      const cplx c = cmul(R[j], __ldg(g.chirp + j));
      R[j] = make_double2(c.x * invM, c.y * invM);
    }
    __syncthreads();
  }
  // ---- store ----------------------------------------------------------------
  const long long obase = l0 * g.out_s0 + l1 * g.out_s1;
  for (int j = threadIdx.x; j < g.keep; j += blockDim.x) {
    const cplx v = R[j];
    const long long o = obase + (long long)j * g.out_se;
    if (g.out_mode == 0)
      reinterpret_cast<cplx*>(g.out)[o] = make_double2(v.x * g.scale, v.y * g.scale);
    else {
      reinterpret_cast<double*>(g.out)[o] =
        apply_transform(v.x * g.scale, g.transform, g.tparam);
      if (g.out_mode == 2 && g.out2)
        reinterpret_cast<double*>(g.out2)[o] = v.y * g.scale;
    }
  }
}

// ---- elementwise kernels of the generic route --------------------------------
struct GenGeom
{
  int dim;
  long long n[3], m0; // torus sizes, stored entries along axis 0
  long long half;     // ceil(L / 2) of the last axis (Philox block layout)
};

// torus covariance as a complex array C[x2][x1][x0] (matrix.hh:624-648)
__global__ void
k_generic_cov(GenGeom g, CovParams cov, cplx* __restrict__ C, long long total)
{
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const unsigned x0 = (unsigned)(i % g.n[0]);
    const unsigned x1 = (unsigned)((i / g.n[0]) % g.n[1]);
    const unsigned x2 = (unsigned)(i / (g.n[0] * g.n[1]));
    C[i] = make_double2(cov_torus(cov, x0, x1, x2), 0.);
  }
}

// the same from a caller-supplied table (custom covariance)
__global__ void
k_generic_cov_table(const double* __restrict__ table, cplx* __restrict__ C,
                    long long total)
{
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x)
    C[i] = make_double2(table[i], 0.);
}

// eigenvalues: Re / N of the transformed covariance on the stored index set
// (k0 < m0), census over the entries the reference's matrix backend stores
// (compact: k_i <= n_i / 2 for i >= 1; matrix.hh:283-306), clamp
__global__ void
k_generic_lambda(GenGeom g, const cplx* __restrict__ C, double* __restrict__ tab,
                 double scale, double threshold, int compact,
                 unsigned long long* census, unsigned long long* smallest_bits,
                 long long total)
{
  unsigned nsmall = 0, nzero = 0, nneg = 0;
  double smallest = 1.7976931348623157e308;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long k0 = i % g.m0, k1 = (i / g.m0) % g.n[1], k2 = i / (g.m0 * g.n[1]);
    double val = C[k0 + g.n[0] * (k1 + g.n[1] * k2)].x * scale;
    const bool counted = !compact || (2 * k1 <= g.n[1] && 2 * k2 <= g.n[2]);
    if (counted) {
      if (val < smallest)
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

// natural table -> user layout (PF_LAMBDA_NATURAL / PF_LAMBDA_REFERENCE)
__global__ void
k_generic_gather(GenGeom g, const double* __restrict__ tab, double* __restrict__ out,
                 long long ob0, long long ob1, long long ob2, int transposed,
                 long long total)
{
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    long long k0 = i % ob0, r = i / ob0, k1, k2;
    if (g.dim == 3 && transposed) {
      k2 = r % ob2;
      k1 = r / ob2;
    } else {
      k1 = r % ob1;
      k2 = r / ob1;
    }
    out[i] = tab[k0 + g.m0 * (k1 + g.n[1] * k2)];
  }
}

// white noise * sigma on the stored spectrum X[k2][k1][k0 < m0]
// (Matrix::generateField matrix.hh:396-421, FieldBackend::set
// r2cfieldbackend.hh:305-325 / dftfieldbackend.hh:276-280); the table holds
// sqrt(lambda) / sqrt(2) (Hermitian rule) or sqrt(lambda) (DFT backend, 1D)
template <int NOISE>
__global__ void
k_generic_noise(GenGeom g, const double* __restrict__ sig, PhiloxKeys keys,
                unsigned long long realization, const double2* __restrict__ noise,
                int hermitian, int corrected, cplx* __restrict__ X, long long total)
{
  constexpr double SQRT2 = 1.41421356237309504880168872420969808;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long k0 = i % g.m0, k1 = (i / g.m0) % g.n[1], k2 = i / (g.m0 * g.n[1]);
    const long long e = g.dim == 3 ? k2 : (g.dim == 2 ? k1 : k0);
    double2 nz;
    if constexpr (NOISE == NOISE_INJECTED) {
      // reference stream order: k0 fastest, then k2, then k1 in 3D
      const long long s = g.dim == 3 ? k0 + g.m0 * (k2 + g.n[2] * k1) : i;
      nz = __ldg(noise + s);
    } else {
      unsigned long long c;
      if (g.dim == 1)
        c = (unsigned long long)(e % g.half);
      else if (g.dim == 2)
        c = (unsigned long long)k0 + (unsigned long long)g.m0 * (e % g.half);
      else
        c = (unsigned long long)k0 + (unsigned long long)g.m0 * (e % g.half) +
            (unsigned long long)g.m0 * g.half * k1;
      const bool hi = e >= g.half;
      if constexpr (NOISE == NOISE_PHILOX_FAST) {
        const uint4 w = philox4x32_r<7>(c, realization, keys);
        nz = hi ? box_muller_fast(w.z, w.w) : box_muller_fast(w.x, w.y);
      } else {
        const uint4 w = philox4x32_r<10>(c, realization, keys);
        nz = hi ? box_muller_f64(w.z, w.w) : box_muller_f64(w.x, w.y);
      }
    }
    const double s = sig[i];
    cplx v = make_double2(s * nz.x, s * nz.y);
    if (hermitian) {
      const bool self0 = (2 * k0) % g.n[0] == 0;
      const bool selfc = self0 && (2 * k1) % g.n[1] == 0 && (2 * k2) % g.n[2] == 0;
      const bool plane = corrected && self0;
      if (plane)
        v = make_double2(v.x * SQRT2, v.y * SQRT2);
      if (selfc)
        v = make_double2(v.x * (plane ? 1.0 : SQRT2), 0.);
    }
    X[i] = v;
  }
}

// spectral product: X *= g(lambda) * scale (fieldBackend.mult, matrix.hh:1495-1513)
__global__ void
k_generic_mult(cplx* __restrict__ X, const double* __restrict__ tab, double scale,
               long long total)
{
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const double f = tab[i] * scale;
    X[i] = make_double2(X[i].x * f, X[i].y * f);
  }
}

} // namespace pf
