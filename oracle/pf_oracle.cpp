// ---------------------------------------------------------------------------
// pf_oracle.cpp -- CPU restatement of the parafields-core circulant-embedding
// hot path.  TEST INFRASTRUCTURE ONLY.
//
// This file is the parity oracle and the CPU baseline ("port") for the CUDA
// product path.  Nothing in the product (dune-parafields_b200/, include/)
// links, loads or calls it.  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may use it.
//
// Parity status: PINNED BY THE REFERENCE ITSELF.  The reference ships no
// golden vectors, but its path compiles from its own UNMODIFIED headers once
// the absent third-party libraries (FFTW3-MPI, MPI, dune-common) are replaced
// by single-rank stand-ins (oracle/shim/, oracle/ref_driver.cpp ->
// oracle/_ref/libpf_ref.so, built by oracle/Makefile).  This restatement is
// checked against that build live (tests/test_reference_live.py) and through
// committed fixtures it produced (tests/golden/ref_*.npz, tests/test_golden.py),
// in addition to first-principles identities (naive O(N^2) DFTs, dense
// circulant algebra, numpy/pocketfft) and the libstdc++ RNG known-answer
// values of SURVEY.md section 8(a) row a11.  What stays unpinned is FFTW's own
// rounding (library absent) -- see DESIGN.md section 4.
//
// Every function cites the reference file:line it follows.  Paths are
// relative to /root/reference/ext/parafields-core/include/parafields/.
// The FFTW3 calls of the reference (third party, version unpinned,
// backends/fftwwrapper.hh:230-415) are replaced by a plain in-file FFT; the
// transform conventions (sign, normalisation, c2r semantics on non-Hermitian
// input, transposed-out layout) follow FFTW's published definitions.
// ---------------------------------------------------------------------------
#include <algorithm>
#include <array>
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstring>
#include <limits>
#include <random>
#include <stdexcept>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

using cd = std::complex<double>;
typedef unsigned int Index; // fieldtraits.hh:80

// --------------------------------------------------------------------------
// configuration mirror (the subset of ParameterTree keys on the hot path,
// fieldtraits.hh:166-206, matrix.hh:283-284,496-508)
// --------------------------------------------------------------------------
struct OrcCfg
{
  int dim;              // 1,2,3
  unsigned cells[3];    // grid.cells
  double ext[3];        // grid.extensions
  unsigned factor;      // embedding.factor (periodic => 1)
  int cov;              // covariance id, see enum below
  double variance;      // stochastic.variance
  double p0, p1;        // expGamma | cauchyAlpha , cauchyBeta
  int aniso;            // 0 none, 1 axiparallel, 2 geometric
  double corr[9];       // corrLength: 1 | dim | dim*dim numbers (row major)
  double threshold;     // embedding.threshold (1e-14)
  int approximate;      // embedding.approximate
  int matrix_backend;   // 0 DCT, 1 R2C, 2 DFT (matrix.hh:1668-1767)
  int field_backend;    // 0 R2C, 1 DFT
  int transposed;       // fftw.transposed after the demotion rules
                        // (fieldtraits.hh:226-264): 3D -> 1, else 0
  unsigned cg_iterations;
  int threads;          // OpenMP threads for the FFT loops (1 = reference)
  // embedding.periodization = merge (matrix.hh:530-572,662-787)
  int periodization;    // 0 classical, 1 merge
  int sigmoid;          // embedding.sigmoid: 0 smoothstep, 1 smooth
  int sigmoid_combine;  // stochastic.sigmoidCombine: 0 radialPlus, 1 radial, 2 tensor
  double sigmoid_start; // embedding.sigmoidStart (1)
  double sigmoid_end;   // embedding.sigmoidEnd (factor - 1)
  unsigned recursions;  // embedding.mergeRecursions after the 99 heuristic
};

enum
{
  COV_EXPONENTIAL = 0,
  COV_GAUSSIAN = 1,
  COV_SPHERICAL = 2,
  COV_SEPARABLE_EXP = 3,
  COV_MATERN32 = 4,
  COV_MATERN52 = 5,
  COV_DAMPED_OSC = 6,
  COV_GAMMA_EXP = 7,
  COV_CAUCHY = 8,
  COV_GEN_CAUCHY = 9,
  COV_CUBIC = 10,
  COV_WHITE_NOISE = 11,
  COV_MATERN = 12
};

struct Stats
{
  int small_, negative, zero;
  double smallest;
};

// --------------------------------------------------------------------------
// plain FFT (stand-in for FFTW3): iterative radix-2 for powers of two,
// Bluestein otherwise.  sign = -1 forward, +1 backward, unnormalised, which
// is FFTW's definition of FFTW_FORWARD / FFTW_BACKWARD.
// --------------------------------------------------------------------------
struct FFT1D
{
  size_t n;
  bool pow2;
  std::vector<cd> tw;       // e^{-2 pi i k / n}, k < n/2 (pow2)
  std::vector<unsigned> rev;
  // Bluestein
  size_t m;
  std::vector<cd> chirp;    // e^{-i pi k^2 / n}
  std::vector<cd> bfft;     // FFT_m of conj chirp kernel (forward sign)
  FFT1D* sub;

  explicit FFT1D(size_t n_) : n(n_), m(0), sub(nullptr)
  {
    pow2 = (n & (n - 1)) == 0;
    if (pow2) {
      tw.resize(n / 2 ? n / 2 : 1);
      for (size_t k = 0; k < n / 2; k++) {
        long double a = -2.0L * M_PIl * (long double)k / (long double)n;
        tw[k] = cd((double)cosl(a), (double)sinl(a));
      }
      rev.resize(n);
      unsigned bits = 0;
      while ((size_t(1) << bits) < n)
        bits++;
      for (size_t i = 0; i < n; i++) {
        unsigned r = 0;
        for (unsigned b = 0; b < bits; b++)
          if (i & (size_t(1) << b))
            r |= 1u << (bits - 1 - b);
        rev[i] = r;
      }
    } else {
      m = 1;
      while (m < 2 * n - 1)
        m <<= 1;
      sub = new FFT1D(m);
      chirp.resize(n);
      for (size_t k = 0; k < n; k++) {
        // k^2 mod 2n keeps the argument small
        unsigned long long k2 = ((unsigned long long)k * k) % (2ull * n);
        long double a = -M_PIl * (long double)k2 / (long double)n;
        chirp[k] = cd((double)cosl(a), (double)sinl(a));
      }
      bfft.assign(m, cd(0, 0));
      bfft[0] = std::conj(chirp[0]);
      for (size_t k = 1; k < n; k++)
        bfft[k] = bfft[m - k] = std::conj(chirp[k]);
      sub->exec(bfft.data(), -1);
    }
  }
  ~FFT1D() { delete sub; }

  void exec(cd* x, int sign) const
  {
    if (n <= 1)
      return;
    if (pow2) {
      for (size_t i = 0; i < n; i++)
        if (rev[i] > i)
          std::swap(x[i], x[rev[i]]);
      for (size_t len = 2; len <= n; len <<= 1) {
        const size_t half = len / 2, step = n / len;
        for (size_t i = 0; i < n; i += len)
          for (size_t k = 0; k < half; k++) {
            cd w = tw[k * step];
            if (sign > 0)
              w = std::conj(w);
            const cd u = x[i + k];
            const cd v = x[i + k + half] * w;
            x[i + k] = u + v;
            x[i + k + half] = u - v;
          }
      }
    } else {
      // Bluestein: X_k = c_k * sum_j (x_j c_j) conj(c_{k-j}), forward sign;
      // backward = conj(forward(conj x))
      std::vector<cd> a(m, cd(0, 0));
      for (size_t k = 0; k < n; k++) {
        const cd xin = sign > 0 ? std::conj(x[k]) : x[k];
        a[k] = xin * chirp[k];
      }
      sub->exec(a.data(), -1);
      for (size_t k = 0; k < m; k++)
        a[k] *= bfft[k];
      sub->exec(a.data(), +1);
      const double inv = 1.0 / (double)m;
      for (size_t k = 0; k < n; k++) {
        const cd r = a[k] * inv * chirp[k];
        x[k] = sign > 0 ? std::conj(r) : r;
      }
    }
  }
};

// complex FFT along one axis of a dim-0-fastest array with shape n[0..d)
static void
fft_axis(cd* data, const size_t* n, int d, int axis, int sign, int threads)
{
  size_t stride = 1;
  for (int i = 0; i < axis; i++)
    stride *= n[i];
  size_t outer = 1;
  for (int i = axis + 1; i < d; i++)
    outer *= n[i];
  const size_t len = n[axis];
  FFT1D plan(len);
  const long long lines = (long long)(stride * outer);
  (void)threads;
#pragma omp parallel num_threads(threads > 0 ? threads : 1)
  {
    std::vector<cd> buf(len);
#pragma omp for schedule(static)
    for (long long l = 0; l < lines; l++) {
      const size_t in = (size_t)l % stride, out = (size_t)l / stride;
      cd* base = data + in + out * stride * len;
      if (stride == 1) {
        plan.exec(base, sign);
      } else {
        for (size_t j = 0; j < len; j++)
          buf[j] = base[j * stride];
        plan.exec(buf.data(), sign);
        for (size_t j = 0; j < len; j++)
          base[j * stride] = buf[j];
      }
    }
  }
}

// --------------------------------------------------------------------------
// index helpers, fieldtraits.hh:397-427 (dimension 0 is the fastest)
// --------------------------------------------------------------------------
static inline size_t
indicesToIndex(const size_t* idx, const size_t* bound, int d)
{
  size_t r = 0;
  for (int i = d - 1; i >= 0; i--)
    r = r * bound[i] + idx[i];
  return r;
}
static inline void
indexToIndices(size_t index, size_t* idx, const size_t* bound, int d)
{
  for (int i = 0; i < d; i++) {
    idx[i] = index % bound[i];
    index /= bound[i];
  }
}

// --------------------------------------------------------------------------
// geometry transforms, covariance.hh:17-136
// --------------------------------------------------------------------------
struct Geometry
{
  int dim, kind;
  double m[9];
  Geometry(const OrcCfg& c) : dim(c.dim), kind(c.aniso)
  {
    if (kind == 0) {
      m[0] = 1. / c.corr[0]; // covariance.hh:31-35
    } else if (kind == 1) {
      for (int i = 0; i < dim; i++)
        m[i] = 1. / c.corr[i]; // covariance.hh:60-65
    } else {
      // covariance.hh:92-125 (adjugate / determinant, same operation order)
      double copy[9];
      for (int i = 0; i < dim * dim; i++)
        copy[i] = m[i] = c.corr[i];
      if (dim == 3) {
        m[0] = copy[4] * copy[8] - copy[5] * copy[7];
        m[1] = copy[2] * copy[7] - copy[1] * copy[8];
        m[2] = copy[1] * copy[5] - copy[2] * copy[4];
        m[3] = copy[5] * copy[6] - copy[3] * copy[8];
        m[4] = copy[0] * copy[8] - copy[2] * copy[6];
        m[5] = copy[2] * copy[3] - copy[0] * copy[5];
        m[6] = copy[3] * copy[7] - copy[4] * copy[6];
        m[7] = copy[1] * copy[6] - copy[0] * copy[7];
        m[8] = copy[0] * copy[4] - copy[1] * copy[3];
        const double det = copy[0] * (copy[4] * copy[8] - copy[5] * copy[7]) -
                           copy[1] * (copy[3] * copy[8] - copy[5] * copy[6]) +
                           copy[2] * (copy[3] * copy[7] - copy[4] * copy[6]);
        for (int i = 0; i < 9; i++)
          m[i] /= det;
      } else if (dim == 2) {
        m[0] = copy[3];
        m[1] = -copy[1];
        m[2] = -copy[2];
        m[3] = copy[0];
        const double det = copy[0] * copy[3] - copy[1] * copy[2];
        for (int i = 0; i < 4; i++)
          m[i] /= det;
      } else {
        m[0] = 1. / m[0];
      }
    }
  }
  void transform(const double* x, double* xt) const
  {
    if (kind == 0)
      for (int i = 0; i < dim; i++)
        xt[i] = m[0] * x[i];
    else if (kind == 1)
      for (int i = 0; i < dim; i++)
        xt[i] = m[i] * x[i];
    else
      for (int i = 0; i < dim; i++) {
        xt[i] = 0.;
        for (int j = 0; j < dim; j++)
          xt[i] += m[i * dim + j] * x[j];
      }
  }
};

// --------------------------------------------------------------------------
// covariance functors, covariance.hh:294-890 (verbatim arithmetic)
// --------------------------------------------------------------------------
static double
covariance(const OrcCfg& c, const double* x)
{
  const int dim = c.dim;
  const double variance = c.variance;
  if (c.cov == COV_SEPARABLE_EXP) { // covariance.hh:483-506
    double sum = 0.;
    for (int i = 0; i < dim; i++)
      sum += std::abs(x[i]);
    return variance * std::exp(-sum);
  }
  if (c.cov == COV_WHITE_NOISE) { // covariance.hh:859-880
    for (int i = 0; i < dim; i++)
      if (std::abs(x[i]) > 1e-10)
        return 0.;
    return variance;
  }
  double sum = 0.;
  for (int i = 0; i < dim; i++)
    sum += x[i] * x[i];
  const double h_eff = std::sqrt(sum);
  switch (c.cov) {
    case COV_EXPONENTIAL: // :355-385
      return variance * std::exp(-h_eff);
    case COV_GAUSSIAN: // :444-474
      return variance * std::exp(-h_eff * h_eff);
    case COV_SPHERICAL: // :294-346
      if (dim == 3) {
        if (h_eff > 1.)
          return 0.;
        return variance * (1. - 1.5 * h_eff + 0.5 * std::pow(h_eff, 3));
      } else if (dim == 2) {
        if (h_eff > 1.)
          return 0.;
        return variance *
               (1. - 2. / M_PI * (h_eff * std::sqrt(1. - std::pow(h_eff, 2)) +
                                  std::asin(h_eff)));
      } else {
        if (h_eff > 1.)
          return 0.;
        return variance * (1. - h_eff);
      }
    case COV_MATERN32: // :591-623
      return variance * (1. + std::sqrt(3.) * h_eff) *
             std::exp(-std::sqrt(3.) * h_eff);
    case COV_MATERN52: // :632-664
      return variance * (1. + std::sqrt(5.) * h_eff + 5. / 3. * h_eff * h_eff) *
             std::exp(-std::sqrt(5.) * h_eff);
    case COV_DAMPED_OSC: // :672-704
      if (dim == 3)
        return variance * std::exp(-h_eff) * std::cos(h_eff / std::sqrt(3.));
      return variance * std::exp(-h_eff) * std::cos(h_eff);
    case COV_GAMMA_EXP: // :396-436
      return variance * std::exp(-std::pow(h_eff, c.p0));
    case COV_CAUCHY: // :713-744
      return variance * std::pow(1. + std::pow(h_eff, 2), -3);
    case COV_GEN_CAUCHY: // :753-797
      return variance * std::pow(1. + std::pow(h_eff, c.p0), -c.p1);
    case COV_CUBIC: // :806-850
      if (dim == 3)
        throw std::runtime_error{
          "cubic covariance only applicable in 1D or 2D"
        };
      if (h_eff > 1.)
        return 0.;
      return variance *
             (1. - 7. * std::pow(h_eff, 2) + 8.75 * std::pow(h_eff, 3) -
              3.5 * std::pow(h_eff, 5) + 0.75 * std::pow(h_eff, 7));
  }
  if (c.cov == COV_MATERN) { // covariance.hh:527-581, general nu = c.p0
    // The reference calls gsl_sf_bessel_Knu (GSL, unpinned version, absent
    // here); libstdc++'s std::cyl_bessel_k is the stand-in for it, and
    // tests/test_oracle.py checks this functor against scipy.special.kv.
    const double nu = c.p0;
    if (nu < 0.)
      throw std::runtime_error{ "matern nu has to be positive" };
    if (h_eff < 1e-10)
      return variance;
    const double sqrtTwoNu = std::sqrt(2. * nu);
    const double z = sqrtTwoNu * h_eff;
    if (z > 700.) // K_nu underflows; std::cyl_bessel_k may raise instead
      return 0.;
    return variance * std::pow(2., 1. - nu) / std::tgamma(nu) *
           std::pow(z, nu) * std::cyl_bessel_k(nu, z);
  }
  throw std::runtime_error{ "covariance structure not known" };
}

// --------------------------------------------------------------------------
// sizes, fieldtraits.hh:214-311 (single rank)
// --------------------------------------------------------------------------
struct Sizes
{
  int dim;
  size_t c[3], n[3];  // cells, extendedCells
  double h[3];        // meshsize
  size_t N, D;        // extendedDomainSize, domainSize
  size_t m0;          // n0/2+1
  Sizes(const OrcCfg& cfg) : dim(cfg.dim)
  {
    N = D = 1;
    for (int i = 0; i < 3; i++) {
      c[i] = n[i] = 1;
      h[i] = 0;
    }
    for (int i = 0; i < dim; i++) {
      c[i] = cfg.cells[i];
      n[i] = (size_t)cfg.factor * cfg.cells[i]; // :268
      h[i] = cfg.ext[i] / cfg.cells[i];         // :267
      N *= n[i];
      D *= c[i];
    }
    m0 = n[0] / 2 + 1;
  }
};

// SmoothSigmoid (covariance.hh:145-176), ClampingSigmoid (:202-218),
// SmoothstepSigmoid (:244-275)
static inline double
sigmoid_value(int kind, double oneEdge, double zeroEdge, double x,
              unsigned recursions)
{
  if (kind == 1) {
    if (std::abs(x) < oneEdge)
      return 1.;
    if (std::abs(x) > zeroEdge)
      return 0.;
    const double x1 = (zeroEdge - std::abs(x)) / (zeroEdge - oneEdge);
    const double x2 = (std::abs(x) - oneEdge) / (zeroEdge - oneEdge);
    const double t1 = x1 > 0. ? std::exp(-1. / x1) : 0.;
    const double t2 = x2 > 0. ? std::exp(-1. / x2) : 0.;
    return t1 / (t1 + t2);
  }
  double x1;
  if (std::abs(x) < oneEdge)
    x1 = 1.;
  else if (std::abs(x) > zeroEdge)
    x1 = 0.;
  else
    x1 = (zeroEdge - std::abs(x)) / (zeroEdge - oneEdge);
  const double x2 = 2. * x1 - 1.;
  double normalization = 1.;
  double addition = 1.;
  double output = 1.;
  for (unsigned int i = 1; i <= recursions; i++) {
    output *= 2. * i;
    addition *= 1 - x2 * x2;
    output += addition;
    output /= 2. * i + 1.;
    normalization *= 2. * i / (2. * i + 1.);
  }
  output *= x2 / normalization;
  output += 1.;
  output /= 2.;
  return output;
}

// Matrix::computeMatrixEntriesWithMerge, matrix.hh:662-787: 2^dim mirrored
// copies of the covariance, each tapered by a sigmoid, the rest filled with
// the value at the far corner (radialPlus)
static inline double
torus_cov_merge(const OrcCfg& cfg, const Sizes& s, const Geometry& g,
                const size_t* idx)
{
  const int dim = s.dim;
  const bool radial = cfg.sigmoid_combine != 2;
  double trueCoord[3], mirrorCoord[3], transCoord[3];
  double constValue = 0.;
  if (cfg.sigmoid_combine == 0) {
    for (int i = 0; i < dim; i++)
      trueCoord[i] = cfg.ext[i] * cfg.factor / 2.;
    g.transform(trueCoord, transCoord);
    constValue = covariance(cfg, transCoord);
  }
  for (int i = 0; i < dim; i++)
    trueCoord[i] = (unsigned)idx[i] * s.h[i];
  double value = 0.;
  double remainder = 1.;
  for (unsigned int j = 0; j < (1u << dim); j++) {
    for (int i = 0; i < dim; i++)
      mirrorCoord[i] =
        (j & (1u << i)) ? cfg.ext[i] * cfg.factor - trueCoord[i] : trueCoord[i];
    double dampening = 1.;
    const double eps = 1e-10;
    const double sqrtDim = std::sqrt(double(dim));
    const unsigned int factor = cfg.factor;
    if (radial && sqrtDim < factor / 2.) {
      double radius = 0.;
      for (int i = 0; i < dim; i++)
        radius += (mirrorCoord[i] * mirrorCoord[i]) / (cfg.ext[i] * cfg.ext[i]);
      radius = std::sqrt(radius);
      dampening = sigmoid_value(cfg.sigmoid, cfg.sigmoid_start * sqrtDim - eps,
                                factor - (factor - cfg.sigmoid_end) * sqrtDim + eps,
                                radius, cfg.recursions);
    } else {
      for (int i = 0; i < dim; i++)
        dampening *= sigmoid_value(cfg.sigmoid, cfg.ext[i] * cfg.sigmoid_start - eps,
                                   cfg.ext[i] * cfg.sigmoid_end + eps,
                                   mirrorCoord[i], cfg.recursions);
    }
    if (dampening < eps)
      continue;
    remainder -= dampening;
    g.transform(mirrorCoord, transCoord);
    value = value + covariance(cfg, transCoord) * dampening;
  }
  return value + constValue * remainder;
}

// torus covariance value at torus cell idx (matrix.hh:637-646)
static inline double
torus_cov(const OrcCfg& cfg, const Sizes& s, const Geometry& g,
          const size_t* idx)
{
  if (cfg.periodization == 1)
    return torus_cov_merge(cfg, s, g, idx);
  double coord[3], tc[3];
  for (int i = 0; i < s.dim; i++) {
    coord[i] = (unsigned)idx[i] * s.h[i];
    if (coord[i] > 0.5 * cfg.ext[i] * cfg.factor)
      coord[i] -= cfg.ext[i] * cfg.factor;
  }
  g.transform(coord, tc);
  return covariance(cfg, tc);
}

// --------------------------------------------------------------------------
// Eigenvalues.  matrix.hh:232-331 + the three matrix backends.
// Output layout = the reference's stored layout after finalize():
//   DCT  (dctmatrixbackend.hh:240-272,346-355): compact prod(n_i/2+1) reals,
//        dim0 fastest; in 3D with transposed=1 the last two dims are swapped
//        (storage bounds (d0,d2,d1), value at [k0 + d0*(k2 + d2*k1)]).
//   R2C  (r2cmatrixbackend.hh:254-285,412-421): m0 * n1 (* n2) reals (real
//        parts of the half spectrum), transposed likewise in 3D.
//   DFT  (dftmatrixbackend.hh:209-238): n0 (* n1 * n2) reals (real parts).
// The census follows matrix.hh:283-306 on the stored entries; for R2C the
// reference walks the interleaved re/im doubles of the first half of the
// entries (r2cmatrixbackend.hh:164-170,378-382) -- reproduced when
// census_quirk != 0.
// --------------------------------------------------------------------------
static void
census_value(double value, double threshold, Stats& st)
{
  if (value < st.smallest)
    st.smallest = value;
  if (value < 1e-6) {
    if (value < threshold) {
      if (value > -threshold)
        st.zero++;
      else
        st.negative++;
    } else
      st.small_++;
  }
}

static size_t
lambda_size(const OrcCfg& cfg)
{
  Sizes s(cfg);
  size_t r = 1;
  for (int i = 0; i < s.dim; i++) {
    if (cfg.matrix_backend == 0)
      r *= s.n[i] / 2 + 1;
    else if (cfg.matrix_backend == 1)
      r *= (i == 0 ? s.m0 : s.n[i]);
    else
      r *= s.n[i];
  }
  return r;
}

static void
eigenvalues(const OrcCfg& cfg, double* lambda, Stats& st, int census_quirk)
{
  Sizes s(cfg);
  Geometry g(cfg);
  const int d = s.dim;
  st.small_ = st.negative = st.zero = 0;
  st.smallest = std::numeric_limits<double>::max();

  // full torus covariance as complex array (the DCT-I of the even corner,
  // dctmatrixbackend.hh:252-266, equals the full DFT of the even extension;
  // the r2c transform, r2cmatrixbackend.hh:266-277, equals the full DFT
  // restricted to k0 <= n0/2)
  std::vector<cd> full(s.N);
  size_t bound[3] = { s.n[0], s.n[1], s.n[2] };
  const bool dct = cfg.matrix_backend == 0;
#pragma omp parallel for num_threads(cfg.threads > 0 ? cfg.threads : 1)
  for (long long index = 0; index < (long long)s.N; index++) {
    size_t idx[3] = { 0, 0, 0 };
    indexToIndices((size_t)index, idx, bound, d);
    if (dct) {
      // only the corner is evaluated by the reference
      // (matrix.hh:634-647 over localDCTCells); the even extension supplies
      // the rest: value at mirrored index
      size_t mi[3];
      for (int i = 0; i < d; i++)
        mi[i] = idx[i] > s.n[i] / 2 ? s.n[i] - idx[i] : idx[i];
      full[index] = torus_cov(cfg, s, g, mi);
    } else
      full[index] = torus_cov(cfg, s, g, idx);
  }
  for (int ax = 0; ax < d; ax++)
    fft_axis(full.data(), bound, d, ax, -1, cfg.threads);
  const double Nd = (double)s.N; // "/= extendedDomainSize"

  // stored bounds (natural order) and the transposed storage permutation
  size_t sb[3] = { 1, 1, 1 };
  for (int i = 0; i < d; i++)
    sb[i] = dct ? s.n[i] / 2 + 1
                : (cfg.matrix_backend == 1 && i == 0 ? s.m0 : s.n[i]);
  const bool tr = cfg.transposed && d == 3;
  const size_t total = sb[0] * sb[1] * sb[2];
  std::vector<cd> stored(total);
  for (size_t k2 = 0; k2 < sb[2]; k2++)
    for (size_t k1 = 0; k1 < sb[1]; k1++)
      for (size_t k0 = 0; k0 < sb[0]; k0++) {
        const cd v = full[k0 + s.n[0] * (k1 + s.n[1] * k2)] / Nd;
        size_t o;
        if (tr)
          o = k0 + sb[0] * (k2 + sb[2] * k1);
        else
          o = k0 + sb[0] * (k1 + sb[1] * k2);
        stored[o] = v;
      }

  // census + clamp, matrix.hh:289-306
  if (cfg.matrix_backend == 1 && census_quirk) {
    // interleaved walk over the first `total` doubles of the complex array
    double* raw = reinterpret_cast<double*>(stored.data());
    for (size_t i = 0; i < total; i++) {
      census_value(raw[i], cfg.threshold, st);
      if (raw[i] < 0.)
        raw[i] = 0.;
    }
  } else {
    for (size_t i = 0; i < total; i++) {
      double v = stored[i].real();
      census_value(v, cfg.threshold, st);
      if (v < 0.)
        stored[i] = cd(0., stored[i].imag());
    }
  }
  for (size_t i = 0; i < total; i++)
    lambda[i] = stored[i].real();
}

// lambda lookup for a half-spectrum / full index tuple in NATURAL (k0,k1,k2)
// order.  dctmatrixbackend.hh:346-355 and r2cmatrixbackend.hh:354-362 mirror
// each index into the stored range; here the storage transposition is undone.
static inline double
lambda_at(const OrcCfg& cfg, const Sizes& s, const double* lambda,
          const size_t* k)
{
  const int d = s.dim;
  size_t sb[3] = { 1, 1, 1 }, kk[3] = { 0, 0, 0 };
  for (int i = 0; i < d; i++) {
    if (cfg.matrix_backend == 0) {
      sb[i] = s.n[i] / 2 + 1;
      kk[i] = k[i] >= sb[i] ? s.n[i] - k[i] : k[i];
    } else if (cfg.matrix_backend == 1) {
      sb[i] = i == 0 ? s.m0 : s.n[i];
      kk[i] = k[i] >= sb[i] ? s.n[i] - k[i] : k[i];
    } else {
      sb[i] = s.n[i];
      kk[i] = k[i];
    }
  }
  if (cfg.transposed && d == 3)
    return lambda[kk[0] + sb[0] * (kk[2] + sb[2] * kk[1])];
  return lambda[kk[0] + sb[0] * (kk[1] + sb[1] * kk[2])];
}

// --------------------------------------------------------------------------
// c2r with FFTW semantics on a (possibly non-Hermitian) half spectrum held in
// natural order X[k0 + m0*(k1 + n1*k2)]: complex backward transforms along
// axes >= 1, then a 1-D half-complex -> real transform along axis 0 that
// ignores Im at k0 in {0, n0/2}.  (r2cfieldbackend.hh:250-274; FFTW manual,
// "Multi-dimensional DFTs of real data".)  Output real[x0 + n0*(x1 + n1*x2)].
// --------------------------------------------------------------------------
static void
c2r(const Sizes& s, std::vector<cd>& X, std::vector<double>& real, int threads)
{
  const int d = s.dim;
  size_t hb[3] = { s.m0, s.n[1], s.n[2] };
  for (int ax = 1; ax < d; ax++)
    fft_axis(X.data(), hb, d, ax, +1, threads);
  const size_t n0 = s.n[0], m0 = s.m0, lines = s.N / n0;
  real.assign(s.N, 0.);
  FFT1D plan(n0);
#pragma omp parallel num_threads(threads > 0 ? threads : 1)
  {
    std::vector<cd> buf(n0);
#pragma omp for schedule(static)
    for (long long l = 0; l < (long long)lines; l++) {
      const cd* in = X.data() + (size_t)l * m0;
      buf[0] = cd(in[0].real(), 0.);
      for (size_t k = 1; k < m0; k++) {
        if (2 * k == n0)
          buf[k] = cd(in[k].real(), 0.);
        else {
          buf[k] = in[k];
          buf[n0 - k] = std::conj(in[k]);
        }
      }
      plan.exec(buf.data(), +1);
      double* out = real.data() + (size_t)l * n0;
      for (size_t j = 0; j < n0; j++)
        out[j] = buf[j].real();
    }
  }
}

// r2c (forward, unnormalised) of a real torus array into the natural-order
// half spectrum (r2cfieldbackend.hh:210-239 before the division)
static void
r2c(const Sizes& s, const std::vector<double>& real, std::vector<cd>& X,
    int threads)
{
  const int d = s.dim;
  const size_t n0 = s.n[0], m0 = s.m0, lines = s.N / n0;
  X.assign(m0 * lines, cd(0, 0));
  FFT1D plan(n0);
#pragma omp parallel num_threads(threads > 0 ? threads : 1)
  {
    std::vector<cd> buf(n0);
#pragma omp for schedule(static)
    for (long long l = 0; l < (long long)lines; l++) {
      const double* in = real.data() + (size_t)l * n0;
      for (size_t j = 0; j < n0; j++)
        buf[j] = cd(in[j], 0.);
      plan.exec(buf.data(), -1);
      for (size_t k = 0; k < m0; k++)
        X[(size_t)l * m0 + k] = buf[k];
    }
  }
  size_t hb[3] = { m0, s.n[1], s.n[2] };
  for (int ax = 1; ax < d; ax++)
    fft_axis(X.data(), hb, d, ax, -1, threads);
}

// --------------------------------------------------------------------------
// generateField from a given noise stream, matrix.hh:343-438.
// noise = r1,r2 pairs in the reference's stream order:
//   R2C backend: s = k0 + m0*k1 (2D), s = k0 + m0*(k2 + n2*k1) (3D
//   transposed, r2cfieldbackend.hh:190-199), natural order otherwise;
//   DFT backend (1D default): s = k0.
// out  = field (D doubles); spare (D doubles, DFT backend only, may be null)
// --------------------------------------------------------------------------
static void
generate_from_noise(const OrcCfg& cfg, const double* lambda,
                    const double* noise, double* out, double* spare)
{
  Sizes s(cfg);
  const int d = s.dim;
  if (cfg.field_backend == 1) {
    // DFTFieldBackend::set, dftfieldbackend.hh:276-280
    std::vector<cd> Z(s.N);
    size_t bound[3] = { s.n[0], s.n[1], s.n[2] };
    const bool tr = cfg.transposed && d == 3;
    for (size_t idx = 0; idx < s.N; idx++) {
      size_t k[3] = { 0, 0, 0 };
      if (tr) {
        size_t b[3] = { s.n[0], s.n[2], s.n[1] }, t[3];
        indexToIndices(idx, t, b, 3);
        k[0] = t[0];
        k[1] = t[2];
        k[2] = t[1];
      } else
        indexToIndices(idx, k, bound, d);
      const double lam = std::sqrt(lambda_at(cfg, s, lambda, k));
      const size_t nat = k[0] + s.n[0] * (k[1] + s.n[1] * k[2]);
      Z[nat] = cd(lam * noise[2 * idx], lam * noise[2 * idx + 1]);
    }
    for (int ax = 0; ax < d; ax++)
      fft_axis(Z.data(), bound, d, ax, +1, cfg.threads);
    // extendedFieldToField, dftfieldbackend.hh:380-393
    for (size_t i2 = 0; i2 < s.c[2]; i2++)
      for (size_t i1 = 0; i1 < s.c[1]; i1++)
        for (size_t i0 = 0; i0 < s.c[0]; i0++) {
          const cd v = Z[i0 + s.n[0] * (i1 + s.n[1] * i2)];
          const size_t o = i0 + s.c[0] * (i1 + s.c[1] * i2);
          out[o] = v.real();
          if (spare)
            spare[o] = v.imag();
        }
    return;
  }

  // R2CFieldBackend
  static const double sqrtTwo = std::sqrt(2.);
  const size_t m0 = s.m0;
  const size_t S = m0 * s.n[1] * s.n[2];
  std::vector<cd> X(S);
  const bool tr = cfg.transposed && d == 3;
#pragma omp parallel for num_threads(cfg.threads > 0 ? cfg.threads : 1)
  for (long long idx_ = 0; idx_ < (long long)S; idx_++) {
    const size_t idx = (size_t)idx_;
    size_t k[3] = { 0, 0, 0 };
    if (tr) {
      size_t b[3] = { m0, s.n[2], s.n[1] }, t[3];
      indexToIndices(idx, t, b, 3);
      k[0] = t[0];
      k[1] = t[2];
      k[2] = t[1];
    } else {
      size_t b[3] = { m0, s.n[1], s.n[2] };
      indexToIndices(idx, k, b, d);
    }
    const double lam = std::sqrt(lambda_at(cfg, s, lambda, k)); // matrix.hh:414
    const double rand1 = noise[2 * idx], rand2 = noise[2 * idx + 1];
    // R2CFieldBackend::set, r2cfieldbackend.hh:305-325
    bool allMultiple = true;
    for (int i = 0; i < d; i++)
      if ((2 * k[i]) % s.n[i] != 0)
        allMultiple = false;
    cd v;
    if (allMultiple)
      v = cd(lam * rand1, 0.);
    else
      v = cd(lam / sqrtTwo * rand1, lam / sqrtTwo * rand2);
    X[k[0] + m0 * (k[1] + s.n[1] * k[2])] = v;
  }
  std::vector<double> real;
  c2r(s, X, real, cfg.threads);
  // extendedFieldToField, r2cfieldbackend.hh:423-441
#pragma omp parallel for num_threads(cfg.threads > 0 ? cfg.threads : 1)
  for (long long i2_ = 0; i2_ < (long long)s.c[2]; i2_++) {
    const size_t i2 = (size_t)i2_;
    for (size_t i1 = 0; i1 < s.c[1]; i1++)
      for (size_t i0 = 0; i0 < s.c[0]; i0++)
        out[i0 + s.c[0] * (i1 + s.c[1] * i2)] =
          real[i0 + s.n[0] * (i1 + s.n[1] * i2)];
  }
}

static size_t
noise_entries(const OrcCfg& cfg)
{
  Sizes s(cfg);
  if (cfg.field_backend == 1)
    return s.N;
  return s.m0 * s.n[1] * s.n[2];
}

// --------------------------------------------------------------------------
// multiplyExtended / multiplyRootExtended / multiplyInverseExtended,
// matrix.hh:1459-1662.  op: 0 lambda, 1 sqrt(lambda), 2 1/lambda.
// --------------------------------------------------------------------------
static void
multiply_ext(const OrcCfg& cfg, const double* lambda, int op, const double* in,
             double* out)
{
  Sizes s(cfg);
  const int d = s.dim;
  const double Nd = (double)s.N;
  if (cfg.field_backend == 1) {
    std::vector<cd> Z(s.N, cd(0, 0));
    size_t bound[3] = { s.n[0], s.n[1], s.n[2] };
    // fieldToExtendedField, dftfieldbackend.hh:307-325
    for (size_t i2 = 0; i2 < s.c[2]; i2++)
      for (size_t i1 = 0; i1 < s.c[1]; i1++)
        for (size_t i0 = 0; i0 < s.c[0]; i0++)
          Z[i0 + s.n[0] * (i1 + s.n[1] * i2)] =
            cd(in[i0 + s.c[0] * (i1 + s.c[1] * i2)], 0.);
    for (int ax = 0; ax < d; ax++)
      fft_axis(Z.data(), bound, d, ax, -1, cfg.threads);
    for (size_t idx = 0; idx < s.N; idx++) {
      size_t k[3] = { 0, 0, 0 };
      indexToIndices(idx, k, bound, d);
      double lam = lambda_at(cfg, s, lambda, k);
      if (op == 1)
        lam = std::sqrt(lam);
      else if (op == 2)
        lam = 1. / lam;
      cd v = Z[idx];
      v = cd(v.real() / Nd, v.imag() / Nd); // dftfieldbackend.hh:forward
      Z[idx] = cd(v.real() * lam, v.imag() * lam);
    }
    for (int ax = 0; ax < d; ax++)
      fft_axis(Z.data(), bound, d, ax, +1, cfg.threads);
    for (size_t i2 = 0; i2 < s.c[2]; i2++)
      for (size_t i1 = 0; i1 < s.c[1]; i1++)
        for (size_t i0 = 0; i0 < s.c[0]; i0++)
          out[i0 + s.c[0] * (i1 + s.c[1] * i2)] =
            Z[i0 + s.n[0] * (i1 + s.n[1] * i2)].real();
    return;
  }
  // fieldToExtendedField, r2cfieldbackend.hh:352-368
  std::vector<double> real(s.N, 0.);
  for (size_t i2 = 0; i2 < s.c[2]; i2++)
    for (size_t i1 = 0; i1 < s.c[1]; i1++)
      for (size_t i0 = 0; i0 < s.c[0]; i0++)
        real[i0 + s.n[0] * (i1 + s.n[1] * i2)] =
          in[i0 + s.c[0] * (i1 + s.c[1] * i2)];
  std::vector<cd> X;
  r2c(s, real, X, cfg.threads);
  const size_t m0 = s.m0;
  for (size_t k2 = 0; k2 < s.n[2]; k2++)
    for (size_t k1 = 0; k1 < s.n[1]; k1++)
      for (size_t k0 = 0; k0 < m0; k0++) {
        size_t k[3] = { k0, k1, k2 };
        double lam = lambda_at(cfg, s, lambda, k);
        if (op == 1)
          lam = std::sqrt(lam);
        else if (op == 2)
          lam = 1. / lam;
        cd& v = X[k0 + m0 * (k1 + s.n[1] * k2)];
        v = cd(v.real() / Nd, v.imag() / Nd); // r2cfieldbackend.hh:233-236
        v = cd(v.real() * lam, v.imag() * lam); // :337-341
      }
  c2r(s, X, real, cfg.threads);
  for (size_t i2 = 0; i2 < s.c[2]; i2++)
    for (size_t i1 = 0; i1 < s.c[1]; i1++)
      for (size_t i0 = 0; i0 < s.c[0]; i0++)
        out[i0 + s.c[0] * (i1 + s.c[1] * i2)] =
          real[i0 + s.n[0] * (i1 + s.n[1] * i2)];
}

// --------------------------------------------------------------------------
// multiplyInverse + innerCG, matrix.hh:198-215,1331-1448 (single rank: the
// MPI_Allreduce calls are identities)
// --------------------------------------------------------------------------
static unsigned
multiply_inverse(const OrcCfg& cfg, const double* lambda, const double* input,
                 double* output)
{
  Sizes s(cfg);
  const size_t D = s.D;
  std::vector<double> solution(input, input + D);
  std::vector<double> iter(input, input + D);
  bool fieldZero = true;
  for (size_t i = 0; i < D; i++)
    if (std::abs(input[i]) > 1e-10)
      fieldZero = false;
  if (fieldZero) {
    std::copy(iter.begin(), iter.end(), output);
    return 0;
  }
  multiply_ext(cfg, lambda, 2, iter.data(), iter.data());

  std::vector<double> matrixTimesSolution(D), matrixTimesIter(D), residual(D),
    precResidual(D), direction(D), matrixTimesDirection(D);
  double scalarProd, scalarProd2, alphaDenominator, alpha, beta;
  multiply_ext(cfg, lambda, 0, solution.data(), matrixTimesSolution.data());
  multiply_ext(cfg, lambda, 0, iter.data(), matrixTimesIter.data());
  for (size_t i = 0; i < D; i++)
    residual[i] = solution[i] - matrixTimesIter[i];
  multiply_ext(cfg, lambda, 2, residual.data(), precResidual.data());
  direction = precResidual;
  bool converged = false;
  scalarProd = 0.;
  for (size_t i = 0; i < D; i++)
    scalarProd += precResidual[i] * residual[i];
  scalarProd2 = 0.;
  for (size_t i = 0; i < D; i++)
    scalarProd2 += residual[i] * residual[i];
  if (std::sqrt(std::abs(scalarProd2)) < 1e-6)
    converged = true;
  double firstValue = 0.;
  for (size_t i = 0; i < D; i++)
    firstValue += iter[i] * (0.5 * matrixTimesIter[i] - solution[i]);
  unsigned count = 0;
  while (!converged && count < cfg.cg_iterations) {
    multiply_ext(cfg, lambda, 0, direction.data(),
                 matrixTimesDirection.data());
    alphaDenominator = 0.;
    for (size_t i = 0; i < D; i++)
      alphaDenominator += direction[i] * matrixTimesDirection[i];
    alpha = scalarProd / alphaDenominator;
    double oldValue = 0.;
    for (size_t i = 0; i < D; i++)
      oldValue += iter[i] * (0.5 * matrixTimesIter[i] - solution[i]);
    for (size_t i = 0; i < D; i++) {
      iter[i] += alpha * direction[i];
      matrixTimesIter[i] += alpha * matrixTimesDirection[i];
    }
    double value = 0.;
    for (size_t i = 0; i < D; i++)
      value += iter[i] * (0.5 * matrixTimesIter[i] - solution[i]);
    for (size_t i = 0; i < D; i++)
      residual[i] = solution[i] - matrixTimesIter[i];
    multiply_ext(cfg, lambda, 2, residual.data(), precResidual.data());
    beta = 1. / scalarProd;
    scalarProd = 0.;
    for (size_t i = 0; i < D; i++)
      scalarProd += precResidual[i] * residual[i];
    beta *= scalarProd;
    for (size_t i = 0; i < D; i++)
      direction[i] = precResidual[i] + beta * direction[i];
    if (value != firstValue) {
      if (std::abs(value - oldValue) / std::abs(value - firstValue) < 1e-16)
        converged = true;
    }
    count++;
  }
  std::copy(iter.begin(), iter.end(), output);
  return count;
}

thread_local std::string g_err;

} // namespace

// ===========================================================================
// C entry points (ctypes)
// ===========================================================================
#define ORC_TRY try {
#define ORC_CATCH                                                              \
  }                                                                            \
  catch (const std::exception& e)                                              \
  {                                                                            \
    g_err = e.what();                                                          \
    return 1;                                                                  \
  }                                                                            \
  return 0;

extern "C" {

const char*
orc_last_error()
{
  return g_err.c_str();
}

size_t
orc_lambda_size(const OrcCfg* cfg)
{
  return lambda_size(*cfg);
}

size_t
orc_noise_entries(const OrcCfg* cfg)
{
  return noise_entries(*cfg);
}

// stats: [small, zero, negative] ints, smallest double
int
orc_eigenvalues(const OrcCfg* cfg, double* lambda, int* counts,
                double* smallest, int census_quirk)
{
  ORC_TRY
  Stats st;
  eigenvalues(*cfg, lambda, st, census_quirk);
  counts[0] = st.small_;
  counts[1] = st.zero;
  counts[2] = st.negative;
  *smallest = st.smallest;
  if (st.negative > 0 && !cfg->approximate) // matrix.hh:322-328
    throw std::runtime_error{ "negative eigenvalues in covariance matrix" };
  ORC_CATCH
}

// lambda in natural order on the half spectrum (m0,n1,n2) / full (DFT), for
// comparisons that should not depend on the storage transposition
int
orc_lambda_natural(const OrcCfg* cfg, const double* lambda, double* out)
{
  ORC_TRY
  Sizes s(*cfg);
  const size_t b0 = cfg->field_backend == 1 ? s.n[0] : s.m0;
  for (size_t k2 = 0; k2 < s.n[2]; k2++)
    for (size_t k1 = 0; k1 < s.n[1]; k1++)
      for (size_t k0 = 0; k0 < b0; k0++) {
        size_t k[3] = { k0, k1, k2 };
        out[k0 + b0 * (k1 + s.n[1] * k2)] = lambda_at(*cfg, s, lambda, k);
      }
  ORC_CATCH
}

int
orc_generate_from_noise(const OrcCfg* cfg, const double* lambda,
                        const double* noise, double* out, double* spare)
{
  ORC_TRY
  generate_from_noise(*cfg, lambda, noise, out, spare);
  ORC_CATCH
}

// the reference's noise stream: CppRNGBackend (backends/cpprngbackend.hh:49-58)
// = std::default_random_engine + std::normal_distribution<double>(0,1),
// reseeded with `seed` (+rank, rank = 0) at every generateField call
// (matrix.hh:353-354); two draws per stored entry (matrix.hh:416-417).
int
orc_reference_noise(unsigned seed, size_t entries, double* noise)
{
  ORC_TRY
  std::default_random_engine generator;
  std::normal_distribution<double> normalDist(0., 1.);
  generator.seed(seed);
  for (size_t i = 0; i < 2 * entries; i++)
    noise[i] = normalDist(generator);
  ORC_CATCH
}

// The same stream as produced by `ranks` MPI ranks of the reference: rank r
// seeds with seed + r (matrix.hh:353) and fills its slab, which is a
// contiguous chunk of the stream because the distributed dimension is the
// slowest one (fieldtraits.hh:271-282).  Used by the all-core CPU baseline.
int
orc_reference_noise_ranks(unsigned seed, size_t entries, int ranks,
                          double* noise)
{
  ORC_TRY
  if (ranks < 1 || entries % (size_t)ranks != 0)
    throw std::runtime_error{
      "number of cells in last dimension has to be multiple of numProc"
    };
  const size_t chunk = entries / (size_t)ranks;
#pragma omp parallel for num_threads(ranks) schedule(static, 1)
  for (int r = 0; r < ranks; r++) {
    std::default_random_engine generator;
    std::normal_distribution<double> normalDist(0., 1.);
    generator.seed(seed + (unsigned)r);
    double* dst = noise + 2 * chunk * (size_t)r;
    for (size_t i = 0; i < 2 * chunk; i++)
      dst[i] = normalDist(generator);
  }
  ORC_CATCH
}

// generateUncorrelatedField, matrix.hh:450-462
int
orc_generate_uncorrelated(unsigned seed, size_t n, double* out)
{
  ORC_TRY
  std::default_random_engine generator(seed);
  std::normal_distribution<double> normalDist(0., 1.);
  for (size_t i = 0; i < n; i++)
    out[i] = normalDist(generator);
  ORC_CATCH
}

int
orc_multiply(const OrcCfg* cfg, const double* lambda, int op, const double* in,
             double* out)
{
  ORC_TRY
  multiply_ext(*cfg, lambda, op, in, out);
  ORC_CATCH
}

int
orc_multiply_inverse(const OrcCfg* cfg, const double* lambda, const double* in,
                     double* out, unsigned* iterations)
{
  ORC_TRY
  *iterations = multiply_inverse(*cfg, lambda, in, out);
  ORC_CATCH
}

// ValueTransform::apply, mutators.hh:36-118
// kind: 0 none, 1 logNormal, 2 foldedNormal, 3 sign, 4 boxCox
int
orc_transform(int kind, double boxcox_lambda, size_t n, double* v)
{
  ORC_TRY
  for (size_t i = 0; i < n; i++) {
    double& value = v[i];
    switch (kind) {
      case 0:
        break;
      case 1:
        value = std::exp(value);
        break;
      case 2:
        value = std::abs(value);
        break;
      case 3:
        if (value < 0.)
          value = -1.;
        else
          value = 1.;
        break;
      case 4:
        value = std::pow(value * boxcox_lambda + 1, -boxcox_lambda);
        break;
      default:
        throw std::runtime_error{ "transform type not known" };
    }
  }
  ORC_CATCH
}

// torus covariance (matrix.hh:624-648) on the full torus, natural order; used
// by tests that rebuild the eigenvalues with an independent FFT
int
orc_torus_covariance(const OrcCfg* cfg, double* out)
{
  ORC_TRY
  Sizes s(*cfg);
  Geometry g(*cfg);
  size_t bound[3] = { s.n[0], s.n[1], s.n[2] };
  for (size_t index = 0; index < s.N; index++) {
    size_t idx[3] = { 0, 0, 0 };
    indexToIndices(index, idx, bound, s.dim);
    out[index] = torus_cov(*cfg, s, g, idx);
  }
  ORC_CATCH
}

// StochasticPart::refine (stochastic.hh:597-670): every cell of the old
// level becomes 2^dim cells with the same value.  cells = OLD local cells.
int
orc_field_refine(int dim, const unsigned* cells, const double* in, double* out)
{
  ORC_TRY
  if (dim == 1)
    throw std::runtime_error{ "not implemented" }; // stochastic.hh:664-665
  if (dim != 2 && dim != 3)
    throw std::runtime_error{ "dimension of field has to be 1, 2 or 3" };
  const size_t o0 = cells[0], o1 = cells[1], o2 = dim == 3 ? cells[2] : 1;
  const size_t n0 = 2 * o0, n1 = 2 * o1;
  for (size_t i2 = 0; i2 < o2; i2++)
    for (size_t i1 = 0; i1 < o1; i1++)
      for (size_t i0 = 0; i0 < o0; i0++) {
        const double v = in[i0 + o0 * (i1 + o1 * i2)];
        const size_t base =
          2 * i0 + n0 * (2 * i1 + n1 * (dim == 3 ? 2 * i2 : 0));
        out[base] = v;
        out[base + 1] = v;
        out[base + n0] = v;
        out[base + n0 + 1] = v;
        if (dim == 3) {
          out[base + n1 * n0] = v;
          out[base + n1 * n0 + 1] = v;
          out[base + n1 * n0 + n0] = v;
          out[base + n1 * n0 + n0 + 1] = v;
        }
      }
  ORC_CATCH
}

// StochasticPart::coarsen (stochastic.hh:679-755): mean of the 2^dim fine
// cells, summed in the reference's order.  cells = OLD (fine) local cells;
// odd counts are refused by the traits (fieldtraits.hh:336-339).
int
orc_field_coarsen(int dim, const unsigned* cells, const double* in, double* out)
{
  ORC_TRY
  if (dim == 1)
    throw std::runtime_error{ "not implemented" }; // stochastic.hh:749-750
  if (dim != 2 && dim != 3)
    throw std::runtime_error{ "dimension of field has to be 1, 2 or 3" };
  for (int i = 0; i < dim; i++)
    if (cells[i] % 2 != 0)
      throw std::runtime_error{ "cannot coarsen odd number of cells" };
  const size_t o0 = cells[0], o1 = cells[1];
  const size_t n0 = o0 / 2, n1 = o1 / 2, n2 = dim == 3 ? cells[2] / 2 : 1;
  for (size_t i2 = 0; i2 < n2; i2++)
    for (size_t i1 = 0; i1 < n1; i1++)
      for (size_t i0 = 0; i0 < n0; i0++) {
        const size_t old =
          2 * i0 + o0 * (2 * i1 + o1 * (dim == 3 ? 2 * i2 : 0));
        double v = 0.;
        v += in[old];
        v += in[old + 1];
        v += in[old + o0];
        v += in[old + o0 + 1];
        if (dim == 3) {
          v += in[old + o1 * o0];
          v += in[old + o1 * o0 + 1];
          v += in[old + o1 * o0 + o0];
          v += in[old + o1 * o0 + o0 + 1];
        }
        out[i0 + n0 * (i1 + n1 * i2)] = v / (dim == 3 ? 8 : 4);
      }
  ORC_CATCH
}

// StochasticPart::localize (stochastic.hh:798-817) with indicesToCoords
// (fieldtraits.hh:461-469): data *= (2*3.14159)^(-dim/2) exp(-|x-c|^2/(2 r^2)).
// cells / offset describe the local block, gcells the whole grid.
int
orc_field_localize(int dim, const unsigned* cells, const unsigned* offset,
                   const unsigned* gcells, const double* extensions,
                   const double* center, double radius, double* data)
{
  ORC_TRY
  size_t bound[3] = { 1, 1, 1 };
  size_t n = 1;
  for (int i = 0; i < dim; i++) {
    bound[i] = cells[i];
    n *= cells[i];
  }
  const double factor = std::pow(2. * 3.14159, -(dim / 2.));
  for (size_t index = 0; index < n; index++) {
    size_t idx[3] = { 0, 0, 0 };
    indexToIndices(index, idx, bound, dim);
    double distSquared = 0.;
    for (int j = 0; j < dim; j++) {
      const unsigned g = (unsigned)idx[j] + offset[j];
      const double loc = (g + 0.5) * extensions[j] / gcells[j];
      distSquared += (loc - center[j]) * (loc - center[j]);
    }
    data[index] *= factor * std::exp(-0.5 * distSquared / (radius * radius));
  }
  ORC_CATCH
}

// StochasticPart::evaluate (stochastic.hh:480-574) on one rank, where
// dataToEval is a copy and the 2*dim overlap arrays stay zero-filled
// (stochastic.hh:838-852), followed by RandomField::evaluate's value
// transform (randomfield.hh:585-595; no trend part).  coordsToIndices:
// fieldtraits.hh:440-448.  A location on the upper boundary of exactly one
// dimension reads the (zero) overlap; on an edge or corner the reference
// steps back into the last cell.  Locations outside the closed box are
// undefined behaviour in the reference and refused here (return 1).
int
orc_field_evaluate(int dim, const unsigned* cells, const double* extensions,
                   const double* field, size_t npoints, const double* loc,
                   int transform, double boxcox_lambda, double* out)
{
  ORC_TRY
  for (size_t p = 0; p < npoints; p++) {
    unsigned idx[3] = { 0, 0, 0 };
    int outside = 0;
    for (int i = 0; i < dim; i++) {
      const double x = loc[p * dim + i];
      if (!(x >= 0.))
        throw std::runtime_error{ "location outside the domain" };
      idx[i] = (unsigned)(x * (cells[i] + 1e-6) / extensions[i]);
      if (idx[i] > cells[i])
        throw std::runtime_error{ "location outside the domain" };
      if (idx[i] == cells[i])
        outside++;
    }
    double v;
    if (outside == 1 && dim <= 3)
      v = 0.; // overlap[2i+1], never filled on one rank
    else {
      for (int i = 0; i < dim; i++)
        if (idx[i] == cells[i])
          idx[i]--;
      size_t index = 0, stride = 1;
      for (int i = 0; i < dim; i++) {
        index += stride * idx[i];
        stride *= cells[i];
      }
      v = field[index];
    }
    out[p] = v;
  }
  if (transform != 0) {
    int rc = orc_transform(transform, boxcox_lambda, npoints, out);
    if (rc != 0)
      return rc;
  }
  ORC_CATCH
}

// TrendComponent::evaluate (trend.hh:601-634) summed by TrendPart::evaluate
// (trend.hh:1491-1505).  A component is (kind, coeff[7]) with coeff =
// meanVector + shiftVector; kinds 0 mean, 1 slope, 2 disk, 3 block.  The image
// component (libpng) is out of scope.
struct OrcTrend
{
  int kind;
  double coeff[7];
};

static double
trend_value(int dim, const double* extensions, int ntrend,
            const OrcTrend* trend, const double* x)
{
  double output = 0.;
  for (int t = 0; t < ntrend; t++) {
    const double* c = trend[t].coeff;
    double comp = 0.;
    switch (trend[t].kind) {
      case 0:
        comp = c[0];
        break;
      case 1:
        for (int i = 0; i < dim; i++)
          comp += c[i] * (x[i] - extensions[i] / 2.);
        break;
      case 2: {
        double distSquared = 0.;
        for (int i = 0; i < dim; i++)
          distSquared += std::pow(x[i] - c[i], 2);
        comp = std::exp(-distSquared / std::pow(c[dim], 2) * c[dim + 1]);
        break;
      }
      case 3: {
        bool inside = true;
        for (int i = 0; i < dim; i++)
          if (std::abs(x[i] - c[i]) > 0.5 * c[dim + i])
            inside = false;
        comp = inside ? c[2 * dim] : 0.;
        break;
      }
      default:
        throw std::runtime_error{ "Trend component type not found!" };
    }
    output += comp;
  }
  return output;
}

// RandomField::evaluate (randomfield.hh:585-595): T(stochastic + trend)
int
orc_field_evaluate_trend(int dim, const unsigned* cells,
                         const double* extensions, const double* field,
                         size_t npoints, const double* loc, int ntrend,
                         const OrcTrend* trend, int transform,
                         double boxcox_lambda, double* out)
{
  int rc = orc_field_evaluate(dim, cells, extensions, field, npoints, loc, 0,
                              0., out);
  if (rc != 0)
    return rc;
  ORC_TRY
  for (size_t p = 0; p < npoints; p++)
    out[p] = out[p] + trend_value(dim, extensions, ntrend, trend, loc + p * dim);
  if (transform != 0)
    return orc_transform(transform, boxcox_lambda, npoints, out);
  ORC_CATCH
}

// RandomField::bulkEvaluate (randomfield.hh:597-631): evaluate at the centre
// of every cell, locations built as (i + 0.5) * (extensions / cells)
int
orc_field_bulk_evaluate(int dim, const unsigned* cells,
                        const double* extensions, const double* field,
                        int ntrend, const OrcTrend* trend, int transform,
                        double boxcox_lambda, double* out)
{
  ORC_TRY
  size_t bound[3] = { 1, 1, 1 };
  size_t n = 1;
  for (int i = 0; i < dim; i++) {
    bound[i] = cells[i];
    n *= cells[i];
  }
  std::vector<double> loc(n * dim);
  for (size_t index = 0; index < n; index++) {
    size_t idx[3] = { 0, 0, 0 };
    indexToIndices(index, idx, bound, dim);
    for (int j = 0; j < dim; j++) {
      const double meshsize = extensions[j] / static_cast<double>(cells[j]);
      loc[index * dim + j] = (static_cast<double>(idx[j]) + 0.5) * meshsize;
    }
  }
  return orc_field_evaluate_trend(dim, cells, extensions, field, n, loc.data(),
                                  ntrend, trend, transform, boxcox_lambda, out);
  ORC_CATCH
}

} // extern "C"

// --------------------------------------------------------------------------
// Device white noise of the CUDA product path, restated on the CPU so that the
// benchmarked variant (counter -> Philox -> Box-Muller -> sigma rule -> c2r)
// has an oracle of its own.  The reference has no counterpart (its
// RNGBackend::sample, backends/cpprngbackend.hh:49-58, is a sequential
// stream); what is restated here is the SPECIFICATION the library documents
// (include/parafields_cuda.h, "device noise"), written from the published
// algorithm -- Salmon, Moraes, Dror, Shaw: "Parallel random numbers: as easy
// as 1, 2, 3", SC'11, Philox4x32-R with the Random123 multipliers and Weyl
// key increments -- and not from the CUDA source:
//   counter = (c lo, c hi, realization lo, realization hi), key = (seed lo,
//   seed hi); c = k0 + m0 * (e mod H) [+ m0 * H * k1 in 3D], H = ceil(L/2), where e is
//   the index along the LAST frequency axis (length L) and one Philox block of
//   four words serves the two entries e and e + L/2: words (0,1) -> entry
//   e < L/2, words (2,3) -> entry e + L/2.  Box-Muller on a word pair (w0,w1):
//   u = (w0 + 1/2) 2^-32, radius sqrt(-2 ln u), angle pi (w1 2^-31 - 1);
//   r1 = radius cos, r2 = radius sin.  mode 1: fp64, 10 rounds; mode 0
//   ("fast"): 7 rounds, the same formula in fp32 (the device uses the MUFU
//   approximations of lg2 / sqrt / sin / cos there, so agreement is ~1e-6).
// The normals are written in the reference's stream order, so
// that generate_from_noise() above finishes the job.
// --------------------------------------------------------------------------
static void
philox4x32(int rounds, const uint32_t ctr[4], const uint32_t key[2],
           uint32_t out[4])
{
  const uint64_t M0 = 0xD2511F53ull, M1 = 0xCD9E8D57ull;
  const uint32_t W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
  uint32_t c[4] = { ctr[0], ctr[1], ctr[2], ctr[3] };
  uint32_t k[2] = { key[0], key[1] };
  for (int r = 0; r < rounds; r++) {
    if (r > 0) { // bump the key between rounds
      k[0] += W0;
      k[1] += W1;
    }
    const uint64_t p0 = M0 * c[0], p1 = M1 * c[2];
    const uint32_t n[4] = { (uint32_t)(p1 >> 32) ^ c[1] ^ k[0], (uint32_t)p1,
                            (uint32_t)(p0 >> 32) ^ c[3] ^ k[1], (uint32_t)p0 };
    for (int i = 0; i < 4; i++)
      c[i] = n[i];
  }
  for (int i = 0; i < 4; i++)
    out[i] = c[i];
}

static void
box_muller(int mode, uint32_t w0, uint32_t w1, double* r1, double* r2)
{
  if (mode == 1) {
    const double u = ((double)w0 + 0.5) * 2.3283064365386963e-10; // 2^-32
    const double radius = std::sqrt(-2.0 * std::log(u));
    const double turn = (double)w1 * 4.6566128730773926e-10 - 1.0; // angle/pi
    const double ang = 3.14159265358979323846264338327950288 * turn;
    *r1 = radius * std::cos(ang);
    *r2 = radius * std::sin(ang);
  } else {
    const float u = std::fmaf((float)w0, 2.3283064365386963e-10f,
                              1.1641532182693481e-10f);
    const float radius = std::sqrt(-1.3862943611198906f * std::log2(u));
    const float ang =
      std::fmaf((float)w1, 1.4629180792671596e-09f, -3.14159265358979f);
    *r1 = (double)(radius * std::cos(ang));
    *r2 = (double)(radius * std::sin(ang));
  }
}

static void
philox_noise(const OrcCfg& cfg, uint64_t seed, uint64_t realization, int mode,
             double* noise)
{
  Sizes s(cfg);
  const int d = s.dim;
  const int rounds = mode == 1 ? 10 : 7;
  const uint32_t key[2] = { (uint32_t)seed, (uint32_t)(seed >> 32) };
  const bool dft = cfg.field_backend == 1;
  const size_t b0 = dft ? s.n[0] : s.m0;
  const size_t L = s.n[d - 1], half = (L + 1) / 2; // odd L (1D only): ceil
  const bool tr = cfg.transposed && d == 3;
  const size_t S = b0 * (d > 1 ? s.n[1] : 1) * (d > 2 ? s.n[2] : 1);
  for (size_t idx = 0; idx < S; idx++) {
    size_t k[3] = { 0, 0, 0 };
    if (tr) {
      size_t b[3] = { b0, s.n[2], s.n[1] }, t[3];
      indexToIndices(idx, t, b, 3);
      k[0] = t[0];
      k[1] = t[2];
      k[2] = t[1];
    } else {
      size_t b[3] = { b0, s.n[1], s.n[2] };
      indexToIndices(idx, k, b, d);
    }
    const size_t e = k[d - 1];
    uint64_t c;
    if (d == 1)
      c = e % half;
    else if (d == 2)
      c = k[0] + (uint64_t)b0 * (e % half);
    else
      c = k[0] + (uint64_t)b0 * (e % half) + (uint64_t)b0 * half * k[1];
    const uint32_t ctr[4] = { (uint32_t)c, (uint32_t)(c >> 32),
                              (uint32_t)realization,
                              (uint32_t)(realization >> 32) };
    uint32_t w[4];
    philox4x32(rounds, ctr, key, w);
    const int o = e >= half ? 2 : 0;
    box_muller(mode, w[o], w[o + 1], &noise[2 * idx], &noise[2 * idx + 1]);
  }
}

extern "C" {

// one Philox4x32-R block (known-answer tests against the Random123 vectors)
int
orc_philox4x32(int rounds, const uint32_t* ctr, const uint32_t* key,
               uint32_t* out)
{
  ORC_TRY
  philox4x32(rounds, ctr, key, out);
  ORC_CATCH
}

// the product path's device noise for (seed, realization) in the reference's
// stream order: 2 * orc_noise_entries(cfg) doubles; mode 0 fast, 1 fp64
int
orc_philox_noise(const OrcCfg* cfg, uint64_t seed, uint64_t realization,
                 int mode, double* noise)
{
  ORC_TRY
  philox_noise(*cfg, seed, realization, mode, noise);
  ORC_CATCH
}

// pf_generate_uncorrelated's specification: block q (counter q0 + q,
// realization word = 2^64 - 1, 10 rounds, fp64 Box-Muller) fills cells
// 4q .. 4q+3 with (r1, r2) of words (0,1) and (r1, r2) of words (2,3)
int
orc_philox_uncorrelated(uint64_t seed, size_t n, uint64_t q0, double* out)
{
  ORC_TRY
  const uint32_t key[2] = { (uint32_t)seed, (uint32_t)(seed >> 32) };
  for (size_t q = 0; 4 * q < n; q++) {
    const uint64_t c = q0 + q;
    const uint32_t ctr[4] = { (uint32_t)c, (uint32_t)(c >> 32), 0xFFFFFFFFu,
                              0xFFFFFFFFu };
    uint32_t w[4];
    philox4x32(10, ctr, key, w);
    double v[4];
    box_muller(1, w[0], w[1], &v[0], &v[1]);
    box_muller(1, w[2], w[3], &v[2], &v[3]);
    for (size_t j = 0; j < 4 && 4 * q + j < n; j++)
      out[4 * q + j] = v[j];
  }
  ORC_CATCH
}

} // extern "C"

extern "C" {

int
orc_max_threads()
{
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

} // extern "C"
