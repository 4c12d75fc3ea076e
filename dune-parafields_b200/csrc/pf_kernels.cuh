// pf_kernels.cuh -- the pass kernels of the circulant-embedding hot path.
//
// Every kernel processes "lines" of N complex doubles with the warp-level
// FFT of pf_fft.cuh and differs only in how a line is produced (loader) and
// consumed (storer).  Layout cycle (3D):
//
//   generate : sigma[k1][k0][k2] --A--> Y[k0][x2][k1] --B--> Z[x2][x1][k0]
//              --C--> field[x2][x1][x0]
//   forward  : field[x2][x1][x0] --F0--> U[x2][k0][x1] --F1--> V[k1][k0][x2]
//              --F2 (x g(lambda)) + A--> Y[k0][x2][k1] --B--> --C-->
//
// Reads are always whole contiguous lines (one warp reads 512 B per
// instruction); writes towards the next pass's contiguous axis go through a
// shared-memory tile so that every store instruction covers LPC*16
// contiguous bytes ("tile store").
#pragma once
#include "pf_args.h"
#include "pf_fft.cuh"
#include "pf_rng.cuh"

namespace pf {

constexpr int PF_THREADS = 128;

template <int N>
struct KCfg
{
  using C = LineCfg<N>;
  static constexpr int E = C::E;
  static constexpr int TPL = C::TPL;
  static constexpr int LPC = PF_THREADS / TPL; // lines (FFTs) per CTA
  static constexpr int SP = LPC + 1;           // stage pitch (odd)
  static constexpr int ZP = N + 1;             // r2c stage pitch (odd)
  static constexpr size_t exch_bytes = sizeof(cplx) * LPC * C::PITCH;
  static __host__ __device__ constexpr size_t smem_tile(int keep)
  {
    size_t st = sizeof(cplx) * (size_t)keep * SP;
    return st > exch_bytes ? st : exch_bytes;
  }
  static __host__ __device__ constexpr size_t smem_r2c()
  {
    size_t st = sizeof(cplx) * (size_t)LPC * ZP;
    return st > exch_bytes ? st : exch_bytes;
  }
};

// ---------------------------------------------------------------------------
// value transforms, mutators.hh:36-118
// ---------------------------------------------------------------------------
__device__ __forceinline__ double
apply_transform(double v, int kind, double lambda)
{
  switch (kind) {
    case 1:
      return exp(v);
    case 2:
      return fabs(v);
    case 3:
      return v < 0. ? -1. : 1.;
    case 4:
      return pow(v * lambda + 1, -lambda);
    default:
      return v;
  }
}

// compile-time variant + dispatcher: keeps the switch out of unrolled store
// loops (one specialised loop per transform kind)
template <int KIND>
__device__ __forceinline__ double
apply_transform_c(double v, double lambda)
{
  if constexpr (KIND == 1)
    return exp(v);
  else if constexpr (KIND == 2)
    return fabs(v);
  else if constexpr (KIND == 3)
    return v < 0. ? -1. : 1.;
  else if constexpr (KIND == 4)
    return pow(v * lambda + 1, -lambda);
  else
    return v;
}
template <class F>
__device__ __forceinline__ void
dispatch_transform(int kind, F&& f)
{
  switch (kind) {
    case 1:
      f(std::integral_constant<int, 1>{});
      break;
    case 2:
      f(std::integral_constant<int, 2>{});
      break;
    case 3:
      f(std::integral_constant<int, 3>{});
      break;
    case 4:
      f(std::integral_constant<int, 4>{});
      break;
    default:
      f(std::integral_constant<int, 0>{});
      break;
  }
}

// ---------------------------------------------------------------------------
// modified Bessel function of the second kind K_nu(x), x > 0, real order
// nu = mu + nl with |mu| <= 1/2 -- what covariance.hh:565-573 takes from
// gsl_sf_bessel_Knu (GSL, absent here).  Published algorithm: Temme's series
// (J. Comput. Phys. 19 (1975) 324) for x < 2, Steed's continued fraction CF2
// in the Thompson-Barnett form for x >= 2, then the upward recurrence
// K_{v+1} = K_{v-1} + (2 v / x) K_v (stable for K).  The Gamma-function
// constants of mu come from the host (CovParams::mt).
// ---------------------------------------------------------------------------
static __device__ __noinline__ double
bessel_k_nu(const double* mt, double x)
{
  const double mu = mt[2], mu2 = mu * mu;
  const int nl = (int)mt[3];
  const double EPS = 1e-16;
  double kmu, kmu1;
  if (x < 2.) {
    const double b = 0.5 * x, lg = -log(b), sg = mu * lg;
    const double pimu = 3.14159265358979323846 * mu;
    const double fsin = fabs(pimu) < 1e-8 ? 1. : pimu / sin(pimu);
    const double fsinh = fabs(sg) < 1e-8 ? 1. : sinh(sg) / sg;
    double f = fsin * (mt[6] * cosh(sg) + mt[7] * fsinh * lg);
    const double ex = exp(sg);
    double p = 0.5 * ex / mt[4], q = 0.5 / (ex * mt[5]);
    double c = 1., sum = f, sum1 = p;
    const double b2 = b * b;
    for (int i = 1; i < 500; i++) {
      f = (i * f + p + q) / (i * (double)i - mu2);
      c *= b2 / i;
      p /= i - mu;
      q /= i + mu;
      const double del = c * f;
      sum += del;
      sum1 += c * (p - i * f);
      if (fabs(del) < fabs(sum) * EPS)
        break;
    }
    kmu = sum;
    kmu1 = sum1 * (2. / x);
  } else {
    double b = 2. * (1. + x), d = 1. / b, h = d, delh = d;
    double q1 = 0., q2 = 1.;
    const double a1 = 0.25 - mu2;
    double q = a1, c = a1, a = -a1;
    double s = 1. + q * delh;
    for (int i = 2; i < 500; i++) {
      a -= 2 * (i - 1);
      c = -a * c / i;
      const double qn = (q1 - b * q2) / a;
      q1 = q2;
      q2 = qn;
      q += c * qn;
      b += 2.;
      d = 1. / (b + a * d);
      delh = (b * d - 1.) * delh;
      h += delh;
      const double dels = q * delh;
      s += dels;
      if (fabs(dels) < fabs(s) * EPS)
        break;
    }
    h = a1 * h;
    kmu = sqrt(3.14159265358979323846 / (2. * x)) * exp(-x) / s;
    kmu1 = kmu * (mu + x + 0.5 - h) / x;
  }
  const double xi2 = 2. / x;
  for (int i = 1; i <= nl; i++) {
    const double kn = (mu + i) * xi2 * kmu1 + kmu;
    kmu = kmu1;
    kmu1 = kn;
  }
  return kmu;
}

// ---------------------------------------------------------------------------
// covariance on the torus, matrix.hh:624-648 + covariance.hh (same arithmetic
// as the reference functors)
// ---------------------------------------------------------------------------
__device__ __forceinline__ double
cov_function(const CovParams& c, const double* x)
{
  const int dim = c.dim;
  if (c.cov == COV_SEPARABLE_EXP) {
    double sum = 0.;
    for (int i = 0; i < dim; i++)
      sum += fabs(x[i]);
    return c.variance * exp(-sum);
  }
  if (c.cov == COV_WHITE_NOISE) {
    for (int i = 0; i < dim; i++)
      if (fabs(x[i]) > 1e-10)
        return 0.;
    return c.variance;
  }
  double sum = 0.;
  for (int i = 0; i < dim; i++)
    sum += x[i] * x[i];
  const double h = sqrt(sum);
  switch (c.cov) {
    case COV_EXPONENTIAL:
      return c.variance * exp(-h);
    case COV_GAUSSIAN:
      return c.variance * exp(-h * h);
    case COV_SPHERICAL:
      if (h > 1.)
        return 0.;
      if (dim == 3)
        return c.variance * (1. - 1.5 * h + 0.5 * (h * h * h));
      if (dim == 2)
        return c.variance *
               (1. - 2. / 3.14159265358979323846 *
                       (h * sqrt(1. - h * h) + asin(h)));
      return c.variance * (1. - h);
    case COV_MATERN32:
      return c.variance * (1. + sqrt(3.) * h) * exp(-sqrt(3.) * h);
    case COV_MATERN52:
      return c.variance * (1. + sqrt(5.) * h + 5. / 3. * h * h) *
             exp(-sqrt(5.) * h);
    case COV_DAMPED_OSC:
      if (dim == 3)
        return c.variance * exp(-h) * cos(h / sqrt(3.));
      return c.variance * exp(-h) * cos(h);
    case COV_GAMMA_EXP:
      return c.variance * exp(-pow(h, c.p0));
    case COV_CAUCHY: {
      const double b = 1. + h * h;
      return c.variance / (b * b * b);
    }
    case COV_GEN_CAUCHY:
      return c.variance * pow(1. + pow(h, c.p0), -c.p1);
    case COV_CUBIC: {
      if (h > 1.)
        return 0.;
      const double h2 = h * h, h3 = h2 * h, h5 = h3 * h2, h7 = h5 * h2;
      return c.variance *
             (1. - 7. * h2 + 8.75 * h3 - 3.5 * h5 + 0.75 * h7);
    }
    case COV_MATERN: { // covariance.hh:565-573
      if (h < 1e-10)
        return c.variance;
      const double z = c.mt[0] * h;
      return c.variance * c.mt[1] * pow(z, c.p0) * bessel_k_nu(c.mt, z);
    }
  }
  return 0.;
}

// geometry transform h' = A h (covariance.hh:17-136), A already inverted
__device__ __forceinline__ void
geo_transform(const CovParams& c, const double* coord, double* tc)
{
  if (c.aniso == 0) {
    for (int i = 0; i < c.dim; i++)
      tc[i] = c.g[0] * coord[i];
  } else if (c.aniso == 1) {
    for (int i = 0; i < c.dim; i++)
      tc[i] = c.g[i] * coord[i];
  } else {
    for (int i = 0; i < c.dim; i++) {
      tc[i] = 0.;
      for (int j = 0; j < c.dim; j++)
        tc[i] += c.g[i * c.dim + j] * coord[j];
    }
  }
}

// SmoothSigmoid (covariance.hh:145-176) / SmoothstepSigmoid on
// ClampingSigmoid (covariance.hh:202-275)
static __device__ __noinline__ double
sigmoid_value(const CovParams& c, double oneEdge, double zeroEdge, double x)
{
  const double ax = fabs(x);
  if (c.sigmoid == 1) {
    if (ax < oneEdge)
      return 1.;
    if (ax > zeroEdge)
      return 0.;
    const double x1 = (zeroEdge - ax) / (zeroEdge - oneEdge);
    const double x2 = (ax - oneEdge) / (zeroEdge - oneEdge);
    const double t1 = x1 > 0. ? exp(-1. / x1) : 0.;
    const double t2 = x2 > 0. ? exp(-1. / x2) : 0.;
    return t1 / (t1 + t2);
  }
  double x1;
  if (ax < oneEdge)
    x1 = 1.;
  else if (ax > zeroEdge)
    x1 = 0.;
  else
    x1 = (zeroEdge - ax) / (zeroEdge - oneEdge);
  const double x2 = 2. * x1 - 1.;
  double normalization = 1., addition = 1., output = 1.;
  for (unsigned i = 1; i <= c.recursions; i++) {
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

// Matrix::computeMatrixEntriesWithMerge (matrix.hh:662-787): 2^dim mirrored
// copies of the covariance, each tapered by a sigmoid; what the tapers leave
// is filled with the covariance of the far corner (radialPlus)
static __device__ __noinline__ double
cov_torus_merge(const CovParams& c, unsigned i0, unsigned i1, unsigned i2)
{
  const unsigned idx[3] = { i0, i1, i2 };
  double trueCoord[3], mirrorCoord[3], tc[3];
  double constValue = 0.;
  if (c.combine == 0) {
    for (int i = 0; i < c.dim; i++)
      trueCoord[i] = c.ext[i] * c.factor / 2.;
    geo_transform(c, trueCoord, tc);
    constValue = cov_function(c, tc);
  }
  for (int i = 0; i < c.dim; i++)
    trueCoord[i] = idx[i] * c.h[i];
  const double eps = 1e-10;
  const double sqrtDim = sqrt((double)c.dim);
  const bool radial = c.combine != 2 && sqrtDim < c.factor / 2.;
  double value = 0., remainder = 1.;
  for (unsigned j = 0; j < (1u << c.dim); j++) {
    for (int i = 0; i < c.dim; i++)
      mirrorCoord[i] =
        (j & (1u << i)) ? c.ext[i] * c.factor - trueCoord[i] : trueCoord[i];
    double dampening = 1.;
    if (radial) {
      double radius = 0.;
      for (int i = 0; i < c.dim; i++)
        radius += (mirrorCoord[i] * mirrorCoord[i]) / (c.ext[i] * c.ext[i]);
      radius = sqrt(radius);
      dampening = sigmoid_value(c, c.s_start * sqrtDim - eps,
                                c.factor - (c.factor - c.s_end) * sqrtDim + eps,
                                radius);
    } else {
      for (int i = 0; i < c.dim; i++)
        dampening *= sigmoid_value(c, c.ext[i] * c.s_start - eps,
                                   c.ext[i] * c.s_end + eps, mirrorCoord[i]);
    }
    if (dampening < eps)
      continue;
    remainder -= dampening;
    geo_transform(c, mirrorCoord, tc);
    value += cov_function(c, tc) * dampening;
  }
  return value + constValue * remainder;
}

// torus cell (i0,i1,i2) -> covariance value
__device__ __forceinline__ double
cov_torus(const CovParams& c, unsigned i0, unsigned i1, unsigned i2)
{
  if (c.periodization == 1)
    return cov_torus_merge(c, i0, i1, i2);
  double coord[3], tc[3];
  const unsigned idx[3] = { i0, i1, i2 };
  for (int i = 0; i < c.dim; i++) {
    coord[i] = idx[i] * c.h[i];
    if (coord[i] > c.half[i])
      coord[i] -= c.period[i];
  }
  if (c.aniso == 0) {
    for (int i = 0; i < c.dim; i++)
      tc[i] = c.g[0] * coord[i];
  } else if (c.aniso == 1) {
    for (int i = 0; i < c.dim; i++)
      tc[i] = c.g[i] * coord[i];
  } else {
    for (int i = 0; i < c.dim; i++) {
      tc[i] = 0.;
      for (int j = 0; j < c.dim; j++)
        tc[i] += c.g[i * c.dim + j] * coord[j];
    }
  }
  return cov_function(c, tc);
}

// ---------------------------------------------------------------------------
// loaders
// ---------------------------------------------------------------------------
__device__ __forceinline__ const double*
table_row(const PassArgs& p, int a, int b)
{
  const int am = (p.tab_mirror_a && a > p.tab_mirror_a / 2) ? p.tab_mirror_a - a
                                                            : a;
  return p.tab + ((long long)am * p.rma + (long long)b * p.rmb) * p.tab_pitch;
}

// white noise scaled by sigma.  Follows Matrix::generateField's loop
// (matrix.hh:396-421) and FieldBackend::set (r2cfieldbackend.hh:305-325,
// dftfieldbackend.hh:276-280); the table holds sqrt(lambda)/sqrt(2)
// (hermitian) or sqrt(lambda) (DFT backend rule).  The sigma row is fetched
// up front (registers are free before the line is populated) and the rare
// cases (self-conjugate entries, `corrected` planes) sit outside the
// unrolled loop.
template <int N, int NOISE>
__device__ __forceinline__ void
load_noise(cplx (&v)[LineCfg<N>::E], const PassArgs& p, int a, int b, int t,
             bool valid)
{
  using C = LineCfg<N>;
  constexpr int H = C::E / 2;
  constexpr double SQRT2 = 1.41421356237309504880168872420969808;
  if (!valid) {
#pragma unroll
    for (int q = 0; q < C::E; q++)
      v[q] = make_double2(0., 0.);
    return;
  }
  const double* row = table_row(p, a, b);
  // second half of the line: e + N/2, or its mirror N/2 - e (compact table)
  const double* row2 = row + N / 2;
  const int dir2 = p.tab_mirror_e ? -1 : 1;
  const unsigned long long cbase =
    (unsigned long long)a * p.cma + (unsigned long long)b * p.cmb;
  const unsigned long long rea = p.realization + (unsigned long long)b * p.real_b;
  const long long ibase = (long long)a * p.ima + (long long)b * p.imb;
  double s1[H], s2[H];
#pragma unroll
  for (int q = 0; q < H; q++) {
    const int e = t + C::TPL * q; // < N/2
    s1[q] = __ldg(row + e);
    s2[q] = __ldg(row2 + dir2 * e);
  }
#pragma unroll
  for (int q = 0; q < H; q++) {
    const int e = t + C::TPL * q;
    double2 n1, n2;
    if constexpr (NOISE == NOISE_INJECTED) {
      n1 = __ldg(p.noise + ibase + (long long)e * p.ime);
      n2 = __ldg(p.noise + ibase + (long long)(e + N / 2) * p.ime);
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
    v[q] = make_double2(s1[q] * n1.x, s1[q] * n1.y);
    v[q + H] = make_double2(s2[q] * n2.x, s2[q] * n2.y);
  }
  if (p.hermitian) {
    const bool a_self = (2 * a) % p.na_self == 0;
    const bool b_self = (2 * b) % p.nb_self == 0;
    const bool plane = p.corrected && (p.plane_axis == 0 ? a_self : b_self);
    if (plane) {
#pragma unroll
      for (int q = 0; q < C::E; q++)
        v[q] = make_double2(v[q].x * SQRT2, v[q].y * SQRT2);
    }
    if (t == 0 && a_self && b_self) {
      // self-conjugate entries e = 0 and e = N/2: real, full weight
      const double f = plane ? 1.0 : SQRT2;
      v[0] = make_double2(v[0].x * f, 0.);
      v[H] = make_double2(v[H].x * f, 0.);
    }
  }
}

// load_noise split in two for kernels that receive the sigma row through
// shared memory (kp_gen3): unit normals first, the scaling afterwards.  Same
// products as load_noise, so the results are bit-identical.
template <int N, int NOISE>
__device__ __forceinline__ void
gen_noise(cplx (&v)[LineCfg<N>::E], const PassArgs& p, int a, int b, int t)
{
  using C = LineCfg<N>;
  constexpr int H = C::E / 2;
  const unsigned long long cbase =
    (unsigned long long)a * p.cma + (unsigned long long)b * p.cmb;
  const unsigned long long rea = p.realization + (unsigned long long)b * p.real_b;
  const long long ibase = (long long)a * p.ima + (long long)b * p.imb;
#pragma unroll
  for (int q = 0; q < H; q++) {
    const int e = t + C::TPL * q;
    double2 n1, n2;
    if constexpr (NOISE == NOISE_INJECTED) {
      n1 = __ldg(p.noise + ibase + (long long)e * p.ime);
      n2 = __ldg(p.noise + ibase + (long long)(e + N / 2) * p.ime);
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
    v[q] = n1;
    v[q + H] = n2;
  }
}

// srow: the line's sigma row (shared memory), indexed like table_row()
template <int N>
__device__ __forceinline__ void
scale_noise(cplx (&v)[LineCfg<N>::E], const PassArgs& p, const double* srow,
            int a, int b, int t)
{
  using C = LineCfg<N>;
  constexpr int H = C::E / 2;
  constexpr double SQRT2 = 1.41421356237309504880168872420969808;
  const int dir2 = p.tab_mirror_e ? -1 : 1;
  const double* srow2 = srow + N / 2;
#pragma unroll
  for (int q = 0; q < H; q++) {
    const int e = t + C::TPL * q;
    const double s1 = srow[e], s2 = srow2[dir2 * e];
    v[q] = make_double2(s1 * v[q].x, s1 * v[q].y);
    v[q + H] = make_double2(s2 * v[q + H].x, s2 * v[q + H].y);
  }
  if (p.hermitian) {
    const bool a_self = (2 * a) % p.na_self == 0;
    const bool b_self = (2 * b) % p.nb_self == 0;
    const bool plane = p.corrected && (p.plane_axis == 0 ? a_self : b_self);
    if (plane) {
#pragma unroll
      for (int q = 0; q < C::E; q++)
        v[q] = make_double2(v[q].x * SQRT2, v[q].y * SQRT2);
    }
    if (t == 0 && a_self && b_self) {
      const double f = plane ? 1.0 : SQRT2;
      v[0] = make_double2(v[0].x * f, 0.);
      v[H] = make_double2(v[H].x * f, 0.);
    }
  }
}

template <int N>
__device__ __forceinline__ void
load_contig(cplx (&v)[LineCfg<N>::E], const cplx* __restrict__ line, int t,
            int n_valid, bool valid)
{
  using C = LineCfg<N>;
#pragma unroll
  for (int q = 0; q < C::E; q++) {
    const int e = t + C::TPL * q;
    v[q] = (valid && e < n_valid) ? __ldg(line + e) : make_double2(0., 0.);
  }
}

// ---------------------------------------------------------------------------
// storers
// ---------------------------------------------------------------------------
// tile store: out[base + e*ose + l], l = line in tile (contiguous), e < keep.
// The stage holds rows of LPC complex numbers (one per line of the tile);
// the 16-byte chunks of a row pair are XOR-swizzled so that both the
// per-line writes (stride one row) and the row-major reads are free of bank
// conflicts for LPC = 4 (N = 1024) and LPC = 8 (N = 512).
template <int N>
__device__ __forceinline__ int
stage_slot(int e, int l)
{
  using K = KCfg<N>;
  if constexpr (K::LPC == 4)
    return e * 4 + (l ^ ((e >> 1) & 3));
  else if constexpr (K::LPC == 8)
    return e * 8 + (l ^ (e & 7));
  else
    return e * K::SP + l;
}

// row pitch of the stage in complex numbers and the slot of (e, l); for the
// swizzled layouts the swizzle term depends on e only through bits that a
// step of PF_THREADS / LPC rows leaves unchanged, so a thread's slots (and
// output addresses) advance by a constant per step
template <int N>
__device__ __forceinline__ constexpr int
stage_pitch()
{
  using K = KCfg<N>;
  return (K::LPC == 4 || K::LPC == 8) ? K::LPC : K::SP;
}

// HALF: keep == N / 2 and scale == 1 known at compile time -- the upper half of
// the transform's outputs is dead code (its last butterflies go with it), no
// bound checks, one stage address per thread (the swizzle term of stage_slot
// does not change over a step of TPL rows)
template <int N, bool HALF = false>
__device__ __forceinline__ void
tile_store(cplx (&v)[LineCfg<N>::E], cplx* stage, int lic, int t, int keep,
           cplx* __restrict__ obase, long long ose, int lines_valid,
           double scale)
{
  using K = KCfg<N>;
  __syncthreads(); // exchange area is about to be reused as the stage
  if constexpr (HALF) {
    keep = N / 2;
    cplx* sp0 = stage + stage_slot<N>(t, lic);
#pragma unroll
    for (int q = 0; q < K::E / 2; q++)
      sp0[q * (K::TPL * stage_pitch<N>())] = v[q];
  } else if (scale == 1.0) {
#pragma unroll
    for (int q = 0; q < K::E; q++) {
      const int e = t + K::TPL * q;
      if (e < keep)
        stage[stage_slot<N>(e, lic)] = v[q];
    }
  } else {
#pragma unroll
    for (int q = 0; q < K::E; q++) {
      const int e = t + K::TPL * q;
      if (e < keep)
        stage[stage_slot<N>(e, lic)] =
          make_double2(v[q].x * scale, v[q].y * scale);
    }
  }
  __syncthreads();
  // read-out: thread -> (row e0 + i*STEP, line l); loads batched four deep so
  // that the shared-memory latency is paid once per batch
  constexpr int STEP = PF_THREADS / K::LPC;
  constexpr int SSTEP = STEP * stage_pitch<N>();
  static_assert(PF_THREADS % K::LPC == 0, "tile read-out mapping");
  const int l = threadIdx.x % K::LPC, e0 = threadIdx.x / K::LPC;
  if (l >= lines_valid)
    return;
  const cplx* sp = stage + stage_slot<N>(e0, l);
  cplx* op = obase + (long long)e0 * ose + l;
  const long long ostep = (long long)STEP * ose;
  int e = e0;
  for (; e + 3 * STEP < keep; e += 4 * STEP) {
    const cplx a0 = sp[0], a1 = sp[SSTEP], a2 = sp[2 * SSTEP],
               a3 = sp[3 * SSTEP];
    op[0] = a0;
    op[ostep] = a1;
    op[2 * ostep] = a2;
    op[3 * ostep] = a3;
    sp += 4 * SSTEP;
    op += 4 * ostep;
  }
  for (; e < keep; e += STEP) {
    op[0] = sp[0];
    sp += SSTEP;
    op += ostep;
  }
}

// output base pointer of a tile, materialised early so that the kernel
// parameters it needs are not fetched on the critical path after the FFT
__device__ __forceinline__ cplx*
tile_base(const PassArgs& p, int b, int first_line)
{
  cplx* o = reinterpret_cast<cplx*>(p.out) + (long long)b * p.osb + first_line;
  asm volatile("" : "+l"(o));
  return o;
}

// ===========================================================================
// kernels
// ===========================================================================

// tile store of pass A in slab mode: rows are scattered to the ranks that own
// them (the FFTW-MPI transpose of the reference, fused into the producing
// pass as stores to peer memory over NVLink)
template <int N>
__device__ __forceinline__ void
tile_store_peers(cplx (&v)[LineCfg<N>::E], cplx* stage, int lic, int t,
                 const PassArgs& p, long long base_off, int lines_valid, int b,
                 int a0, double scale)
{
  using K = KCfg<N>;
  __syncthreads();
#pragma unroll
  for (int q = 0; q < K::E; q++) {
    const int e = t + K::TPL * q;
    if (e < p.keep)
      stage[stage_slot<N>(e, lic)] =
        make_double2(v[q].x * scale, v[q].y * scale);
  }
  __syncthreads();
  const int total = p.keep * K::LPC;
  for (int idx = threadIdx.x; idx < total; idx += PF_THREADS) {
    const int l = idx % K::LPC, e = idx / K::LPC;
    if (l < lines_valid) {
      const int owner = e / p.rows_per_peer;
      const int el = e - owner * p.rows_per_peer;
      cplx* o = reinterpret_cast<cplx*>(p.peer_out[owner]);
      if (p.blk_T > 0) {
        const int ag = a0 + l; // global line
        o[(((long long)b * p.blk_T + (ag >> 3)) * p.blk_rows + el) * 8 +
          (ag & 7)] = stage[stage_slot<N>(e, l)];
      } else
        o[base_off + (long long)el * p.ose + l] = stage[stage_slot<N>(e, l)];
    }
  }
}

// pass A: noise -> inverse FFT -> tile
#ifndef PF_GEN_MINB
#define PF_GEN_MINB 3
#endif
template <int N, int NOISE, bool HALF = false>
__global__ void __launch_bounds__(PF_THREADS, PF_GEN_MINB)
k_gen_tile(const PassArgs p, const cplx* __restrict__ tw2)
{
  using K = KCfg<N>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* sm = reinterpret_cast<cplx*>(smem_raw);
  const int lic = threadIdx.x / K::TPL, t = threadIdx.x % K::TPL;
  const int ntA = (p.A + K::LPC - 1) / K::LPC;
  const int at = blockIdx.x % ntA, b = blockIdx.x / ntA;
  const int al = at * K::LPC + lic; // local line, global line = a_begin + al
  cplx* obase = tile_base(p, b, p.a_begin + at * K::LPC);
  long long ose = p.ose;
  asm volatile("" : "+l"(ose));
  cplx v[K::E];
  load_noise<N, NOISE>(v, p, p.a_begin + al, b, t, al < p.A);
  line_fft<N, +1>(v, sm + lic * K::C::PITCH, t, tw2);
  const int lv = min(K::LPC, p.A - at * K::LPC);
  if (p.n_peers > 1)
    tile_store_peers<N>(v, sm, lic, t, p,
                        (long long)b * p.osb + p.a_begin + at * K::LPC, lv, b,
                        p.a_begin + at * K::LPC, p.scale);
  else
    tile_store<N, HALF>(v, sm, lic, t, p.keep, obase, ose, lv, p.scale);
}

// 1D: noise -> inverse FFT -> real part / imaginary part lines
template <int N, int NOISE>
__global__ void __launch_bounds__(PF_THREADS, 3)
k_gen_reim(const PassArgs p, const cplx* __restrict__ tw2)
{
  using K = KCfg<N>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* sm = reinterpret_cast<cplx*>(smem_raw);
  const int lic = threadIdx.x / K::TPL, t = threadIdx.x % K::TPL;
  const int ntA = (p.A + K::LPC - 1) / K::LPC;
  const int at = blockIdx.x % ntA, b = blockIdx.x / ntA;
  const int a = at * K::LPC + lic;
  const bool valid = a < p.A;
  cplx v[K::E];
  load_noise<N, NOISE>(v, p, a, b, t, valid);
  line_fft<N, +1>(v, sm + lic * K::C::PITCH, t, tw2);
  if (!valid)
    return;
  const long long line = (long long)b * p.A + a;
  double* o1 = reinterpret_cast<double*>(p.out);
  double* o2 = reinterpret_cast<double*>(p.out2);
  dispatch_transform(p.transform, [&](auto trc) {
    constexpr int TR = decltype(trc)::value;
#pragma unroll
    for (int q = 0; q < K::E; q++) {
      const int e = t + K::TPL * q;
      if (e < p.keep) {
        if (o1)
          o1[line * p.out_pitch + e] =
            apply_transform_c<TR>(v[q].x * p.scale, p.tparam);
        if (o2)
          o2[line * p.out_pitch + e] = v[q].y * p.scale;
      }
    }
  });
}

// contiguous complex lines -> FFT -> tile (pass B, F1, E1)
template <int N, int DIR>
__global__ void __launch_bounds__(PF_THREADS, 3)
k_c2t(const PassArgs p, const cplx* __restrict__ tw2)
{
  using K = KCfg<N>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* sm = reinterpret_cast<cplx*>(smem_raw);
  const int lic = threadIdx.x / K::TPL, t = threadIdx.x % K::TPL;
  const int ntA = (p.A + K::LPC - 1) / K::LPC;
  const int at = blockIdx.x % ntA, b = blockIdx.x / ntA;
  const int a = at * K::LPC + lic;
  cplx v[K::E];
  if (p.blk_T > 0) {
    // blocked exchange layout: line (a = plane, b = row), element e = 8*tile+j
    const cplx* base = reinterpret_cast<const cplx*>(p.in) +
                       ((long long)a * p.blk_T * p.blk_rows + b) * 8;
#pragma unroll
    for (int q = 0; q < K::E; q++) {
      const int e = t + K::TPL * q;
      v[q] = (a < p.A && e < p.n_valid)
               ? __ldg(base + (long long)(e >> 3) * p.blk_rows * 8 + (e & 7))
               : make_double2(0., 0.);
    }
  } else {
    const cplx* line = reinterpret_cast<const cplx*>(p.in) +
                       (long long)a * p.isa + (long long)b * p.isb;
    load_contig<N>(v, line, t, p.n_valid, a < p.A);
  }
  line_fft<N, DIR>(v, sm + lic * K::C::PITCH, t, tw2);
  const int lv = min(K::LPC, p.A - at * K::LPC);
  tile_store<N>(v, sm, lic, t, p.keep, tile_base(p, b, at * K::LPC), p.ose, lv,
                p.scale);
}

// epilogue of pass C: registers (x = line L0, y = line L1, thread t owns
// elements t + TPL*q) -> two real output lines, restricted to e < keep, value
// transform fused.  Pointers and bounds are materialised once; the common case
// (both lines valid, scale 1) is a run of predicated 8-byte stores.
template <int N, int TR, bool SCALE>
__device__ __forceinline__ void
store_real_pair_impl(const cplx (&v)[LineCfg<N>::E], double* __restrict__ o0,
                     double* __restrict__ o1, bool v0, bool v1, int t, int keep,
                     double scale, double tparam)
{
  using C = LineCfg<N>;
  if (v0 && v1) {
#pragma unroll
    for (int q = 0; q < C::E; q++) {
      if (t + C::TPL * q < keep) {
        double x = v[q].x, y = v[q].y;
        if constexpr (SCALE) {
          x *= scale;
          y *= scale;
        }
        o0[C::TPL * q] = apply_transform_c<TR>(x, tparam);
        o1[C::TPL * q] = apply_transform_c<TR>(y, tparam);
      }
    }
  } else if (v0) {
#pragma unroll
    for (int q = 0; q < C::E; q++) {
      if (t + C::TPL * q < keep) {
        double x = v[q].x;
        if constexpr (SCALE)
          x *= scale;
        o0[C::TPL * q] = apply_transform_c<TR>(x, tparam);
      }
    }
  }
}

template <int N>
__device__ __forceinline__ void
store_real_pair(const cplx (&v)[LineCfg<N>::E], const PassArgs& p, long long L0,
                bool v0, bool v1, int t)
{
  double* out = reinterpret_cast<double*>(p.out);
  double* o0 = out + L0 * p.out_pitch + t;
  double* o1 = o0 + p.out_pitch;
  const int keep = p.keep;
  const double scale = p.scale, tparam = p.tparam;
  dispatch_transform(p.transform, [&](auto trc) {
    constexpr int TR = decltype(trc)::value;
    if (scale == 1.0)
      store_real_pair_impl<N, TR, false>(v, o0, o1, v0, v1, t, keep, scale,
                                         tparam);
    else
      store_real_pair_impl<N, TR, true>(v, o0, o1, v0, v1, t, keep, scale,
                                        tparam);
  });
}

// pass C: two half-spectrum lines (m0 = N/2+1 complex each) -> one complex
// inverse FFT -> two real lines, restricted to e < keep, value transform
// fused.  c2r semantics of FFTW on non-Hermitian input: Im at k0 in {0,N/2}
// is ignored (r2cfieldbackend.hh:250-274).
template <int N>
__global__ void __launch_bounds__(PF_THREADS, 3)
k_c2r(const PassArgs p, const cplx* __restrict__ tw2)
{
  using K = KCfg<N>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* sm = reinterpret_cast<cplx*>(smem_raw);
  const int lic = threadIdx.x / K::TPL, t = threadIdx.x % K::TPL;
  const long long f = (long long)blockIdx.x * K::LPC + lic;
  const long long L0 = 2 * f, L1 = 2 * f + 1;
  const long long nl = (long long)p.A * p.B;
  const bool v0 = L0 < nl, v1 = L1 < nl;
  const cplx* in = reinterpret_cast<const cplx*>(p.in);
  const cplx* la = in + L0 * p.isa;
  const cplx* lb = in + L1 * p.isa;
  cplx v[K::E];
#pragma unroll
  for (int q = 0; q < K::E; q++) {
    const int e = t + K::TPL * q;
    const int kk = e <= N / 2 ? e : N - e;
    const cplx xa = v0 ? __ldg(la + kk) : make_double2(0., 0.);
    const cplx xb = v1 ? __ldg(lb + kk) : make_double2(0., 0.);
    if (e == 0 || 2 * e == N)
      v[q] = make_double2(xa.x, xb.x);
    else if (e < N / 2)
      v[q] = make_double2(xa.x - xb.y, xa.y + xb.x);
    else
      v[q] = make_double2(xa.x + xb.y, xb.x - xa.y);
  }
  line_fft<N, +1>(v, sm + lic * K::C::PITCH, t, tw2);
  store_real_pair<N>(v, p, L0, v0, v1, t);
}

// Long c2r axis (n0 = N1 * N > 1024), step 1 of 2.  The half-spectrum rows
// arrive PERMUTED: column k1 * (N/2+1) + k2 holds k0 = k1 + N1 * k2 (pass A
// writes them that way), so that sub-transform k1 -- the N-point inverse FFT
// over k2' of V[k1 + N1 k2'], V = A + iB Hermitian-extended as in k_c2r --
// reads two contiguous rows: k1 forwards and its mirror N1 - k1 backwards.
// Output U[pair][k1][j2] * W_n0^{+ j2 k1}; step 2 (k_long_c2r2, pf_api.cu)
// finishes with an N1-point transform over k1.
// Lines: a = k1 < A = N1, b = pair; p.n_valid = number of real lines.
template <int N>
__global__ void __launch_bounds__(PF_THREADS, 3)
k_long_c2r1(const PassArgs p, const cplx* __restrict__ tw2)
{
  using K = KCfg<N>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* sm = reinterpret_cast<cplx*>(smem_raw);
  const int lic = threadIdx.x / K::TPL, t = threadIdx.x % K::TPL;
  constexpr int M = N / 2 + 1;
  const int N1 = p.long_n1;
  const int ntA = (N1 + K::LPC - 1) / K::LPC;
  const int at = blockIdx.x % ntA, b = blockIdx.x / ntA;
  const int k1 = at * K::LPC + lic;
  const bool valid = k1 < N1;
  const long long L0 = 2ll * b, L1 = L0 + 1;
  const bool v0 = valid && L0 < p.n_valid, v1 = valid && L1 < p.n_valid;
  const cplx* in = reinterpret_cast<const cplx*>(p.in);
  const int kr = valid ? (N1 - k1) % N1 : 0;
  const cplx* a_dir = in + L0 * p.isa + (long long)(valid ? k1 : 0) * M;
  const cplx* a_mir = in + L0 * p.isa + (long long)kr * M;
  const cplx* b_dir = a_dir + p.isa;
  const cplx* b_mir = a_mir + p.isa;
  cplx v[K::E];
#pragma unroll
  for (int q = 0; q < K::E; q++) {
    const int e = t + K::TPL * q;
    const bool direct = (2 * e < N) || (k1 == 0 && 2 * e == N);
    const int col = direct ? e : (k1 == 0 ? N - e : N - 1 - e);
    const cplx xa = v0 ? __ldg((direct ? a_dir : a_mir) + col)
                       : make_double2(0., 0.);
    const cplx xb = v1 ? __ldg((direct ? b_dir : b_mir) + col)
                       : make_double2(0., 0.);
    if (k1 == 0 && (e == 0 || 2 * e == N))
      v[q] = make_double2(xa.x, xb.x);
    else if (direct)
      v[q] = make_double2(xa.x - xb.y, xa.y + xb.x);
    else
      v[q] = make_double2(xa.x + xb.y, xb.x - xa.y);
  }
  line_fft<N, +1>(v, sm + lic * K::C::PITCH, t, tw2);
  if (!valid)
    return;
  cplx* out = reinterpret_cast<cplx*>(p.out) + ((long long)b * N1 + k1) * N;
#pragma unroll
  for (int q = 0; q < K::E; q++) {
    const int j2 = t + K::TPL * q;
    out[j2] = cmulc(v[q], __ldg(p.long_tw + (long long)k1 * N + j2));
  }
}

// F0 / E0: two real lines (loaded or covariance-evaluated) -> one complex
// forward FFT -> split into the two half spectra -> tile store.
// Lines a (tile axis, pairs a, a+1), b; element e = x0.
template <int N, bool COV>
__global__ void __launch_bounds__(PF_THREADS, 3)
k_r2c(const PassArgs p, const cplx* __restrict__ tw2)
{
  using K = KCfg<N>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* sm = reinterpret_cast<cplx*>(smem_raw);
  const int lic = threadIdx.x / K::TPL, t = threadIdx.x % K::TPL;
  constexpr int LT = 2 * K::LPC; // real lines per tile
  const int ntA = (p.A + LT - 1) / LT;
  const int at = blockIdx.x % ntA, b = blockIdx.x / ntA;
  const int a0 = at * LT + 2 * lic, a1 = a0 + 1;
  const bool v0 = a0 < p.A, v1 = a1 < p.A;
  cplx v[K::E];
  if constexpr (COV) {
    unsigned ia[3] = { 0, 0, 0 }, ib[3] = { 0, 0, 0 };
    if (p.cov_axis_a >= 0) {
      ia[p.cov_axis_a] = a0;
      ib[p.cov_axis_a] = a1;
    }
    if (p.cov_axis_b >= 0) {
      ia[p.cov_axis_b] = b;
      ib[p.cov_axis_b] = b;
    }
#pragma unroll 4
    for (int q = 0; q < K::E; q++) {
      const int e = t + K::TPL * q;
      v[q].x = v0 ? cov_torus(p.cov, e, ia[1], ia[2]) : 0.;
      v[q].y = v1 ? cov_torus(p.cov, e, ib[1], ib[2]) : 0.;
    }
  } else {
    const double* in = reinterpret_cast<const double*>(p.in);
    const double* l0 = in + (long long)a0 * p.isa + (long long)b * p.isb;
    const double* l1 = in + (long long)a1 * p.isa + (long long)b * p.isb;
#pragma unroll
    for (int q = 0; q < K::E; q++) {
      const int e = t + K::TPL * q;
      v[q].x = (v0 && e < p.n_valid) ? __ldg(l0 + e) : 0.;
      v[q].y = (v1 && e < p.n_valid) ? __ldg(l1 + e) : 0.;
    }
  }
  line_fft<N, -1>(v, sm + lic * K::C::PITCH, t, tw2);
  __syncthreads();
#pragma unroll
  for (int q = 0; q < K::E; q++)
    sm[lic * K::ZP + t + K::TPL * q] = v[q];
  __syncthreads();
  cplx* out = reinterpret_cast<cplx*>(p.out);
  const long long base = (long long)b * p.osb + (long long)at * LT;
  const int lv = min(LT, p.A - at * LT);
  const int total = p.keep * LT;
  for (int idx = threadIdx.x; idx < total; idx += PF_THREADS) {
    const int l = idx % LT, k = idx / LT;
    if (l < lv) {
      const cplx zk = sm[(l >> 1) * K::ZP + k];
      const cplx zm = sm[(l >> 1) * K::ZP + ((N - k) & (N - 1))];
      cplx r;
      if ((l & 1) == 0)
        r = make_double2(0.5 * (zk.x + zm.x), 0.5 * (zk.y - zm.y));
      else
        r = make_double2(0.5 * (zk.y + zm.y), 0.5 * (zm.x - zk.x));
      out[base + (long long)k * p.ose + l] =
        make_double2(r.x * p.scale, r.y * p.scale);
    }
  }
}

__device__ __forceinline__ unsigned long long
ordered_bits(double v)
{
  const unsigned long long u = (unsigned long long)__double_as_longlong(v);
  return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}

// E2: contiguous complex lines -> forward FFT -> eigenvalues: real part,
// scale 1/N_ext, census and clamp (matrix.hh:283-306), contiguous store of
// the first `keep` entries.  COV1D: the line is evaluated from the covariance
// (1D DFT matrix backend, dftmatrixbackend.hh:209-238).
template <int N, bool COV1D>
__global__ void __launch_bounds__(PF_THREADS, 3)
k_lambda(const PassArgs p, const cplx* __restrict__ tw2)
{
  using K = KCfg<N>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* sm = reinterpret_cast<cplx*>(smem_raw);
  const int lic = threadIdx.x / K::TPL, t = threadIdx.x % K::TPL;
  const long long line = (long long)blockIdx.x * K::LPC + lic;
  const long long nl = (long long)p.A * p.B;
  const bool valid = line < nl;
  cplx v[K::E];
  if constexpr (COV1D) {
#pragma unroll 4
    for (int q = 0; q < K::E; q++) {
      const int e = t + K::TPL * q;
      v[q] = make_double2(valid ? cov_torus(p.cov, e, 0, 0) : 0., 0.);
    }
  } else {
    load_contig<N>(v, reinterpret_cast<const cplx*>(p.in) + line * p.isa, t,
                   p.n_valid, valid);
  }
  line_fft<N, -1>(v, sm + lic * K::C::PITCH, t, tw2);
  unsigned nsmall = 0, nzero = 0, nneg = 0;
  double smallest = 1.7976931348623157e308;
  double* out = reinterpret_cast<double*>(p.out);
#pragma unroll
  for (int q = 0; q < K::E; q++) {
    const int e = t + K::TPL * q;
    if (valid && e < p.keep) {
      double val = v[q].x * p.scale;
      if (val < smallest)
        smallest = val;
      if (val < 1e-6) {
        if (val < p.threshold) {
          if (val > -p.threshold)
            nzero++;
          else
            nneg++;
        } else
          nsmall++;
      }
      if (val < 0.)
        val = 0.;
      out[line * p.out_pitch + e] = val;
    }
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
      atomicAdd(p.census + 0, (unsigned long long)nsmall);
    if (nzero)
      atomicAdd(p.census + 1, (unsigned long long)nzero);
    if (nneg)
      atomicAdd(p.census + 2, (unsigned long long)nneg);
    atomicMin(p.smallest_bits, ordered_bits(smallest));
  }
}

// F2 + A fused: contiguous complex line -> forward FFT -> * g(lambda)*scale
// -> inverse FFT -> tile store (fieldBackend.mult between forwardTransform
// and backwardTransform, matrix.hh:1495-1513).  REAL1D: 1D variant, real
// input line, real output line (DFT field backend, component 0).
// HALF (tile variant, one rank): n_valid == keep == N / 2 known at compile
// time -- the zero half of the input folds away in the first butterflies of
// the forward transform, the dropped half of the output in the last ones of
// the inverse
template <int N, bool REAL1D, bool HALF = false>
__global__ void __launch_bounds__(PF_THREADS, 3)
k_spec(const PassArgs p, const cplx* __restrict__ tw2)
{
  using K = KCfg<N>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* sm = reinterpret_cast<cplx*>(smem_raw);
  const int lic = threadIdx.x / K::TPL, t = threadIdx.x % K::TPL;
  const int ntA = (p.A + K::LPC - 1) / K::LPC;
  const int at = blockIdx.x % ntA, b = blockIdx.x / ntA;
  const int a = at * K::LPC + lic;
  const bool valid = a < p.A;
  cplx v[K::E];
  if constexpr (REAL1D) {
    const double* line = reinterpret_cast<const double*>(p.in) +
                         (long long)a * p.isa + (long long)b * p.isb;
#pragma unroll
    for (int q = 0; q < K::E; q++) {
      const int e = t + K::TPL * q;
      v[q] = make_double2((valid && e < p.n_valid) ? __ldg(line + e) : 0., 0.);
    }
  } else {
    const cplx* line = reinterpret_cast<const cplx*>(p.in) +
                       (long long)a * p.isa + (long long)b * p.isb;
    if constexpr (HALF) {
#pragma unroll
      for (int q = 0; q < K::E; q++)
        v[q] = (q < K::E / 2 && valid) ? __ldg(line + t + K::TPL * q)
                                       : make_double2(0., 0.);
    } else
      load_contig<N>(v, line, t, p.n_valid, valid);
  }
  line_fft<N, -1>(v, sm + lic * K::C::PITCH, t, tw2);
  if (valid) {
    // slab mode: this rank holds the lines a_begin .. a_begin + A - 1
    const double* row = table_row(p, p.a_begin + a, b);
#pragma unroll
    for (int q = 0; q < K::E; q++) {
      const int e = t + K::TPL * q;
      const int em = (p.tab_mirror_e && e > N / 2) ? N - e : e;
      const double g = __ldg(row + em) * p.scale;
      v[q] = make_double2(v[q].x * g, v[q].y * g);
    }
  }
  line_fft<N, +1>(v, sm + lic * K::C::PITCH, t, tw2);
  if constexpr (REAL1D) {
    if (valid) {
      double* o1 = reinterpret_cast<double*>(p.out);
      const long long line = (long long)b * p.A + a;
#pragma unroll
      for (int q = 0; q < K::E; q++) {
        const int e = t + K::TPL * q;
        if (e < p.keep)
          o1[line * p.out_pitch + e] = v[q].x;
      }
    }
  } else {
    const int lv = min(K::LPC, p.A - at * K::LPC);
    if (p.n_peers > 1)
      tile_store_peers<N>(v, sm, lic, t, p,
                          (long long)b * p.osb + p.a_begin + at * K::LPC, lv, b,
                          p.a_begin + at * K::LPC, 1.0);
    else
      tile_store<N, HALF>(v, sm, lic, t, p.keep, tile_base(p, b, at * K::LPC),
                          p.ose, lv, 1.0);
  }
}

} // namespace pf
