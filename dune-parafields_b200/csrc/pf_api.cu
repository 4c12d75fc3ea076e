// pf_api.cu -- plan object, pass orchestration and the C ABI of
// libparafields_cuda.so (include/parafields_cuda.h).
//
// Host-side counterpart of parafields::Matrix (matrix.hh:59-1663): it owns the
// eigenvalue tables and work buffers in HBM and strings the pass kernels of
// pf_kernels.cuh together.  No CPU fallback exists: every compute entry point
// needs a CUDA device.
#include "../../include/parafields_cuda.h"
#include "pf_args.h"

#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <complex>
#include <cstdlib>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

using pf::PassArgs;
typedef double2 cplx;

// launchers instantiated per line length in pf_inst.cu
#define PF_DECL(n)                                                             \
  extern "C" int pf_launch_n##n(int, const PassArgs*, const void*, cudaStream_t);
PF_DECL(2)
PF_DECL(4)
PF_DECL(8)
PF_DECL(16)
PF_DECL(32)
PF_DECL(64)
PF_DECL(128)
PF_DECL(256)
PF_DECL(512)
PF_DECL(1024)
// persistent TMA-pipelined variants (pf_pipe.cuh), N >= 256
#define PF_DECL_PIPE(n)                                                        \
  extern "C" int pf_pipe_launch_n##n(int, const PassArgs*, const void*,        \
                                     const void*, int, cudaStream_t);
PF_DECL_PIPE(256)
PF_DECL_PIPE(512)
PF_DECL_PIPE(1024)
// streaming route for large cubic 3D tori (pf_stream.cuh), N = 512 / 1024
extern "C" int pf_stream_launch_n512(int, const void*, const void*, const void*,
                                     int, cudaStream_t);
extern "C" int pf_stream_launch_n1024(int, const void*, const void*, const void*,
                                      int, cudaStream_t);

namespace {

thread_local std::string g_err;

int
fail(int code, const std::string& msg)
{
  g_err = msg;
  return code;
}

#define PF_CUDA(call)                                                          \
  do {                                                                         \
    cudaError_t e__ = (call);                                                  \
    if (e__ != cudaSuccess)                                                    \
      return fail(PF_ERR_CUDA, std::string(#call) + ": " +                     \
                                 cudaGetErrorString(e__));                     \
  } while (0)

#define PF_TRY(call)                                                           \
  do {                                                                         \
    int rc__ = (call);                                                         \
    if (rc__ != 0)                                                             \
      return rc__;                                                             \
  } while (0)

typedef int (*launch_fn)(int, const PassArgs*, const void*, cudaStream_t);

launch_fn
launcher_for(long long n)
{
  switch (n) {
    case 2:
      return pf_launch_n2;
    case 4:
      return pf_launch_n4;
    case 8:
      return pf_launch_n8;
    case 16:
      return pf_launch_n16;
    case 32:
      return pf_launch_n32;
    case 64:
      return pf_launch_n64;
    case 128:
      return pf_launch_n128;
    case 256:
      return pf_launch_n256;
    case 512:
      return pf_launch_n512;
    case 1024:
      return pf_launch_n1024;
  }
  return nullptr;
}

typedef int (*pipe_fn)(int, const PassArgs*, const void*, const void*, int,
                       cudaStream_t);
pipe_fn
pipe_launcher_for(long long n)
{
  switch (n) {
    case 256:
      return pf_pipe_launch_n256;
    case 512:
      return pf_pipe_launch_n512;
    case 1024:
      return pf_pipe_launch_n1024;
  }
  return nullptr;
}

// cuTensorMapEncodeTiled through the runtime (no link-time libcuda dependency)
typedef CUresult (*encode_tiled_fn)(
  CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

encode_tiled_fn
get_encode_tiled()
{
  static encode_tiled_fn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault,
                                &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<encode_tiled_fn>(p);
    else
      cudaGetLastError();
  }
  return fn;
}

bool
is_device_ptr(const void* p)
{
  if (!p)
    return false;
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

// ---------------------------------------------------------------------------
// small elementwise / reduction kernels
// ---------------------------------------------------------------------------
__global__ void
k_derive_table(const double* __restrict__ lam, double* __restrict__ out,
               size_t n, int mode)
{
  // mode 0: sqrt(l)/sqrt(2)   1: sqrt(l)   2: 1/l
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  const double l = lam[i];
  double r;
  if (mode == 0)
    r = sqrt(l) / sqrt(2.);
  else if (mode == 1)
    r = sqrt(l);
  else
    r = 1. / l;
  out[i] = r;
}

// same with a padded output pitch (rows of P values at pitch Pg >= P, padding
// zeroed): the sigma table of generateField keeps its rows 16-byte aligned so
// that a row can travel by one bulk copy (kp_gen3)
__global__ void
k_derive_table_padded(const double* __restrict__ lam, double* __restrict__ out,
                      size_t rows, long long P, long long Pg, int mode)
{
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * (size_t)Pg)
    return;
  const size_t r = i / (size_t)Pg;
  const long long e = (long long)(i % (size_t)Pg);
  double v = 0.;
  if (e < P) {
    const double l = lam[r * (size_t)P + e];
    v = mode == 0 ? sqrt(l) / sqrt(2.) : (mode == 1 ? sqrt(l) : 1. / l);
  }
  out[i] = v;
}

struct GatherArgs
{
  int dim;
  long long n[3];       // extended cells
  long long ob[3];      // output bounds
  int transposed;       // output order k0,(k2,k1) instead of k0,(k1,k2)
  int compact;          // table mirrors a / e
  long long m0, P;      // table geometry
};

// eigenvalue table -> user layout (PF_LAMBDA_NATURAL / PF_LAMBDA_REFERENCE)
__global__ void
k_gather_lambda(const double* __restrict__ lam, double* __restrict__ out,
                GatherArgs g, size_t total)
{
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total)
    return;
  long long k[3] = { 0, 0, 0 };
  size_t r = i;
  if (g.dim == 3 && g.transposed) {
    k[0] = r % g.ob[0];
    r /= g.ob[0];
    k[2] = r % g.ob[2];
    r /= g.ob[2];
    k[1] = r;
  } else {
    for (int d = 0; d < g.dim; d++) {
      k[d] = r % g.ob[d];
      r /= g.ob[d];
    }
  }
  // mirror into the stored range
  long long km[3];
  for (int d = 0; d < 3; d++)
    km[d] = (g.compact && k[d] > g.n[d] / 2) ? g.n[d] - k[d] : k[d];
  double v;
  if (g.dim == 3)
    v = lam[(km[1] * g.m0 + k[0]) * g.P + km[2]];
  else if (g.dim == 2)
    v = lam[k[0] * g.P + km[1]];
  else
    v = lam[k[0]];
  out[i] = v;
}

__global__ void
k_axpy(size_t n, double alpha, const double* __restrict__ x,
       double* __restrict__ y)
{
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x)
    y[i] += alpha * x[i];
}
__global__ void
k_scale(size_t n, double alpha, double* __restrict__ x)
{
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x)
    x[i] *= alpha;
}
// ValueTransform::apply (mutators.hh:36-162) on one value
__device__ __forceinline__ double
transform_value(double v, int kind, double lambda)
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

// StochasticPart::refine (stochastic.hh:597-670): blockIdx.y / z walk the old
// rows and planes (no index division), threads walk the old cells of a row;
// a thread reads one old value and stores it as one 16-byte pair into each
// of the 2^(dim-1) new rows, so a warp writes 512 contiguous bytes per row.
// o* = old cells; new rows have o0 pairs.
template <int DIM>
__global__ void
k_refine(unsigned o0, unsigned o1, unsigned o2, const double* __restrict__ in,
         double2* __restrict__ out)
{
  for (unsigned i2 = blockIdx.z; i2 < o2; i2 += gridDim.z)
    for (unsigned i1 = blockIdx.y; i1 < o1; i1 += gridDim.y) {
      const double* __restrict__ src = in + (size_t)o0 * (i1 + (size_t)o1 * i2);
      double2* __restrict__ r0 =
        out + (size_t)o0 * (2 * (size_t)i1 + 2 * (size_t)o1 * (DIM == 3 ? 2 * (size_t)i2 : 0));
      double2* __restrict__ r1 = r0 + (size_t)o0 * 2 * o1; // plane 2 i2 + 1
      for (unsigned i0 = blockIdx.x * blockDim.x + threadIdx.x; i0 < o0;
           i0 += gridDim.x * blockDim.x) {
        const double v = src[i0];
        const double2 w = make_double2(v, v);
        r0[i0] = w;
        r0[i0 + o0] = w;
        if (DIM == 3) {
          r1[i0] = w;
          r1[i0 + o0] = w;
        }
      }
    }
}

// StochasticPart::coarsen (stochastic.hh:679-755): one thread per new cell,
// 16-byte loads of the 2^(dim-1) fine row pairs, summed in the reference's
// order (x0, then x1, then x2) so the result is bit-identical.
// n* = new cells; the fine rows have 2*n0 entries.
template <int DIM>
__global__ void
k_coarsen(unsigned n0, unsigned n1, size_t total,
          const double2* __restrict__ in, double* __restrict__ out)
{
  for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (size_t)gridDim.x * blockDim.x) {
    const size_t row = t / n0;
    const unsigned i0 = (unsigned)(t - row * n0);
    const size_t i2 = DIM == 3 ? row / n1 : 0;
    const size_t i1 = row - i2 * n1;
    // index in units of double2 (fine row length n0 pairs)
    const size_t base = i0 + (size_t)n0 * (2 * i1 + 2 * (size_t)n1 * 2 * i2);
    const double2 a = in[base], b = in[base + n0];
    double v = 0.;
    v = __dadd_rn(v, a.x);
    v = __dadd_rn(v, a.y);
    v = __dadd_rn(v, b.x);
    v = __dadd_rn(v, b.y);
    if (DIM == 3) {
      const size_t up = base + (size_t)n0 * 2 * n1;
      const double2 c = in[up], d = in[up + n0];
      v = __dadd_rn(v, c.x);
      v = __dadd_rn(v, c.y);
      v = __dadd_rn(v, d.x);
      v = __dadd_rn(v, d.y);
    }
    out[t] = v / (DIM == 3 ? 8. : 4.);
  }
}

struct FieldGeom
{
  int dim;
  unsigned cells[3];  // local block
  unsigned offset[3]; // first global index of the block
  unsigned gcells[3]; // whole grid
  double ext[3];
  double center[3];
};

// StochasticPart::localize (stochastic.hh:798-817) in two steps.  The
// Gaussian weight factor * exp(-|x - c|^2 / (2 r^2)) is separable, so
// k_localize_axes evaluates, per axis j and local index i,
// exp(-0.5 (loc_j(i) - center_j)^2 / r^2) with loc by indicesToCoords
// (fieldtraits.hh:461-469) into one small table (cells[0] + cells[1] +
// cells[2] doubles; factor folded into axis 0), and k_localize multiplies the
// field with the product of the three entries: a pure HBM sweep.  Against the
// reference's single exp() of the summed distance this differs by a few ulp
// (tests: 1e-14 relative to the max norm).
__global__ void
k_localize_axes(FieldGeom g, double factor, double rr, double* __restrict__ table)
{
  const unsigned total = g.cells[0] + (g.dim > 1 ? g.cells[1] : 0) +
                         (g.dim > 2 ? g.cells[2] : 0);
  for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += gridDim.x * blockDim.x) {
    int j = 0;
    unsigned i = t;
    if (i >= g.cells[0]) {
      i -= g.cells[0];
      j = 1;
      if (i >= g.cells[1]) {
        i -= g.cells[1];
        j = 2;
      }
    }
    const double loc = __ddiv_rn(
      __dmul_rn((double)(i + g.offset[j]) + 0.5, g.ext[j]), (double)g.gcells[j]);
    const double d = __dadd_rn(loc, -g.center[j]);
    const double w = exp(__ddiv_rn(__dmul_rn(-0.5, __dmul_rn(d, d)), rr));
    table[t] = j == 0 ? __dmul_rn(factor, w) : w;
  }
}

template <int DIM>
__global__ void
k_localize(unsigned c0, unsigned c1, unsigned c2,
           const double* __restrict__ table, double* __restrict__ data)
{
  constexpr unsigned ROWS = 4; // rows per block step: t0 is loaded once for them
  const double* __restrict__ t0 = table;
  const double* __restrict__ t1 = table + c0;
  const double* __restrict__ t2 = t1 + c1;
  for (unsigned i2 = blockIdx.z; i2 < c2; i2 += gridDim.z)
    for (unsigned b1 = blockIdx.y * ROWS; b1 < c1; b1 += gridDim.y * ROWS) {
      const double w2 = DIM > 2 ? t2[i2] : 1.;
      for (unsigned i0 = blockIdx.x * blockDim.x + threadIdx.x; i0 < c0;
           i0 += gridDim.x * blockDim.x) {
        const double w0 = t0[i0];
        double v[ROWS];
#pragma unroll
        for (unsigned r = 0; r < ROWS; r++)
          if (b1 + r < c1)
            v[r] = data[(size_t)c0 * (b1 + r + (size_t)c1 * i2) + i0];
#pragma unroll
        for (unsigned r = 0; r < ROWS; r++)
          if (b1 + r < c1) {
            double w = w0;
            if (DIM > 1)
              w = __dmul_rn(w, t1[b1 + r]);
            if (DIM > 2)
              w = __dmul_rn(w, w2);
            data[(size_t)c0 * (b1 + r + (size_t)c1 * i2) + i0] =
              __dmul_rn(v[r], w);
          }
      }
    }
}

// TrendPart::evaluate (trend.hh:1491-1505) over TrendComponent::evaluate
// (trend.hh:601-634); coeff = meanVector + shiftVector.  Explicit _rn
// operations reproduce the reference's evaluation order without contractions;
// the disk's exp() may differ from libm by an ulp.
struct TrendSet
{
  int n;
  int kind[PF_MAX_TREND_COMPONENTS];
  double coeff[PF_MAX_TREND_COMPONENTS][7];
};

template <int DIM>
__device__ __forceinline__ double
trend_value(const TrendSet& tr, const double* ext, const double* x)
{
  double output = 0.;
  for (int t = 0; t < tr.n; t++) {
    const double* c = tr.coeff[t];
    double comp = 0.;
    switch (tr.kind[t]) {
      case PF_TREND_MEAN:
        comp = c[0];
        break;
      case PF_TREND_SLOPE:
#pragma unroll
        for (int i = 0; i < DIM; i++)
          comp = __dadd_rn(
            comp, __dmul_rn(c[i], __dadd_rn(x[i], -__dmul_rn(ext[i], 0.5))));
        break;
      case PF_TREND_DISK: {
        double dist = 0.;
#pragma unroll
        for (int i = 0; i < DIM; i++) {
          const double d = __dadd_rn(x[i], -c[i]);
          dist = __dadd_rn(dist, __dmul_rn(d, d));
        }
        comp = exp(__dmul_rn(__ddiv_rn(-dist, __dmul_rn(c[DIM], c[DIM])),
                             c[DIM + 1]));
        break;
      }
      default: { // PF_TREND_BLOCK
        bool inside = true;
#pragma unroll
        for (int i = 0; i < DIM; i++)
          if (fabs(__dadd_rn(x[i], -c[i])) > __dmul_rn(0.5, c[DIM + i]))
            inside = false;
        comp = inside ? c[2 * DIM] : 0.;
      }
    }
    output = __dadd_rn(output, comp);
  }
  return output;
}

// RandomField::bulkEvaluate (randomfield.hh:597-631): one fused sweep
// out = T(field + trend(centre)), centre_j = (i_j + 0.5) * mesh_j with
// mesh_j = extensions_j / cells_j.  The reference sends every centre through
// coordsToIndices again; (i + 0.5) (1 + 1e-6 / cells) lies strictly inside
// (i, i + 1), so that lookup returns cell i itself and is skipped here.
// Per-axis parts of the trend components at the cell centres (the terms of
// TrendComponent::evaluate, trend.hh:601-634, that depend on one coordinate
// only): slope c_j (x_j - ext_j / 2), disk (x_j - c_j)^2, block 1 / 0 for
// |x_j - c_j| <= w_j / 2.  Layout [component][c0 + c1 + c2].  Summing them in
// k_bulk_evaluate in the reference's order gives the per-cell values bit for
// bit (same operations, same order) without recomputing them per cell.
__global__ void
k_trend_axes(FieldGeom g, TrendSet tr, double* __restrict__ table)
{
  const unsigned L = g.cells[0] + g.cells[1] + g.cells[2];
  const unsigned total = L * (unsigned)tr.n;
  for (unsigned e = blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += gridDim.x * blockDim.x) {
    const unsigned t = e / L;
    unsigned i = e - t * L;
    int j = 0;
    if (i >= g.cells[0]) {
      i -= g.cells[0];
      j = 1;
      if (i >= g.cells[1]) {
        i -= g.cells[1];
        j = 2;
      }
    }
    const double* c = tr.coeff[t];
    const double x = __dmul_rn((double)i + 0.5, g.center[j]); // center[] = mesh
    double v = 0.;
    if (j < g.dim) {
      if (tr.kind[t] == PF_TREND_SLOPE)
        v = __dmul_rn(c[j], __dadd_rn(x, -__dmul_rn(g.ext[j], 0.5)));
      else if (tr.kind[t] == PF_TREND_DISK) {
        const double d = __dadd_rn(x, -c[j]);
        v = __dmul_rn(d, d);
      } else if (tr.kind[t] == PF_TREND_BLOCK)
        v = fabs(__dadd_rn(x, -c[j])) > __dmul_rn(0.5, c[g.dim + j]) ? 0. : 1.;
    } else if (tr.kind[t] == PF_TREND_BLOCK)
      v = 1.;
    table[e] = v;
  }
}

template <int DIM, bool TREND>
__global__ void
k_bulk_evaluate(FieldGeom g, TrendSet tr, const double* __restrict__ table,
                const double* __restrict__ field, int kind, double lambda,
                double* __restrict__ out)
{
  constexpr unsigned ROWS = 4; // rows per block step: loads of all rows first
  const unsigned c0 = g.cells[0], c1 = DIM > 1 ? g.cells[1] : 1,
                 c2 = DIM > 2 ? g.cells[2] : 1;
  const unsigned L = g.cells[0] + g.cells[1] + g.cells[2];
  for (unsigned i2 = blockIdx.z; i2 < c2; i2 += gridDim.z)
    for (unsigned b1 = blockIdx.y * ROWS; b1 < c1; b1 += gridDim.y * ROWS) {
      for (unsigned i0 = blockIdx.x * blockDim.x + threadIdx.x; i0 < c0;
           i0 += gridDim.x * blockDim.x) {
        double v[ROWS], tsum[ROWS];
#pragma unroll
        for (unsigned r = 0; r < ROWS; r++) {
          tsum[r] = 0.;
          if (b1 + r < c1)
            v[r] = field[(size_t)c0 * (b1 + r + (size_t)c1 * i2) + i0];
        }
        if (TREND) {
          for (int t = 0; t < tr.n; t++) {
            const double* __restrict__ tab = table + (size_t)t * L;
            const double* c = tr.coeff[t];
            const int k = tr.kind[t];
            const double a0 = tab[i0];
            const double a2 = DIM > 2 ? tab[g.cells[0] + g.cells[1] + i2] : 0.;
#pragma unroll
            for (unsigned r = 0; r < ROWS; r++) {
              const unsigned i1 = min(b1 + r, c1 - 1);
              const double a1 = DIM > 1 ? tab[g.cells[0] + i1] : 0.;
              double comp;
              if (k == PF_TREND_MEAN)
                comp = c[0];
              else if (k == PF_TREND_BLOCK) {
                bool inside = a0 != 0.;
                if (DIM > 1)
                  inside = inside && a1 != 0.;
                if (DIM > 2)
                  inside = inside && a2 != 0.;
                comp = inside ? c[2 * DIM] : 0.;
              } else {
                double sum = a0; // 0 + a0
                if (DIM > 1)
                  sum = __dadd_rn(sum, a1);
                if (DIM > 2)
                  sum = __dadd_rn(sum, a2);
                comp = k == PF_TREND_SLOPE
                         ? sum
                         : exp(__dmul_rn(
                             __ddiv_rn(-sum, __dmul_rn(c[DIM], c[DIM])),
                             c[DIM + 1]));
              }
              tsum[r] = __dadd_rn(tsum[r], comp);
            }
          }
        }
#pragma unroll
        for (unsigned r = 0; r < ROWS; r++)
          if (b1 + r < c1) {
            const double w = TREND ? __dadd_rn(v[r], tsum[r]) : v[r];
            out[(size_t)c0 * (b1 + r + (size_t)c1 * i2) + i0] =
              transform_value(w, kind, lambda);
          }
      }
    }
}

// StochasticPart::evaluate (stochastic.hh:480-574) for a batch of locations
// on one rank + RandomField::evaluate's value transform (randomfield.hh:
// 585-595).  coordsToIndices: fieldtraits.hh:440-448.  A location on the
// upper face of exactly one dimension reads the (zero-filled) overlap, on an
// edge or corner it steps back into the last cell; anything outside the
// closed box raises *bad.
template <int DIM, bool TREND>
__global__ void
k_evaluate(FieldGeom g, TrendSet tr, const double* __restrict__ field,
           size_t npoints, const double* __restrict__ loc, int kind,
           double lambda, double* __restrict__ out, int* __restrict__ bad)
{
  // U points per thread and step: all coordinate loads are issued before the
  // first dependent gather, which keeps enough bytes in flight per SM
  constexpr int U = 4;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t p0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p0 < npoints;
       p0 += U * stride) {
    double x[U][DIM];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const size_t p = p0 + u * stride;
#pragma unroll
      for (int i = 0; i < DIM; i++)
        x[u][i] = p < npoints ? loc[p * DIM + i] : 0.;
    }
    size_t index[U];
    int state[U]; // 0 inside, 1 overlap (zero), 2 invalid
#pragma unroll
    for (int u = 0; u < U; u++) {
      unsigned idx[DIM];
      int outside = 0;
      bool invalid = false;
#pragma unroll
      for (int i = 0; i < DIM; i++) {
        // Index(x * (cells + 1e-6) / ext): the product with the reciprocal
        // is within 2 ulp of the quotient, so both truncate alike unless the
        // value sits next to an integer -- only then the division is done
        const double num = __dmul_rn(x[u][i], (double)g.cells[i] + 1e-6);
        double q = __dmul_rn(num, g.center[i]); // center[] holds 1 / ext here
        if (fabs(q - rint(q)) <= 1e-9 * (1. + fabs(q)))
          q = __ddiv_rn(num, g.ext[i]);
        idx[i] = 0;
        if (!(x[u][i] >= 0.) || !(q < 4294967295.))
          invalid = true;
        else {
          idx[i] = (unsigned)q; // truncation, like the reference's Index(...)
          if (idx[i] > g.cells[i])
            invalid = true;
          else if (idx[i] == g.cells[i])
            outside++;
        }
      }
      size_t lin = 0, step = 1;
#pragma unroll
      for (int i = 0; i < DIM; i++) {
        const unsigned k = idx[i] == g.cells[i] ? idx[i] - 1 : idx[i];
        lin += step * k;
        step *= g.cells[i];
      }
      index[u] = lin;
      state[u] = invalid ? 2 : (outside == 1 ? 1 : 0);
    }
    double v[U];
#pragma unroll
    for (int u = 0; u < U; u++)
      v[u] = (state[u] == 0 && p0 + u * stride < npoints) ? field[index[u]] : 0.;
#pragma unroll
    for (int u = 0; u < U; u++) {
      const size_t p = p0 + u * stride;
      if (p < npoints) {
        if (state[u] == 2) {
          *bad = 1;
          out[p] = nan("");
        } else {
          // RandomField::evaluate: T(stochastic + trend)
          const double w =
            TREND ? __dadd_rn(v[u], trend_value<DIM>(tr, g.ext, x[u])) : v[u];
          out[p] = transform_value(w, kind, lambda);
        }
      }
    }
  }
}

__global__ void
k_fill(size_t n, double v, double* __restrict__ x)
{
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x)
    x[i] = v;
}
// r = b - Mx
__global__ void
k_residual(size_t n, const double* __restrict__ b, const double* __restrict__ Mx,
           double* __restrict__ r)
{
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x)
    r[i] = b[i] - Mx[i];
}
// ---- innerCG with device-resident scalars (matrix.hh:1385-1444) ------------
// S[0] z.r   S[1] r.r   S[2] first energy   S[3] d.Md   S[4] old energy
// S[5] energy   S[6] new z.r   S[7] converged (0 / 1)   S[8] iterations done
// Every update is guarded by S[7], so an iteration enqueued after convergence
// (the host learns about it one iteration late) changes nothing.
__global__ void
k_cg_init(double* S)
{
  S[7] = sqrt(fabs(S[1])) < 1e-6 ? 1. : 0.;
  S[8] = 0.;
  // the energy BEFORE an update is the energy after the previous one (same
  // vectors, same reduction): carried over instead of being reduced again
  S[4] = S[2];
}
// x += alpha d, Mx += alpha Md with alpha = (z.r) / (d.Md): k_cg_update_fused
// d = z + beta d with beta = (1 / old z.r) * new z.r  (the reference's order)
__global__ void
k_cg_direction(size_t n, const double* __restrict__ S,
               const double* __restrict__ z, double* __restrict__ d)
{
  if (S[7] != 0.)
    return;
  double beta = 1. / S[0];
  beta *= S[6];
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x)
    d[i] = z[i] + beta * d[i];
}
// end of an iteration: shift z.r, energy stagnation test, count
__global__ void
k_cg_scalars(double* S)
{
  if (S[7] != 0.)
    return;
  S[0] = S[6];
  const double value = S[5], oldValue = S[4], firstValue = S[2];
  if (value != firstValue)
    if (fabs(value - oldValue) / fabs(value - firstValue) < 1e-16)
      S[7] = 1.;
  S[4] = value; // "old energy" of the next iteration
  S[8] += 1.;
}

__global__ void
k_transform(size_t n, const double* __restrict__ in, double* __restrict__ out,
            int kind, double lambda)
{
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x) {
    double v = in[i];
    switch (kind) {
      case 1:
        v = exp(v);
        break;
      case 2:
        v = fabs(v);
        break;
      case 3:
        v = v < 0. ? -1. : 1.;
        break;
      case 4:
        v = pow(v * lambda + 1, -lambda);
        break;
    }
    out[i] = v;
  }
}

// reductions: mode 0 dot(x,y), 1 sum|x|, 2 max|x|, 3 sum x*(0.5*y - z),
// 4 max|x| > 1e-10 flag (as max)
constexpr int RED_BLOCKS = 592; // 4 per SM
constexpr int RED_THREADS = 256;

__device__ __forceinline__ double
red_combine(double a, double b, int mode)
{
  return (mode == 2) ? fmax(a, b) : a + b;
}

__global__ void
k_reduce1(size_t n, const double* __restrict__ x, const double* __restrict__ y,
          const double* __restrict__ z, int mode, double* __restrict__ partial)
{
  double acc = 0.;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x) {
    double t;
    if (mode == 0)
      t = x[i] * y[i];
    else if (mode == 1)
      t = fabs(x[i]);
    else if (mode == 2)
      t = fabs(x[i]);
    else if (mode == 4)
      t = (x[i] != y[i]) ? 1. : 0.; // number of differing entries
    else
      t = x[i] * (0.5 * y[i] - z[i]);
    acc = red_combine(acc, t, mode);
  }
  __shared__ double sh[RED_THREADS / 32];
  for (int o = 16; o > 0; o >>= 1)
    acc = red_combine(acc, __shfl_xor_sync(0xffffffffu, acc, o), mode);
  if ((threadIdx.x & 31) == 0)
    sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double r = sh[0];
    for (int w = 1; w < RED_THREADS / 32; w++)
      r = red_combine(r, sh[w], mode);
    partial[blockIdx.x] = r;
  }
}
// k_cg_update, k_residual and the energy reduction x.(Mx/2 - b) that follows them
// in ONE sweep (4 of the 22 vector passes of a PCG iteration less): same
// expressions, same grid and summation order as the separate kernels, so the
// iterates are bit-identical.  partial[] as in k_reduce1.
__global__ void
k_cg_update_fused(size_t n, const double* __restrict__ S,
                  const double* __restrict__ d, const double* __restrict__ Md,
                  double* __restrict__ x, double* __restrict__ Mx,
                  const double* __restrict__ b, double* __restrict__ r,
                  double* __restrict__ partial)
{
  const bool live = S[7] == 0.;
  const double alpha = live ? S[0] / S[3] : 0.;
  double acc = 0.;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x) {
    double xi = x[i], mi = Mx[i];
    if (live) {
      xi += alpha * d[i];
      mi += alpha * Md[i];
      x[i] = xi;
      Mx[i] = mi;
    }
    const double bi = b[i];
    r[i] = bi - mi;
    acc += xi * (0.5 * mi - bi);
  }
  __shared__ double sh[RED_THREADS / 32];
  for (int o = 16; o > 0; o >>= 1)
    acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0)
    sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = sh[0];
    for (int w = 1; w < RED_THREADS / 32; w++)
      t += sh[w];
    partial[blockIdx.x] = t;
  }
}

__global__ void
k_reduce2(int nb, const double* __restrict__ partial, int mode,
          double* __restrict__ result)
{
  // single block, fixed order: deterministic
  __shared__ double sh[RED_THREADS];
  double acc = 0.;
  for (int i = threadIdx.x; i < nb; i += RED_THREADS)
    acc = red_combine(acc, partial[i], mode);
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s = RED_THREADS / 2; s > 0; s >>= 1) {
    if (threadIdx.x < s)
      sh[threadIdx.x] = red_combine(sh[threadIdx.x], sh[threadIdx.x + s], mode);
    __syncthreads();
  }
  if (threadIdx.x == 0)
    result[0] = sh[0];
}

// iid N(0,1) per cell, device Philox (Matrix::generateUncorrelatedField)
__global__ void
k_uncorrelated(size_t n, unsigned long long seed, unsigned long long q0,
               double* __restrict__ out);

} // namespace

#include "pf_rng.cuh"
#include "pf_long.cuh"
#include "pf_stream.cuh"
#include "pf_generic.cuh"

namespace {

__global__ void
k_uncorrelated(size_t n, unsigned long long seed, unsigned long long q0,
               double* __restrict__ out)
{
  // one Philox call -> two (r1,r2) pairs -> four cells
  const size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t i = 4 * q;
  if (i >= n)
    return;
  const uint4 w = pf::philox4x32_10(q0 + q, 0xFFFFFFFFFFFFFFFFull, seed);
  const double2 a = pf::box_muller_f64(w.x, w.y);
  const double2 b = pf::box_muller_f64(w.z, w.w);
  out[i] = a.x;
  if (i + 1 < n)
    out[i + 1] = a.y;
  if (i + 2 < n)
    out[i + 2] = b.x;
  if (i + 3 < n)
    out[i + 3] = b.y;
}

// ---- slab mode: flags in peer memory ---------------------------------------
// thread i publishes `epoch` in slot `me` of rank i's flag array
__global__ void
k_slab_signal(unsigned long long* const* flags, int n, int me,
              unsigned long long epoch)
{
  const int i = threadIdx.x;
  if (i < n) {
    __threadfence_system();
    volatile unsigned long long* f = flags[i] + me;
    *f = epoch;
    __threadfence_system();
  }
}
// thread i waits until rank i has published at least `epoch` here.  The spin
// is bounded (limit_ns of globaltimer, PF_SLAB_TIMEOUT_S, default 120 s) so
// that a lost peer cannot hang the device for ever; a timeout is FATAL for the
// plan: the flag (pinned host memory) is sticky, every entry point that
// synchronises reports PF_ERR_CUDA once it is set, and nothing computed after
// it may be used.
__device__ __forceinline__ unsigned long long
global_ns()
{
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__global__ void
k_slab_wait(const unsigned long long* mine, int n, unsigned long long epoch,
            volatile int* timeout_flag, unsigned long long limit_ns)
{
  const int i = threadIdx.x;
  if (i < n) {
    const volatile unsigned long long* f = mine + i;
    const unsigned long long t0 = global_ns();
    unsigned spins = 0;
    while (*f < epoch) {
      if ((++spins & 1023u) == 0 &&
          (*timeout_flag || global_ns() - t0 > limit_ns)) {
        *timeout_flag = 1;
        break;
      }
    }
    __threadfence_system();
  }
}

// cross-rank combination of one scalar (CG dot products, norms): thread i
// publishes this rank's value in slot [seq & 1][me] of rank i, then waits for
// rank i's value here; thread 0 combines the n values in rank order, so every
// rank gets the same bits.  Slots live behind the flag arrays (16 u64 in).
__global__ void
k_slab_reduce(unsigned long long* const* flags, unsigned long long* mine, int n,
              int me, unsigned long long seq, int mode, double* value,
              volatile int* timeout_flag, unsigned long long limit_ns)
{
  __shared__ double vals[8];
  const int i = threadIdx.x;
  const int par = (int)(seq & 1);
  if (i < n) {
    volatile unsigned long long* slot = flags[i] + 16 + 2 * (par * 8 + me);
    slot[1] = (unsigned long long)__double_as_longlong(*value);
    __threadfence_system();
    slot[0] = seq;
    __threadfence_system();
    const volatile unsigned long long* in = mine + 16 + 2 * (par * 8 + i);
    const unsigned long long t0 = global_ns();
    unsigned spins = 0;
    while (in[0] < seq) {
      if ((++spins & 1023u) == 0 &&
          (*timeout_flag || global_ns() - t0 > limit_ns)) {
        *timeout_flag = 1;
        break;
      }
    }
    __threadfence_system();
    vals[i] = __longlong_as_double((long long)in[1]);
  }
  __syncthreads();
  if (i == 0) {
    double r = vals[0];
    for (int q = 1; q < n; q++)
      r = red_combine(r, vals[q], mode);
    *value = r;
  }
}

inline unsigned
grid_for(size_t n, int threads = 256)
{
  size_t b = (n + threads - 1) / threads;
  return (unsigned)std::min<size_t>(b, 148 * 16);
}

} // namespace

// embedding.mergeRecursions with the magic value 99 (matrix.hh:691-719);
// ~0u = "magic recursions value 99 not implemented for choice of covariance"
static unsigned
merge_recursions(const pf_desc& d, int dim)
{
  if (d.periodization != 1)
    return 0;
  if (d.merge_recursions != 99)
    return d.merge_recursions;
  switch (d.covariance) {
    case PF_COV_MATERN:
      return (unsigned)std::floor(d.cov_param[0] * 2. + double(dim) / 2.);
    case PF_COV_GAMMA_EXPONENTIAL:
    case PF_COV_GENERALIZED_CAUCHY:
      return (unsigned)std::floor(d.cov_param[0] + double(dim) / 2.);
    case PF_COV_EXPONENTIAL:
      return 2;
    case PF_COV_MATERN32:
      return 4;
    case PF_COV_MATERN52:
      return 6;
    case PF_COV_CAUCHY:
      return 3;
    default:
      return ~0u;
  }
}

// ===========================================================================
// plan
// ===========================================================================
struct pf_plan
{
  pf_desc d;
  int dim = 0;
  long long c[3] = { 1, 1, 1 }, n[3] = { 1, 1, 1 };
  long long m0 = 1, m0p = 1;
  unsigned factor = 2;
  bool compact = false;   // DCT matrix backend semantics (iso / axiparallel)
  bool hermitian = false; // R2C field backend (dim > 1); 1D uses DFT backends
  long long rows = 0, P = 0; // table rows and row pitch
  long long Pg = 0;          // pitch of sig_gen (P rounded up to even)
  size_t tab_elems = 0;
  size_t D = 1, Next = 1;

  double* lam = nullptr;
  double* sig_gen = nullptr;
  double* sig = nullptr;
  double* inv = nullptr;
  std::map<long long, cplx*> tw; // twiddle tables per line length
  // arbitrary extended sizes (pf_generic.cuh): natural-order arrays, one
  // shared-memory transform per line (Bluestein for non powers of two)
  bool generic = false;
  struct GenTab
  {
    int M = 0;
    cplx *chirp = nullptr, *bhat = nullptr;
  };
  std::map<std::pair<long long, int>, GenTab> gen_tabs; // (n, dir)
  std::map<long long, cplx*> gen_tw;                    // per M
  cplx* GW = nullptr; // stored spectrum m0 * n1 * n2
  cplx* LO[4] = { nullptr, nullptr, nullptr, nullptr }; // operators on long lines: scratch
  // custom covariance (custom-iso / custom-aniso, matrix.hh:270-279): the torus
  // covariance evaluated by the caller, natural order [x2][x1][x0]
  double* cov_table = nullptr;
  // 2D plans with an axis longer than 1024 (pf_long.cuh): n[i] = LN1[i] * LN2[i]
  bool long2d = false;
  long long LN1[2] = { 1, 1 }, LN2[2] = { 1, 1 };
  long long mcols = 0;            // columns of Z: LN1[0] * (LN2[0]/2 + 1)
  int ly_batch = 1;               // realizations LY holds (generate_long2d)
  int long_batch = 1;             // realizations per generate_long2d call (batch API)
  std::map<long long, cplx*> twl; // [a][b] = exp(-2 pi i a b / n) per long line length
  cplx *LY = nullptr, *LU = nullptr; // step buffers of the long passes
  cplx* Y = nullptr;
  cplx* Z = nullptr;
  cplx* Zb = nullptr;      // batched 2D generate: Z of a whole batch
  size_t zb_elems = 0;
  cplx* V = nullptr;
  // streaming route (pf_stream.cuh): ring of pass-A planes (L2-resident),
  // ticket + per-plane completion counters
  cplx* ring = nullptr;
  unsigned* stream_sync = nullptr;
  int stream_slots = 0, stream_lag = 0; // 0: chosen from the problem size
  // layout of the pass-B output of the streaming route: 0 = Zt[k0][x2][x1]
  // (contiguous line stores, transposing pass C kp_c2rt), 1 = Z[x2][x1][k0]
  // (16-byte stores merged by L2, contiguous-line pass C kp_c2r); PF_ZT
  int stream_zt = 0;
  // one GPU: off by default -- measured on B200 at K3 (profiles/r2c): the fused
  // kernel takes 4.24 ms where passes A + B take 3.74 ms (one more barrier per
  // item and a larger instruction footprint outweigh the saved HBM traffic);
  // PF_STREAM=1 selects it.  It is the default of the multi-GPU route, where
  // it removes the exchange pass.
  bool stream_on = false;
  double* spare = nullptr;
  // 1D: pl->spare holds the imaginary part of transform (spare_seed, spare_t)
  bool spare_valid = false;
  uint64_t spare_seed = 0, spare_t = 0;
  int spare_rng = 0;
  double* dev_io[2] = { nullptr, nullptr }; // staging for host fields
  double* axis_tab = nullptr; // pf_field_localize: per-axis squared distances
  size_t axis_cap = 0;
  int* eval_flag = nullptr;   // pf_field_evaluate: "location outside" flag
  double* dev_in = nullptr;
  double* cg[6] = { nullptr, nullptr, nullptr, nullptr, nullptr, nullptr };
  double* red_partial = nullptr; // RED_BLOCKS + 8 doubles
  double* cg_scalars = nullptr;  // device scalars of innerCG (k_cg_*)
  double* cg_host = nullptr;     // pinned copies of (converged, count), 4 slots
  cudaEvent_t cg_ev[4] = {};
  unsigned long long* census = nullptr; // 4 x u64
  pf_eig_stats stats{};
  bool eig_valid = false;
  cudaStream_t stream = nullptr;
  cudaStream_t copy_stream = nullptr;
  cudaStream_t d2h_stream = nullptr; // slab mode, host output: field copies
  // slab exchange by DMA: one stream per destination so that several copy
  // engines drive NVLink at once (PF_SLAB_STREAMS limits the fan-out)
  cudaStream_t peer_stream[8] = {};
  cudaEvent_t ev_gate = nullptr, ev_peer[8] = {};
  int peer_streams = 0;
  cudaEvent_t ev_done[2] = { nullptr, nullptr }, ev_copied[2] = { nullptr, nullptr };
  unsigned long long launches = 0;
  // optional per-pass timing (CUDA events on the plan's stream)
  bool profiling = false;
  std::vector<std::pair<int, std::pair<cudaEvent_t, cudaEvent_t>>> prof_events;
  double prof_ms[16] = { 0 };
  unsigned long long prof_count[16] = { 0 };
  pf::CovParams cov{};
  int device = 0;
  int sm_count = 148;
  // slab mode, kernel exchange (peer stores): pass A of realization r+1 and
  // the tail (passes B, C) of realization r run CONCURRENTLY on disjoint sets
  // of SMs -- pass A is NVLink-bound there and leaves SMs idle -- by capping
  // the grids of the persistent kernels: sm_cap (0 = all SMs) applies to the
  // launches issued while it is set
  int sm_cap = 0;
  int split_a = 0;               // SMs given to pass A (0: no overlap)
  // one GPU, 3D, pf_generate_batch into device memory (experiment, off by
  // default): pass A of realization r+1 on split_one SMs, writing the other Y
  // buffer, beside the tail of realization r on the remaining SMs
  // (PF_SPLIT=<SMs>).  Measured on B200 at K3 (profiles/r02g): results
  // bit-identical, but SLOWER (206/s at 100 SMs against 233/s sequential):
  // passes B and C need all 148 SMs to reach their 6 TB/s -- their time scales
  // with 1/SMs like pass A's, so there is no idle resource to overlap into.
  int split_one = 0;
  cplx* Y2 = nullptr;            // second pass A output for that pipeline
  cplx* y_override = nullptr;    // generate_dev: pass A output / tail input
  int gen_phase = 0;             // generate_dev: 0 both, 1 pass A only, 2 tail only
  cudaStream_t tail_stream = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  // ---- slab mode (one process per GPU) -------------------------------------
  int nranks = 1, rank = 0;
  long long cl = 1;          // local layers of the last dimension
  size_t Dl = 1;             // cells of the local slab
  int a_begin = 0, a_count = 0; // pass A lines owned by this rank
  size_t xbuf_elems = 0;     // complex elements of one exchange buffer
  long long a_chunk = 0, a_total = 0; // rank q owns lines [q*a_chunk, ...) of a_total
  size_t fwd_elems = 0;      // forward-exchange region of the operators
  // streaming multi-GPU route (pf_stream.cuh): this rank computes the planes
  // [plane0, plane0 + planes) of k0 through BOTH first axes and stores pass
  // B's lines into the owners' Zt buffers; two buffers (parity of the epoch)
  // of m0 * cl * c1 complex numbers behind the forward region of the block
  bool stream_slab = false;
  int plane0 = 0, planes = 0;
  size_t zt_elems = 0;
  unsigned long long red_seq = 0; // cross-rank reductions issued so far
  unsigned char* slab_block = nullptr; // [2 exchange buffers][flags]
  void* peer_block[8] = { nullptr };
  unsigned long long** d_flagsA = nullptr; // device arrays of peer pointers
  unsigned long long** d_flagsB = nullptr;
  // sticky "a wait on a peer timed out" flag in pinned, device-mapped host
  // memory: the kernels set it, slab_timed_out() reads it after a stream
  // synchronisation without another copy
  int* h_timeout = nullptr;
  int* d_timeout = nullptr;
  unsigned long long slab_limit_ns = 120ull * 1000000000ull;
  bool connected = false;
  unsigned long long epoch = 0;
  // exchange by DMA (default): pass A writes the rows of other ranks into
  // local staging buffers with the geometry of the peers' buffers, the copy
  // engines move them over NVLink (contiguous row segments) while the next
  // realization's pass A runs.  PF_SLAB_EXCHANGE=peer selects direct stores
  // to peer memory from the kernel instead.
  bool slab_dma = true;
  cplx* send_stage[2] = { nullptr, nullptr }; // [(P-1) buffers] per parity
  cudaEvent_t ev_a[2] = { nullptr, nullptr }, ev_x[2] = { nullptr, nullptr };
  // bit mask of pass kinds routed to the persistent TMA kernels (N >= 256);
  // PF_PIPE=0 in the environment selects the plain kernels everywhere
  // (bit K_GEN_TILE, the 8-warp persistent pass A kp_gen, is off by default:
  // pass A is dispatch-bound and needs 12 FFT warps per SM; kp_gen remains
  // the slab-mode pass A, where it stores to peer memory)
  bool gen3_force = false; // tests: take kp_gen3 whenever it is applicable
  bool long_pipe = true;   // PF_LONG_PIPE=0: plain k_long_c2r1
  // bit K_GEN3: grouped persistent pass A (kp_gen3), 12 FFT warps per SM
  unsigned pipe_mask = (1u << pf::K_C2T_INV) | (1u << pf::K_C2T_FWD) |
                       (1u << pf::K_C2R) | (1u << pf::K_GEN3);
};

namespace {

// slab mode: call after the plan's stream has been synchronised
int
slab_check(const pf_plan* pl)
{
  if (pl->h_timeout && *(volatile int*)pl->h_timeout)
    return fail(PF_ERR_CUDA,
                "slab mode: timed out waiting for a peer rank (all ranks must "
                "issue the same calls within PF_SLAB_TIMEOUT_S seconds of each "
                "other); the plan is unusable");
  return 0;
}

template <class T>
int
dev_alloc(T** p, size_t count)
{
  if (*p)
    return 0;
  void* q = nullptr;
  cudaError_t e = cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T));
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(PF_ERR_NOMEM, std::string("cudaMalloc of ") +
                                std::to_string(count * sizeof(T)) +
                                " bytes failed: " + cudaGetErrorString(e));
  }
  *p = reinterpret_cast<T*>(q);
  return 0;
}
template <class T>
void
dev_free(T** p)
{
  if (*p)
    cudaFree(*p);
  *p = nullptr;
}

void
free_sized(pf_plan* pl)
{
  dev_free(&pl->lam);
  dev_free(&pl->sig_gen);
  dev_free(&pl->sig);
  dev_free(&pl->inv);
  dev_free(&pl->Y);
  dev_free(&pl->ring);
  dev_free(&pl->stream_sync);
  dev_free(&pl->Y2);
  dev_free(&pl->Z);
  dev_free(&pl->Zb);
  pl->zb_elems = 0;
  dev_free(&pl->V);
  dev_free(&pl->LY);
  dev_free(&pl->LU);
  dev_free(&pl->GW);
  for (auto& q : pl->LO)
    dev_free(&q);
  dev_free(&pl->spare);
  pl->spare_valid = false;
  pl->eig_valid = false;
}

bool
supported_n(long long n)
{
  return launcher_for(n) != nullptr;
}

// extended sizes this build has kernels for
// generic route: any even length (any length in 1D) whose shared-memory
// transform fits: powers of two up to 4096, otherwise 2 n - 1 <= 4096
bool
generic_n(const pf_plan* pl, long long n)
{
  if (pl->nranks > 1 || n < 2 || (pl->dim > 1 && n % 2 != 0))
    return false;
  if ((n & (n - 1)) == 0)
    return n <= 4096;
  return 2 * n - 1 <= 4096;
}

bool
size_supported(const pf_plan* pl, long long n)
{
  const bool long_ok = pl->dim == 2 && pl->nranks <= 1 && n > 1024 &&
                       n <= 16384 && (n & (n - 1)) == 0;
  return supported_n(n) || long_ok || generic_n(pl, n);
}

// sizes, fieldtraits.hh:214-311 (single device = single rank)
int
plan_resize(pf_plan* pl)
{
  free_sized(pl);
  const pf_desc& d = pl->d;
  pl->D = 1;
  pl->Next = 1;
  for (int i = 0; i < 3; i++) {
    pl->c[i] = pl->n[i] = 1;
  }
  for (int i = 0; i < pl->dim; i++) {
    pl->c[i] = d.cells[i];
    pl->n[i] = (long long)pl->factor * d.cells[i];
    if (!size_supported(pl, pl->n[i]))
      return fail(PF_ERR_INVALID,
                  "unsupported extended size " + std::to_string(pl->n[i]) +
                    " in dimension " + std::to_string(i) +
                    " (this build supports powers of two up to 1024 per axis "
                    "[16384 in 2D] on the specialised kernels, and on one GPU any "
                    "even size up to 2048 [any size in 1D, powers of two up to "
                    "4096] on the generic route)");
    pl->D *= pl->c[i];
    pl->Next *= pl->n[i];
  }
  pl->hermitian = pl->dim > 1;
  pl->compact = pl->dim > 1 && d.anisotropy != PF_ANISO_GEOMETRIC;
  pl->m0 = pl->hermitian ? pl->n[0] / 2 + 1 : pl->n[0];
  pl->m0p = (pl->m0 + 7) / 8 * 8;
  pl->long2d = false;
  // generic route: some axis has no specialised kernel (not a power of two, or
  // longer than the line kernels in 3D / mixed with one in 2D)
  pl->generic = false;
  for (int i = 0; i < pl->dim; i++) {
    const long long ni = pl->n[i];
    const bool long_ok = pl->dim == 2 && pl->nranks <= 1 && ni > 1024 &&
                         ni <= 16384 && (ni & (ni - 1)) == 0;
    if (!supported_n(ni) && !long_ok)
      pl->generic = true;
  }
  if (d.covariance == PF_COV_CUSTOM && (pl->dim == 1 || pl->n[0] > 1024 || pl->n[1] > 1024))
    pl->generic = true; // the table-reading eigenvalue passes exist for short 2D / 3D lines
  if (pl->generic)
    for (int i = 0; i < pl->dim; i++)
      if (!generic_n(pl, pl->n[i]))
        return fail(PF_ERR_INVALID,
                    "unsupported combination of extended sizes (generic route: "
                    "every axis even and <= 2048, or a power of two <= 4096)");
  if (!pl->generic && pl->dim == 2 && (pl->n[0] > 1024 || pl->n[1] > 1024)) {
    pl->long2d = true;
    for (int i = 0; i < 2; i++) {
      pl->LN2[i] = std::min<long long>(pl->n[i], 1024);
      pl->LN1[i] = pl->n[i] / pl->LN2[i];
    }
    // Z columns in the permuted order of the long c2r axis (pf_long.cuh)
    pl->mcols = pl->LN1[0] * (pl->LN2[0] / 2 + 1);
    pl->m0p = (pl->mcols + 7) / 8 * 8;
  }
  if (pl->dim == 3) {
    pl->rows = (pl->compact ? pl->n[1] / 2 + 1 : pl->n[1]) * pl->m0;
    pl->P = pl->compact ? pl->n[2] / 2 + 1 : pl->n[2];
  } else if (pl->dim == 2) {
    pl->rows = pl->m0;
    pl->P = pl->compact ? pl->n[1] / 2 + 1 : pl->n[1];
  } else {
    pl->rows = 1;
    pl->P = pl->n[0];
  }
  pl->tab_elems = (size_t)pl->rows * pl->P;
  if (pl->generic) // natural order [k2][k1][k0 < m0]
    pl->tab_elems = (size_t)pl->m0 * pl->n[1] * pl->n[2];
  pl->Pg = (pl->P + 1) & ~1LL;

  // covariance parameters (matrix.hh:624-648, covariance.hh:17-136)
  pf::CovParams& cv = pl->cov;
  memset(&cv, 0, sizeof(cv));
  cv.dim = pl->dim;
  cv.cov = d.covariance;
  cv.aniso = d.anisotropy;
  cv.variance = d.variance;
  cv.p0 = d.cov_param[0];
  cv.p1 = d.cov_param[1];
  if (d.covariance == PF_COV_MATERN) {
    // constants of MaternCovariance (covariance.hh:541-545) and the
    // Gamma-function terms of Temme's series for K_nu (pf_kernels.cuh)
    const double nu = d.cov_param[0];
    const double nl = std::floor(nu + 0.5), mu = nu - nl;
    const double gampl = 1. / std::tgamma(1. + mu);
    const double gammi = 1. / std::tgamma(1. - mu);
    double gam1;
    if (std::abs(mu) < 0.01) {
      // odd part of the Taylor series of 1/Gamma(1+z) (Abramowitz & Stegun 6.1.34)
      const double m2 = mu * mu;
      gam1 = -(0.5772156649015329 +
               m2 * (-0.0420026350340952 +
                     m2 * (-0.0421977345555443 + m2 * 0.0072189432466630)));
    } else
      gam1 = (gammi - gampl) / (2. * mu);
    cv.mt[0] = std::sqrt(2. * nu);
    cv.mt[1] = std::pow(2., 1. - nu) / std::tgamma(nu);
    cv.mt[2] = mu;
    cv.mt[3] = nl;
    cv.mt[4] = gampl;
    cv.mt[5] = gammi;
    cv.mt[6] = gam1;
    cv.mt[7] = 0.5 * (gammi + gampl);
  }
  for (int i = 0; i < pl->dim; i++) {
    cv.h[i] = d.extensions[i] / d.cells[i];
    cv.half[i] = 0.5 * d.extensions[i] * pl->factor;
    cv.period[i] = d.extensions[i] * pl->factor;
    cv.ext[i] = d.extensions[i];
  }
  // embedding.periodization = merge, matrix.hh:662-719
  cv.periodization = d.periodization;
  cv.sigmoid = d.sigmoid;
  cv.combine = d.sigmoid_combine;
  cv.factor = pl->factor;
  cv.s_start = d.sigmoid_start;
  cv.s_end = d.sigmoid_end < 0. ? pl->factor - 1. : d.sigmoid_end;
  cv.recursions = merge_recursions(d, pl->dim);
  const int dim = pl->dim;
  if (d.anisotropy == PF_ANISO_NONE) {
    cv.g[0] = 1. / d.corr_length[0];
  } else if (d.anisotropy == PF_ANISO_AXIPARALLEL) {
    for (int i = 0; i < dim; i++)
      cv.g[i] = 1. / d.corr_length[i];
  } else {
    // GeneralMatrix, covariance.hh:92-125: adjugate / determinant, same
    // operation order as the reference
    const double* copy = d.corr_length;
    double* m = cv.g;
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
      m[0] = 1. / copy[0];
    }
  }
  // slab decomposition (fieldtraits.hh:217-224,271-282: the last dimension
  // is split evenly over the ranks)
  pl->cl = pl->c[pl->dim - 1];
  pl->Dl = pl->D;
  pl->a_begin = 0;
  pl->a_count = 0;
  pl->stream_slab = false;
  pl->zt_elems = 0;
  if (pl->nranks > 1) {
    const int P = pl->nranks;
    if (pl->dim < 2)
      return fail(PF_ERR_INVALID, "slab mode needs dim >= 2");
    if (pl->c[pl->dim - 1] % P != 0)
      return fail(PF_ERR_INVALID, "number of cells in last dimension has to be "
                                  "multiple of numProc"); // fieldtraits.hh:219
    pl->cl = pl->c[pl->dim - 1] / P;
    pl->Dl = pl->D / P;
    // pass A lines (3D: k1, 2D: k0) in chunks aligned to the tile width
    const long long A = pl->dim == 3 ? pl->n[1] : pl->m0;
    const long long N = pl->dim == 3 ? pl->n[2] : pl->n[1];
    const long long lpc = std::max<long long>(8, std::min<long long>(128, 4096 / N));
    long long chunk = (A + P - 1) / P;
    chunk = (chunk + lpc - 1) / lpc * lpc;
    const long long b0 = std::min<long long>(A, chunk * pl->rank);
    const long long b1 = std::min<long long>(A, chunk * (pl->rank + 1));
    pl->a_begin = (int)b0;
    pl->a_count = (int)(b1 - b0);
    pl->a_chunk = chunk;
    pl->a_total = A;
    // operators: rank q receives its lines with the FULL last axis
    pl->fwd_elems = (size_t)std::min<long long>(A, chunk) *
                    (pl->dim == 3 ? pl->m0 : 1) * pl->c[pl->dim - 1];
    pl->xbuf_elems = pl->dim == 3 ? (size_t)pl->m0 * pl->cl * pl->n[1]
                                  : (size_t)pl->cl * pl->m0p;
    // streaming route: cubic tori of 512 / 1024, factor 2, whole B-items per
    // owner (PF_STREAM=0: exchange route)
    const long long N0 = pl->n[0];
    const int lpg = N0 == 1024 ? 4 : 8;
    bool want = pl->dim == 3 && pl->factor == 2 && (N0 == 512 || N0 == 1024) &&
                pl->n[1] == N0 && pl->n[2] == N0 && pl->cl % lpg == 0 &&
                pl->m0 >= P;
    if (const char* env = getenv("PF_STREAM"))
      want = want && atoi(env) != 0;
    pl->stream_slab = want;
    if (want) {
      const long long base = pl->m0 / P, extra = pl->m0 % P;
      pl->planes = (int)(base + (pl->rank < extra ? 1 : 0));
      pl->plane0 = (int)(base * pl->rank + std::min<long long>(pl->rank, extra));
      pl->zt_elems = (size_t)pl->m0 * pl->cl * pl->c[1];
    }
  }
  return 0;
}

// per-N inter-pass twiddle table [R2][32] = exp(-2 pi i r k / N)
int
twiddles(pf_plan* pl, long long N, const cplx** out)
{
  auto it = pl->tw.find(N);
  if (it != pl->tw.end()) {
    *out = it->second;
    return 0;
  }
  cplx* dptr = nullptr;
  const long long R2 = N >= 64 ? N / 32 : 0;
  std::vector<cplx> h((size_t)std::max<long long>(R2 * 32, 1));
  h[0] = make_double2(1., 0.);
  for (long long r = 0; r < R2; r++)
    for (long long k = 0; k < 32; k++) {
      const long double a =
        -2.0L * 3.14159265358979323846264338327950288L * (long double)(r * k) /
        (long double)N;
      h[r * 32 + k] = make_double2((double)cosl(a), (double)sinl(a));
    }
  PF_TRY(dev_alloc(&dptr, h.size()));
  PF_CUDA(cudaMemcpyAsync(dptr, h.data(), h.size() * sizeof(cplx),
                          cudaMemcpyHostToDevice, pl->stream));
  PF_CUDA(cudaStreamSynchronize(pl->stream));
  pl->tw[N] = dptr;
  *out = dptr;
  return 0;
}

// transposed-tile output of a pass as a 3D tensor map: dim0 = 2*A doubles
// (the tile axis), dim1 = element row e < keep, dim2 = b
int
make_tile_map_at(void* base, long long cols, long long rows, long long planes,
                 long long ose, long long osb, int boxr, CUtensorMap* tm,
                 int box_cols = 8)
{
  encode_tiled_fn enc = get_encode_tiled();
  if (!enc)
    return fail(PF_ERR_CUDA, "cuTensorMapEncodeTiled is not available");
  const cuuint64_t gdim[3] = { (cuuint64_t)2 * cols, (cuuint64_t)rows,
                               (cuuint64_t)planes };
  const cuuint64_t row = (cuuint64_t)ose * 16;
  const cuuint64_t plane =
    planes > 1 ? (cuuint64_t)osb * 16 : row * (cuuint64_t)rows;
  const cuuint64_t gstr[2] = { row, plane };
  // box_cols complex columns per row: 8 -> 128-byte rows, 4 -> 64-byte rows,
  // staged with the matching swizzle
  const cuuint32_t box[3] = { (cuuint32_t)(2 * box_cols), (cuuint32_t)boxr, 1 };
  const cuuint32_t estr[3] = { 1, 1, 1 };
  const CUresult r =
    enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, base, gdim, gstr, box, estr,
        CU_TENSOR_MAP_INTERLEAVE_NONE,
        box_cols == 4 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(PF_ERR_CUDA, "cuTensorMapEncodeTiled failed with code " +
                               std::to_string((int)r));
  return 0;
}

int
make_tile_map(const PassArgs& a, int boxr, CUtensorMap* tm)
{
  return make_tile_map_at(a.out, a.A, a.keep, a.B, a.ose, a.osb, boxr, tm);
}

struct TileMapsHost
{
  CUtensorMap m[8];
};
struct C2TMapsHost
{
  CUtensorMap out, in;
};

// blocked exchange buffer [plane][tile][row][8 complex] as a 4D tensor of
// doubles; box = {16, box_rows, box_tiles, 1}
int
make_blocked_map(void* base, long long rows, long long tiles, long long planes,
                 int box_rows, int box_tiles, bool swizzle, CUtensorMap* tm)
{
  encode_tiled_fn enc = get_encode_tiled();
  if (!enc)
    return fail(PF_ERR_CUDA, "cuTensorMapEncodeTiled is not available");
  const cuuint64_t gdim[4] = { 16, (cuuint64_t)rows, (cuuint64_t)tiles,
                               (cuuint64_t)planes };
  const cuuint64_t gstr[3] = { 128, (cuuint64_t)rows * 128,
                               (cuuint64_t)rows * tiles * 128 };
  const cuuint32_t box[4] = { 16, (cuuint32_t)box_rows, (cuuint32_t)box_tiles, 1 };
  const cuuint32_t estr[4] = { 1, 1, 1, 1 };
  const CUresult r = enc(
    tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 4, base, gdim, gstr, box, estr,
    CU_TENSOR_MAP_INTERLEAVE_NONE,
    swizzle ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
    CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(PF_ERR_CUDA, "cuTensorMapEncodeTiled (blocked) failed with code " +
                               std::to_string((int)r));
  return 0;
}

// rows per TMA box of the persistent tile kernels (PCfg<N>::BOXR)
int
pipe_box_rows(long long N)
{
  const int tpl = (int)(N / 32), lt = 8 * (32 / tpl);
  const int sr = 32768 / (lt * 16);
  return sr < 256 ? sr : 256;
}

bool
use_pipe(const pf_plan* pl, long long N, int kind, const PassArgs& a)
{
  if (N < 256 || !(pl->pipe_mask & (1u << kind)) || !pipe_launcher_for(N))
    return false;
  if (kind == pf::K_C2R)
    return a.B == 1 && (a.isa * 16) % 16 == 0;
  if (kind == pf::K_GEN_TILE && a.n_peers > 1) {
    // slab mode: rows per rank must be a whole number of tensor-store boxes
    const int rpp = a.rows_per_peer;
    if (rpp < 8 || (rpp & (rpp - 1)) != 0 || a.a_begin % 8 != 0)
      return false;
  }
  if (a.blk_T > 256)
    return false;
  // tile kernels: plain scale, 16-byte aligned lines, whole-line bulk copies
  if (a.scale != 1.0)
    return false;
  if (kind != pf::K_GEN_TILE && (a.n_valid <= 0 || a.n_valid > N))
    return false;
  return a.keep >= 1;
}

int
run_pass(pf_plan* pl, long long N, int kind, const PassArgs& a)
{
  launch_fn fn = launcher_for(N);
  if (!fn)
    return fail(PF_ERR_INVALID, "no kernel for line length " + std::to_string(N));
  const cplx* tw = nullptr;
  PF_TRY(twiddles(pl, N, &tw));
  // slab mode pass A always goes through the persistent kernel when it can:
  // its tensor stores write 128-byte rows to peer memory
  const bool slab_gen =
    kind == pf::K_GEN_TILE && a.n_peers > 1 && !pl->slab_dma;
  bool pipe = use_pipe(pl, N, kind, a);
  if (slab_gen && !pipe && N >= 256 && pipe_launcher_for(N) &&
      !getenv("PF_SLAB_PLAIN")) {
    const unsigned saved = pl->pipe_mask;
    pl->pipe_mask |= 1u << pf::K_GEN_TILE;
    pipe = use_pipe(pl, N, kind, a);
    pl->pipe_mask = saved;
  }
  PassArgs a2 = a;
  TileMapsHost tms;
  C2TMapsHost c2t;
  const void* tmptr = &c2t;
  int grid = 0;
  int launch_kind = kind;
  // grouped persistent pass A: single rank, lines of 512 / 1024, at most
  // half of the line kept, sigma rows 16-byte aligned
  // (measured on B200: ahead of the per-tile kernel only for N = 1024 and
  // grids of many waves; K2's N = 512 and small 2D problems stay plain)
  const bool gen3_size =
    pl->gen3_force ? (N == 512 || N == 1024)
                   : (N == 1024 &&
                      (long long)((a.A + 3) / 4) * a.B >= 8LL * pl->sm_count);
  const bool gen3 =
    kind == pf::K_GEN_TILE && a.n_peers <= 1 && gen3_size &&
    (pl->pipe_mask & (1u << pf::K_GEN3)) && !(pl->pipe_mask & 1u) &&
    a.keep >= 1 && a.keep <= N / 2 && a.scale == 1.0 && a.a_begin == 0 &&
    a.tab_pitch % 2 == 0 && (reinterpret_cast<uintptr_t>(a.tab) & 15) == 0 &&
    (long long)(a.tab_mirror_e ? N / 2 + 2 : N) <= a.tab_pitch &&
    pipe_launcher_for(N);
  if (gen3) {
    const int lpc = N == 1024 ? 4 : 8;
    const long long ntiles = (long long)((a.A + lpc - 1) / lpc) * a.B;
    const int boxr = (int)std::min<long long>(256, N / 2);
    PF_TRY(make_tile_map_at(a.out, a.A, a.keep, a.B, a.ose, a.osb, boxr,
                            &tms.m[0], lpc));
    tmptr = &tms;
    grid = (int)std::min<long long>((ntiles + 2) / 3,
                                    pl->sm_cap > 0 ? pl->sm_cap : pl->sm_count);
    launch_kind = pf::K_GEN3;
    pipe = true;
  } else if (pipe) {
    const int tpl = (int)(N / 32), lt = 8 * (32 / tpl);
    long long ntiles;
    if (kind == pf::K_C2R) {
      const long long nl = (long long)a.A * a.B;
      ntiles = ((nl + 1) / 2 + lt - 1) / lt;
    } else if (kind == pf::K_GEN_TILE) {
      ntiles = (long long)((a.A + lt - 1) / lt) * a.B;
      int boxr = pipe_box_rows(N);
      if (a.n_peers > 1) {
        boxr = std::min(boxr, a.rows_per_peer);
        // global column extent: the peers' buffers hold all lines
        const long long cols = a.ose; // row pitch = number of columns
        for (int q = 0; q < a.n_peers; q++) {
          if (a.blk_T > 0)
            PF_TRY(make_blocked_map(a.peer_out[q], a.blk_rows, a.blk_T, a.B,
                                    boxr, 1, true, &tms.m[q]));
          else
            PF_TRY(make_tile_map_at(a.peer_out[q], cols, a.rows_per_peer, a.B,
                                    a.ose, a.osb, boxr, &tms.m[q]));
        }
      } else {
        PF_TRY(make_tile_map(a, boxr, &tms.m[0]));
      }
      a2.box_rows = boxr;
      tmptr = &tms;
    } else {
      ntiles = (long long)((a.A + lt - 1) / lt) * a.B;
      PF_TRY(make_tile_map(a, pipe_box_rows(N), &c2t.out));
      c2t.in = c2t.out;
      if (a.blk_T > 0)
        PF_TRY(make_blocked_map(const_cast<void*>(a.in), a.blk_rows, a.blk_T,
                                a.A, 1, a.blk_T, false, &c2t.in));
      tmptr = &c2t;
    }
    grid = (int)std::min<long long>(ntiles, pl->sm_cap > 0 ? pl->sm_cap : pl->sm_count);
  }
  // long c2r axis, step 1: persistent pair kernel (pf_pipe.cuh)
  if (kind == pf::K_LONG_C2R1 && N == 1024 && pl->long_pipe && pipe_launcher_for(N)) {
    const long long items = (long long)(a.long_n1 / 2 + 1) * a.B;
    pipe = true;
    launch_kind = kind;
    grid = (int)std::min<long long>((items + 5) / 6,
                                    pl->sm_cap > 0 ? pl->sm_cap : pl->sm_count);
  }
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (pl->profiling) {
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, pl->stream);
  }
  const int e = pipe ? pipe_launcher_for(N)(launch_kind, &a2, tw, tmptr, grid, pl->stream)
                     : fn(kind, &a, tw, pl->stream);
  if (pl->profiling) {
    cudaEventRecord(e1, pl->stream);
    pl->prof_events.push_back({ kind, { e0, e1 } });
  }
  if (e != 0)
    return fail(PF_ERR_CUDA, std::string("kernel launch failed: ") +
                               cudaGetErrorString((cudaError_t)e));
  pl->launches++;
  return 0;
}

void
prof_collect(pf_plan* pl)
{
  for (auto& pe : pl->prof_events) {
    cudaEventSynchronize(pe.second.second);
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, pe.second.first, pe.second.second) ==
        cudaSuccess) {
      pl->prof_ms[pe.first & 15] += ms;
      pl->prof_count[pe.first & 15]++;
    }
    cudaEventDestroy(pe.second.first);
    cudaEventDestroy(pe.second.second);
  }
  pl->prof_events.clear();
}

PassArgs
blank_args()
{
  PassArgs a;
  memset(&a, 0, sizeof(a));
  a.scale = 1.0;
  a.na_self = a.nb_self = 1;
  a.cov_axis_a = a.cov_axis_b = -1;
  return a;
}

// ---------------------------------------------------------------------------
// long lines (2D, n up to 16384): host side of pf_long.cuh
// ---------------------------------------------------------------------------
int
long_twiddles(pf_plan* pl, long long n, const cplx** out)
{
  auto it = pl->twl.find(n);
  if (it != pl->twl.end()) {
    *out = it->second;
    return 0;
  }
  cplx* d = nullptr;
  PF_TRY(dev_alloc(&d, (size_t)n));
  pf::k_long_twiddles<<<(unsigned)((n + 255) / 256), 256, 0, pl->stream>>>(
    d, n, (int)std::min<long long>(n, 1024));
  PF_CUDA(cudaGetLastError());
  pl->launches++;
  pl->twl[n] = d;
  *out = d;
  return 0;
}

template <int SRC, int DIR, bool BATCH = false>
int
launch_long1(pf_plan* pl, int N1, const pf::LongArgs& g, const PassArgs& p,
             long long lines)
{
  const dim3 grid((unsigned)((g.N2 + 255) / 256), (unsigned)lines);
  switch (N1) {
    case 1:
      pf::k_long1<1, SRC, DIR, BATCH><<<grid, 256, 0, pl->stream>>>(g, p);
      break;
    case 2:
      pf::k_long1<2, SRC, DIR, BATCH><<<grid, 256, 0, pl->stream>>>(g, p);
      break;
    case 4:
      pf::k_long1<4, SRC, DIR, BATCH><<<grid, 256, 0, pl->stream>>>(g, p);
      break;
    case 8:
      pf::k_long1<8, SRC, DIR, BATCH><<<grid, 256, 0, pl->stream>>>(g, p);
      break;
    case 16:
      pf::k_long1<16, SRC, DIR, BATCH><<<grid, 256, 0, pl->stream>>>(g, p);
      break;
    default:
      return fail(PF_ERR_INVALID, "long line: unsupported split");
  }
  PF_CUDA(cudaGetLastError());
  pl->launches++;
  return 0;
}

// forward FFT of `lines` contiguous lines of n = N1*N2 complex numbers:
// step 1 (covariance evaluation or complex input) into Y, step 2 = k_c2t with
// (a = j1, b = line), whose tile store interleaves the sub-lines back into
// natural order: out[line*out_pitch + j1 + N1*e], e < keep
// dir = +1: the backward transform (conjugate twiddles, kp_c2t<+1>);
// in_valid / in_real: zero padding and real input of step 1 (LongArgs)
int
long_fft_forward(pf_plan* pl, int axis, bool from_cov, const cplx* in,
                 long long in_pitch, long long lines, cplx* Y, cplx* out,
                 long long out_pitch, long long keep, int dir = -1,
                 long long in_valid = 0, int in_real = 0, long long herm_n = 0)
{
  const long long n = pl->n[axis];
  const int N1 = (int)pl->LN1[axis], N2 = (int)pl->LN2[axis];
  pf::LongArgs g;
  memset(&g, 0, sizeof(g));
  g.N2 = N2;
  g.n = n;
  PF_TRY(long_twiddles(pl, n, &g.tw));
  g.in = in;
  g.in_pitch = in_pitch;
  g.in_valid = in_valid;
  g.in_real = in_real;
  g.herm_n = herm_n;
  g.out = Y;
  PassArgs p = blank_args();
  p.cov = pl->cov;
  if (from_cov)
    PF_TRY((launch_long1<pf::LSRC_COV, -1>(pl, N1, g, p, lines)));
  else if (dir > 0)
    PF_TRY((launch_long1<pf::LSRC_CPLX, +1>(pl, N1, g, p, lines)));
  else
    PF_TRY((launch_long1<pf::LSRC_CPLX, -1>(pl, N1, g, p, lines)));
  PassArgs a = blank_args();
  a.A = N1;
  a.B = (int)lines;
  a.in = Y;
  a.isa = N2;
  a.isb = n;
  a.n_valid = N2;
  a.out = out;
  a.osb = out_pitch;
  a.ose = N1;
  a.keep = (int)keep;
  return run_pass(pl, N2, dir > 0 ? pf::K_C2T_INV : pf::K_C2T_FWD, a);
}

// multiplyExtended / multiplyRootExtended / multiplyInverseExtended
// (matrix.hh:1459-1662) for 2D plans with long lines: full complex transforms
// of the zero-padded field with the two-step primitive, rows first, transposed,
// rows again; g(lambda) / N in between; the same backwards, pruned to the
// physical sub-box.
int
apply_long2d(pf_plan* pl, const double* tab, const double* in_dev, double* out_dev)
{
  const long long* n = pl->n;
  const long long* c = pl->c;
  const long long m0 = pl->m0;
  const long long N1a = pl->LN1[0], N2a = pl->LN2[0]; // axis 0: n0 = N1a * N2a
  const long long N1b = pl->LN1[1], N2b = pl->LN2[1];
  // the field is real: after the transform along axis 0 only k0 <= n0 / 2 is
  // kept (sub-line outputs e <= N2a / 2, i.e. w0 >= m0 columns) and carried
  // through axis 1, the multiplication and back; the last transform rebuilds
  // the other half as conjugates while it loads (LongArgs::herm_n)
  const long long h0 = N2a / 2 + 1, w0 = N1a * h0;
  const long long keep0 = (c[0] + N1a - 1) / N1a, p0 = N1a * keep0; // >= c0
  const long long keep1 = (c[1] + N1b - 1) / N1b, p1 = N1b * keep1; // >= c1
  // scratch kept with the plan (7 GB at K4's size: a PCG would otherwise pay
  // the allocations on every product)
  cplx *&Y = pl->LO[0], *&A = pl->LO[1], *&At = pl->LO[2], *&S = pl->LO[3];
  auto transpose = [&](const cplx* in, long long ipitch, cplx* out, long long opitch,
                       long long rows, long long cols) -> int {
    const dim3 grid((unsigned)((cols + 31) / 32), (unsigned)((rows + 31) / 32));
    pf::k_long_transpose<<<grid, dim3(32, 8), 0, pl->stream>>>(in, ipitch, out, opitch,
                                                             rows, cols);
    pl->launches++;
    return cudaGetLastError() == cudaSuccess ? 0 : fail(PF_ERR_CUDA, "transpose");
  };
  int rc = dev_alloc(&Y, (size_t)std::max(c[1] * n[0], w0 * n[1]));
  if (rc == 0)
    rc = dev_alloc(&A, (size_t)c[1] * w0);
  if (rc == 0)
    rc = dev_alloc(&At, (size_t)w0 * std::max(c[1], p1));
  if (rc == 0)
    rc = dev_alloc(&S, (size_t)std::max(w0 * n[1], c[1] * p0));
  // forward: rows x1 < c1 along axis 0 (real input, zero padded beyond c0),
  // half spectrum kept
  if (rc == 0)
    rc = long_fft_forward(pl, 0, false, reinterpret_cast<const cplx*>(in_dev), c[0],
                          c[1], Y, A, w0, h0, -1, c[0], 1);
  if (rc == 0)
    rc = transpose(A, w0, At, c[1], c[1], w0); // At[k0 < w0][x1 < c1]
  // rows k0 <= n0 / 2 along axis 1 (zero padded beyond c1)
  if (rc == 0)
    rc = long_fft_forward(pl, 1, false, At, c[1], m0, Y, S, n[1], N2b, -1, c[1], 0);
  if (rc == 0) {
    const long long total = m0 * n[1];
    pf::k_long_spec<<<(unsigned)((total + 255) / 256), 256, 0, pl->stream>>>(
      S, n[0], n[1], tab, m0, pl->P, pl->compact ? 1 : 0, 1.0 / (double)pl->Next);
    pl->launches++;
    rc = cudaGetLastError() == cudaSuccess ? 0 : fail(PF_ERR_CUDA, "spectral multiply");
  }
  // backward along axis 1, kept x1 < p1 (>= c1)
  if (rc == 0)
    rc = long_fft_forward(pl, 1, false, S, n[1], m0, Y, At, p1, keep1, +1);
  if (rc == 0)
    rc = transpose(At, p1, A, w0, m0, c[1]); // A[x1 < c1][k0 < m0]
  // backward along axis 0 from the half spectrum, kept x0 < p0
  if (rc == 0)
    rc = long_fft_forward(pl, 0, false, A, w0, c[1], Y, S, p0, keep0, +1, 0, 0, n[0]);
  if (rc == 0) {
    const long long total = c[1] * c[0];
    pf::k_long_real<<<(unsigned)((total + 255) / 256), 256, 0, pl->stream>>>(
      S, p0, out_dev, c[1], c[0]);
    pl->launches++;
    rc = cudaGetLastError() == cudaSuccess ? 0 : fail(PF_ERR_CUDA, "real part");
  }
  return rc;
}

// eigenvalues of a 2D plan with long lines: full complex transforms (one-time
// cost), axis 0 first, transpose, axis 1, then real part / N with the census
int
eigen_long2d(pf_plan* pl)
{
  const long long* n = pl->n;
  const long long m0 = pl->m0;
  const double invN = 1.0 / (double)pl->Next;
  // k0 <= n0/2 needs sub-line outputs e <= N2/2 (a few columns more than m0)
  const long long keep0 = pl->LN2[0] / 2 + 1;
  const long long w0 = pl->LN1[0] * keep0; // >= m0 columns per row
  cplx *Y = nullptr, *C1 = nullptr, *C1T = nullptr, *C2 = nullptr;
  const unsigned saved_mask = pl->pipe_mask;
  pl->pipe_mask = 0; // one-time work: plain kernels
  int rc = dev_alloc(&Y, (size_t)n[0] * n[1]);
  if (rc == 0)
    rc = dev_alloc(&C1, (size_t)n[1] * w0);
  if (rc == 0) // axis 0: lines x1, covariance evaluated in step 1
    rc = long_fft_forward(pl, 0, true, nullptr, 0, n[1], Y, C1, w0, keep0);
  if (rc == 0)
    rc = dev_alloc(&C1T, (size_t)m0 * n[1]);
  if (rc == 0) {
    const dim3 grid((unsigned)((m0 + 31) / 32), (unsigned)((n[1] + 31) / 32));
    pf::k_long_transpose<<<grid, dim3(32, 8), 0, pl->stream>>>(C1, w0, C1T, n[1],
                                                             n[1], m0);
    rc = cudaGetLastError() == cudaSuccess ? 0 : fail(PF_ERR_CUDA, "transpose");
    pl->launches++;
  }
  if (rc == 0)
    rc = dev_alloc(&C2, (size_t)m0 * n[1]);
  if (rc == 0) // axis 1: lines k0
    rc = long_fft_forward(pl, 1, false, C1T, n[1], m0, Y, C2, n[1], pl->LN2[1]);
  if (rc == 0) {
    const long long total = m0 * pl->P;
    pf::k_long_lambda<<<(unsigned)((total + 255) / 256), 256, 0, pl->stream>>>(
      C2, n[1], pl->lam, m0, pl->P, invN, pl->d.threshold, pl->census,
      pl->census + 3);
    rc = cudaGetLastError() == cudaSuccess ? 0 : fail(PF_ERR_CUDA, "lambda");
    pl->launches++;
  }
  pl->pipe_mask = saved_mask;
  cudaStreamSynchronize(pl->stream);
  cudaFree(Y);
  cudaFree(C1);
  cudaFree(C1T);
  cudaFree(C2);
  return rc;
}

// ---------------------------------------------------------------------------
// generic route (pf_generic.cuh): host side
// ---------------------------------------------------------------------------
// tables of one (length, direction): the shared-memory transform length M, the
// twiddles of M, and for Bluestein the chirp and the spectrum of its conjugate
int
gen_tables(pf_plan* pl, long long n, int dir, pf::GenLineArgs& g)
{
  const bool pow2 = (n & (n - 1)) == 0;
  int M = (int)n;
  if (!pow2) {
    M = 1;
    while (M < 2 * n - 1)
      M <<= 1;
  }
  g.n = (int)n;
  g.M = M;
  g.dir = dir;
  auto tw = pl->gen_tw.find(M);
  if (tw == pl->gen_tw.end()) {
    std::vector<cplx> h((size_t)std::max(M / 2, 1));
    for (int m = 0; m < M / 2; m++) {
      const long double a = -2.0L * 3.14159265358979323846264338327950288L * m / M;
      h[m] = make_double2((double)cosl(a), (double)sinl(a));
    }
    cplx* d = nullptr;
    PF_TRY(dev_alloc(&d, h.size()));
    PF_CUDA(cudaMemcpy(d, h.data(), h.size() * sizeof(cplx), cudaMemcpyHostToDevice));
    tw = pl->gen_tw.emplace((long long)M, d).first;
  }
  g.twM = tw->second;
  g.chirp = nullptr;
  g.bhat = nullptr;
  if (pow2)
    return 0;
  auto key = std::make_pair(n, dir);
  auto it = pl->gen_tabs.find(key);
  if (it == pl->gen_tabs.end()) {
    // w_j = exp(dir pi i j^2 / n), j^2 reduced modulo 2n for accuracy
    std::vector<cplx> w((size_t)n);
    std::vector<std::complex<long double>> b((size_t)M, { 0, 0 });
    for (long long j = 0; j < n; j++) {
      const long long q = (j * j) % (2 * n);
      const long double a = dir * 3.14159265358979323846264338327950288L * q / n;
      w[j] = make_double2((double)cosl(a), (double)sinl(a));
      const std::complex<long double> c(cosl(a), -sinl(a)); // conj(w_j)
      b[j] = c;
      if (j > 0)
        b[M - j] = c;
    }
    // forward FFT of b on the host (long double, recursive radix 2)
    std::vector<std::complex<long double>> tmp(b.size());
    for (int len = 1; len < M; len <<= 1) {
      const int T = M / 2, tstep = T / len;
      for (int j = 0; j < T; j++) {
        const int k = j & (len - 1);
        const long double a = -2.0L * 3.14159265358979323846264338327950288L * (k * tstep) / M;
        const std::complex<long double> tww(cosl(a), sinl(a));
        const auto u0 = b[j], u1 = b[j + T] * tww;
        const int j0 = ((j - k) << 1) + k;
        tmp[j0] = u0 + u1;
        tmp[j0 + len] = u0 - u1;
      }
      b.swap(tmp);
    }
    std::vector<cplx> bh((size_t)M);
    for (int j = 0; j < M; j++)
      bh[j] = make_double2((double)b[j].real(), (double)b[j].imag());
    pf_plan::GenTab t;
    t.M = M;
    PF_TRY(dev_alloc(&t.chirp, w.size()));
    PF_TRY(dev_alloc(&t.bhat, bh.size()));
    PF_CUDA(cudaMemcpy(t.chirp, w.data(), w.size() * sizeof(cplx), cudaMemcpyHostToDevice));
    PF_CUDA(cudaMemcpy(t.bhat, bh.data(), bh.size() * sizeof(cplx), cudaMemcpyHostToDevice));
    it = pl->gen_tabs.emplace(key, t).first;
  }
  g.chirp = it->second.chirp;
  g.bhat = it->second.bhat;
  return 0;
}

// one batch of lines through k_generic_lines; `g` carries addressing and modes
int
gen_lines(pf_plan* pl, long long n, int dir, pf::GenLineArgs g)
{
  PF_TRY(gen_tables(pl, n, dir, g));
  const long long lines = g.L0 * g.L1;
  if (lines <= 0)
    return 0;
  if (lines > 0x7fffffffLL)
    return fail(PF_ERR_INVALID, "generic route: too many lines");
  const size_t smem = 2 * (size_t)g.M * sizeof(cplx);
  static bool attr_set = false;
  if (!attr_set) {
    PF_CUDA(cudaFuncSetAttribute(pf::k_generic_lines,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 2 * 4096 * (int)sizeof(cplx)));
    attr_set = true;
  }
  pf::k_generic_lines<<<(unsigned)lines, 256, smem, pl->stream>>>(g);
  PF_CUDA(cudaGetLastError());
  pl->launches++;
  return 0;
}

pf::GenGeom
gen_geom(const pf_plan* pl)
{
  pf::GenGeom g;
  g.dim = pl->dim;
  for (int i = 0; i < 3; i++)
    g.n[i] = pl->n[i];
  g.m0 = pl->m0;
  g.half = (pl->n[pl->dim - 1] + 1) / 2;
  return g;
}

pf::GenLineArgs
gen_blank()
{
  pf::GenLineArgs g;
  memset(&g, 0, sizeof(g));
  g.L0 = g.L1 = 1;
  g.scale = 1.0;
  return g;
}

// inverse transform of the stored spectrum W[k2][k1][k0 < m0] into a field
// (backwardTransform + extendedFieldToField, r2cfieldbackend.hh:250-274,423-441;
// 1D: dftfieldbackend.hh:223-247,380-393), pruned axis by axis
int
generic_inverse(pf_plan* pl, cplx* W, double* out_dev, double* spare_dev,
                int transform)
{
  const long long* n = pl->n;
  const long long* c = pl->c;
  const long long m0 = pl->m0;
  if (pl->dim == 1) {
    pf::GenLineArgs g = gen_blank();
    g.in_mode = 0;
    g.in = W;
    g.in_se = 1;
    g.n_valid = (int)n[0];
    g.out_mode = 2;
    g.out = out_dev;
    g.out2 = spare_dev;
    g.out_se = 1;
    g.keep = (int)c[0];
    g.transform = transform;
    g.tparam = pl->d.boxcox_lambda;
    return gen_lines(pl, n[0], +1, g);
  }
  if (pl->dim == 3) { // axis 2: lines (k0, k1), keep x2 < c2
    pf::GenLineArgs g = gen_blank();
    g.L0 = m0;
    g.L1 = n[1];
    g.in = g.out = W;
    g.in_s0 = g.out_s0 = 1;
    g.in_s1 = g.out_s1 = m0;
    g.in_se = g.out_se = m0 * n[1];
    g.n_valid = (int)n[2];
    g.keep = (int)c[2];
    PF_TRY(gen_lines(pl, n[2], +1, g));
  }
  { // axis 1: lines (k0, x2 < c2), keep x1 < c1
    pf::GenLineArgs g = gen_blank();
    g.L0 = m0;
    g.L1 = pl->dim == 3 ? c[2] : 1;
    g.in = g.out = W;
    g.in_s0 = g.out_s0 = 1;
    g.in_s1 = g.out_s1 = m0 * n[1];
    g.in_se = g.out_se = m0;
    g.n_valid = (int)n[1];
    g.keep = (int)c[1];
    PF_TRY(gen_lines(pl, n[1], +1, g));
  }
  { // axis 0, c2r: lines (x1 < c1, x2 < c2) -> real field, keep x0 < c0
    pf::GenLineArgs g = gen_blank();
    g.L0 = c[1];
    g.L1 = pl->dim == 3 ? c[2] : 1;
    g.in_mode = 2;
    g.in = W;
    g.in_s0 = m0;
    g.in_s1 = m0 * n[1];
    g.in_se = 1;
    g.out_mode = 1;
    g.out = out_dev;
    g.out_s0 = c[0];
    g.out_s1 = c[0] * c[1];
    g.out_se = 1;
    g.keep = (int)c[0];
    g.transform = transform;
    g.tparam = pl->d.boxcox_lambda;
    PF_TRY(gen_lines(pl, n[0], +1, g));
  }
  return 0;
}

// Matrix::fillTransformedMatrix on the generic route: full complex transforms
// of the torus covariance (one-time cost), table in natural order
int
eigen_generic(pf_plan* pl)
{
  const long long* n = pl->n;
  const long long N = (long long)pl->Next;
  cplx* C = nullptr;
  PF_TRY(dev_alloc(&C, (size_t)N));
  const pf::GenGeom gg = gen_geom(pl);
  if (pl->cov_table)
    pf::k_generic_cov_table<<<grid_for((size_t)N), 256, 0, pl->stream>>>(pl->cov_table, C, N);
  else
    pf::k_generic_cov<<<grid_for((size_t)N), 256, 0, pl->stream>>>(gg, pl->cov, C, N);
  pl->launches++;
  int rc = cudaGetLastError() == cudaSuccess ? 0 : fail(PF_ERR_CUDA, "k_generic_cov");
  const long long str[3] = { 1, n[0], n[0] * n[1] };
  for (int ax = 0; ax < pl->dim && rc == 0; ax++) {
    pf::GenLineArgs g = gen_blank();
    const int o0 = ax == 0 ? 1 : 0, o1 = ax == 2 ? 1 : 2;
    g.L0 = n[o0];
    g.L1 = n[o1];
    g.in = g.out = C;
    g.in_s0 = g.out_s0 = str[o0];
    g.in_s1 = g.out_s1 = str[o1];
    g.in_se = g.out_se = str[ax];
    g.n_valid = g.keep = (int)n[ax];
    rc = gen_lines(pl, n[ax], -1, g);
  }
  if (rc == 0) {
    const long long total = (long long)pl->tab_elems;
    pf::k_generic_lambda<<<grid_for((size_t)total), 256, 0, pl->stream>>>(
      gg, C, pl->lam, 1.0 / (double)pl->Next, pl->d.threshold, pl->compact ? 1 : 0,
      pl->census, pl->census + 3, total);
    pl->launches++;
    rc = cudaGetLastError() == cudaSuccess ? 0 : fail(PF_ERR_CUDA, "k_generic_lambda");
  }
  cudaStreamSynchronize(pl->stream);
  cudaFree(C);
  return rc;
}

// ---------------------------------------------------------------------------
// eigenvalues: Matrix::fillTransformedMatrix, matrix.hh:232-331
// ---------------------------------------------------------------------------
int
eigen_once(pf_plan* pl)
{
  const long long* n = pl->n;
  const long long m0 = pl->m0;
  PF_TRY(dev_alloc(&pl->lam, pl->tab_elems));
  PF_TRY(dev_alloc(&pl->census, 4));
  {
    unsigned long long init[4] = { 0, 0, 0, 0xFFFFFFFFFFFFFFFFull };
    PF_CUDA(cudaMemcpyAsync(pl->census, init, sizeof(init),
                            cudaMemcpyHostToDevice, pl->stream));
  }
  const double invN = 1.0 / (double)pl->Next;
  if (pl->d.covariance == PF_COV_CUSTOM && !pl->cov_table)
    return fail(PF_ERR_INVALID, // matrix.hh:236-239
                "you need to call fillMatrix with your covariance class as parameter");
  if (pl->generic) {
    PF_TRY(eigen_generic(pl));
  } else if (pl->dim == 1) {
    PassArgs a = blank_args();
    a.A = 1;
    a.B = 1;
    a.cov = pl->cov;
    a.out = pl->lam;
    a.out_pitch = n[0];
    a.keep = (int)n[0];
    a.scale = invN;
    a.census = pl->census;
    a.smallest_bits = pl->census + 3;
    a.threshold = pl->d.threshold;
    PF_TRY(run_pass(pl, n[0], pf::K_LAMBDA_COV, a));
  } else if (pl->dim == 2 && pl->long2d) {
    PF_TRY(eigen_long2d(pl));
  } else if (pl->dim == 2) {
    cplx* U = nullptr;
    PF_TRY(dev_alloc(&U, (size_t)m0 * n[1]));
    PassArgs a = blank_args();
    a.A = (int)n[1];
    a.B = 1;
    a.cov = pl->cov;
    a.cov_axis_a = 1;
    a.out = U;
    a.osb = 0;
    a.ose = n[1];
    a.keep = (int)m0;
    if (pl->cov_table) { // custom covariance: rows of the uploaded torus table
      a.in = pl->cov_table;
      a.isa = n[0];
      a.n_valid = (int)n[0];
    }
    int rc = run_pass(pl, n[0], pl->cov_table ? pf::K_R2C_LOAD : pf::K_R2C_COV, a);
    if (rc == 0) {
      PassArgs e = blank_args();
      e.A = (int)m0;
      e.B = 1;
      e.in = U;
      e.isa = n[1];
      e.n_valid = (int)n[1];
      e.out = pl->lam;
      e.out_pitch = pl->P;
      e.keep = (int)pl->P;
      e.scale = invN;
      e.census = pl->census;
      e.smallest_bits = pl->census + 3;
      e.threshold = pl->d.threshold;
      rc = run_pass(pl, n[1], pf::K_LAMBDA, e);
    }
    cudaStreamSynchronize(pl->stream);
    cudaFree(U);
    PF_TRY(rc);
  } else {
    const long long R1 = pl->compact ? n[1] / 2 + 1 : n[1];
    cplx *U = nullptr, *V = nullptr;
    PF_TRY(dev_alloc(&U, (size_t)n[2] * m0 * n[1]));
    int rc = dev_alloc(&V, (size_t)R1 * m0 * n[2]);
    if (rc == 0) {
      PassArgs a = blank_args(); // E0: x0 lines, (a,b) = (x1,x2)
      a.A = (int)n[1];
      a.B = (int)n[2];
      a.cov = pl->cov;
      a.cov_axis_a = 1;
      a.cov_axis_b = 2;
      a.out = U; // U[x2][k0][x1]
      a.osb = m0 * n[1];
      a.ose = n[1];
      a.keep = (int)m0;
      if (pl->cov_table) { // custom covariance: rows of the uploaded torus table
        a.in = pl->cov_table;
        a.isa = n[0];
        a.isb = n[0] * n[1];
        a.n_valid = (int)n[0];
      }
      rc = run_pass(pl, n[0], pl->cov_table ? pf::K_R2C_LOAD : pf::K_R2C_COV, a);
    }
    if (rc == 0) {
      PassArgs a = blank_args(); // E1: x1 lines, (a,b) = (x2,k0)
      a.A = (int)n[2];
      a.B = (int)m0;
      a.in = U;
      a.isa = m0 * n[1];
      a.isb = n[1];
      a.n_valid = (int)n[1];
      a.out = V; // V[k1][k0][x2]
      a.osb = n[2];
      a.ose = m0 * n[2];
      a.keep = (int)R1;
      rc = run_pass(pl, n[1], pf::K_C2T_FWD, a);
    }
    if (rc == 0) {
      PassArgs e = blank_args(); // E2: x2 lines
      e.A = (int)(R1 * m0);
      e.B = 1;
      e.in = V;
      e.isa = n[2];
      e.n_valid = (int)n[2];
      e.out = pl->lam;
      e.out_pitch = pl->P;
      e.keep = (int)pl->P;
      e.scale = invN;
      e.census = pl->census;
      e.smallest_bits = pl->census + 3;
      e.threshold = pl->d.threshold;
      rc = run_pass(pl, n[2], pf::K_LAMBDA, e);
    }
    cudaStreamSynchronize(pl->stream);
    cudaFree(U);
    cudaFree(V);
    PF_TRY(rc);
  }
  unsigned long long res[4];
  PF_CUDA(cudaMemcpyAsync(res, pl->census, sizeof(res), cudaMemcpyDeviceToHost,
                          pl->stream));
  PF_CUDA(cudaStreamSynchronize(pl->stream));
  pl->stats.small_count = res[0];
  pl->stats.zero_count = res[1];
  pl->stats.negative_count = res[2];
  unsigned long long bits = res[3];
  bits = (bits >> 63) ? (bits & 0x7FFFFFFFFFFFFFFFull) : ~bits;
  double sm;
  memcpy(&sm, &bits, sizeof(sm));
  pl->stats.smallest = sm;
  pl->stats.factor_used = pl->factor;
  return 0;
}

// would plan_resize accept this embedding factor?  (checked BEFORE the retry
// touches the plan, so that an unsupported doubled size leaves the plan
// intact and the caller sees the negative-eigenvalue error with its census)
bool
factor_supported(const pf_plan* pl, unsigned factor)
{
  for (int i = 0; i < pl->dim; i++) {
    const long long n = (long long)factor * pl->d.cells[i];
    if (!size_supported(pl, n))
      return false;
  }
  return true;
}

int
ensure_eigen(pf_plan* pl)
{
  if (pl->eig_valid)
    return 0;
  for (;;) {
    PF_TRY(eigen_once(pl));
    if (pl->stats.negative_count > 0 && !pl->d.approximate) {
      if (pl->d.max_factor && pl->factor * 2 <= pl->d.max_factor &&
          !pl->d.periodic && factor_supported(pl, pl->factor * 2)) {
        pl->factor *= 2; // embedding-factor retry (new behaviour, off by
                         // default; the reference only throws)
        PF_TRY(plan_resize(pl));
        continue;
      }
      dev_free(&pl->lam);
      return fail(PF_ERR_NEGATIVE_EIGENVALUES,
                  "negative eigenvalues in covariance matrix");
    }
    break;
  }
  pl->eig_valid = true;
  return 0;
}

int
ensure_table(pf_plan* pl, double** tab, int mode)
{
  PF_TRY(ensure_eigen(pl));
  if (*tab)
    return 0;
  PF_TRY(dev_alloc(tab, pl->tab_elems));
  k_derive_table<<<(unsigned)((pl->tab_elems + 255) / 256), 256, 0, pl->stream>>>(
    pl->lam, *tab, pl->tab_elems, mode);
  PF_CUDA(cudaGetLastError());
  pl->launches++;
  return 0;
}

// sigma table of generateField: padded pitch Pg
int
ensure_sig_gen(pf_plan* pl, int mode)
{
  PF_TRY(ensure_eigen(pl));
  if (pl->sig_gen)
    return 0;
  const size_t total = (size_t)pl->rows * pl->Pg;
  PF_TRY(dev_alloc(&pl->sig_gen, total));
  k_derive_table_padded<<<(unsigned)((total + 255) / 256), 256, 0, pl->stream>>>(
    pl->lam, pl->sig_gen, (size_t)pl->rows, pl->P, pl->Pg, mode);
  PF_CUDA(cudaGetLastError());
  pl->launches++;
  return 0;
}

int
ensure_work(pf_plan* pl, bool need_y = true)
{
  if (pl->dim == 3) {
    if (need_y)
      PF_TRY(dev_alloc(&pl->Y, (size_t)pl->m0 * pl->c[2] * pl->n[1]));
    PF_TRY(dev_alloc(&pl->Z, (size_t)pl->c[2] * pl->c[1] * pl->m0p));
  } else if (pl->dim == 2) {
    PF_TRY(dev_alloc(&pl->Z, (size_t)pl->c[1] * pl->m0p));
  }
  return 0;
}

// table addressing shared by pass A and the fused spectral pass
void
set_table(const pf_plan* pl, PassArgs& a, const double* tab)
{
  a.tab = tab;
  a.tab_pitch = pl->P;
  if (pl->dim == 3) {
    a.rma = pl->m0;
    a.rmb = 1;
    a.tab_mirror_a = pl->compact ? (int)pl->n[1] : 0;
    a.tab_mirror_e = pl->compact ? 1 : 0;
  } else if (pl->dim == 2) {
    a.rma = 1;
    a.rmb = 0;
    a.tab_mirror_a = 0;
    a.tab_mirror_e = pl->compact ? 1 : 0;
  } else {
    a.rma = a.rmb = 0;
    a.tab_mirror_a = 0;
    a.tab_mirror_e = 0;
  }
}

// inverse passes B and C (3D) / C (2D) from Y / Z into a device field;
// `layers` = extent of the last physical dimension held here (all of it, or
// this rank's slab)
int
inverse_tail(pf_plan* pl, const cplx* Yin, long long layers, double* out_dev,
             int transform, bool blocked = false)
{
  const long long* n = pl->n;
  const long long* c = pl->c;
  if (pl->dim == 3) {
    PassArgs b = blank_args(); // pass B: k1 lines, (a,b) = (k0, x2)
    b.A = (int)pl->m0;
    b.B = (int)layers;
    b.in = Yin;
    b.isa = layers * n[1];
    b.isb = n[1];
    b.n_valid = (int)n[1];
    if (blocked) { // slab mode: Y[k0][k1/8][x2 local][k1%8]
      b.blk_T = (int)(n[1] / 8);
      b.blk_rows = (int)layers;
    }
    b.out = pl->Z; // Z[x2][x1][k0]
    b.osb = c[1] * pl->m0p;
    b.ose = pl->m0p;
    b.keep = (int)c[1];
    PF_TRY(run_pass(pl, n[1], pf::K_C2T_INV, b));
  }
  PassArgs cc = blank_args(); // pass C
  cc.A = (int)(pl->dim == 3 ? c[1] * layers : layers);
  cc.B = 1;
  cc.in = pl->dim == 3 ? pl->Z : Yin;
  cc.isa = pl->m0p;
  cc.out = out_dev;
  cc.out_pitch = c[0];
  cc.keep = (int)c[0];
  cc.transform = transform;
  cc.tparam = pl->d.boxcox_lambda;
  PF_TRY(run_pass(pl, n[0], pf::K_C2R, cc));
  return 0;
}

// generateField for a 2D plan with long lines (pf_long.cuh); `a` carries the
// noise source, keys and sigma table as prepared by generate_dev
int
generate_long2d(pf_plan* pl, PassArgs a, double* out_dev, int transform,
                int batch = 1)
{
  const long long* n = pl->n;
  const long long* c = pl->c;
  const long long m0 = pl->m0;
  const int N1a = (int)pl->LN1[1], N2a = (int)pl->LN2[1]; // pass A axis (k1)
  const int N1c = (int)pl->LN1[0], N2c = (int)pl->LN2[0]; // c2r axis (k0)
  // `batch` realizations (a.realization + r) share step 1: the sigma table is
  // read once for all of them (BASELINE config K4 is a batch of realizations);
  // the later steps run per realization on its Y
  const size_t ysize = (size_t)pl->mcols * n[1];
  if (pl->ly_batch < batch) {
    PF_CUDA(cudaStreamSynchronize(pl->stream));
    dev_free(&pl->LY);
    pl->ly_batch = batch;
  }
  PF_TRY(dev_alloc(&pl->LY, ysize * (size_t)pl->ly_batch));
  // ---- pass A, step 1: noise * sigma -> DFT over k1 -> twiddle -> Y --------
  a.cma = 1; // counter / injected index of (k0, e): k0 + m0 * e
  a.cmb = 0;
  a.cme = m0;
  a.ima = 1;
  a.imb = 0;
  a.ime = m0;
  a.na_self = (int)n[0];
  a.nb_self = 1;
  a.plane_axis = 0;
  pf::LongArgs g;
  memset(&g, 0, sizeof(g));
  g.N2 = N2a;
  g.n = n[1];
  g.n_half = n[1] / 2;
  PF_TRY(long_twiddles(pl, n[1], &g.tw));
  g.perm_n1 = N1c;
  g.perm_m = N2c / 2 + 1;
  g.k0_max = n[0] / 2;
  g.out = pl->LY;
  g.batch = batch;
  g.batch_stride = (long long)ysize;
  if (batch > 1 && a.noise_kind == pf::NOISE_PHILOX_FAST)
    PF_TRY((launch_long1<pf::LSRC_NOISE_FAST, +1, true>(pl, N1a, g, a, pl->mcols)));
  else if (batch > 1 && a.noise_kind == pf::NOISE_PHILOX_F64)
    PF_TRY((launch_long1<pf::LSRC_NOISE_F64, +1, true>(pl, N1a, g, a, pl->mcols)));
  else if (batch > 1)
    return fail(PF_ERR_INVALID, "long lines: batches need device noise");
  else if (a.noise_kind == pf::NOISE_PHILOX_FAST)
    PF_TRY((launch_long1<pf::LSRC_NOISE_FAST, +1>(pl, N1a, g, a, pl->mcols)));
  else if (a.noise_kind == pf::NOISE_PHILOX_F64)
    PF_TRY((launch_long1<pf::LSRC_NOISE_F64, +1>(pl, N1a, g, a, pl->mcols)));
  else
    PF_TRY((launch_long1<pf::LSRC_NOISE_INJECTED, +1>(pl, N1a, g, a, pl->mcols)));
  auto later_steps = [&](const cplx* Yr, double* out_r) -> int {
    // ---- pass A, step 2: sub-lines (a = column, b = j1) -> Z[x1][column],
    //      x1 = j1 + N1a * e, restricted to x1 < c1 -----------------------------
    PassArgs b = blank_args();
    b.A = (int)pl->mcols;
    b.B = N1a;
    b.in = Yr;
    b.isa = n[1];
    b.isb = N2a;
    b.n_valid = N2a;
    b.out = pl->Z;
    b.osb = pl->m0p;
    b.ose = (long long)N1a * pl->m0p;
    b.keep = (int)((c[1] + N1a - 1) / N1a);
    PF_TRY(run_pass(pl, N2a, pf::K_C2T_INV, b));
    // ---- pass C ---------------------------------------------------------------
    if (N1c == 1) {
      PassArgs cc = blank_args(); // natural column order: the short-line kernel
      cc.A = (int)c[1];
      cc.B = 1;
      cc.in = pl->Z;
      cc.isa = pl->m0p;
      cc.out = out_r;
      cc.out_pitch = c[0];
      cc.keep = (int)c[0];
      cc.transform = transform;
      cc.tparam = pl->d.boxcox_lambda;
      return run_pass(pl, n[0], pf::K_C2R, cc);
    }
    const long long pairs = (c[1] + 1) / 2;
    PF_TRY(dev_alloc(&pl->LU, (size_t)pairs * n[0]));
    PassArgs u = blank_args();
    u.A = N1c;
    u.B = (int)pairs;
    u.in = pl->Z;
    u.isa = pl->m0p;
    u.n_valid = (int)c[1];
    u.out = pl->LU;
    u.long_n1 = N1c;
    PF_TRY(long_twiddles(pl, n[0], &u.long_tw));
    PF_TRY(run_pass(pl, N2c, pf::K_LONG_C2R1, u));
    const dim3 grid((unsigned)((N2c + 255) / 256), (unsigned)pairs);
    const double tp = pl->d.boxcox_lambda;
    switch (N1c) {
      case 2:
        pf::k_long_c2r2<2><<<grid, 256, 0, pl->stream>>>(pl->LU, out_r, N2c, c[1], c[0], (int)c[0], 1.0, transform, tp);
        break;
      case 4:
        pf::k_long_c2r2<4><<<grid, 256, 0, pl->stream>>>(pl->LU, out_r, N2c, c[1], c[0], (int)c[0], 1.0, transform, tp);
        break;
      case 8:
        pf::k_long_c2r2<8><<<grid, 256, 0, pl->stream>>>(pl->LU, out_r, N2c, c[1], c[0], (int)c[0], 1.0, transform, tp);
        break;
      case 16:
        pf::k_long_c2r2<16><<<grid, 256, 0, pl->stream>>>(pl->LU, out_r, N2c, c[1], c[0], (int)c[0], 1.0, transform, tp);
        break;
      default:
        return fail(PF_ERR_INVALID, "long line: unsupported split");
    }
    PF_CUDA(cudaGetLastError());
    pl->launches++;
    return 0;
  };
  for (int r = 0; r < batch; r++)
    PF_TRY(later_steps(pl->LY + (size_t)r * ysize, out_dev + (size_t)r * pl->Dl));
  return 0;
}

int
generate_slab_dev(pf_plan* pl, int noise_kind, unsigned long long seed,
                  unsigned long long realization, const double2* noise_dev,
                  double* out_dev, int transform);

// `count` realizations of a 2D field in TWO launches: pass A takes the
// realization as its second line index (b), pass C sees count * c1 lines.
// The sigma table is read once per batch (L2-resident for the small grids
// where this matters: BASELINE config K1 is launch-bound one by one).
int
generate2d_batch(pf_plan* pl, int noise_kind, unsigned long long seed,
                 unsigned long long first, size_t count, double* out_dev,
                 int transform)
{
  PF_TRY(ensure_sig_gen(pl, 0));
  const long long* n = pl->n;
  const long long* c = pl->c;
  const long long m0 = pl->m0;
  const size_t zper = (size_t)c[1] * pl->m0p;
  // batch buffer: at most 1 GiB of Z
  const size_t chunk = std::max<size_t>(1, std::min<size_t>(count, ((size_t)1 << 26) / zper));
  if (pl->zb_elems < chunk * zper) {
    PF_CUDA(cudaStreamSynchronize(pl->stream));
    dev_free(&pl->Zb);
    PF_TRY(dev_alloc(&pl->Zb, chunk * zper));
    pl->zb_elems = chunk * zper;
  }
  for (size_t r0 = 0; r0 < count; r0 += chunk) {
    const size_t nb = std::min(chunk, count - r0);
    PassArgs a = blank_args();
    set_table(pl, a, pl->sig_gen);
    a.tab_pitch = pl->Pg;
    a.noise_kind = noise_kind;
    a.keys = pf::philox_keys(seed);
    a.realization = first + r0;
    a.real_b = 1;
    a.hermitian = 1;
    a.corrected = pl->d.parity_mode == 1;
    a.A = (int)m0; // a = k0, b = realization, e = k1
    a.B = (int)nb;
    a.cma = 1;
    a.cmb = 0;
    a.cme = m0;
    a.na_self = (int)n[0];
    a.nb_self = 1;
    a.plane_axis = 0;
    a.out = pl->Zb; // Zb[realization][x1][k0]
    a.osb = (long long)zper;
    a.ose = pl->m0p;
    a.keep = (int)c[1];
    PF_TRY(run_pass(pl, n[1], pf::K_GEN_TILE, a));
    PassArgs cc = blank_args(); // pass C over all lines of the batch
    cc.A = (int)(c[1] * (long long)nb);
    cc.B = 1;
    cc.in = pl->Zb;
    cc.isa = pl->m0p;
    cc.out = out_dev + r0 * pl->D;
    cc.out_pitch = c[0];
    cc.keep = (int)c[0];
    cc.transform = transform;
    cc.tparam = pl->d.boxcox_lambda;
    PF_TRY(run_pass(pl, n[0], pf::K_C2R, cc));
  }
  return 0;
}

// ---------------------------------------------------------------------------
// streaming route (pf_stream.cuh): large cubic 3D tori on one GPU
// ---------------------------------------------------------------------------
bool
stream_route(const pf_plan* pl)
{
  if (!pl->stream_on || pl->dim != 3 || pl->nranks > 1 || pl->factor != 2)
    return false;
  const long long N = pl->n[0];
  if ((N != 512 && N != 1024) || pl->n[1] != N || pl->n[2] != N)
    return false;
  if (pl->split_one || pl->gen3_force || !(pl->pipe_mask & (1u << pf::K_GEN3)))
    return false; // experiments / tests that pin the three-pass kernels
  return true;
}

int
encode_map3(void* base, cuuint64_t d0, cuuint64_t d1, cuuint64_t d2,
            cuuint64_t s1_bytes, cuuint64_t s2_bytes, cuuint32_t b0,
            cuuint32_t b1, cuuint32_t b2, CUtensorMapSwizzle swz, CUtensorMap* tm)
{
  encode_tiled_fn enc = get_encode_tiled();
  if (!enc)
    return fail(PF_ERR_CUDA, "cuTensorMapEncodeTiled is not available");
  const cuuint64_t gdim[3] = { d0, d1, d2 };
  const cuuint64_t gstr[2] = { s1_bytes, s2_bytes };
  const cuuint32_t box[3] = { b0, b1, b2 };
  const cuuint32_t estr[3] = { 1, 1, 1 };
  const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, base, gdim, gstr,
                         box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(PF_ERR_CUDA, "cuTensorMapEncodeTiled (stream) failed with code " +
                               std::to_string((int)r));
  return 0;
}

// last axis of the streaming route: Zt[k0][x2 < layers][x1] -> field
int
stream_last_axis(pf_plan* pl, cplx* zt, long long layers, double* out_dev,
                 int transform)
{
  const long long N = pl->n[0];
  const long long* c = pl->c;
  const cplx* tw = nullptr;
  PF_TRY(twiddles(pl, N, &tw));
  auto fn = N == 1024 ? pf_stream_launch_n1024 : pf_stream_launch_n512;
  const int grid = pl->sm_cap > 0 ? pl->sm_cap : pl->sm_count;
  CUtensorMap maps[2];
  PassArgs cc = blank_args();
  cc.A = (int)c[1];
  cc.B = (int)layers;
  cc.out = out_dev;
  cc.out_pitch = c[0];
  cc.keep = (int)c[0];
  cc.transform = transform;
  cc.tparam = pl->d.boxcox_lambda;
  PF_TRY(encode_map3(zt, 2 * (cuuint64_t)c[1], (cuuint64_t)layers, (cuuint64_t)pl->m0,
                     (cuuint64_t)c[1] * 16, (cuuint64_t)layers * c[1] * 16, 16, 1, 256,
                     CU_TENSOR_MAP_SWIZZLE_128B, &maps[0]));
  PF_TRY(encode_map3(zt, 2 * (cuuint64_t)c[1], (cuuint64_t)layers, (cuuint64_t)pl->m0,
                     (cuuint64_t)c[1] * 16, (cuuint64_t)layers * c[1] * 16, 16, 1, 8,
                     CU_TENSOR_MAP_SWIZZLE_128B, &maps[1]));
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (pl->profiling) {
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, pl->stream);
  }
  const int e = fn(1, &cc, tw, maps, grid, pl->stream);
  if (e != 0)
    return fail(PF_ERR_CUDA, std::string("kernel launch failed (kp_c2rt): ") +
                               cudaGetErrorString((cudaError_t)e));
  pl->launches++;
  if (pl->profiling) {
    cudaEventRecord(e1, pl->stream);
    pl->prof_events.push_back({ pf::K_C2R, { e0, e1 } });
  }
  return 0;
}

// first two axes of the streaming route for the planes [plane0, plane0 +
// planes) of k0.  `a`: pass A as prepared by generate_dev / slab_issue_a
// (table, noise, counters, Hermitian rule; lines a = k1, b = k0, elements k2).
// peer_zt != nullptr: pass B's lines go to the owners' buffers (multi-GPU).
int
stream_first_axes(pf_plan* pl, const PassArgs& a, int plane0, int planes,
                  cplx* zt, cplx* const* peer_zt, long long layers_per_peer)
{
  const long long N = pl->n[0];
  const long long* c = pl->c;
  const int lpg = N == 1024 ? 4 : 8;
  const int grid = pl->sm_cap > 0 ? pl->sm_cap : pl->sm_count;
  // B-items run `lag` planes behind the A-lines: the groups in flight hold
  // 3 * grid consecutive tickets, i.e. that many / (items per plane) planes,
  // and a plane's A-lines must be long finished when its B-items are drawn;
  // the ring holds the planes in between plus slack
  if (pl->stream_lag <= 0) {
    const long long per_plane = N / lpg + c[2] / lpg;
    pl->stream_lag = (int)((3LL * pl->sm_count + per_plane - 1) / per_plane) + 2;
  }
  if (pl->stream_slots < pl->stream_lag + 3) {
    dev_free(&pl->ring);
    pl->stream_slots = pl->stream_lag + 3;
  }
  const int slots = pl->stream_slots;
  PF_TRY(dev_alloc(&pl->ring, (size_t)slots * N * c[2]));
  const size_t nsync = 16 + 2 * (size_t)pl->m0;
  PF_TRY(dev_alloc(&pl->stream_sync, nsync));
  const cplx* tw = nullptr;
  PF_TRY(twiddles(pl, N, &tw));
  auto fn = N == 1024 ? pf_stream_launch_n1024 : pf_stream_launch_n512;
  PF_CUDA(cudaMemsetAsync(pl->stream_sync, 0, nsync * sizeof(unsigned), pl->stream));
  pf::StreamArgs sa;
  memset(&sa, 0, sizeof(sa));
  sa.a = a;
  sa.a.B = planes;
  sa.a.keep = (int)c[2];
  sa.ring = pl->ring;
  sa.slots = slots;
  sa.zt = zt;
  sa.plane0 = plane0;
  if (peer_zt) {
    sa.n_peers = pl->nranks;
    sa.rows_per_peer = (int)layers_per_peer;
    for (int q = 0; q < pl->nranks; q++)
      sa.peer_zt[q] = peer_zt[q];
    sa.zs_k0 = layers_per_peer * c[1];
    sa.zs_x2 = c[1];
    sa.zs_x1 = 1;
  } else if (pl->stream_zt == 1) { // Z[x2][x1][k0], row pitch m0p
    sa.zs_k0 = 1;
    sa.zs_x1 = pl->m0p;
    sa.zs_x2 = c[1] * pl->m0p;
  } else {                         // Zt[k0][x2][x1]
    sa.zs_k0 = c[2] * c[1];
    sa.zs_x2 = c[1];
    sa.zs_x1 = 1;
  }
  sa.c1 = (int)c[1];
  sa.c2 = (int)c[2];
  sa.lag = pl->stream_lag;
  sa.sync = pl->stream_sync;
  CUtensorMap map;
  // ring[slot][k1][x2]: box = lpg x2 (64 / 128 bytes) x 256 k1
  PF_TRY(encode_map3(pl->ring, 2 * (cuuint64_t)c[2], (cuuint64_t)N, (cuuint64_t)slots,
                     (cuuint64_t)c[2] * 16, (cuuint64_t)N * c[2] * 16,
                     (cuuint32_t)(2 * lpg), 256, 1,
                     lpg == 4 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                     &map));
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (pl->profiling) {
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, pl->stream);
  }
  const int e = fn(0, &sa, tw, &map, grid, pl->stream);
  if (e != 0)
    return fail(PF_ERR_CUDA, std::string("kernel launch failed (kp_ab): ") +
                               cudaGetErrorString((cudaError_t)e));
  pl->launches++;
  if (pl->profiling) {
    cudaEventRecord(e1, pl->stream);
    pl->prof_events.push_back({ 12, { e0, e1 } }); // fused first two axes
  }
  return 0;
}

// one GPU: both kernels back to back
int
generate_stream(pf_plan* pl, const PassArgs& a, double* out_dev, int transform)
{
  const long long N = pl->n[0];
  const long long* c = pl->c;
  PF_TRY(stream_first_axes(pl, a, 0, (int)pl->m0, pl->Z, nullptr, 0));
  if (pl->stream_zt == 1) {
    // contiguous-line pass C of the three-pass route on Z[x2][x1][k0]
    PassArgs c3 = blank_args();
    c3.A = (int)(c[1] * c[2]);
    c3.B = 1;
    c3.in = pl->Z;
    c3.isa = pl->m0p;
    c3.out = out_dev;
    c3.out_pitch = c[0];
    c3.keep = (int)c[0];
    c3.transform = transform;
    c3.tparam = pl->d.boxcox_lambda;
    return run_pass(pl, N, pf::K_C2R, c3);
  }
  return stream_last_axis(pl, pl->Z, c[2], out_dev, transform);
}

// generateField and the covariance operators on the generic route
int
generate_generic(pf_plan* pl, int noise_kind, unsigned long long seed,
                 unsigned long long realization, const double2* noise_dev,
                 double* out_dev, double* spare_dev, int transform)
{
  PF_TRY(ensure_table(pl, &pl->sig_gen, pl->hermitian ? 0 : 1));
  PF_TRY(dev_alloc(&pl->GW, pl->tab_elems));
  const pf::GenGeom gg = gen_geom(pl);
  const long long total = (long long)pl->tab_elems;
  const pf::PhiloxKeys keys = pf::philox_keys(seed);
  const unsigned grid = grid_for((size_t)total);
  const int herm = pl->hermitian ? 1 : 0, corr = pl->d.parity_mode == 1;
  if (noise_kind == pf::NOISE_PHILOX_FAST)
    pf::k_generic_noise<pf::NOISE_PHILOX_FAST><<<grid, 256, 0, pl->stream>>>(
      gg, pl->sig_gen, keys, realization, noise_dev, herm, corr, pl->GW, total);
  else if (noise_kind == pf::NOISE_PHILOX_F64)
    pf::k_generic_noise<pf::NOISE_PHILOX_F64><<<grid, 256, 0, pl->stream>>>(
      gg, pl->sig_gen, keys, realization, noise_dev, herm, corr, pl->GW, total);
  else
    pf::k_generic_noise<pf::NOISE_INJECTED><<<grid, 256, 0, pl->stream>>>(
      gg, pl->sig_gen, keys, realization, noise_dev, herm, corr, pl->GW, total);
  PF_CUDA(cudaGetLastError());
  pl->launches++;
  return generic_inverse(pl, pl->GW, out_dev, spare_dev, transform);
}

int
apply_generic(pf_plan* pl, const double* tab, const double* in_dev, double* out_dev)
{
  const long long* n = pl->n;
  const long long* c = pl->c;
  const long long m0 = pl->m0;
  PF_TRY(dev_alloc(&pl->GW, pl->tab_elems));
  cplx* W = pl->GW;
  { // axis 0 (r2c; 1D: complex transform of the real line), zero padding implicit
    pf::GenLineArgs g = gen_blank();
    g.L0 = pl->dim > 1 ? c[1] : 1;
    g.L1 = pl->dim == 3 ? c[2] : 1;
    g.in_mode = 1;
    g.in = in_dev;
    g.in_s0 = c[0];
    g.in_s1 = c[0] * c[1];
    g.in_se = 1;
    g.n_valid = (int)c[0];
    g.out = W;
    g.out_s0 = m0;
    g.out_s1 = m0 * n[1];
    g.out_se = 1;
    g.keep = (int)m0;
    PF_TRY(gen_lines(pl, n[0], -1, g));
  }
  if (pl->dim >= 2) { // axis 1: lines (k0, x2 < c2), input x1 < c1
    pf::GenLineArgs g = gen_blank();
    g.L0 = m0;
    g.L1 = pl->dim == 3 ? c[2] : 1;
    g.in = g.out = W;
    g.in_s0 = g.out_s0 = 1;
    g.in_s1 = g.out_s1 = m0 * n[1];
    g.in_se = g.out_se = m0;
    g.n_valid = (int)c[1];
    g.keep = (int)n[1];
    PF_TRY(gen_lines(pl, n[1], -1, g));
  }
  if (pl->dim == 3) { // axis 2: lines (k0, k1), input x2 < c2
    pf::GenLineArgs g = gen_blank();
    g.L0 = m0;
    g.L1 = n[1];
    g.in = g.out = W;
    g.in_s0 = g.out_s0 = 1;
    g.in_s1 = g.out_s1 = m0;
    g.in_se = g.out_se = m0 * n[1];
    g.n_valid = (int)c[2];
    g.keep = (int)n[2];
    PF_TRY(gen_lines(pl, n[2], -1, g));
  }
  const long long total = (long long)pl->tab_elems;
  pf::k_generic_mult<<<grid_for((size_t)total), 256, 0, pl->stream>>>(
    W, tab, 1.0 / (double)pl->Next, total);
  PF_CUDA(cudaGetLastError());
  pl->launches++;
  return generic_inverse(pl, W, out_dev, nullptr, 0);
}

// Matrix::generateField (matrix.hh:343-438) into device memory
int
generate_dev(pf_plan* pl, int noise_kind, unsigned long long seed,
             unsigned long long realization, const double2* noise_dev,
             double* out_dev, double* spare_dev, int transform)
{
  if (pl->nranks > 1)
    return generate_slab_dev(pl, noise_kind, seed, realization, noise_dev,
                             out_dev, transform);
  if (pl->generic)
    return generate_generic(pl, noise_kind, seed, realization, noise_dev, out_dev,
                            spare_dev, transform);
  PF_TRY(ensure_sig_gen(pl, pl->hermitian ? 0 : 1));
  const bool stream = stream_route(pl) && pl->gen_phase == 0 && !pl->y_override &&
                      (pl->Pg % 2 == 0);
  PF_TRY(ensure_work(pl, !stream));
  const long long* n = pl->n;
  const long long* c = pl->c;
  const long long m0 = pl->m0;
  PassArgs a = blank_args();
  set_table(pl, a, pl->sig_gen);
  a.tab_pitch = pl->Pg;
  a.noise_kind = noise_kind;
  a.keys = pf::philox_keys(seed);
  a.realization = realization;
  a.noise = noise_dev;
  a.hermitian = pl->hermitian ? 1 : 0;
  a.corrected = pl->d.parity_mode == 1;
  if (pl->dim == 3) {
    a.A = (int)n[1]; // a = k1, b = k0, e = k2
    a.B = (int)m0;
    a.cma = m0 * (n[2] / 2);
    a.cmb = 1;
    a.cme = m0;
    a.ima = m0 * n[2]; // s = k0 + m0*(k2 + n2*k1), r2cfieldbackend.hh:190-199
    a.imb = 1;
    a.ime = m0;
    a.na_self = (int)n[1];
    a.nb_self = (int)n[0];
    a.plane_axis = 1;
    if (stream)
      return generate_stream(pl, a, out_dev, transform);
    cplx* Yb = pl->y_override ? pl->y_override : pl->Y;
    a.out = Yb; // Y[k0][x2][k1]
    a.osb = c[2] * n[1];
    a.ose = n[1];
    a.keep = (int)c[2];
    if (pl->gen_phase != 2)
      PF_TRY(run_pass(pl, n[2], pf::K_GEN_TILE, a));
    if (pl->gen_phase != 1)
      PF_TRY(inverse_tail(pl, Yb, c[2], out_dev, transform));
  } else if (pl->dim == 2 && pl->long2d) {
    PF_TRY(generate_long2d(pl, a, out_dev, transform, pl->long_batch));
  } else if (pl->dim == 2) {
    a.A = (int)m0; // a = k0, e = k1
    a.B = 1;
    a.cma = 1;
    a.cmb = 0;
    a.cme = m0;
    a.ima = 1; // s = k0 + m0*k1
    a.imb = 0;
    a.ime = m0;
    a.na_self = (int)n[0];
    a.nb_self = 1;
    a.plane_axis = 0;
    a.out = pl->Z; // Z[x1][k0]
    a.osb = 0;
    a.ose = pl->m0p;
    a.keep = (int)c[1];
    PF_TRY(run_pass(pl, n[1], pf::K_GEN_TILE, a));
    PF_TRY(inverse_tail(pl, pl->Z, c[1], out_dev, transform));
  } else {
    a.A = 1;
    a.B = 1;
    a.cme = 1;
    a.ime = 1;
    a.out = out_dev;
    a.out2 = spare_dev;
    a.out_pitch = c[0];
    a.keep = (int)c[0];
    a.transform = transform;
    a.tparam = pl->d.boxcox_lambda;
    PF_TRY(run_pass(pl, n[0], pf::K_GEN_REIM, a));
  }
  return 0;
}

// ---------------------------------------------------------------------------
// slab mode: generateField over the GPUs of one node.  Rank r runs pass A for
// its share of the (a) lines and stores every transposed tile row directly
// into the memory of the rank that owns that layer of the last dimension
// (peer stores over NVLink); flags in peer memory order the passes.  After
// the exchange the layout is the reference's field distribution
// (fieldtraits.hh:281-282), so passes B and C are local.
// ---------------------------------------------------------------------------
cplx*
slab_buffer(const pf_plan* pl, void* block, int which)
{
  return reinterpret_cast<cplx*>(static_cast<unsigned char*>(block)) +
         (size_t)which * pl->xbuf_elems;
}
unsigned long long*
slab_flags(const pf_plan* pl, void* block, int which)
{
  return reinterpret_cast<unsigned long long*>(
           static_cast<unsigned char*>(block) +
           2 * pl->xbuf_elems * sizeof(cplx)) +
         8 * which;
}

cplx*
slab_fwd(const pf_plan* pl, void* block)
{
  return reinterpret_cast<cplx*>(static_cast<unsigned char*>(block) +
                                 2 * pl->xbuf_elems * sizeof(cplx) + 512);
}
// streaming route: buffer `which` of pass B's output Zt[k0][x2 local][x1]
cplx*
slab_zt(const pf_plan* pl, void* block, int which)
{
  return slab_fwd(pl, block) + pl->fwd_elems + (size_t)which * pl->zt_elems;
}

// first half of a slab-mode realization: pass A + start of the exchange.
// Returns the epoch of this realization in *epoch_out.
int
slab_issue_a(pf_plan* pl, int noise_kind, unsigned long long seed,
             unsigned long long realization, const double2* noise_dev,
             unsigned long long* epoch_out)
{
  if (!pl->connected)
    return fail(PF_ERR_INVALID,
                "slab mode: call pf_slab_connect on every rank first");
  PF_TRY(ensure_sig_gen(pl, 0));
  if (pl->dim == 3)
    PF_TRY(dev_alloc(&pl->Z, (size_t)pl->cl * pl->c[1] * pl->m0p));
  const long long* n = pl->n;
  const long long* c = pl->c;
  const long long m0 = pl->m0;
  const int P = pl->nranks;
  const unsigned long long e = ++pl->epoch;
  const int s = (int)(e & 1);
  *epoch_out = e;
  if (pl->stream_slab) {
    // streaming route: this rank takes its planes of k0 through BOTH first
    // axes (kp_ab) and stores pass B's lines straight into the owners' Zt
    // buffers -- no exchange pass, half the NVLink volume of the transposed
    // transform (the lines are pruned twice before they travel)
    if (e > 2) { // the owners have finished reading buffer s (epoch e - 2)
      k_slab_wait<<<1, 32, 0, pl->stream>>>(slab_flags(pl, pl->slab_block, 1), P,
                                            e - 2, pl->d_timeout, pl->slab_limit_ns);
      pl->launches++;
    }
    PassArgs a = blank_args();
    set_table(pl, a, pl->sig_gen);
    a.tab_pitch = pl->Pg;
    a.noise_kind = noise_kind;
    a.keys = pf::philox_keys(seed);
    a.realization = realization;
    a.noise = noise_dev;
    a.hermitian = 1;
    a.corrected = pl->d.parity_mode == 1;
    a.A = (int)n[1]; // a = k1, b = k0 (this rank's planes), e = k2
    a.cma = m0 * (n[2] / 2);
    a.cmb = 1;
    a.cme = m0;
    a.ima = m0 * n[2];
    a.imb = 1;
    a.ime = m0;
    a.na_self = (int)n[1];
    a.nb_self = (int)n[0];
    a.plane_axis = 1;
    cplx* peers[8] = { nullptr };
    for (int q = 0; q < P; q++)
      peers[q] = slab_zt(pl, pl->peer_block[q], s);
    PF_TRY(stream_first_axes(pl, a, pl->plane0, pl->planes, peers[pl->rank], peers,
                             pl->cl));
    k_slab_signal<<<1, 32, 0, pl->stream>>>(pl->d_flagsA, P, pl->rank, e);
    pl->launches++;
    PF_CUDA(cudaGetLastError());
    return 0;
  }
  if (pl->slab_dma) {
    for (int i = 0; i < 2; i++)
      PF_TRY(dev_alloc(&pl->send_stage[i], (size_t)(P - 1) * pl->xbuf_elems));
    // the copies out of this staging buffer (two calls ago) must be done
    PF_CUDA(cudaStreamWaitEvent(pl->stream, pl->ev_x[s], 0));
  } else if (e > 2) {
    // peers must have consumed what we stored into their buffer s
    k_slab_wait<<<1, 32, 0, pl->stream>>>(slab_flags(pl, pl->slab_block, 1), P,
                                          e - 2, pl->d_timeout, pl->slab_limit_ns);
    pl->launches++;
  }
  PassArgs a = blank_args();
  set_table(pl, a, pl->sig_gen);
  a.tab_pitch = pl->Pg;
  a.noise_kind = noise_kind;
  a.keys = pf::philox_keys(seed);
  a.realization = realization;
  a.noise = noise_dev;
  a.hermitian = 1;
  a.corrected = pl->d.parity_mode == 1;
  a.a_begin = pl->a_begin;
  a.A = pl->a_count;
  a.n_peers = P;
  a.rows_per_peer = (int)pl->cl;
  for (int q = 0, k = 0; q < P; q++) {
    if (q == pl->rank || !pl->slab_dma)
      a.peer_out[q] = slab_buffer(pl, pl->peer_block[q], s);
    else
      a.peer_out[q] = pl->send_stage[s] + (size_t)(k++) * pl->xbuf_elems;
  }
  a.out = a.peer_out[pl->rank];
  // geometry of the rows this rank contributes to one exchange buffer, in
  // complex units: `rows` runs of `xwidth` elements at offset `xoff`
  long long rows, pitch, xoff = 0, xwidth = 0;
  if (pl->dim == 3) {
    a.B = (int)m0; // a = k1 (split over ranks), b = k0, e = k2
    a.cma = m0 * (n[2] / 2);
    a.cmb = 1;
    a.cme = m0;
    a.ima = m0 * n[2];
    a.imb = 1;
    a.ime = m0;
    a.na_self = (int)n[1];
    a.nb_self = (int)n[0];
    a.plane_axis = 1;
    a.osb = pl->cl * n[1]; // Y_q[k0][k1/8][x2 local][k1%8] (blocked)
    a.ose = n[1];
    a.keep = (int)c[2];
    if (n[1] % 8 == 0) {
      a.blk_T = (int)(n[1] / 8);
      a.blk_rows = (int)pl->cl;
      // per plane k0 this rank's tiles are one contiguous run
      rows = m0;
      pitch = n[1] * pl->cl;
      xoff = (long long)(pl->a_begin / 8) * pl->cl * 8;
      xwidth = (long long)((pl->a_count + 7) / 8) * pl->cl * 8;
    } else {
      rows = m0 * pl->cl;
      pitch = n[1];
      xoff = pl->a_begin;
      xwidth = pl->a_count;
    }
    if (a.A > 0)
      PF_TRY(run_pass(pl, n[2], pf::K_GEN_TILE, a));
  } else {
    a.B = 1; // a = k0 (split over ranks), e = k1
    a.cma = 1;
    a.cmb = 0;
    a.cme = m0;
    a.ima = 1;
    a.imb = 0;
    a.ime = m0;
    a.na_self = (int)n[0];
    a.nb_self = 1;
    a.plane_axis = 0;
    a.osb = 0; // Z_q[x1 local][k0]
    a.ose = pl->m0p;
    a.keep = (int)c[1];
    rows = pl->cl;
    pitch = pl->m0p;
    xoff = pl->a_begin;
    xwidth = pl->a_count;
    if (a.A > 0)
      PF_TRY(run_pass(pl, n[1], pf::K_GEN_TILE, a));
  }
  if (pl->slab_dma) {
    PF_CUDA(cudaEventRecord(pl->ev_a[s], pl->stream));
    PF_CUDA(cudaStreamWaitEvent(pl->copy_stream, pl->ev_a[s], 0));
    if (e > 2) {
      k_slab_wait<<<1, 32, 0, pl->copy_stream>>>(
        slab_flags(pl, pl->slab_block, 1), P, e - 2, pl->d_timeout, pl->slab_limit_ns);
      pl->launches++;
    }
    if (pl->a_count > 0) {
      const int ns = pl->peer_streams;
      if (ns > 1)
        PF_CUDA(cudaEventRecord(pl->ev_gate, pl->copy_stream));
      // destinations in ring order starting behind this rank, so that at any
      // time the ranks target different peers
      for (int i = 1; i < P; i++) {
        const int q = (pl->rank + i) % P;
        const int k = q < pl->rank ? q : q - 1; // staging slot of peer q
        const cplx* src = pl->send_stage[s] + (size_t)k * pl->xbuf_elems + xoff;
        cplx* dst = slab_buffer(pl, pl->peer_block[q], s) + xoff;
        cudaStream_t st = ns > 1 ? pl->peer_stream[i % ns] : pl->copy_stream;
        if (ns > 1 && i <= ns)
          PF_CUDA(cudaStreamWaitEvent(st, pl->ev_gate, 0));
        PF_CUDA(cudaMemcpy2DAsync(dst, (size_t)pitch * sizeof(cplx), src,
                                  (size_t)pitch * sizeof(cplx),
                                  (size_t)xwidth * sizeof(cplx),
                                  (size_t)rows, cudaMemcpyDeviceToDevice, st));
      }
      for (int j = 0; j < ns && ns > 1; j++) {
        PF_CUDA(cudaEventRecord(pl->ev_peer[j], pl->peer_stream[j]));
        PF_CUDA(cudaStreamWaitEvent(pl->copy_stream, pl->ev_peer[j], 0));
      }
    }
    k_slab_signal<<<1, 32, 0, pl->copy_stream>>>(pl->d_flagsA, P, pl->rank, e);
    pl->launches++;
    PF_CUDA(cudaEventRecord(pl->ev_x[s], pl->copy_stream));
  } else {
    k_slab_signal<<<1, 32, 0, pl->stream>>>(pl->d_flagsA, P, pl->rank, e);
    pl->launches++;
  }
  PF_CUDA(cudaGetLastError());
  return 0;
}

// second half: wait for every rank's rows, passes B and C on the local slab
int
slab_issue_tail(pf_plan* pl, unsigned long long e, double* out_dev,
                int transform, bool from_zt = false)
{
  const int P = pl->nranks;
  const int s = (int)(e & 1);
  k_slab_wait<<<1, 32, 0, pl->stream>>>(slab_flags(pl, pl->slab_block, 0), P, e,
                                        pl->d_timeout, pl->slab_limit_ns);
  pl->launches++;
  if (from_zt)
    PF_TRY(stream_last_axis(pl, slab_zt(pl, pl->slab_block, s), pl->cl, out_dev,
                            transform));
  else
    PF_TRY(inverse_tail(pl, slab_buffer(pl, pl->slab_block, s), pl->cl, out_dev,
                        transform, pl->dim == 3 && pl->n[1] % 8 == 0));
  // publish "buffer s consumed"
  k_slab_signal<<<1, 32, 0, pl->stream>>>(pl->d_flagsB, P, pl->rank, e);
  pl->launches++;
  PF_CUDA(cudaGetLastError());
  return 0;
}

int
generate_slab_dev(pf_plan* pl, int noise_kind, unsigned long long seed,
                  unsigned long long realization, const double2* noise_dev,
                  double* out_dev, int transform)
{
  unsigned long long e = 0;
  PF_TRY(slab_issue_a(pl, noise_kind, seed, realization, noise_dev, &e));
  return slab_issue_tail(pl, e, out_dev, transform, pl->stream_slab);
}

// The same operators in slab mode (P ranks, fields distributed over the last
// dimension as in fieldtraits.hh:281-282).  Two exchanges, mirroring the two
// FFTW-MPI transposes of the reference's forward and backward transforms:
//   forward : after the local F0 / F1 passes every rank sends, for the lines
//             of rank q, its cl samples of the last axis into q's forward
//             region (strided DMA copies) -- epoch e1;
//   backward: the fused spectral pass writes its transposed tiles into the
//             owners' exchange buffers (peer stores, as pass A) -- epoch e2;
// then passes B and C run on the local layers.
int
apply_slab_dev(pf_plan* pl, int op, const double* in_dev, double* out_dev)
{
  if (!pl->connected)
    return fail(PF_ERR_INVALID,
                "slab mode: call pf_slab_connect on every rank first");
  double** tabp = op == PF_OP_MATRIX ? &pl->lam
                  : op == PF_OP_ROOT ? &pl->sig
                                     : &pl->inv;
  if (op == PF_OP_MATRIX)
    PF_TRY(ensure_eigen(pl));
  else
    PF_TRY(ensure_table(pl, tabp, op == PF_OP_ROOT ? 1 : 2));
  const long long* n = pl->n;
  const long long* c = pl->c;
  const long long m0 = pl->m0, cl = pl->cl;
  const int P = pl->nranks;
  const double invN = 1.0 / (double)pl->Next;
  const long long clast = c[pl->dim - 1];
  const long long rows_per_line = pl->dim == 3 ? m0 : 1; // rows per a-line
  if (pl->dim == 3)
    PF_TRY(dev_alloc(&pl->Z, (size_t)cl * c[1] * pl->m0p));
  PF_TRY(dev_alloc(&pl->V, (size_t)pl->a_total * rows_per_line * cl));
  // ---- local forward passes -------------------------------------------------
  if (pl->dim == 3) {
    cplx* Up = pl->Z; // pruned U[cl][m0][c1]
    PassArgs f0 = blank_args(); // F0: x0 lines, (a,b) = (x1, x2 local)
    f0.A = (int)c[1];
    f0.B = (int)cl;
    f0.in = in_dev;
    f0.isa = c[0];
    f0.isb = c[1] * c[0];
    f0.n_valid = (int)c[0];
    f0.out = Up;
    f0.osb = m0 * c[1];
    f0.ose = c[1];
    f0.keep = (int)m0;
    PF_TRY(run_pass(pl, n[0], pf::K_R2C_LOAD, f0));
    PassArgs f1 = blank_args(); // F1: x1 lines, (a,b) = (x2 local, k0)
    f1.A = (int)cl;
    f1.B = (int)m0;
    f1.in = Up;
    f1.isa = m0 * c[1];
    f1.isb = c[1];
    f1.n_valid = (int)c[1];
    f1.out = pl->V; // V[k1][k0][x2 local]
    f1.osb = cl;
    f1.ose = m0 * cl;
    f1.keep = (int)n[1];
    PF_TRY(run_pass(pl, n[1], pf::K_C2T_FWD, f1));
  } else {
    PassArgs f0 = blank_args(); // F0: x0 lines, a = x1 local
    f0.A = (int)cl;
    f0.B = 1;
    f0.in = in_dev;
    f0.isa = c[0];
    f0.n_valid = (int)c[0];
    f0.out = pl->V; // V[k0][x1 local]
    f0.osb = 0;
    f0.ose = cl;
    f0.keep = (int)m0;
    PF_TRY(run_pass(pl, n[0], pf::K_R2C_LOAD, f0));
  }
  // ---- forward exchange, epoch e1 ---------------------------------------------
  const unsigned long long e1 = ++pl->epoch;
  if (e1 > 2) {
    k_slab_wait<<<1, 32, 0, pl->stream>>>(slab_flags(pl, pl->slab_block, 1), P,
                                          e1 - 2, pl->d_timeout, pl->slab_limit_ns);
    pl->launches++;
  }
  for (int i = 0; i < P; i++) {
    const int q = (pl->rank + i) % P;
    const long long qb = std::min<long long>(pl->a_total, pl->a_chunk * q);
    const long long qe = std::min<long long>(pl->a_total, pl->a_chunk * (q + 1));
    const long long rows = (qe - qb) * rows_per_line;
    if (rows <= 0)
      continue;
    const cplx* src = pl->V + (size_t)qb * rows_per_line * cl;
    cplx* dst = slab_fwd(pl, pl->peer_block[q]) + (size_t)pl->rank * cl;
    PF_CUDA(cudaMemcpy2DAsync(dst, (size_t)clast * sizeof(cplx), src,
                              (size_t)cl * sizeof(cplx), (size_t)cl * sizeof(cplx),
                              (size_t)rows, cudaMemcpyDeviceToDevice, pl->stream));
  }
  k_slab_signal<<<1, 32, 0, pl->stream>>>(pl->d_flagsA, P, pl->rank, e1);
  k_slab_wait<<<1, 32, 0, pl->stream>>>(slab_flags(pl, pl->slab_block, 0), P, e1,
                                        pl->d_timeout, pl->slab_limit_ns);
  pl->launches += 2;
  // ---- fused spectral pass, tiles stored into the owners' buffers, epoch e2 ---
  const unsigned long long e2 = ++pl->epoch;
  const int s2 = (int)(e2 & 1);
  if (e2 > 2) {
    k_slab_wait<<<1, 32, 0, pl->stream>>>(slab_flags(pl, pl->slab_block, 1), P,
                                          e2 - 2, pl->d_timeout, pl->slab_limit_ns);
    pl->launches++;
  }
  PassArgs s = blank_args();
  set_table(pl, s, *tabp);
  s.a_begin = pl->a_begin;
  s.A = pl->a_count;
  s.in = slab_fwd(pl, pl->slab_block);
  s.n_valid = (int)clast;
  s.scale = invN;
  s.keep = (int)clast;
  s.n_peers = P;
  s.rows_per_peer = (int)cl;
  for (int q = 0; q < P; q++)
    s.peer_out[q] = slab_buffer(pl, pl->peer_block[q], s2);
  s.out = s.peer_out[pl->rank];
  if (pl->dim == 3) {
    s.B = (int)m0; // lines (a = k1 local, b = k0) along x2
    s.isa = m0 * clast;
    s.isb = clast;
    s.osb = cl * n[1];
    s.ose = n[1];
    if (n[1] % 8 == 0) {
      s.blk_T = (int)(n[1] / 8);
      s.blk_rows = (int)cl;
    }
    if (s.A > 0)
      PF_TRY(run_pass(pl, n[2], pf::K_SPEC_TILE, s));
  } else {
    s.B = 1; // lines a = k0 local along x1
    s.isa = clast;
    s.osb = 0;
    s.ose = pl->m0p;
    if (s.A > 0)
      PF_TRY(run_pass(pl, n[1], pf::K_SPEC_TILE, s));
  }
  // forward region consumed (e1), tiles delivered (e2)
  k_slab_signal<<<1, 32, 0, pl->stream>>>(pl->d_flagsB, P, pl->rank, e1);
  k_slab_signal<<<1, 32, 0, pl->stream>>>(pl->d_flagsA, P, pl->rank, e2);
  pl->launches += 2;
  PF_CUDA(cudaGetLastError());
  return slab_issue_tail(pl, e2, out_dev, 0);
}

// ---------------------------------------------------------------------------
// slab <-> block redistribution with overlap (StochasticPart::dataToEval /
// evalToData / exchangeOverlap, stochastic.hh:830-1127; block layout of
// DefaultLoadBalance, fieldtraits.hh:483-562, offsets stochastic.hh:237-257)
// ---------------------------------------------------------------------------
struct BlockGeom
{
  int dim, P, rank;
  unsigned cl;     // layers of the last dimension per slab
  unsigned c[3];   // global cells
  unsigned ppd[3]; // processors per dimension
  unsigned le[3];  // localEvalCells
  unsigned lo[3];  // localEvalOffset of this rank
  const double* peer[8]; // staged arrays of all ranks (peer memory)
};

__device__ __forceinline__ double
slab_value(const BlockGeom& g, unsigned x, unsigned y, unsigned z)
{
  // the slab decomposition splits the LAST dimension
  const unsigned last = g.dim == 3 ? z : (g.dim == 2 ? y : x);
  const unsigned owner = last / g.cl, ll = last - owner * g.cl;
  size_t idx;
  if (g.dim == 3)
    idx = x + (size_t)g.c[0] * (y + (size_t)g.c[1] * ll);
  else if (g.dim == 2)
    idx = x + (size_t)g.c[0] * ll;
  else
    idx = ll;
  return g.peer[owner][idx];
}

// evalVector (block of this rank, natural order) followed by the 2 * dim
// overlap faces: [lower halo of dim 0][upper halo of dim 0][lower of dim 1] ...
__global__ void
k_slab_to_blocks(BlockGeom g, double* __restrict__ eval, double* __restrict__ overlap,
                 size_t Dl, size_t total)
{
  for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (size_t)gridDim.x * blockDim.x) {
    if (t < Dl) {
      const unsigned e0 = (unsigned)(t % g.le[0]);
      const unsigned e1 = (unsigned)((t / g.le[0]) % g.le[1]);
      const unsigned e2 = (unsigned)(t / ((size_t)g.le[0] * g.le[1]));
      eval[t] = slab_value(g, g.lo[0] + e0, g.lo[1] + e1, g.lo[2] + e2);
      continue;
    }
    size_t r = t - Dl;
    for (int i = 0; i < g.dim; i++) {
      const size_t F = Dl / g.le[i];
      if (r >= 2 * F) {
        r -= 2 * F;
        continue;
      }
      const bool upper = r >= F;
      const size_t f = upper ? r - F : r;
      unsigned e[3] = { 0, 0, 0 };
      if (g.dim == 3) {
        const int in = (i + 1) % 3, inn = (i + 2) % 3;
        e[in] = (unsigned)(f % g.le[in]);
        e[inn] = (unsigned)(f / g.le[in]);
      } else if (g.dim == 2)
        e[(i + 1) % 2] = (unsigned)f;
      unsigned x[3] = { g.lo[0] + e[0], g.lo[1] + e[1], g.lo[2] + e[2] };
      // periodic neighbours (exchangeOverlap wraps the processor grid)
      x[i] = upper ? (g.lo[i] + g.le[i]) % g.c[i] : (g.lo[i] + g.c[i] - 1) % g.c[i];
      overlap[t - Dl] = slab_value(g, x[0], x[1], x[2]);
      break;
    }
  }
}

// evalToData: the slab of this rank gathered from the blocks of all ranks
__global__ void
k_blocks_to_slab(BlockGeom g, double* __restrict__ slab, size_t Dl)
{
  for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < Dl;
       t += (size_t)gridDim.x * blockDim.x) {
    unsigned x[3] = { 0, 0, 0 };
    if (g.dim == 3) {
      x[0] = (unsigned)(t % g.c[0]);
      x[1] = (unsigned)((t / g.c[0]) % g.c[1]);
      x[2] = (unsigned)(t / ((size_t)g.c[0] * g.c[1])) + g.rank * g.cl;
    } else if (g.dim == 2) {
      x[0] = (unsigned)(t % g.c[0]);
      x[1] = (unsigned)(t / g.c[0]) + g.rank * g.cl;
    } else
      x[0] = (unsigned)t + g.rank * g.cl;
    unsigned b[3], e[3];
    for (int i = 0; i < 3; i++) {
      b[i] = x[i] / g.le[i];
      e[i] = x[i] - b[i] * g.le[i];
    }
    const unsigned owner = b[0] + g.ppd[0] * (b[1] + g.ppd[1] * b[2]);
    slab[t] = g.peer[owner][e[0] + (size_t)g.le[0] * (e[1] + (size_t)g.le[1] * e[2])];
  }
}

int
block_geom(pf_plan* pl, const int* ppd, BlockGeom& g)
{
  if (pl->nranks <= 1 || !pl->connected)
    return fail(PF_ERR_INVALID, "slab <-> block redistribution needs a connected "
                                "multi-GPU plan (pf_slab_connect)");
  memset(&g, 0, sizeof(g));
  g.dim = pl->dim;
  g.P = pl->nranks;
  g.rank = pl->rank;
  g.cl = (unsigned)pl->cl;
  long long prod = 1;
  for (int i = 0; i < 3; i++) {
    g.c[i] = (unsigned)pl->c[i];
    g.ppd[i] = i < pl->dim ? (unsigned)ppd[i] : 1u;
    if (g.ppd[i] < 1 || g.c[i] % g.ppd[i] != 0) // stochastic.hh:239-243
      return fail(PF_ERR_INVALID, "cells in dimension not divisable by numProcs");
    g.le[i] = g.c[i] / g.ppd[i];
    prod *= g.ppd[i];
  }
  if (prod != pl->nranks)
    return fail(PF_ERR_INVALID, "procPerDim does not multiply to the number of ranks");
  // stochastic.hh:244-257
  const unsigned r = (unsigned)pl->rank;
  if (pl->dim == 3) {
    g.lo[0] = (r % (g.ppd[0] * g.ppd[1])) % g.ppd[0] * g.le[0];
    g.lo[1] = (r % (g.ppd[0] * g.ppd[1])) / g.ppd[0] * g.le[1];
    g.lo[2] = r / (g.ppd[0] * g.ppd[1]) * g.le[2];
  } else if (pl->dim == 2) {
    g.lo[0] = r % g.ppd[0] * g.le[0];
    g.lo[1] = r / g.ppd[0] * g.le[1];
  } else
    g.lo[0] = r * g.le[0];
  return 0;
}

// stage `src` (Dl doubles) in this rank's exchange buffer, tell the peers, wait
// for all of them; returns the epoch and fills g.peer[]
int
block_stage(pf_plan* pl, const double* src_dev, BlockGeom& g, unsigned long long* e_out)
{
  const int P = pl->nranks;
  const unsigned long long e = ++pl->epoch;
  const int s = (int)(e & 1);
  if (e > 2) {
    k_slab_wait<<<1, 32, 0, pl->stream>>>(slab_flags(pl, pl->slab_block, 1), P, e - 2,
                                          pl->d_timeout, pl->slab_limit_ns);
    pl->launches++;
  }
  if (pl->Dl * sizeof(double) > pl->xbuf_elems * sizeof(cplx))
    return fail(PF_ERR_INVALID, "redistribution: exchange buffer too small");
  PF_CUDA(cudaMemcpyAsync(slab_buffer(pl, pl->slab_block, s), src_dev,
                          pl->Dl * sizeof(double), cudaMemcpyDeviceToDevice, pl->stream));
  k_slab_signal<<<1, 32, 0, pl->stream>>>(pl->d_flagsA, P, pl->rank, e);
  k_slab_wait<<<1, 32, 0, pl->stream>>>(slab_flags(pl, pl->slab_block, 0), P, e,
                                        pl->d_timeout, pl->slab_limit_ns);
  pl->launches += 2;
  for (int q = 0; q < P; q++)
    g.peer[q] = reinterpret_cast<const double*>(slab_buffer(pl, pl->peer_block[q], s));
  *e_out = e;
  return 0;
}

// multiplyExtended / multiplyRootExtended / multiplyInverseExtended
// (matrix.hh:1459-1662) on device fields; in_dev may equal out_dev
int
apply_dev(pf_plan* pl, int op, const double* in_dev, double* out_dev)
{
  if (pl->nranks > 1)
    return apply_slab_dev(pl, op, in_dev, out_dev);
  double** tabp = op == PF_OP_MATRIX ? &pl->lam
                  : op == PF_OP_ROOT ? &pl->sig
                                     : &pl->inv;
  if (op == PF_OP_MATRIX)
    PF_TRY(ensure_eigen(pl));
  else
    PF_TRY(ensure_table(pl, tabp, op == PF_OP_ROOT ? 1 : 2));
  if (pl->generic)
    return apply_generic(pl, *tabp, in_dev, out_dev);
  if (pl->long2d)
    return apply_long2d(pl, *tabp, in_dev, out_dev);
  PF_TRY(ensure_work(pl));
  const double* tab = *tabp;
  const long long* n = pl->n;
  const long long* c = pl->c;
  const long long m0 = pl->m0;
  const double invN = 1.0 / (double)pl->Next;
  if (pl->dim == 3) {
    PF_TRY(dev_alloc(&pl->V, (size_t)n[1] * m0 * c[2]));
    cplx* Up = pl->Z; // pruned U[c2][m0][c1] fits in Z (c2*c1*m0p)
    PassArgs f0 = blank_args(); // F0: x0 lines, (a,b) = (x1,x2)
    f0.A = (int)c[1];
    f0.B = (int)c[2];
    f0.in = in_dev;
    f0.isa = c[0];
    f0.isb = c[1] * c[0];
    f0.n_valid = (int)c[0];
    f0.out = Up;
    f0.osb = m0 * c[1];
    f0.ose = c[1];
    f0.keep = (int)m0;
    PF_TRY(run_pass(pl, n[0], pf::K_R2C_LOAD, f0));
    PassArgs f1 = blank_args(); // F1: x1 lines, (a,b) = (x2,k0)
    f1.A = (int)c[2];
    f1.B = (int)m0;
    f1.in = Up;
    f1.isa = m0 * c[1];
    f1.isb = c[1];
    f1.n_valid = (int)c[1];
    f1.out = pl->V; // V[k1][k0][x2<c2]
    f1.osb = c[2];
    f1.ose = m0 * c[2];
    f1.keep = (int)n[1];
    PF_TRY(run_pass(pl, n[1], pf::K_C2T_FWD, f1));
    PassArgs s = blank_args(); // F2 * g + A: x2 lines, (a,b) = (k1,k0)
    set_table(pl, s, tab);
    s.A = (int)n[1];
    s.B = (int)m0;
    s.in = pl->V;
    s.isa = m0 * c[2];
    s.isb = c[2];
    s.n_valid = (int)c[2];
    s.scale = invN;
    s.out = pl->Y;
    s.osb = c[2] * n[1];
    s.ose = n[1];
    s.keep = (int)c[2];
    PF_TRY(run_pass(pl, n[2], pf::K_SPEC_TILE, s));
    PF_TRY(inverse_tail(pl, pl->Y, c[2], out_dev, 0));
  } else if (pl->dim == 2) {
    PF_TRY(dev_alloc(&pl->V, (size_t)m0 * c[1]));
    PassArgs f0 = blank_args(); // F0: x0 lines, a = x1
    f0.A = (int)c[1];
    f0.B = 1;
    f0.in = in_dev;
    f0.isa = c[0];
    f0.n_valid = (int)c[0];
    f0.out = pl->V; // U[k0][x1<c1]
    f0.osb = 0;
    f0.ose = c[1];
    f0.keep = (int)m0;
    PF_TRY(run_pass(pl, n[0], pf::K_R2C_LOAD, f0));
    PassArgs s = blank_args(); // x1 lines, a = k0
    set_table(pl, s, tab);
    s.A = (int)m0;
    s.B = 1;
    s.in = pl->V;
    s.isa = c[1];
    s.n_valid = (int)c[1];
    s.scale = invN;
    s.out = pl->Z;
    s.osb = 0;
    s.ose = pl->m0p;
    s.keep = (int)c[1];
    PF_TRY(run_pass(pl, n[1], pf::K_SPEC_TILE, s));
    PF_TRY(inverse_tail(pl, pl->Z, c[1], out_dev, 0));
  } else {
    PassArgs s = blank_args();
    set_table(pl, s, tab);
    s.A = 1;
    s.B = 1;
    s.in = in_dev;
    s.n_valid = (int)c[0];
    s.scale = invN;
    s.out = out_dev;
    s.out_pitch = c[0];
    s.keep = (int)c[0];
    PF_TRY(run_pass(pl, n[0], pf::K_SPEC_REIM, s));
  }
  return 0;
}

int
reduce_dev(pf_plan* pl, size_t nn, const double* x, const double* y,
           const double* z, int mode, double* result_host)
{
  PF_TRY(dev_alloc(&pl->red_partial, RED_BLOCKS + 8));
  const int nb = (int)std::min<size_t>(RED_BLOCKS, (nn + RED_THREADS - 1) / RED_THREADS);
  k_reduce1<<<std::max(nb, 1), RED_THREADS, 0, pl->stream>>>(nn, x, y, z, mode,
                                                           pl->red_partial);
  k_reduce2<<<1, RED_THREADS, 0, pl->stream>>>(std::max(nb, 1), pl->red_partial,
                                               mode, pl->red_partial + RED_BLOCKS);
  if (pl->nranks > 1) {
    k_slab_reduce<<<1, 32, 0, pl->stream>>>(
      pl->d_flagsA, slab_flags(pl, pl->slab_block, 0), pl->nranks, pl->rank,
      ++pl->red_seq, mode, pl->red_partial + RED_BLOCKS, pl->d_timeout,
      pl->slab_limit_ns);
    pl->launches++;
  }
  PF_CUDA(cudaGetLastError());
  pl->launches += 2;
  PF_CUDA(cudaMemcpyAsync(result_host, pl->red_partial + RED_BLOCKS,
                          sizeof(double), cudaMemcpyDeviceToHost, pl->stream));
  PF_CUDA(cudaStreamSynchronize(pl->stream));
  return slab_check(pl);
}

// the same reduction with the result left in device memory (`dst`): no host
// round trip, no synchronisation
int
reduce_to(pf_plan* pl, size_t nn, const double* x, const double* y,
          const double* z, int mode, double* dst)
{
  PF_TRY(dev_alloc(&pl->red_partial, RED_BLOCKS + 8));
  const int nb = (int)std::min<size_t>(RED_BLOCKS, (nn + RED_THREADS - 1) / RED_THREADS);
  k_reduce1<<<std::max(nb, 1), RED_THREADS, 0, pl->stream>>>(nn, x, y, z, mode,
                                                           pl->red_partial);
  k_reduce2<<<1, RED_THREADS, 0, pl->stream>>>(std::max(nb, 1), pl->red_partial,
                                               mode, dst);
  pl->launches += 2;
  if (pl->nranks > 1) {
    k_slab_reduce<<<1, 32, 0, pl->stream>>>(
      pl->d_flagsA, slab_flags(pl, pl->slab_block, 0), pl->nranks, pl->rank,
      ++pl->red_seq, mode, dst, pl->d_timeout, pl->slab_limit_ns);
    pl->launches++;
  }
  PF_CUDA(cudaGetLastError());
  return 0;
}

// staging of host arrays
struct Staged
{
  pf_plan* pl;
  const double* dev = nullptr;
  double* owned = nullptr;
  ~Staged()
  {
    if (owned)
      cudaFree(owned);
  }
};

int
stage_in(pf_plan* pl, const double* p, size_t nn, Staged& s)
{
  s.pl = pl;
  if (is_device_ptr(p)) {
    s.dev = p;
    return 0;
  }
  PF_TRY(dev_alloc(&s.owned, nn));
  PF_CUDA(cudaMemcpyAsync(s.owned, p, nn * sizeof(double),
                          cudaMemcpyHostToDevice, pl->stream));
  s.dev = s.owned;
  return 0;
}

// output (or in/out) array: the caller's pointer when it is device memory,
// otherwise a device scratch that finish() copies back to the host
struct StagedOut
{
  pf_plan* pl = nullptr;
  double* dev = nullptr;
  double* owned = nullptr;
  double* host = nullptr;
  size_t count = 0;
  ~StagedOut()
  {
    if (owned)
      cudaFree(owned);
  }
  int finish()
  {
    if (owned)
      PF_CUDA(cudaMemcpyAsync(host, owned, count * sizeof(double),
                              cudaMemcpyDeviceToHost, pl->stream));
    PF_CUDA(cudaStreamSynchronize(pl->stream));
    return slab_check(pl);
  }
};

int
stage_out(pf_plan* pl, double* p, size_t nn, bool copy_in, StagedOut& s)
{
  s.pl = pl;
  s.count = nn;
  if (is_device_ptr(p)) {
    s.dev = p;
    return 0;
  }
  PF_TRY(dev_alloc(&s.owned, nn));
  s.host = p;
  s.dev = s.owned;
  if (copy_in)
    PF_CUDA(cudaMemcpyAsync(s.owned, p, nn * sizeof(double),
                            cudaMemcpyHostToDevice, pl->stream));
  return 0;
}

int
check_plan(const pf_plan* pl)
{
  if (!pl)
    return fail(PF_ERR_INVALID, "null plan");
  cudaError_t e = cudaSetDevice(pl->device);
  if (e != cudaSuccess)
    return fail(PF_ERR_CUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
  // a plan whose peer wait timed out stays failed (fatal)
  return slab_check(pl);
}

} // namespace

// ===========================================================================
// C ABI
// ===========================================================================
extern "C" {

const char*
pf_version(void)
{
  return "parafields-b200 0.1 (sm_100a)";
}

const char*
pf_last_error(void)
{
  return g_err.c_str();
}

int
pf_device_count(void)
{
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

int
pf_device_info(int device, char* name, int* sm_count, size_t* mem_bytes)
{
  cudaDeviceProp prop;
  PF_CUDA(cudaGetDeviceProperties(&prop, device));
  if (name) {
    strncpy(name, prop.name, 255);
    name[255] = 0;
  }
  if (sm_count)
    *sm_count = prop.multiProcessorCount;
  if (mem_bytes)
    *mem_bytes = prop.totalGlobalMem;
  return 0;
}

void
pf_desc_default(pf_desc* d)
{
  memset(d, 0, sizeof(*d));
  d->dim = 2;
  for (int i = 0; i < 3; i++) {
    d->cells[i] = 1;
    d->extensions[i] = 1.0;
  }
  d->embedding_factor = 2;     // fieldtraits.hh:182
  d->covariance = PF_COV_EXPONENTIAL;
  d->variance = 1.0;
  d->anisotropy = PF_ANISO_NONE;
  d->corr_length[0] = 0.1;
  d->threshold = 1e-14;        // matrix.hh:284
  d->cg_iterations = 100;      // fieldtraits.hh:178
  d->device = -1;
  d->sigmoid_start = 1.;       // matrix.hh:687-688
  d->sigmoid_end = -1.;        // factor - 1, matrix.hh:689-690
  d->merge_recursions = 99;    // matrix.hh:691-692
}

int
pf_plan_create(const pf_desc* desc, pf_plan** out)
{
  if (!desc || !out)
    return fail(PF_ERR_INVALID, "null argument");
  *out = nullptr;
  if (desc->dim < 1 || desc->dim > 3)
    return fail(PF_ERR_INVALID, "dimension must be 1, 2 or 3");
  if (desc->covariance < 0 || desc->covariance > PF_COV_CUSTOM)
    return fail(PF_ERR_INVALID, "covariance structure not known");
  if (desc->covariance == PF_COV_CUSTOM &&
      (desc->periodization != 0 || desc->max_factor != 0 || desc->n_ranks > 1))
    return fail(PF_ERR_INVALID, "custom covariance tables: classical periodisation, "
                                "no embedding-factor retry, one GPU");
  // covariance.hh:547-548 (nu = 0 divides by Gamma(0) there; rejected too)
  if (desc->covariance == PF_COV_MATERN && !(desc->cov_param[0] > 0.))
    return fail(PF_ERR_INVALID, "matern nu has to be positive");
  if (desc->anisotropy < 0 || desc->anisotropy > 2)
    return fail(PF_ERR_INVALID,
                "stochastic.anisotropy must be \"none\", \"axiparallel\" or "
                "\"geometric\"");
  if (desc->transform < 0 || desc->transform > 4)
    return fail(PF_ERR_INVALID, "transform type not known");
  // matrix.hh:530-590, 671-680, 715-718
  if (desc->periodization < 0 || desc->periodization > 1)
    return fail(PF_ERR_INVALID,
                "embedding.periodization must be \"classical\" or \"merge\" "
                "(fold and cofold need GSL quadrature and are not built)");
  if (desc->periodization == 1) {
    if (desc->sigmoid < 0 || desc->sigmoid > 1)
      return fail(PF_ERR_INVALID,
                  "embedding.sigmoid must be \"smooth\" or \"smoothstep\"");
    if (desc->sigmoid_combine < 0 || desc->sigmoid_combine > 2)
      return fail(PF_ERR_INVALID,
                  "sigmoidCombine must be tensor, radial or radialPlus");
    if (merge_recursions(*desc, desc->dim) == ~0u)
      return fail(PF_ERR_INVALID, "magic recursions value 99 not implemented "
                                  "for choice of covariance");
  }
  // covariance.hh argument checks
  if (desc->covariance == PF_COV_CUBIC && desc->dim == 3)
    return fail(PF_ERR_INVALID, "cubic covariance only applicable in 1D or 2D");
  if (desc->covariance == PF_COV_GAMMA_EXPONENTIAL &&
      (desc->cov_param[0] < 0. || desc->cov_param[0] > 2.))
    return fail(PF_ERR_INVALID, "exponent gamma has to be between 0 and 2");
  if (desc->covariance == PF_COV_GENERALIZED_CAUCHY) {
    if (desc->cov_param[0] <= 0. || desc->cov_param[0] > 2.)
      return fail(PF_ERR_INVALID,
                  "generalized Cauchy alpha has to be in range (0,2]");
    if (desc->cov_param[1] <= 0.)
      return fail(PF_ERR_INVALID, "generalized Cauchy beta has to be positive");
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return fail(PF_ERR_CUDA, "no CUDA device available (there is no CPU "
                             "fallback in libparafields_cuda)");
  }
  pf_plan* pl = new pf_plan();
  pl->d = *desc;
  pl->dim = desc->dim;
  if (desc->device >= 0)
    pl->device = desc->device;
  else
    cudaGetDevice(&pl->device);
  if (cudaSetDevice(pl->device) != cudaSuccess) {
    delete pl;
    return fail(PF_ERR_CUDA, "cudaSetDevice failed");
  }
  cudaDeviceGetAttribute(&pl->sm_count, cudaDevAttrMultiProcessorCount,
                         pl->device);
  if (const char* env = getenv("PF_PIPE"))
    pl->pipe_mask = (unsigned)strtoul(env, nullptr, 0);
  if (const char* env = getenv("PF_LONG_PIPE"))
    pl->long_pipe = atoi(env) != 0;
  if (const char* env = getenv("PF_GEN3")) // "force": skip the size heuristic
    pl->gen3_force = std::string(env) == "force";
  if (const char* env = getenv("PF_SPLIT")) // SMs for pass A in the one-GPU pipeline
    pl->split_one = atoi(env);
  if (const char* env = getenv("PF_STREAM")) // 0: three-pass route everywhere
    pl->stream_on = atoi(env) != 0;
  if (const char* env = getenv("PF_RING")) // planes in the L2-resident ring
    pl->stream_slots = std::max(2, std::min(64, atoi(env)));
  if (const char* env = getenv("PF_LAG")) // planes pass B runs behind pass A
    pl->stream_lag = std::max(1, std::min(32, atoi(env)));
  if (const char* env = getenv("PF_ZT"))
    pl->stream_zt = atoi(env);
  pl->nranks = desc->n_ranks > 1 ? desc->n_ranks : 1;
  pl->rank = pl->nranks > 1 ? desc->rank : 0;
  if (pl->nranks > 8 || pl->rank < 0 || pl->rank >= pl->nranks) {
    delete pl;
    return fail(PF_ERR_INVALID, "slab mode supports 1..8 ranks, 0 <= rank < n_ranks");
  }
  if (pl->nranks > 1 && desc->max_factor) {
    delete pl;
    return fail(PF_ERR_INVALID, "the embedding-factor retry is not available in slab mode");
  }
  pl->factor = desc->embedding_factor ? desc->embedding_factor : 2;
  if (desc->periodic)
    pl->factor = 1; // fieldtraits.hh:196-202
  int rc = plan_resize(pl);
  if (rc != 0) {
    delete pl;
    return rc;
  }
  if (pl->nranks > 1) {
    // exchange block: two buffers + flag arrays, one cudaMalloc so that one
    // IPC handle describes it
    // [buffer 0][buffer 1][flags A, B, reduction slots: 512 bytes][forward
    // exchange region of the covariance operators]
    const size_t bytes = 2 * pl->xbuf_elems * sizeof(cplx) + 512 +
                         pl->fwd_elems * sizeof(cplx) +
                         2 * pl->zt_elems * sizeof(cplx);
    void* blk = nullptr;
    if (cudaMalloc(&blk, bytes) != cudaSuccess ||
        cudaMemset(blk, 0, bytes) != cudaSuccess ||
        cudaMalloc((void**)&pl->d_flagsA, 8 * sizeof(void*)) != cudaSuccess ||
        cudaMalloc((void**)&pl->d_flagsB, 8 * sizeof(void*)) != cudaSuccess ||
        cudaHostAlloc((void**)&pl->h_timeout, sizeof(int),
                      cudaHostAllocMapped) != cudaSuccess ||
        cudaHostGetDevicePointer((void**)&pl->d_timeout, pl->h_timeout, 0) !=
          cudaSuccess) {
      cudaGetLastError();
      delete pl;
      return fail(PF_ERR_NOMEM, "slab mode: allocating the exchange block failed");
    }
    *pl->h_timeout = 0;
    if (const char* env = getenv("PF_SLAB_TIMEOUT_S")) {
      const double sec = atof(env);
      if (sec > 0.)
        pl->slab_limit_ns = (unsigned long long)(sec * 1e9);
    }
    pl->slab_block = static_cast<unsigned char*>(blk);
    pl->peer_block[pl->rank] = blk;
    // measured on B200 (profiles/r01q, r01r): from 4 ranks on, stores to
    // peer memory straight from pass A, overlapped with the previous tail on
    // the other SMs, beat the staged DMA exchange; with 2 ranks the DMA
    // exchange wins
    pl->slab_dma = pl->nranks < 4;
    if (const char* env = getenv("PF_SLAB_EXCHANGE"))
      pl->slab_dma = strcmp(env, "peer") != 0;
    for (int i = 0; i < 2; i++) {
      cudaEventCreateWithFlags(&pl->ev_a[i], cudaEventDisableTiming);
      cudaEventCreateWithFlags(&pl->ev_x[i], cudaEventDisableTiming);
    }
    cudaDeviceSynchronize();
  }
  cudaStreamCreateWithFlags(&pl->copy_stream, cudaStreamNonBlocking);
  if (pl->nranks > 1) {
    cudaStreamCreateWithFlags(&pl->tail_stream, cudaStreamNonBlocking);
    cudaEventCreateWithFlags(&pl->ev_fork, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&pl->ev_join, cudaEventDisableTiming);
    // default split for the kernel exchange: pass A is NVLink-bound, the tail
    // HBM-bound; about 55 % of the SMs for pass A balances the two on B200
    // (the tail needs SM time ~ 1/P, the NVLink-bound pass A a time ~
    // (P-1)/P^2: about 2/3 of the SMs for pass A at 4 ranks, 4/7 at 8)
    if (!pl->slab_dma)
      pl->split_a = (pl->nranks >= 8 ? pl->sm_count * 4 / 7
                                     : pl->sm_count * 2 / 3) & ~1;
    // streaming route: the fused kernel is long, the last axis short; they
    // simply alternate on one stream (kp_ab of realization r+1, then kp_c2rt of
    // realization r, whose inputs have long arrived)
    if (pl->stream_slab)
      pl->split_a = 0;
    if (const char* env = getenv("PF_SLAB_SPLIT"))
      pl->split_a = std::max(0, std::min(pl->sm_count - 8, atoi(env)));
    pl->peer_streams = pl->nranks - 1;
    if (const char* env = getenv("PF_SLAB_STREAMS"))
      pl->peer_streams = std::max(1, std::min(pl->nranks - 1, atoi(env)));
    if (pl->peer_streams > 1) {
      cudaEventCreateWithFlags(&pl->ev_gate, cudaEventDisableTiming);
      for (int i = 0; i < pl->peer_streams; i++) {
        cudaStreamCreateWithFlags(&pl->peer_stream[i], cudaStreamNonBlocking);
        cudaEventCreateWithFlags(&pl->ev_peer[i], cudaEventDisableTiming);
      }
    }
  }
  for (int i = 0; i < 2; i++) {
    cudaEventCreateWithFlags(&pl->ev_done[i], cudaEventDisableTiming);
    cudaEventCreateWithFlags(&pl->ev_copied[i], cudaEventDisableTiming);
  }
  *out = pl;
  return 0;
}

int
pf_plan_destroy(pf_plan* pl)
{
  if (!pl)
    return 0;
  cudaSetDevice(pl->device);
  cudaDeviceSynchronize();
  prof_collect(pl);
  free_sized(pl);
  for (auto& kv : pl->tw)
    cudaFree(kv.second);
  for (auto& kv : pl->twl)
    cudaFree(kv.second);
  for (auto& kv : pl->gen_tw)
    cudaFree(kv.second);
  for (auto& kv : pl->gen_tabs) {
    cudaFree(kv.second.chirp);
    cudaFree(kv.second.bhat);
  }
  dev_free(&pl->dev_io[0]);
  dev_free(&pl->dev_io[1]);
  dev_free(&pl->axis_tab);
  dev_free(&pl->eval_flag);
  dev_free(&pl->dev_in);
  for (int i = 0; i < 6; i++)
    dev_free(&pl->cg[i]);
  dev_free(&pl->red_partial);
  dev_free(&pl->cov_table);
  dev_free(&pl->cg_scalars);
  if (pl->cg_host)
    cudaFreeHost(pl->cg_host);
  for (int i = 0; i < 4; i++)
    if (pl->cg_ev[i])
      cudaEventDestroy(pl->cg_ev[i]);
  dev_free(&pl->census);
  for (int q = 0; q < pl->nranks; q++)
    if (q != pl->rank && pl->peer_block[q])
      cudaIpcCloseMemHandle(pl->peer_block[q]);
  if (pl->slab_block)
    cudaFree(pl->slab_block);
  for (int i = 0; i < 2; i++) {
    dev_free(&pl->send_stage[i]);
    if (pl->ev_a[i])
      cudaEventDestroy(pl->ev_a[i]);
    if (pl->ev_x[i])
      cudaEventDestroy(pl->ev_x[i]);
  }
  if (pl->d_flagsA)
    cudaFree(pl->d_flagsA);
  if (pl->d_flagsB)
    cudaFree(pl->d_flagsB);
  if (pl->h_timeout)
    cudaFreeHost(pl->h_timeout);
  if (pl->copy_stream)
    cudaStreamDestroy(pl->copy_stream);
  if (pl->d2h_stream)
    cudaStreamDestroy(pl->d2h_stream);
  if (pl->tail_stream)
    cudaStreamDestroy(pl->tail_stream);
  if (pl->ev_fork)
    cudaEventDestroy(pl->ev_fork);
  if (pl->ev_join)
    cudaEventDestroy(pl->ev_join);
  for (int i = 0; i < 8; i++) {
    if (pl->peer_stream[i])
      cudaStreamDestroy(pl->peer_stream[i]);
    if (pl->ev_peer[i])
      cudaEventDestroy(pl->ev_peer[i]);
  }
  if (pl->ev_gate)
    cudaEventDestroy(pl->ev_gate);
  for (int i = 0; i < 2; i++) {
    if (pl->ev_done[i])
      cudaEventDestroy(pl->ev_done[i]);
    if (pl->ev_copied[i])
      cudaEventDestroy(pl->ev_copied[i]);
  }
  delete pl;
  return 0;
}

int
pf_plan_set_stream(pf_plan* pl, void* stream)
{
  PF_TRY(check_plan(pl));
  pl->stream = reinterpret_cast<cudaStream_t>(stream);
  return 0;
}

int
pf_plan_synchronize(pf_plan* pl)
{
  PF_TRY(check_plan(pl));
  PF_CUDA(cudaStreamSynchronize(pl->stream));
  return slab_check(pl);
}

unsigned long long
pf_plan_launch_count(const pf_plan* pl)
{
  return pl ? pl->launches : 0;
}

int
pf_plan_set_profiling(pf_plan* pl, int on)
{
  PF_TRY(check_plan(pl));
  prof_collect(pl);
  pl->profiling = on != 0;
  for (int i = 0; i < 16; i++) {
    pl->prof_ms[i] = 0.;
    pl->prof_count[i] = 0;
  }
  return 0;
}

int
pf_plan_get_pass_times(pf_plan* pl, double* ms, unsigned long long* counts)
{
  PF_TRY(check_plan(pl));
  prof_collect(pl);
  for (int i = 0; i < 16; i++) {
    if (ms)
      ms[i] = pl->prof_ms[i];
    if (counts)
      counts[i] = pl->prof_count[i];
  }
  return 0;
}

size_t
pf_domain_size(const pf_plan* pl)
{
  return pl ? pl->D : 0;
}
size_t
pf_local_domain_size(const pf_plan* pl)
{
  return pl ? pl->Dl : 0;
}

int
pf_slab_handle(pf_plan* pl, void* handle_out)
{
  PF_TRY(check_plan(pl));
  static_assert(sizeof(cudaIpcMemHandle_t) == PF_SLAB_HANDLE_BYTES,
                "IPC handle size");
  if (pl->nranks <= 1 || !handle_out)
    return fail(PF_ERR_INVALID, "pf_slab_handle: plan is not in slab mode");
  cudaIpcMemHandle_t h;
  PF_CUDA(cudaIpcGetMemHandle(&h, pl->slab_block));
  memcpy(handle_out, &h, sizeof(h));
  return 0;
}

int
pf_slab_connect(pf_plan* pl, const void* handles, int count)
{
  PF_TRY(check_plan(pl));
  if (pl->nranks <= 1 || count != pl->nranks || !handles)
    return fail(PF_ERR_INVALID, "pf_slab_connect: need one handle per rank");
  const unsigned char* hb = static_cast<const unsigned char*>(handles);
  for (int q = 0; q < pl->nranks; q++) {
    if (q == pl->rank || pl->peer_block[q])
      continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, hb + (size_t)q * PF_SLAB_HANDLE_BYTES, sizeof(h));
    void* ptr = nullptr;
    PF_CUDA(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
    pl->peer_block[q] = ptr;
  }
  unsigned long long* fa[8] = { nullptr };
  unsigned long long* fb[8] = { nullptr };
  for (int q = 0; q < pl->nranks; q++) {
    fa[q] = slab_flags(pl, pl->peer_block[q], 0);
    fb[q] = slab_flags(pl, pl->peer_block[q], 1);
  }
  PF_CUDA(cudaMemcpy(pl->d_flagsA, fa, sizeof(fa), cudaMemcpyHostToDevice));
  PF_CUDA(cudaMemcpy(pl->d_flagsB, fb, sizeof(fb), cudaMemcpyHostToDevice));
  pl->connected = true;
  return 0;
}

size_t
pf_extended_size(const pf_plan* pl)
{
  return pl ? pl->Next : 0;
}
size_t
pf_noise_entries(const pf_plan* pl)
{
  if (!pl)
    return 0;
  return (size_t)pl->m0 * pl->n[1] * pl->n[2];
}

int
pf_set_covariance_table(pf_plan* pl, const double* values, size_t count)
{
  PF_TRY(check_plan(pl));
  if (pl->d.covariance != PF_COV_CUSTOM)
    return fail(PF_ERR_INVALID, // matrix.hh:272-276
                "you can't fill the matrix explicitly if a default covariance was chosen");
  if (!values || count != pl->Next)
    return fail(PF_ERR_INVALID, "pf_set_covariance_table: need one value per cell of the extended domain");
  PF_CUDA(cudaStreamSynchronize(pl->stream));
  PF_TRY(dev_alloc(&pl->cov_table, pl->Next));
  PF_CUDA(cudaMemcpyAsync(pl->cov_table, values, pl->Next * sizeof(double),
                          cudaMemcpyDefault, pl->stream));
  PF_CUDA(cudaStreamSynchronize(pl->stream));
  // eigenvalues and everything derived from them are stale
  pl->eig_valid = false;
  dev_free(&pl->sig_gen);
  dev_free(&pl->sig);
  dev_free(&pl->inv);
  return 0;
}

int
pf_compute_eigenvalues(pf_plan* pl, pf_eig_stats* stats)
{
  PF_TRY(check_plan(pl));
  int rc = ensure_eigen(pl);
  if (stats)
    *stats = pl->stats;
  return rc;
}

size_t
pf_eigenvalue_count(const pf_plan* pl, int layout)
{
  if (!pl)
    return 0;
  size_t r = 1;
  for (int i = 0; i < pl->dim; i++) {
    if (layout == PF_LAMBDA_NATURAL || !pl->compact)
      r *= (i == 0 ? pl->m0 : pl->n[i]);
    else
      r *= pl->n[i] / 2 + 1;
  }
  return r;
}

int
pf_get_eigenvalues(pf_plan* pl, double* out, size_t count, int layout)
{
  PF_TRY(check_plan(pl));
  PF_TRY(ensure_eigen(pl));
  const size_t total = pf_eigenvalue_count(pl, layout);
  if (count < total)
    return fail(PF_ERR_INVALID, "eigenvalue buffer too small");
  GatherArgs g;
  memset(&g, 0, sizeof(g));
  g.dim = pl->dim;
  for (int i = 0; i < 3; i++) {
    g.n[i] = pl->n[i];
    g.ob[i] = 1;
  }
  for (int i = 0; i < pl->dim; i++) {
    if (layout == PF_LAMBDA_NATURAL || !pl->compact)
      g.ob[i] = (i == 0 ? pl->m0 : pl->n[i]);
    else
      g.ob[i] = pl->n[i] / 2 + 1;
  }
  // fftw.transposed: storage has the last two dims swapped in 3D
  // (dctmatrixbackend.hh:221-231, r2cmatrixbackend.hh transposeIfNeeded)
  g.transposed = (layout == PF_LAMBDA_REFERENCE && pl->dim == 3) ? 1 : 0;
  g.compact = pl->compact ? 1 : 0;
  g.m0 = pl->m0;
  g.P = pl->P;
  const bool dev = is_device_ptr(out);
  double* dst = out;
  double* tmp = nullptr;
  if (!dev) {
    PF_TRY(dev_alloc(&tmp, total));
    dst = tmp;
  }
  if (pl->generic)
    pf::k_generic_gather<<<grid_for(total), 256, 0, pl->stream>>>(
      gen_geom(pl), pl->lam, dst, g.ob[0], g.ob[1], g.ob[2], g.transposed,
      (long long)total);
  else
    k_gather_lambda<<<(unsigned)((total + 255) / 256), 256, 0, pl->stream>>>(
      pl->lam, dst, g, total);
  cudaError_t e = cudaGetLastError();
  pl->launches++;
  if (e == cudaSuccess && !dev)
    e = cudaMemcpyAsync(out, tmp, total * sizeof(double),
                        cudaMemcpyDeviceToHost, pl->stream);
  if (e == cudaSuccess && !dev)
    e = cudaStreamSynchronize(pl->stream);
  if (tmp)
    cudaFree(tmp);
  if (e != cudaSuccess)
    return fail(PF_ERR_CUDA, cudaGetErrorString(e));
  return 0;
}

// does pf_generate_batch pipeline pass A against the tails on one GPU?  Only
// where pass A is the grouped persistent kernel (run_pass: 3D, lines of 1024,
// many waves), because the overlap works by capping persistent grids
static bool
overlap_one(const pf_plan* pl)
{
  if (pl->nranks > 1 || pl->dim != 3 || pl->n[2] != 1024 || pl->gen3_force)
    return false;
  if (pl->split_one < 8 || pl->split_one > pl->sm_count - 8)
    return false;
  if (!(pl->pipe_mask & (1u << pf::K_GEN3)) || (pl->pipe_mask & 1u))
    return false;
  return (long long)((pl->n[1] + 3) / 4) * pl->m0 >= 8LL * pl->sm_count;
}

int
pf_generate_batch(pf_plan* pl, uint64_t seed, uint64_t first, size_t count,
                  double* out, int apply_transform)
{
  PF_TRY(check_plan(pl));
  if (!out)
    return fail(PF_ERR_INVALID, "null output");
  const int tr = apply_transform ? pl->d.transform : 0;
  const int nk = pl->d.rng_mode == 1 ? pf::NOISE_PHILOX_F64 : pf::NOISE_PHILOX_FAST;
  const bool dev = is_device_ptr(out);
  if (pl->dim == 1) {
    // DFT backend: one transform yields two fields (matrix.hh:428-437).  The
    // C ABI is stateless: realization r is the real (r even) or imaginary
    // (r odd) part of transform r / 2; the reference's "every second call
    // returns the spare field" protocol lives in the host plugin
    // (cudamatrix.hh).  The plan merely caches the other half of the last
    // transform so that the pair (2t, 2t + 1) costs one FFT.
    PF_TRY(dev_alloc(&pl->spare, pl->Dl));
    PF_TRY(dev_alloc(&pl->dev_io[0], pl->Dl));
    PF_TRY(dev_alloc(&pl->dev_io[1], pl->Dl));
    for (size_t r = 0; r < count; r++) {
      const uint64_t idx = first + r, t = idx >> 1;
      const bool imag = (idx & 1) != 0;
      double* o = out + r * pl->Dl;
      double* dst = dev ? o : pl->dev_io[0];
      const bool cached = pl->spare_valid && pl->spare_seed == seed &&
                          pl->spare_t == t && pl->spare_rng == nk;
      if (!imag) {
        // the real part; the (untransformed) imaginary part goes to the cache
        PF_TRY(generate_dev(pl, nk, seed, t, nullptr, dst, pl->spare, tr));
        pl->spare_valid = true;
        pl->spare_seed = seed;
        pl->spare_t = t;
        pl->spare_rng = nk;
      } else {
        if (!cached) {
          PF_TRY(generate_dev(pl, nk, seed, t, nullptr, pl->dev_io[1], pl->spare, 0));
          pl->spare_valid = true;
          pl->spare_seed = seed;
          pl->spare_t = t;
          pl->spare_rng = nk;
        }
        if (tr) {
          k_transform<<<grid_for(pl->Dl), 256, 0, pl->stream>>>(
            pl->Dl, pl->spare, dst, tr, pl->d.boxcox_lambda);
          pl->launches++;
        } else
          PF_CUDA(cudaMemcpyAsync(dst, pl->spare, pl->Dl * sizeof(double),
                                  cudaMemcpyDeviceToDevice, pl->stream));
      }
      if (!dev) {
        PF_CUDA(cudaMemcpyAsync(o, dst, pl->Dl * sizeof(double),
                                cudaMemcpyDeviceToHost, pl->stream));
        PF_CUDA(cudaStreamSynchronize(pl->stream));
      }
    }
    return 0;
  }
  if (pl->nranks > 1) {
    // slab mode, software pipelined: pass A (and the start of the exchange)
    // of realization r+1 is issued before the tail of realization r, so the
    // NVLink transfer overlaps compute.  Host output: the tails write two
    // device buffers in turn and a third stream copies them out, so the D2H
    // transfer of realization r overlaps the kernels of realization r+1.
    unsigned long long e_prev = 0, e = 0;
    // kernel exchange with an SM split: pass A stays on the plan's stream,
    // the tails go to a second stream and both run on their share of the SMs
    const bool split = pl->split_a > 0 && !pl->slab_dma && count > 1;
    cudaStream_t main_stream = pl->stream;
    if (!dev) {
      PF_TRY(dev_alloc(&pl->dev_io[0], pl->Dl));
      if (count > 1)
        PF_TRY(dev_alloc(&pl->dev_io[1], pl->Dl));
      if (!pl->d2h_stream)
        PF_CUDA(cudaStreamCreateWithFlags(&pl->d2h_stream, cudaStreamNonBlocking));
    }
    if (split) {
      PF_CUDA(cudaEventRecord(pl->ev_fork, main_stream));
      PF_CUDA(cudaStreamWaitEvent(pl->tail_stream, pl->ev_fork, 0));
    }
    auto tail = [&](unsigned long long ep, size_t r) -> int {
      cudaStream_t ts = split ? pl->tail_stream : main_stream;
      const int b = (int)(r & 1);
      double* o = dev ? out + r * pl->Dl : pl->dev_io[b];
      if (!dev && r >= 2) // the copy of realization r-2 has left this buffer
        PF_CUDA(cudaStreamWaitEvent(ts, pl->ev_copied[b], 0));
      if (split) {
        pl->stream = pl->tail_stream;
        pl->sm_cap = pl->sm_count - pl->split_a;
      }
      const int rc = slab_issue_tail(pl, ep, o, tr, pl->stream_slab);
      pl->stream = main_stream;
      pl->sm_cap = 0;
      PF_TRY(rc);
      if (!dev) {
        PF_CUDA(cudaEventRecord(pl->ev_done[b], ts));
        PF_CUDA(cudaStreamWaitEvent(pl->d2h_stream, pl->ev_done[b], 0));
        PF_CUDA(cudaMemcpyAsync(out + r * pl->Dl, o, pl->Dl * sizeof(double),
                                cudaMemcpyDeviceToHost, pl->d2h_stream));
        PF_CUDA(cudaEventRecord(pl->ev_copied[b], pl->d2h_stream));
      }
      return 0;
    };
    for (size_t r = 0; r < count; r++) {
      pl->sm_cap = split ? pl->split_a : 0;
      const int rc = slab_issue_a(pl, nk, seed, first + r, nullptr, &e);
      pl->sm_cap = 0;
      PF_TRY(rc);
      if (r > 0)
        PF_TRY(tail(e_prev, r - 1));
      e_prev = e;
    }
    if (count > 0)
      PF_TRY(tail(e_prev, count - 1));
    if (split) {
      PF_CUDA(cudaEventRecord(pl->ev_join, pl->tail_stream));
      PF_CUDA(cudaStreamWaitEvent(main_stream, pl->ev_join, 0));
    }
    if (!dev) {
      PF_CUDA(cudaStreamSynchronize(pl->d2h_stream));
      PF_CUDA(cudaStreamSynchronize(main_stream));
      return slab_check(pl);
    }
    return 0;
  }
  if (dev && count > 1 && overlap_one(pl)) {
    // one GPU, 3D: pass A of realization r+1 (dispatch-bound, split_one SMs,
    // plan stream) runs beside the tail of realization r (HBM-bound, the other
    // SMs, tail stream); two Y buffers, events order producer and consumer
    PF_TRY(ensure_work(pl));
    PF_TRY(dev_alloc(&pl->Y2, (size_t)pl->m0 * pl->c[2] * pl->n[1]));
    if (!pl->tail_stream) {
      PF_CUDA(cudaStreamCreateWithFlags(&pl->tail_stream, cudaStreamNonBlocking));
      PF_CUDA(cudaEventCreateWithFlags(&pl->ev_fork, cudaEventDisableTiming));
      PF_CUDA(cudaEventCreateWithFlags(&pl->ev_join, cudaEventDisableTiming));
    }
    for (int i = 0; i < 2; i++) {
      if (!pl->ev_a[i])
        PF_CUDA(cudaEventCreateWithFlags(&pl->ev_a[i], cudaEventDisableTiming));
      if (!pl->ev_x[i])
        PF_CUDA(cudaEventCreateWithFlags(&pl->ev_x[i], cudaEventDisableTiming));
    }
    cudaStream_t main_stream = pl->stream;
    PF_CUDA(cudaEventRecord(pl->ev_fork, main_stream));
    PF_CUDA(cudaStreamWaitEvent(pl->tail_stream, pl->ev_fork, 0));
    int rc = 0;
    for (size_t r = 0; r < count && rc == 0; r++) {
      const int b = (int)(r & 1);
      cplx* Yb = b ? pl->Y2 : pl->Y;
      if (r >= 2) // the tail of realization r-2 has read this buffer
        PF_CUDA(cudaStreamWaitEvent(main_stream, pl->ev_x[b], 0));
      pl->y_override = Yb;
      pl->gen_phase = 1;
      pl->sm_cap = r == 0 ? 0 : pl->split_one; // nothing runs beside the first
      rc = generate_dev(pl, nk, seed, first + r, nullptr, nullptr, nullptr, tr);
      if (rc == 0 && cudaEventRecord(pl->ev_a[b], main_stream) == cudaSuccess &&
          cudaStreamWaitEvent(pl->tail_stream, pl->ev_a[b], 0) == cudaSuccess) {
        pl->stream = pl->tail_stream;
        pl->gen_phase = 2;
        pl->sm_cap = r + 1 == count ? 0 : pl->sm_count - pl->split_one;
        rc = generate_dev(pl, nk, seed, first + r, nullptr, out + r * pl->Dl,
                          nullptr, tr);
        if (rc == 0 && cudaEventRecord(pl->ev_x[b], pl->tail_stream) != cudaSuccess)
          rc = fail(PF_ERR_CUDA, "event record failed");
      } else if (rc == 0)
        rc = fail(PF_ERR_CUDA, "event record failed");
      pl->stream = main_stream;
      pl->y_override = nullptr;
      pl->gen_phase = 0;
      pl->sm_cap = 0;
    }
    PF_CUDA(cudaEventRecord(pl->ev_join, pl->tail_stream));
    PF_CUDA(cudaStreamWaitEvent(main_stream, pl->ev_join, 0));
    return rc;
  }
  if (dev && pl->dim == 2 && !pl->long2d && !pl->generic && count > 1 && pl->c[1] * (long long)count < 0x7fffffffLL)
    return generate2d_batch(pl, nk, seed, first, count, out, tr);
  if (dev && pl->dim == 2 && pl->long2d && !pl->generic && nk != pf::NOISE_INJECTED) {
    // long lines: step 1 of pass A serves up to four realizations from one read
    // of the sigma table (BASELINE config K4 is a batch of realizations)
    int rc = 0;
    for (size_t r = 0; r < count && rc == 0; r += 4) {
      pl->long_batch = (int)std::min<size_t>(4, count - r);
      rc = generate_dev(pl, nk, seed, first + r, nullptr, out + r * pl->Dl, nullptr, tr);
    }
    pl->long_batch = 1;
    return rc;
  }
  if (dev) {
    for (size_t r = 0; r < count; r++)
      PF_TRY(generate_dev(pl, nk, seed, first + r, nullptr, out + r * pl->Dl,
                          nullptr, tr));
    return 0;
  }
  // host output: double-buffered so that the D2H copy of realization r
  // overlaps the kernels of realization r+1 (needs pinned host memory to
  // overlap; pageable memory still works, serialised by the runtime)
  PF_TRY(dev_alloc(&pl->dev_io[0], pl->Dl));
  if (count > 1)
    PF_TRY(dev_alloc(&pl->dev_io[1], pl->Dl));
  for (size_t r = 0; r < count; r++) {
    const int s = (int)(r & 1);
    if (r >= 2)
      PF_CUDA(cudaStreamWaitEvent(pl->stream, pl->ev_copied[s], 0));
    PF_TRY(generate_dev(pl, nk, seed, first + r, nullptr, pl->dev_io[s], nullptr, tr));
    PF_CUDA(cudaEventRecord(pl->ev_done[s], pl->stream));
    PF_CUDA(cudaStreamWaitEvent(pl->copy_stream, pl->ev_done[s], 0));
    PF_CUDA(cudaMemcpyAsync(out + r * pl->Dl, pl->dev_io[s],
                            pl->Dl * sizeof(double), cudaMemcpyDeviceToHost,
                            pl->copy_stream));
    PF_CUDA(cudaEventRecord(pl->ev_copied[s], pl->copy_stream));
  }
  PF_CUDA(cudaStreamSynchronize(pl->copy_stream));
  PF_CUDA(cudaStreamSynchronize(pl->stream));
  return slab_check(pl);
}

int
pf_generate(pf_plan* pl, uint64_t seed, uint64_t realization, double* out,
            int apply_transform)
{
  return pf_generate_batch(pl, seed, realization, 1, out, apply_transform);
}

int
pf_generate_from_noise(pf_plan* pl, const double* noise, double* out,
                       double* spare)
{
  PF_TRY(check_plan(pl));
  if (!noise || !out)
    return fail(PF_ERR_INVALID, "null argument");
  const size_t ne = pf_noise_entries(pl);
  Staged sn;
  PF_TRY(stage_in(pl, noise, 2 * ne, sn));
  const bool dev = is_device_ptr(out);
  const bool sdev = is_device_ptr(spare);
  PF_TRY(dev_alloc(&pl->dev_io[0], pl->Dl));
  double* dst = dev ? out : pl->dev_io[0];
  double* sp = nullptr;
  if (pl->dim == 1) {
    PF_TRY(dev_alloc(&pl->spare, pl->Dl));
    sp = (spare && sdev) ? spare : pl->spare;
    pl->spare_valid = false; // the Philox cache no longer describes pl->spare
  }
  PF_TRY(generate_dev(pl, pf::NOISE_INJECTED, 0, 0,
                      reinterpret_cast<const double2*>(sn.dev), dst, sp, 0));
  if (!dev)
    PF_CUDA(cudaMemcpyAsync(out, dst, pl->Dl * sizeof(double),
                            cudaMemcpyDeviceToHost, pl->stream));
  if (spare && !sdev && pl->dim == 1)
    PF_CUDA(cudaMemcpyAsync(spare, sp, pl->Dl * sizeof(double),
                            cudaMemcpyDeviceToHost, pl->stream));
  PF_CUDA(cudaStreamSynchronize(pl->stream));
  return slab_check(pl);
}

int
pf_generate_uncorrelated(pf_plan* pl, uint64_t seed, double* out)
{
  PF_TRY(check_plan(pl));
  const bool dev = is_device_ptr(out);
  PF_TRY(dev_alloc(&pl->dev_io[0], pl->Dl));
  double* dst = dev ? out : pl->dev_io[0];
  const size_t q = (pl->Dl + 3) / 4;
  // slab mode: the Philox counter is the global cell-quad index
  k_uncorrelated<<<(unsigned)((q + 255) / 256), 256, 0, pl->stream>>>(
    pl->Dl, seed, (unsigned long long)pl->rank * q, dst);
  PF_CUDA(cudaGetLastError());
  pl->launches++;
  if (!dev) {
    PF_CUDA(cudaMemcpyAsync(out, dst, pl->Dl * sizeof(double),
                            cudaMemcpyDeviceToHost, pl->stream));
    PF_CUDA(cudaStreamSynchronize(pl->stream));
  }
  return 0;
}

int
pf_set_variance_as_field(pf_plan* pl, double* out)
{
  PF_TRY(check_plan(pl));
  if (is_device_ptr(out)) {
    k_fill<<<grid_for(pl->Dl), 256, 0, pl->stream>>>(pl->Dl, pl->d.variance, out);
    PF_CUDA(cudaGetLastError());
    pl->launches++;
  } else {
    for (size_t i = 0; i < pl->Dl; i++)
      out[i] = pl->d.variance;
  }
  return 0;
}

int
pf_apply(pf_plan* pl, int op, const double* in, double* out)
{
  PF_TRY(check_plan(pl));
  if (op < 0 || op > 2 || !in || !out)
    return fail(PF_ERR_INVALID, "bad argument");
  // slab mode: `in` / `out` are this rank's layers (Dl values)
  Staged si;
  PF_TRY(stage_in(pl, in, pl->Dl, si));
  const bool dev = is_device_ptr(out);
  PF_TRY(dev_alloc(&pl->dev_io[0], pl->Dl));
  double* dst = dev ? out : pl->dev_io[0];
  PF_TRY(apply_dev(pl, op, si.dev, dst));
  if (!dev)
    PF_CUDA(cudaMemcpyAsync(out, dst, pl->Dl * sizeof(double),
                            cudaMemcpyDeviceToHost, pl->stream));
  PF_CUDA(cudaStreamSynchronize(pl->stream));
  return slab_check(pl);
}

// Matrix::multiplyInverse + innerCG (matrix.hh:198-215,1331-1448).  Vectors AND
// the scalars of the recurrence stay in HBM.
int
pf_solve(pf_plan* pl, const double* in, double* out, unsigned* iterations)
{
  PF_TRY(check_plan(pl));
  if (!in || !out)
    return fail(PF_ERR_INVALID, "null argument");
  // slab mode: vectors are this rank's layers, scalar products are combined
  // over the ranks in rank order (k_slab_reduce)
  const size_t D = pl->Dl;
  Staged si;
  PF_TRY(stage_in(pl, in, D, si));
  const double* b = si.dev; // "solution"
  for (int i = 0; i < 6; i++)
    PF_TRY(dev_alloc(&pl->cg[i], D));
  double* x = pl->cg[0];   // iter
  double* Mx = pl->cg[1];  // matrixTimesIter
  double* r = pl->cg[2];   // residual
  double* z = pl->cg[3];   // precResidual
  double* dir = pl->cg[4]; // direction
  double* Md = pl->cg[5];  // matrixTimesDirection
  unsigned count = 0;
  double maxabs = 0.;
  PF_TRY(reduce_dev(pl, D, b, nullptr, nullptr, 2, &maxabs));
  const unsigned g = grid_for(D);
  if (!(maxabs > 1e-10)) {
    // fieldZero: return the input unchanged (matrix.hh:202-207)
    PF_CUDA(cudaMemcpyAsync(x, b, D * sizeof(double), cudaMemcpyDeviceToDevice,
                            pl->stream));
  } else {
    PF_TRY(apply_dev(pl, PF_OP_INVERSE_EXT, b, x));
    PF_TRY(apply_dev(pl, PF_OP_MATRIX, x, Mx));
    k_residual<<<g, 256, 0, pl->stream>>>(D, b, Mx, r);
    pl->launches++;
    PF_TRY(apply_dev(pl, PF_OP_INVERSE_EXT, r, z));
    PF_CUDA(cudaMemcpyAsync(dir, z, D * sizeof(double), cudaMemcpyDeviceToDevice,
                            pl->stream));
    // scalars of the recurrence stay on the device (k_cg_*); the host only
    // learns "converged?" -- one iteration late, through a pinned copy -- to
    // decide whether to enqueue another iteration.  The decision for iteration
    // i + 1 uses the flag after iteration i - 1 (a blocking wait on that copy),
    // so it is deterministic and identical on all ranks of a slab plan; the
    // one iteration that may be enqueued after convergence is a no-op.
    PF_TRY(dev_alloc(&pl->cg_scalars, 16));
    if (!pl->cg_host) {
      PF_CUDA(cudaHostAlloc((void**)&pl->cg_host, 4 * 2 * sizeof(double),
                            cudaHostAllocDefault));
      for (int i = 0; i < 4; i++)
        PF_CUDA(cudaEventCreateWithFlags(&pl->cg_ev[i], cudaEventDisableTiming));
    }
    double* S = pl->cg_scalars;
    PF_TRY(reduce_to(pl, D, z, r, nullptr, 0, S + 0));
    PF_TRY(reduce_to(pl, D, r, r, nullptr, 0, S + 1));
    PF_TRY(reduce_to(pl, D, x, Mx, b, 3, S + 2));
    k_cg_init<<<1, 1, 0, pl->stream>>>(S);
    pl->launches++;
    auto publish = [&](int slot) -> int {
      PF_CUDA(cudaMemcpyAsync(pl->cg_host + 2 * slot, S + 7, 2 * sizeof(double),
                              cudaMemcpyDeviceToHost, pl->stream));
      PF_CUDA(cudaEventRecord(pl->cg_ev[slot], pl->stream));
      return 0;
    };
    PF_TRY(publish(3)); // state before the first iteration
    PF_CUDA(cudaEventSynchronize(pl->cg_ev[3]));
    bool converged = pl->cg_host[2 * 3] != 0.;
    const unsigned maxit = pl->d.cg_iterations;
    unsigned enq = 0; // iterations enqueued
    while (!converged && enq < maxit) {
      PF_TRY(apply_dev(pl, PF_OP_MATRIX, dir, Md));
      PF_TRY(reduce_to(pl, D, dir, Md, nullptr, 0, S + 3));
      { // update, residual and the new energy in one sweep
        const int nb = (int)std::max<size_t>(
          1, std::min<size_t>(RED_BLOCKS, (D + RED_THREADS - 1) / RED_THREADS));
        k_cg_update_fused<<<nb, RED_THREADS, 0, pl->stream>>>(D, S, dir, Md, x, Mx, b, r,
                                                             pl->red_partial);
        k_reduce2<<<1, RED_THREADS, 0, pl->stream>>>(nb, pl->red_partial, 3, S + 5);
        pl->launches += 2;
        if (pl->nranks > 1) {
          k_slab_reduce<<<1, 32, 0, pl->stream>>>(
            pl->d_flagsA, slab_flags(pl, pl->slab_block, 0), pl->nranks, pl->rank,
            ++pl->red_seq, 3, S + 5, pl->d_timeout, pl->slab_limit_ns);
          pl->launches++;
        }
      }
      PF_TRY(apply_dev(pl, PF_OP_INVERSE_EXT, r, z));
      PF_TRY(reduce_to(pl, D, z, r, nullptr, 0, S + 6));
      k_cg_direction<<<g, 256, 0, pl->stream>>>(D, S, z, dir);
      k_cg_scalars<<<1, 1, 0, pl->stream>>>(S);
      pl->launches += 2;
      PF_TRY(publish((int)(enq % 3)));
      enq++;
      if (enq >= 2) { // flag after iteration enq - 2
        const int slot = (int)((enq - 2) % 3);
        PF_CUDA(cudaEventSynchronize(pl->cg_ev[slot]));
        converged = pl->cg_host[2 * slot] != 0.;
      }
    }
    // final state: the number of iterations that really ran
    PF_TRY(publish(3));
    PF_CUDA(cudaEventSynchronize(pl->cg_ev[3]));
    count = (unsigned)pl->cg_host[2 * 3 + 1];
  }
  if (iterations)
    *iterations = count;
  PF_CUDA(cudaMemcpyAsync(out, x, D * sizeof(double),
                          is_device_ptr(out) ? cudaMemcpyDeviceToDevice
                                             : cudaMemcpyDeviceToHost,
                          pl->stream));
  PF_CUDA(cudaStreamSynchronize(pl->stream));
  return slab_check(pl);
}

int
pf_transform(pf_plan* pl, const double* in, double* out, size_t nn)
{
  PF_TRY(check_plan(pl));
  Staged si;
  PF_TRY(stage_in(pl, in, nn, si));
  const bool dev = is_device_ptr(out);
  double* tmp = nullptr;
  if (!dev)
    PF_TRY(dev_alloc(&tmp, nn));
  double* dst = dev ? out : tmp;
  k_transform<<<grid_for(nn), 256, 0, pl->stream>>>(nn, si.dev, dst,
                                                   pl->d.transform,
                                                   pl->d.boxcox_lambda);
  cudaError_t e = cudaGetLastError();
  pl->launches++;
  if (e == cudaSuccess && !dev)
    e = cudaMemcpyAsync(out, dst, nn * sizeof(double), cudaMemcpyDeviceToHost,
                        pl->stream);
  if (e == cudaSuccess)
    e = cudaStreamSynchronize(pl->stream);
  if (tmp)
    cudaFree(tmp);
  if (e != cudaSuccess)
    return fail(PF_ERR_CUDA, cudaGetErrorString(e));
  return 0;
}

int
pf_vec_axpy(pf_plan* pl, size_t nn, double alpha, const double* x, double* y)
{
  PF_TRY(check_plan(pl));
  const bool dev = is_device_ptr(y);
  Staged sx;
  StagedOut sy;
  PF_TRY(stage_in(pl, x, nn, sx));
  PF_TRY(stage_out(pl, y, nn, true, sy));
  k_axpy<<<grid_for(nn), 256, 0, pl->stream>>>(nn, alpha, sx.dev, sy.dev);
  PF_CUDA(cudaGetLastError());
  pl->launches++;
  if (dev && !sx.owned)
    return 0; // device arrays: asynchronous on the plan's stream
  return sy.finish();
}

int
pf_vec_scale(pf_plan* pl, size_t nn, double alpha, double* x)
{
  PF_TRY(check_plan(pl));
  StagedOut sx;
  PF_TRY(stage_out(pl, x, nn, true, sx));
  k_scale<<<grid_for(nn), 256, 0, pl->stream>>>(nn, alpha, sx.dev);
  PF_CUDA(cudaGetLastError());
  pl->launches++;
  if (!sx.owned)
    return 0;
  return sx.finish();
}

static int
vec_reduce(pf_plan* pl, size_t nn, const double* x, const double* y, int mode,
           double* result)
{
  PF_TRY(check_plan(pl));
  Staged sx, sy;
  PF_TRY(stage_in(pl, x, nn, sx));
  if (y)
    PF_TRY(stage_in(pl, y, nn, sy));
  return reduce_dev(pl, nn, sx.dev, y ? sy.dev : nullptr, nullptr, mode, result);
}

int
pf_vec_dot(pf_plan* pl, size_t nn, const double* x, const double* y,
           double* result)
{
  return vec_reduce(pl, nn, x, y, 0, result);
}
int
pf_vec_norm1(pf_plan* pl, size_t nn, const double* x, double* result)
{
  return vec_reduce(pl, nn, x, nullptr, 1, result);
}
int
pf_vec_norminf(pf_plan* pl, size_t nn, const double* x, double* result)
{
  return vec_reduce(pl, nn, x, nullptr, 2, result);
}
int
pf_vec_equal(pf_plan* pl, size_t nn, const double* x, const double* y,
             int* same)
{
  if (!same)
    return fail(PF_ERR_INVALID, "pf_vec_equal: null result");
  double differing = 0.;
  PF_TRY(vec_reduce(pl, nn, x, y, 4, &differing));
  *same = differing == 0. ? 1 : 0;
  return 0;
}
int
pf_vec_fill(pf_plan* pl, size_t nn, double value, double* x)
{
  PF_TRY(check_plan(pl));
  StagedOut sx;
  PF_TRY(stage_out(pl, x, nn, false, sx));
  k_fill<<<grid_for(nn), 256, 0, pl->stream>>>(nn, value, sx.dev);
  PF_CUDA(cudaGetLastError());
  pl->launches++;
  if (!sx.owned)
    return 0;
  return sx.finish();
}

} // extern "C"

// ---------------------------------------------------------------------------
// pf_vec_dot_static: the scalar product of pf_vec_dot with its scratch (the
// per-block partial sums of k_reduce1 / k_reduce2) in a file-scope device
// array instead of plan memory from cudaMalloc.  No counterpart in the
// reference; added for callers that
// cannot allocate (CUDA graph capture, serving loops).  Every GPU that loads
// the library holds the array (4800 bytes) until exit; every call shares it,
// so calls that overlap on one GPU race.  Nothing in it has to start at zero:
// each of the nb blocks overwrites its partial before k_reduce2 reads the
// first nb entries.
// ---------------------------------------------------------------------------
namespace {
constexpr size_t STATIC_DOT_MAX = size_t(1) << 28; // elements per call
static __device__ __align__(256) double g_static_red[RED_BLOCKS + 8];
constexpr int STATIC_MAX_DEVICES = 64;
double* g_static_red_ptr[STATIC_MAX_DEVICES] = {};
std::mutex g_static_red_guard;

// device address of g_static_red on `device` (which is current): the name
// itself is a host-side stand-in in host code
int
static_red_address(int device, double** ptr)
{
  if (device < 0 || device >= STATIC_MAX_DEVICES)
    return fail(PF_ERR_INVALID, "pf_vec_dot_static: device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_static_red_guard);
  if (!g_static_red_ptr[device]) {
    double* p = nullptr;
    PF_CUDA(cudaGetSymbolAddress((void**)&p, g_static_red));
    g_static_red_ptr[device] = p;
  }
  *ptr = g_static_red_ptr[device];
  return 0;
}
} // namespace

extern "C" {

int
pf_vec_dot_static(pf_plan* pl, size_t nn, const double* x, const double* y,
                  double* result)
{
  PF_TRY(check_plan(pl));
  if (!x || !y || !result)
    return fail(PF_ERR_INVALID, "pf_vec_dot_static: null argument");
  if (nn > STATIC_DOT_MAX)
    return fail(PF_ERR_INVALID,
                "pf_vec_dot_static: n = " + std::to_string(nn) +
                  " exceeds the maximum of " + std::to_string(STATIC_DOT_MAX) +
                  " elements (use pf_vec_dot)");
  if (pl->nranks > 1)
    return fail(PF_ERR_INVALID,
                "pf_vec_dot_static: not available on a connected multi-GPU plan");
  // host arrays would need a staging allocation
  if (!is_device_ptr(x) || !is_device_ptr(y))
    return fail(PF_ERR_INVALID,
                "pf_vec_dot_static: x and y must be device arrays");
  double* scratch = nullptr;
  PF_TRY(static_red_address(pl->device, &scratch));
  // same grid and summation order as pf_vec_dot (reduce_dev): bit-identical
  const int nb =
    std::max(1, (int)std::min<size_t>(RED_BLOCKS, (nn + RED_THREADS - 1) / RED_THREADS));
  const bool dev_result = is_device_ptr(result);
  double* dst = dev_result ? result : scratch + RED_BLOCKS;
  k_reduce1<<<nb, RED_THREADS, 0, pl->stream>>>(nn, x, y, nullptr, 0, scratch);
  k_reduce2<<<1, RED_THREADS, 0, pl->stream>>>(nb, scratch, 0, dst);
  PF_CUDA(cudaGetLastError());
  pl->launches += 2;
  if (dev_result)
    return 0; // asynchronous on the plan's stream
  PF_CUDA(cudaMemcpyAsync(result, dst, sizeof(double), cudaMemcpyDeviceToHost,
                          pl->stream));
  PF_CUDA(cudaStreamSynchronize(pl->stream));
  return 0;
}

// ---------------------------------------------------------------------------
// resolution transfer, localisation, point evaluation
// ---------------------------------------------------------------------------
namespace {
// one block row per field row: x covers the cells of a row, y / z the rows
// and planes (grid-strided beyond the launch limits)
dim3
row_grid(unsigned c0, unsigned c1, unsigned c2)
{
  return dim3(std::min<unsigned>((c0 + 255) / 256, 1024),
              std::min<unsigned>(c1, 65535), std::min<unsigned>(c2, 65535));
}

int
field_dims(const char* who, int dim, const unsigned* cells, int min_dim,
           size_t* count)
{
  if (dim < 1 || dim > 3)
    return fail(PF_ERR_INVALID, "dimension of field has to be 1, 2 or 3");
  if (dim < min_dim)
    return fail(PF_ERR_INVALID, "not implemented"); // stochastic.hh:664,749
  if (!cells)
    return fail(PF_ERR_INVALID, std::string(who) + ": null cells");
  size_t n = 1;
  for (int i = 0; i < dim; i++) {
    if (cells[i] == 0)
      return fail(PF_ERR_INVALID, std::string(who) + ": empty dimension");
    n *= cells[i];
  }
  *count = n;
  return 0;
}
} // namespace

int
pf_field_refine(pf_plan* pl, int dim, const unsigned* cells, const double* in,
                double* out)
{
  PF_TRY(check_plan(pl));
  size_t n_old = 0;
  PF_TRY(field_dims("pf_field_refine", dim, cells, 2, &n_old));
  const size_t n_new = n_old << dim;
  Staged si;
  StagedOut so;
  PF_TRY(stage_in(pl, in, n_old, si));
  PF_TRY(stage_out(pl, out, n_new, false, so));
  if (reinterpret_cast<uintptr_t>(so.dev) & 15)
    return fail(PF_ERR_INVALID, "pf_field_refine: out must be 16-byte aligned");
  const unsigned o2 = dim == 3 ? cells[2] : 1;
  const dim3 grid = row_grid(cells[0], cells[1], o2);
  if (dim == 3)
    k_refine<3><<<grid, 256, 0, pl->stream>>>(
      cells[0], cells[1], o2, si.dev, reinterpret_cast<double2*>(so.dev));
  else
    k_refine<2><<<grid, 256, 0, pl->stream>>>(
      cells[0], cells[1], o2, si.dev, reinterpret_cast<double2*>(so.dev));
  PF_CUDA(cudaGetLastError());
  pl->launches++;
  return so.finish();
}

int
pf_field_coarsen(pf_plan* pl, int dim, const unsigned* cells, const double* in,
                 double* out)
{
  PF_TRY(check_plan(pl));
  size_t n_old = 0;
  PF_TRY(field_dims("pf_field_coarsen", dim, cells, 2, &n_old));
  for (int i = 0; i < dim; i++)
    if (cells[i] % 2 != 0) // fieldtraits.hh:336-339
      return fail(PF_ERR_INVALID, "cannot coarsen odd number of cells");
  const size_t n_new = n_old >> dim;
  Staged si;
  StagedOut so;
  PF_TRY(stage_in(pl, in, n_old, si));
  PF_TRY(stage_out(pl, out, n_new, false, so));
  if (reinterpret_cast<uintptr_t>(si.dev) & 15)
    return fail(PF_ERR_INVALID, "pf_field_coarsen: in must be 16-byte aligned");
  const double2* src = reinterpret_cast<const double2*>(si.dev);
  if (dim == 3)
    k_coarsen<3><<<grid_for(n_new), 256, 0, pl->stream>>>(
      cells[0] / 2, cells[1] / 2, n_new, src, so.dev);
  else
    k_coarsen<2><<<grid_for(n_new), 256, 0, pl->stream>>>(
      cells[0] / 2, cells[1] / 2, n_new, src, so.dev);
  PF_CUDA(cudaGetLastError());
  pl->launches++;
  return so.finish();
}

int
pf_field_localize(pf_plan* pl, int dim, const unsigned* cells,
                  const unsigned* offset, const unsigned* global_cells,
                  const double* extensions, const double* center,
                  double radius, double* data)
{
  PF_TRY(check_plan(pl));
  size_t n = 0;
  PF_TRY(field_dims("pf_field_localize", dim, cells, 1, &n));
  if (!extensions || !center)
    return fail(PF_ERR_INVALID, "pf_field_localize: null geometry");
  FieldGeom g{};
  g.dim = dim;
  for (int i = 0; i < dim; i++) {
    g.cells[i] = cells[i];
    g.offset[i] = offset ? offset[i] : 0;
    g.gcells[i] = global_cells ? global_cells[i] : cells[i];
    g.ext[i] = extensions[i];
    g.center[i] = center[i];
  }
  StagedOut sd;
  PF_TRY(stage_out(pl, data, n, true, sd));
  const double factor = std::pow(2. * 3.14159, -(dim / 2.));
  const unsigned c1 = dim > 1 ? cells[1] : 1, c2 = dim > 2 ? cells[2] : 1;
  const size_t entries = (size_t)cells[0] + (dim > 1 ? c1 : 0) + (dim > 2 ? c2 : 0);
  if (entries > pl->axis_cap) {
    PF_CUDA(cudaStreamSynchronize(pl->stream));
    dev_free(&pl->axis_tab);
    PF_TRY(dev_alloc(&pl->axis_tab, entries));
    pl->axis_cap = entries;
  }
  k_localize_axes<<<grid_for(entries), 256, 0, pl->stream>>>(g, factor, radius * radius,
                                                              pl->axis_tab);
  const dim3 grid = row_grid(cells[0], (c1 + 3) / 4, c2);
  if (dim == 3)
    k_localize<3><<<grid, 256, 0, pl->stream>>>(cells[0], c1, c2, pl->axis_tab,
                                                sd.dev);
  else if (dim == 2)
    k_localize<2><<<grid, 256, 0, pl->stream>>>(cells[0], c1, c2, pl->axis_tab,
                                                sd.dev);
  else
    k_localize<1><<<grid, 256, 0, pl->stream>>>(cells[0], c1, c2, pl->axis_tab,
                                                sd.dev);
  PF_CUDA(cudaGetLastError());
  pl->launches += 2;
  return sd.finish();
}

namespace {
int
make_trend(const char* who, int dim, int ntrend,
           const pf_trend_component* trend, TrendSet& tr)
{
  tr.n = 0;
  if (ntrend < 0 || ntrend > PF_MAX_TREND_COMPONENTS)
    return fail(PF_ERR_INVALID, std::string(who) + ": at most " +
                                  std::to_string(PF_MAX_TREND_COMPONENTS) +
                                  " trend components");
  if (ntrend > 0 && !trend)
    return fail(PF_ERR_INVALID, std::string(who) + ": null trend");
  for (int t = 0; t < ntrend; t++) {
    if (trend[t].kind < PF_TREND_MEAN || trend[t].kind > PF_TREND_BLOCK)
      return fail(PF_ERR_INVALID, "Trend component type not found!"); // trend.hh:633
    tr.kind[t] = trend[t].kind;
    for (int i = 0; i < 7; i++)
      tr.coeff[t][i] = i < 2 * dim + 1 ? trend[t].coeff[i] : 0.;
  }
  tr.n = ntrend;
  return 0;
}
} // namespace

int
pf_field_evaluate_trend(pf_plan* pl, int dim, const unsigned* cells,
                        const double* extensions, const double* field,
                        size_t npoints, const double* locations, int ntrend,
                        const pf_trend_component* trend, int apply_transform,
                        double* out)
{
  PF_TRY(check_plan(pl));
  size_t n = 0;
  PF_TRY(field_dims("pf_field_evaluate", dim, cells, 1, &n));
  if (!extensions)
    return fail(PF_ERR_INVALID, "pf_field_evaluate: null extensions");
  TrendSet tr;
  PF_TRY(make_trend("pf_field_evaluate", dim, ntrend, trend, tr));
  if (npoints == 0)
    return 0;
  FieldGeom g{};
  g.dim = dim;
  for (int i = 0; i < dim; i++) {
    g.cells[i] = cells[i];
    g.gcells[i] = cells[i];
    g.ext[i] = extensions[i];
    g.center[i] = 1. / extensions[i]; // reciprocal for the index fast path
  }
  Staged sf, sl;
  StagedOut so;
  PF_TRY(stage_in(pl, field, n, sf));
  PF_TRY(stage_in(pl, locations, npoints * dim, sl));
  PF_TRY(stage_out(pl, out, npoints, false, so));
  PF_TRY(dev_alloc(&pl->eval_flag, 1));
  PF_CUDA(cudaMemsetAsync(pl->eval_flag, 0, sizeof(int), pl->stream));
  const int kind = apply_transform ? pl->d.transform : 0;
  const unsigned blocks = grid_for((npoints + 3) / 4);
#define PF_EVAL(D, T)                                                          \
  k_evaluate<D, T><<<blocks, 256, 0, pl->stream>>>(                            \
    g, tr, sf.dev, npoints, sl.dev, kind, pl->d.boxcox_lambda, so.dev,         \
    pl->eval_flag)
  if (dim == 3) {
    if (tr.n) PF_EVAL(3, true); else PF_EVAL(3, false);
  } else if (dim == 2) {
    if (tr.n) PF_EVAL(2, true); else PF_EVAL(2, false);
  } else {
    if (tr.n) PF_EVAL(1, true); else PF_EVAL(1, false);
  }
#undef PF_EVAL
  PF_CUDA(cudaGetLastError());
  pl->launches++;
  int flag = 0;
  PF_CUDA(cudaMemcpyAsync(&flag, pl->eval_flag, sizeof(int),
                          cudaMemcpyDeviceToHost, pl->stream));
  PF_TRY(so.finish());
  if (flag)
    return fail(PF_ERR_INVALID, "pf_field_evaluate: location outside the domain");
  return 0;
}

int
pf_field_evaluate(pf_plan* pl, int dim, const unsigned* cells,
                  const double* extensions, const double* field,
                  size_t npoints, const double* locations,
                  int apply_transform, double* out)
{
  return pf_field_evaluate_trend(pl, dim, cells, extensions, field, npoints,
                                 locations, 0, nullptr, apply_transform, out);
}

int
pf_field_bulk_evaluate(pf_plan* pl, int dim, const unsigned* cells,
                       const double* extensions, const double* field,
                       int ntrend, const pf_trend_component* trend,
                       int apply_transform, double* out)
{
  PF_TRY(check_plan(pl));
  size_t n = 0;
  PF_TRY(field_dims("pf_field_bulk_evaluate", dim, cells, 1, &n));
  if (!extensions)
    return fail(PF_ERR_INVALID, "pf_field_bulk_evaluate: null extensions");
  TrendSet tr;
  PF_TRY(make_trend("pf_field_bulk_evaluate", dim, ntrend, trend, tr));
  FieldGeom g{};
  g.dim = dim;
  for (int i = 0; i < 3; i++) {
    g.cells[i] = i < dim ? cells[i] : 1;
    g.gcells[i] = g.cells[i];
    g.ext[i] = i < dim ? extensions[i] : 1.;
    // meshsize of bulkEvaluate, randomfield.hh:622
    g.center[i] = g.ext[i] / static_cast<double>(g.cells[i]);
  }
  Staged sf;
  StagedOut so;
  PF_TRY(stage_in(pl, field, n, sf));
  PF_TRY(stage_out(pl, out, n, false, so));
  const int kind = apply_transform ? pl->d.transform : 0;
  const dim3 grid = row_grid(g.cells[0], (g.cells[1] + 3) / 4, g.cells[2]);
  if (tr.n) {
    const size_t entries =
      (size_t)tr.n * ((size_t)g.cells[0] + g.cells[1] + g.cells[2]);
    if (entries > pl->axis_cap) {
      PF_CUDA(cudaStreamSynchronize(pl->stream));
      dev_free(&pl->axis_tab);
      PF_TRY(dev_alloc(&pl->axis_tab, entries));
      pl->axis_cap = entries;
    }
    k_trend_axes<<<grid_for(entries), 256, 0, pl->stream>>>(g, tr, pl->axis_tab);
    pl->launches++;
  }
#define PF_BULK(D, T)                                                          \
  k_bulk_evaluate<D, T><<<grid, 256, 0, pl->stream>>>(                         \
    g, tr, pl->axis_tab, sf.dev, kind, pl->d.boxcox_lambda, so.dev)
  if (dim == 3) {
    if (tr.n) PF_BULK(3, true); else PF_BULK(3, false);
  } else if (dim == 2) {
    if (tr.n) PF_BULK(2, true); else PF_BULK(2, false);
  } else {
    if (tr.n) PF_BULK(1, true); else PF_BULK(1, false);
  }
#undef PF_BULK
  PF_CUDA(cudaGetLastError());
  pl->launches++;
  return so.finish();
}

size_t
pf_field_overlap_size(const pf_plan* pl, const int* ppd)
{
  if (!pl || !ppd)
    return 0;
  size_t total = 0;
  for (int i = 0; i < pl->dim; i++) {
    if (ppd[i] < 1 || pl->c[i] % ppd[i] != 0)
      return 0;
    total += 2 * (pl->Dl / (size_t)(pl->c[i] / ppd[i]));
  }
  return total;
}

int
pf_field_slab_to_blocks(pf_plan* pl, const int* ppd, const double* slab,
                        double* eval, double* overlap)
{
  PF_TRY(check_plan(pl));
  if (!ppd || !slab || !eval)
    return fail(PF_ERR_INVALID, "null argument");
  BlockGeom g;
  PF_TRY(block_geom(pl, ppd, g));
  const size_t nov = pf_field_overlap_size(pl, ppd);
  Staged si;
  StagedOut se, so;
  PF_TRY(stage_in(pl, slab, pl->Dl, si));
  PF_TRY(stage_out(pl, eval, pl->Dl, false, se));
  if (overlap)
    PF_TRY(stage_out(pl, overlap, nov, false, so));
  unsigned long long e = 0;
  PF_TRY(block_stage(pl, si.dev, g, &e));
  const size_t total = pl->Dl + (overlap ? nov : 0);
  k_slab_to_blocks<<<grid_for(total), 256, 0, pl->stream>>>(g, se.dev, so.dev, pl->Dl, total);
  k_slab_signal<<<1, 32, 0, pl->stream>>>(pl->d_flagsB, pl->nranks, pl->rank, e);
  PF_CUDA(cudaGetLastError());
  pl->launches += 2;
  if (overlap)
    PF_TRY(so.finish());
  return se.finish();
}

int
pf_field_blocks_to_slab(pf_plan* pl, const int* ppd, const double* eval, double* slab)
{
  PF_TRY(check_plan(pl));
  if (!ppd || !slab || !eval)
    return fail(PF_ERR_INVALID, "null argument");
  BlockGeom g;
  PF_TRY(block_geom(pl, ppd, g));
  Staged si;
  StagedOut so;
  PF_TRY(stage_in(pl, eval, pl->Dl, si));
  PF_TRY(stage_out(pl, slab, pl->Dl, false, so));
  unsigned long long e = 0;
  PF_TRY(block_stage(pl, si.dev, g, &e));
  k_blocks_to_slab<<<grid_for(pl->Dl), 256, 0, pl->stream>>>(g, so.dev, pl->Dl);
  k_slab_signal<<<1, 32, 0, pl->stream>>>(pl->d_flagsB, pl->nranks, pl->rank, e);
  PF_CUDA(cudaGetLastError());
  pl->launches += 2;
  return so.finish();
}

int
pf_device_alloc(pf_plan* pl, size_t bytes, void** ptr)
{
  PF_TRY(check_plan(pl));
  cudaError_t e = cudaMalloc(ptr, bytes);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(PF_ERR_NOMEM, cudaGetErrorString(e));
  }
  return 0;
}

int
pf_device_free(pf_plan* pl, void* ptr)
{
  // plan may be NULL (a buffer that outlives the plan it was made with):
  // cudaFree waits for the work that uses the buffer by itself
  if (pl) {
    PF_TRY(check_plan(pl));
    PF_CUDA(cudaStreamSynchronize(pl->stream));
  }
  PF_CUDA(cudaFree(ptr));
  return 0;
}

int
pf_memcpy(pf_plan* pl, void* dst, const void* src, size_t bytes)
{
  PF_TRY(check_plan(pl));
  PF_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, pl->stream));
  PF_CUDA(cudaStreamSynchronize(pl->stream));
  return 0;
}

} // extern "C"
