// oracle/shim/fftw_shim.cpp -- single-rank implementation of the MPI and
// FFTW3(-MPI) entry points declared in this directory.  TEST INFRASTRUCTURE:
// linked only into oracle/_ref/libpf_ref.so, which runs the UNMODIFIED
// reference headers from /root/reference to pin the oracle and to generate
// tests/golden/*.  Nothing in the product path uses it.
//
// Transform definitions (FFTW manual, section "What FFTW Really Computes"):
//   c2c   Y_k = sum_j X_j exp(sign 2 pi i j k / n), unnormalised, row major
//   r2c   forward c2c of real data, last dimension cut to n/2+1, in-place
//         layout padded to 2(n/2+1) reals per row
//   c2r   inverse; complex transforms along the leading dimensions, then a
//         half-complex -> real transform along the last one (the imaginary
//         parts of the DC and Nyquist entries of each line are not read)
//   REDFT00  Y_k = X_0 + (-1)^k X_{n-1} + 2 sum_{j=1}^{n-2} X_j cos(pi j k/(n-1))
//   FFTW_MPI_TRANSPOSED_OUT / _IN: the first two dimensions of the output /
//         input array are stored swapped (n1 x n0 x ...)
#include "fftw3-mpi.h"

#include <cmath>
#include <complex>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <stdexcept>
#include <vector>

typedef std::complex<double> cd;

// ---------------------------------------------------------------------------
// MPI, one rank
// ---------------------------------------------------------------------------
namespace {
struct Msg
{
  int tag;
  std::vector<char> data;
};
std::deque<Msg> g_queue;
}

extern "C" {
int MPI_Init(int*, char***) { return 0; }
int MPI_Finalize(void) { return 0; }
int MPI_Comm_rank(MPI_Comm, int* r) { *r = 0; return 0; }
int MPI_Comm_size(MPI_Comm, int* s) { *s = 1; return 0; }
int MPI_Barrier(MPI_Comm) { return 0; }
int MPI_Allreduce(const void* in, void* out, int count, MPI_Datatype t, MPI_Op,
                  MPI_Comm)
{
  if (in != out)
    memmove(out, in, (size_t)count * (size_t)t);
  return 0;
}
int MPI_Isend(const void* buf, int count, MPI_Datatype t, int, int tag,
              MPI_Comm, MPI_Request* req)
{
  Msg m;
  m.tag = tag;
  m.data.assign((const char*)buf, (const char*)buf + (size_t)count * t);
  g_queue.push_back(std::move(m));
  if (req)
    *req = 0;
  return 0;
}
int MPI_Recv(void* buf, int count, MPI_Datatype t, int, int tag, MPI_Comm,
             MPI_Status*)
{
  for (auto it = g_queue.begin(); it != g_queue.end(); ++it)
    if (it->tag == tag) {
      memcpy(buf, it->data.data(),
             std::min(it->data.size(), (size_t)count * t));
      g_queue.erase(it);
      return 0;
    }
  throw std::runtime_error("shim MPI_Recv: no matching message");
}
int MPI_Wait(MPI_Request*, MPI_Status*) { return 0; }
int MPI_Waitall(int, MPI_Request*, MPI_Status*) { return 0; }
}

// ---------------------------------------------------------------------------
// 1D transforms
// ---------------------------------------------------------------------------
namespace {

const long double PI_L = 3.141592653589793238462643383279502884L;

std::vector<cd>
roots(size_t n, int sign)
{
  std::vector<cd> w(n);
  for (size_t k = 0; k < n; k++) {
    const long double a = sign * 2.0L * PI_L * (long double)k / (long double)n;
    w[k] = cd((double)cosl(a), (double)sinl(a));
  }
  return w;
}

void
fft1d(cd* x, size_t n, int sign)
{
  if (n <= 1)
    return;
  const std::vector<cd> w = roots(n, sign);
  if ((n & (n - 1)) == 0) {
    // iterative radix 2, decimation in time
    for (size_t i = 1, j = 0; i < n; i++) {
      size_t bit = n >> 1;
      for (; j & bit; bit >>= 1)
        j ^= bit;
      j ^= bit;
      if (i < j)
        std::swap(x[i], x[j]);
    }
    for (size_t len = 2; len <= n; len <<= 1) {
      const size_t step = n / len;
      for (size_t i = 0; i < n; i += len)
        for (size_t k = 0; k < len / 2; k++) {
          const cd u = x[i + k], v = x[i + k + len / 2] * w[k * step];
          x[i + k] = u + v;
          x[i + k + len / 2] = u - v;
        }
    }
    return;
  }
  std::vector<cd> y(n);
  for (size_t k = 0; k < n; k++) {
    cd s = 0.;
    for (size_t j = 0; j < n; j++)
      s += x[j] * w[(j * k) % n];
    y[k] = s;
  }
  std::copy(y.begin(), y.end(), x);
}

void
redft00_1d(cd* x, size_t m)
{
  // acts on the real parts; m = number of points
  if (m <= 1)
    return;
  const size_t L = 2 * (m - 1);
  std::vector<cd> e(L);
  for (size_t j = 0; j < m; j++)
    e[j] = cd(x[j].real(), 0.);
  for (size_t j = 1; j + 1 < m; j++)
    e[L - j] = cd(x[j].real(), 0.);
  fft1d(e.data(), L, -1);
  for (size_t k = 0; k < m; k++)
    x[k] = cd(e[k].real(), 0.);
}

// half-complex (n/2+1 entries) -> n real values; Im of DC / Nyquist ignored
void
hc2r_1d(const cd* X, double* out, size_t n)
{
  std::vector<cd> f(n);
  const size_t h = n / 2;
  f[0] = cd(X[0].real(), 0.);
  for (size_t k = 1; k <= h; k++) {
    if (2 * k == n)
      f[k] = cd(X[k].real(), 0.);
    else {
      f[k] = X[k];
      f[n - k] = std::conj(X[k]);
    }
  }
  fft1d(f.data(), n, +1);
  for (size_t j = 0; j < n; j++)
    out[j] = f[j].real();
}

struct Dims
{
  std::vector<size_t> n; // logical, row major
  size_t total() const
  {
    size_t t = 1;
    for (size_t v : n)
      t *= v;
    return t;
  }
};

enum Kind1D
{
  K_FFT_FWD,
  K_FFT_BWD,
  K_REDFT00
};

void
along_axis(std::vector<cd>& a, const Dims& d, size_t axis, Kind1D kind)
{
  const size_t len = d.n[axis];
  size_t inner = 1;
  for (size_t i = axis + 1; i < d.n.size(); i++)
    inner *= d.n[i];
  const size_t outer = d.total() / (len * inner);
  std::vector<cd> line(len);
  for (size_t o = 0; o < outer; o++)
    for (size_t in = 0; in < inner; in++) {
      cd* base = a.data() + o * len * inner + in;
      for (size_t j = 0; j < len; j++)
        line[j] = base[j * inner];
      if (kind == K_REDFT00)
        redft00_1d(line.data(), len);
      else
        fft1d(line.data(), len, kind == K_FFT_FWD ? -1 : +1);
      for (size_t j = 0; j < len; j++)
        base[j * inner] = line[j];
    }
}

// offset of logical index (i0, i1, rest) in an array stored with the first
// two dimensions swapped
inline size_t
swapped_offset(size_t lin, const Dims& d)
{
  if (d.n.size() < 2)
    return lin;
  size_t rest = 1;
  for (size_t i = 2; i < d.n.size(); i++)
    rest *= d.n[i];
  const size_t r = lin % rest;
  const size_t i1 = (lin / rest) % d.n[1];
  const size_t i0 = lin / (rest * d.n[1]);
  return (i1 * d.n[0] + i0) * rest + r;
}

} // namespace

struct pf_shim_plan
{
  int kind; // 0 c2c, 1 r2c, 2 c2r, 3 r2r REDFT00
  int sign;
  unsigned flags;
  std::vector<size_t> n;
  void *in, *out;
};

extern "C" {

double*
fftw_alloc_real(size_t n)
{
  return (double*)malloc(sizeof(double) * (n ? n : 1));
}
fftw_complex*
fftw_alloc_complex(size_t n)
{
  return (fftw_complex*)malloc(sizeof(fftw_complex) * (n ? n : 1));
}
void
fftw_free(void* p)
{
  free(p);
}
int
fftw_import_wisdom_from_filename(const char*)
{
  return 0;
}
int
fftw_export_wisdom_to_filename(const char*)
{
  return 0;
}
void
fftw_mpi_init(void)
{}
void
fftw_mpi_broadcast_wisdom(MPI_Comm)
{}
void
fftw_mpi_gather_wisdom(MPI_Comm)
{}

ptrdiff_t
fftw_mpi_local_size(int rnk, const ptrdiff_t* n, MPI_Comm, ptrdiff_t* local_n0,
                    ptrdiff_t* local_0_start)
{
  ptrdiff_t t = 1;
  for (int i = 0; i < rnk; i++)
    t *= n[i];
  *local_n0 = n[0];
  *local_0_start = 0;
  return t;
}
ptrdiff_t
fftw_mpi_local_size_1d(ptrdiff_t n0, MPI_Comm, int, unsigned,
                       ptrdiff_t* local_ni, ptrdiff_t* local_i_start,
                       ptrdiff_t* local_no, ptrdiff_t* local_o_start)
{
  *local_ni = *local_no = n0;
  *local_i_start = *local_o_start = 0;
  return n0;
}
ptrdiff_t
fftw_mpi_local_size_transposed(int rnk, const ptrdiff_t* n, MPI_Comm comm,
                               ptrdiff_t* local_n0, ptrdiff_t* local_0_start,
                               ptrdiff_t* local_n1, ptrdiff_t* local_1_start)
{
  const ptrdiff_t t = fftw_mpi_local_size(rnk, n, comm, local_n0, local_0_start);
  *local_n1 = rnk > 1 ? n[1] : n[0];
  *local_1_start = 0;
  return t;
}
ptrdiff_t
fftw_mpi_local_size_many_transposed(int rnk, const ptrdiff_t* n,
                                    ptrdiff_t howmany, ptrdiff_t, ptrdiff_t,
                                    MPI_Comm comm, ptrdiff_t* local_n0,
                                    ptrdiff_t* local_0_start,
                                    ptrdiff_t* local_n1,
                                    ptrdiff_t* local_1_start)
{
  return howmany * fftw_mpi_local_size_transposed(rnk, n, comm, local_n0,
                                                  local_0_start, local_n1,
                                                  local_1_start);
}

static fftw_plan
make_plan(int kind, int sign, unsigned flags, int rnk, const ptrdiff_t* n,
          void* in, void* out)
{
  pf_shim_plan* p = new pf_shim_plan;
  p->kind = kind;
  p->sign = sign;
  p->flags = flags;
  p->n.assign(n, n + rnk);
  p->in = in;
  p->out = out;
  return p;
}

fftw_plan
fftw_mpi_plan_dft(int rnk, const ptrdiff_t* n, fftw_complex* in,
                  fftw_complex* out, MPI_Comm, int sign, unsigned flags)
{
  return make_plan(0, sign, flags, rnk, n, in, out);
}
fftw_plan
fftw_mpi_plan_dft_r2c(int rnk, const ptrdiff_t* n, double* in,
                      fftw_complex* out, MPI_Comm, unsigned flags)
{
  return make_plan(1, -1, flags, rnk, n, in, out);
}
fftw_plan
fftw_mpi_plan_dft_c2r(int rnk, const ptrdiff_t* n, fftw_complex* in,
                      double* out, MPI_Comm, unsigned flags)
{
  return make_plan(2, +1, flags, rnk, n, in, out);
}
fftw_plan
fftw_mpi_plan_r2r(int rnk, const ptrdiff_t* n, double* in, double* out,
                  MPI_Comm, const fftw_r2r_kind* kind, unsigned flags)
{
  for (int i = 0; i < rnk; i++)
    if (kind[i] != FFTW_REDFT00)
      return nullptr; // only the DCT-I of the matrix backend is provided
  return make_plan(3, -1, flags, rnk, n, in, out);
}
fftw_plan
fftw_mpi_plan_many_r2r(int, const ptrdiff_t*, ptrdiff_t, ptrdiff_t, ptrdiff_t,
                       double*, double*, MPI_Comm, const fftw_r2r_kind*,
                       unsigned)
{
  return nullptr; // DCTDSTFieldBackend is out of scope (never a default)
}

void
fftw_destroy_plan(fftw_plan p)
{
  delete p;
}

void
fftw_execute(const fftw_plan p)
{
  const size_t rnk = p->n.size();
  const bool tin = (p->flags & FFTW_MPI_TRANSPOSED_IN) && rnk > 1;
  const bool tout = (p->flags & FFTW_MPI_TRANSPOSED_OUT) && rnk > 1;
  if (p->kind == 0) {
    Dims d{ p->n };
    const size_t T = d.total();
    std::vector<cd> a(T);
    const cd* in = reinterpret_cast<const cd*>(p->in);
    for (size_t i = 0; i < T; i++)
      a[i] = in[tin ? swapped_offset(i, d) : i];
    for (size_t ax = 0; ax < rnk; ax++)
      along_axis(a, d, ax, p->sign < 0 ? K_FFT_FWD : K_FFT_BWD);
    cd* out = reinterpret_cast<cd*>(p->out);
    for (size_t i = 0; i < T; i++)
      out[tout ? swapped_offset(i, d) : i] = a[i];
  } else if (p->kind == 1) {
    Dims d{ p->n };
    const size_t nl = d.n[rnk - 1], h = nl / 2 + 1, pitch = 2 * h;
    const size_t rows = d.total() / nl;
    std::vector<cd> a(d.total());
    const double* in = reinterpret_cast<const double*>(p->in);
    for (size_t r = 0; r < rows; r++)
      for (size_t j = 0; j < nl; j++)
        a[r * nl + j] = cd(in[r * pitch + j], 0.);
    for (size_t ax = 0; ax < rnk; ax++)
      along_axis(a, d, ax, K_FFT_FWD);
    Dims dh{ p->n };
    dh.n[rnk - 1] = h;
    cd* out = reinterpret_cast<cd*>(p->out);
    for (size_t r = 0; r < rows; r++)
      for (size_t k = 0; k < h; k++) {
        const size_t lin = r * h + k;
        out[tout ? swapped_offset(lin, dh) : lin] = a[r * nl + k];
      }
  } else if (p->kind == 2) {
    Dims dh{ p->n };
    const size_t nl = p->n[rnk - 1], h = nl / 2 + 1, pitch = 2 * h;
    dh.n[rnk - 1] = h;
    const size_t T = dh.total(), rows = T / h;
    std::vector<cd> a(T);
    const cd* in = reinterpret_cast<const cd*>(p->in);
    for (size_t i = 0; i < T; i++)
      a[i] = in[tin ? swapped_offset(i, dh) : i];
    for (size_t ax = 0; ax + 1 < rnk; ax++)
      along_axis(a, dh, ax, K_FFT_BWD);
    double* out = reinterpret_cast<double*>(p->out);
    std::vector<double> line(nl);
    for (size_t r = 0; r < rows; r++) {
      hc2r_1d(a.data() + r * h, line.data(), nl);
      for (size_t j = 0; j < nl; j++)
        out[r * pitch + j] = line[j];
    }
  } else {
    Dims d{ p->n };
    const size_t T = d.total();
    std::vector<cd> a(T);
    const double* in = reinterpret_cast<const double*>(p->in);
    for (size_t i = 0; i < T; i++)
      a[i] = cd(in[tin ? swapped_offset(i, d) : i], 0.);
    for (size_t ax = 0; ax < rnk; ax++)
      along_axis(a, d, ax, K_REDFT00);
    double* out = reinterpret_cast<double*>(p->out);
    for (size_t i = 0; i < T; i++)
      out[tout ? swapped_offset(i, d) : i] = a[i].real();
  }
}

} // extern "C"
