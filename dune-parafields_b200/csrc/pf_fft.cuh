// pf_fft.cuh -- fp64 complex FFT building blocks for sm_100a.
//
// Design (see DESIGN.md "FFT core"):
//  * a line of N complex doubles is held in registers, E = min(N,32) elements
//    per thread, TPL = N/E threads per line (so a 1024-point line is exactly
//    one warp, shorter lines share a warp), thread t owning elements
//    t + TPL*q, q < E -- on input AND on output;
//  * Stockham autosort passes: radix-32 in registers, one padded shared-memory
//    exchange, radix-N/32 in registers (N <= 1024);
//  * twiddles inside a register butterfly are compile-time constants, the
//    inter-pass twiddles come from a per-N table laid out [r][lane] so that a
//    warp reads 512 contiguous bytes.
//
// Replaces the FFTW3 plans behind backends/fftwwrapper.hh:230-415 of the
// reference (fftw_mpi_plan_dft / _r2c / _c2r); transform conventions are
// FFTW's: DIR=-1 forward, DIR=+1 backward, both unnormalised.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <type_traits>
#include <utility>

#define PF_HD __host__ __device__ __forceinline__

namespace pf {

typedef double2 cplx;

template <class F, int... I>
PF_HD void
static_for_impl(F&& f, std::integer_sequence<int, I...>)
{
  (f(std::integral_constant<int, I>{}), ...);
}
template <int N, class F>
PF_HD void
static_for(F&& f)
{
  static_for_impl(f, std::make_integer_sequence<int, N>{});
}

PF_HD cplx
cadd(cplx a, cplx b)
{
  return make_double2(a.x + b.x, a.y + b.y);
}
PF_HD cplx
csub(cplx a, cplx b)
{
  return make_double2(a.x - b.x, a.y - b.y);
}
PF_HD cplx
cmul(cplx a, cplx w)
{
  return make_double2(fma(a.x, w.x, -(a.y * w.y)), fma(a.x, w.y, a.y * w.x));
}
// a * conj(w)
PF_HD cplx
cmulc(cplx a, cplx w)
{
  return make_double2(fma(a.x, w.x, a.y * w.y), fma(a.y, w.x, -(a.x * w.y)));
}

// ---- compile-time 32nd roots of unity -------------------------------------
__host__ __device__ constexpr double
cos32_tab(int k)
{
  return k == 0   ? 1.0
         : k == 1 ? 0.98078528040323044912618223613424
         : k == 2 ? 0.92387953251128675612818318939679
         : k == 3 ? 0.83146961230254523707878837761791
         : k == 4 ? 0.70710678118654752440084436210485
         : k == 5 ? 0.55557023301960222474283081394853
         : k == 6 ? 0.38268343236508977172845998403040
         : k == 7 ? 0.19509032201612826784828486847702
                  : 0.0;
}
__host__ __device__ constexpr double
cos32(int k)
{
  k = ((k % 32) + 32) % 32;
  if (k > 16)
    k = 32 - k;
  return k <= 8 ? cos32_tab(k) : -cos32_tab(16 - k);
}
__host__ __device__ constexpr double
sin32(int k)
{
  return cos32(k - 8);
}

// a * W32^K with W32 = exp(DIR * 2 pi i / 32)
template <int K, int DIR>
PF_HD cplx
mul_w32(cplx a)
{
  constexpr int k = ((K % 32) + 32) % 32;
  if constexpr (k == 0)
    return a;
  else if constexpr (k == 16)
    return make_double2(-a.x, -a.y);
  else if constexpr (k == 8)
    return DIR > 0 ? make_double2(-a.y, a.x) : make_double2(a.y, -a.x);
  else if constexpr (k == 24)
    return DIR > 0 ? make_double2(a.y, -a.x) : make_double2(-a.y, a.x);
  else if constexpr (k % 8 == 4) {
    constexpr double c = cos32(k), s = DIR * sin32(k);
    constexpr double h = 0.70710678118654752440084436210485;
    const double re = (c > 0 ? a.x : -a.x) - (s > 0 ? a.y : -a.y);
    const double im = (s > 0 ? a.x : -a.x) + (c > 0 ? a.y : -a.y);
    return make_double2(h * re, h * im);
  } else {
    constexpr double c = cos32(k), s = DIR * sin32(k);
    return make_double2(fma(a.x, c, -(a.y * s)), fma(a.x, s, a.y * c));
  }
}

// ---- register butterflies: in-place DFT of v[0], v[S], ..., v[(R-1)S],
//      natural-order output ---------------------------------------------------
template <int R, int DIR, int S>
struct Dft;

template <int DIR, int S>
struct Dft<1, DIR, S>
{
  static PF_HD void run(cplx*) {}
};

template <int DIR, int S>
struct Dft<2, DIR, S>
{
  static PF_HD void run(cplx* v)
  {
    const cplx a = v[0], b = v[S];
    v[0] = cadd(a, b);
    v[S] = csub(a, b);
  }
};

template <int DIR, int S>
struct Dft<4, DIR, S>
{
  static PF_HD void run(cplx* v)
  {
    const cplx t0 = cadd(v[0], v[2 * S]), t1 = csub(v[0], v[2 * S]);
    const cplx t2 = cadd(v[S], v[3 * S]);
    const cplx t3 = mul_w32<8, DIR>(csub(v[S], v[3 * S]));
    v[0] = cadd(t0, t2);
    v[S] = cadd(t1, t3);
    v[2 * S] = csub(t0, t2);
    v[3 * S] = csub(t1, t3);
  }
};

template <int DIR, int S>
struct Dft<8, DIR, S>
{
  static PF_HD void run(cplx* v)
  {
    // even / odd halves
    cplx e0 = v[0], e1 = v[2 * S], e2 = v[4 * S], e3 = v[6 * S];
    cplx o0 = v[S], o1 = v[3 * S], o2 = v[5 * S], o3 = v[7 * S];
    {
      const cplx t0 = cadd(e0, e2), t1 = csub(e0, e2), t2 = cadd(e1, e3);
      const cplx t3 = mul_w32<8, DIR>(csub(e1, e3));
      e0 = cadd(t0, t2);
      e1 = cadd(t1, t3);
      e2 = csub(t0, t2);
      e3 = csub(t1, t3);
    }
    {
      const cplx t0 = cadd(o0, o2), t1 = csub(o0, o2), t2 = cadd(o1, o3);
      const cplx t3 = mul_w32<8, DIR>(csub(o1, o3));
      o0 = cadd(t0, t2);
      o1 = mul_w32<4, DIR>(cadd(t1, t3));
      o2 = mul_w32<8, DIR>(csub(t0, t2));
      o3 = mul_w32<12, DIR>(csub(t1, t3));
    }
    v[0] = cadd(e0, o0);
    v[4 * S] = csub(e0, o0);
    v[S] = cadd(e1, o1);
    v[5 * S] = csub(e1, o1);
    v[2 * S] = cadd(e2, o2);
    v[6 * S] = csub(e2, o2);
    v[3 * S] = cadd(e3, o3);
    v[7 * S] = csub(e3, o3);
  }
};

// Cooley-Tukey composition R = R1*R2 on registers (n = R2*n1 + n2)
template <int R1, int R2, int DIR, int S>
struct DftCT
{
  static constexpr int R = R1 * R2;
  static PF_HD void run(cplx* v)
  {
    static_for<R2>([&](auto n2c) {
      constexpr int n2 = decltype(n2c)::value;
      Dft<R1, DIR, S * R2>::run(v + S * n2);
    });
    static_for<R1>([&](auto k1c) {
      constexpr int k1 = decltype(k1c)::value;
      static_for<R2>([&](auto n2c) {
        constexpr int n2 = decltype(n2c)::value;
        if constexpr (k1 > 0 && n2 > 0)
          v[S * (R2 * k1 + n2)] =
            mul_w32<(32 / R) * n2 * k1, DIR>(v[S * (R2 * k1 + n2)]);
      });
    });
    static_for<R1>([&](auto k1c) {
      constexpr int k1 = decltype(k1c)::value;
      Dft<R2, DIR, S>::run(v + S * R2 * k1);
    });
    // X[k1 + R1*k2] sits at position R2*k1 + k2: undo the digit reversal
    cplx tmp[R];
    static_for<R>([&](auto ic) {
      constexpr int i = decltype(ic)::value;
      tmp[i] = v[S * i];
    });
    static_for<R1>([&](auto k1c) {
      constexpr int k1 = decltype(k1c)::value;
      static_for<R2>([&](auto k2c) {
        constexpr int k2 = decltype(k2c)::value;
        v[S * (k1 + R1 * k2)] = tmp[R2 * k1 + k2];
      });
    });
  }
};

template <int DIR, int S>
struct Dft<16, DIR, S>
{
  static PF_HD void run(cplx* v)
  {
    DftCT<4, 4, DIR, S>::run(v);
  }
};
template <int DIR, int S>
struct Dft<32, DIR, S>
{
  static PF_HD void run(cplx* v)
  {
    DftCT<4, 8, DIR, S>::run(v);
  }
};

// ---- line configuration -----------------------------------------------------
template <int N>
struct LineCfg
{
  static_assert(N >= 2 && N <= 1024 && (N & (N - 1)) == 0,
                "warp-level line FFT supports N = 2..1024, power of two");
  static constexpr int E = N >= 32 ? 32 : N; // elements per thread
  static constexpr int TPL = N / E;          // threads per line
  static constexpr int R1 = E;
  static constexpr int R2 = N / R1;          // 1 (single pass) .. 32
  static constexpr int NB2 = R2 > 1 ? 32 / R2 : 0; // pass-2 butterflies/thread
  static constexpr int PITCH = N + N / 32;   // padded smem line (complex)
  static constexpr int TW2 = R2 > 1 ? R2 * 32 : 0; // twiddle table entries
};

PF_HD int
pidx(int i)
{
  return i + (i >> 5);
}

// In-place FFT of one line distributed over TPL threads of a warp.
//   v   : E registers, thread t holds x[t + TPL*q]  (input AND output)
//   s   : this line's shared-memory exchange area (PITCH complex)
//   tw2 : table [R2][32], W_N^{r*k} forward (exp(-2 pi i r k / N))
// Split in two phases so that a host harness can emulate the warp
// (tests/host/fft_emul.cu); on the device line_fft() runs both with a
// __syncwarp() in between.  All lanes of the warp must call it together.
template <int N, int DIR>
PF_HD void
line_fft_phase1(cplx (&v)[LineCfg<N>::E], cplx* s, int t)
{
  using C = LineCfg<N>;
  Dft<C::R1, DIR, 1>::run(v);
  if constexpr (C::R2 > 1) {
    // exchange: pass-1 butterfly j = t writes out[j*32 + r]
#pragma unroll
    for (int r = 0; r < 32; r++)
      s[pidx(t * 32 + r)] = v[r];
  }
}

// TWS: the twiddle table lives in shared memory (plain loads) instead of
// global memory read through the non-coherent path
template <int N, int DIR, bool TWS = false>
PF_HD void
line_fft_phase2(cplx (&v)[LineCfg<N>::E], const cplx* s, int t,
                const cplx* __restrict__ tw2)
{
  using C = LineCfg<N>;
  if constexpr (C::R2 > 1) {
#pragma unroll
    for (int i = 0; i < C::NB2; i++) {
      const int jb = t + C::TPL * i; // butterfly index, < 32
      cplx u[C::R2];
#pragma unroll
      for (int r = 0; r < C::R2; r++)
        u[r] = s[pidx(jb + 32 * r)];
#pragma unroll
      for (int r = 1; r < C::R2; r++) {
#ifdef __CUDA_ARCH__
        const cplx w = TWS ? tw2[r * 32 + jb] : __ldg(&tw2[r * 32 + jb]);
#else
        const cplx w = tw2[r * 32 + jb];
#endif
        u[r] = DIR < 0 ? cmul(u[r], w) : cmulc(u[r], w);
      }
      Dft<C::R2, DIR, 1>::run(u);
#pragma unroll
      for (int r = 0; r < C::R2; r++)
        v[i + C::NB2 * r] = u[r];
    }
  }
}

// Phase 2 split in two so that a kernel can release the exchange area (and
// start refilling it asynchronously) before the second register pass:
//   load    : all butterflies' operands, v[i + NB2*r] = s[jb_i + 32 r]
//   compute : twiddle + radix-R2 butterflies in place (register stride NB2)
template <int N>
PF_HD void
line_fft_phase2_load(cplx (&v)[LineCfg<N>::E], const cplx* s, int t)
{
  using C = LineCfg<N>;
  static_assert(C::R2 > 1, "two-pass lines only");
#pragma unroll
  for (int i = 0; i < C::NB2; i++)
#pragma unroll
    for (int r = 0; r < C::R2; r++)
      v[i + C::NB2 * r] = s[pidx(t + C::TPL * i + 32 * r)];
}

template <int N, int DIR, bool TWS = false>
PF_HD void
line_fft_phase2_compute(cplx (&v)[LineCfg<N>::E], int t,
                        const cplx* __restrict__ tw2)
{
  using C = LineCfg<N>;
#pragma unroll
  for (int i = 0; i < C::NB2; i++) {
    const int jb = t + C::TPL * i;
#pragma unroll
    for (int r = 1; r < C::R2; r++) {
#ifdef __CUDA_ARCH__
      const cplx w = TWS ? tw2[r * 32 + jb] : __ldg(&tw2[r * 32 + jb]);
#else
      const cplx w = tw2[r * 32 + jb];
#endif
      v[i + C::NB2 * r] =
        DIR < 0 ? cmul(v[i + C::NB2 * r], w) : cmulc(v[i + C::NB2 * r], w);
    }
    Dft<C::R2, DIR, C::NB2>::run(v + i);
  }
}

#ifdef __CUDACC__
template <int N, int DIR, bool TWS = false>
__device__ __forceinline__ void
line_fft(cplx (&v)[LineCfg<N>::E], cplx* s, int t, const cplx* __restrict__ tw2)
{
  line_fft_phase1<N, DIR>(v, s, t);
  if constexpr (LineCfg<N>::R2 > 1) {
    __syncwarp();
    line_fft_phase2<N, DIR, TWS>(v, s, t, tw2);
    __syncwarp();
  }
}
#endif

} // namespace pf
