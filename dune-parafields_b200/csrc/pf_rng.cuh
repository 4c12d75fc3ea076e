// pf_rng.cuh -- counter-based white noise for the product path.
//
// Philox4x32 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3",
// SC'11; constants of Random123) keyed by the user seed, counter =
// (global frequency index, realization).  Two modes (pf_desc.rng_mode):
//   fp64 (1): 10 rounds (the Random123 / cuRAND default) + fp64 Box-Muller;
//   fast (0):  7 rounds (the smallest count Salmon et al. report as passing
//              BigCrush) + fp32 Box-Muller on the MUFU units (lg2 / sqrt / sin
//              / cos .approx, ~1e-6 absolute error per normal, 24-bit angle).
// Both are pinned by known-answer tests and by a CPU restatement of
// counter -> normals -> field (tests/test_philox.py, tests/test_gpu_noise.py).  The counter is a pure function of
// the frequency coordinates, so a field does not depend on the launch
// geometry or on how the spectrum is sharded over GPUs.
//
// Replaces RNGBackend::sample (backends/cpprngbackend.hh:49-58,
// backends/gslrngbackend.hh:107-130), which is a sequential stream and cannot
// be parallelised; bit-level parity with the reference stream is provided by
// the injected-noise entry point instead.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "pf_args.h"

namespace pf {

#define PF_PHILOX_M0 0xD2511F53u
#define PF_PHILOX_M1 0xCD9E8D57u
#define PF_PHILOX_W0 0x9E3779B9u
#define PF_PHILOX_W1 0xBB67AE85u

// round keys k_r = key + r * Weyl, r < 10 (computed once per launch on the host
// so that the device loop has no key arithmetic)
__host__ __device__ __forceinline__ PhiloxKeys
philox_keys(unsigned long long seed)
{
  PhiloxKeys k;
  uint32_t a = (uint32_t)seed, b = (uint32_t)(seed >> 32);
  for (int r = 0; r < 10; r++) {
    k.k0[r] = a;
    k.k1[r] = b;
    a += PF_PHILOX_W0;
    b += PF_PHILOX_W1;
  }
  return k;
}

// Philox4x32-R with precomputed round keys.  R = 10 is the Random123 / cuRAND
// default; R = 7 is the smallest round count Salmon et al. report as passing
// BigCrush (used by the "fast" noise mode).
template <int ROUNDS>
__host__ __device__ __forceinline__ uint4
philox4x32_r(unsigned long long ctr, unsigned long long realization,
             const PhiloxKeys& k)
{
  uint32_t c0 = (uint32_t)ctr, c1 = (uint32_t)(ctr >> 32);
  uint32_t c2 = (uint32_t)realization, c3 = (uint32_t)(realization >> 32);
#pragma unroll
  for (int r = 0; r < ROUNDS; r++) {
    // one 32x32->64 multiply per lane pair (IMAD.WIDE.U32 on the device)
    const uint64_t p0 = (uint64_t)PF_PHILOX_M0 * c0;
    const uint64_t p1 = (uint64_t)PF_PHILOX_M1 * c2;
    const uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
    const uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
    const uint32_t n0 = hi1 ^ c1 ^ k.k0[r];
    const uint32_t n2 = hi0 ^ c3 ^ k.k1[r];
    c0 = n0;
    c1 = lo1;
    c2 = n2;
    c3 = lo0;
  }
  return make_uint4(c0, c1, c2, c3);
}

__host__ __device__ __forceinline__ uint4
philox4x32_10(unsigned long long ctr, unsigned long long realization,
              unsigned long long seed)
{
  return philox4x32_r<10>(ctr, realization, philox_keys(seed));
}

#ifdef __CUDACC__
// Box-Muller on two 32-bit words.  u1 = (w0 + 0.5) 2^-32 in (0,1),
// theta = 2 pi (w1 2^-32 - 1/2) in [-pi, pi).
// fast: fp32 special-function units (lg2, sqrt, sin, cos approx), promoted to
// fp64 afterwards; absolute error ~1e-6 on an N(0,1) sample.
__device__ __forceinline__ double2
box_muller_fast(uint32_t w0, uint32_t w1)
{
  // (w0 + 0.5) 2^-32: exact for small w0, so the tail keeps its resolution
  const float u1 = fmaf(__uint2float_rn(w0), 2.3283064365386963e-10f,
                        1.1641532182693481e-10f);
  float l, r, s, c;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(u1));
  // -2 ln u1 = -2 ln2 * log2(u1)
  const float r2 = -1.3862943611198906f * l;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(r2));
  const float th =
    fmaf(__uint2float_rn(w1), 1.4629180792671596e-09f, -3.14159265358979f);
  asm("sin.approx.ftz.f32 %0, %1;" : "=f"(s) : "f"(th));
  asm("cos.approx.ftz.f32 %0, %1;" : "=f"(c) : "f"(th));
  return make_double2((double)(r * c), (double)(r * s));
}

__device__ __forceinline__ double2
box_muller_f64(uint32_t w0, uint32_t w1)
{
  const double u1 = ((double)w0 + 0.5) * 2.3283064365386963e-10;
  const double r = sqrt(-2.0 * log(u1));
  // theta / pi = 2 (w1 2^-32) - 1
  const double tp = (double)w1 * 4.6566128730773926e-10 - 1.0;
  double s, c;
  sincospi(tp, &s, &c);
  return make_double2(r * c, r * s);
}
#endif

} // namespace pf
