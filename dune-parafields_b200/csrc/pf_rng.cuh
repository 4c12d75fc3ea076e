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

// fp64 Box-Muller by the library routines (log, sqrt, sincospi): the definition
// the oracle restates with libm; kept for the tests of the routines below
__device__ __forceinline__ double2
box_muller_f64_lib(uint32_t w0, uint32_t w1)
{
  const double u1 = ((double)w0 + 0.5) * 2.3283064365386963e-10;
  const double r = sqrt(-2.0 * log(u1));
  // theta / pi = 2 (w1 2^-32) - 1
  const double tp = (double)w1 * 4.6566128730773926e-10 - 1.0;
  double s, c;
  sincospi(tp, &s, &c);
  return make_double2(r * c, r * s);
}

// The same two numbers from routines written for THESE arguments (the library
// ones cost ~220 instructions per complex normal, three times the FFT the
// normals feed): the uniform is the dyadic rational (2 w0 + 1) 2^-33 and the
// angle the dyadic fraction w1 2^-32 of a turn, so both range reductions are
// exact integer work.  Checked against long double libm over 2 * 10^7 draws and
// the edge words: -2 ln u to 3.0e-16 relative, sin / cos to 1.8e-16 absolute
// (tests/host/box_muller_f64.cu), i.e. the results agree with the oracle's
// libm restatement to rounding.
//
// {1 / c_i rounded to double, -ln of that double [- ln 2 for i >= 53]},
// c_i = 1 + (i + 1/2) / 128 (c_0 = 1, c_127 = 2): ln m = -ln(1/c_i) + ln(1 + z),
// z = m / c_i - 1, |z| <= 2^-7; from i = 53 (m ~ sqrt 2) on the entry counts
// as one more power of two, so that ln u has no cancellation next to u = 1
static __device__ const double PF_LOG_TAB[128][2] = {
  { 0x1.0000000000000p+0, 0x0.0p+0 },
  { 0x1.fa11caa01fa12p-1, 0x1.7dc475f810a69p-7 },
  { 0x1.f6310aca0dbb5p-1, 0x1.3cea44346a584p-6 },
  { 0x1.f25f644230ab5p-1, 0x1.b9fc027af919ap-6 },
  { 0x1.ee9c7f8458e02p-1, 0x1.1b0d98923d97fp-5 },
  { 0x1.eae807aba01ebp-1, 0x1.58a5bafc8e4d3p-5 },
  { 0x1.e741aa59750e4p-1, 0x1.95c830ec8e3f2p-5 },
  { 0x1.e3a9179dc1a73p-1, 0x1.d276b8adb0b56p-5 },
  { 0x1.e01e01e01e01ep-1, 0x1.075983598e471p-4 },
  { 0x1.dca01dca01dcap-1, 0x1.253f62f0a1417p-4 },
  { 0x1.d92f2231e7f8ap-1, 0x1.42edcbea646eep-4 },
  { 0x1.d5cac807572b2p-1, 0x1.60658a93750c4p-4 },
  { 0x1.d272ca3fc5b1ap-1, 0x1.7da766d7b12d0p-4 },
  { 0x1.cf26e5c44bfc6p-1, 0x1.9ab42462033aep-4 },
  { 0x1.cbe6d9601cbe7p-1, 0x1.b78c82bb0eda0p-4 },
  { 0x1.c8b265afb8a42p-1, 0x1.d4313d66cb35dp-4 },
  { 0x1.c5894d10d4986p-1, 0x1.f0a30c01162a4p-4 },
  { 0x1.c26b5392ea01cp-1, 0x1.0671512ca596fp-3 },
  { 0x1.bf583ee868d8bp-1, 0x1.14785846742acp-3 },
  { 0x1.bc4fd65883e7bp-1, 0x1.2266f190a5acdp-3 },
  { 0x1.b951e2b18ff23p-1, 0x1.303d718e47fd5p-3 },
  { 0x1.b65e2e3beee05p-1, 0x1.3dfc2b0ecc62ap-3 },
  { 0x1.b37484ad806cep-1, 0x1.4ba36f39a55e5p-3 },
  { 0x1.b094b31d922a4p-1, 0x1.59338d9982085p-3 },
  { 0x1.adbe87f94905ep-1, 0x1.66acd4272ad51p-3 },
  { 0x1.aaf1d2f87ebfdp-1, 0x1.740f8f54037a3p-3 },
  { 0x1.a82e65130e159p-1, 0x1.815c0a14357e9p-3 },
  { 0x1.a574107688a4ap-1, 0x1.8e928de886d41p-3 },
  { 0x1.a2c2a87c51ca0p-1, 0x1.9bb362e7dfb85p-3 },
  { 0x1.a01a01a01a01ap-1, 0x1.a8becfc882f19p-3 },
  { 0x1.9d79f176b682dp-1, 0x1.b5b519e8fb5a6p-3 },
  { 0x1.9ae24ea5510dap-1, 0x1.c2968558c18c2p-3 },
  { 0x1.9852f0d8ec0ffp-1, 0x1.cf6354e09c5ddp-3 },
  { 0x1.95cbb0be377aep-1, 0x1.dc1bca0abec7bp-3 },
  { 0x1.934c67f9b2ce6p-1, 0x1.e8c0252aa5a60p-3 },
  { 0x1.90d4f120190d5p-1, 0x1.f550a564b7b37p-3 },
  { 0x1.8e6527af1373fp-1, 0x1.00e6c45ad501dp-2 },
  { 0x1.8bfce8062ff3ap-1, 0x1.071b85fcd590dp-2 },
  { 0x1.899c0f601899cp-1, 0x1.0d46b579ab74bp-2 },
  { 0x1.87427bcc092b9p-1, 0x1.136870293a8b0p-2 },
  { 0x1.84f00c2780614p-1, 0x1.1980d2dd4236fp-2 },
  { 0x1.82a4a0182a4a0p-1, 0x1.1f8ff9e48a2f3p-2 },
  { 0x1.8060180601806p-1, 0x1.2596010df763ap-2 },
  { 0x1.7e225515a4f1dp-1, 0x1.2b9303ab89d25p-2 },
  { 0x1.7beb3922e017cp-1, 0x1.31871c9544185p-2 },
  { 0x1.79baa6bb6398bp-1, 0x1.3772662bfd85cp-2 },
  { 0x1.77908119ac60dp-1, 0x1.3d54fa5c1f710p-2 },
  { 0x1.756cac201756dp-1, 0x1.432ef2a04e813p-2 },
  { 0x1.734f0c541fe8dp-1, 0x1.49006804009d0p-2 },
  { 0x1.713786d9c7c09p-1, 0x1.4ec9732600269p-2 },
  { 0x1.6f26016f26017p-1, 0x1.548a2c3add263p-2 },
  { 0x1.6d1a62681c861p-1, 0x1.5a42ab0f4cfe2p-2 },
  { 0x1.6b1490aa31a3dp-1, 0x1.5ff3070a793d4p-2 },
  { 0x1.691473a88d0c0p-1, -0x1.602d08af091ecp-2 },
  { 0x1.6719f3601671ap-1, -0x1.5a8cadbbedfa1p-2 },
  { 0x1.6524f853b4aa3p-1, -0x1.54f431b7be1a8p-2 },
  { 0x1.63356b88ac0dep-1, -0x1.4f637ebba9810p-2 },
  { 0x1.614b36831ae94p-1, -0x1.49da7f3bcc420p-2 },
  { 0x1.5f66434292dfcp-1, -0x1.44591e0539f49p-2 },
  { 0x1.5d867c3ece2a5p-1, -0x1.3edf463c1683ep-2 },
  { 0x1.5babcc647fa91p-1, -0x1.396ce359bbf53p-2 },
  { 0x1.59d61f123ccaap-1, -0x1.3401e12aecba0p-2 },
  { 0x1.5805601580560p-1, -0x1.2e9e2bce12286p-2 },
  { 0x1.56397ba7c52e2p-1, -0x1.2941afb186b7cp-2 },
  { 0x1.54725e6bb82fep-1, -0x1.23ec5991eba49p-2 },
  { 0x1.52aff56a8054bp-1, -0x1.1e9e1678899f5p-2 },
  { 0x1.50f22e111c4c5p-1, -0x1.1956d3b9bc2f9p-2 },
  { 0x1.4f38f62dd4c9bp-1, -0x1.14167ef367784p-2 },
  { 0x1.4d843bedc2c4cp-1, -0x1.0edd060b78082p-2 },
  { 0x1.4bd3edda68fe1p-1, -0x1.09aa572e6c6d4p-2 },
  { 0x1.4a27fad76014ap-1, -0x1.047e60cde83b7p-2 },
  { 0x1.4880522014880p-1, -0x1.feb2233ea07cbp-3 },
  { 0x1.46dce34596066p-1, -0x1.f474b134df228p-3 },
  { 0x1.453d9e2c776cap-1, -0x1.ea4449f04aaf5p-3 },
  { 0x1.43a2730abee4dp-1, -0x1.e020cc6235ab5p-3 },
  { 0x1.420b5265e5951p-1, -0x1.d60a17f903514p-3 },
  { 0x1.40782d10e6566p-1, -0x1.cc000c9db3c52p-3 },
  { 0x1.3ee8f42a5af07p-1, -0x1.c2028ab17f9b5p-3 },
  { 0x1.3d5d991aa75c6p-1, -0x1.b811730b823d4p-3 },
  { 0x1.3bd60d9232955p-1, -0x1.ae2ca6f672bd8p-3 },
  { 0x1.3a524387ac822p-1, -0x1.a454082e6ab03p-3 },
  { 0x1.38d22d366088ep-1, -0x1.9a8778debaa3ap-3 },
  { 0x1.3755bd1c945eep-1, -0x1.90c6db9fcbcdbp-3 },
  { 0x1.35dce5f9f2af8p-1, -0x1.871213750e994p-3 },
  { 0x1.34679ace01346p-1, -0x1.7d6903caf5acdp-3 },
  { 0x1.32f5ced6a1dfap-1, -0x1.73cb9074fd14dp-3 },
  { 0x1.3187758e9ebb6p-1, -0x1.6a399dabbd383p-3 },
  { 0x1.301c82ac40260p-1, -0x1.60b3100b09474p-3 },
  { 0x1.2eb4ea1fed14bp-1, -0x1.5737cc9018cddp-3 },
  { 0x1.2d50a012d50a0p-1, -0x1.4dc7b897bc1c7p-3 },
  { 0x1.2bef98e5a3711p-1, -0x1.4462b9dc9b3dcp-3 },
  { 0x1.2a91c92f3c105p-1, -0x1.3b08b6757f2a7p-3 },
  { 0x1.293725bb804a5p-1, -0x1.31b994d3a4f86p-3 },
  { 0x1.27dfa38a1ce4dp-1, -0x1.28753bc11aba2p-3 },
  { 0x1.268b37cd60127p-1, -0x1.1f3b925f25d44p-3 },
  { 0x1.2539d7e9177b2p-1, -0x1.160c8024b27b0p-3 },
  { 0x1.23eb79717605bp-1, -0x1.0ce7ecdccc28bp-3 },
  { 0x1.22a0122a0122ap-1, -0x1.03cdc0a51ec0dp-3 },
  { 0x1.21579804855e6p-1, -0x1.f57bc7d9005dbp-4 },
  { 0x1.2012012012012p-1, -0x1.e3707ee30487bp-4 },
  { 0x1.1ecf43c7fb84cp-1, -0x1.d179788219362p-4 },
  { 0x1.1d8f5672e4abdp-1, -0x1.bf968769fca18p-4 },
  { 0x1.1c522fc1ce059p-1, -0x1.adc77ee5aea8ep-4 },
  { 0x1.1b17c67f2bae3p-1, -0x1.9c0c32d4d254dp-4 },
  { 0x1.19e0119e0119ep-1, -0x1.8a6477a91dc29p-4 },
  { 0x1.18ab083902bdbp-1, -0x1.78d02263d82d7p-4 },
  { 0x1.1778a191bd684p-1, -0x1.674f089365a78p-4 },
  { 0x1.1648d50fc3201p-1, -0x1.55e10050e0382p-4 },
  { 0x1.151b9a3fdd5c9p-1, -0x1.4485e03dbdfb0p-4 },
  { 0x1.13f0e8d344724p-1, -0x1.333d7f8183f4ap-4 },
  { 0x1.12c8b89edc0acp-1, -0x1.2207b5c7854a1p-4 },
  { 0x1.11a3019a74826p-1, -0x1.10e45b3cae829p-4 },
  { 0x1.107fbbe011080p-1, -0x1.ffa6911ab9309p-5 },
  { 0x1.0f5edfab325a2p-1, -0x1.dda8adc67ee59p-5 },
  { 0x1.0e40655826011p-1, -0x1.bbcebfc68f424p-5 },
  { 0x1.0d24456359e3ap-1, -0x1.9a187b573de81p-5 },
  { 0x1.0c0a7868b4171p-1, -0x1.788595a3577c8p-5 },
  { 0x1.0af2f722eecb5p-1, -0x1.5715c4c03cee1p-5 },
  { 0x1.09ddba6af8360p-1, -0x1.35c8bfaa13069p-5 },
  { 0x1.08cabb37565e2p-1, -0x1.149e3e4005a8dp-5 },
  { 0x1.07b9f29b8eae2p-1, -0x1.e72bf2813ce6ap-6 },
  { 0x1.06ab59c7912fbp-1, -0x1.a55f548c5c427p-6 },
  { 0x1.059eea0727586p-1, -0x1.63d6178690bbep-6 },
  { 0x1.04949cc1664c5p-1, -0x1.228fb1fea2e0ap-6 },
  { 0x1.038c6b78247fcp-1, -0x1.c317384c75f0dp-7 },
  { 0x1.02864fc7729e9p-1, -0x1.41929f968330cp-7 },
  { 0x1.0182436517a37p-1, -0x1.8121214586b02p-8 },
  { 0x1.0000000000000p-1, 0x0.0p+0 },
};

// -2 ln u, u = (w0 + 1/2) 2^-32
__device__ __forceinline__ double
neg2_log_uniform(uint32_t w0)
{
  const double x = fma((double)w0, 2.0, 1.0); // 2 w0 + 1 < 2^33, exact
  const int hi = __double2hiint(x);
  const int i = (hi >> 13) & 127;
  const int k = (hi >> 20) - 1023 - 33 + (i >= 53 ? 1 : 0);
  const double m = __hiloint2double((hi & 0x000fffff) | 0x3ff00000, __double2loint(x));
  const double z = fma(m, __ldg(&PF_LOG_TAB[i][0]), -1.0);
  double p = -1.0 / 8;
  p = fma(p, z, 1.0 / 7);
  p = fma(p, z, -1.0 / 6);
  p = fma(p, z, 1.0 / 5);
  p = fma(p, z, -1.0 / 4);
  p = fma(p, z, 1.0 / 3);
  p = fma(p, z, -0.5);
  p = fma(p * z, z, z); // ln(1 + z)
  const double kd = (double)k;
  const double lo = fma(kd, 1.90821492927058770002e-10, p) + __ldg(&PF_LOG_TAB[i][1]);
  return -2.0 * fma(kd, 6.93147180369123816490e-01, lo);
}

// sin and cos of pi (w1 2^-31 - 1) = 2 pi phi - pi, phi = w1 2^-32
__device__ __forceinline__ void
sincos_turn_fraction(uint32_t w1, double* sn, double* cs)
{
  const uint32_t j = (w1 + 0x20000000u) >> 30;       // nearest quarter turn
  const int r = (int)(w1 - (j << 30));               // [-2^29, 2^29)
  const double a = (double)r * 0x1.921fb54442d18p-30; // 2 pi / 2^32: |a| <= pi / 4
  const double a2 = a * a;
  double s = -1.0 / 1307674368000.0;
  s = fma(s, a2, 1.0 / 6227020800.0);
  s = fma(s, a2, -1.0 / 39916800.0);
  s = fma(s, a2, 1.0 / 362880.0);
  s = fma(s, a2, -1.0 / 5040.0);
  s = fma(s, a2, 1.0 / 120.0);
  s = fma(s, a2, -1.0 / 6.0);
  s = fma(s * a2, a, a);
  double c = 1.0 / 20922789888000.0;
  c = fma(c, a2, -1.0 / 87178291200.0);
  c = fma(c, a2, 1.0 / 479001600.0);
  c = fma(c, a2, -1.0 / 3628800.0);
  c = fma(c, a2, 1.0 / 40320.0);
  c = fma(c, a2, -1.0 / 720.0);
  c = fma(c, a2, 1.0 / 24.0);
  c = fma(c, a2, -0.5);
  c = fma(c, a2, 1.0);
  // rotate by j quarter turns, then by the half turn of "- pi"
  const double ss = (j & 1) ? c : s, cc = (j & 1) ? s : c;
  const bool neg_s = ((j + 2) & 2) != 0;       // sin(2 pi phi - pi) = -sin(2 pi phi)
  const bool neg_c = ((j + 1 + 2) & 2) != 0;   // cos: sign pattern +,-,-,+ shifted
  *sn = neg_s ? -ss : ss;
  *cs = neg_c ? -cc : cc;
}

// sqrt of 2.3e-10 <= x <= 46 (no special cases): rsqrt seed + two Newton steps
__device__ __forceinline__ double
sqrt_positive(double x)
{
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double h = 0.5 * y;
  double r = x * y;
  r = fma(fma(-r, r, x), h, r);
  r = fma(fma(-r, r, x), h, r);
  return r;
}

__device__ __forceinline__ double2
box_muller_f64(uint32_t w0, uint32_t w1)
{
  const double r = sqrt_positive(neg2_log_uniform(w0));
  double s, c;
  sincos_turn_fraction(w1, &s, &c);
  return make_double2(r * c, r * s);
}
#endif

} // namespace pf
