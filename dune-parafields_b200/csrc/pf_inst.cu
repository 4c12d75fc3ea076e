// pf_inst.cu -- instantiates every pass kernel for one line length PF_N
// (compiled once per N: -DPF_N=2 ... 1024) and exports one launcher.
#include "pf_kernels.cuh"
#if PF_N >= 256
#include "pf_pipe.cuh"
#endif
#if PF_N >= 512
#include "pf_stream.cuh"
#endif

#ifndef PF_N
#error "compile with -DPF_N=<line length>"
#endif

#define PF_CAT2(a, b) a##b
#define PF_CAT(a, b) PF_CAT2(a, b)

namespace pf {

template <class Kern>
static cudaError_t
launch(Kern kern, const PassArgs& a, const cplx* tw2, long long blocks,
       size_t smem, cudaStream_t st)
{
  if (blocks <= 0)
    return cudaSuccess;
  if (blocks > 0x7fffffffLL)
    return cudaErrorInvalidConfiguration;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(
      kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess)
      return e;
  }
  kern<<<(unsigned)blocks, PF_THREADS, smem, st>>>(a, tw2);
  return cudaGetLastError();
}

} // namespace pf

using namespace pf;

extern "C" int PF_CAT(pf_launch_n, PF_N)(int kind, const PassArgs* ap,
                                         const void* tw2v, cudaStream_t st)
{
  constexpr int N = PF_N;
  using K = KCfg<N>;
  const PassArgs& a = *ap;
  const cplx* tw2 = reinterpret_cast<const cplx*>(tw2v);
  const long long ntA = (a.A + K::LPC - 1) / K::LPC;
  const long long ntA2 = (a.A + 2 * K::LPC - 1) / (2 * K::LPC);
  const long long nlines = (long long)a.A * a.B;
  cudaError_t e = cudaErrorInvalidValue;
  switch (kind) {
    case K_GEN_TILE: {
      const size_t sm = K::smem_tile(a.keep);
      const long long nb = ntA * a.B;
#if PF_N >= 256
      // embedding factor 2, one rank: half of the line kept, known at compile time
      if (a.noise_kind == NOISE_PHILOX_FAST && a.keep == N / 2 && a.scale == 1.0 &&
          a.n_peers <= 1)
        e = launch(k_gen_tile<N, NOISE_PHILOX_FAST, true>, a, tw2, nb, sm, st);
      else
#endif
      if (a.noise_kind == NOISE_PHILOX_FAST)
        e = launch(k_gen_tile<N, NOISE_PHILOX_FAST>, a, tw2, nb, sm, st);
      else if (a.noise_kind == NOISE_PHILOX_F64)
        e = launch(k_gen_tile<N, NOISE_PHILOX_F64>, a, tw2, nb, sm, st);
      else
        e = launch(k_gen_tile<N, NOISE_INJECTED>, a, tw2, nb, sm, st);
      break;
    }
    case K_GEN_REIM: {
      const size_t sm = K::exch_bytes;
      const long long nb = ntA * a.B;
      if (a.noise_kind == NOISE_PHILOX_FAST)
        e = launch(k_gen_reim<N, NOISE_PHILOX_FAST>, a, tw2, nb, sm, st);
      else if (a.noise_kind == NOISE_PHILOX_F64)
        e = launch(k_gen_reim<N, NOISE_PHILOX_F64>, a, tw2, nb, sm, st);
      else
        e = launch(k_gen_reim<N, NOISE_INJECTED>, a, tw2, nb, sm, st);
      break;
    }
    case K_C2T_INV:
      e = launch(k_c2t<N, +1>, a, tw2, ntA * a.B, K::smem_tile(a.keep), st);
      break;
    case K_C2T_FWD:
      e = launch(k_c2t<N, -1>, a, tw2, ntA * a.B, K::smem_tile(a.keep), st);
      break;
    case K_C2R:
      e = launch(k_c2r<N>, a, tw2,
                 ((nlines + 1) / 2 + K::LPC - 1) / K::LPC, K::exch_bytes, st);
      break;
    case K_R2C_LOAD:
      e = launch(k_r2c<N, false>, a, tw2, ntA2 * a.B, K::smem_r2c(), st);
      break;
    case K_R2C_COV:
      e = launch(k_r2c<N, true>, a, tw2, ntA2 * a.B, K::smem_r2c(), st);
      break;
    case K_LAMBDA:
      e = launch(k_lambda<N, false>, a, tw2, (nlines + K::LPC - 1) / K::LPC,
                 K::exch_bytes, st);
      break;
    case K_LAMBDA_COV:
      e = launch(k_lambda<N, true>, a, tw2, (nlines + K::LPC - 1) / K::LPC,
                 K::exch_bytes, st);
      break;
    case K_SPEC_TILE:
#if PF_N >= 256
      if (a.n_valid == N / 2 && a.keep == N / 2 && a.n_peers <= 1)
        e = launch(k_spec<N, false, true>, a, tw2, ntA * a.B,
                   K::smem_tile(a.keep), st);
      else
#endif
      e = launch(k_spec<N, false>, a, tw2, ntA * a.B, K::smem_tile(a.keep),
                 st);
      break;
    case K_LONG_C2R1: {
      const long long nt = (a.long_n1 + K::LPC - 1) / K::LPC;
      e = launch(k_long_c2r1<N>, a, tw2, nt * a.B, K::exch_bytes, st);
      break;
    }
    case K_SPEC_REIM:
      e = launch(k_spec<N, true>, a, tw2, ntA * a.B, K::exch_bytes, st);
      break;
    default:
      break;
  }
  return (int)e;
}

#if PF_N >= 256
// persistent TMA-pipelined variants (pf_pipe.cuh); `tmap` is a host pointer to
// the CUtensorMap of the transposed output (unused for K_C2R)
namespace pf {
template <class Kern, class... Extra>
static cudaError_t
launch_pipe(Kern kern, int threads, size_t smem, int grid, cudaStream_t st,
            const PassArgs& a, const cplx* tw2, Extra... extra)
{
  cudaError_t e = cudaFuncSetAttribute(
    kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess)
    return e;
  kern<<<grid, threads, smem, st>>>(a, tw2, extra...);
  return cudaGetLastError();
}
} // namespace pf

extern "C" int PF_CAT(pf_pipe_launch_n, PF_N)(int kind, const PassArgs* ap,
                                              const void* tw2v,
                                              const void* tmap, int grid,
                                              cudaStream_t st)
{
  constexpr int N = PF_N;
  using P = PCfg<N>;
  const PassArgs& a = *ap;
  const cplx* tw2 = reinterpret_cast<const cplx*>(tw2v);
  cudaError_t e = cudaErrorInvalidValue;
  if (grid <= 0)
    return (int)cudaSuccess;
  switch (kind) {
    case K_GEN_TILE: {
      // `tmap` points to a TileMaps (one map per destination rank)
      const TileMaps tms = *reinterpret_cast<const TileMaps*>(tmap);
      if (a.noise_kind == NOISE_PHILOX_FAST)
        e = launch_pipe(kp_gen<N, NOISE_PHILOX_FAST>, P::THREADS, P::SMEM_TILE,
                        grid, st, a, tw2, tms);
      else if (a.noise_kind == NOISE_PHILOX_F64)
        e = launch_pipe(kp_gen<N, NOISE_PHILOX_F64>, P::THREADS, P::SMEM_TILE,
                        grid, st, a, tw2, tms);
      else
        e = launch_pipe(kp_gen<N, NOISE_INJECTED>, P::THREADS, P::SMEM_TILE,
                        grid, st, a, tw2, tms);
      break;
    }
#if PF_N >= 512
    case K_GEN3: {
      // grouped pass A: `tmap` points to a TileMaps whose m[0] describes the
      // output with LPC*16-byte boxes in the matching swizzle mode
      using G = G3Cfg<N>;
      const TileMaps tms = *reinterpret_cast<const TileMaps*>(tmap);
#if PF_N == 1024
      // embedding factor 2: half of the line is kept (compile-time pruning)
      if (a.keep == N / 2 && a.noise_kind == NOISE_PHILOX_FAST)
        e = launch_pipe(kp_gen3<N, NOISE_PHILOX_FAST, true>, G::THREADS, G::SMEM,
                        grid, st, a, tw2, tms);
      else if (a.keep == N / 2 && a.noise_kind == NOISE_PHILOX_F64)
        e = launch_pipe(kp_gen3<N, NOISE_PHILOX_F64, true>, G::THREADS, G::SMEM,
                        grid, st, a, tw2, tms);
      else
#endif
      if (a.noise_kind == NOISE_PHILOX_FAST)
        e = launch_pipe(kp_gen3<N, NOISE_PHILOX_FAST>, G::THREADS, G::SMEM, grid,
                        st, a, tw2, tms);
      else if (a.noise_kind == NOISE_PHILOX_F64)
        e = launch_pipe(kp_gen3<N, NOISE_PHILOX_F64>, G::THREADS, G::SMEM, grid,
                        st, a, tw2, tms);
      else
        e = launch_pipe(kp_gen3<N, NOISE_INJECTED>, G::THREADS, G::SMEM, grid,
                        st, a, tw2, tms);
      break;
    }
#endif
    case K_C2T_INV: {
      const C2TMaps tm = *reinterpret_cast<const C2TMaps*>(tmap);
      e = launch_pipe(kp_c2t<N, +1>, P::THREADS, P::SMEM_TILE, grid, st, a, tw2,
                      tm);
      break;
    }
    case K_C2T_FWD: {
      const C2TMaps tm = *reinterpret_cast<const C2TMaps*>(tmap);
      e = launch_pipe(kp_c2t<N, -1>, P::THREADS, P::SMEM_TILE, grid, st, a, tw2,
                      tm);
      break;
    }
    case K_C2R:
      e = launch_pipe(kp_c2r<N>, P::THREADS - 32, P::SMEM_LINES, grid, st, a,
                      tw2);
      break;
#if PF_N == 1024
    case K_LONG_C2R1:
      e = launch_pipe(kp_long_c2r1<N>, LC1Cfg<N>::THREADS, LC1Cfg<N>::SMEM, grid, st,
                      a, tw2);
      break;
#endif
    default:
      break;
  }
  return (int)e;
}
#endif

#if PF_N >= 512
namespace pf {
template <class Kern, class... Args>
static cudaError_t
launch_stream(Kern kern, int threads, size_t smem, int grid, cudaStream_t st,
              Args... args)
{
  cudaError_t e = cudaFuncSetAttribute(
    kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess)
    return e;
  kern<<<grid, threads, smem, st>>>(args...);
  return cudaGetLastError();
}
} // namespace pf

// streaming route (pf_stream.cuh): which = 0 fused first two axes (kp_ab),
// `maps` -> one CUtensorMap (ring tiles); which = 1 last axis (kp_c2rt),
// `maps` -> two CUtensorMaps (256-row and 8-row boxes of Zt)
extern "C" int PF_CAT(pf_stream_launch_n, PF_N)(int which, const void* args,
                                                const void* tw2v,
                                                const void* maps, int grid,
                                                cudaStream_t st)
{
  constexpr int N = PF_N;
  using S = SCfg<N>;
  const cplx* tw2 = reinterpret_cast<const cplx*>(tw2v);
  const CUtensorMap* tm = reinterpret_cast<const CUtensorMap*>(maps);
  cudaError_t e = cudaErrorInvalidValue;
  if (grid <= 0)
    return (int)cudaSuccess;
  if (which == 0) {
    const StreamArgs& sa = *reinterpret_cast<const StreamArgs*>(args);
    if (sa.a.noise_kind == NOISE_PHILOX_FAST && sa.c1 == N / 2 && sa.c2 == N / 2)
      e = launch_stream(kp_ab<N, NOISE_PHILOX_FAST, true>, S::THREADS, S::SMEM, grid,
                        st, sa, tw2, tm[0]);
    else if (sa.a.noise_kind == NOISE_PHILOX_FAST)
      e = launch_stream(kp_ab<N, NOISE_PHILOX_FAST>, S::THREADS, S::SMEM, grid, st,
                      sa, tw2, tm[0]);
    else if (sa.a.noise_kind == NOISE_PHILOX_F64)
      e = launch_stream(kp_ab<N, NOISE_PHILOX_F64>, S::THREADS, S::SMEM, grid, st,
                      sa, tw2, tm[0]);
    else
      e = launch_stream(kp_ab<N, NOISE_INJECTED>, S::THREADS, S::SMEM, grid, st, sa,
                      tw2, tm[0]);
  } else {
    const PassArgs& a = *reinterpret_cast<const PassArgs*>(args);
    e = launch_stream(kp_c2rt<N>, S::THREADS, S::SMEM, grid, st, a, tw2, tm[0],
                    tm[1]);
  }
  return (int)e;
}
#endif
