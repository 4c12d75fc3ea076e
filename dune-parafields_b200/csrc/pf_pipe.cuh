// pf_pipe.cuh -- persistent, warp-specialised versions of the three passes of
// the generate pipeline (A: noise -> inverse FFT -> tile, B: contiguous lines
// -> FFT -> tile, C: half-spectrum line pairs -> real lines) for long lines
// (N >= 256).  Same arithmetic as pf_kernels.cuh (k_gen_tile / k_c2t / k_c2r);
// what changes is how data moves:
//
//  * one CTA per SM, 8 FFT warps + 1 store warp, looping over tiles;
//  * input lines arrive by 1D bulk copies (cp.async.bulk, completion on an
//    mbarrier) straight into the line's exchange buffer; the copy of the
//    NEXT tile's line is issued as soon as the current line has left shared
//    memory for the second register pass, so HBM latency hides behind the
//    fp64 butterflies;
//  * the transposed output tile ([element][line], 8 lines = 128 contiguous
//    bytes per element row) is staged in shared memory with the 128-byte
//    swizzle and written by one TMA tensor store per 256 rows
//    (cp.async.bulk.tensor), issued by the store warp, so FFT warps never
//    wait for global stores.
//
// Shared memory (N=1024): 8 x 16.5 KB exchange/landing buffers + 64 KB stage.
#pragma once
#include <cuda.h>

#include "pf_kernels.cuh"

namespace pf {

template <int N>
struct PCfg
{
  using C = LineCfg<N>;
  static constexpr int E = C::E;
  static constexpr int TPL = C::TPL;
  static constexpr int CW = 8;              // FFT (consumer) warps per CTA
  static constexpr int LPW = 32 / TPL;      // lines per warp
  static constexpr int LT = CW * LPW;       // lines per tile
  static constexpr int SUBS = LT / 8;       // 8-line (128 B) sub-stages
  static constexpr int THREADS = 32 * (CW + 1);
  static constexpr int SLOT_BYTES = C::PITCH * 16;
  static constexpr int SLOTS_BYTES = LT * SLOT_BYTES;
  static constexpr int NSTAGE = 2;           // double-buffered output stage
  static constexpr int STAGE_BYTES = 32768;  // per stage
  static constexpr int SR = STAGE_BYTES / (LT * 16); // stage rows per chunk
  static constexpr int BOXR = SR < 256 ? SR : 256;   // TMA box rows
  static constexpr int SUB_BYTES = SR * 128;
  static constexpr int BAR_BYTES = 128;
  // inter-pass twiddles copied to shared memory behind the barriers (with
  // ~200 KB of shared memory in use the remaining L1 keeps missing on the
  // table).  Measured on B200 (profiles/r2h): kp_c2t N = 512 0.194 -> 0.159 ms,
  // N = 1024 1.073 -> 1.088 ms (kept on the global table); kp_c2r N = 1024
  // 0.532 -> 0.499 ms, N = 512 unchanged
  static constexpr bool TWS_C2T = N <= 512;
  static constexpr bool TWS_C2R = true;
  static constexpr int TWB = C::TW2 * 16;
  static constexpr int SMEM_TILE =
    SLOTS_BYTES + NSTAGE * STAGE_BYTES + BAR_BYTES + (TWS_C2T ? TWB : 0) + 1024;
  static constexpr int SMEM_LINES =
    SLOTS_BYTES + BAR_BYTES + (TWS_C2R ? TWB : 0) + 1024;
  static_assert(SLOTS_BYTES % 1024 == 0, "stage must stay 1024-byte aligned");
  static_assert(N + 2 <= C::PITCH, "c2r landing zone needs 2*(N/2+1) entries");
};

// ---- PTX wrappers -----------------------------------------------------------
__device__ __forceinline__ uint32_t
smem_u32(const void* p)
{
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void
mbar_init(uint32_t bar, uint32_t count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)
               : "memory");
}
__device__ __forceinline__ void
mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void
mbar_arrive(uint32_t bar)
{
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void
mbar_wait(uint32_t bar, uint32_t parity)
{
  asm volatile("{\n"
               ".reg .pred P1;\n"
               "LAB_WAIT:\n"
               "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
               "@P1 bra DONE;\n"
               "bra LAB_WAIT;\n"
               "DONE:\n"
               "}" ::"r"(bar),
               "r"(parity)
               : "memory");
}
__device__ __forceinline__ void
fence_proxy_async()
{
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// global -> shared bulk copy, bytes % 16 == 0, completion counted on `bar`
__device__ __forceinline__ void
bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes "
               "[%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
// same with an L2 evict-last hint: rows that a mirrored line reads again soon
__device__ __forceinline__ void
bulk_load_keep(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar)
{
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes"
               ".L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar), "l"(pol)
               : "memory");
}
// shared -> global tensor store of one box at (c0, c1, c2)
__device__ __forceinline__ void
tma_store_3d(const CUtensorMap* tm, uint32_t src, int c0, int c1, int c2)
{
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group "
               "[%0, {%2, %3, %4}], [%1];" ::"l"(tm),
               "r"(src), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
__device__ __forceinline__ void
tma_store_4d(const CUtensorMap* tm, uint32_t src, int c0, int c1, int c2, int c3)
{
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group "
               "[%0, {%2, %3, %4, %5}], [%1];" ::"l"(tm),
               "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
// global -> shared tensor load of one box, completion counted on `bar`
__device__ __forceinline__ void
tma_load_4d(const CUtensorMap* tm, uint32_t dst, uint32_t bar, int c0, int c1,
            int c2, int c3)
{
  asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::"
               "complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
               "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
__device__ __forceinline__ void
bulk_commit()
{
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void
bulk_wait_read0()
{
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void
bulk_wait0()
{
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

__device__ __forceinline__ unsigned char*
align1024(unsigned char* p)
{
  const uint32_t a = smem_u32(p);
  return p + ((1024u - (a & 1023u)) & 1023u);
}

// shared-memory carve-up of the tile kernels
template <int N>
struct PSmem
{
  using P = PCfg<N>;
  unsigned char* base;
  __device__ __forceinline__ explicit PSmem(unsigned char* raw)
    : base(align1024(raw))
  {}
  __device__ __forceinline__ cplx* slot(int s) const
  {
    return reinterpret_cast<cplx*>(base + (size_t)s * P::SLOT_BYTES);
  }
  __device__ __forceinline__ unsigned char* stage(int i) const
  {
    return base + P::SLOTS_BYTES + (size_t)i * P::STAGE_BYTES;
  }
  // barriers: [0..CW) input per warp, [CW + i] stage i full,
  // [CW + 2 + i] stage i free
  __device__ __forceinline__ uint32_t bar(int i, bool tile) const
  {
    return smem_u32(base + P::SLOTS_BYTES +
                    (tile ? P::NSTAGE * P::STAGE_BYTES : 0)) +
           8u * i;
  }
};

__device__ __forceinline__ void
bulk_wait_read1()
{
  asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
}

// the store warp's loop: wait for a full stage, write it with TMA, release
// the PREVIOUS stage once its bulk group has been read out of shared memory
template <int N>
__device__ __forceinline__ void
store_warp_loop(const PassArgs& p, const CUtensorMap* tm, const PSmem<N>& sm,
                int ntA, long long ntiles)
{
  using P = PCfg<N>;
  if ((threadIdx.x & 31) != 0)
    return;
  const int nchunk = (p.keep + P::SR - 1) / P::SR;
  uint32_t it = 0;
  for (long long T = blockIdx.x; T < ntiles; T += gridDim.x) {
    const int at = (int)(T % ntA), b = (int)(T / ntA);
    for (int c = 0; c < nchunk; c++, it++) {
      const int s = it & 1;
      mbar_wait(sm.bar(P::CW + s, true), (it >> 1) & 1);
      const uint32_t st = smem_u32(sm.stage(s));
      const int rows = min(P::SR, p.keep - c * P::SR);
#pragma unroll
      for (int sub = 0; sub < P::SUBS; sub++) {
        if (at * P::LT + sub * 8 < p.A) {
          for (int r0 = 0; r0 < rows; r0 += P::BOXR)
            tma_store_3d(tm, st + sub * P::SUB_BYTES + r0 * 128,
                         2 * (at * P::LT + sub * 8), c * P::SR + r0, b);
        }
      }
      bulk_commit();
      if (it > 0) {
        bulk_wait_read1();
        mbar_arrive(sm.bar(P::CW + 2 + (s ^ 1), true));
      }
    }
  }
  bulk_wait0();
}

// FFT warps: write registers (thread t of line `slot` owns e = t + TPL*q) to
// the swizzled stage, one chunk of SR rows at a time
template <int N>
__device__ __forceinline__ void
stage_out(const cplx (&v)[LineCfg<N>::E], const PSmem<N>& sm, int slot, int t,
          int keep, uint32_t& it)
{
  using P = PCfg<N>;
  // chunk c holds element rows [c*SR, (c+1)*SR): registers QPC*c .. QPC*c+QPC-1
  // of every thread, row r = t + TPL*j inside the chunk, so the swizzle term
  // (r & 7 == t & 7) and all offsets but the thread's own are compile-time
  constexpr int QPC = P::SR / P::TPL;
  constexpr int NCH = P::E / QPC;
  static_assert(P::SR % P::TPL == 0 && P::TPL % 8 == 0 && P::E % QPC == 0,
                "stage_out chunk mapping");
  const int l8 = slot & 7;
  const int nchunk = (keep + P::SR - 1) / P::SR;
  const uint32_t toff =
    (uint32_t)(slot >> 3) * P::SUB_BYTES + t * 128 + ((l8 ^ (t & 7)) << 4);
#pragma unroll
  for (int c = 0; c < NCH; c++) {
    if (c < nchunk) {
      const int s = it & 1;
      mbar_wait(sm.bar(P::CW + 2 + s, true), ((it >> 1) & 1) ^ 1);
      unsigned char* st = sm.stage(s) + toff;
#pragma unroll
      for (int j = 0; j < QPC; j++) {
        if (t + P::TPL * (QPC * c + j) < keep)
          *reinterpret_cast<cplx*>(st + j * P::TPL * 128) = v[QPC * c + j];
      }
      fence_proxy_async();
      __syncwarp();
      if ((threadIdx.x & 31) == 0)
        mbar_arrive(sm.bar(P::CW + s, true));
      it++;
    }
  }
}

template <int N>
__device__ __forceinline__ void
pipe_init(const PSmem<N>& sm, bool tile)
{
  using P = PCfg<N>;
  if (threadIdx.x == 0) {
    for (int w = 0; w < P::CW; w++)
      mbar_init(sm.bar(w, tile), 1);
    if (tile) {
      for (int i = 0; i < P::NSTAGE; i++) {
        mbar_init(sm.bar(P::CW + i, true), P::CW);
        mbar_init(sm.bar(P::CW + 2 + i, true), 1);
      }
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_proxy_async();
  }
  __syncthreads();
}

// ===========================================================================
// pass A: noise -> inverse FFT -> transposed tile (k_gen_tile)
// ===========================================================================
// one tensor map per destination rank (slab mode); m[0] alone otherwise
struct TileMaps
{
  CUtensorMap m[8];
};

// store warp of pass A: like store_warp_loop, but every box of `box_rows`
// element rows goes to the rank that owns those rows (tensor stores to peer
// memory over NVLink in slab mode)
template <int N>
__device__ __forceinline__ void
store_warp_loop_gen(const PassArgs& p, const TileMaps* tms, const PSmem<N>& sm,
                    int ntA, long long ntiles)
{
  using P = PCfg<N>;
  if ((threadIdx.x & 31) != 0)
    return;
  const int nchunk = (p.keep + P::SR - 1) / P::SR;
  const int boxr = p.box_rows;
  const int rpp = p.n_peers > 1 ? p.rows_per_peer : 0x7fffffff;
  uint32_t it = 0;
  for (long long T = blockIdx.x; T < ntiles; T += gridDim.x) {
    const int at = (int)(T % ntA), b = (int)(T / ntA);
    for (int c = 0; c < nchunk; c++, it++) {
      const int s = it & 1;
      mbar_wait(sm.bar(P::CW + s, true), (it >> 1) & 1);
      const uint32_t st = smem_u32(sm.stage(s));
      const int rows = min(P::SR, p.keep - c * P::SR);
#pragma unroll
      for (int sub = 0; sub < P::SUBS; sub++) {
        if (at * P::LT + sub * 8 < p.A) {
          for (int r0 = 0; r0 < rows; r0 += boxr) {
            const int g = c * P::SR + r0;
            const int owner = g / rpp;
            if (p.blk_T > 0)
              tma_store_4d(&tms->m[owner], st + sub * P::SUB_BYTES + r0 * 128, 0,
                           g - owner * rpp,
                           (p.a_begin + at * P::LT + sub * 8) >> 3, b);
            else
              tma_store_3d(&tms->m[owner], st + sub * P::SUB_BYTES + r0 * 128,
                           2 * (p.a_begin + at * P::LT + sub * 8),
                           g - owner * rpp, b);
          }
        }
      }
      bulk_commit();
      if (it > 0) {
        bulk_wait_read1();
        mbar_arrive(sm.bar(P::CW + 2 + (s ^ 1), true));
      }
    }
  }
  bulk_wait0();
}

template <int N, int NOISE>
__global__ void __launch_bounds__(PCfg<N>::THREADS, 1)
kp_gen(const PassArgs p, const cplx* __restrict__ tw2,
       const __grid_constant__ TileMaps tms)
{
  using P = PCfg<N>;
  extern __shared__ unsigned char smem_raw[];
  PSmem<N> sm(smem_raw);
  pipe_init<N>(sm, true);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ntA = (p.A + P::LT - 1) / P::LT;
  const long long ntiles = (long long)ntA * p.B;
  if (warp == P::CW) {
    store_warp_loop_gen<N>(p, &tms, sm, ntA, ntiles);
    return;
  }
  const int slot = warp * P::LPW + lane / P::TPL, t = lane % P::TPL;
  cplx* buf = sm.slot(slot);
  uint32_t it = 0;
  for (long long T = blockIdx.x; T < ntiles; T += gridDim.x) {
    const int at = (int)(T % ntA), b = (int)(T / ntA);
    const int a = at * P::LT + slot;
    cplx v[P::E];
    // lines beyond A recompute the last valid line; the tensor store clips
    // them (A ends either at the global extent or on an 8-line boundary)
    load_noise<N, NOISE>(v, p, p.a_begin + min(a, p.A - 1), b, t, true);
    line_fft<N, +1>(v, buf, t, tw2);
    stage_out<N>(v, sm, slot, t, p.keep, it);
  }
}

// ===========================================================================
// pass A, grouped form (kp_gen3): one CTA per SM, three independent groups of
// four warps (the register file holds twelve 168-register FFT warps and no
// more, so no warp is spent on a store role).  Per tile of LPC lines a group
//   1. starts bulk copies of the lines' sigma rows into its landing zone,
//   2. generates the unit normals in registers (Philox + Box-Muller),
//   3. waits for the rows, scales, runs the line FFTs,
//   4. stages the transposed tile with the TMA swizzle and hands it to ONE
//      tensor store, which drains while the group is already generating the
//      next tile's noise (the stage is only needed again at the next FFT).
// Shared memory per group = the exchange buffers of its lines (66 KB), reused
// as stage [0, 32 KB) and sigma landing zone [32 KB, 64 KB).
// ===========================================================================
template <int N>
struct G3Cfg
{
  using K = KCfg<N>;
  using C = LineCfg<N>;
  static constexpr int GROUPS = 3;
  static constexpr int THREADS = GROUPS * PF_THREADS;
  static constexpr int ROWB = K::LPC * 16;          // bytes per stage row
  static constexpr int REGION = (int)K::exch_bytes; // per group
  static constexpr int MAXKEEP = N / 2;
  static constexpr int STAGE_BYTES = MAXKEEP * ROWB;
  static constexpr int SIG_OFF = STAGE_BYTES;
  static constexpr int SIG_LINE = N * 8;
  static constexpr int BOXR = MAXKEEP < 256 ? MAXKEEP : 256;
  // the table of inter-pass twiddles is copied to shared memory: with ~200 KB
  // of shared memory in use the remaining L1 (~20 KB) misses on 10 % of the
  // table loads (ncu, profiles/r2a), each a trip to L2
  static constexpr int TWB = C::TW2 * 16;
  static constexpr int SMEM = GROUPS * REGION + 64 + TWB + 1024;
  static_assert(K::LPC == 4 || K::LPC == 8, "64- or 128-byte stage rows");
  static_assert(REGION % 1024 == 0, "swizzle atoms need aligned regions");
  static_assert(SIG_OFF + K::LPC * SIG_LINE <= REGION, "landing zone");
  static_assert(MAXKEEP % BOXR == 0, "whole boxes inside the stage");
};

__device__ __forceinline__ void
group_sync(int g)
{
  asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "r"(PF_THREADS) : "memory");
}

// HALF: keep == N / 2 (embedding factor 2, the usual case) is known at compile
// time, so the upper half of the transform's outputs is dead code: the last
// butterflies that would produce it and the per-element bound checks go
// (ncu source view, profiles/r2k: 192 of the 263 instructions between the FFT
// and the tensor store were swizzle addresses and bound checks)
template <int N, int NOISE, bool HALF = false>
__global__ void __launch_bounds__(G3Cfg<N>::THREADS, 1)
kp_gen3(const PassArgs p, const cplx* __restrict__ tw2,
        const __grid_constant__ TileMaps tms)
{
  using G = G3Cfg<N>;
  using K = KCfg<N>;
  using C = LineCfg<N>;
  extern __shared__ unsigned char smem_raw[];
  unsigned char* base = align1024(smem_raw);
  const int g = threadIdx.x / PF_THREADS, tg = threadIdx.x % PF_THREADS;
  const int lic = tg / C::TPL, t = tg % C::TPL;
  unsigned char* region = base + g * G::REGION;
  cplx* stage = reinterpret_cast<cplx*>(region);
  cplx* xbuf = stage + lic * C::PITCH;
  const double* srow =
    reinterpret_cast<const double*>(region + G::SIG_OFF + lic * G::SIG_LINE);
  const uint32_t bar = smem_u32(base + G::GROUPS * G::REGION) + 8u * g;
  cplx* tws = reinterpret_cast<cplx*>(base + G::GROUPS * G::REGION + 64);
  for (int i = threadIdx.x; i < C::TW2; i += G::THREADS)
    tws[i] = __ldg(tw2 + i);
  if (tg == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_proxy_async();
  }
  __syncthreads();
  const int ntA = (p.A + K::LPC - 1) / K::LPC;
  const long long ntiles = (long long)ntA * p.B;
  const uint32_t sig_bytes = (uint32_t)(p.tab_mirror_e ? N / 2 + 2 : N) * 8u;
  const int a_last = p.a_begin + p.A - 1;
  const int keep = p.keep;
  uint32_t ph = 0;
  bool pending = false;
  for (long long T = (long long)blockIdx.x * G::GROUPS + g; T < ntiles;
       T += (long long)gridDim.x * G::GROUPS) {
    const int at = (int)(T % ntA), b = (int)(T / ntA);
    const int a0 = p.a_begin + at * K::LPC;
    // lines beyond A recompute the last valid line; the tensor store clips
    const int a = min(a0 + lic, a_last);
    // issued by the second warp: the first one issues the tensor stores
    if (tg == 32) {
      mbar_expect_tx(bar, K::LPC * sig_bytes);
#pragma unroll
      for (int l = 0; l < K::LPC; l++)
        bulk_load_keep(smem_u32(region + G::SIG_OFF + l * G::SIG_LINE),
                       table_row(p, min(a0 + l, a_last), b), sig_bytes, bar);
    }
    cplx v[C::E];
    gen_noise<N, NOISE>(v, p, a, b, t);
    mbar_wait(bar, ph);
    ph ^= 1;
    scale_noise<N>(v, p, srow, a, b, t);
    // the previous tile's tensor store must have read the stage before the
    // exchange buffers (same memory) are written again
    if (tg == 0 && pending)
      bulk_wait_read0();
    group_sync(g);
    line_fft<N, +1, true>(v, xbuf, t, tws);
    group_sync(g);
    {
      // element e = t + TPL q of line lic: the swizzle term of stage_slot
      // does not change with q, the slot advances by TPL * LPC per step
      cplx* sp = stage + stage_slot<N>(t, lic);
#pragma unroll
      for (int q = 0; q < (HALF ? C::E / 2 : C::E); q++)
        if (HALF || t + C::TPL * q < keep)
          sp[q * (C::TPL * K::LPC)] = v[q];
    }
    fence_proxy_async();
    // only the storing thread has to see the whole tile: the other warps
    // announce their rows and go on to the next tile's noise (registers only).
    // Own barrier id: a warp that runs ahead waits at the next group_sync,
    // which must not be taken for the missing arrivals of this one.
    if (tg < 32)
      asm volatile("bar.sync %0, %1;" ::"r"(g + 1 + G::GROUPS), "r"(PF_THREADS)
                   : "memory");
    else
      asm volatile("bar.arrive %0, %1;" ::"r"(g + 1 + G::GROUPS), "r"(PF_THREADS)
                   : "memory");
    if (tg == 0) {
#pragma unroll
      for (int r0 = 0; r0 < G::MAXKEEP; r0 += G::BOXR)
        if (HALF || r0 < keep)
          tma_store_3d(&tms.m[0], smem_u32(region) + r0 * G::ROWB,
                       2 * at * K::LPC, r0, b);
      bulk_commit();
      pending = true;
    }
  }
  if (tg == 0 && pending)
    bulk_wait0();
}

// ===========================================================================
// pass B / F1 / E1: contiguous complex lines -> FFT -> transposed tile (k_c2t)
// ===========================================================================
// output tile map and (blocked exchange layout only) input line map
struct C2TMaps
{
  CUtensorMap out, in;
};

template <int N, int DIR>
__global__ void __launch_bounds__(PCfg<N>::THREADS, 1)
kp_c2t(const PassArgs p, const cplx* __restrict__ tw2,
       const __grid_constant__ C2TMaps maps)
{
  const CUtensorMap& tm = maps.out;
  using P = PCfg<N>;
  extern __shared__ unsigned char smem_raw[];
  PSmem<N> sm(smem_raw);
  if constexpr (P::TWS_C2T) {
    cplx* tws = reinterpret_cast<cplx*>(sm.base + P::SLOTS_BYTES +
                                        P::NSTAGE * P::STAGE_BYTES + P::BAR_BYTES);
    for (int i = threadIdx.x; i < P::C::TW2; i += P::THREADS)
      tws[i] = __ldg(tw2 + i);
    tw2 = tws;
  }
  pipe_init<N>(sm, true);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ntA = (p.A + P::LT - 1) / P::LT;
  const long long ntiles = (long long)ntA * p.B;
  if (warp == P::CW) {
    store_warp_loop<N>(p, &tm, sm, ntA, ntiles);
    return;
  }
  const int slot = warp * P::LPW + lane / P::TPL, t = lane % P::TPL;
  cplx* buf = sm.slot(slot);
  const uint32_t inbar = sm.bar(warp, true);
  const cplx* in = reinterpret_cast<const cplx*>(p.in);
  const uint32_t line_bytes = (uint32_t)p.n_valid * 16u;
  // lane 0 arms the warp's barrier and copies the warp's lines of tile T
  auto issue = [&](long long T) {
    const int at = (int)(T % ntA), b = (int)(T / ntA);
    const int a0 = at * P::LT + warp * P::LPW;
    const int nv = max(0, min(P::LPW, p.A - a0));
    if (lane == 0) {
      fence_proxy_async();
      mbar_expect_tx(inbar, (uint32_t)nv * line_bytes);
      for (int l = 0; l < nv; l++) {
        if (p.blk_T > 0) // line (plane a, row b): blk_T segments of 128 bytes
          tma_load_4d(&maps.in, smem_u32(sm.slot(warp * P::LPW + l)), inbar, 0,
                      b, 0, a0 + l);
        else
          bulk_load(smem_u32(sm.slot(warp * P::LPW + l)),
                    in + (long long)(a0 + l) * p.isa + (long long)b * p.isb,
                    line_bytes, inbar);
      }
    }
  };
  uint32_t it = 0, ph = 0;
  long long T = blockIdx.x;
  if (T < ntiles)
    issue(T);
  for (; T < ntiles; T += gridDim.x) {
    const int at = (int)(T % ntA);
    const int a = at * P::LT + slot;
    const bool valid = a < p.A;
    cplx v[P::E];
    mbar_wait(inbar, ph);
    ph ^= 1;
#pragma unroll
    for (int q = 0; q < P::E; q++) {
      const int e = t + P::TPL * q;
      v[q] = (valid && e < p.n_valid) ? buf[e] : make_double2(0., 0.);
    }
    __syncwarp();
    line_fft_phase1<N, DIR>(v, buf, t);
    __syncwarp();
    line_fft_phase2_load<N>(v, buf, t);
    __syncwarp();
    // the exchange area is free again: start the next tile's copy (generic
    // accesses of every lane ordered before the async-proxy write)
    fence_proxy_async();
    __syncwarp();
    const long long Tn = T + gridDim.x;
    if (Tn < ntiles)
      issue(Tn);
    line_fft_phase2_compute<N, DIR, P::TWS_C2T>(v, t, tw2);
    stage_out<N>(v, sm, slot, t, p.keep, it);
  }
}

// ===========================================================================
// pass C: pairs of half-spectrum lines -> one complex inverse FFT -> two real
// lines, restricted, value transform fused (k_c2r)
// ===========================================================================
template <int N>
__global__ void __launch_bounds__(PCfg<N>::THREADS - 32, 1)
kp_c2r(const PassArgs p, const cplx* __restrict__ tw2)
{
  using P = PCfg<N>;
  constexpr int M0 = N / 2 + 1;
  extern __shared__ unsigned char smem_raw[];
  PSmem<N> sm(smem_raw);
  if constexpr (P::TWS_C2R) {
    cplx* tws = reinterpret_cast<cplx*>(sm.base + P::SLOTS_BYTES + P::BAR_BYTES);
    for (int i = threadIdx.x; i < P::C::TW2; i += P::THREADS - 32)
      tws[i] = __ldg(tw2 + i);
    tw2 = tws;
  }
  pipe_init<N>(sm, false);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long nl = (long long)p.A * p.B;
  const long long npairs = (nl + 1) / 2;
  const long long ntiles = (npairs + P::LT - 1) / P::LT;
  const int slot = warp * P::LPW + lane / P::TPL, t = lane % P::TPL;
  cplx* buf = sm.slot(slot);
  const uint32_t inbar = sm.bar(warp, false);
  const cplx* in = reinterpret_cast<const cplx*>(p.in);
  constexpr uint32_t line_bytes = (uint32_t)M0 * 16u;
  auto issue = [&](long long T) {
    const long long L = 2 * (T * P::LT + warp * P::LPW); // first line
    const long long nv = max(0ll, min((long long)2 * P::LPW, nl - L));
    if (lane == 0) {
      fence_proxy_async();
      mbar_expect_tx(inbar, (uint32_t)nv * line_bytes);
      for (int l = 0; l < (int)nv; l++)
        bulk_load(smem_u32(sm.slot(warp * P::LPW + (l >> 1))) +
                    (l & 1) * line_bytes,
                  in + (L + l) * p.isa, line_bytes, inbar);
    }
  };
  uint32_t ph = 0;
  long long T = blockIdx.x;
  if (T < ntiles)
    issue(T);
  for (; T < ntiles; T += gridDim.x) {
    const long long L0 = 2 * (T * P::LT + slot), L1 = L0 + 1;
    const bool v0 = L0 < nl, v1 = L1 < nl;
    cplx v[P::E];
    mbar_wait(inbar, ph);
    ph ^= 1;
#pragma unroll
    for (int q = 0; q < P::E; q++) {
      const int e = t + P::TPL * q;
      const int kk = e <= N / 2 ? e : N - e;
      const cplx xa = v0 ? buf[kk] : make_double2(0., 0.);
      const cplx xb = v1 ? buf[M0 + kk] : make_double2(0., 0.);
      if (e == 0 || 2 * e == N)
        v[q] = make_double2(xa.x, xb.x);
      else if (e < N / 2)
        v[q] = make_double2(xa.x - xb.y, xa.y + xb.x);
      else
        v[q] = make_double2(xa.x + xb.y, xb.x - xa.y);
    }
    __syncwarp();
    line_fft_phase1<N, +1>(v, buf, t);
    __syncwarp();
    line_fft_phase2_load<N>(v, buf, t);
    fence_proxy_async();
    __syncwarp();
    const long long Tn = T + gridDim.x;
    if (Tn < ntiles)
      issue(Tn);
    line_fft_phase2_compute<N, +1, P::TWS_C2R>(v, t, tw2);
    store_real_pair<N>(v, p, L0, v0, v1, t);
  }
}

// ===========================================================================
// long c2r axis, step 1 (k_long_c2r1 of pf_kernels.cuh) as a persistent kernel:
// the plain kernel leaves its loads exposed (3.2 TB/s at K4, ncu:
// profiles/r2o).  Here a PAIR of warps takes the two mirror lines k1 and
// N1 - k1 of one line pair (L0, L1): both are assembled from the same four rows
// of the permuted half spectra, which arrive by bulk copies into a region that
// afterwards serves as the two warps' exchange buffers; the next item's rows
// are requested as soon as both warps hold their second-phase operands.
// Six pairs (twelve 168-register warps) per SM.
// ===========================================================================
template <int N>
struct LC1Cfg
{
  using C = LineCfg<N>;
  static constexpr int PAIRS = 6;
  static constexpr int THREADS = PAIRS * 64;
  static constexpr int M = N / 2 + 1;
  static constexpr int ROWB = M * 16;
  static constexpr int XB = C::PITCH * 16;
  static constexpr int REGION =
    ((4 * ROWB > 2 * XB ? 4 * ROWB : 2 * XB) + 127) / 128 * 128;
  static constexpr int SMEM = PAIRS * REGION + 64 + 128;
  static_assert(C::TPL == 32, "one warp per line");
};

template <int N>
__global__ void __launch_bounds__(LC1Cfg<N>::THREADS, 1)
kp_long_c2r1(const PassArgs p, const cplx* __restrict__ tw2)
{
  using L = LC1Cfg<N>;
  using C = LineCfg<N>;
  extern __shared__ unsigned char smem_raw[];
  unsigned char* base = reinterpret_cast<unsigned char*>(
    (reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  const int warp = threadIdx.x >> 5, t = threadIdx.x & 31;
  const int pair = warp >> 1, w = warp & 1;
  unsigned char* region = base + pair * L::REGION;
  cplx* xbuf = reinterpret_cast<cplx*>(region + w * L::XB);
  const uint32_t bar = smem_u32(base + L::PAIRS * L::REGION) + 8u * pair;
  if (threadIdx.x == 0) {
    for (int i = 0; i < L::PAIRS; i++)
      mbar_init(smem_u32(base + L::PAIRS * L::REGION) + 8u * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_proxy_async();
  }
  __syncthreads();
  const int N1 = p.long_n1;
  const int per_b = N1 / 2 + 1;                    // items per line pair
  const long long nitems = (long long)per_b * p.B;
  const cplx* in = reinterpret_cast<const cplx*>(p.in);
  auto pair_sync = [&]() {
    asm volatile("bar.sync %0, %1;" ::"r"(pair + 1), "r"(64) : "memory");
  };
  // rows 0, 1: line L0 at k1 and at kr; rows 2, 3: line L1
  auto issue = [&](long long T) {
    const long long b = T / per_b;
    const int k1 = (int)(T % per_b), kr = (N1 - k1) % N1;
    const long long L0 = 2 * b, L1 = L0 + 1;
    const bool v1 = L1 < p.n_valid, two = kr != k1;
    const uint32_t rows = (v1 ? 2u : 1u) * (two ? 2u : 1u);
    fence_proxy_async();
    mbar_expect_tx(bar, rows * (uint32_t)L::ROWB);
    bulk_load(smem_u32(region), in + L0 * p.isa + (long long)k1 * L::M, L::ROWB, bar);
    if (two)
      bulk_load(smem_u32(region) + L::ROWB, in + L0 * p.isa + (long long)kr * L::M,
                L::ROWB, bar);
    if (v1) {
      bulk_load(smem_u32(region) + 2 * L::ROWB, in + L1 * p.isa + (long long)k1 * L::M,
                L::ROWB, bar);
      if (two)
        bulk_load(smem_u32(region) + 3 * L::ROWB,
                  in + L1 * p.isa + (long long)kr * L::M, L::ROWB, bar);
    }
  };
  uint32_t ph = 0;
  long long T = (long long)blockIdx.x * L::PAIRS + pair;
  const long long dT = (long long)gridDim.x * L::PAIRS;
  if (T < nitems && w == 0 && t == 0)
    issue(T);
  for (; T < nitems; T += dT) {
    const long long b = T / per_b;
    const int k1 = (int)(T % per_b), kr = (N1 - k1) % N1;
    const bool two = kr != k1;
    const int k = w == 0 ? k1 : kr;       // this warp's line
    const bool mine = w == 0 || two;      // self-mirror items have one line
    const bool v1 = 2 * b + 1 < p.n_valid;
    // direct rows hold k, mirror rows hold N1 - k
    const cplx* a_dir = reinterpret_cast<const cplx*>(region + (w == 0 ? 0 : 1) * L::ROWB);
    const cplx* a_mir =
      reinterpret_cast<const cplx*>(region + ((w == 0 && two) ? 1 : 0) * L::ROWB);
    const cplx* b_dir = a_dir + 2 * L::M;
    const cplx* b_mir = a_mir + 2 * L::M;
    cplx v[C::E];
    mbar_wait(bar, ph);
    ph ^= 1;
    if (mine) {
#pragma unroll
      for (int q = 0; q < C::E; q++) {
        const int e = t + C::TPL * q;
        const bool direct = (2 * e < N) || (k == 0 && 2 * e == N);
        const int col = direct ? e : (k == 0 ? N - e : N - 1 - e);
        const cplx xa = (direct ? a_dir : a_mir)[col];
        const cplx xb = v1 ? (direct ? b_dir : b_mir)[col] : make_double2(0., 0.);
        if (k == 0 && (e == 0 || 2 * e == N))
          v[q] = make_double2(xa.x, xb.x);
        else if (direct)
          v[q] = make_double2(xa.x - xb.y, xa.y + xb.x);
        else
          v[q] = make_double2(xa.x + xb.y, xb.x - xa.y);
      }
    }
    pair_sync(); // both warps hold their lines: the rows become exchange space
    if (mine) {
      line_fft_phase1<N, +1>(v, xbuf, t);
      __syncwarp();
      line_fft_phase2_load<N>(v, xbuf, t);
    }
    fence_proxy_async(); // generic accesses of the region before the bulk copies
    pair_sync();         // the region is free for the next item's rows
    if (w == 0 && t == 0 && T + dT < nitems)
      issue(T + dT);
    if (mine) {
      line_fft_phase2_compute<N, +1, false>(v, t, tw2);
      cplx* out = reinterpret_cast<cplx*>(p.out) + (b * N1 + k) * N;
      const cplx* ltw = p.long_tw + (long long)k * N;
#pragma unroll
      for (int q = 0; q < C::E; q++) {
        const int j2 = t + C::TPL * q;
        out[j2] = cmulc(v[q], __ldg(ltw + j2));
      }
    }
  }
}

} // namespace pf
