// pf_stream.cuh -- the "streaming" form of generateField for large cubic 3D
// tori (n0 = n1 = n2 = N in {512, 1024}, embedding factor 2; BASELINE configs
// K3 and K2/K5): the backward transform of r2cfieldbackend.hh:250-274 as
//
//   kp_ab  : ONE persistent kernel for the first two axes.
//            A-lines (k0 = plane p, k1): Philox noise * sigma -> inverse FFT
//            along k2 -> the kept half x2 < c2 is written CONTIGUOUSLY into a
//            ring of R planes  ring[p % R][k1][x2]  that lives in the 126 MB
//            L2 (R * 8.4 MB at K3); every warp owns its lines, no CTA-level
//            barrier, no staging tile.
//            B-items (plane p - LAG, four/eight lines x2): a TMA box load
//            gathers the transposed tile [k1][x2 block] from the ring (64/128
//            byte rows, hardware swizzle), inverse FFT along k1, the kept half
//            x1 < c1 is written contiguously to  Zt[k0][x2][x1].
//            Work items are handed out by a global ticket in an order that
//            interleaves plane p's A-lines with plane (p - LAG)'s B-items;
//            per-plane completion counters in global memory order producer
//            and consumer (an item only ever waits for items with smaller
//            tickets, so the scheme cannot deadlock whatever the residency).
//            The ring is rewritten before L2 evicts it: the 2 x 4.3 GB of
//            HBM traffic between the first two passes disappear, and the
//            memory-bound pass B hides behind the issue-bound pass A.
//   kp_c2rt: last axis.  A TMA box load gathers [k0][x1 block] from Zt (128
//            byte segments), two half-spectrum lines per complex FFT (FFTW's
//            c2r semantics on non-Hermitian input), restriction and value
//            transform fused, contiguous stores.
//
// Arithmetic per line is that of pf_kernels.cuh (gen_noise / scale_noise /
// line_fft / store_real_pair), so fields are bit-identical to the three-pass
// route; only the data movement differs.
#pragma once
#include <cuda.h>

#include "pf_pipe.cuh"

namespace pf {

template <int N>
struct SCfg
{
  using C = LineCfg<N>;
  static constexpr int E = C::E;
  static constexpr int TPL = C::TPL;
  static constexpr int LPW = 32 / TPL;       // lines per warp (1 or 2)
  static constexpr int GW = 4;               // warps per group
  static constexpr int GROUPS = 3;           // groups per CTA
  static constexpr int LPG = GW * LPW;       // lines (FFTs) per group item
  static constexpr int THREADS = GROUPS * GW * 32;
  static constexpr int XBUF = C::PITCH * 16; // exchange buffer of one line
  static constexpr int WARP_BYTES = LPW * XBUF;
  static constexpr int REGION = GW * WARP_BYTES; // per group
  static constexpr int ROWB = LPG * 16;      // row of a transposed B tile
  static constexpr int M0 = N / 2 + 1;
  static constexpr int BOXR = 256;           // rows per TMA box
  static constexpr int CROWS = (M0 + 7) / 8 * 8; // rows of a C tile
  static constexpr int CTILES = LPG / 4;     // x1 octets per C item
  static constexpr int CTRL = 512;           // barriers + ticket slots
  static constexpr int TWB = C::TW2 * 16;    // inter-pass twiddles, shared copy
  static constexpr int SMEM = GROUPS * REGION + CTRL + TWB + 1024;
  static_assert(N == 512 || N == 1024, "streaming route: lines of 512 / 1024");
  static_assert(REGION % 1024 == 0, "swizzled tiles need 1024-byte regions");
  static_assert(N * ROWB <= REGION, "B tile must fit the group's region");
  static_assert(CTILES * CROWS * 128 <= REGION, "C tiles must fit the region");
};

// arguments of kp_ab; `a` is pass A as generate_dev describes it (lines
// a = k1 < A, b = k0 < B, elements k2; table, noise, Hermitian rule)
struct StreamArgs
{
  PassArgs a;
  cplx* ring;          // [slots][n1][c2]
  int slots;
  cplx* zt;            // pass B output, element (k0, x2, x1) at
  long long zs_k0, zs_x2, zs_x1; //   k0 * zs_k0 + x2 * zs_x2 + x1 * zs_x1
  int c1, c2;
  // multi-GPU (one process per GPU): this rank computes the planes
  // plane0 .. plane0 + planes - 1 of k0 (a.B = planes); the lines x2 of pass
  // B's output belong to rank x2 / rows_per_peer and are stored into that
  // rank's buffer peer_zt[] (peer memory over NVLink, x2 taken modulo
  // rows_per_peer) -- the exchange of FFTW-MPI's transposed transform fused
  // into the producing pass, after BOTH prunings.  n_peers <= 1: zt alone.
  int plane0;
  int n_peers, rows_per_peer;
  cplx* peer_zt[8];
  int lag;             // B-items run `lag` planes behind the A-lines
  unsigned* sync;      // [0] ticket, [16 + p] A-lines done, [16 + m0 + p] B-items done
};

__device__ __forceinline__ uint64_t
l2_policy_evict_last()
{
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ uint64_t
l2_policy_evict_first()
{
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void
st_hint(cplx* p, const cplx& v, uint64_t pol)
{
  asm volatile("st.global.L2::cache_hint.v2.f64 [%0], {%1, %2}, %3;" ::"l"(p),
               "d"(v.x), "d"(v.y), "l"(pol)
               : "memory");
}
__device__ __forceinline__ unsigned
ld_acquire(const unsigned* p)
{
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void
spin_until(const unsigned* p, unsigned target)
{
  unsigned ns = 32;
  while (ld_acquire(p) < target) {
    __nanosleep(ns);
    if (ns < 1024)
      ns *= 2;
  }
}
// global -> shared tensor load of one 3D box, completion counted on `bar`
__device__ __forceinline__ void
tma_load_3d(const CUtensorMap* tm, uint32_t dst, uint32_t bar, int c0, int c1,
            int c2)
{
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::"
               "complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
               "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
__device__ __forceinline__ void
tma_load_3d_keep(const CUtensorMap* tm, uint32_t dst, uint32_t bar, int c0,
                 int c1, int c2, uint64_t pol)
{
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::"
               "complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4, %5}], [%2], %6;"
               ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "l"(pol)
               : "memory");
}

// byte offset of (row, 16-byte chunk) in a TMA-swizzled tile with rows of
// ROWB bytes (64: SWIZZLE_64B, 128: SWIZZLE_128B)
template <int ROWB>
__device__ __forceinline__ uint32_t
swz_off(int row, int chunk)
{
  if constexpr (ROWB == 64)
    return (uint32_t)row * 64u + (uint32_t)((chunk ^ ((row >> 1) & 3)) << 4);
  else
    return (uint32_t)row * 128u + (uint32_t)((chunk ^ (row & 7)) << 4);
}

// phase layout of the ticket order: phase p holds plane p's A-items and plane
// (p - lag)'s B-items, interleaved evenly (Bresenham)
struct StreamItem
{
  int plane;   // plane of the item, < 0: nothing to do
  int index;   // A: first line / LPG, B: x2 block
  bool is_b;
};
__device__ __forceinline__ StreamItem
stream_item(unsigned long long T, int tA, int tB, int m0, int lag)
{
  const int PH = tA + tB;
  const int p = (int)(T / (unsigned)PH), r = (int)(T % (unsigned)PH);
  const int jb0 = (int)(((long long)r * tB) / PH);
  const int jb1 = (int)(((long long)(r + 1) * tB) / PH);
  StreamItem it;
  it.is_b = jb1 > jb0;
  if (it.is_b) {
    it.plane = p - lag;
    it.index = jb0;
    if (it.plane < 0 || it.plane >= m0)
      it.plane = -1;
  } else {
    it.plane = p < m0 ? p : -1;
    it.index = r - jb0;
  }
  return it;
}

// non-blocking dependency test of an item (see the ticket order above)
__device__ __forceinline__ bool
stream_ready(const StreamItem& w, const unsigned* doneA, const unsigned* doneB,
             int n1, int tB, int slots)
{
  if (w.plane < 0)
    return true;
  if (w.is_b)
    return ld_acquire(doneA + w.plane) >= (unsigned)n1;
  if (w.plane >= slots)
    return ld_acquire(doneB + w.plane - slots) >= (unsigned)tB;
  return true;
}

// HALF: c1 == c2 == N / 2 (embedding factor 2) known at compile time -- the
// upper half of the outputs and the bound checks are dead code (as kp_gen3)
template <int N, int NOISE, bool HALF = false>
__global__ void __launch_bounds__(SCfg<N>::THREADS, 1)
kp_ab(const __grid_constant__ StreamArgs s, const cplx* __restrict__ tw2,
      const __grid_constant__ CUtensorMap ring_map)
{
  using S = SCfg<N>;
  using C = LineCfg<N>;
  extern __shared__ unsigned char smem_raw[];
  unsigned char* base = align1024(smem_raw);
  const PassArgs& p = s.a;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = warp / S::GW, wg = warp % S::GW;  // group, warp in group
  const int tg = threadIdx.x % (S::GW * 32);
  const int lw = lane / S::TPL, t = lane % S::TPL; // line in warp, thread in line
  const int lic = wg * S::LPW + lw;                 // line in group item
  unsigned char* region = base + g * S::REGION;
  cplx* xbuf = reinterpret_cast<cplx*>(region + lic * S::XBUF);
  unsigned char* ctrl = base + S::GROUPS * S::REGION;
  const uint32_t wbar = smem_u32(ctrl) + 8u * warp;        // sigma rows of a warp
  const uint32_t gbar = smem_u32(ctrl) + 128u + 8u * g;    // B tile of a group
  volatile long long* tslot = reinterpret_cast<volatile long long*>(ctrl + 256) + 2 * g;
  // the 16 KB table of inter-pass twiddles: with ~200 KB of shared memory in
  // use the L1 that remains (~20 KB) keeps missing on it, a shared copy does not
  cplx* tws = reinterpret_cast<cplx*>(ctrl + S::CTRL);
  for (int i = threadIdx.x; i < C::TW2; i += S::THREADS)
    tws[i] = __ldg(tw2 + i);
  if (threadIdx.x == 0) {
    for (int i = 0; i < S::GROUPS * S::GW; i++)
      mbar_init(smem_u32(ctrl) + 8u * i, 1);
    for (int i = 0; i < S::GROUPS; i++)
      mbar_init(smem_u32(ctrl) + 128u + 8u * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_proxy_async();
  }
  __syncthreads();
  const int n1 = p.A, m0 = p.B, c1 = s.c1, c2 = s.c2;
  const int tA = n1 / S::LPG, tB = c2 / S::LPG;
  const long long total = (long long)(m0 + s.lag) * (tA + tB);
  unsigned* ticket = s.sync;
  unsigned* doneA = s.sync + 16;
  unsigned* doneB = s.sync + 16 + m0;
  const uint32_t sig_bytes = (uint32_t)(p.tab_mirror_e ? N / 2 + 2 : N) * 8u;
  const long long plane_elems = (long long)n1 * c2;
  const uint64_t keep_pol = l2_policy_evict_last();
  const uint64_t stream_pol = l2_policy_evict_first();
  constexpr long long FLUSH = -2;
  uint32_t wph = 0, gph = 0, it = 0;
  // state of the ticket thread (tg == 0): the ticket in hand and the plane of
  // the group's previous A-item, whose completion is published after the next
  // barrier (every warp's stores precede its arrival there)
  long long T_mine = 0;
  int pendA = -1;
  if (tg == 0)
    T_mine = (long long)atomicAdd(ticket, 1u);
  for (;; it++) {
    if (tg == 0) {
      long long code = T_mine;
      if (T_mine < total) {
        const StreamItem w = stream_item((unsigned long long)T_mine, tA, tB, m0, s.lag);
        if (!stream_ready(w, doneA, doneB, n1, tB, s.slots)) {
          if (pendA >= 0)
            code = FLUSH; // never wait while holding back a completion signal
          else {
            unsigned ns = 32;
            while (!stream_ready(w, doneA, doneB, n1, tB, s.slots)) {
              __nanosleep(ns);
              if (ns < 1024)
                ns *= 2;
            }
          }
        }
      }
      tslot[it & 1] = code;
    }
    // generic accesses of the previous item before async-proxy writes of this one
    fence_proxy_async();
    group_sync(g);
    const long long T = tslot[it & 1];
    if (tg == 0 && pendA >= 0) {
      // every warp of the group issued its stores before the barrier: one
      // gpu-scope fence publishes the whole item
      __threadfence();
      atomicAdd(doneA + pendA, (unsigned)S::LPG);
      pendA = -1;
    }
    if (T == FLUSH)
      continue;
    if (T >= total)
      break;
    const StreamItem w = stream_item((unsigned long long)T, tA, tB, m0, s.lag);
    if (tg == 0) // next ticket: the atomic's latency hides behind this item
      T_mine = (long long)atomicAdd(ticket, 1u);
    if (w.plane < 0)
      continue;
    cplx v[C::E];
    cplx* out;
    long long ostep;
    int keep;
    uint64_t pol;
    if (!w.is_b) {
      // ================= A: noise * sigma (k0 = plane, k1 = a) ==============
      const int b = s.plane0 + w.plane;        // global k0
      const int a = w.index * S::LPG + lic;
      if (lane == 0) {
        mbar_expect_tx(wbar, S::LPW * sig_bytes);
#pragma unroll
        for (int l = 0; l < S::LPW; l++)
          bulk_load(smem_u32(region + (wg * S::LPW + l) * S::XBUF),
                    table_row(p, w.index * S::LPG + wg * S::LPW + l, b),
                    sig_bytes, wbar);
      }
      gen_noise<N, NOISE>(v, p, a, b, t);
      mbar_wait(wbar, wph);
      wph ^= 1;
      scale_noise<N>(v, p, reinterpret_cast<const double*>(xbuf), a, b, t);
      out = s.ring + (long long)(w.plane % s.slots) * plane_elems +
            (long long)a * c2 + t;
      ostep = S::TPL;
      keep = c2;
      pol = keep_pol;
      if (tg == 0)
        pendA = w.plane;
    } else {
      // ================= B: transposed tile of the ring (k0 = plane, x2) ====
      const int pb = w.plane;
      const int x2 = w.index * S::LPG + lic;
      if (tg == 0) {
        fence_proxy_async();                   // generic acquire -> async reads
        asm volatile("fence.proxy.async.global;" ::: "memory");
        mbar_expect_tx(gbar, (uint32_t)N * S::ROWB);
#pragma unroll
        for (int r0 = 0; r0 < N; r0 += S::BOXR)
          tma_load_3d_keep(&ring_map, smem_u32(region) + r0 * S::ROWB, gbar,
                           2 * w.index * S::LPG, r0, pb % s.slots, keep_pol);
      }
      mbar_wait(gbar, gph);
      gph ^= 1;
      {
        // row e = t + TPL q: the swizzle term does not change with q
        const unsigned char* src = region + swz_off<S::ROWB>(t, lic);
#pragma unroll
        for (int q = 0; q < C::E; q++)
          v[q] = *reinterpret_cast<const cplx*>(src + q * (S::TPL * S::ROWB));
      }
      group_sync(g); // every warp holds its lines: the region becomes exchange space
      if (tg == 0)
        atomicAdd(doneB + pb, 1u);
      if (s.n_peers > 1) {
        const int owner = x2 / s.rows_per_peer;
        out = s.peer_zt[owner] + (long long)(s.plane0 + pb) * s.zs_k0 +
              (long long)(x2 - owner * s.rows_per_peer) * s.zs_x2 +
              (long long)t * s.zs_x1;
      } else
        out = s.zt + (long long)pb * s.zs_k0 + (long long)x2 * s.zs_x2 +
              (long long)t * s.zs_x1;
      ostep = (long long)S::TPL * s.zs_x1;
      keep = c1;
      pol = stream_pol;
    }
    // ================= common: inverse FFT of the line, kept half out ========
    __syncwarp();
    line_fft<N, +1, true>(v, xbuf, t, tws);
#pragma unroll
    for (int q = 0; q < (HALF ? C::E / 2 : C::E); q++) {
      if (HALF || t + S::TPL * q < keep)
        st_hint(out, v[q], pol);
      out += ostep;
    }
  }
}

// last axis: Zt[k0][x2][x1] -> field[x2][x1][x0 < keep]
template <int N>
__global__ void __launch_bounds__(SCfg<N>::THREADS, 1)
kp_c2rt(const PassArgs p, const cplx* __restrict__ tw2,
        const __grid_constant__ CUtensorMap map_full,
        const __grid_constant__ CUtensorMap map_tail)
{
  using S = SCfg<N>;
  using C = LineCfg<N>;
  constexpr int M0 = S::M0;
  extern __shared__ unsigned char smem_raw[];
  unsigned char* base = align1024(smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = warp / S::GW, wg = warp % S::GW;
  const int tg = threadIdx.x % (S::GW * 32);
  const int lw = lane / S::TPL, t = lane % S::TPL;
  const int lic = wg * S::LPW + lw;   // FFT (= pair of real lines) in the item
  unsigned char* region = base + g * S::REGION;
  cplx* xbuf = reinterpret_cast<cplx*>(region + lic * S::XBUF);
  unsigned char* ctrl = base + S::GROUPS * S::REGION;
  const uint32_t gbar = smem_u32(ctrl) + 8u * g;
  cplx* tws = reinterpret_cast<cplx*>(ctrl + S::CTRL);
  for (int i = threadIdx.x; i < C::TW2; i += S::THREADS)
    tws[i] = __ldg(tw2 + i);
  if (threadIdx.x == 0) {
    for (int i = 0; i < S::GROUPS; i++)
      mbar_init(smem_u32(ctrl) + 8u * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    fence_proxy_async();
  }
  __syncthreads();
  // p.A = c1 (x1), p.B = c2 (x2); items: (x2, block of 2 * LPG lines x1)
  const int c1 = p.A, c2 = p.B;
  const int bpr = c1 / (2 * S::LPG);              // items per x2 row
  const long long nitems = (long long)bpr * c2;
  const long long stride = (long long)gridDim.x * S::GROUPS;
  constexpr int NFULL = (M0 - 1) / S::BOXR;       // full boxes of 256 rows
  constexpr uint32_t TILE_BYTES = S::CROWS * 128;
  auto issue = [&](long long I) {
    const int x2 = (int)(I / bpr), xb = (int)(I % bpr);
    fence_proxy_async();
    mbar_expect_tx(gbar, S::CTILES * (NFULL * S::BOXR + 8) * 128u);
#pragma unroll
    for (int u = 0; u < S::CTILES; u++) {
      const int col = 2 * (xb * 2 * S::LPG + 8 * u); // doubles
#pragma unroll
      for (int f = 0; f < NFULL; f++)
        tma_load_3d(&map_full, smem_u32(region) + u * TILE_BYTES + f * S::BOXR * 128,
                    gbar, col, x2, f * S::BOXR);
      tma_load_3d(&map_tail, smem_u32(region) + u * TILE_BYTES + NFULL * S::BOXR * 128,
                  gbar, col, x2, NFULL * S::BOXR);
    }
  };
  uint32_t gph = 0;
  long long I = (long long)blockIdx.x * S::GROUPS + g;
  if (I < nitems && tg == 0)
    issue(I);
  for (; I < nitems; I += stride) {
    const int x2 = (int)(I / bpr), xb = (int)(I % bpr);
    const unsigned char* tile = region + (lic / 4) * TILE_BYTES;
    const int ca = 2 * (lic % 4), cb = ca + 1; // chunks of the two real lines
    cplx v[C::E];
    mbar_wait(gbar, gph);
    gph ^= 1;
#pragma unroll
    for (int q = 0; q < C::E; q++) {
      const int e = t + S::TPL * q;
      const int kk = e <= N / 2 ? e : N - e;
      const cplx xa = *reinterpret_cast<const cplx*>(tile + swz_off<128>(kk, ca));
      const cplx xb2 = *reinterpret_cast<const cplx*>(tile + swz_off<128>(kk, cb));
      if (e == 0 || 2 * e == N)
        v[q] = make_double2(xa.x, xb2.x);
      else if (e < N / 2)
        v[q] = make_double2(xa.x - xb2.y, xa.y + xb2.x);
      else
        v[q] = make_double2(xa.x + xb2.y, xb2.x - xa.y);
    }
    group_sync(g); // tile consumed: the region becomes exchange space
    line_fft_phase1<N, +1>(v, xbuf, t);
    __syncwarp();
    line_fft_phase2_load<N>(v, xbuf, t);
    fence_proxy_async();
    group_sync(g); // exchange space free again: fetch the next tile
    const long long In = I + stride;
    if (In < nitems && tg == 0)
      issue(In);
    line_fft_phase2_compute<N, +1, true>(v, t, tws);
    const long long L0 = (long long)x2 * c1 + (long long)xb * 2 * S::LPG + 2 * lic;
    store_real_pair<N>(v, p, L0, true, true, t);
  }
}

} // namespace pf
