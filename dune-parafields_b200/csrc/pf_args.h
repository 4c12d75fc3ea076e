// pf_args.h -- argument block shared by all pass kernels (host + device).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pf {

// covariance ids = the reference's names (matrix.hh:241-266)
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
  COV_MATERN = 12 // general nu, covariance.hh:527-581
};

struct CovParams
{
  int dim;
  int cov;
  int aniso;         // 0 none 1 axiparallel 2 geometric
  double variance;
  double p0, p1;     // expGamma | cauchyAlpha, cauchyBeta | maternNu
  // general Matern (host-prepared, pf_api.cu): [0] sqrt(2 nu), [1] 2^(1-nu) /
  // Gamma(nu), [2] mu = nu - round(nu), [3] round(nu), [4] 1/Gamma(1+mu),
  // [5] 1/Gamma(1-mu), [6] Temme's Gamma_1(mu), [7] Gamma_2(mu)
  double mt[8];
  double h[3];       // meshsize
  double half[3];    // 0.5 * ext * factor  (mirror threshold, matrix.hh:640)
  double period[3];  // ext * factor
  double g[9];       // geometry matrix already inverted (covariance.hh:17-136)
  // merge periodisation (matrix.hh:662-787)
  int periodization; // 0 classical, 1 merge
  int sigmoid;       // 0 smoothstep, 1 smooth
  int combine;       // 0 radialPlus, 1 radial, 2 tensor
  unsigned recursions;
  unsigned factor;
  double s_start, s_end;
  double ext[3];
};

// kernel kinds (dispatch ids of pf_launch)
enum
{
  K_GEN_TILE = 0,   // noise*sigma -> inverse FFT -> tile store      (pass A)
  K_GEN_REIM = 1,   // noise*sigma -> inverse FFT -> re / im lines   (1D)
  K_C2T_INV = 2,    // contiguous complex -> inverse FFT -> tile     (pass B)
  K_C2T_FWD = 3,    // contiguous complex -> forward FFT -> tile     (F1/E1)
  K_C2R = 4,        // two half-spectrum lines -> two real lines     (pass C)
  K_R2C_LOAD = 5,   // two real lines -> half spectra -> tile        (F0)
  K_R2C_COV = 6,    // covariance eval -> half spectra -> tile       (E0)
  K_LAMBDA = 7,     // contiguous complex -> fwd FFT -> eigenvalues  (E2)
  K_LAMBDA_COV = 8, // covariance eval (1D) -> fwd FFT -> eigenvalues
  K_SPEC_TILE = 9,  // contiguous -> fwd FFT * g(lambda) inv FFT -> tile (F2+A)
  K_SPEC_REIM = 10, // real line (1D) -> fwd * g inv -> real line
  K_LONG_C2R1 = 11, // long c2r axis, step 1: permuted half spectra -> sub-FFT
  K_GEN3 = 16       // pass A, grouped persistent form (pipe launcher only)
};

// noise sources of K_GEN_*
enum
{
  NOISE_PHILOX_FAST = 0, // Philox4x32-7 + fp32 (MUFU) Box-Muller
  NOISE_PHILOX_F64 = 1,  // Philox4x32-10 + fp64 Box-Muller
  NOISE_INJECTED = 2     // caller supplied (r1,r2) pairs, reference order
};

// Philox round keys k_r = key + r * Weyl (pf_rng.cuh), computed on the host
struct PhiloxKeys
{
  uint32_t k0[10], k1[10];
};

struct PassArgs
{
  // ---- line space: lines (a, b), a in [0,A) is the tile axis -------------
  int A, B;
  // slab mode (pass A only): this rank owns the global lines
  // a_begin .. a_begin + A - 1; noise counters, table rows and output columns
  // use the global index
  int a_begin;

  // ---- input --------------------------------------------------------------
  const void* in;       // complex or real lines
  long long isa, isb;   // line base = a*isa + b*isb   (elements of in type)
  int n_valid;          // input elements present, zero beyond (pruned input)

  // ---- tile store: out[b*osb + e*ose + a], e < keep -----------------------
  void* out;
  long long osb, ose;
  int keep;
  double scale;

  // slab mode (pass A only): row e of the tile belongs to rank e /
  // rows_per_peer and is stored into that rank's buffer (peer memory mapped
  // through CUDA IPC) at local row e % rows_per_peer; n_peers <= 1: off
  int n_peers, rows_per_peer;
  void* peer_out[8];
  int box_rows; // persistent pass A: element rows per tensor store
  // blocked exchange layout (3D slab mode), blk_T > 0: the buffer between
  // pass A and pass B is [b][tile = a/8][row][a%8] with blk_T tiles per
  // plane and blk_rows rows per tile, so that the rows a rank sends to one
  // peer form contiguous blocks (long NVLink bursts)
  int blk_T, blk_rows;

  // ---- contiguous line stores (re/im, lambda): line*pitch + e ------------
  void* out2;           // second line set (imag part / spare field)
  long long out_pitch;

  // ---- eigenvalue / sigma tables: row = am*rma + b*rmb, elem e (mirrored) -
  const double* tab;
  long long tab_pitch;
  long long rma, rmb;
  int tab_mirror_a;     // 0: am = a; else am = a <= n/2 ? a : n - a  (n given)
  int tab_mirror_e;     // 0: em = e; else em = e <= N/2 ? e : N - e

  // ---- noise ----------------------------------------------------------------
  int noise_kind;
  PhiloxKeys keys;         // round keys of the user seed
  unsigned long long realization;
  // batched 2D launches: line index b is the realization within the batch,
  // realization word of the counter = realization + b * real_b (0 otherwise)
  unsigned long long real_b;
  long long cma, cmb, cme; // philox counter = a*cma + b*cmb + (e mod N/2)*cme
  const double2* noise;    // injected
  long long ima, imb, ime; // injected index = a*ima + b*imb + e*ime
  int na_self, nb_self;    // self-conjugate test: 2a % na_self == 0 etc.
  int hermitian;           // 1: R2C field backend rule, 0: DFT backend rule
  int corrected;           // parity_mode corrected
  int plane_axis;          // which of a(0)/b(1) is k0 (for `corrected`)

  // ---- value transform (mutators.hh) ---------------------------------------
  int transform;
  double tparam;

  // ---- covariance evaluation ------------------------------------------------
  CovParams cov;
  int cov_axis_a, cov_axis_b; // torus axis indexed by a / b (-1: none)

  // ---- long lines (n = long_n1 * N > 1024, 2D): see pf_long in pf_api.cu -----
  int long_n1;             // sub-transforms per long line
  const double2* long_tw;  // [k1][j2] = exp(-2 pi i k1 j2 / n)

  // ---- eigenvalue census (matrix.hh:283-306) --------------------------------
  unsigned long long* census; // [small, zero, negative]
  unsigned long long* smallest_bits; // ordered-bits atomic min
  double threshold;
};

} // namespace pf
