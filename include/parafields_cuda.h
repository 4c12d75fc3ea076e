/* parafields_cuda.h -- C ABI of libparafields_cuda.so
 *
 * B200-native (sm_100a) replacement for the sampling hot path of
 * parafields-core: the covariance operator `parafields::Matrix`
 * (ext/parafields-core/include/parafields/matrix.hh:59-1663) together with the
 * matrix / field / RNG backends it drives
 * (.../backends/{dct,r2c,dft}matrixbackend.hh, {r2c,dft}fieldbackend.hh,
 * cpprngbackend.hh, fftwwrapper.hh).
 *
 * Conventions
 *  - every function returns 0 on success, a pf_status otherwise; nothing
 *    throws across the boundary; pf_last_error() returns the message of the
 *    last failure on the calling thread.  Messages reuse the reference's
 *    exception texts where one exists (e.g. "negative eigenvalues in
 *    covariance matrix", matrix.hh:327).
 *  - `double*` data arguments may point to host OR device memory (detected
 *    with cudaPointerGetAttributes); fields are dense, dimension 0 fastest:
 *    index = i0 + c0*(i1 + c1*i2)  (fieldtraits.hh:397-405).
 *  - a plan is not thread safe; calls are synchronous on return unless the
 *    data pointers are device pointers, in which case work is only enqueued
 *    on the plan's stream (pf_plan_set_stream / pf_plan_synchronize).
 *  - there is no CPU fallback: without a CUDA device every compute entry
 *    point fails with PF_ERR_CUDA.
 */
#ifndef PARAFIELDS_CUDA_H
#define PARAFIELDS_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pf_plan pf_plan;

typedef enum pf_status
{
  PF_OK = 0,
  PF_ERR_INVALID = 1,              /* bad argument / unsupported size      */
  PF_ERR_NEGATIVE_EIGENVALUES = 2, /* matrix.hh:322-328                    */
  PF_ERR_CUDA = 3,                 /* CUDA runtime failure / no device     */
  PF_ERR_NOMEM = 4
} pf_status;

/* stochastic.covariance, names as in matrix.hh:241-266 */
typedef enum pf_covariance
{
  PF_COV_EXPONENTIAL = 0,
  PF_COV_GAUSSIAN = 1,
  PF_COV_SPHERICAL = 2,
  PF_COV_SEPARABLE_EXPONENTIAL = 3,
  PF_COV_MATERN32 = 4,
  PF_COV_MATERN52 = 5,
  PF_COV_DAMPED_OSCILLATION = 6,
  PF_COV_GAMMA_EXPONENTIAL = 7,
  PF_COV_CAUCHY = 8,
  PF_COV_GENERALIZED_CAUCHY = 9,
  PF_COV_CUBIC = 10,
  PF_COV_WHITE_NOISE = 11,
  PF_COV_MATERN = 12, /* general nu (covariance.hh:527-581; the reference needs
                         GSL's K_nu for it, here a device function)       */
  PF_COV_CUSTOM = 13  /* custom-iso / custom-aniso (matrix.hh:236,270-279): the
                         covariance is a host functor of the caller; its values
                         on the extended domain arrive through
                         pf_set_covariance_table.  anisotropy NONE = the
                         symmetric (DCT matrix backend) storage of custom-iso,
                         GEOMETRIC = the half-spectrum storage of custom-aniso */
} pf_covariance;

/* stochastic.anisotropy, matrix.hh:496-508 */
typedef enum pf_anisotropy
{
  PF_ANISO_NONE = 0,
  PF_ANISO_AXIPARALLEL = 1,
  PF_ANISO_GEOMETRIC = 2
} pf_anisotropy;

/* randomField.transform, mutators.hh:137-155 */
typedef enum pf_transform_kind
{
  PF_TRANSFORM_NONE = 0,
  PF_TRANSFORM_LOG_NORMAL = 1,
  PF_TRANSFORM_FOLDED_NORMAL = 2,
  PF_TRANSFORM_SIGN = 3,
  PF_TRANSFORM_BOX_COX = 4
} pf_transform_kind;

typedef enum pf_op
{
  PF_OP_MATRIX = 0,      /* Matrix::operator*,   matrix.hh:151,1459-1517  */
  PF_OP_ROOT = 1,        /* multiplyRoot,        matrix.hh:176,1530-1589  */
  PF_OP_INVERSE_EXT = 2  /* multiplyInverseExtended, matrix.hh:1603-1662  */
} pf_op;

typedef enum pf_lambda_layout
{
  PF_LAMBDA_NATURAL = 0, /* field-backend index set, natural order:
                            (m0,n1,n2) half spectrum, k0 fastest; n0 in 1D */
  PF_LAMBDA_REFERENCE = 1 /* as stored by the reference's matrix backend
                            after finalize(): DCT compact / R2C half spectrum
                            (last two dims swapped in 3D) / DFT full        */
} pf_lambda_layout;

/* Configuration = the ini keys the hot path reads
 * (fieldtraits.hh:166-206, matrix.hh:283-284,496-508, covariance.hh,
 * mutators.hh:137-155).  Fill with pf_desc_default() first. */
typedef struct pf_desc
{
  int dim;                  /* 1..3 = number of grid.extensions entries      */
  unsigned cells[3];        /* grid.cells                                     */
  double extensions[3];     /* grid.extensions                                */
  unsigned embedding_factor; /* embedding.factor (default 2)                  */
  int covariance;           /* pf_covariance                                  */
  double variance;          /* stochastic.variance                            */
  double cov_param[2];      /* expGamma | cauchyAlpha, cauchyBeta | maternNu  */
  int anisotropy;           /* pf_anisotropy                                  */
  double corr_length[9];    /* stochastic.corrLength: 1 | dim | dim*dim
                               numbers, row major (NOT inverted)              */
  double threshold;         /* embedding.threshold (1e-14)                    */
  int approximate;          /* embedding.approximate                          */
  int periodic;             /* randomField.periodic (forces factor 1)         */
  int transform;            /* pf_transform_kind                              */
  double boxcox_lambda;     /* transform.boxCoxLambda                         */
  unsigned cg_iterations;   /* randomField.cgIterations (100)                 */
  int parity_mode;          /* 0 reference sampler, 1 corrected  */
  int rng_mode;             /* 0 Philox + fp32 Box-Muller, 1 fp64 Box-Muller  */
  int census_quirk;         /* reserved (accepted, ignored): the reference's
                               R2C census walks interleaved re/im of half the
                               entries; this library
                               always scans and clamps every real part, which
                               is identical for valid embeddings             */
  unsigned max_factor;      /* embedding-factor retry bound, 0 = no retry     */
  int device;               /* CUDA device ordinal, -1 = current              */
  /* slab decomposition over the GPUs of one node, one process per GPU
     (replaces the FFTW-MPI slab layout of fieldtraits.hh:271-282): the field
     is split over its last dimension, rank r owning cells[dim-1]/n_ranks
     layers; 0 or 1 = single GPU */
  int n_ranks;
  int rank;
  /* embedding.periodization (matrix.hh:530-572): 0 classical circulant
     embedding (mirroring, matrix.hh:624-648), 1 merge = 2^dim mirrored copies
     tapered by sigmoids (computeMatrixEntriesWithMerge, matrix.hh:662-787).
     fold / cofold need GSL quadrature and are not built. */
  int periodization;
  int sigmoid;              /* embedding.sigmoid: 0 smoothstep (default),
                               1 smooth (covariance.hh:145-275)              */
  int sigmoid_combine;      /* stochastic.sigmoidCombine: 0 radialPlus
                               (default), 1 radial, 2 tensor                  */
  double sigmoid_start;     /* embedding.sigmoidStart (1)                     */
  double sigmoid_end;       /* embedding.sigmoidEnd; < 0 = factor - 1         */
  unsigned merge_recursions; /* embedding.mergeRecursions; 99 = the
                               reference's per-covariance heuristic
                               (matrix.hh:693-719)                            */
} pf_desc;

/* eigenvalue census, matrix.hh:283-320 */
typedef struct pf_eig_stats
{
  unsigned long long small_count;    /* thr <= v < 1e-6   */
  unsigned long long zero_count;     /* |v| < thr         */
  unsigned long long negative_count; /* v <= -thr         */
  double smallest;
  unsigned factor_used;              /* after retries     */
} pf_eig_stats;

const char* pf_version(void);
const char* pf_last_error(void);
int pf_device_count(void);
/* name (<=256 bytes), SM count, memory bytes of a device */
int pf_device_info(int device, char* name, int* sm_count, size_t* mem_bytes);

void pf_desc_default(pf_desc* desc);

/* Matrix(traits) + update(), matrix.hh:106-142: sizes, tables, streams. */
int pf_plan_create(const pf_desc* desc, pf_plan** plan);
int pf_plan_destroy(pf_plan* plan);
/* cudaStream_t as void*; NULL = default stream */
int pf_plan_set_stream(pf_plan* plan, void* cuda_stream);
int pf_plan_synchronize(pf_plan* plan);
/* number of kernels this plan has launched so far */
unsigned long long pf_plan_launch_count(const pf_plan* plan);
/* per-pass device timing with CUDA events on the plan's stream (off by
 * default).  pf_plan_get_pass_times fills 16 slots indexed by pass kind
 * (0 noise+inverse axis pass "A", 2 inverse middle pass "B", 3 forward middle
 * pass, 4 c2r pass "C", 5/6 r2c pass, 7/8 eigenvalue pass, 9 fused spectral
 * pass, 1/10 the 1D variants): accumulated milliseconds and launch counts. */
int pf_plan_set_profiling(pf_plan* plan, int on);
int pf_plan_get_pass_times(pf_plan* plan, double* ms16,
                           unsigned long long* counts16);

/* sizes: Traits::{domainSize, extendedDomainSize} (fieldtraits.hh:284-296),
 * FieldBackend::localFieldSize() (r2cfieldbackend.hh:151) */
size_t pf_domain_size(const pf_plan* plan);
size_t pf_extended_size(const pf_plan* plan);
size_t pf_noise_entries(const pf_plan* plan);
/* cells of this rank's slab (= pf_domain_size without slab decomposition),
 * Traits::localDomainSize */
size_t pf_local_domain_size(const pf_plan* plan);

/* Slab mode wiring (one process per GPU).  Every rank exports PF_SLAB_HANDLE_BYTES
 * bytes describing its exchange buffer (a CUDA IPC memory handle), the host
 * program gathers the handles of all ranks in rank order (e.g. with
 * torch.distributed / MPI allgather) and passes them to pf_slab_connect,
 * which maps the peers' buffers.  After that pf_generate* scatters the
 * transposed tiles of the first inverse pass straight into the owning
 * rank's memory over NVLink (the FFTW-MPI transpose of
 * r2cfieldbackend.hh:250-274 fused into the producing kernel) and
 * synchronises the ranks with flags in peer memory; `out` receives the local
 * slab.  pf_apply and pf_solve work on the distributed field too (`in` / `out`
 * = this rank's layers; forward exchange by DMA, backward exchange by peer
 * stores, scalar products combined over the ranks in rank order).  All ranks
 * must issue the same sequence of pf_generate*, pf_apply and pf_solve calls,
 * and no rank may fall behind its peers by more than the wait bound
 * (PF_SLAB_TIMEOUT_S in the environment, default 120 s: eigenvalue setup,
 * allocation or I/O between calls count).  A wait that exceeds the bound is
 * FATAL for the plan: the call that synchronises next (pf_plan_synchronize,
 * pf_apply, pf_solve, pf_generate* with host buffers, every later call)
 * returns PF_ERR_CUDA and nothing computed after the timeout may be used. */
#define PF_SLAB_HANDLE_BYTES 64
int pf_slab_handle(pf_plan* plan, void* handle_out);
int pf_slab_connect(pf_plan* plan, const void* handles, int count);

/* Matrix::fillTransformedMatrix<Covariance>, matrix.hh:270-279 (custom-iso /
 * custom-aniso): the torus covariance evaluated by the caller, one value per
 * cell of the EXTENDED domain in natural order (index 0 fastest), cell j at
 * coordinate j_i * meshsize_i folded to (-L_i, L_i] as in matrix.hh:637-642;
 * host or device memory, copied.  The eigenvalues are recomputed from it. */
int pf_set_covariance_table(pf_plan* plan, const double* values, size_t count);

/* Matrix::fillTransformedMatrix, matrix.hh:232-331 (lazy in every other
 * entry point, like matrixBackend.valid()).  Returns
 * PF_ERR_NEGATIVE_EIGENVALUES exactly when the reference throws. */
int pf_compute_eigenvalues(pf_plan* plan, pf_eig_stats* stats);
size_t pf_eigenvalue_count(const pf_plan* plan, int layout);
int pf_get_eigenvalues(pf_plan* plan, double* out, size_t count, int layout);

/* Matrix::generateField, matrix.hh:343-438, with device Philox noise keyed by
 * (seed, realization): the result is a pure function of the plan's
 * configuration, seed and realization (no hidden state; launch geometry and
 * slab decomposition do not enter).  apply_transform fuses
 * ValueTransform::apply (randomfield.hh:600) into the last pass.
 *
 * Device noise, specification (restated on the CPU in oracle/pf_oracle.cpp and
 * checked by tests/test_gpu_noise.py): one Philox4x32 block per pair of
 * frequency entries, counter = (c, realization), key = seed, with
 * c = k0 + m0 (e mod L/2) [+ m0 (L/2) k1 in 3D], e = index along the LAST
 * frequency axis of length L; words (0,1) serve entry e < L/2, words (2,3)
 * entry e + L/2; Box-Muller: u = (w0 + 1/2) 2^-32, radius sqrt(-2 ln u),
 * angle pi (w1 2^-31 - 1), (r1, r2) = radius (cos, sin).  rng_mode 1: 10
 * rounds, fp64; rng_mode 0: 7 rounds, the same formula in fp32.
 *
 * 1D (DFT field backend, matrix.hh:349,428-437: one complex transform yields
 * two fields): realization r is the real part (r even) or the imaginary part
 * (r odd) of transform r / 2.  The reference's protocol "every second
 * generate() returns the spare field and ignores its seed" is kept by the
 * host plugin (cudamatrix.hh), not by hidden state in the plan. */
int pf_generate(pf_plan* plan, uint64_t seed, uint64_t realization,
                double* out, int apply_transform);
/* `count` realizations first..first+count-1 into out[count][domain] */
int pf_generate_batch(pf_plan* plan, uint64_t seed, uint64_t first,
                      size_t count, double* out, int apply_transform);
/* same arithmetic with caller-supplied white noise: 2*pf_noise_entries()
 * doubles (r1,r2 interleaved) in the reference's stream order
 * (SURVEY Appendix A.5).  spare may be NULL (1D: imaginary part). */
int pf_generate_from_noise(pf_plan* plan, const double* noise, double* out,
                           double* spare);
/* Matrix::generateUncorrelatedField, matrix.hh:450-462 (device Philox) */
int pf_generate_uncorrelated(pf_plan* plan, uint64_t seed, double* out);
/* Matrix::setVarianceAsField, matrix.hh:473-479 */
int pf_set_variance_as_field(pf_plan* plan, double* out);

/* multiplyExtended / multiplyRootExtended / multiplyInverseExtended */
int pf_apply(pf_plan* plan, int op, const double* in, double* out);
/* Matrix::multiplyInverse + innerCG, matrix.hh:198-215,1331-1448 */
int pf_solve(pf_plan* plan, const double* in, double* out,
             unsigned* iterations);

/* ValueTransform::apply over n values (mutators.hh:161) */
int pf_transform(pf_plan* plan, const double* in, double* out, size_t n);

/* StochasticPart vector space operations (stochastic.hh:320-467,762-787) on
 * device or host arrays of n doubles */
int pf_vec_axpy(pf_plan* plan, size_t n, double alpha, const double* x,
                double* y);                       /* y += alpha x */
int pf_vec_scale(pf_plan* plan, size_t n, double alpha, double* x);
int pf_vec_dot(pf_plan* plan, size_t n, const double* x, const double* y,
               double* result);
int pf_vec_norm1(pf_plan* plan, size_t n, const double* x, double* result);
int pf_vec_norminf(pf_plan* plan, size_t n, const double* x, double* result);
/* StochasticPart::operator== (stochastic.hh:438-452): *same = 1 when every
 * entry of x equals the entry of y (NaN differs from everything) */
int pf_vec_equal(pf_plan* plan, size_t n, const double* x, const double* y,
                 int* same);
/* StochasticPart::zero (stochastic.hh:581-587) and constant fields */
int pf_vec_fill(pf_plan* plan, size_t n, double value, double* x);
/* pf_vec_dot without any allocation.  No counterpart in the reference: added
 * for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  The per-block partial sums live in a file-scope
 * `static __device__` array of 600 doubles (4800 bytes, held by every GPU that
 * loads the library until exit) instead of plan memory; the call performs no
 * cudaMalloc / cudaFree.  x and y must be DEVICE arrays of n doubles (host
 * arrays would need a staging buffer: PF_ERR_INVALID), n <= PF_STATIC_DOT_MAX
 * (larger n: PF_ERR_INVALID, use pf_vec_dot), single-GPU plans only.  result
 * may be a device pointer (written asynchronously on the plan's stream) or a
 * host pointer (the call synchronises).  Same grid and summation order as
 * pf_vec_dot: the value is bit-identical.  All calls on one GPU share the
 * array (it keeps the last call's partial sums; nothing has to start at zero
 * because every block overwrites its entry before it is read): calls that
 * overlap on the GPU -- different plans or streams -- race. */
#define PF_STATIC_DOT_MAX ((size_t)1 << 28)
int pf_vec_dot_static(pf_plan* plan, size_t n, const double* x,
                      const double* y, double* result);

/* Resolution transfer, localisation and point evaluation of a field
 * (SURVEY 8(f) rows 1 and 3).  Arrays may be device or host memory (host
 * arrays are staged through the device, nothing is computed on the CPU);
 * cells[] has dim entries, index 0 fastest.
 *
 * StochasticPart::refine, stochastic.hh:597-670: cells = OLD local cells,
 * out holds 2^dim * prod(cells) doubles (16-byte aligned).  dim 2 or 3; 1D
 * fails with the reference's "not implemented". */
int pf_field_refine(pf_plan* plan, int dim, const unsigned* cells,
                    const double* in, double* out);
/* StochasticPart::coarsen, stochastic.hh:679-755: cells = OLD (fine) local
 * cells, all even ("cannot coarsen odd number of cells",
 * fieldtraits.hh:336-339); out holds prod(cells)/2^dim doubles; sums in the
 * reference's order, bit-identical. */
int pf_field_coarsen(pf_plan* plan, int dim, const unsigned* cells,
                     const double* in, double* out);
/* StochasticPart::localize, stochastic.hh:798-817 (cell centres by
 * indicesToCoords, fieldtraits.hh:461-469), in place.  offset (first global
 * index of the local block) and global_cells may be NULL on one rank. */
int pf_field_localize(pf_plan* plan, int dim, const unsigned* cells,
                      const unsigned* offset, const unsigned* global_cells,
                      const double* extensions, const double* center,
                      double radius, double* data);
/* StochasticPart::evaluate (stochastic.hh:480-574, coordsToIndices
 * fieldtraits.hh:440-448) for npoints locations (npoints x dim doubles) on
 * one rank, followed by RandomField::evaluate's value transform
 * (randomfield.hh:585-595) when apply_transform != 0.  A location on the
 * upper face of exactly one dimension yields the reference's zero-filled
 * overlap value; locations outside the closed box fail with PF_ERR_INVALID
 * (undefined behaviour in the reference). */
int pf_field_evaluate(pf_plan* plan, int dim, const unsigned* cells,
                      const double* extensions, const double* field,
                      size_t npoints, const double* locations,
                      int apply_transform, double* out);

/* One TrendComponent (trend.hh:601-634): coeff = meanVector + shiftVector.
 * mean: {m}; slope: {s_0..s_dim-1}, value sum s_i (x_i - ext_i/2);
 * disk: {c_0..c_dim-1, radius, height}, value exp(-|x-c|^2 / radius^2 * height);
 * block: {c_0..c_dim-1, w_0..w_dim-1, value} inside |x_i - c_i| <= w_i/2.
 * The image component (libpng) is not supported. */
typedef enum pf_trend_kind
{
  PF_TREND_MEAN = 0,
  PF_TREND_SLOPE = 1,
  PF_TREND_DISK = 2,
  PF_TREND_BLOCK = 3
} pf_trend_kind;
typedef struct pf_trend_component
{
  int kind;        /* pf_trend_kind */
  double coeff[7]; /* up to 2*dim+1 coefficients */
} pf_trend_component;
#define PF_MAX_TREND_COMPONENTS 16

/* RandomField::evaluate (randomfield.hh:585-595) for npoints locations:
 * out = T(stochastic + trend), the trend summed over ntrend components in
 * order (TrendPart::evaluate, trend.hh:1491-1505).  ntrend = 0 equals
 * pf_field_evaluate. */
int pf_field_evaluate_trend(pf_plan* plan, int dim, const unsigned* cells,
                            const double* extensions, const double* field,
                            size_t npoints, const double* locations,
                            int ntrend, const pf_trend_component* trend,
                            int apply_transform, double* out);
/* RandomField::bulkEvaluate (randomfield.hh:597-631): the same at the centre
 * of every cell, (i + 0.5) * (extensions / cells), in one fused sweep
 * (field read once, no location array); out holds prod(cells) doubles and
 * may alias field. */
int pf_field_bulk_evaluate(pf_plan* plan, int dim, const unsigned* cells,
                           const double* extensions, const double* field,
                           int ntrend, const pf_trend_component* trend,
                           int apply_transform, double* out);

/* Slab <-> block redistribution with overlap on a connected multi-GPU plan
 * (StochasticPart::dataToEval / evalToData / exchangeOverlap,
 * stochastic.hh:830-1127).  proc_per_dim[dim] = the processor grid of the
 * reference's load balancer (DefaultLoadBalance, fieldtraits.hh:483-562;
 * product = n_ranks, cells divisible); rank r owns the block at
 * localEvalOffset of stochastic.hh:244-257.  slab = this rank's layers of the
 * last dimension (pf_local_domain_size doubles, what pf_generate returns);
 * eval = its block in natural order (same count); overlap (may be NULL) =
 * pf_field_overlap_size doubles: for every dimension i the lower halo (the
 * lower neighbour's last layer) followed by the upper halo, neighbours
 * periodic, face index as in stochastic.hh:1025-1040.  Collective: all ranks
 * call it in the same order; the data moves as loads from the peers' exchange
 * buffers over NVLink (no host round trip).  Arrays may be host or device. */
size_t pf_field_overlap_size(const pf_plan* plan, const int* proc_per_dim);
int pf_field_slab_to_blocks(pf_plan* plan, const int* proc_per_dim,
                            const double* slab, double* eval, double* overlap);
int pf_field_blocks_to_slab(pf_plan* plan, const int* proc_per_dim,
                            const double* eval, double* slab);

/* device memory helpers so that a host-only caller (C++, cgo, ctypes) can
 * keep fields resident in HBM */
int pf_device_alloc(pf_plan* plan, size_t bytes, void** ptr);
/* plan may be NULL here: a buffer may outlive the plan it was made with */
int pf_device_free(pf_plan* plan, void* ptr);
int pf_memcpy(pf_plan* plan, void* dst, const void* src, size_t bytes);

#ifdef __cplusplus
}
#endif
#endif /* PARAFIELDS_CUDA_H */
