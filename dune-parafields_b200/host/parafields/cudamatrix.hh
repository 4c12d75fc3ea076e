// parafields/cudamatrix.hh -- the B200 covariance operator as a drop-in for
// parafields::Matrix (reference: include/parafields/matrix.hh:59-1663).
//
//   using Field = parafields::RandomField<GridTraits,
//                                         parafields::CudaMatrix,   // IsoMatrix
//                                         parafields::CudaMatrix>;  // AnisoMatrix
//
// CudaMatrix<Traits> implements the members RandomField uses from Matrix
// (matrix.hh:106,130,151,176,198,232,343,450,473) on top of the C ABI of
// libparafields_cuda.so (include/parafields_cuda.h).  It works both with the
// reference's own randomfield.hh / fieldtraits.hh / stochastic.hh (it is named
// a friend there through the IsoMatrix / AnisoMatrix template parameters,
// fieldtraits.hh:102-109, stochastic.hh:37-38) and with the stand-alone
// headers shipped next to this file.
//
// Extra ini keys (all optional, section [cuda]):
//   cuda.noise      philox (default): device Philox4x32 keyed by the seed
//                   reference: the reference's CppRNGBackend stream
//                     (std::default_random_engine + std::normal_distribution,
//                     cpprngbackend.hh:49-58) drawn on the host and injected,
//                     so generate(seed) reproduces the reference's field
//   cuda.rng        fp64 (default: 10 Philox rounds, fp64 Box-Muller -- normals
//                   of the reference's precision) | fast (7 rounds, fp32
//                   Box-Muller on the special-function units; what bench.py
//                   times by default)
//   cuda.sampler    reference (default) | corrected   
//   cuda.maxFactor  embedding-factor retry bound (0 = off, the reference throws)
//   cuda.device     CUDA device ordinal (-1 = current)
#ifndef PARAFIELDS_CUDAMATRIX_HH
#define PARAFIELDS_CUDAMATRIX_HH

#include <array>
#include <cstdint>
#include <iostream>
#include <memory>
#include <random>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include <parafields_cuda.h>

// device-resident StochasticPart for RandomField<GridTraits, CudaMatrix,
// CudaMatrix> (and the plan registry both sides use)
#include <parafields/cudastochastic.hh>

namespace parafields {

template<typename Traits>
class StochasticPart;

namespace cuda_detail {
template<typename Part, typename = void>
struct IsDeviceResident : std::false_type
{};
template<typename Part>
struct IsDeviceResident<Part, std::void_t<typename Part::DeviceResidentTag>>
  : std::true_type
{};
} // namespace cuda_detail

template<typename Traits>
class CudaMatrix
{
public:
  // named as friends by the traits class; no per-element backends exist here
  struct MatrixBackendType {};
  struct FieldBackendType {};
  struct RNGBackendType {};

  using StochasticPartType = StochasticPart<Traits>;

private:
  using RF = typename Traits::RF;
  enum { dim = Traits::dim };
  static_assert(std::is_same<RF, double>::value,
                "the CUDA path computes in fp64 (RangeField must be double)");
  // dim 4 compiles (the reference CLI instantiates it) and reports at run time

  const std::shared_ptr<Traits> traits;
  mutable pf_plan* plan = nullptr;
  mutable bool filled = false;
  bool referenceNoise = false;

  static void check(int rc)
  {
    if (rc != 0)
      throw std::runtime_error{ pf_last_error() };
  }

  static int covarianceId(const std::string& name)
  {
    static const std::pair<const char*, int> table[] = {
      { "exponential", PF_COV_EXPONENTIAL },
      { "gaussian", PF_COV_GAUSSIAN },
      { "spherical", PF_COV_SPHERICAL },
      { "separableExponential", PF_COV_SEPARABLE_EXPONENTIAL },
      { "matern32", PF_COV_MATERN32 },
      { "matern52", PF_COV_MATERN52 },
      { "matern", PF_COV_MATERN },
      { "dampedOscillation", PF_COV_DAMPED_OSCILLATION },
      { "gammaExponential", PF_COV_GAMMA_EXPONENTIAL },
      { "cauchy", PF_COV_CAUCHY },
      { "generalizedCauchy", PF_COV_GENERALIZED_CAUCHY },
      { "cubic", PF_COV_CUBIC },
      { "whiteNoise", PF_COV_WHITE_NOISE }
    };
    for (const auto& e : table)
      if (name == e.first)
        return e.second;
    // host functors of the caller: evaluated once on the host into a table
    // (fillTransformedMatrix<Covariance>, pf_set_covariance_table)
    if (name == "custom-iso" || name == "custom-aniso")
      return PF_COV_CUSTOM;
    // same text as matrix.hh:267-268
    throw std::runtime_error{ "covariance structure " + name + " not known" };
  }

public:
  explicit CudaMatrix(const std::shared_ptr<Traits>& traits_)
    : traits(traits_)
  {
    update();
  }

  CudaMatrix(const CudaMatrix&) = delete;
  CudaMatrix& operator=(const CudaMatrix&) = delete;

  ~CudaMatrix()
  {
    if (plan) {
      cuda_detail::PlanRegistry::withdraw(traits.get(), plan);
      cuda_detail::BufferPool::drain(traits.get(), plan);
      pf_plan_destroy(plan);
    }
  }

  // Matrix::update, matrix.hh:130-142: (re)read sizes after construction or
  // refinement; the eigenvalues are recomputed lazily
  void update()
  {
    if (plan) {
      cuda_detail::PlanRegistry::withdraw(traits.get(), plan);
      cuda_detail::BufferPool::drain(traits.get(), plan);
      pf_plan_destroy(plan);
      plan = nullptr;
    }
    filled = false;
    spare.clear();
    spareSeedValid = false;
    if (dim < 1 || dim > 3)
      throw std::runtime_error{ "CudaMatrix supports dim 1, 2, 3" };
    const auto& config = (*traits).config;
    pf_desc d;
    pf_desc_default(&d);
    d.dim = dim;
    for (unsigned int i = 0; i < (dim <= 3 ? dim : 3); i++) {
      d.cells[i] = (*traits).cells[i];
      d.extensions[i] = (*traits).extensions[i];
    }
    d.embedding_factor = (*traits).embeddingFactor;
    d.periodic = (*traits).periodic ? 1 : 0;
    d.approximate = (*traits).approximate ? 1 : 0;
    d.variance = (*traits).variance;
    d.covariance = covarianceId((*traits).covariance);
    d.cg_iterations = (*traits).cgIterations;
    d.threshold = config.template get<RF>("embedding.threshold", 1e-14);

    // stochastic.anisotropy / corrLength, matrix.hh:496-508 and
    // covariance.hh:32,61,94 (arity follows the anisotropy)
    const std::string anisotropy =
      config.template get<std::string>("stochastic.anisotropy", "none");
    if (d.covariance == PF_COV_CUSTOM) {
      // always evaluated with ScaledIdentityMatrix (matrix.hh:277-278);
      // custom-iso keeps the symmetric storage, custom-aniso the half spectrum
      d.anisotropy = (*traits).covariance == "custom-aniso" ? PF_ANISO_GEOMETRIC
                                                            : PF_ANISO_NONE;
      d.corr_length[0] = config.template get<RF>("stochastic.corrLength");
      if (d.anisotropy == PF_ANISO_GEOMETRIC)
        for (unsigned int i = 0; i < dim; i++)
          for (unsigned int j = 0; j < dim; j++)
            d.corr_length[i * dim + j] = i == j ? d.corr_length[0] : 0.;
    } else if (anisotropy == "none") {
      d.anisotropy = PF_ANISO_NONE;
      d.corr_length[0] = config.template get<RF>("stochastic.corrLength");
    } else if (anisotropy == "axiparallel") {
      d.anisotropy = PF_ANISO_AXIPARALLEL;
      const auto c =
        config.template get<std::array<RF, dim>>("stochastic.corrLength");
      for (unsigned int i = 0; i < dim; i++)
        d.corr_length[i] = c[i];
    } else if (anisotropy == "geometric") {
      d.anisotropy = PF_ANISO_GEOMETRIC;
      const auto c =
        config.template get<std::array<RF, dim * dim>>("stochastic.corrLength");
      for (unsigned int i = 0; i < dim * dim; i++)
        d.corr_length[i] = c[i];
    } else
      throw std::runtime_error{
        "stochastic.anisotropy must be \"none\", \"axiparallel\" or \"geometric\""
      };

    if (d.covariance == PF_COV_GAMMA_EXPONENTIAL)
      d.cov_param[0] = config.template get<RF>("stochastic.expGamma");
    // covariance.hh:541 (the reference needs GSL for this functor; the CUDA
    // path evaluates K_nu on the device)
    if (d.covariance == PF_COV_MATERN)
      d.cov_param[0] = config.template get<double>("stochastic.maternNu");
    if (d.covariance == PF_COV_GENERALIZED_CAUCHY) {
      d.cov_param[0] = config.template get<RF>("stochastic.cauchyAlpha");
      d.cov_param[1] = config.template get<RF>("stochastic.cauchyBeta");
    }

    // randomField.transform is applied at evaluation time by the host
    // (ValueTransform, randomfield.hh:600); the plan only needs it for the
    // fused variant used by the batch API
    const std::string transform =
      config.template get<std::string>("randomField.transform", "none");
    if (transform == "logNormal")
      d.transform = PF_TRANSFORM_LOG_NORMAL;
    else if (transform == "foldedNormal")
      d.transform = PF_TRANSFORM_FOLDED_NORMAL;
    else if (transform == "sign")
      d.transform = PF_TRANSFORM_SIGN;
    else if (transform == "boxCox") {
      d.transform = PF_TRANSFORM_BOX_COX;
      d.boxcox_lambda = config.template get<RF>("transform.boxCoxLambda");
    }

    const std::string noise =
      config.template get<std::string>("cuda.noise", "philox");
    if (noise != "philox" && noise != "reference")
      throw std::runtime_error{ "cuda.noise must be \"philox\" or \"reference\"" };
    referenceNoise = (noise == "reference");
    d.rng_mode =
      config.template get<std::string>("cuda.rng", "fp64") == "fast" ? 0 : 1;
    d.parity_mode =
      config.template get<std::string>("cuda.sampler", "reference") ==
          "corrected"
        ? 1
        : 0;
    d.max_factor = config.template get<unsigned int>("cuda.maxFactor", 0);
    d.device = config.template get<int>("cuda.device", -1);

    // embedding.periodization / embedding.sigmoid / stochastic.sigmoidCombine
    // (matrix.hh:530-590, 669-692); fold and cofold need GSL quadrature
    const std::string periodization =
      config.template get<std::string>("embedding.periodization", "classical");
    if (periodization == "classical")
      d.periodization = 0;
    else if (periodization == "merge")
      d.periodization = 1;
    else if (periodization == "fold" || periodization == "cofold")
      throw std::runtime_error{
        "CudaMatrix: embedding.periodization fold / cofold is not built "
        "(GSL quadrature); use classical or merge"
      };
    else
      throw std::runtime_error{
        "stochastic.periodization must be \"classical\", \"merge\", \"fold\", "
        "or \"cofold\""
      }; // matrix.hh:588-591
    if (d.periodization == 1) {
      const std::string sigmoid =
        config.template get<std::string>("embedding.sigmoid", "smoothstep");
      if (sigmoid == "smooth")
        d.sigmoid = 1;
      else if (sigmoid == "smoothstep")
        d.sigmoid = 0;
      else
        throw std::runtime_error{
          "embedding.sigmoid must be \"smooth\" or \"smoothstep\""
        };
      const std::string combine = config.template get<std::string>(
        "stochastic.sigmoidCombine", "radialPlus");
      if (combine == "radialPlus")
        d.sigmoid_combine = 0;
      else if (combine == "radial")
        d.sigmoid_combine = 1;
      else if (combine == "tensor")
        d.sigmoid_combine = 2;
      else
        throw std::runtime_error{
          "sigmoidCombine must be tensor, radial or radialPlus"
        };
      d.sigmoid_start = config.template get<RF>("embedding.sigmoidStart", 1.);
      d.sigmoid_end = config.template get<RF>(
        "embedding.sigmoidEnd", (*traits).embeddingFactor - 1.);
      d.merge_recursions =
        config.template get<unsigned int>("embedding.mergeRecursions", 99);
    }
    // the reference's R2C-matrix census quirk is
    // kept so that the clamp pattern matches the reference
    d.census_quirk = (d.anisotropy == PF_ANISO_GEOMETRIC) ? 1 : 0;

    if ((*traits).commSize != 1)
      throw std::runtime_error{
        "CudaMatrix: one process drives the GPUs of a plan; MPI ranks > 1 "
        "are not supported (use the multi-GPU plan instead)"
      };

    check(pf_plan_create(&d, &plan));
    // the fields of these traits find their plan here (cudastochastic.hh)
    cuda_detail::PlanRegistry::announce(traits.get(), plan);
  }

  // Matrix::operator*, matrix.hh:151-160
  StochasticPartType operator*(const StochasticPartType& input) const
  {
    return apply(PF_OP_MATRIX, input);
  }

  // Matrix::multiplyRoot, matrix.hh:176-185
  StochasticPartType multiplyRoot(const StochasticPartType& input) const
  {
    return apply(PF_OP_ROOT, input);
  }

  // Matrix::multiplyInverse, matrix.hh:198-215 (PCG of matrix.hh:1331-1448)
  StochasticPartType multiplyInverse(const StochasticPartType& input) const
  {
    ensureFilled();
    if (isEmpty(input))
      return input;
    StochasticPartType output = likeInput(input);
    unsigned int iterations = 0;
    check(pf_solve(plan, readPointer(input),
                   writePointer(output, fieldSize(input)), &iterations));
    if ((*traits).verbose && (*traits).rank == 0)
      std::cout << iterations << " iterations" << std::endl;
    return output;
  }

  // Matrix::fillTransformedMatrix, matrix.hh:232-331
  template<typename Covariance = void>
  void fillTransformedMatrix() const
  {
    if constexpr (!std::is_same<Covariance, void>::value) {
      // matrix.hh:270-279: only with custom-iso / custom-aniso
      if ((*traits).covariance != "custom-iso" &&
          (*traits).covariance != "custom-aniso")
        throw std::runtime_error{
          "you can't fill the matrix explicitly if a default "
          "covariance was chosen"
        };
      uploadCustomCovariance<Covariance>();
    }
    pf_eig_stats stats{};
    const int rc = pf_compute_eigenvalues(plan, &stats);
    if ((*traits).verbose && (*traits).rank == 0)
      std::cout << stats.small_count << " small, " << stats.zero_count
                << " approx. zero and " << stats.negative_count
                << " large negative eigenvalues in covariance matrix, smallest "
                << stats.smallest << std::endl;
    if (rc == PF_ERR_NEGATIVE_EIGENVALUES && (*traits).rank == 0)
      std::cerr << "negative eigenvalues in covariance matrix, "
                << "consider increasing embeddingFactor, or alternatively "
                << "allow generation of approximate samples" << std::endl;
    check(rc);
    filled = true;
  }

  // Matrix::generateField, matrix.hh:343-438
  void generateField(unsigned int seed, StochasticPartType& stochasticPart) const
  {
    ensureFilled();
    const std::size_t count = (*traits).localDomainSize;
    seed += (*traits).rank; // different seed for each processor (:352)
    if (referenceNoise) {
      if (dim == 1 && spare.size() == count) {
        // DFT field backend: every second call returns the imaginary part
        // and ignores its seed (matrix.hh:349,433-437)
        assignHost(stochasticPart, spare);
        spare.clear();
      } else {
        const size_t entries = pf_noise_entries(plan);
        std::vector<double> noise(2 * entries);
        std::default_random_engine generator(seed);
        std::normal_distribution<RF> normalDist(0., 1.);
        for (auto& v : noise)
          v = normalDist(generator);
        if (dim == 1)
          spare.resize(count);
        check(pf_generate_from_noise(plan, noise.data(),
                                     writePointer(stochasticPart, count),
                                     dim == 1 ? spare.data() : nullptr));
      }
    } else if (dim == 1) {
      // DFT field backend (matrix.hh:349,428-437): the first call returns the
      // real part of one complex transform, the next call its imaginary part
      // WHATEVER seed it is given.  The C ABI is stateless (realization 2t /
      // 2t+1 = real / imaginary part of transform t), so the protocol is kept
      // here.
      if (spareSeedValid) {
        check(pf_generate(plan, spareSeed, 1,
                          writePointer(stochasticPart, count), 0));
        spareSeedValid = false;
      } else {
        check(pf_generate(plan, seed, 0, writePointer(stochasticPart, count), 0));
        spareSeed = seed;
        spareSeedValid = true;
      }
    } else
      check(pf_generate(plan, seed, 0, writePointer(stochasticPart, count), 0));
  }

  // Matrix::generateUncorrelatedField, matrix.hh:450-462
  void generateUncorrelatedField(unsigned int seed,
                                 StochasticPartType& stochasticPart) const
  {
    const std::size_t count = (*traits).localDomainSize;
    seed += (*traits).rank;
    if (referenceNoise) {
      std::vector<RF> drawn(count);
      std::default_random_engine generator(seed);
      std::normal_distribution<RF> normalDist(0., 1.);
      for (auto& v : drawn)
        v = normalDist(generator);
      assignHost(stochasticPart, drawn);
    } else
      check(pf_generate_uncorrelated(plan, seed,
                                     writePointer(stochasticPart, count)));
  }

  // Matrix::setVarianceAsField, matrix.hh:473-479
  void setVarianceAsField(StochasticPartType& stochasticPart) const
  {
    check(pf_set_variance_as_field(
      plan, writePointer(stochasticPart, (*traits).localDomainSize)));
  }

  // access for callers that want to keep data on the device
  pf_plan* cudaPlan() const { return plan; }

private:
  mutable std::vector<RF> spare;
  mutable unsigned int spareSeed = 0;
  mutable bool spareSeedValid = false;

  void ensureFilled() const
  {
    if (!filled)
      fillTransformedMatrix();
  }

  // Matrix::computeMatrixEntriesWithMirroring<Covariance, ScaledIdentityMatrix>
  // (matrix.hh:624-648) on the host over the WHOLE extended domain: the
  // caller's functor cannot run on the device, its values can
  template<typename Covariance>
  void uploadCustomCovariance() const
  {
    const auto& config = (*traits).config;
    const Covariance covariance(config);
    // ScaledIdentityMatrix (covariance.hh:31-41): value = 1 / corrLength
    const RF scale = 1. / config.template get<RF>("stochastic.corrLength");
    const RF variance = (*traits).variance;
    const unsigned int factor = (*traits).embeddingFactor;
    std::array<unsigned int, dim> n;
    size_t total = 1;
    for (unsigned int i = 0; i < dim; i++) {
      n[i] = factor * (*traits).cells[i];
      total *= n[i];
    }
    std::vector<double> table(total);
    std::array<RF, dim> coord, transCoord;
    std::array<unsigned int, dim> idx{};
    for (size_t index = 0; index < total; index++) {
      size_t r = index;
      for (unsigned int i = 0; i < dim; i++) {
        idx[i] = r % n[i];
        r /= n[i];
        const RF ext = (*traits).extensions[i];
        coord[i] = idx[i] * (ext / (*traits).cells[i]);
        if (coord[i] > 0.5 * ext * factor)
          coord[i] -= ext * factor;
        transCoord[i] = scale * coord[i];
      }
      table[index] = covariance(variance, transCoord);
    }
    check(pf_set_covariance_table(plan, table.data(), table.size()));
  }

  StochasticPartType apply(int op, const StochasticPartType& input) const
  {
    ensureFilled();
    if (isEmpty(input))
      return input; // a field that was never filled is the zero field
    StochasticPartType output = likeInput(input);
    check(pf_apply(plan, op, readPointer(input),
                   writePointer(output, fieldSize(input))));
    return output;
  }

  // Where the values of a StochasticPart are.  The reference's class
  // (stochastic.hh:54, this class is its friend through the traits) keeps
  // them in a host vector that pf_* stages through the device; the
  // specialisation of cudastochastic.hh hands out device pointers, and
  // nothing is copied.
  static constexpr bool deviceResident()
  {
    return cuda_detail::IsDeviceResident<StochasticPartType>::value;
  }
  static bool isEmpty(const StochasticPartType& part)
  {
    if constexpr (deviceResident())
      return part.values.empty();
    else
      return part.dataVector.empty();
  }
  static std::size_t fieldSize(const StochasticPartType& part)
  {
    if constexpr (deviceResident())
      return part.values.size();
    else
      return part.dataVector.size();
  }
  static const double* readPointer(const StochasticPartType& part)
  {
    if constexpr (deviceResident())
      return part.values.deviceRead();
    else
      return part.dataVector.data();
  }
  // the whole field is about to be overwritten
  static double* writePointer(StochasticPartType& part, std::size_t count)
  {
    if constexpr (deviceResident()) {
      part.values.resize(count);
      return part.values.deviceOverwrite();
    } else {
      part.dataVector.resize(count);
      part.evalValid = false;
      return part.dataVector.data();
    }
  }
  static void assignHost(StochasticPartType& part, const std::vector<RF>& from)
  {
    if constexpr (deviceResident()) {
      part.values.resize(from.size());
      part.values.hostOverwrite() = from;
    } else {
      part.dataVector = from;
      part.evalValid = false;
    }
  }
  // a field with the geometry of `input` whose values will be overwritten:
  // the device-resident class skips the copy of the values
  static StochasticPartType likeInput(const StochasticPartType& input)
  {
    if constexpr (deviceResident())
      return StochasticPartType(input, typename StochasticPartType::GeometryOnly{});
    else
      return StochasticPartType(input);
  }
};

} // namespace parafields

#endif // PARAFIELDS_CUDAMATRIX_HH
