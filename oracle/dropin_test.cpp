// oracle/dropin_test.cpp -- drop-in proof.  TEST INFRASTRUCTURE.
//
// The UNMODIFIED reference parafields::RandomField (from /root/reference) is
// instantiated twice in one process:
//   RandomField<GridTraits>                           reference CPU matrices
//                                                     (FFTW via oracle/shim)
//   RandomField<GridTraits, CudaMatrix, CudaMatrix>   this repository's plugin
// and both are driven through the reference's own public API with the same
// configuration and seed.  With cuda.noise = reference the plugin consumes the
// reference's RNG stream, so every result must agree to 1e-10 (fp64).
// Built by oracle/Makefile into oracle/_ref/dropin_test (needs the reference
// checkout at build time, a B200 at run time; run by tests/test_dropin.py).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

#include <dune/common/fvector.hh>
#include <dune/common/parametertree.hh>
#include <dune/common/parametertreeparser.hh>

#include <parafields/cudamatrix.hh>
#include <parafields/randomfield.hh>

template<typename DF, typename RF, unsigned int dimension>
class GridTraits
{
public:
  enum
  {
    dim = dimension
  };
  using RangeField = RF;
  using Scalar = Dune::FieldVector<RF, 1>;
  using DomainField = DF;
  using Domain = Dune::FieldVector<DF, dim>;
};

static int failures = 0;

// compares the raw cell values (StochasticPart::dataVector; this test file is
// compiled with -fno-access-control).  bulkEvaluate is not used here because
// the reference's spare-field branch (matrix.hh:433-437) leaves evalValid set,
// so it would return the previous field.
// the reference's host class exposes dataVector; the device-resident class
// of cudastochastic.hh hands out its host mirror
template<class Part>
auto
cellValues(const Part& part, int) -> decltype(part.hostData())
{
  return part.hostData();
}
template<class Part>
const std::vector<double>&
cellValues(const Part& part, long)
{
  return part.dataVector;
}

template<class A, class B>
double
relDiff(const A& a, const B& b)
{
  const auto& va = cellValues(a.stochasticPart, 0);
  const auto& vb = cellValues(b.stochasticPart, 0);
  double num = 0., den = 0.;
  for (size_t i = 0; i < va.size() && i < vb.size(); i++) {
    num = std::max(num, std::abs(va[i] - vb[i]));
    den = std::max(den, std::abs(va[i]));
  }
  return va.size() == vb.size() && den > 0. ? num / den : 1e300;
}

static void
report(const std::string& what, double err, double tol)
{
  const bool ok = err < tol;
  if (!ok)
    failures++;
  std::printf("  %-34s rel err %.3e  %s\n", what.c_str(), err, ok ? "ok" : "FAIL");
}

template<unsigned int dim>
void
run(const std::string& name, const std::string& ini)
{
  using GT = GridTraits<double, double, dim>;
  using RefField = parafields::RandomField<GT>;
  using CudaField =
    parafields::RandomField<GT, parafields::CudaMatrix, parafields::CudaMatrix>;
  std::printf("%s\n", name.c_str());
  Dune::ParameterTree config;
  std::istringstream in(ini);
  Dune::ParameterTreeParser::readINITree(in, config);
  Dune::ParameterTree cconfig = config;
  cconfig["cuda.noise"] = "reference";

  RefField ref(config);
  CudaField gpu(cconfig);
  ref.generate(20240611u);
  gpu.generate(20240611u);
  report("generate(seed)", relDiff(ref, gpu), 1e-10);
  if (dim == 1) { // spare field of the DFT backend
    ref.generate(1u);
    gpu.generate(1u);
    report("generate (spare field)", relDiff(ref, gpu), 1e-10);
  }
  {
    RefField a(ref);
    CudaField b(gpu);
    a.timesMatrix();
    b.timesMatrix();
    report("timesMatrix", relDiff(a, b), 1e-10);
  }
  {
    RefField a(ref);
    CudaField b(gpu);
    a.timesMatrixRoot();
    b.timesMatrixRoot();
    report("timesMatrixRoot", relDiff(a, b), 1e-10);
  }
  {
    RefField a(ref);
    CudaField b(gpu);
    a.timesInverseMatrix();
    b.timesInverseMatrix();
    report("timesInverseMatrix (pinned CG)", relDiff(a, b), 1e-8);
  }
  if (dim > 1) { // fieldtest.cc:103-130: M^-1 M x ~ x (CG count is pinned)
    CudaField b(gpu);
    b.timesMatrix();
    b.timesInverseMatrix();
    b -= gpu;
    report("|M^-1 M x - x| / |x| (plugin)",
           std::sqrt(b * b) / std::sqrt(gpu * gpu), 1e-2);
  }
  { // vector space operations of the reference run on plugin fields unchanged
    CudaField b(gpu);
    b *= 2.;
    b.axpy(gpu, -1.);
    b -= gpu;
    report("(2x - x) - x == 0", std::sqrt(b * b), 1e-12);
  }
  { // the same operations against the reference's host loops, value by value
    RefField a(ref);
    CudaField b(gpu);
    a *= 0.75;
    b *= 0.75;
    a.axpy(ref, 0.3);
    b.axpy(gpu, 0.3);
    a += ref;
    b += gpu;
    report("*=, axpy, += (values)", relDiff(a, b), 1e-10);
    report("scalar product", std::abs(a * ref - b * gpu) / std::abs(a * ref), 1e-12);
    report("oneNorm", std::abs(a.oneNorm() - b.oneNorm()) / a.oneNorm(), 1e-12);
    report("infNorm", std::abs(a.infNorm() - b.infNorm()) / a.infNorm(), 1e-14);
    CudaField c(b);
    report("operator== (copy)", c == b ? 0. : 1., 0.5);
    c *= 1.0000001;
    report("operator!= (scaled copy)", c != b ? 0. : 1., 0.5);
#ifndef PARAFIELDS_CUDA_HOST_STOCHASTIC
    typename RefField::Traits::DomainType center;
    for (unsigned int i = 0; i < dim; i++)
      center[i] = 0.4 + 0.1 * i;
    // the reference's own localize does not compile (stochastic.hh:806 calls
    // the non-static indicesToCoords without an object), so its loop
    // (stochastic.hh:798-817, cell centres by fieldtraits.hh:461-469) is
    // applied to the reference's values here
    {
      auto& va = a.stochasticPart.dataVector;
      const auto& tr = *a.traits;
      const double factor = std::pow(2. * 3.14159, -(dim / 2.));
      for (size_t i = 0; i < va.size(); i++) {
        size_t r = i;
        double dist = 0.;
        for (unsigned int j = 0; j < dim; j++) {
          const double x = (r % tr.localCells[j] + 0.5) * tr.extensions[j] / tr.cells[j];
          r /= tr.localCells[j];
          dist += (x - center[j]) * (x - center[j]);
        }
        va[i] *= factor * std::exp(-0.5 * dist / (0.3 * 0.3));
      }
    }
    b.localize(center, 0.3);
    report("localize", relDiff(a, b), 1e-12);
#endif
    a.zero();
    b.zero();
    report("zero", a.oneNorm() + b.oneNorm(), 1e-300);
  }
#ifndef PARAFIELDS_CUDA_HOST_STOCHASTIC
  { // the plugin's field stays in HBM through an operator chain: nothing is
    // staged through the host until values are asked for
    CudaField b(gpu);
    b.timesMatrix();
    b.timesInverseMatrix();
    b *= 2.;
    const bool resident = b.stochasticPart.residentOnDevice() &&
                          !b.stochasticPart.values.onHost();
    report("operator chain stays on the device", resident ? 0. : 1., 0.5);
  }
#endif
  { // evaluation-time value transform through the reference's bulkEvaluate
    std::vector<typename RefField::Traits::RangeType> va, vb;
    typename RefField::Traits::Indices sa, sb;
    RefField a(ref);
    CudaField b(gpu);
    a.bulkEvaluate(va, sa);
    b.bulkEvaluate(vb, sb);
    double num = 0., den = 0.;
    for (size_t i = 0; i < va.size(); i++) {
      num = std::max(num, std::abs(va[i][0] - vb[i][0]));
      den = std::max(den, std::abs(va[i][0]));
    }
    report("bulkEvaluate (transform)", num / den, 1e-10);
  }
  if (dim > 1) { // RandomField::refineMatrix / refine / coarsenMatrix / coarsen
    // (randomfield.hh:829-962) on fields of their own (the traits are shared
    // by copies, and refineMatrix changes them)
    RefField a(config);
    CudaField b(cconfig);
    a.generate(99u);
    b.generate(99u);
    CudaField before(b);
    bool referenceRefines = true;
    try {
      a.refineMatrix();
      a.refine();
    } catch (const std::runtime_error& e) {
      // the reference's DCT matrix backend refuses ("matrix is finalized")
      referenceRefines = false;
      std::printf("  (reference cannot refine this configuration: %s)\n", e.what());
    }
    b.refineMatrix();
    b.refine();
    if (referenceRefines)
      report("refineMatrix + refine", relDiff(a, b), 1e-10);
    if (referenceRefines) {
      a.coarsenMatrix();
      a.coarsen();
    }
    b.coarsenMatrix();
    b.coarsen();
    if (referenceRefines)
      report("coarsenMatrix + coarsen", relDiff(a, b), 1e-10);
    // coarsen(refine(x)) averages 2^dim equal values: up to the rounding of the
    // sums (stochastic.hh:725,745)
    b -= before;
    report("coarsen(refine(x)) == x (plugin)", std::sqrt(b * b) / std::sqrt(before * before), 1e-14);
  }
  { // device RNG: different from the reference stream but a valid field
    Dune::ParameterTree pconfig = config;
    CudaField p(pconfig);
    p.generate(7u);
    const double n2 = p * p / p.dofs();
    report("philox field: |mean square - 1|", std::abs(n2 - 1.), 0.6);
  }
}

// custom-iso / custom-aniso (randomfield.hh:445-452, matrix.hh:270-279): a
// covariance class of the CALLER, installed with fillMatrix<Cov>() on both sides
struct UserCovariance
{
  explicit UserCovariance(const Dune::ParameterTree&) {}
  template<typename RF, long unsigned int dim>
  RF operator()(const RF variance, const std::array<RF, dim>& x) const
  {
    RF sum = 0.;
    for (unsigned int i = 0; i < dim; i++)
      sum += x[i] * x[i];
    const RF h = std::sqrt(sum);
    return variance * (1. + 0.5 * h) * std::exp(-1.5 * h);
  }
};
struct SkewCovariance // not symmetric per dimension: needs custom-aniso
{
  explicit SkewCovariance(const Dune::ParameterTree&) {}
  template<typename RF, long unsigned int dim>
  RF operator()(const RF variance, const std::array<RF, dim>& x) const
  {
    RF sum = 0.;
    for (unsigned int i = 0; i < dim; i++)
      sum += x[i] * x[i];
    if (dim > 1)
      sum += 0.8 * x[0] * x[1];
    return variance * std::exp(-std::sqrt(sum));
  }
};

template<unsigned int dim, typename Cov>
void
runCustom(const std::string& name, const std::string& ini)
{
  using GT = GridTraits<double, double, dim>;
  using RefField = parafields::RandomField<GT>;
  using CudaField =
    parafields::RandomField<GT, parafields::CudaMatrix, parafields::CudaMatrix>;
  std::printf("%s\n", name.c_str());
  Dune::ParameterTree config;
  std::istringstream in(ini);
  Dune::ParameterTreeParser::readINITree(in, config);
  Dune::ParameterTree cconfig = config;
  cconfig["cuda.noise"] = "reference";
  RefField ref(config);
  CudaField gpu(cconfig);
  bool threw = false;
  try {
    gpu.generate(1u); // matrix.hh:236-239: the class has to be supplied first
  } catch (const std::runtime_error&) {
    threw = true;
  }
  report("generate before fillMatrix<Cov> throws", threw ? 0. : 1., 0.5);
  ref.template fillMatrix<Cov>();
  gpu.template fillMatrix<Cov>();
  ref.generate(20240611u);
  gpu.generate(20240611u);
  report("generate(seed), custom covariance", relDiff(ref, gpu), 1e-10);
  {
    RefField a(ref);
    CudaField b(gpu);
    a.timesMatrix();
    b.timesMatrix();
    report("timesMatrix, custom covariance", relDiff(a, b), 1e-10);
  }
  {
    RefField a(ref);
    CudaField b(gpu);
    a.timesMatrixRoot();
    b.timesMatrixRoot();
    report("timesMatrixRoot, custom covariance", relDiff(a, b), 1e-10);
  }
}

// `dropin_test chain [cells] [repeats]`: K5's operator chain
// timesMatrix -> timesInverseMatrix through the reference's RandomField on the
// plugin alone (3D cells^3 exponential, caches off, CG count pinned), timed
// with the host clock around synchronised calls.  Compiled a second time with
// PARAFIELDS_CUDA_HOST_STOCHASTIC (the reference's host StochasticPart, every
// operand staged through PCIe) for the figure beside it.
#include <chrono>
static int
chain(unsigned int cells, int repeats)
{
  using GT = GridTraits<double, double, 3>;
  using CudaField =
    parafields::RandomField<GT, parafields::CudaMatrix, parafields::CudaMatrix>;
  const std::string c = std::to_string(cells);
  const std::string ini =
    "[grid]\nextensions = 1 1 1\ncells = " + c + " " + c + " " + c +
    "\n[stochastic]\nvariance = 1.\ncovariance = exponential\ncorrLength = 0.05\n"
    "[randomField]\ncgIterations = 10\ncacheInvMatvec = 0\ncacheInvRootMatvec = 0\n"
    "[cuda]\nrng = fast\n";
  Dune::ParameterTree config;
  std::istringstream in(ini);
  Dune::ParameterTreeParser::readINITree(in, config);
  CudaField field(config);
  field.generate(42u);
  pf_plan* plan = field.isoMatrix->cudaPlan();
  auto now = [&] {
    pf_plan_synchronize(plan);
    return std::chrono::steady_clock::now();
  };
  auto ms = [](auto a, auto b) {
    return std::chrono::duration<double, std::milli>(b - a).count();
  };
  double tMat = 0., tInv = 0., tVec = 0.;
  CudaField other(field);
  if (std::getenv("PF_CHAIN_DETAIL")) { // where the time of one call goes
    for (int r = 0; r < 3; r++) {
      const auto t0 = now();
      auto product = field.isoMatrix->multiplyInverse(field.stochasticPart);
      const auto t1 = now();
      field.stochasticPart = product;
      const auto t2 = now();
      field.trendPart.timesInverseMatrix();
      const auto t3 = now();
      std::fprintf(stderr, "multiplyInverse %.3f ms, assignment %.3f ms, trend %.3f ms\n",
                   ms(t0, t1), ms(t1, t2), ms(t2, t3));
    }
  }
  for (int r = -1; r < repeats; r++) { // one untimed round first
    const auto t0 = now();
    field.timesMatrix();
    const auto t1 = now();
    field.timesInverseMatrix();
    const auto t2 = now();
    field.axpy(other, 0.5);
    field *= 0.5;
    const double dot = field * other;
    const auto t3 = now();
    if (!(dot == dot))
      return 2;
    if (r >= 0) {
      tMat += ms(t0, t1);
      tInv += ms(t1, t2);
      tVec += ms(t2, t3);
    }
  }
#ifdef PARAFIELDS_CUDA_HOST_STOCHASTIC
  const char* where = "host StochasticPart of the reference (values staged per call)";
#else
  const char* where = field.stochasticPart.values.onHost()
                        ? "device-resident StochasticPart, HOST MIRROR TOUCHED"
                        : "device-resident StochasticPart, host mirror never filled";
#endif
  std::printf("{\"chain\": \"K5 timesMatrix -> timesInverseMatrix (10 CG iterations) -> axpy, *=, "
              "scalar product through RandomField\", \"cells\": %u, \"repeats\": %d, "
              "\"timesMatrix_ms\": %.4f, \"timesInverseMatrix_ms\": %.4f, \"vector_ops_ms\": %.4f, "
              "\"chain_ms\": %.4f, \"fields\": \"%s\"}\n",
              cells, repeats, tMat / repeats, tInv / repeats, tVec / repeats,
              (tMat + tInv + tVec) / repeats, where);
  return 0;
}

int
main(int argc, char** argv)
{
  if (argc > 1 && std::string(argv[1]) == "chain")
    return chain(argc > 2 ? std::stoul(argv[2]) : 256u,
                 argc > 3 ? std::stoi(argv[3]) : 5);
  const std::string tail = "[stochastic]\nvariance = 1.\n";
  run<1>("1D 32, exponential (test/randomfield1d.ini)",
         "[grid]\nextensions = 1\ncells = 32\n" + tail +
           "covariance = exponential\ncorrLength = 0.05\n");
  run<2>("2D 16x16, exponential (test/randomfield2d.ini)",
         "[grid]\nextensions = 1 1\ncells = 16 16\n" + tail +
           "covariance = exponential\ncorrLength = 0.05\n[randomField]\ncgIterations = 4\n");
  run<3>("3D 8x8x8, spherical (test/randomfield3d.ini)",
         "[grid]\nextensions = 1 1 1\ncells = 8 8 8\n" + tail +
           "covariance = spherical\ncorrLength = 0.05\n[randomField]\ncgIterations = 4\n");
  run<3>("3D 16x8x32, gaussian axiparallel, logNormal (K2 scaled)",
         "[grid]\nextensions = 1 1 1\ncells = 16 8 32\n" + tail +
           "covariance = gaussian\nanisotropy = axiparallel\ncorrLength = 0.05 0.03 0.02\n"
           "[embedding]\napproximate = 1\n[randomField]\ntransform = logNormal\ncgIterations = 4\n");
  run<2>("2D 32x64, cubic geometric (K4 scaled)",
         "[grid]\nextensions = 1 1\ncells = 32 64\n" + tail +
           "covariance = cubic\nanisotropy = geometric\ncorrLength = 0.08 0.03 0.03 0.05\n"
           "[embedding]\napproximate = 1\n[randomField]\ncgIterations = 4\n");
  run<2>("2D 128x128, matern32 with caches (K5 scaled)",
         "[grid]\nextensions = 1 1\ncells = 128 128\n" + tail +
           "covariance = matern32\ncorrLength = 0.05\n"
           "[randomField]\ncgIterations = 4\ncacheInvMatvec = 1\ncacheInvRootMatvec = 1\n");
  const std::string capp = "[embedding]\napproximate = 1\n";
  runCustom<1, UserCovariance>("1D 64, custom-iso (user functor; generic route)",
                               "[grid]\nextensions = 1\ncells = 64\n" + tail +
                                 "covariance = custom-iso\ncorrLength = 0.05\n" + capp);
  runCustom<2, UserCovariance>("2D 32x16, custom-iso (user functor)",
                               "[grid]\nextensions = 1 1\ncells = 32 16\n" + tail +
                                 "covariance = custom-iso\ncorrLength = 0.05\n" + capp);
  runCustom<3, UserCovariance>("3D 8x16x8, custom-iso (user functor)",
                               "[grid]\nextensions = 1 1 1\ncells = 8 16 8\n" + tail +
                                 "covariance = custom-iso\ncorrLength = 0.1\n" + capp);
  runCustom<2, SkewCovariance>("2D 16x32, custom-aniso (asymmetric user functor)",
                               "[grid]\nextensions = 1 1\ncells = 16 32\n" + tail +
                                 "covariance = custom-aniso\ncorrLength = 0.05\n" + capp);
  runCustom<2, UserCovariance>("2D 50x30, custom-iso on a non power-of-two grid",
                               "[grid]\nextensions = 1 1\ncells = 50 30\n" + tail +
                                 "covariance = custom-iso\ncorrLength = 0.05\n" + capp);
  std::printf("%s (%d failure%s)\n", failures ? "DROP-IN FAILED" : "DROP-IN OK",
              failures, failures == 1 ? "" : "s");
  return failures ? 1 : 0;
}
