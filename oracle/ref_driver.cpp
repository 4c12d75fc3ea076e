// oracle/ref_driver.cpp -- C entry points around the UNMODIFIED reference
// headers (/root/reference/ext/parafields-core/include/parafields/*.hh).
// TEST INFRASTRUCTURE ONLY: built into oracle/_ref/libpf_ref.so by
// oracle/Makefile when /root/reference is present.  It is used to
//   (i)  validate the CPU restatement in oracle/pf_oracle.cpp,
//   (ii) generate the golden vectors under tests/golden/ (tools in
//        tests/golden/make_golden.py), and
//   (iii) optionally serve as the timed CPU baseline of bench.py
//        (cpu_baseline.kind = "reference").
// The reference's third-party dependencies that are absent from this image
// are replaced by single-rank stand-ins: FFTW3(-MPI) and MPI by oracle/shim/,
// dune-common by dune-parafields_b200/host/dune/common/.  Everything between
// those boundaries -- index math, covariance functors, mirroring, Hermitian
// noise rule, RNG backend, CG, cache state machine -- is the reference's own
// code.  No reference source is copied into this repository.
#include <array>
#include <cstring>
#include <iostream>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>
#include <random>
#include <map>
#include <fstream>
#include <cmath>
#include <complex>
#include <algorithm>
#include <functional>
#include <numeric>
#include <chrono>

#include <dune/common/fvector.hh>
#include <dune/common/parametertree.hh>
#include <dune/common/parametertreeparser.hh>

// white-box access to the matrix / field backends of the reference: this
// file (test code only) is compiled with -fno-access-control
#include <parafields/randomfield.hh>

namespace {

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

thread_local std::string g_err;

struct RefBase
{
  virtual ~RefBase() = default;
  virtual void generate(unsigned seed) = 0;
  virtual void generate_uncorrelated(unsigned seed) = 0;
  virtual size_t size() = 0;
  virtual void get(double* out) = 0;
  virtual void set(const double* in) = 0;
  virtual void times(int op) = 0;
  virtual size_t lambda_size() = 0;
  virtual void lambda(double* out) = 0;
  virtual size_t stored_lambda_size() = 0;
  virtual void stored_lambda(double* out) = 0;
  virtual void bulk(double* out) = 0;
  virtual double dot_self() = 0;
  virtual void refine() = 0;
  virtual void coarsen() = 0;
  virtual void evaluate(size_t n, const double* loc, double* out) = 0;
  virtual void write(const std::string& base) = 0;
};

template<unsigned int dim>
struct Ref : RefBase
{
  using GT = GridTraits<double, double, dim>;
  using Field = parafields::RandomField<GT>;
  using Traits = typename Field::Traits;
  std::unique_ptr<Field> field;

  explicit Ref(const Dune::ParameterTree& config)
    : field(new Field(config))
  {}
  // RandomField(fileName): config from <base>.field, values from
  // <base>.stoch.h5 / <base>.trend (randomfield.hh:190-225)
  explicit Ref(const std::string& base)
    : field(new Field(base))
  {}
  // RandomField::writeToFile (randomfield.hh:656-663)
  void write(const std::string& base) override { field->writeToFile(base); }

  void generate(unsigned seed) override { field->generate(seed); }
  void generate_uncorrelated(unsigned seed) override
  {
    field->generateUncorrelated(seed);
  }
  size_t size() override { return field->stochasticPart.dataVector.size(); }
  void get(double* out) override
  {
    const auto& v = field->stochasticPart.dataVector;
    std::copy(v.begin(), v.end(), out);
  }
  void set(const double* in) override
  {
    auto& sp = field->stochasticPart;
    sp.dataVector.assign(in, in + (*field->traits).localDomainSize);
    sp.evalValid = false;
    field->invMatvecValid = false;
    field->invRootMatvecValid = false;
  }
  void times(int op) override
  {
    switch (op) {
      case 0:
        field->timesMatrix();
        break;
      case 1:
        field->timesMatrixRoot();
        break;
      case 2:
        field->timesInverseMatrix();
        break;
      case 3:
        field->timesInvMatRoot();
        break;
      default:
        throw std::runtime_error("unknown op");
    }
  }

  template<class Matrix>
  void lambda_from(Matrix& m, double* out, size_t* count)
  {
    if (!m.matrixBackend.valid())
      m.fillTransformedMatrix();
    auto& fb = m.fieldBackend;
    fb.transposeIfNeeded();
    const auto n = fb.localFieldSize();
    if (out) {
      typename Traits::Indices indices;
      for (typename Traits::Index index = 0; index < n; index++) {
        Traits::indexToIndices(index, indices, fb.localFieldCells());
        out[index] = m.matrixBackend.eval(indices);
      }
    }
    fb.transposeIfNeeded(); // toggle back
    *count = n;
  }
  // eigenvalues in the order generateField consumes them (matrix.hh:396-421)
  size_t lambda_size() override
  {
    size_t n = 0;
    if (field->useAnisoMatrix)
      lambda_from(*field->anisoMatrix, nullptr, &n);
    else
      lambda_from(*field->isoMatrix, nullptr, &n);
    return n;
  }
  void lambda(double* out) override
  {
    size_t n = 0;
    if (field->useAnisoMatrix)
      lambda_from(*field->anisoMatrix, out, &n);
    else
      lambda_from(*field->isoMatrix, out, &n);
  }
  template<class Matrix>
  size_t stored_from(Matrix& m, double* out)
  {
    if (!m.matrixBackend.valid())
      m.fillTransformedMatrix();
    size_t n = m.matrixBackend.localMatrixSize();
    try {
      (void)m.matrixBackend.get(0);
      if (out)
        for (size_t i = 0; i < n; i++)
          out[i] = m.matrixBackend.get(i);
    } catch (const std::runtime_error&) {
      // R2CMatrixBackend: get() is closed after finalize(); the real parts
      // are addressed by eval(Index) over the half-spectrum entries
      n = m.fieldBackend.localFieldSize();
      if (out)
        for (size_t i = 0; i < n; i++)
          out[i] = m.matrixBackend.eval((typename Traits::Index)i);
    }
    return n;
  }
  // eigenvalues as stored by the matrix backend after finalize()
  size_t stored_lambda_size() override
  {
    return field->useAnisoMatrix ? stored_from(*field->anisoMatrix, nullptr)
                                 : stored_from(*field->isoMatrix, nullptr);
  }
  void stored_lambda(double* out) override
  {
    if (field->useAnisoMatrix)
      stored_from(*field->anisoMatrix, out);
    else
      stored_from(*field->isoMatrix, out);
  }
  void bulk(double* out) override
  {
    std::vector<typename Traits::RangeType> v;
    typename Traits::Indices sizes;
    field->bulkEvaluate(v, sizes);
    for (size_t i = 0; i < v.size(); i++)
      out[i] = v[i][0];
  }
  double dot_self() override { return (*field) * (*field); }
  // RandomField::refineMatrix + refine (randomfield.hh:825-880)
  void refine() override
  {
    field->refineMatrix();
    field->refine();
  }
  void coarsen() override
  {
    field->coarsenMatrix();
    field->coarsen();
  }
  // RandomField::localize cannot be instantiated: stochastic.hh:806 calls the
  // non-static Traits::indicesToCoords without an object (compile error), so
  // the oracle's localize stays pinned by its closed form only.
  // RandomField::evaluate (randomfield.hh:585-595), point by point
  void evaluate(size_t n, const double* loc, double* out) override
  {
    typename Traits::DomainType x;
    typename Traits::RangeType v;
    for (size_t p = 0; p < n; p++) {
      for (unsigned i = 0; i < dim; i++)
        x[i] = loc[p * dim + i];
      field->evaluate(x, v);
      out[p] = v[0];
    }
  }
};

template<class F>
int
guarded(F&& f)
{
  try {
    f();
    return 0;
  } catch (const std::exception& e) {
    g_err = e.what();
    return 1;
  } catch (...) {
    g_err = "unknown exception";
    return 1;
  }
}

} // namespace

extern "C" {

const char*
ref_last_error(void)
{
  return g_err.c_str();
}

// ini_text: the ini file contents; dimension = number of grid.extensions
void*
ref_create(const char* ini_text)
{
  RefBase* r = nullptr;
  const int rc = guarded([&] {
    Dune::ParameterTree config;
    std::istringstream in(ini_text);
    Dune::ParameterTreeParser::readINITree(in, config);
    const auto ext = config.get<std::vector<double>>("grid.extensions");
    switch (ext.size()) {
      case 1:
        r = new Ref<1>(config);
        break;
      case 2:
        r = new Ref<2>(config);
        break;
      case 3:
        r = new Ref<3>(config);
        break;
      default:
        throw std::runtime_error("dimension must be 1, 2 or 3");
    }
  });
  return rc == 0 ? r : nullptr;
}

// file-reading constructor of RandomField; dim = dimension of the stored field
void*
ref_read(const char* base, int dim)
{
  RefBase* r = nullptr;
  const int rc = guarded([&] {
    switch (dim) {
      case 1:
        r = new Ref<1>(std::string(base));
        break;
      case 2:
        r = new Ref<2>(std::string(base));
        break;
      case 3:
        r = new Ref<3>(std::string(base));
        break;
      default:
        throw std::runtime_error("dimension must be 1, 2 or 3");
    }
  });
  return rc == 0 ? r : nullptr;
}

int
ref_write(void* h, const char* base)
{
  return guarded([&] { static_cast<RefBase*>(h)->write(base); });
}

void
ref_destroy(void* h)
{
  delete static_cast<RefBase*>(h);
}

int
ref_generate(void* h, unsigned seed)
{
  return guarded([&] { static_cast<RefBase*>(h)->generate(seed); });
}
int
ref_generate_uncorrelated(void* h, unsigned seed)
{
  return guarded([&] { static_cast<RefBase*>(h)->generate_uncorrelated(seed); });
}
size_t
ref_size(void* h)
{
  return static_cast<RefBase*>(h)->size();
}
int
ref_get(void* h, double* out)
{
  return guarded([&] { static_cast<RefBase*>(h)->get(out); });
}
int
ref_set(void* h, const double* in)
{
  return guarded([&] { static_cast<RefBase*>(h)->set(in); });
}
int
ref_times(void* h, int op)
{
  return guarded([&] { static_cast<RefBase*>(h)->times(op); });
}
long long
ref_lambda_size(void* h)
{
  long long n = -1;
  guarded([&] { n = (long long)static_cast<RefBase*>(h)->lambda_size(); });
  return n;
}
int
ref_lambda(void* h, double* out)
{
  return guarded([&] { static_cast<RefBase*>(h)->lambda(out); });
}
long long
ref_stored_lambda_size(void* h)
{
  long long n = -1;
  guarded([&] { n = (long long)static_cast<RefBase*>(h)->stored_lambda_size(); });
  return n;
}
int
ref_stored_lambda(void* h, double* out)
{
  return guarded([&] { static_cast<RefBase*>(h)->stored_lambda(out); });
}
int
ref_bulk_evaluate(void* h, double* out)
{
  return guarded([&] { static_cast<RefBase*>(h)->bulk(out); });
}
double
ref_dot_self(void* h)
{
  double d = 0.;
  guarded([&] { d = static_cast<RefBase*>(h)->dot_self(); });
  return d;
}

int
ref_refine(void* h)
{
  return guarded([&] { static_cast<RefBase*>(h)->refine(); });
}
int
ref_coarsen(void* h)
{
  return guarded([&] { static_cast<RefBase*>(h)->coarsen(); });
}
int
ref_evaluate(void* h, size_t n, const double* loc, double* out)
{
  return guarded([&] { static_cast<RefBase*>(h)->evaluate(n, loc, out); });
}

// the reference's fallback RNG backend (backends/cpprngbackend.hh) on its own
int
ref_rng(unsigned seed, size_t n, double* out)
{
  return guarded([&] {
    std::default_random_engine gen(seed);
    std::normal_distribution<double> dist(0., 1.);
    for (size_t i = 0; i < n; i++)
      out[i] = dist(gen);
  });
}

} // extern "C"
