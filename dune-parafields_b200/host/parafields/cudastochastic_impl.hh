// parafields/cudastochastic_impl.hh -- the StochasticPart specialisation
// behind cudastochastic.hh (included from there).  Member list and meaning
// follow the reference's StochasticPart (stochastic.hh:25-1130); what differs
// is where the values are: cuda_detail::FieldBuffer instead of
// std::vector<RF> dataVector.
#ifndef PARAFIELDS_CUDASTOCHASTIC_IMPL_HH
#define PARAFIELDS_CUDASTOCHASTIC_IMPL_HH

namespace parafields {

template<typename GridTraits>
class StochasticPart<RandomFieldTraits<GridTraits, CudaMatrix, CudaMatrix>>
{
public:
  using Traits = RandomFieldTraits<GridTraits, CudaMatrix, CudaMatrix>;
  // seen by CudaMatrix: the values are handed over as device pointers
  struct DeviceResidentTag {};
  // copy of the sizes without the values (result of a matrix product)
  struct GeometryOnly {};

private:
  using RF = typename Traits::RF;
  using Index = typename Traits::Index;
  using Indices = typename Traits::Indices;
  enum { dim = Traits::dim };
  static_assert(std::is_same<RF, double>::value,
                "the CUDA path computes in fp64 (RangeField must be double)");

  friend class CudaMatrix<Traits>;

  std::shared_ptr<Traits> traits;

  int rank = 0;
  std::array<RF, dim> extensions;
  Indices cells;
  unsigned int level = 0;
  Indices localCells;
  Indices localOffset;
  Index localDomainSize = 0;

  cuda_detail::FieldBuffer values;

  static void check(int rc) { cuda_detail::check(rc); }

  std::array<unsigned, dim> cellArray(const Indices& c) const
  {
    std::array<unsigned, dim> a;
    for (unsigned int i = 0; i < dim; i++)
      a[i] = c[i];
    return a;
  }

public:
  // stochastic.hh:80-118: read <fileName>.stoch.h5 or start as the zero field
  StochasticPart(const std::shared_ptr<Traits>& traits_,
                 const std::string& fileName)
    : traits(traits_)
    , values(traits_.get())
  {
    update();

    if (fileName != "") {
#if HAVE_HDF5
      if (!fileExists(fileName + ".stoch.h5"))
        throw std::runtime_error{ "File is missing: " + fileName +
                                  ".stoch.h5" };
      if ((*traits).verbose && rank == 0)
        std::cout << "loading random field from file " << fileName << std::endl;
      std::vector<RF> fromFile;
      readParallelFromHDF5<RF, dim>(fromFile,
                                    localCells,
                                    localOffset,
                                    (*traits).comm,
                                    "/stochastic",
                                    fileName + ".stoch.h5");
      values.resize(fromFile.size());
      values.hostOverwrite() = fromFile;
#else
      throw std::runtime_error{
        "Writing and reading field files requires parallel HDF5 support"
      };
#endif
    } else {
      if ((*traits).verbose && rank == 0)
        std::cout << "generating homogeneous random field" << std::endl;
      zero();
    }
  }

  StochasticPart(const StochasticPart& other, GeometryOnly)
    : traits(other.traits)
    , rank(other.rank)
    , extensions(other.extensions)
    , cells(other.cells)
    , level(other.level)
    , localCells(other.localCells)
    , localOffset(other.localOffset)
    , localDomainSize(other.localDomainSize)
    , values(other.traits.get())
  {
  }

  // stochastic.hh:226-262 for one process (CudaMatrix refuses commSize > 1;
  // a multi-GPU plan is driven by one process)
  void update()
  {
    rank = (*traits).rank;
    extensions = (*traits).extensions;
    cells = (*traits).cells;
    level = (*traits).level;
    localCells = (*traits).localCells;
    localOffset = (*traits).localOffset;
    localDomainSize = (*traits).localDomainSize;
    if ((*traits).commSize != 1)
      throw std::runtime_error{
        "StochasticPart (CUDA): one process drives the GPUs of a plan; MPI "
        "ranks > 1 are not supported"
      };
  }

  // stochastic.hh:272-296
  void writeToFile(const std::string& fileName) const
  {
#if HAVE_HDF5
    if ((*traits).verbose && rank == 0)
      std::cout << "writing random field to file " << fileName << std::endl;
    writeParallelToHDF5<RF, dim>((*traits).cells,
                                 values.hostRead(),
                                 localCells,
                                 localOffset,
                                 (*traits).comm,
                                 "/stochastic",
                                 fileName + ".stoch.h5");
    if (rank == 0)
      writeToXDMF<RF, dim>((*traits).cells, (*traits).extensions, fileName);
#else
    (void)fileName;
    throw std::runtime_error{
      "Writing and reading field files requires parallel HDF5 support"
    };
#endif
  }

  // stochastic.hh:306-313
  unsigned int dofs() const { return localDomainSize; }

  // stochastic.hh:320-336
  StochasticPart& operator+=(const StochasticPart& other)
  {
    return combine(other, 1.);
  }

  // stochastic.hh:345-361
  StochasticPart& operator-=(const StochasticPart& other)
  {
    return combine(other, -1.);
  }

  // stochastic.hh:370-380
  StochasticPart& operator*=(const RF alpha)
  {
    if (!values.empty())
      check(pf_vec_scale(values.plan(), values.size(), alpha,
                         values.deviceUpdate()));
    return *this;
  }

  // stochastic.hh:392-408
  StochasticPart& axpy(const StochasticPart& other, const RF alpha)
  {
    return combine(other, alpha);
  }

  // stochastic.hh:419-429: zero when the sizes differ
  RF operator*(const StochasticPart& other) const
  {
    RF sum = 0.;
    if (values.size() == other.values.size() && !values.empty())
      check(pf_vec_dot(values.plan(), values.size(), values.deviceRead(),
                       other.values.deviceRead(), &sum));
    return sum;
  }

  // stochastic.hh:438-452
  bool operator==(const StochasticPart& other) const
  {
    if (values.size() != other.values.size())
      return false;
    if (values.empty())
      return true;
    int same = 0;
    check(pf_vec_equal(values.plan(), values.size(), values.deviceRead(),
                       other.values.deviceRead(), &same));
    return same != 0;
  }
  bool operator!=(const StochasticPart& other) const
  {
    return !operator==(other);
  }

  // stochastic.hh:480-574 for one process: the block is the whole field
  // (dataToEval, stochastic.hh:848-852) and the overlap rows are the
  // zero-filled vectors of stochastic.hh:838-843, which a location on the
  // upper face of exactly one dimension reads.  The values come from the host
  // mirror, filled once after the field last changed on the device.
  void evaluate(const typename Traits::DomainType& location,
                typename Traits::RangeType& output) const
  {
    if (values.empty()) {
      // stochastic.hh:832-836: an empty field evaluates as zero
      output[0] = 0.;
      return;
    }
    const std::vector<RF>& data = values.hostRead();

    Indices indices, origin;
    origin.fill(0);
    (*traits).coordsToIndices(location, indices, origin);

    unsigned int onFace = 0;
    for (unsigned int i = 0; i < dim; i++)
      if (indices[i] >= localCells[i])
        onFace++;
    if (onFace == 1 && dim <= 3) {
      output[0] = 0.;
      return;
    }
    if (onFace != 0 && dim > 3)
      throw std::runtime_error{
        "overlap only available for dimensions 1, 2 and 3"
      };
    for (unsigned int i = 0; i < dim; i++)
      if (indices[i] >= localCells[i])
        indices[i] = localCells[i] - 1;
    output[0] = data[Traits::indicesToIndex(indices, localCells)];
  }

  // stochastic.hh:581-587 (an empty field stays empty)
  void zero()
  {
    if (!values.empty())
      check(pf_vec_fill(values.plan(), values.size(), 0.,
                        values.deviceOverwrite()));
  }

  // stochastic.hh:597-670
  void refine()
  {
    if (level == (*traits).level)
      return;
    update();
    if (values.empty())
      return;
    Indices oldLocalCells;
    for (unsigned int i = 0; i < dim; i++)
      oldLocalCells[i] = localCells[i] / 2;
    transfer(oldLocalCells, pf_field_refine);
  }

  // stochastic.hh:679-755
  void coarsen()
  {
    if (level == (*traits).level)
      return;
    update();
    if (values.empty())
      return;
    Indices oldLocalCells;
    for (unsigned int i = 0; i < dim; i++)
      oldLocalCells[i] = localCells[i] * 2;
    transfer(oldLocalCells, pf_field_coarsen);
  }

  // stochastic.hh:761-771
  RF oneNorm() const
  {
    RF sum = 0.;
    if (!values.empty())
      check(pf_vec_norm1(values.plan(), values.size(), values.deviceRead(),
                         &sum));
    return sum;
  }

  // stochastic.hh:778-788
  RF infNorm() const
  {
    RF max = 0.;
    if (!values.empty())
      check(pf_vec_norminf(values.plan(), values.size(), values.deviceRead(),
                           &max));
    return max;
  }

  // stochastic.hh:798-817
  void localize(const typename Traits::DomainType& center, const RF radius)
  {
    if (values.empty())
      return;
    const auto c = cellArray(localCells);
    const auto offset = cellArray(localOffset);
    const auto global = cellArray(cells);
    std::array<double, dim> ext, mid;
    for (unsigned int i = 0; i < dim; i++) {
      ext[i] = extensions[i];
      mid[i] = center[i];
    }
    check(pf_field_localize(values.plan(), dim, c.data(), offset.data(),
                            global.data(), ext.data(), mid.data(), radius,
                            values.deviceUpdate()));
  }

  // the values, for callers that keep working on the device (the pointer is
  // valid until the field is next modified) or want them on the host
  const double* deviceData() const
  {
    return values.empty() ? nullptr : values.deviceRead();
  }
  const std::vector<RF>& hostData() const { return values.hostRead(); }
  std::size_t size() const { return values.size(); }
  bool residentOnDevice() const { return values.onDevice(); }

private:
  // this += alpha other with the reference's rule for fields that were never
  // filled (stochastic.hh:322-331): an empty left side takes alpha other
  StochasticPart& combine(const StochasticPart& other, const RF alpha)
  {
    if (other.values.empty())
      return *this;
    if (values.empty()) {
      values.resize(localDomainSize);
      check(pf_vec_fill(values.plan(), values.size(), 0.,
                        values.deviceOverwrite()));
    }
    if (values.size() != other.values.size())
      throw std::runtime_error{
        "StochasticPart (CUDA): fields of different size"
      };
    check(pf_vec_axpy(values.plan(), values.size(), alpha,
                      other.values.deviceRead(), values.deviceUpdate()));
    return *this;
  }

  template<typename Transfer>
  void transfer(const Indices& oldLocalCells, Transfer kernel)
  {
    pf_plan* plan = values.plan();
    void* fresh = nullptr;
    check(pf_device_alloc(plan, std::size_t(localDomainSize) * sizeof(double),
                          &fresh));
    const auto old = cellArray(oldLocalCells);
    const int rc = kernel(plan, dim, old.data(), values.deviceRead(),
                          static_cast<double*>(fresh));
    if (rc != 0) {
      pf_device_free(plan, fresh);
      check(rc);
    }
    values.adoptDevice(static_cast<double*>(fresh), localDomainSize);
  }
};

} // namespace parafields

#endif // PARAFIELDS_CUDASTOCHASTIC_IMPL_HH
