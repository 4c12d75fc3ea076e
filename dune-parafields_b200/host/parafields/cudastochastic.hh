// parafields/cudastochastic.hh -- StochasticPart whose values live in HBM.
//
// The reference keeps the field in `std::vector<RF> dataVector`
// (stochastic.hh:54) and runs every vector operation as a host loop
// (stochastic.hh:320-467, 581-587, 597-817).  With CudaMatrix as both matrix
// parameters of RandomField the field is produced and consumed on the GPU, so
// this header specialises
//
//   StochasticPart<RandomFieldTraits<GridTraits, CudaMatrix, CudaMatrix>>
//
// to own a device buffer next to a host mirror with "which copy is current"
// flags (SURVEY.md 8(b) Ownership).  generate / timesMatrix /
// timesInverseMatrix and += -= *= axpy * == norms zero refine coarsen localize
// run through the C ABI (pf_vec_*, pf_field_*) without the values leaving the
// device; the host mirror is only filled when the caller asks for values on
// the host (evaluate, writeToFile).  It is picked up by including
// parafields/cudamatrix.hh BEFORE parafields/randomfield.hh; define
// PARAFIELDS_CUDA_HOST_STOCHASTIC to keep the reference's host class.
#ifndef PARAFIELDS_CUDASTOCHASTIC_HH
#define PARAFIELDS_CUDASTOCHASTIC_HH

#include <array>
#include <cmath>
#include <cstddef>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include <parafields_cuda.h>

#if HAVE_HDF5
#include <parafields/io.hh>
#endif

namespace parafields {

template<typename GridTraits,
         template<typename> class IsoMatrix,
         template<typename> class AnisoMatrix>
class RandomFieldTraits;

template<typename Traits>
class StochasticPart;

template<typename Traits>
class CudaMatrix;

namespace cuda_detail {

// Which plan serves the fields of a traits object.  RandomField creates its
// matrix and its StochasticPart from the same shared traits
// (randomfield.hh:138-161) but does not hand one to the other, so the matrix
// announces its plan here and the fields look it up when they first need the
// device.
class PlanRegistry
{
  std::mutex guard;
  std::map<const void*, pf_plan*> plans;

  static PlanRegistry& instance()
  {
    static PlanRegistry registry;
    return registry;
  }

public:
  static void announce(const void* traits, pf_plan* plan)
  {
    auto& r = instance();
    std::lock_guard<std::mutex> lock(r.guard);
    r.plans[traits] = plan;
  }
  static void withdraw(const void* traits, pf_plan* plan)
  {
    auto& r = instance();
    std::lock_guard<std::mutex> lock(r.guard);
    auto it = r.plans.find(traits);
    if (it != r.plans.end() && it->second == plan)
      r.plans.erase(it);
  }
  static pf_plan* find(const void* traits)
  {
    auto& r = instance();
    std::lock_guard<std::mutex> lock(r.guard);
    auto it = r.plans.find(traits);
    return it == r.plans.end() ? nullptr : it->second;
  }
};

inline void
check(int rc)
{
  if (rc != 0)
    throw std::runtime_error{ pf_last_error() };
}

// Device arrays given back by fields of one traits object, kept for the next
// field of the same size: every matrix product returns a new StochasticPart
// (matrix.hh:151-215) and the old one is destroyed right after the assignment,
// so without this each product would pay a cudaMalloc and a cudaFree (which
// synchronises the device).  Emptied when the matrix of these traits goes.
class BufferPool
{
  struct Entry
  {
    double* ptr;
    std::size_t capacity;
  };
  std::mutex guard;
  std::map<const void*, std::vector<Entry>> free;
  enum { maxKept = 4 };

  static BufferPool& instance()
  {
    static BufferPool pool;
    return pool;
  }

public:
  // an array of at least n doubles (and not more than twice that), or nullptr
  static double* take(const void* key, std::size_t n, std::size_t& capacity)
  {
    auto& p = instance();
    std::lock_guard<std::mutex> lock(p.guard);
    auto& list = p.free[key];
    for (std::size_t i = 0; i < list.size(); i++)
      if (list[i].capacity >= n && list[i].capacity <= 2 * n) {
        double* ptr = list[i].ptr;
        capacity = list[i].capacity;
        list.erase(list.begin() + i);
        return ptr;
      }
    return nullptr;
  }
  // false: the pool is full, the caller frees the array
  static bool give(const void* key, double* ptr, std::size_t capacity)
  {
    auto& p = instance();
    std::lock_guard<std::mutex> lock(p.guard);
    auto& list = p.free[key];
    if (list.size() >= maxKept)
      return false;
    list.push_back(Entry{ ptr, capacity });
    return true;
  }
  static void drain(const void* key, pf_plan* plan)
  {
    auto& p = instance();
    std::lock_guard<std::mutex> lock(p.guard);
    auto it = p.free.find(key);
    if (it == p.free.end())
      return;
    for (auto& e : it->second)
      pf_device_free(plan, e.ptr);
    p.free.erase(it);
  }
};

// `count` doubles with a device copy and a host copy; at most one of them is
// stale.  count == 0 is the reference's "dataVector.empty()" state (a field
// that was never generated or read, stochastic.hh:322,347,372,394).
class FieldBuffer
{
  const void* key; // the traits object the plan is announced under
  std::size_t count = 0;
  mutable double* dev = nullptr;
  mutable std::size_t devCapacity = 0;
  mutable std::vector<double> host;
  mutable bool devCurrent = false;
  mutable bool hostCurrent = false;

  void reserveDevice(std::size_t n) const
  {
    if (devCapacity >= n && dev)
      return;
    releaseDevice();
    dev = BufferPool::take(key, n, devCapacity);
    if (dev)
      return;
    void* p = nullptr;
    check(pf_device_alloc(plan(), n * sizeof(double), &p));
    dev = static_cast<double*>(p);
    devCapacity = n;
  }
  void releaseDevice() const
  {
    // the pool only serves while the matrix of these traits (and with it the
    // plan whose stream orders the work on the array) exists
    pf_plan* p = PlanRegistry::find(key);
    if (dev && !(p && BufferPool::give(key, dev, devCapacity)))
      pf_device_free(p, dev);
    dev = nullptr;
    devCapacity = 0;
    devCurrent = false;
  }

public:
  explicit FieldBuffer(const void* key_)
    : key(key_)
  {
  }
  FieldBuffer(const FieldBuffer& other)
    : key(other.key)
  {
    *this = other;
  }
  FieldBuffer& operator=(const FieldBuffer& other)
  {
    if (this == &other)
      return *this;
    key = other.key;
    count = other.count;
    devCurrent = hostCurrent = false;
    if (count == 0)
      return *this;
    if (other.devCurrent) {
      // device to device, the host is not involved
      reserveDevice(count);
      check(pf_memcpy(plan(), dev, other.dev, count * sizeof(double)));
      devCurrent = true;
    } else {
      host = other.host;
      hostCurrent = true;
    }
    return *this;
  }
  ~FieldBuffer() { releaseDevice(); }

  pf_plan* plan() const
  {
    pf_plan* p = PlanRegistry::find(key);
    if (!p)
      throw std::runtime_error{
        "StochasticPart (CUDA): no CudaMatrix was built for these traits"
      };
    return p;
  }

  std::size_t size() const { return count; }
  bool empty() const { return count == 0; }
  bool onDevice() const { return devCurrent; }
  bool onHost() const { return hostCurrent; }

  // forget the values; the next access decides where they are written
  void resize(std::size_t n)
  {
    if (n != count) {
      count = n;
      devCurrent = hostCurrent = false;
      host.clear();
    }
  }
  void clear() { resize(0); }

  // device views
  const double* deviceRead() const
  {
    if (!devCurrent) {
      if (!hostCurrent)
        throw std::logic_error{ "FieldBuffer: no current copy" };
      reserveDevice(count);
      check(pf_memcpy(plan(), dev, host.data(), count * sizeof(double)));
      devCurrent = true;
    }
    return dev;
  }
  double* deviceUpdate()
  {
    deviceRead();
    hostCurrent = false;
    return dev;
  }
  double* deviceOverwrite()
  {
    reserveDevice(count);
    devCurrent = true;
    hostCurrent = false;
    return dev;
  }

  // host views
  const std::vector<double>& hostRead() const
  {
    if (!hostCurrent) {
      host.resize(count);
      if (count) {
        if (!devCurrent)
          throw std::logic_error{ "FieldBuffer: no current copy" };
        check(pf_memcpy(plan(), host.data(), dev, count * sizeof(double)));
      }
      hostCurrent = true;
    }
    return host;
  }
  std::vector<double>& hostOverwrite()
  {
    host.resize(count);
    hostCurrent = true;
    devCurrent = false;
    return host;
  }

  // take over a device array of n doubles made by the caller with
  // pf_device_alloc (refine / coarsen write into a new array)
  void adoptDevice(double* p, std::size_t n)
  {
    releaseDevice();
    dev = p;
    devCapacity = n;
    count = n;
    devCurrent = true;
    hostCurrent = false;
    host.clear();
  }
};

} // namespace cuda_detail
} // namespace parafields

#ifndef PARAFIELDS_CUDA_HOST_STOCHASTIC
#include <parafields/cudastochastic_impl.hh>
#endif

#endif // PARAFIELDS_CUDASTOCHASTIC_HH
