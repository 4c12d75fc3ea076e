// Stand-in for Dune::MPIHelper: without MPI this is dune-common's
// "FakeMPIHelper" (single rank).  The GPU path shards over NCCL ranks, not
// MPI ranks; see DESIGN.md "Multi-GPU".
#ifndef PF_STANDIN_DUNE_COMMON_PARALLEL_MPIHELPER_HH
#define PF_STANDIN_DUNE_COMMON_PARALLEL_MPIHELPER_HH
namespace Dune {
class FakeMPIHelper
{
public:
  constexpr static bool isFake = true;
  typedef int MPICommunicator;
  static MPICommunicator getCommunicator() { return 0; }
  static MPICommunicator getLocalCommunicator() { return 0; }
  static FakeMPIHelper& instance(int&, char**&)
  {
    static FakeMPIHelper h;
    return h;
  }
  static FakeMPIHelper& instance()
  {
    static FakeMPIHelper h;
    return h;
  }
  int rank() const { return 0; }
  int size() const { return 1; }
};
#ifndef PF_HAVE_REAL_MPIHELPER
typedef FakeMPIHelper MPIHelper;
#endif
} // namespace Dune
#endif
