// oracle/fieldgenerator_cuda.cpp -- the reference's OWN command line
// (src/fieldgenerator.cc, compiled unmodified from /root/reference) on top of
// this repository's CUDA path.  TEST INFRASTRUCTURE / drop-in proof.
//
// The only change a maintainer makes to switch the CLI over is the default
// matrix of RandomField; here that edit is applied
// from the outside: the class templates are compiled under other names and
// `RandomField<GridTraits>` / `RandomFieldList<GridTraits>` are re-introduced
// as aliases that plug parafields::CudaMatrix in.  Everything else -- the ini
// parser calls, -key value options, seed handling, refinement flags, writers
// -- is the reference's code.
#define RandomField RandomFieldTemplate
#define RandomFieldList RandomFieldListTemplate
#include <parafields/cudamatrix.hh>
#include <parafields/randomfield.hh>
#undef RandomField
#undef RandomFieldList

namespace parafields {
template<typename GridTraits>
using RandomField = RandomFieldTemplate<GridTraits, CudaMatrix, CudaMatrix>;
template<typename GridTraits>
using RandomFieldList =
  RandomFieldListTemplate<GridTraits, CudaMatrix, CudaMatrix, RandomFieldTemplate>;
} // namespace parafields

#include "../../reference/src/fieldgenerator.cc"
