// fieldgenerator_cuda.cpp -- the reference's OWN command line
// (src/fieldgenerator.cc, compiled UNMODIFIED from the reference checkout named
// by PF_REFERENCE_CLI) on top of this repository's CUDA path.
//
// The only change a maintainer makes to switch the CLI over is the default
// matrix of RandomField; here that edit is applied
// from the outside: the class templates are compiled under other names and
// `RandomField<GridTraits>` / `RandomFieldList<GridTraits>` are re-introduced
// as aliases that plug parafields::CudaMatrix in.  Everything else -- the ini
// parser calls, -key value options, seed handling, refinement flags, list
// mode, writers -- is the reference's code.
//
// Linked against standins/fftw_sizes.cpp: the reference's traits class asks
// FFTW-MPI for the slab sizes (fieldtraits.hh:204,368-379); those calls are
// answered, every transform entry point aborts.  So this binary PROVES that no
// CPU FFT runs behind the CLI.
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

#ifndef PF_REFERENCE_CLI
#error "compile with -DPF_REFERENCE_CLI='\"<reference>/src/fieldgenerator.cc\"'"
#endif
#include PF_REFERENCE_CLI
