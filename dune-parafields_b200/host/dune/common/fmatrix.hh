// Stand-in for dune-common's <dune/common/fmatrix.hh> (dune-common is not in
// this image).  The sampling path of parafields-core only includes it; no
// FieldMatrix operation is used without dune-grid.
#ifndef DUNE_COMMON_FMATRIX_HH
#define DUNE_COMMON_FMATRIX_HH
#include <array>
#include <cstddef>

#include <dune/common/fvector.hh>

namespace Dune {
template<class K, int ROWS, int COLS>
class FieldMatrix
{
  std::array<FieldVector<K, COLS>, ROWS> r{};

public:
  FieldVector<K, COLS>& operator[](std::size_t i) { return r[i]; }
  const FieldVector<K, COLS>& operator[](std::size_t i) const { return r[i]; }
};
} // namespace Dune
#endif
