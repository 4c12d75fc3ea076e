// Stand-in for <dune/common/power.hh> (only included, not used, by the field
// code; kept so that user code including it keeps compiling).
#ifndef PF_STANDIN_DUNE_COMMON_POWER_HH
#define PF_STANDIN_DUNE_COMMON_POWER_HH
namespace Dune {
template<class T>
constexpr T
power(T x, unsigned n)
{
  T r = T(1);
  for (unsigned i = 0; i < n; i++) r *= x;
  return r;
}
} // namespace Dune
#endif
