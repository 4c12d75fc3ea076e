// Host build of the product's counter-based generator (csrc/pf_rng.cuh): prints
// philox4x32_r<R>(ctr, realization, keys(seed)) for the argument tuples given
// on stdin ("R c0 c1 c2 c3 k0 k1", hex), one result line each.  Built and run
// by tests/test_philox.py (no GPU needed); compared there with the Random123
// known-answer vectors and with an independent restatement.
#include "pf_rng.cuh"
#include <cstdio>
using namespace pf;

int main()
{
  unsigned r, c[4], k[2];
  while (scanf("%u %x %x %x %x %x %x", &r, &c[0], &c[1], &c[2], &c[3], &k[0], &k[1]) == 7) {
    const unsigned long long ctr = ((unsigned long long)c[1] << 32) | c[0];
    const unsigned long long rea = ((unsigned long long)c[3] << 32) | c[2];
    const unsigned long long seed = ((unsigned long long)k[1] << 32) | k[0];
    const PhiloxKeys keys = philox_keys(seed);
    uint4 w;
    if (r == 7)
      w = philox4x32_r<7>(ctr, rea, keys);
    else if (r == 10)
      w = philox4x32_r<10>(ctr, rea, keys);
    else
      return 2;
    printf("%08x %08x %08x %08x\n", w.x, w.y, w.z, w.w);
  }
  return 0;
}
