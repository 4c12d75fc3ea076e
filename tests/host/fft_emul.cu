// Host emulation of the warp-level line FFT (pf_fft.cuh): runs the two phases
// of line_fft for all "lanes" sequentially and compares with a naive DFT.
// Built and run by tests/test_fft_emulation.py (no GPU needed).
#include "pf_fft.cuh"
#include <cmath>
#include <array>
#include <cstdio>
#include <vector>
using namespace pf;

template <int N, int DIR>
static double check()
{
  using C = LineCfg<N>;
  std::vector<cplx> x(N), ref(N), tw(C::TW2 > 0 ? C::TW2 : 1);
  for (int i = 0; i < N; i++) {
    x[i].x = std::sin(0.37 * i * i + 1.0) + 0.1 * i / N;
    x[i].y = std::cos(1.13 * i + 0.5 * (i % 7));
  }
  for (int k = 0; k < N; k++) {
    long double re = 0, im = 0;
    for (int j = 0; j < N; j++) {
      long double a = DIR * 2.0L * M_PIl * (long double)((long long)j * k % N) / N;
      re += x[j].x * cosl(a) - x[j].y * sinl(a);
      im += x[j].x * sinl(a) + x[j].y * cosl(a);
    }
    ref[k].x = (double)re;
    ref[k].y = (double)im;
  }
  for (int r = 0; r < C::R2 && C::R2 > 1; r++)
    for (int k = 0; k < 32; k++) {
      long double a = -2.0L * M_PIl * (long double)(r * k) / N;
      tw[r * 32 + k].x = (double)cosl(a);
      tw[r * 32 + k].y = (double)sinl(a);
    }
  std::vector<cplx> smem(C::PITCH + 8);
  std::vector<std::array<cplx, C::E>> regs(C::TPL);
  for (int t = 0; t < C::TPL; t++)
    for (int q = 0; q < C::E; q++)
      regs[t][q] = x[t + C::TPL * q];
  for (int t = 0; t < C::TPL; t++) {
    cplx v[C::E];
    for (int q = 0; q < C::E; q++) v[q] = regs[t][q];
    line_fft_phase1<N, DIR>(v, smem.data(), t);
    for (int q = 0; q < C::E; q++) regs[t][q] = v[q];
  }
  for (int t = 0; t < C::TPL; t++) {
    cplx v[C::E];
    for (int q = 0; q < C::E; q++) v[q] = regs[t][q];
    line_fft_phase2<N, DIR>(v, smem.data(), t, tw.data());
    for (int q = 0; q < C::E; q++) regs[t][q] = v[q];
  }
  double err = 0, mx = 0;
  for (int t = 0; t < C::TPL; t++)
    for (int q = 0; q < C::E; q++) {
      const int e = t + C::TPL * q;
      err = std::fmax(err, std::hypot(regs[t][q].x - ref[e].x, regs[t][q].y - ref[e].y));
      mx = std::fmax(mx, std::hypot(ref[e].x, ref[e].y));
    }
  return err / mx;
}

template <int N>
static int run()
{
  double ef = check<N, -1>(), eb = check<N, +1>();
  printf("N=%d fwd_err=%.3e bwd_err=%.3e\n", N, ef, eb);
  return (ef < 1e-13 && eb < 1e-13) ? 0 : 1;
}

int main()
{
  int bad = 0;
  bad += run<2>();
  bad += run<4>();
  bad += run<8>();
  bad += run<16>();
  bad += run<32>();
  bad += run<64>();
  bad += run<128>();
  bad += run<256>();
  bad += run<512>();
  bad += run<1024>();
  printf(bad ? "FAIL\n" : "OK\n");
  return bad;
}
