// Stand-in for Dune::Timer (<dune/common/timer.hh>).
#ifndef PF_STANDIN_DUNE_COMMON_TIMER_HH
#define PF_STANDIN_DUNE_COMMON_TIMER_HH
#include <chrono>
namespace Dune {
class Timer
{
  using clock = std::chrono::steady_clock;

public:
  explicit Timer(bool startImmediately = true) { if (startImmediately) start(); }
  void reset() { _sum = 0.; _start = clock::now(); }
  void start() { _running = true; _start = clock::now(); }
  double stop() { if (_running) _sum += since(); _running = false; return _sum; }
  double elapsed() const { return _running ? _sum + since() : _sum; }
  double lastElapsed() const { return since(); }

private:
  double since() const
  {
    return std::chrono::duration<double>(clock::now() - _start).count();
  }
  clock::time_point _start{};
  double _sum = 0.;
  bool _running = false;
};
} // namespace Dune
#endif
