// timer.h — wall-clock timer. ref: /root/reference/src/utils/Timer.h:14-44 (seconds
// resolution there; microsecond resolution here, same start/stop/timeout surface).
#pragma once
#include <chrono>
namespace ssb {
class Timer {
 public:
  Timer() { start(); }
  void start() { t0_ = clock::now(); running_ = true; }
  double stop() { t1_ = clock::now(); running_ = false; return elapsed(); }
  // seconds since start() (or between start() and stop())
  double elapsed() const {
    auto end = running_ ? clock::now() : t1_;
    return std::chrono::duration<double>(end - t0_).count();
  }
  double elapsed_ms() const { return elapsed() * 1e3; }
  void set_time_span(double seconds) { span_ = seconds; }
  bool timeout() const { return span_ > 0 && elapsed() >= span_; }

 private:
  typedef std::chrono::steady_clock clock;
  clock::time_point t0_, t1_;
  bool running_ = false;
  double span_ = 0;
};
}  // namespace ssb
