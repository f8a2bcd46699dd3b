// daemon_thread.h — background loop thread. ref: /root/reference/src/utils/DaemonThread.h:6-20
// (detached loop-until-flag there; joinable with a stop flag here).
#pragma once
#include <atomic>
#include <chrono>
#include <functional>
#include <thread>
namespace ssb {
class DaemonThread {
 public:
  DaemonThread() = default;
  ~DaemonThread() { stop(); }
  // runs `body` every `period_ms` until stop()
  void start(std::function<void()> body, int period_ms = 0) {
    stop();
    stop_.store(false);
    th_ = std::thread([this, body, period_ms] {
      while (!stop_.load()) {
        body();
        if (period_ms > 0) std::this_thread::sleep_for(std::chrono::milliseconds(period_ms));
      }
    });
  }
  void stop() {
    stop_.store(true);
    if (th_.joinable()) th_.join();
  }
  bool running() const { return th_.joinable() && !stop_.load(); }

 private:
  std::atomic<bool> stop_{true};
  std::thread th_;
};
}  // namespace ssb
