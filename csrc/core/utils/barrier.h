// barrier.h — Barrier, CompBarrier, StateBarrier, CounterBarrier.
// ref: /root/reference/src/utils/Barrier.h:14-135.
// StateBarrier is the framework's only failure detector (start-up timeouts). Differences by
// design: the state flag is set and notified under the mutex (the reference's lost-wakeup
// window, SURVEY.md D7, is gone) and time limits use cv.wait_for instead of a detached thread
// that sleeps the whole timeout.
#pragma once
#include <chrono>
#include <condition_variable>
#include <functional>
#include <mutex>
namespace ssb {

// Gate: block() until unblock() is called `n` times... here: simple open/close gate.
class Barrier {
 public:
  void block() {
    std::unique_lock<std::mutex> lk(m_);
    cv_.wait(lk, [this] { return open_; });
  }
  void unblock() {
    { std::lock_guard<std::mutex> lk(m_); open_ = true; }
    cv_.notify_all();
  }
  void reset() { std::lock_guard<std::mutex> lk(m_); open_ = false; }

 private:
  std::mutex m_;
  std::condition_variable cv_;
  bool open_ = false;
};

// Completion barrier for a fixed party of threads (reusable).
class CompBarrier {
 public:
  explicit CompBarrier(int parties) : parties_(parties) {}
  void wait() {
    std::unique_lock<std::mutex> lk(m_);
    int gen = gen_;
    if (++arrived_ == parties_) {
      arrived_ = 0;
      ++gen_;
      cv_.notify_all();
    } else {
      cv_.wait(lk, [&] { return gen != gen_; });
    }
  }

 private:
  std::mutex m_;
  std::condition_variable cv_;
  int parties_, arrived_ = 0, gen_ = 0;
};

class StateBarrier {
 public:
  void set_state_valid() {
    { std::lock_guard<std::mutex> lk(m_); valid_ = true; }
    cv_.notify_all();
  }
  void try_unblock() { cv_.notify_all(); }
  void unblock() { set_state_valid(); }
  void reset() { std::lock_guard<std::mutex> lk(m_); valid_ = false; }
  bool state_valid() { std::lock_guard<std::mutex> lk(m_); return valid_; }
  // Blocks until the state is valid. If a time limit was set and expires first, the timeout
  // handler runs (once) and block() returns false.
  bool block() {
    std::unique_lock<std::mutex> lk(m_);
    if (limit_ms_ <= 0) {
      cv_.wait(lk, [this] { return valid_; });
      return true;
    }
    bool ok = cv_.wait_for(lk, std::chrono::milliseconds(limit_ms_), [this] { return valid_; });
    if (!ok) {
      auto h = handler_;
      lk.unlock();
      if (h) h();
    }
    return ok;
  }
  void time_limit(long ms, std::function<void()> timeout_handler) {
    std::lock_guard<std::mutex> lk(m_);
    limit_ms_ = ms;
    handler_ = std::move(timeout_handler);
  }

 private:
  std::mutex m_;
  std::condition_variable cv_;
  bool valid_ = false;
  long limit_ms_ = 0;
  std::function<void()> handler_;
};

// Counts down from n; block() returns when the count reaches zero.
class CounterBarrier {
 public:
  explicit CounterBarrier(int n = 0) : count_(n) {}
  void reset(int n) { std::lock_guard<std::mutex> lk(m_); count_ = n; }
  void count_down() {
    std::lock_guard<std::mutex> lk(m_);
    if (--count_ <= 0) cv_.notify_all();
  }
  bool block(long timeout_ms = 0) {
    std::unique_lock<std::mutex> lk(m_);
    if (timeout_ms <= 0) { cv_.wait(lk, [this] { return count_ <= 0; }); return true; }
    return cv_.wait_for(lk, std::chrono::milliseconds(timeout_ms), [this] { return count_ <= 0; });
  }

 private:
  std::mutex m_;
  std::condition_variable cv_;
  int count_;
};

}  // namespace ssb
