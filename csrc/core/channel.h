// channel.h — closable MPMC task queue. ref: /root/reference/src/core/BasicChannel.h:37-112.
// pop() returns false once the channel is closed AND drained; close() wakes every waiter.
#pragma once
#include <condition_variable>
#include <deque>
#include <mutex>
namespace ssb {

template <typename T>
class BasicChannel {
 public:
  bool push(T v) {
    {
      std::lock_guard<std::mutex> lk(m_);
      if (closed_) return false;
      q_.push_back(std::move(v));
    }
    cv_.notify_one();
    return true;
  }
  bool pop(T& out) {
    std::unique_lock<std::mutex> lk(m_);
    cv_.wait(lk, [this] { return closed_ || !q_.empty(); });
    if (q_.empty()) return false;
    out = std::move(q_.front());
    q_.pop_front();
    return true;
  }
  void close() {
    { std::lock_guard<std::mutex> lk(m_); closed_ = true; }
    cv_.notify_all();
  }
  bool closed() const { std::lock_guard<std::mutex> lk(m_); return closed_; }
  size_t size() const { std::lock_guard<std::mutex> lk(m_); return q_.size(); }

 private:
  mutable std::mutex m_;
  std::condition_variable cv_;
  std::deque<T> q_;
  bool closed_ = false;
};

}  // namespace ssb
