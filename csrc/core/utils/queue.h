// queue.h — threadsafe_queue and the bounded producer/consumer queue.
// ref: /root/reference/src/utils/queue.h:15-111.
#pragma once
#include <condition_variable>
#include <deque>
#include <memory>
#include <mutex>
namespace ssb {

template <typename T>
class threadsafe_queue {
 public:
  void push(T v) {
    { std::lock_guard<std::mutex> lk(m_); q_.push_back(std::move(v)); }
    cv_.notify_one();
  }
  void wait_and_pop(T& out) {
    std::unique_lock<std::mutex> lk(m_);
    cv_.wait(lk, [this] { return !q_.empty(); });
    out = std::move(q_.front());
    q_.pop_front();
  }
  std::shared_ptr<T> wait_and_pop() {
    std::unique_lock<std::mutex> lk(m_);
    cv_.wait(lk, [this] { return !q_.empty(); });
    auto r = std::make_shared<T>(std::move(q_.front()));
    q_.pop_front();
    return r;
  }
  bool try_pop(T& out) {
    std::lock_guard<std::mutex> lk(m_);
    if (q_.empty()) return false;
    out = std::move(q_.front());
    q_.pop_front();
    return true;
  }
  bool empty() const { std::lock_guard<std::mutex> lk(m_); return q_.empty(); }
  size_t size() const { std::lock_guard<std::mutex> lk(m_); return q_.size(); }

 private:
  mutable std::mutex m_;
  std::condition_variable cv_;
  std::deque<T> q_;
};

// Bounded queue for data pipelines. end_input(n, poison) pushes n poison values so n consumers
// terminate (ref: queue.h:76-111).
template <typename T>
class queue_with_capacity {
 public:
  explicit queue_with_capacity(size_t capacity = 1024) : cap_(capacity) {}
  void push(T v) {
    std::unique_lock<std::mutex> lk(m_);
    not_full_.wait(lk, [this] { return q_.size() < cap_; });
    q_.push_back(std::move(v));
    lk.unlock();
    not_empty_.notify_one();
  }
  void wait_and_pop(T& out) {
    std::unique_lock<std::mutex> lk(m_);
    not_empty_.wait(lk, [this] { return !q_.empty(); });
    out = std::move(q_.front());
    q_.pop_front();
    lk.unlock();
    not_full_.notify_one();
  }
  void end_input(int n_consumers, const T& poison) {
    for (int i = 0; i < n_consumers; ++i) push(poison);
  }
  size_t size() const { std::lock_guard<std::mutex> lk(m_); return q_.size(); }
  size_t capacity() const { return cap_; }

 private:
  mutable std::mutex m_;
  std::condition_variable not_full_, not_empty_;
  std::deque<T> q_;
  size_t cap_;
};

}  // namespace ssb
