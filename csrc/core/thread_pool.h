// thread_pool.h — typed pool: items of type T are pushed, Func processes them.
// ref: /root/reference/src/core/ThreadPool.h:14-71 (never used by the reference; kept for parity).
#pragma once
#include <thread>
#include <vector>
#include "channel.h"
namespace ssb {

template <typename T, typename Func>
class ThreadPool {
 public:
  ThreadPool(int thread_num, Func f) : f_(f) {
    for (int i = 0; i < thread_num; ++i)
      threads_.emplace_back([this] {
        T item;
        while (ch_.pop(item)) f_(item);
      });
  }
  ~ThreadPool() { join(); }
  bool push(T item) { return ch_.push(std::move(item)); }
  void join() {
    ch_.close();
    for (auto& t : threads_)
      if (t.joinable()) t.join();
  }

 private:
  Func f_;
  BasicChannel<T> ch_;
  std::vector<std::thread> threads_;
};

}  // namespace ssb
