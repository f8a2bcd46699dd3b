// locks.h — SpinLock and RWLock with scoped guards.
// ref: /root/reference/src/utils/SpinLock.h:42-53, utils/RWLock.h:14-49.
// RWLock wraps std::shared_mutex (the reference wraps pthread_rwlock_t by hand).
#pragma once
#include <atomic>
#include <mutex>
#include <shared_mutex>
#include <thread>
namespace ssb {

class SpinLock {
 public:
  void lock() {
    int spins = 0;
    while (flag_.test_and_set(std::memory_order_acquire))
      if (++spins > 64) { std::this_thread::yield(); spins = 0; }
  }
  bool try_lock() { return !flag_.test_and_set(std::memory_order_acquire); }
  void unlock() { flag_.clear(std::memory_order_release); }

 private:
  std::atomic_flag flag_ = ATOMIC_FLAG_INIT;
};

class RWLock {
 public:
  void rdlock() { m_.lock_shared(); }
  void wrlock() { m_.lock(); }
  void rdunlock() { m_.unlock_shared(); }
  void wrunlock() { m_.unlock(); }
  std::shared_mutex& native() { return m_; }

 private:
  std::shared_mutex m_;
};

class rwlock_read_guard {
 public:
  explicit rwlock_read_guard(RWLock& l) : l_(l) { l_.rdlock(); }
  ~rwlock_read_guard() { l_.rdunlock(); }
  rwlock_read_guard(const rwlock_read_guard&) = delete;

 private:
  RWLock& l_;
};
class rwlock_write_guard {
 public:
  explicit rwlock_write_guard(RWLock& l) : l_(l) { l_.wrlock(); }
  ~rwlock_write_guard() { l_.wrunlock(); }
  rwlock_write_guard(const rwlock_write_guard&) = delete;

 private:
  RWLock& l_;
};

}  // namespace ssb
