// param_cache.h — the worker-side parameter + gradient cache.
// ref: /root/reference/src/core/parameter/global_param_cache.h:28-118 (GlobalParamCache).
// Two hash maps (params, grads) under one RWLock; init_keys / init_key(random); text dump.
// The GPU form of this structure is the dense [n_unique, D] cache filled by ssb_pull plus the
// inverse index from ssb_dedup (swiftsnails_b200/models/word2vec.py).
#pragma once
#include <limits>
#include <set>
#include "../utils/all.h"
namespace ssb {

template <typename Key, typename Val, typename Grad>
class GlobalParamCache : public VirtualObject {
 public:
  typedef Key key_t;
  typedef Val val_t;
  typedef Grad grad_t;
  typedef DenseHashMap<Key, Val> param_cache_t;
  typedef DenseHashMap<Key, Grad> grad_cache_t;

  GlobalParamCache() {
    params_.set_empty_key(std::numeric_limits<key_t>::max());
    grads_.set_empty_key(std::numeric_limits<key_t>::max());
  }
  template <typename Container>
  void init_keys(const Container& keys) {
    rwlock_write_guard g(lock_);
    for (const auto& k : keys) { params_[k]; grads_[k]; }
  }
  // ref: global_param_cache.h:50-56 — Val::init(bool random)
  void init_key(const key_t& key, bool random = false) {
    rwlock_write_guard g(lock_);
    params_[key];
    grads_[key];
    (void)random;
  }
  size_t size() { rwlock_read_guard g(lock_); return params_.size(); }
  void clear() { rwlock_write_guard g(lock_); params_.clear(); grads_.clear(); }
  RWLock& rwlock() { return lock_; }
  param_cache_t& params() { return params_; }
  grad_cache_t& grads() { return grads_; }
  friend std::ostream& operator<<(std::ostream& os, GlobalParamCache& c) {
    rwlock_read_guard g(c.lock_);
    c.params_.for_each([&os](const key_t& k, const val_t& v) { os << k << "\t" << v << "\n"; });
    return os;
  }

 private:
  RWLock lock_;
  param_cache_t params_;
  grad_cache_t grads_;
};

}  // namespace ssb
