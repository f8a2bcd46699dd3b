// hashmap.h — open-addressing hash maps with an explicit empty key.
// ref: /root/reference/src/utils/hashmap.h:39-87 (wrappers over google sparsehash) and the
// direct dense_hash_map uses in sparsetable.h / global_param_cache.h. sparsehash is not on
// this image; DenseHashMap below is an own flat linear-probing table with the same contract
// (set_empty_key, find/insert/operator[], iteration, rehash invalidates references).
#pragma once
#include <cstddef>
#include <cstdint>
#include <functional>
#include <limits>
#include <utility>
#include <vector>
#include "hash.h"
namespace ssb {

template <typename K, typename V, typename Hash = std::hash<K>>
class DenseHashMap {
 public:
  typedef std::pair<K, V> value_type;
  explicit DenseHashMap(size_t initial = 32) { rebuild(round_up(initial)); }
  void set_empty_key(const K& k) {
    empty_ = k;
    for (auto& s : slots_) s.first = empty_;
    size_ = 0;
  }
  const K& empty_key() const { return empty_; }
  size_t size() const { return size_; }
  bool empty() const { return size_ == 0; }
  size_t bucket_count() const { return slots_.size(); }

  V* find(const K& k) {
    size_t i = probe(k);
    return slots_[i].first == k ? &slots_[i].second : nullptr;
  }
  const V* find(const K& k) const { return const_cast<DenseHashMap*>(this)->find(k); }
  bool count(const K& k) const { return find(k) != nullptr; }
  V& operator[](const K& k) {
    size_t i = probe(k);
    if (slots_[i].first == k) return slots_[i].second;
    if ((size_ + 1) * 10 > slots_.size() * 7) {
      rebuild(slots_.size() * 2);
      i = probe(k);
    }
    slots_[i].first = k;
    slots_[i].second = V();
    ++size_;
    return slots_[i].second;
  }
  // returns true if inserted (false: key existed, value untouched — like map::insert)
  bool insert(const K& k, V v) {
    size_t before = size_;
    V& ref = (*this)[k];
    if (size_ != before) { ref = std::move(v); return true; }
    return false;
  }
  void clear() {
    for (auto& s : slots_) { s.first = empty_; s.second = V(); }
    size_ = 0;
  }
  template <typename F>
  void for_each(F f) {
    for (auto& s : slots_)
      if (!(s.first == empty_)) f(s.first, s.second);
  }
  template <typename F>
  void for_each(F f) const {
    for (auto& s : slots_)
      if (!(s.first == empty_)) f(s.first, s.second);
  }

 private:
  static size_t round_up(size_t n) { size_t p = 8; while (p < n) p <<= 1; return p; }
  size_t probe(const K& k) const {
    size_t mask = slots_.size() - 1;
    size_t i = (size_t)get_hash_code((uint64_t)Hash()(k)) & mask;
    while (!(slots_[i].first == k) && !(slots_[i].first == empty_)) i = (i + 1) & mask;
    return i;
  }
  void rebuild(size_t n) {
    std::vector<value_type> old;
    old.swap(slots_);
    slots_.assign(n, value_type(empty_, V()));
    size_ = 0;
    for (auto& s : old)
      if (!(s.first == empty_)) {
        size_t i = probe(s.first);
        slots_[i] = std::move(s);
        ++size_;
      }
  }
  std::vector<value_type> slots_;
  K empty_ = std::numeric_limits<K>::max();
  size_t size_ = 0;
};

// The reference has a "sparse" (memory-lean) and a "dense" (fast) wrapper; both map onto the
// same flat table here.
template <typename K, typename V>
using SparseHashMap = DenseHashMap<K, V>;

}  // namespace ssb
