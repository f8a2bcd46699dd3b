// sparsetable.h — the CPU sharded sparse hash table and the pull / push access agents.
// ref: /root/reference/src/core/parameter/sparsetable.h:5-204.
// This is the CPU fallback data plane (BASELINE config 1) and the oracle for the HBM table
// (csrc/kernels/table.cu). shard = fmix64(key) % shard_num (sparsetable.h:115); each shard is
// a flat hash map + RWLock; text dump is `key \t value \n` per entry, shard by shard
// (sparsetable.h:49-56,100-104).
// Differences by design: apply() runs the update under the shard's write lock, so no pointer
// into the map ever escapes a lock (the reference returns one and mutates it lock-free, D5).
#pragma once
#include <limits>
#include "access_method.h"
namespace ssb {

template <typename Key, typename Value>
class alignas(64) SparseTableShard : public VirtualObject {
 public:
  typedef Key key_t;
  typedef Value value_t;
  typedef DenseHashMap<Key, Value> map_t;

  SparseTableShard() { data_.set_empty_key(std::numeric_limits<key_t>::max()); }

  bool find(const key_t& key, value_t& val) {
    rwlock_read_guard g(lock_);
    const value_t* p = data_.find(key);
    if (!p) return false;
    val = *p;
    return true;
  }
  void assign(const key_t& key, const value_t& val) {
    rwlock_write_guard g(lock_);
    data_[key] = val;
  }
  // find-or-create then mutate, all under the write lock
  template <typename Init, typename Fn>
  void upsert(const key_t& key, Init init, Fn fn) {
    rwlock_write_guard g(lock_);
    size_t before = data_.size();
    value_t& v = data_[key];
    if (data_.size() != before) init(v);
    fn(v);
  }
  // mutate an existing entry; returns false if the key is absent
  template <typename Fn>
  bool apply(const key_t& key, Fn fn) {
    rwlock_write_guard g(lock_);
    value_t* p = data_.find(key);
    if (!p) return false;
    fn(*p);
    return true;
  }
  size_t size() {
    rwlock_read_guard g(lock_);
    return data_.size();
  }
  void set_shard_id(int id) { shard_id_ = id; }
  int shard_id() const { return shard_id_; }
  // text dump. ref: sparsetable.h:49-56
  void output(std::ostream& os) {
    rwlock_read_guard g(lock_);
    data_.for_each([&os](const key_t& k, const value_t& v) { os << k << "\t" << v << "\n"; });
  }
  template <typename F>
  void for_each(F f) {
    rwlock_read_guard g(lock_);
    data_.for_each(f);
  }

 private:
  map_t data_;
  RWLock lock_;
  int shard_id_ = -1;
};

template <typename Key, typename Value>
class SparseTable : public VirtualObject {
 public:
  typedef Key key_t;
  typedef Value value_t;
  typedef SparseTableShard<Key, Value> shard_t;

  // shard_num from the config. ref: sparsetable.h:77
  SparseTable() : SparseTable(global_config().get_config("shard_num", "8").to_int32()) {}
  explicit SparseTable(int shard_num) : shard_num_(shard_num) {
    CHECK_GT(shard_num_, 0);
    shards_.reset(new shard_t[shard_num_]);
    for (int i = 0; i < shard_num_; ++i) shards_[i].set_shard_id(i);
  }
  int to_shard_id(const key_t& key) const { return (int)(get_hash_code((uint64_t)key) % (uint64_t)shard_num_); }
  shard_t& shard(int i) { return shards_[i]; }
  shard_t& shard_of(const key_t& key) { return shards_[to_shard_id(key)]; }
  bool find(const key_t& key, value_t& val) { return shard_of(key).find(key, val); }
  void assign(const key_t& key, const value_t& val) { shard_of(key).assign(key, val); }
  int shard_num() const { return shard_num_; }
  size_t size() {
    size_t n = 0;
    for (int i = 0; i < shard_num_; ++i) n += shards_[i].size();
    return n;
  }
  // dump every shard. ref: sparsetable.h:100-104 (to stdout there; any stream here)
  void output(std::ostream& os = std::cout) {
    for (int i = 0; i < shard_num_; ++i) shards_[i].output(os);
    os.flush();
  }

 private:
  int shard_num_;
  std::unique_ptr<shard_t[]> shards_;
};

// Server-side pull: lazily create, then produce the value sent to the worker.
// ref: sparsetable.h:123-160
template <typename Table, typename AccessMethod>
class PullAccessAgent {
 public:
  typedef typename Table::key_t key_t;
  typedef typename AccessMethod::pull_val_t pull_val_t;
  typedef typename AccessMethod::pull_param_t pull_param_t;
  PullAccessAgent(Table& table, AccessMethod& method) : table_(table), method_(method) {}
  void get_pull_value(const key_t& key, pull_val_t& val) {
    table_.shard_of(key).upsert(
        key, [&](pull_param_t& p) { method_.init_param(key, p); },
        [&](pull_param_t& p) { method_.get_pull_value(key, p, val); });
  }

 private:
  Table& table_;
  AccessMethod& method_;
};

// Server-side push: apply the pushed gradient to the stored parameter.
// ref: sparsetable.h:162-198
template <typename Table, typename AccessMethod>
class PushAccessAgent {
 public:
  typedef typename Table::key_t key_t;
  typedef typename AccessMethod::push_val_t push_val_t;
  typedef typename AccessMethod::push_param_t push_param_t;
  PushAccessAgent(Table& table, AccessMethod& method) : table_(table), method_(method) {}
  void merge_push_value(const key_t& key, push_val_t& a, const push_val_t& b) { method_.merge_push_value(key, a, b); }
  // returns false when the key was never pulled ("new key should be inited before",
  // sparsetable.h:184 aborts; here the caller decides)
  bool apply_push_value(const key_t& key, const push_val_t& grad) {
    return table_.shard_of(key).apply(key, [&](push_param_t& p) { method_.apply_push_value(key, p, grad); });
  }

 private:
  Table& table_;
  AccessMethod& method_;
};

template <typename Key, typename Value>
inline SparseTable<Key, Value>& global_sparse_table() {
  static SparseTable<Key, Value> table;
  return table;
}

template <typename Table, typename AccessMethod>
inline std::unique_ptr<PullAccessAgent<Table, AccessMethod>> make_pull_access(Table& t, AccessMethod& m) {
  return std::unique_ptr<PullAccessAgent<Table, AccessMethod>>(new PullAccessAgent<Table, AccessMethod>(t, m));
}
template <typename Table, typename AccessMethod>
inline std::unique_ptr<PushAccessAgent<Table, AccessMethod>> make_push_access(Table& t, AccessMethod& m) {
  return std::unique_ptr<PushAccessAgent<Table, AccessMethod>>(new PushAccessAgent<Table, AccessMethod>(t, m));
}

}  // namespace ssb
