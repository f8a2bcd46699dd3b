// access_method.h — the user hooks that define what a server does on pull and on push, plus
// the stock SGD / AdaGrad implementations.
// ref: /root/reference/src/core/parameter/sparse_access_method.h:10-48 (PullAccessMethod /
// PushAccessMethod). The reference leaves SGD/AdaGrad to the (absent) apps; they are provided
// here because the device kernels (csrc/kernels/table.cu push_kernel) implement the same two
// update rules and the CPU path is their oracle.
#pragma once
#include <cmath>
#include "../utils/all.h"
namespace ssb {

template <typename Table, typename PullVal, typename PullParam>
class PullAccessMethod {
 public:
  typedef Table table_t;
  typedef typename Table::key_t key_t;
  typedef typename Table::value_t value_t;
  typedef PullVal pull_val_t;      // what travels to the worker
  typedef PullParam pull_param_t;  // what the server stores
  virtual ~PullAccessMethod() = default;
  // lazily create the server-side parameter of a key seen for the first time
  virtual void init_param(const key_t& key, pull_param_t& param) = 0;
  // fill the value returned to the worker from the stored parameter
  virtual void get_pull_value(const key_t& key, const pull_param_t& param, pull_val_t& val) = 0;
};

template <typename Table, typename PushVal, typename PushParam>
class PushAccessMethod {
 public:
  typedef Table table_t;
  typedef typename Table::key_t key_t;
  typedef typename Table::value_t value_t;
  typedef PushVal push_val_t;
  typedef PushParam push_param_t;
  virtual ~PushAccessMethod() = default;
  virtual void merge_push_value(const key_t& key, push_val_t& push_val, const push_val_t& other) = 0;
  // apply one pushed gradient to the stored parameter (SGD, AdaGrad, …)
  virtual void apply_push_value(const key_t& key, push_param_t& param, const push_val_t& push_val) = 0;
};

// ---- stock row types ---------------------------------------------------------------------------
// Server-side parameter: the row plus (optionally) its AdaGrad accumulator. Only `val` is sent
// to workers; the accumulator stays on the server (the Param-vs-Val split of the reference).
struct RowParam {
  VecF val;
  VecF acc;  // empty for SGD
  friend std::ostream& operator<<(std::ostream& os, const RowParam& p) { return os << p.val; }
};

// Deterministic init shared with the device table: element j of key's row.
inline float row_init_value(uint32_t key, uint32_t j, uint64_t seed, float scale) {
  uint64_t h = get_hash_code((((uint64_t)key << 32) | (uint64_t)j) ^ seed);
  float u = (float)(uint32_t)(h >> 40) * (1.0f / 16777216.0f);
  return (u - 0.5f) * scale;
}

template <typename Table>
class RowPullMethod : public PullAccessMethod<Table, VecF, RowParam> {
 public:
  RowPullMethod(int dim, bool with_acc, uint64_t seed, float init_scale)
      : dim_(dim), with_acc_(with_acc), seed_(seed), scale_(init_scale) {}
  void init_param(const typename Table::key_t& key, RowParam& param) override {
    param.val.init(dim_);
    for (int j = 0; j < dim_; ++j) param.val[j] = row_init_value((uint32_t)key, (uint32_t)j, seed_, scale_);
    if (with_acc_) param.acc.init(dim_);
  }
  void get_pull_value(const typename Table::key_t&, const RowParam& param, VecF& val) override { val = param.val; }

 private:
  int dim_;
  bool with_acc_;
  uint64_t seed_;
  float scale_;
};

template <typename Table>
class SgdPushMethod : public PushAccessMethod<Table, VecF, RowParam> {
 public:
  explicit SgdPushMethod(float lr) : lr_(lr) {}
  void merge_push_value(const typename Table::key_t&, VecF& a, const VecF& b) override { a += b; }
  void apply_push_value(const typename Table::key_t&, RowParam& param, const VecF& grad) override {
    const size_t n = param.val.size();
    for (size_t j = 0; j < n; ++j) param.val[j] -= lr_ * grad[j];
  }

 private:
  float lr_;
};

template <typename Table>
class AdaGradPushMethod : public PushAccessMethod<Table, VecF, RowParam> {
 public:
  AdaGradPushMethod(float lr, float eps) : lr_(lr), eps_(eps) {}
  void merge_push_value(const typename Table::key_t&, VecF& a, const VecF& b) override { a += b; }
  void apply_push_value(const typename Table::key_t&, RowParam& param, const VecF& grad) override {
    const size_t n = param.val.size();
    if (param.acc.size() != n) param.acc.init(n);
    for (size_t j = 0; j < n; ++j) {
      float g = grad[j];
      param.acc[j] += g * g;
      param.val[j] -= lr_ * g / std::sqrt(param.acc[j] + eps_);
    }
  }

 private:
  float lr_, eps_;
};

}  // namespace ssb
