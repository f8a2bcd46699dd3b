// vec.h — dense vector math for CPU-side rows (the reference apps' embedding-row type).
// ref: /root/reference/src/utils/vec1.h:6-252 (Vec of double). Here the element type is a
// template parameter (float rows match the HBM table; double keeps the reference's precision
// for oracles). Same operations: init(size, random), dot, element-wise + - * / with scalar or
// Vec, outer, sqrt, randInit = (rand/RAND_MAX - 0.5)/size, text form "Vec:\t v0 v1 … ".
#pragma once
#include <cmath>
#include <cstdlib>
#include <iostream>
#include <vector>
#include "buffer.h"
#include "common.h"
namespace ssb {

template <typename T>
class VecT {
 public:
  typedef T value_type;
  VecT() = default;
  explicit VecT(size_t size, bool random = false) { init(size, random); }
  VecT(std::initializer_list<T> l) : d_(l) {}

  void init(size_t size, bool random = false) {
    d_.assign(size, T(0));
    if (random) randInit();
  }
  // ref: vec1.h:223-226
  void randInit() {
    for (auto& x : d_) x = (T)((std::rand() / (double)RAND_MAX - 0.5) / (double)d_.size());
  }
  size_t size() const { return d_.size(); }
  bool empty() const { return d_.empty(); }
  T& operator[](size_t i) { return d_[i]; }
  const T& operator[](size_t i) const { return d_[i]; }
  T* data() { return d_.data(); }
  const T* data() const { return d_.data(); }
  void reset() { std::fill(d_.begin(), d_.end(), T(0)); }  // Grad::reset() contract

  T dot(const VecT& o) const {
    CHECK_EQ(size(), o.size());
    T s = 0;
    for (size_t i = 0; i < d_.size(); ++i) s += d_[i] * o.d_[i];
    return s;
  }
  T sum() const { T s = 0; for (auto x : d_) s += x; return s; }
  T norm2() const { return std::sqrt(dot(*this)); }
  // outer product as a row-major matrix
  std::vector<VecT> outer(const VecT& o) const {
    std::vector<VecT> m(size(), VecT(o.size()));
    for (size_t i = 0; i < size(); ++i)
      for (size_t j = 0; j < o.size(); ++j) m[i][j] = d_[i] * o.d_[j];
    return m;
  }

#define SSB_VEC_OP(op)                                                             \
  VecT& operator op##=(const VecT& o) {                                            \
    CHECK_EQ(size(), o.size());                                                    \
    for (size_t i = 0; i < d_.size(); ++i) d_[i] op## = o.d_[i];                   \
    return *this;                                                                  \
  }                                                                                \
  VecT& operator op##=(T s) {                                                      \
    for (auto& x : d_) x op## = s;                                                 \
    return *this;                                                                  \
  }                                                                                \
  friend VecT operator op(VecT a, const VecT& b) { a op## = b; return a; }         \
  friend VecT operator op(VecT a, T s) { a op## = s; return a; }
  SSB_VEC_OP(+)
  SSB_VEC_OP(-)
  SSB_VEC_OP(*)
  SSB_VEC_OP(/)
#undef SSB_VEC_OP
  friend VecT operator*(T s, VecT a) { a *= s; return a; }
  friend VecT operator-(VecT a) { for (auto& x : a.d_) x = -x; return a; }

  friend VecT sqrt(VecT a) { for (auto& x : a.d_) x = std::sqrt(x); return a; }

  // text form, ref: vec1.h:106-121 ("Vec:\t" then each element followed by a space)
  friend std::ostream& operator<<(std::ostream& os, const VecT& v) {
    os << "Vec:\t";
    for (auto x : v.d_) os << x << " ";
    return os;
  }
  // wire form: uint32 size + raw elements
  friend BinaryBuffer& operator<<(BinaryBuffer& bb, const VecT& v) {
    bb << (uint32_t)v.size();
    bb.put_array(v.d_.data(), v.d_.size());
    return bb;
  }
  friend BinaryBuffer& operator>>(BinaryBuffer& bb, VecT& v) {
    uint32_t n = 0;
    bb >> n;
    v.d_.resize(n);
    bb.get_array(v.d_.data(), n);
    return bb;
  }

 private:
  std::vector<T> d_;
};

typedef VecT<double> Vec;    // the reference's Vec
typedef VecT<float> VecF;

}  // namespace ssb
