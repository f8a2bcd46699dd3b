// buffer.h — growable byte buffer and the binary / text serialisers used on the wire.
// ref: /root/reference/src/utils/Buffer.h:15-324 (BasicBuffer, BinaryBuffer, TextBuffer).
// Same wire behaviour: host-endian, unframed, unaligned raw memcpy of each primitive
// (Buffer.h:200-232) so messages interoperate with the reference's formats.
// Differences by design: storage is a std::vector (no delete/new[] mismatch, D8), move leaves
// the source empty (D9), set() accepts zero bytes, bulk put/get for arrays of PODs.
#pragma once
#include <cstdint>
#include <cstring>
#include <sstream>
#include <string>
#include <type_traits>
#include <vector>
#include "common.h"
namespace ssb {

class BasicBuffer {
 public:
  BasicBuffer() { data_.reserve(1024); }  // ref: Buffer.h:158 initial capacity
  BasicBuffer(const BasicBuffer&) = default;
  BasicBuffer& operator=(const BasicBuffer&) = default;
  BasicBuffer(BasicBuffer&& o) noexcept : data_(std::move(o.data_)), cursor_(o.cursor_) { o.clear(); }
  BasicBuffer& operator=(BasicBuffer&& o) noexcept {
    if (this != &o) { data_ = std::move(o.data_); cursor_ = o.cursor_; o.clear(); }
    return *this;
  }
  // copy `size` bytes in and rewind the read cursor (ref: Buffer.h:30-36)
  void set(const char* buf, size_t size) {
    data_.assign(buf, buf + size);
    cursor_ = 0;
  }
  void clear() { data_.clear(); cursor_ = 0; }
  char* buffer() { return data_.data(); }
  const char* buffer() const { return data_.data(); }
  size_t size() const { return data_.size(); }
  size_t capacity() const { return data_.capacity(); }
  size_t status() const { return cursor_; }          // read position
  bool read_finished() const { return cursor_ >= data_.size(); }
  void reset_cursor() { cursor_ = 0; }
  void reserve(size_t n) { data_.reserve(n); }
  const char* cursor() const { return data_.data() + cursor_; }

 protected:
  void put_raw(const void* p, size_t n) {
    const char* c = static_cast<const char*>(p);
    data_.insert(data_.end(), c, c + n);
  }
  void get_raw(void* p, size_t n) {
    CHECK(cursor_ + n <= data_.size()) << "buffer underrun: need " << n << " have " << (data_.size() - cursor_);
    std::memcpy(p, data_.data() + cursor_, n);
    cursor_ += n;
  }
  std::vector<char> data_;
  size_t cursor_ = 0;
};

class BinaryBuffer : public BasicBuffer {
 public:
  template <typename T>
  typename std::enable_if<std::is_arithmetic<T>::value, BinaryBuffer&>::type operator<<(const T& v) {
    put_raw(&v, sizeof(T));
    return *this;
  }
  template <typename T>
  typename std::enable_if<std::is_arithmetic<T>::value, BinaryBuffer&>::type operator>>(T& v) {
    get_raw(&v, sizeof(T));
    return *this;
  }
  // bulk POD arrays (not in the reference: saves a call per element on the hot path)
  template <typename T>
  void put_array(const T* p, size_t n) {
    static_assert(std::is_trivially_copyable<T>::value, "POD only");
    put_raw(p, n * sizeof(T));
  }
  template <typename T>
  void get_array(T* p, size_t n) {
    static_assert(std::is_trivially_copyable<T>::value, "POD only");
    get_raw(p, n * sizeof(T));
  }
};

// Text form: whitespace-separated tokens. ref: Buffer.h:240-322 (unused by the reference's
// product code; kept for API parity, minus the stray debug print at Buffer.h:286).
class TextBuffer : public BasicBuffer {
 public:
  template <typename T>
  TextBuffer& operator<<(const T& v) {
    std::ostringstream os;
    os << v;
    std::string s = os.str();
    if (!data_.empty()) data_.push_back(' ');
    put_raw(s.data(), s.size());
    return *this;
  }
  template <typename T>
  TextBuffer& operator>>(T& v) {
    while (cursor_ < data_.size() && std::isspace((unsigned char)data_[cursor_])) ++cursor_;
    size_t b = cursor_;
    while (cursor_ < data_.size() && !std::isspace((unsigned char)data_[cursor_])) ++cursor_;
    std::istringstream is(std::string(data_.data() + b, cursor_ - b));
    is >> v;
    return *this;
  }
  std::string str() const { return std::string(data_.begin(), data_.end()); }
};

}  // namespace ssb
