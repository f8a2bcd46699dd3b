// string_util.h — trim / split / key_value_split / headswith / format_string / LineFileReader.
// ref: /root/reference/src/utils/string.h:14-114.
#pragma once
#include <cstdarg>
#include <cstdio>
#include <string>
#include <utility>
#include <vector>
#include "common.h"
namespace ssb {

inline std::string trim(const std::string& s, const std::string& ws = " \t\r\n") {
  size_t b = s.find_first_not_of(ws);
  if (b == std::string::npos) return "";
  size_t e = s.find_last_not_of(ws);
  return s.substr(b, e - b + 1);
}

// Split on any character of `delims`; empty fields are dropped.
inline std::vector<std::string> split(const std::string& s, const std::string& delims) {
  std::vector<std::string> out;
  size_t i = 0;
  while (i < s.size()) {
    size_t b = s.find_first_not_of(delims, i);
    if (b == std::string::npos) break;
    size_t e = s.find_first_of(delims, b);
    if (e == std::string::npos) e = s.size();
    out.emplace_back(s.substr(b, e - b));
    i = e;
  }
  return out;
}

// Split at the FIRST delimiter only, so values such as tcp://host:port survive (string.h:48-61).
inline std::pair<std::string, std::string> key_value_split(const std::string& s, const std::string& delims) {
  size_t p = s.find_first_of(delims);
  if (p == std::string::npos) return {trim(s), ""};
  return {trim(s.substr(0, p)), trim(s.substr(p + 1))};
}

inline bool headswith(const std::string& s, const std::string& head) {
  return s.size() >= head.size() && s.compare(0, head.size(), head) == 0;
}

inline std::string format_string(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  va_list ap2;
  va_copy(ap2, ap);
  int n = std::vsnprintf(nullptr, 0, fmt, ap);
  va_end(ap);
  std::string out(n > 0 ? n : 0, '\0');
  if (n > 0) std::vsnprintf(&out[0], n + 1, fmt, ap2);
  va_end(ap2);
  return out;
}

// getdelim-based line reader that reuses one buffer (string.h:89-114).
class LineFileReader {
 public:
  LineFileReader() = default;
  LineFileReader(const LineFileReader&) = delete;
  ~LineFileReader() { std::free(buf_); }
  // Returns nullptr at EOF; the trailing delimiter is stripped.
  char* getline(FILE* f) { return getdelim(f, '\n'); }
  char* getdelim(FILE* f, char delim) {
    ssize_t n = ::getdelim(&buf_, &cap_, delim, f);
    if (n < 0) { len_ = 0; return nullptr; }
    if (n > 0 && buf_[n - 1] == delim) buf_[--n] = '\0';
    len_ = (size_t)n;
    return buf_;
  }
  char* get() const { return buf_; }
  size_t length() const { return len_; }

 private:
  char* buf_ = nullptr;
  size_t cap_ = 0, len_ = 0;
};

}  // namespace ssb
