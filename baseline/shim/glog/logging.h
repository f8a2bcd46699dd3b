// Minimal glog stand-in so the UNMODIFIED reference headers compile on this image (glog is not
// installed). Only what the reference uses: LOG, DLOG (no-op: the reference
// defines NDEBUG), CHECK family, PCHECK, CHECK_NOTNULL.
#pragma once
#include <cerrno>
#include <functional>
#include <algorithm>
#include <limits>
#include <climits>
#include <unistd.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <sstream>
namespace ssb_glog_shim {
enum { INFO = 0, WARNING = 1, ERROR = 2, FATAL = 3 };
inline int min_level() { static int l = std::getenv("GLOG_minloglevel") ? std::atoi(std::getenv("GLOG_minloglevel")) : 1; return l; }
class Msg {
 public:
  Msg(const char* f, int l, int sev, bool perr = false) : sev_(sev), perr_(perr), err_(errno) { os_ << "IWEF"[sev] << ' ' << f << ':' << l << "] "; }
  ~Msg() {
    if (perr_) os_ << ": " << std::strerror(err_);
    if (sev_ >= min_level() || sev_ == FATAL) { os_ << '\n'; std::string s = os_.str(); std::fwrite(s.data(), 1, s.size(), stderr); }
    if (sev_ == FATAL) std::abort();
  }
  std::ostream& stream() { return os_; }
 private:
  std::ostringstream os_; int sev_; bool perr_; int err_;
};
struct Voidify { void operator&(std::ostream&) {} };
template <typename T> T* check_notnull(const char* f, int l, const char* expr, T* p) { if (!p) { Msg(f, l, FATAL).stream() << "'" << expr << "' must be non NULL"; } return p; }
}  // namespace ssb_glog_shim
#define LOG(sev) ::ssb_glog_shim::Msg(__FILE__, __LINE__, ::ssb_glog_shim::sev).stream()
#define DLOG(sev) true ? (void)0 : ::ssb_glog_shim::Voidify() & LOG(sev)
#define CHECK(c) (c) ? (void)0 : ::ssb_glog_shim::Voidify() & ::ssb_glog_shim::Msg(__FILE__, __LINE__, ::ssb_glog_shim::FATAL).stream() << "Check failed: " #c " "
#define PCHECK(c) (c) ? (void)0 : ::ssb_glog_shim::Voidify() & ::ssb_glog_shim::Msg(__FILE__, __LINE__, ::ssb_glog_shim::FATAL, true).stream() << "Check failed: " #c " "
#define SSB_SHIM_OP(a, b, op) ((a)op(b)) ? (void)0 : ::ssb_glog_shim::Voidify() & ::ssb_glog_shim::Msg(__FILE__, __LINE__, ::ssb_glog_shim::FATAL).stream() << "Check failed: " #a " " #op " " #b " "
#define CHECK_EQ(a, b) SSB_SHIM_OP(a, b, ==)
#define CHECK_NE(a, b) SSB_SHIM_OP(a, b, !=)
#define CHECK_GT(a, b) SSB_SHIM_OP(a, b, >)
#define CHECK_GE(a, b) SSB_SHIM_OP(a, b, >=)
#define CHECK_LT(a, b) SSB_SHIM_OP(a, b, <)
#define CHECK_LE(a, b) SSB_SHIM_OP(a, b, <=)
#define CHECK_NOTNULL(p) ::ssb_glog_shim::check_notnull(__FILE__, __LINE__, #p, (p))
