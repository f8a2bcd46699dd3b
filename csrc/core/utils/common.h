// common.h — prelude of the host core: fixed-width key type, logging / CHECK macros,
// fork-safe popen.  ref: /root/reference/src/utils/common.h:1-76.
// Differences by design: no redefinition of int64_t/uint64_t (the reference's typedefs collide
// with glibc); logging is a ~60-line stream logger instead of glog; DLOG/DCHECK
// are active unless SSB_NDEBUG is defined.
#pragma once
#include <atomic>
#include <cerrno>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <sstream>
#include <string>
#include <thread>
#include <unordered_map>
#include <utility>
#include <vector>

namespace ssb {

typedef uint32_t index_t;  // ref: utils/common.h:45 — the only key type the reference supports
typedef std::function<void()> voidf_t;

// ---- logging -----------------------------------------------------------------------------
enum LogSeverity { LOG_INFO = 0, LOG_WARNING = 1, LOG_ERROR = 2, LOG_FATAL = 3 };

inline std::atomic<int>& min_log_level() {
  static std::atomic<int> lvl{[] {
    const char* e = std::getenv("SSB200_LOG_LEVEL");
    return e ? std::atoi(e) : (int)LOG_WARNING;
  }()};
  return lvl;
}

class LogMessage {
 public:
  LogMessage(const char* file, int line, int sev, bool with_errno = false)
      : sev_(sev), errno_(with_errno ? errno : 0), with_errno_(with_errno) {
    static const char tag[] = {'I', 'W', 'E', 'F'};
    const char* base = std::strrchr(file, '/');
    os_ << tag[sev] << ' ' << (base ? base + 1 : file) << ':' << line << "] ";
  }
  ~LogMessage() noexcept(false) {
    if (with_errno_) os_ << ": " << std::strerror(errno_) << " [" << errno_ << "]";
    if (sev_ >= min_log_level().load() || sev_ == LOG_FATAL) {
      os_ << '\n';
      std::string s = os_.str();
      std::fwrite(s.data(), 1, s.size(), stderr);
      std::fflush(stderr);
    }
    if (sev_ == LOG_FATAL) std::abort();
  }
  std::ostream& stream() { return os_; }

 private:
  std::ostringstream os_;
  int sev_, errno_;
  bool with_errno_;
};

struct LogVoidify {
  void operator&(std::ostream&) {}
};

#define SSB_LOG_INFO ::ssb::LogMessage(__FILE__, __LINE__, ::ssb::LOG_INFO).stream()
#define SSB_LOG_WARNING ::ssb::LogMessage(__FILE__, __LINE__, ::ssb::LOG_WARNING).stream()
#define SSB_LOG_ERROR ::ssb::LogMessage(__FILE__, __LINE__, ::ssb::LOG_ERROR).stream()
#define SSB_LOG_FATAL ::ssb::LogMessage(__FILE__, __LINE__, ::ssb::LOG_FATAL).stream()
#define LOG(sev) SSB_LOG_##sev
#define CHECK(cond) \
  (cond) ? (void)0 : ::ssb::LogVoidify() & ::ssb::LogMessage(__FILE__, __LINE__, ::ssb::LOG_FATAL).stream() << "Check failed: " #cond " "
#define PCHECK(cond) \
  (cond) ? (void)0 : ::ssb::LogVoidify() & ::ssb::LogMessage(__FILE__, __LINE__, ::ssb::LOG_FATAL, true).stream() << "Check failed: " #cond " "
#define SSB_CHECK_OP(a, b, op) \
  ((a)op(b)) ? (void)0 : ::ssb::LogVoidify() & ::ssb::LogMessage(__FILE__, __LINE__, ::ssb::LOG_FATAL).stream() << "Check failed: " #a " " #op " " #b " (" << (a) << " vs " << (b) << ") "
#define CHECK_EQ(a, b) SSB_CHECK_OP(a, b, ==)
#define CHECK_NE(a, b) SSB_CHECK_OP(a, b, !=)
#define CHECK_GT(a, b) SSB_CHECK_OP(a, b, >)
#define CHECK_GE(a, b) SSB_CHECK_OP(a, b, >=)
#define CHECK_LT(a, b) SSB_CHECK_OP(a, b, <)
#define CHECK_LE(a, b) SSB_CHECK_OP(a, b, <=)
#ifdef SSB_NDEBUG
#define DLOG(sev) true ? (void)0 : ::ssb::LogVoidify() & LOG(sev)
#define DCHECK(cond) true ? (void)0 : ::ssb::LogVoidify() & LOG(FATAL)
#else
#define DLOG(sev) LOG(sev)
#define DCHECK(cond) CHECK(cond)
#endif

// ---- fork-safe popen (ref: utils/common.h:57-71) -------------------------------------------
inline std::mutex& global_fork_mutex() {
  static std::mutex m;
  return m;
}
inline FILE* guarded_popen(const char* command, const char* type) {
  std::lock_guard<std::mutex> lock(global_fork_mutex());
  return popen(command, type);
}
inline int guarded_pclose(FILE* stream) {
  std::lock_guard<std::mutex> lock(global_fork_mutex());
  return pclose(stream);
}

// Non-copyable mixin. ref: utils/VirtualObject.h:14-20
class VirtualObject {
 public:
  VirtualObject() = default;
  VirtualObject(const VirtualObject&) = delete;
  VirtualObject& operator=(const VirtualObject&) = delete;
  virtual ~VirtualObject() = default;
};

}  // namespace ssb
