// net.h — network helpers and the Addr type.
// ref: /root/reference/src/core/common.h:27-170 (ignore_signal_call, zmq_bind_random_port,
// zmq_send_push_once, thread_guard, get_local_ip, Addr).
#pragma once
#include <arpa/inet.h>
#include <ifaddrs.h>
#include <netinet/in.h>
#include <random>
#include "utils/all.h"
#include "zmq_api.h"
namespace ssb {

// Retry a call while it fails with EINTR. ref: core/common.h:27-38
template <typename F, typename... Args>
inline auto ignore_signal_call(F&& f, Args&&... args) -> decltype(f(args...)) {
  for (;;) {
    auto r = f(args...);
    if (r != -1 || errno != EINTR) return r;
  }
}

// Bind to a random free port on `ip`; returns the port. ref: core/common.h:40-55.
// Uses the wildcard-port form and reads the endpoint back instead of guessing ports.
inline int zmq_bind_random_port(const std::string& ip, void* socket, std::string& addr_out, int min_port = 1024,
                                int max_port = 65535) {
  (void)min_port; (void)max_port;
  std::string ep = "tcp://" + ip + ":*";
  PCHECK(zmq_bind(socket, ep.c_str()) == 0) << "zmq_bind " << ep << ": " << zmq_strerror(zmq_errno());
  char buf[256];
  size_t len = sizeof(buf);
  PCHECK(zmq_getsockopt(socket, ZMQ_LAST_ENDPOINT, buf, &len) == 0);
  addr_out = buf;
  size_t colon = addr_out.find_last_of(':');
  return std::stoi(addr_out.substr(colon + 1));
}

// Send one message through an ephemeral PUSH socket (used for listener poison pills).
// ref: core/common.h:57-64
inline void zmq_send_push_once(void* zmq_ctx, zmq_msg_t* msg, const std::string& addr) {
  void* sender = zmq_socket(zmq_ctx, ZMQ_PUSH);
  PCHECK(sender != nullptr);
  int linger = 1000;
  zmq_setsockopt(sender, ZMQ_LINGER, &linger, sizeof(linger));
  PCHECK(ignore_signal_call(zmq_connect, sender, addr.c_str()) == 0);
  PCHECK(ignore_signal_call(zmq_msg_send, msg, sender, 0) >= 0);
  PCHECK(zmq_close(sender) == 0);
}

// Joins the thread on scope exit. ref: core/common.h:67-85
class thread_guard {
 public:
  explicit thread_guard(std::thread& t) : t_(t) {}
  ~thread_guard() { if (t_.joinable()) t_.join(); }
  thread_guard(const thread_guard&) = delete;
  thread_guard& operator=(const thread_guard&) = delete;

 private:
  std::thread& t_;
};

// Last non-loopback IPv4 address, else 127.0.0.1. ref: core/common.h:87-113.
// SSB200_LOCAL_IP overrides (single-box jobs use 127.0.0.1).
inline std::string get_local_ip() {
  if (const char* e = std::getenv("SSB200_LOCAL_IP")) return e;
  std::string ip = "127.0.0.1";
  struct ifaddrs* ifs = nullptr;
  if (getifaddrs(&ifs) != 0) return ip;
  for (struct ifaddrs* it = ifs; it; it = it->ifa_next) {
    if (!it->ifa_addr || it->ifa_addr->sa_family != AF_INET) continue;
    char buf[INET_ADDRSTRLEN];
    inet_ntop(AF_INET, &((struct sockaddr_in*)it->ifa_addr)->sin_addr, buf, sizeof(buf));
    if (std::strncmp(buf, "127.", 4) != 0) ip = buf;
  }
  freeifaddrs(ifs);
  return ip;
}

// IPv4 address + port; wire form 5 x uint16. ref: core/common.h:115-170
struct Addr {
  uint16_t addr[4] = {0, 0, 0, 0};
  uint16_t port = 0;
  Addr() = default;
  explicit Addr(const std::string& s) { from_string(s); }
  // accepts "a.b.c.d:p" and "tcp://a.b.c.d:p"
  void from_string(const std::string& s0) {
    std::string s = s0;
    if (headswith(s, "tcp://")) s = s.substr(6);
    auto parts = split(s, ".:");
    CHECK_EQ(parts.size(), (size_t)5) << "bad address " << s0;
    for (int i = 0; i < 4; ++i) addr[i] = (uint16_t)std::stoi(parts[i]);
    port = (uint16_t)std::stoi(parts[4]);
  }
  std::string to_string() const {
    return format_string("tcp://%d.%d.%d.%d:%d", addr[0], addr[1], addr[2], addr[3], port);
  }
  bool operator==(const Addr& o) const {
    return std::memcmp(addr, o.addr, sizeof(addr)) == 0 && port == o.port;
  }
  friend BinaryBuffer& operator<<(BinaryBuffer& bb, const Addr& a) {
    for (int i = 0; i < 4; ++i) bb << a.addr[i];
    bb << a.port;
    return bb;
  }
  friend BinaryBuffer& operator>>(BinaryBuffer& bb, Addr& a) {
    for (int i = 0; i < 4; ++i) bb >> a.addr[i];
    bb >> a.port;
    return bb;
  }
  friend std::ostream& operator<<(std::ostream& os, const Addr& a) { return os << a.to_string(); }
};

}  // namespace ssb
