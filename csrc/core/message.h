// message.h — MetaMessage, Message (RAII zmq_msg_t), Package, Request.
// ref: /root/reference/src/core/Message.h:12-200.
// Wire format kept: frame 1 = 24-byte raw MetaMessage
// {int32 message_class @0; uint16 ip[4] @4; uint16 port @12; pad; int32 client_id @16;
//  int32 message_id @20}, frame 2 = content. message_class -1 = response; client_id -1/-2 =
// unregistered worker/server. Differences by design: the meta is read from zmq_msg_data()
// (the reference memcpy's from the zmq_msg_t struct itself, Message.h:156, D1), a Request is
// built from the received frames with one copy, and empty content is legal.
#pragma once
#include "net.h"
namespace ssb {

struct MetaMessage {
  int32_t message_class = 0;
  Addr addr;                 // 10 bytes @4, then 2 bytes padding
  int32_t client_id = -3;
  int32_t message_id = -1;
};
static_assert(sizeof(MetaMessage) == 24, "MetaMessage must be 24 bytes on the wire");

// RAII wrapper over zmq_msg_t
class Message {
 public:
  Message() { PCHECK(zmq_msg_init(&msg_) == 0); }
  explicit Message(size_t size) { PCHECK(zmq_msg_init_size(&msg_, size) == 0); }
  Message(const char* buf, size_t size) {
    PCHECK(zmq_msg_init_size(&msg_, size) == 0);
    if (size) std::memcpy(zmq_msg_data(&msg_), buf, size);
  }
  ~Message() { zmq_msg_close(&msg_); }
  Message(const Message&) = delete;
  Message& operator=(const Message&) = delete;
  Message(Message&& o) noexcept {
    zmq_msg_init(&msg_);
    zmq_msg_move(&msg_, &o.msg_);
  }
  Message& operator=(Message&& o) noexcept {
    if (this != &o) zmq_msg_move(&msg_, &o.msg_);
    return *this;
  }
  void assign(const char* buf, size_t size) {
    zmq_msg_close(&msg_);
    PCHECK(zmq_msg_init_size(&msg_, size) == 0);
    if (size) std::memcpy(zmq_msg_data(&msg_), buf, size);
  }
  size_t size() const { return zmq_msg_size(&msg_); }
  bool empty() const { return size() == 0; }
  char* buffer() { return static_cast<char*>(zmq_msg_data(&msg_)); }
  zmq_msg_t& zmg() { return msg_; }

 private:
  zmq_msg_t msg_;
};

struct Request;

// Two frames ready to be sent. ref: Message.h:130-143
struct Package {
  Package() = default;
  explicit Package(Request& request);
  Message meta;
  Message cont;
};

struct Request {
  typedef std::function<void(std::shared_ptr<Request>)> response_call_back_t;
  Request() = default;
  Request(const Request&) = delete;
  Request& operator=(const Request&) = delete;
  Request(Request&& o) noexcept : meta(o.meta), cont(std::move(o.cont)), call_back_handler(std::move(o.call_back_handler)) {}
  // build from received frames (meta read from the message DATA — D1 fixed)
  explicit Request(Package&& pkg) {
    CHECK_EQ(pkg.meta.size(), sizeof(MetaMessage));
    std::memcpy(&meta, pkg.meta.buffer(), sizeof(MetaMessage));
    cont.set(pkg.cont.buffer(), pkg.cont.size());
  }
  bool is_response() const { return meta.message_class == -1; }       // ref: Message.h:175-176
  void set_msg_id(int id) { meta.message_id = id; }
  int message_id() const { return meta.message_id; }
  bool is_worker_unregistered() const { return meta.client_id == -1; }  // ref: Message.h:180-183
  bool is_server_unregistered() const { return meta.client_id == -2; }

  MetaMessage meta;
  BinaryBuffer cont;
  response_call_back_t call_back_handler;
};

inline Package::Package(Request& request)
    : meta(reinterpret_cast<const char*>(&request.meta), sizeof(MetaMessage)),
      cont(request.cont.buffer(), request.cont.size()) {}

}  // namespace ssb
