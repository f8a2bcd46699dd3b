// listener.h — owns the process's inbox: one ZMQ_PULL socket and its receiver threads.
// ref: /root/reference/src/core/transfer/Listener.h:15-119.
// listen() binds a random port on the local IP, listen(addr) a fixed endpoint;
// service_start() spawns `thread_num` threads running main_loop(); service_end() sends one
// empty poison message per thread to its own address and joins.
#pragma once
#include "../message.h"
namespace ssb {

class Listener : public VirtualObject {
 public:
  explicit Listener(void* zmq_ctx = nullptr) { if (zmq_ctx) init_receiver(zmq_ctx); }
  // create the inbox socket on `zmq_ctx` (the route owns the context, ref: Route.h:23)
  void init_receiver(void* zmq_ctx) {
    CHECK(receiver_ == nullptr);
    zmq_ctx_ = zmq_ctx;
    receiver_ = zmq_socket(zmq_ctx_, ZMQ_PULL);
    PCHECK(receiver_ != nullptr) << "zmq_socket(PULL)";
    int hwm = 0;  // unbounded: a PS inbox must never drop or block senders on HWM
    zmq_setsockopt(receiver_, ZMQ_RCVHWM, &hwm, sizeof(hwm));
  }
  ~Listener() override {
    if (service_running_) service_end();
    if (receiver_) { int linger = 0; zmq_setsockopt(receiver_, ZMQ_LINGER, &linger, sizeof(linger)); zmq_close(receiver_); }
  }
  void set_thread_num(int n) { CHECK_GT(n, 0); thread_num_ = n; }
  int thread_num() const { return thread_num_; }

  // bind a random port on the local ip. ref: Listener.h:80-88
  void listen() {
    std::string ip = get_local_ip();
    recv_port_ = zmq_bind_random_port(ip, receiver_, recv_addr_);
    LOG(INFO) << "listener bound to " << recv_addr_;
  }
  // bind a fixed endpoint. ref: Listener.h:92-97
  void listen(const std::string& addr) {
    PCHECK(ignore_signal_call(zmq_bind, receiver_, addr.c_str()) == 0)
        << "zmq_bind " << addr << ": " << zmq_strerror(zmq_errno());
    recv_addr_ = addr;
    recv_port_ = Addr(addr).port;
  }
  const std::string& recv_addr() const { return recv_addr_; }
  int recv_port() const { return recv_port_; }
  void* receiver() { return receiver_; }
  void* zmq_ctx() { return zmq_ctx_; }
  std::mutex& receiver_mutex() { return receiver_mutex_; }

  void service_start() {
    CHECK(!service_running_);
    CHECK(!recv_addr_.empty()) << "listen() first";
    service_running_ = true;
    for (int i = 0; i < thread_num_; ++i) threads_.emplace_back([this] { main_loop(); });
  }
  void service_end() {
    if (!service_running_) return;
    for (int i = 0; i < thread_num_; ++i) {
      zmq_msg_t poison;
      zmq_msg_init_size(&poison, 0);
      zmq_send_push_once(zmq_ctx_, &poison, recv_addr_);
      zmq_msg_close(&poison);
    }
    for (auto& t : threads_)
      if (t.joinable()) t.join();
    threads_.clear();
    service_running_ = false;
  }
  virtual void main_loop() = 0;

 protected:
  void* zmq_ctx_ = nullptr;
  void* receiver_ = nullptr;
  std::string recv_addr_;
  int recv_port_ = -1;
  int thread_num_ = 1;
  bool service_running_ = false;
  std::vector<std::thread> threads_;
  std::mutex receiver_mutex_;
};

}  // namespace ssb
