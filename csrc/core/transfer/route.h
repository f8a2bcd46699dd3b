// route.h — BaseRoute: the process's zmq context plus id -> address / PUSH socket / mutex maps.
// ref: /root/reference/src/core/transfer/Route.h:20-112.
#pragma once
#include <chrono>
#include <thread>
#include "../message.h"
namespace ssb {

class BaseRoute : public VirtualObject {
 public:
  BaseRoute() {
    zmq_ctx_ = zmq_ctx_new();
    PCHECK(zmq_ctx_ != nullptr) << "zmq_ctx_new";
  }
  ~BaseRoute() override {
    for (auto& kv : senders_) {
      int linger = 200;
      zmq_setsockopt(kv.second, ZMQ_LINGER, &linger, sizeof(linger));
      zmq_close(kv.second);
    }
    senders_.clear();
    if (zmq_ctx_) zmq_ctx_term(zmq_ctx_);
  }
  void* zmq_ctx() { return zmq_ctx_; }

  // create + connect a PUSH socket to `addr` for node `id`. ref: Route.h:31-39,92-96
  void register_node(int id, std::string addr) {
    rwlock_write_guard g(lock_);
    CHECK(addrs_.count(id) == 0) << "node " << id << " already registered";
    void* sender = zmq_socket(zmq_ctx_, ZMQ_PUSH);
    PCHECK(sender != nullptr);
    int hwm = 0;
    zmq_setsockopt(sender, ZMQ_SNDHWM, &hwm, sizeof(hwm));
    PCHECK(ignore_signal_call(zmq_connect, sender, addr.c_str()) == 0) << "zmq_connect " << addr;
    addrs_[id] = std::move(addr);
    senders_[id] = sender;
    send_mutexes_[id] = std::unique_ptr<std::mutex>(new std::mutex);
  }
  void delete_node(int id) {  // ref: Route.h:43-64
    rwlock_write_guard g(lock_);
    auto it = senders_.find(id);
    CHECK(it != senders_.end()) << "no node " << id;
    int linger = 0;
    zmq_setsockopt(it->second, ZMQ_LINGER, &linger, sizeof(linger));
    PCHECK(zmq_close(it->second) == 0);
    senders_.erase(it);
    addrs_.erase(id);
    send_mutexes_.erase(id);
  }
  bool has_node(int id) { rwlock_read_guard g(lock_); return addrs_.count(id) > 0; }
  // The master hands the route table to the nodes one after another, so a worker that already has
  // it can reach a server whose copy is still in flight; the server then has to answer a node it
  // does not know yet. Lookups therefore wait (bounded) for a missing id instead of aborting.
  bool wait_for_node(int id, int timeout_ms) {
    const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
    while (!has_node(id)) {
      if (std::chrono::steady_clock::now() >= deadline) return false;
      std::this_thread::sleep_for(std::chrono::milliseconds(1));
    }
    return true;
  }
  void set_unknown_node_wait_ms(int ms) { unknown_node_wait_ms_ = ms; }
  void* sender(int id) {
    if (!has_node(id)) wait_for_node(id, unknown_node_wait_ms_);
    rwlock_read_guard g(lock_);
    auto it = senders_.find(id);
    CHECK(it != senders_.end()) << "no route to node " << id;
    return it->second;
  }
  std::mutex& send_mutex(int id) {
    if (!has_node(id)) wait_for_node(id, unknown_node_wait_ms_);
    rwlock_read_guard g(lock_);
    auto it = send_mutexes_.find(id);
    CHECK(it != send_mutexes_.end()) << "no route to node " << id;
    return *it->second;
  }
  std::string address(int id) {
    if (!has_node(id)) wait_for_node(id, unknown_node_wait_ms_);
    rwlock_read_guard g(lock_);
    auto it = addrs_.find(id);
    CHECK(it != addrs_.end()) << "no route to node " << id;
    return it->second;
  }
  std::map<int, std::string> addresses() { rwlock_read_guard g(lock_); return addrs_; }
  size_t size() { rwlock_read_guard g(lock_); return addrs_.size(); }
  virtual void update() {}

 protected:
  void* zmq_ctx_ = nullptr;
  int unknown_node_wait_ms_ = 30000;
  RWLock lock_;
  std::map<int, std::string> addrs_;
  std::map<int, void*> senders_;
  std::map<int, std::unique_ptr<std::mutex>> send_mutexes_;
};

}  // namespace ssb
