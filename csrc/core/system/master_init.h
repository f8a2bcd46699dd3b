// master_init.h — master side of bring-up: wait for `expected_node_num` registrations, build
// the hash-fragment map over the servers, then answer every (deferred) registration with the
// full route table; serve fragment-map requests.
// ref: /root/reference/src/core/system/master/init.h:21-171 (MasterTransferInit).
#pragma once
#include "../parameter/hashfrag.h"
#include "message_classes.h"
#include "server_worker_route.h"
namespace ssb {

class MasterTransferInit {
 public:
  MasterTransferInit() : transfer_(ps_transfer()) {
    expected_node_num_ = global_config().get_config("expected_node_num").to_int32();
    CHECK_GT(expected_node_num_, 0);
    // the master registers itself as server id 0. ref: master/init.h:34
    int id = transfer_.route().register_node_(true, transfer_.recv_addr());
    CHECK_EQ(id, 0);
    transfer_.set_client_id(0);
  }
  void operator()() {
    register_message_class();
    wait_for_nodes_register();
    init_hashfrag();
    send_route_to_nodes();
  }
  int server_count() const { return transfer_.route().server_num() - 1; }  // excluding the master
  int worker_count() const { return transfer_.route().worker_num(); }

 private:
  struct Pending { int id; int message_id; };

  void register_message_class() {  // ref: master/init.h:53-61
    transfer_.message_class().add(NODE_INIT_ADDRESS, [this](std::shared_ptr<Request> req, Request& rsp) {
      node_init_address(req, rsp);
    });
    transfer_.message_class().add(NODE_ASKFOR_HASHFRAG, [this](std::shared_ptr<Request>, Request& rsp) {
      std::lock_guard<std::mutex> lk(mutex_);
      shared_hashfrag().serialize(rsp.cont);
    });
  }
  // ref: master/init.h:122-150 — register, remember the message id, leave the response EMPTY
  // (deferred) and unblock the main thread when everyone has arrived.
  void node_init_address(std::shared_ptr<Request> req, Request& /*rsp*/) {
    Addr addr;
    req->cont >> addr;
    bool is_server = req->is_server_unregistered();
    CHECK(is_server || req->is_worker_unregistered()) << "registration from an already-registered node";
    int id = transfer_.route().register_node_(is_server, addr.to_string());
    bool all = false;
    {
      std::lock_guard<std::mutex> lk(mutex_);
      pending_.push_back({id, req->meta.message_id});
      all = (int)pending_.size() == expected_node_num_;
    }
    LOG(INFO) << "master registered " << (is_server ? "server " : "worker ") << id << " at " << addr;
    if (all) barrier_.set_state_valid();
  }
  void wait_for_nodes_register() {  // ref: master/init.h:63-72
    long timeout_ms = 1000L * global_config().get_config("master_time_out", "120").to_int32();
    barrier_.time_limit(timeout_ms, [this] {
      LOG(FATAL) << "master timed out: " << pending_.size() << " of " << expected_node_num_ << " nodes registered";
    });
    barrier_.block();
  }
  void init_hashfrag() {  // ref: master/init.h:101-106
    std::lock_guard<std::mutex> lk(mutex_);
    CHECK_GT(server_count(), 0) << "no server registered";
    shared_hashfrag().init(server_count());
  }
  void send_route_to_nodes() {  // ref: master/init.h:74-99
    std::vector<Pending> pend;
    { std::lock_guard<std::mutex> lk(mutex_); pend = pending_; }
    for (auto& p : pend) {
      Request rsp;
      rsp.meta.message_class = -1;
      rsp.meta.message_id = p.message_id;
      rsp.meta.client_id = p.id;  // the node adopts this id
      transfer_.route().serialize(rsp.cont);
      transfer_.send_response(std::move(rsp), p.id);
    }
  }
  PsTransfer& transfer_;
  int expected_node_num_ = 0;
  std::mutex mutex_;
  std::vector<Pending> pending_;
  StateBarrier barrier_;
};

}  // namespace ssb
