// node_init.h — node side of cluster bring-up: register with the master, adopt the route and
// the assigned id, then fetch the hash-fragment map.
// ref: /root/reference/src/core/system/node_init.h:16-152 (NodeTransferInit, NodeHashfragInit).
#pragma once
#include "../parameter/hashfrag.h"
#include "message_classes.h"
#include "server_worker_route.h"
namespace ssb {

class NodeTransferInit {
 public:
  NodeTransferInit() : transfer_(ps_transfer()) {}
  // is_server selects the unregistered-id marker: -2 server / -1 worker (Message.h:180-183)
  void operator()(bool is_server) {
    register_master();
    register_local_node_to_master(is_server);
    wait_for_route();
  }

 private:
  void register_master() {  // ref: node_init.h:28-32
    std::string master_addr = global_config().get_config("master_addr").to_string();
    if (!transfer_.route().has_node(0)) transfer_.route().register_node(0, master_addr);
  }
  void register_local_node_to_master(bool is_server) {  // ref: node_init.h:35-71
    Request req;
    req.meta.message_class = NODE_INIT_ADDRESS;
    req.meta.client_id = is_server ? -2 : -1;
    Addr self(transfer_.recv_addr());
    req.meta.addr = self;
    req.cont << self;
    req.call_back_handler = [this](std::shared_ptr<Request> rsp) {
      transfer_.route().deserialize(rsp->cont);
      transfer_.set_client_id(rsp->meta.client_id);
      LOG(INFO) << "node registered as id " << rsp->meta.client_id;
      barrier_.set_state_valid();
    };
    transfer_.send(std::move(req), 0);
  }
  void wait_for_route() {  // ref: node_init.h:73-84
    long timeout_ms = 1000L * global_config().get_config("init_timeout", "60").to_int32();
    barrier_.time_limit(timeout_ms, [] { LOG(FATAL) << "timed out waiting for the route table from the master"; });
    barrier_.block();
  }
  PsTransfer& transfer_;
  StateBarrier barrier_;
};

class NodeHashfragInit {
 public:
  NodeHashfragInit() : transfer_(ps_transfer()) {}
  void operator()() {  // ref: node_init.h:105-145
    Request req;
    req.meta.message_class = NODE_ASKFOR_HASHFRAG;
    req.cont << (int32_t)123;  // the protocol's non-empty dummy payload
    req.call_back_handler = [this](std::shared_ptr<Request> rsp) {
      shared_hashfrag().deserialize(rsp->cont);
      barrier_.set_state_valid();
    };
    transfer_.send(std::move(req), 0);
    long timeout_ms = 1000L * global_config().get_config("init_timeout", "60").to_int32();
    barrier_.time_limit(timeout_ms, [] { LOG(FATAL) << "timed out waiting for the hash-fragment map"; });
    barrier_.block();
  }

 private:
  PsTransfer& transfer_;
  StateBarrier barrier_;
};

}  // namespace ssb
