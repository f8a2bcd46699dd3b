// server.h — the server role: hold a shard of the sparse table, answer pulls, apply pushes,
// back the parameters up, dump on termination.
// ref: /root/reference/src/core/framework/SwiftServer.h:17-53.
// Order differs on purpose: the pull/push/terminate handlers are installed BEFORE the node
// registers with the master. The master releases the route only when every node has
// registered, so by the time any worker can address this server its handlers exist — this
// replaces the reference's `sleep(3 s)` in the worker (SwiftWorker.h:103-105, D4).
#pragma once
#include "../system/node_init.h"
#include "../system/server_init.h"
#include "../system/server_terminate.h"
namespace ssb {

template <typename Key, typename Param, typename PullMethod, typename PushMethod>
class SwiftServer {
 public:
  typedef SparseTable<Key, Param> table_t;
  SwiftServer(PullMethod& pull_method, PushMethod& push_method)
      : table_(global_sparse_table<Key, Param>()),
        init_pull_(table_, pull_method),
        init_push_(table_, push_method),
        terminate_(table_) {}
  void operator()() {
    ps_transfer();  // bind + start the listener
    init_pull_();
    init_push_();
    terminate_.register_handler();
    node_transfer_init_(true);
    node_hashfrag_init_();
    LOG(WARNING) << "server " << ps_transfer().client_id() << " is serving";
    terminate_();
    LOG(WARNING) << "server terminated normally";
  }
  table_t& table() { return table_; }

 private:
  table_t& table_;
  ServerInitPullMethod<table_t, PullMethod> init_pull_;
  ServerInitPushMethod<table_t, PushMethod> init_push_;
  ServerTerminate<table_t> terminate_;
  NodeTransferInit node_transfer_init_;
  NodeHashfragInit node_hashfrag_init_;
};

}  // namespace ssb
