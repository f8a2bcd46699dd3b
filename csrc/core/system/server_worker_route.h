// server_worker_route.h — node-id allocation and the serialised route table.
// ref: /root/reference/src/core/system/ServerWorkerRoute.h:14-84.
// ids: master = 0 (registered as a server), servers = 1, 2, …, workers = INT_MAX-1, INT_MAX-2, …
// wire: int32 server_num (incl. master), int32 worker_num, {int32 id, Addr}* — every entry
// except id 0 is registered on deserialise (the node already knows the master).
#pragma once
#include <climits>
#include "../transfer/transfer.h"
namespace ssb {

class ServerWorkerRoute : public BaseRoute {
 public:
  // master side: allocate an id for a new node and register it. ref: ServerWorkerRoute.h:17-31
  int register_node_(bool is_server, std::string addr) {
    std::lock_guard<std::mutex> lk(alloc_mutex_);
    int id = is_server ? server_num_++ : (INT_MAX - 1 - worker_num_++);
    register_node(id, std::move(addr));
    (is_server ? server_ids_ : worker_ids_).push_back(id);
    return id;
  }
  int server_num() const { return server_num_; }  // includes the master (id 0)
  int worker_num() const { return worker_num_; }
  const std::vector<int>& server_ids() const { return server_ids_; }
  const std::vector<int>& worker_ids() const { return worker_ids_; }

  void serialize(BinaryBuffer& bb) {
    auto addrs = addresses();
    bb << (int32_t)server_num_ << (int32_t)worker_num_;
    for (auto& kv : addrs) {
      CHECK(headswith(kv.second, "tcp://")) << "only tcp:// endpoints are routable: " << kv.second;
      bb << (int32_t)kv.first << Addr(kv.second);
    }
  }
  // node side: adopt the table. ref: ServerWorkerRoute.h:53-71 (there server_ids()/worker_ids()
  // are only populated on the master; here on every node).
  void deserialize(BinaryBuffer& bb) {
    int32_t s = 0, w = 0;
    bb >> s >> w;
    server_num_ = s;
    worker_num_ = w;
    server_ids_.clear();
    worker_ids_.clear();
    while (!bb.read_finished()) {
      int32_t id;
      Addr a;
      bb >> id >> a;
      if (id >= 0 && id < s) server_ids_.push_back(id);
      else worker_ids_.push_back(id);
      if (id == 0 || has_node(id)) continue;
      register_node(id, a.to_string());
    }
  }

 private:
  std::mutex alloc_mutex_;
  int server_num_ = 0, worker_num_ = 0;
  std::vector<int> server_ids_, worker_ids_;
};

typedef Transfer<ServerWorkerRoute> PsTransfer;
inline PsTransfer& ps_transfer() { return global_transfer<ServerWorkerRoute>(); }

}  // namespace ssb
