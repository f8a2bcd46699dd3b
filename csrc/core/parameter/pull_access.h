// pull_access.h — worker-side pull: fetch rows for a key set from their owner servers into the
// local cache and block until every server answered.
// ref: /root/reference/src/core/parameter/global_pull_access.h:13-131 (GlobalPullAccess).
// Differences by design: the request carries keys only (the reference ships a full
// default-constructed Val per key, D13); the wait has a timeout (config `rpc_timeout`, s)
// so a dead server is reported instead of hanging the worker forever.
#pragma once
#include "../system/message_classes.h"
#include "../system/server_worker_route.h"
#include "hashfrag.h"
#include "param_cache.h"
namespace ssb {

template <typename Key, typename Val, typename Grad>
class GlobalPullAccess : public VirtualObject {
 public:
  typedef Key key_t;
  typedef Val val_t;
  typedef Grad grad_t;
  typedef GlobalParamCache<Key, Val, Grad> param_cache_t;

  GlobalPullAccess() : transfer_(ps_transfer()) {}

  template <typename KeyContainer>
  void pull_with_barrier(const KeyContainer& keys, param_cache_t& cache) {
    std::map<int, std::vector<key_t>> node_reqs;
    arrange_local_keys(keys, node_reqs);
    if (node_reqs.empty()) return;
    CounterBarrier barrier((int)node_reqs.size());
    for (auto& kv : node_reqs) {
      Request req;
      req.meta.message_class = WORKER_PULL_REQUEST;
      if (reference_wire()) {      // the reference's form: (key, empty value)*  — ref: global_pull_access.h:84-88
        val_t empty;
        for (const key_t& k : kv.second) req.cont << k << empty;
      } else {
        req.cont.reserve(kv.second.size() * sizeof(key_t));
        req.cont.put_array(kv.second.data(), kv.second.size());
      }
      req.call_back_handler = [&cache, &barrier](std::shared_ptr<Request> rsp) {
        {
          rwlock_write_guard g(cache.rwlock());
          auto& params = cache.params();
          auto& grads = cache.grads();
          while (!rsp->cont.read_finished()) {
            key_t key;
            val_t val;
            rsp->cont >> key >> val;
            params[key] = std::move(val);
            grads[key] = grad_t();
          }
        }
        barrier.count_down();
      };
      transfer_.send(std::move(req), kv.first);
    }
    long timeout_ms = 1000L * global_config().get_config("rpc_timeout", "120").to_int32();
    CHECK(barrier.block(timeout_ms)) << "pull timed out after " << timeout_ms << " ms: a server is not answering";
  }

 private:
  // ref: global_pull_access.h:58-72
  template <typename KeyContainer>
  void arrange_local_keys(const KeyContainer& keys, std::map<int, std::vector<key_t>>& node_reqs) {
    auto& hashfrag = shared_hashfrag();
    for (const auto& key : keys) node_reqs[hashfrag.to_node_id((uint64_t)key)].push_back(key);
  }
  PsTransfer& transfer_;
};

template <typename Key, typename Val, typename Grad>
inline GlobalPullAccess<Key, Val, Grad>& global_pull_access() {
  static GlobalPullAccess<Key, Val, Grad> access;
  return access;
}

}  // namespace ssb
