// push_access.h — worker-side push: send (key, grad) for a key set to the owner servers, reset
// the local gradients, block until every server acknowledged.
// ref: /root/reference/src/core/parameter/global_push_access.h:12-159 (GlobalPushAccess).
// Differences by design: the countdown counts only the buckets actually sent (the reference
// counts empty ones too and can hang, D6); the wait has a timeout.
#pragma once
#include "pull_access.h"
namespace ssb {

template <typename Key, typename Val, typename Grad>
class GlobalPushAccess : public VirtualObject {
 public:
  typedef Key key_t;
  typedef Val val_t;
  typedef Grad grad_t;
  typedef GlobalParamCache<Key, Val, Grad> param_cache_t;

  GlobalPushAccess() : transfer_(ps_transfer()) {}

  template <typename KeyContainer>
  void push_with_barrier(const KeyContainer& keys, param_cache_t& cache) {
    std::map<int, BinaryBuffer> node_reqs;
    arrange_local_grads(keys, cache, node_reqs);
    if (node_reqs.empty()) return;
    CounterBarrier barrier((int)node_reqs.size());
    for (auto& kv : node_reqs) {
      Request req;
      req.meta.message_class = WORKER_PUSH_REQUEST;
      req.cont = std::move(kv.second);
      req.call_back_handler = [&barrier](std::shared_ptr<Request>) { barrier.count_down(); };
      transfer_.send(std::move(req), kv.first);
    }
    long timeout_ms = 1000L * global_config().get_config("rpc_timeout", "120").to_int32();
    CHECK(barrier.block(timeout_ms)) << "push timed out after " << timeout_ms << " ms: a server is not answering";
  }

 private:
  // ref: global_push_access.h:80-99 — copy (key, grad) into the owner's bucket, reset the grad
  template <typename KeyContainer>
  void arrange_local_grads(const KeyContainer& keys, param_cache_t& cache, std::map<int, BinaryBuffer>& node_reqs) {
    auto& hashfrag = shared_hashfrag();
    rwlock_write_guard g(cache.rwlock());
    auto& grads = cache.grads();
    for (const auto& key : keys) {
      grad_t* gr = grads.find(key);
      if (!gr) continue;
      BinaryBuffer& bb = node_reqs[hashfrag.to_node_id((uint64_t)key)];
      bb << key << *gr;
      gr->reset();
    }
  }
  PsTransfer& transfer_;
};

template <typename Key, typename Val, typename Grad>
inline GlobalPushAccess<Key, Val, Grad>& global_push_access() {
  static GlobalPushAccess<Key, Val, Grad> access;
  return access;
}

}  // namespace ssb
