// group_exchange.h — the master as the control plane of a GPU job.
//
// ref: /root/reference/src/core/system/master/init.h:21-171 (registration, deferred route reply),
//      system/node_init.h:16-152 (node side), master/terminate.h:15-109 + server/terminate.h:16-54 +
//      worker/terminate.h:17-58 (WORKER_FINISH_WORK -> SERVER_TOLD_TO_TERMINATE -> dump).
// A GPU rank hosts server shard g AND worker g in one process. It registers with the master as a SERVER node
// (NODE_INIT_ADDRESS; the id the master assigns is its rank + 1), adopts the route table from the deferred reply,
// fetches the fragment map (NODE_ASKFOR_HASHFRAG), reports WORKER_FINISH_WORK when its worker half is done and
// dumps when the master says SERVER_TOLD_TO_TERMINATE — the reference's six message classes, unchanged.
// One class is added, NODE_GROUP_EXCHANGE: an all-gather of small byte strings THROUGH the master, answered with
// the same deferred-reply mechanism as the registration (the reply is held until every rank's contribution for
// that sequence number has arrived). The ranks use it for what the reference's nodes never needed: exchanging
// the names of the sockets that carry the VMM file descriptors, barriers around table set-up, and the max-over-
// ranks of a timing. The data plane (NVLink loads / stores / atomics on peer-mapped shards) never touches it.
#pragma once
#include <algorithm>
#include <map>
#include "../parameter/hashfrag.h"
#include "message_classes.h"
#include "node_init.h"
#include "server_worker_route.h"
namespace ssb {

// master side
class MasterGroupExchange {
 public:
  MasterGroupExchange() : transfer_(ps_transfer()) {}
  void register_handler() {
    if (registered_) return;
    registered_ = true;
    transfer_.message_class().add(NODE_GROUP_EXCHANGE, [this](std::shared_ptr<Request> req, Request& /*rsp: deferred*/) {
      int32_t seq = 0, len = 0;
      req->cont >> seq >> len;
      std::string payload((size_t)len, '\0');
      for (int32_t i = 0; i < len; ++i) { char ch; req->cont >> ch; payload[(size_t)i] = ch; }
      std::vector<Part> ready;
      {
        std::lock_guard<std::mutex> lk(mutex_);
        auto& parts = rounds_[seq];
        parts.push_back({req->meta.client_id, req->meta.message_id, std::move(payload)});
        const int group = transfer_.route().server_num() - 1;        // every server except the master itself
        if ((int)parts.size() < group) return;
        ready = std::move(parts);
        rounds_.erase(seq);
      }
      std::sort(ready.begin(), ready.end(), [](const Part& a, const Part& b) { return a.id < b.id; });
      for (auto& to : ready) {
        Request rsp;
        rsp.meta.message_class = -1;
        rsp.meta.message_id = to.message_id;
        rsp.meta.client_id = to.id;
        rsp.cont << (int32_t)ready.size();
        for (auto& p : ready) {
          rsp.cont << (int32_t)p.id << (int32_t)p.payload.size();
          for (char ch : p.payload) rsp.cont << ch;
        }
        transfer_.send_response(std::move(rsp), to.id);
      }
    });
  }

 private:
  struct Part { int id; int message_id; std::string payload; };
  PsTransfer& transfer_;
  std::mutex mutex_;
  std::map<int32_t, std::vector<Part>> rounds_;
  bool registered_ = false;
};

// node side: one per process
class GpuNode {
 public:
  GpuNode() : transfer_(ps_transfer()) {}
  // Register (as a server), adopt the route, fetch the fragment map. Returns the rank (id - 1).
  int start() {
    transfer_.message_class().add(SERVER_TOLD_TO_TERMINATE, [this](std::shared_ptr<Request>, Request& rsp) {
      told_.set_state_valid();
      // the reference's server dumps inside this handler, before the ack (server/terminate.h:27-45); the dump of
      // a GPU shard is made by the caller, so hold the ack until it reports the dump done (bounded)
      long wait_ms = 1000L * global_config().get_config("param_dump_timeout", "600").to_int32();
      dumped_.time_limit(wait_ms, [] { LOG(ERROR) << "dump not reported in time; acknowledging termination anyway"; });
      dumped_.block();
      rsp.cont << (int32_t)1234;
    });
    NodeTransferInit init;
    init(true);
    NodeHashfragInit frag;
    frag();
    return rank();
  }
  int rank() const { return transfer_.client_id() - 1; }
  int world() { return transfer_.route().server_num() - 1; }

  // All-gather of one byte string per rank, through the master. parts[r] = rank r's bytes. false = timed out.
  bool exchange(const std::string& mine, std::vector<std::string>& parts, long timeout_ms) {
    Request req;
    req.meta.message_class = NODE_GROUP_EXCHANGE;
    const int32_t seq = seq_++;
    req.cont << seq << (int32_t)mine.size();
    for (char ch : mine) req.cont << ch;
    auto done = std::make_shared<StateBarrier>();
    auto out = std::make_shared<std::vector<std::string>>();
    req.call_back_handler = [done, out](std::shared_ptr<Request> rsp) {
      int32_t n = 0;
      rsp->cont >> n;
      out->resize((size_t)n);
      for (int32_t i = 0; i < n; ++i) {
        int32_t id = 0, len = 0;
        rsp->cont >> id >> len;
        std::string s((size_t)len, '\0');
        for (int32_t k = 0; k < len; ++k) { char ch; rsp->cont >> ch; s[(size_t)k] = ch; }
        if (id >= 1 && id <= n) (*out)[(size_t)(id - 1)] = std::move(s);
      }
      done->set_state_valid();
    };
    transfer_.send(std::move(req), 0);
    bool timed_out = false;
    if (timeout_ms > 0) done->time_limit(timeout_ms, [&timed_out] { timed_out = true; });
    done->block();
    if (timed_out) return false;
    parts = std::move(*out);
    return true;
  }
  // The worker half is done: WORKER_FINISH_WORK to the master (acked), then wait until the master — having heard
  // from every worker — tells this server to terminate. After this returns the caller dumps its shard and calls
  // dump_done(). false = the master did not answer in time.
  bool finish(long timeout_ms) {
    Request req;
    req.meta.message_class = WORKER_FINISH_WORK;
    req.cont << (int32_t)1234;
    auto acked = std::make_shared<StateBarrier>();
    req.call_back_handler = [acked](std::shared_ptr<Request>) { acked->set_state_valid(); };
    transfer_.send(std::move(req), 0);
    bool timed_out = false;
    if (timeout_ms > 0) {
      acked->time_limit(timeout_ms, [&timed_out] { timed_out = true; });
      told_.time_limit(timeout_ms, [&timed_out] { timed_out = true; });
    }
    acked->block();
    if (timed_out) return false;
    told_.block();
    return !timed_out;
  }
  void dump_done() {
    dumped_.set_state_valid();
    std::this_thread::sleep_for(std::chrono::milliseconds(200));   // let the ack leave before the sockets go down
  }

 private:
  PsTransfer& transfer_;
  StateBarrier told_, dumped_;
  std::atomic<int32_t> seq_{1};
};

}  // namespace ssb
