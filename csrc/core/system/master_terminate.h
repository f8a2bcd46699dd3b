// master_terminate.h — count WORKER_FINISH_WORK from every worker, then tell every server to
// terminate and wait for their acks.
// ref: /root/reference/src/core/system/master/terminate.h:15-109 (MasterTerminate).
#pragma once
#include "message_classes.h"
#include "server_worker_route.h"
namespace ssb {

class MasterTerminate {
 public:
  MasterTerminate() : transfer_(ps_transfer()) {}
  void operator()() {
    wait_workers();
    terminate_servers();
    LOG(WARNING) << "Master terminated normally";
  }
  // must be installed before the route goes out so no finish message is dropped
  void register_handler() {
    if (registered_) return;
    registered_ = true;
    transfer_.message_class().add(WORKER_FINISH_WORK, [this](std::shared_ptr<Request>, Request& rsp) {
      rsp.cont << (int32_t)1234;
      int done = ++finished_workers_;
      if (done >= expected_workers() && expected_workers() > 0) workers_done_.set_state_valid();
    });
  }
  // `colocated_workers: 1` (GPU jobs: every registered server process also hosts a worker, group_exchange.h): a
  // WORKER_FINISH_WORK is expected from every server as well as from every worker node.
  int expected_workers() {
    int n = transfer_.route().worker_num();
    if (global_config().get_config("colocated_workers", "0").to_int32() != 0) n += transfer_.route().server_num() - 1;
    return n;
  }

 private:
  void wait_workers() {  // ref: master/terminate.h:44-62
    register_handler();
    if (expected_workers() == 0 || finished_workers_.load() >= expected_workers()) workers_done_.set_state_valid();
    long alive_s = global_config().get_config("master_longest_alive_duration", "0").to_int32();
    if (alive_s > 0)
      workers_done_.time_limit(1000L * alive_s, [] { LOG(FATAL) << "master_longest_alive_duration exceeded"; });
    workers_done_.block();
  }
  void terminate_servers() {  // ref: master/terminate.h:67-92
    std::vector<int> servers;
    for (int id : transfer_.route().server_ids())
      if (id != 0) servers.push_back(id);
    if (servers.empty()) return;
    CounterBarrier acks((int)servers.size());
    for (int id : servers) {
      Request req;
      req.meta.message_class = SERVER_TOLD_TO_TERMINATE;
      req.cont << (int32_t)1234;
      req.call_back_handler = [&acks](std::shared_ptr<Request>) { acks.count_down(); };
      transfer_.send(std::move(req), id);
    }
    if (!acks.block(60000)) LOG(ERROR) << "not every server acknowledged termination";
  }
  PsTransfer& transfer_;
  std::atomic<int> finished_workers_{0};
  StateBarrier workers_done_;
  bool registered_ = false;
};

}  // namespace ssb
