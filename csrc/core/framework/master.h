// master.h — the master role: bring the cluster up, wait for the workers, shut the servers down.
// ref: /root/reference/src/core/framework/SwiftMaster.h:8-29.
#pragma once
#include "../system/master_init.h"
#include "../system/group_exchange.h"
#include "../system/master_terminate.h"
namespace ssb {

class SwiftMaster {
 public:
  SwiftMaster() = default;
  void operator()() {
    terminate_.register_handler();  // before the route goes out: a fast worker may finish early
    group_.register_handler();      // GPU ranks all-gather through the master (group_exchange.h)
    init_();
    LOG(WARNING) << "master: cluster is up (" << init_.server_count() << " servers, " << init_.worker_count() << " workers)";
    terminate_();
  }

 private:
  MasterTransferInit init_;
  MasterTerminate terminate_;
  MasterGroupExchange group_;
};

}  // namespace ssb
