// worker_terminate.h — tell the master this worker is done and wait for the ack.
// ref: /root/reference/src/core/system/worker/terminate.h:17-58 (ClientTerminate). The
// reference sleeps 5 s first (worker/terminate.h:39) to let in-flight pushes land; pushes here
// are acknowledged before push_with_barrier returns, so no sleep is needed (D4).
#pragma once
#include "message_classes.h"
#include "server_worker_route.h"
namespace ssb {

class ClientTerminate {
 public:
  ClientTerminate() : transfer_(ps_transfer()) {}
  void operator()() {
    Request req;
    req.meta.message_class = WORKER_FINISH_WORK;
    req.cont << (int32_t)1234;
    req.call_back_handler = [this](std::shared_ptr<Request>) { barrier_.set_state_valid(); };
    transfer_.send(std::move(req), 0);
    barrier_.time_limit(60000, [] { LOG(ERROR) << "master did not acknowledge WORKER_FINISH_WORK"; });
    barrier_.block();
    LOG(WARNING) << "worker terminated normally";
  }

 private:
  PsTransfer& transfer_;
  StateBarrier barrier_;
};

}  // namespace ssb
