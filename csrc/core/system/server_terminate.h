// server_terminate.h — on SERVER_TOLD_TO_TERMINATE dump the table (stdout, or `param_dump_path`
// if configured), acknowledge, and release the main thread.
// ref: /root/reference/src/core/system/server/terminate.h:16-54 (ServerTerminate).
#pragma once
#include <fstream>
#include "../parameter/sparsetable.h"
#include "message_classes.h"
#include "server_worker_route.h"
namespace ssb {

template <typename Table>
class ServerTerminate {
 public:
  explicit ServerTerminate(Table& table) : transfer_(ps_transfer()), table_(table) {}
  void register_handler() {
    transfer_.message_class().add(SERVER_TOLD_TO_TERMINATE, [this](std::shared_ptr<Request>, Request& rsp) {
      std::string path = global_config().get_config("param_dump_path", "").to_string();
      if (path.empty()) {
        table_.output(std::cout);  // Hadoop streaming collected stdout as the job output
      } else {
        std::ofstream out(format_string("%s.%d", path.c_str(), transfer_.client_id()));
        table_.output(out);
      }
      rsp.cont << (int32_t)1234;
      barrier_.set_state_valid();
    });
  }
  void operator()() {
    barrier_.block();
    // let the ack leave before the process tears the sockets down
    std::this_thread::sleep_for(std::chrono::milliseconds(200));
  }

 private:
  PsTransfer& transfer_;
  Table& table_;
  StateBarrier barrier_;
};

}  // namespace ssb
