// server_init.h — the server's pull and push request handlers and the periodic parameter
// backup.  ref: /root/reference/src/core/system/server/init.h:27-163
// (ServerInitPullMethod, ServerInitPushMethod).
// pull: parse keys, for each get_pull_value (lazy-creates the row), answer (key, val)*.
// push: parse (key, grad)*, apply_push_value each, answer 1234; every `param_backup_period`
// push requests write <param_backup_root>/param-<node id>-<counter>.txt (the reference names the
// file param-<counter>.txt on every server, so servers sharing a directory overwrite each
// other, D10 — the node id is added here).
#pragma once
#include <fstream>
#include "../parameter/sparsetable.h"
#include "message_classes.h"
#include "server_worker_route.h"
namespace ssb {

template <typename Table, typename PullMethod>
class ServerInitPullMethod {
 public:
  typedef typename Table::key_t key_t;
  typedef typename PullMethod::pull_val_t val_t;
  ServerInitPullMethod(Table& table, PullMethod& method)
      : transfer_(ps_transfer()), agent_(make_pull_access(table, method)) {}
  void operator()() {
    transfer_.message_class().add(WORKER_PULL_REQUEST, [this](std::shared_ptr<Request> req, Request& rsp) {
      // ref: server/init.h:49-69. The reference's worker sends (key, empty value) pairs; this framework's worker
      // sends the bare key array (a third of the bytes for dim-0 values). `wire_compat: reference` makes both ends
      // speak the reference's form, so that nodes of the two code bases can share a cluster (tests/test_interop.py).
      std::vector<key_t> keys;
      if (reference_wire()) {
        while (!req->cont.read_finished()) {
          key_t key;
          val_t unused;
          req->cont >> key >> unused;
          keys.push_back(key);
        }
      } else {
        keys.resize((req->cont.size() - req->cont.status()) / sizeof(key_t));
        req->cont.get_array(keys.data(), keys.size());
      }
      const size_t n = keys.size();
      val_t val;
      for (const key_t& key : keys) {
        agent_->get_pull_value(key, val);
        rsp.cont << key << val;
      }
      if (n == 0) rsp.cont << (int32_t)0;  // never defer an answer
    });
  }

 private:
  PsTransfer& transfer_;
  std::unique_ptr<PullAccessAgent<Table, PullMethod>> agent_;
};

template <typename Table, typename PushMethod>
class ServerInitPushMethod {
 public:
  typedef typename Table::key_t key_t;
  typedef typename PushMethod::push_val_t grad_t;
  ServerInitPushMethod(Table& table, PushMethod& method)
      : transfer_(ps_transfer()), table_(table), agent_(make_push_access(table, method)) {
    backup_period_ = global_config().get_config("param_backup_period", "0").to_int32();
    if (backup_period_ > 0) backup_root_ = global_config().get_config("param_backup_root").to_string();
  }
  void operator()() {
    transfer_.message_class().add(WORKER_PUSH_REQUEST, [this](std::shared_ptr<Request> req, Request& rsp) {
      // ref: server/init.h:116-132
      key_t key;
      grad_t grad;
      while (!req->cont.read_finished()) {
        req->cont >> key >> grad;
        if (!agent_->apply_push_value(key, grad)) LOG(ERROR) << "push for key " << key << " that was never pulled (ignored)";
      }
      rsp.cont << (int32_t)1234;
      long c = ++push_counter_;
      if (backup_period_ > 0 && c % backup_period_ == 0) backup_param(c);
    });
  }
  long push_counter() const { return push_counter_.load(); }

 private:
  void backup_param(long counter) {  // ref: server/init.h:138-149
    std::lock_guard<std::mutex> lk(backup_mutex_);
    std::string path = format_string("%s/param-%d-%ld.txt", backup_root_.c_str(), transfer_.client_id(), counter);
    std::ofstream out(path);
    if (!out.good()) { LOG(ERROR) << "cannot write backup " << path; return; }
    table_.output(out);
    LOG(INFO) << "parameter backup written to " << path;
  }
  PsTransfer& transfer_;
  Table& table_;
  std::unique_ptr<PushAccessAgent<Table, PushMethod>> agent_;
  std::atomic<long> push_counter_{0};
  int backup_period_ = 0;
  std::string backup_root_;
  std::mutex backup_mutex_;
};

}  // namespace ssb
