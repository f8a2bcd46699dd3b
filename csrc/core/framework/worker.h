// worker.h — BaseAlgorithm (the user's training algorithm) and the worker role.
// ref: /root/reference/src/core/framework/SwiftWorker.h:19-153.
// The user subclasses BaseAlgorithm<Key,Val,Grad,Record>, implements train() and
// parse_record(), and inside train() calls global_pull_access<K,V,G>().pull_with_barrier(keys,
// cache) / global_push_access<K,V,G>().push_with_barrier(keys, cache) on a GlobalParamCache it
// owns.  local_train=1 runs the algorithm without any cluster (SwiftWorker.h:115-123).
#pragma once
#include "../parameter/push_access.h"
#include "../system/node_init.h"
#include "../system/worker_terminate.h"
namespace ssb {

template <typename Key, typename Val, typename Grad, typename Record>
class BaseAlgorithm : public VirtualObject {
 public:
  typedef Key key_t;
  typedef Val val_t;
  typedef Grad grad_t;
  typedef Record rcd_t;
  typedef GlobalParamCache<Key, Val, Grad> param_cache_t;

  virtual void train() = 0;
  virtual rcd_t parse_record(const std::string& line) = 0;

  void set_data_path(const std::string& path) { data_path_ = path; }
  const std::string& data_path() const {
    CHECK(!data_path_.empty()) << "set_data_path first";
    return data_path_;
  }
  std::shared_ptr<AsynExec::channel_t>& async_channel() { return async_channel_; }
  void create_async_channel() {  // ref: SwiftWorker.h:43-50
    CHECK(!async_channel_) << "can be called only once";
    async_channel_thread_num_ = global_config().get_config("async_channel_thread_num", "1").to_int32();
    CHECK_GT(async_channel_thread_num_, 0);
    AsynExec as(async_channel_thread_num_);
    async_channel_ = as.open();
  }
  int async_channel_thread_num() const { return async_channel_thread_num_; }
  void set_local_train(bool v) { local_train_ = v; }
  bool local_train() const { return local_train_; }

 protected:
  std::string data_path_;
  int async_channel_thread_num_ = 0;
  std::shared_ptr<AsynExec::channel_t> async_channel_;
  bool local_train_ = false;
};

template <typename Algorithm>
class SwiftWorker {
 public:
  typedef typename Algorithm::key_t key_t;
  typedef typename Algorithm::val_t val_t;
  typedef typename Algorithm::grad_t grad_t;

  explicit SwiftWorker(Algorithm& alg) : alg_(alg) {
    num_iters_ = global_config().get_config("num_iters", "1").to_int32();
    CHECK_GT(num_iters_, 0);
    learning_rate_ = global_config().get_config("learning_rate", "0.01").to_float();
    CHECK_GT(learning_rate_, 0.f);
    local_train_ = global_config().get_config("local_train", "0").to_int32() > 0;
    alg_.set_local_train(local_train_);
  }
  void operator()() {  // ref: SwiftWorker.h:88-124
    if (!local_train_) {
      NodeTransferInit node_transfer_init;
      NodeHashfragInit node_hashfrag_init;
      node_transfer_init(false);
      node_hashfrag_init();
      LOG(WARNING) << "worker " << ps_transfer().client_id() << " begins to train";
      alg_.train();
      ClientTerminate terminate;
      terminate();
    } else {
      alg_.train();
    }
  }
  int num_iters() const { return num_iters_; }
  float learning_rate() const { return learning_rate_; }

 private:
  Algorithm& alg_;
  int num_iters_ = 1;
  float learning_rate_ = 0.01f;
  bool local_train_ = false;
};

}  // namespace ssb
