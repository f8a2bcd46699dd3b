// transfer.h — asynchronous request/response RPC over ZeroMQ PUSH/PULL.
// ref: /root/reference/src/core/transfer/transfer.h:16-298 (MessageClass, Transfer<Route>,
// global_transfer). This is the CONTROL plane and the CPU fallback data plane of the new
// framework; on a B200 box the data plane is peer memory over NVLink (csrc/kernels/table.cu).
//
// Same protocol: send(Request&&, to_id) assigns message_id from an atomic counter, remembers
// the response callback, sends meta + content as two frames under the destination's mutex;
// main_loop() receives both frames, dispatches to handle_response when message_class == -1
// else handle_request; a handler that leaves the response content empty defers the reply
// (the master uses this to hold registrations until every node has arrived).
// Differences by design: MessageClass::get is locked (D12); destruction drains instead of
// asserting; unknown message classes are logged and dropped instead of aborting the process.
#pragma once
#include "../async_exec.h"
#include "listener.h"
#include "route.h"
namespace ssb {

template <typename Handler>
class MessageClass {
 public:
  void add(int id, Handler&& handler) {
    std::lock_guard<SpinLock> lk(lock_);
    CHECK(handlers_.count(id) == 0) << "message class " << id << " registered twice";
    handlers_.emplace(id, std::move(handler));
  }
  void remove(int id) {
    std::lock_guard<SpinLock> lk(lock_);
    handlers_.erase(id);
  }
  bool has(int id) {
    std::lock_guard<SpinLock> lk(lock_);
    return handlers_.count(id) > 0;
  }
  // returns a copy so the caller can run it without holding the lock
  bool get(int id, Handler& out) {
    std::lock_guard<SpinLock> lk(lock_);
    auto it = handlers_.find(id);
    if (it == handlers_.end()) return false;
    out = it->second;
    return true;
  }

 private:
  SpinLock lock_;
  std::map<int, Handler> handlers_;
};

template <typename Route>
class Transfer : public Listener {
 public:
  typedef Request::response_call_back_t msgrsp_handler_t;
  typedef std::function<void(std::shared_ptr<Request>, Request&)> msgcls_handler_t;
  typedef Route route_t;

  Transfer() : Listener(nullptr), route_(new Route) {
    // the route owns the zmq context (ref: Route.h:23); the inbox socket is created on it
    init_receiver(route_->zmq_ctx());
  }
  ~Transfer() override { shutdown(); }

  void shutdown() {
    if (down_) return;
    down_ = true;
    service_end();
    async_channel_.reset();  // close + join the handler pool
    {
      std::lock_guard<SpinLock> lk(handlers_lock_);
      if (!msg_handlers_.empty()) LOG(WARNING) << msg_handlers_.size() << " response callbacks were never consumed";
      msg_handlers_.clear();
    }
    if (receiver_) {
      int linger = 0;
      zmq_setsockopt(receiver_, ZMQ_LINGER, &linger, sizeof(linger));
      zmq_close(receiver_);
      receiver_ = nullptr;
    }
    route_.reset();  // closes the PUSH sockets and terminates the context
  }

  void init_async_channel(int thread_num) {
    CHECK(!async_channel_) << "async channel already started";
    AsynExec as(thread_num);
    async_channel_ = as.open();
  }
  std::shared_ptr<AsynExec::channel_t>& async_channel() { return async_channel_; }
  Route& route() { return *route_; }
  MessageClass<msgcls_handler_t>& message_class() { return message_class_; }
  void set_client_id(int id) { client_id_ = id; }
  int client_id() const { return client_id_; }

  // Fire a request at node `to_id`; `request.call_back_handler` runs when the response arrives.
  // ref: transfer.h:75-112
  void send(Request&& request, int to_id) {
    int msg_id = msg_id_counter_.fetch_add(1);
    request.set_msg_id(msg_id);
    if (request.meta.client_id == -3) request.meta.client_id = client_id_;
    CHECK(request.call_back_handler) << "a request needs a response callback";
    {
      std::lock_guard<SpinLock> lk(handlers_lock_);
      msg_handlers_.emplace(msg_id, std::move(request.call_back_handler));
    }
    send_frames(request, to_id);
  }
  // ref: transfer.h:210-225
  void send_response(Request&& request, int to_id) {
    CHECK(request.meta.message_class == -1);
    send_frames(request, to_id);
  }

  // Receiver thread body. ref: transfer.h:119-150
  void main_loop() override {
    for (;;) {
      Package pkg;
      {
        std::lock_guard<std::mutex> lk(receiver_mutex());
        int r = ignore_signal_call(zmq_msg_recv, &pkg.meta.zmg(), receiver(), 0);
        if (r < 0) { LOG(WARNING) << "zmq_msg_recv: " << zmq_strerror(zmq_errno()); return; }
        if (pkg.meta.size() == 0) return;  // poison pill
        if (!zmq_msg_more(&pkg.meta.zmg())) { LOG(ERROR) << "dropping single-frame message"; continue; }
        PCHECK(ignore_signal_call(zmq_msg_recv, &pkg.cont.zmg(), receiver(), 0) >= 0);
      }
      if (pkg.meta.size() != sizeof(MetaMessage)) { LOG(ERROR) << "dropping message with bad meta size " << pkg.meta.size(); continue; }
      auto request = std::make_shared<Request>(std::move(pkg));
      if (request->is_response()) handle_response(request);
      else handle_request(request);
    }
  }

  size_t pending_responses() {
    std::lock_guard<SpinLock> lk(handlers_lock_);
    return msg_handlers_.size();
  }

 protected:
  void send_frames(Request& request, int to_id) {
    Package pkg(request);
    std::lock_guard<std::mutex> lk(route_->send_mutex(to_id));
    void* s = route_->sender(to_id);
    PCHECK(ignore_signal_call(zmq_msg_send, &pkg.meta.zmg(), s, ZMQ_SNDMORE) >= 0) << "send meta to " << to_id;
    PCHECK(ignore_signal_call(zmq_msg_send, &pkg.cont.zmg(), s, 0) >= 0) << "send content to " << to_id;
  }
  // ref: transfer.h:154-179
  void handle_request(std::shared_ptr<Request> request) {
    msgcls_handler_t handler;
    if (!message_class_.get(request->meta.message_class, handler)) {
      LOG(ERROR) << "no handler for message class " << request->meta.message_class << " (dropped)";
      return;
    }
    // pool threads get the channel as a raw pointer: the member is reset by shutdown() on another
    // thread, and the channel outlives its own worker threads (its destructor joins them)
    AsynExec::channel_t* ch = async_channel_.get();
    ch->push([this, ch, handler, request] {
      auto response = std::make_shared<Request>();
      handler(request, *response);
      response->meta.message_class = -1;
      response->meta.message_id = request->meta.message_id;
      response->meta.client_id = request->meta.client_id;
      // an empty response defers the reply (ref: transfer.h:173-177)
      if (response->cont.size() > 0) deliver_response(ch, response, request->meta.client_id, 0);
    });
  }
  // The requester may not be in this node's route yet: the master hands the route table out node by
  // node, so a worker that already has it can reach a server whose copy is still queued behind this
  // very request. Blocking here could starve that route message of a pool thread, so the reply is
  // re-queued (1 ms apart, ~30 s at most) until the route knows the requester.
  void deliver_response(AsynExec::channel_t* ch, std::shared_ptr<Request> response, int to_id, int attempt) {
    if (!route_->has_node(to_id) && attempt < 30000) {
      std::this_thread::sleep_for(std::chrono::milliseconds(1));
      if (ch->push([this, ch, response, to_id, attempt] { deliver_response(ch, response, to_id, attempt + 1); })) return;
    }
    send_response(std::move(*response), to_id);
  }
  // ref: transfer.h:183-208
  void handle_response(std::shared_ptr<Request> response) {
    msgrsp_handler_t cb;
    {
      std::lock_guard<SpinLock> lk(handlers_lock_);
      auto it = msg_handlers_.find(response->meta.message_id);
      if (it == msg_handlers_.end()) { LOG(ERROR) << "response for unknown message id " << response->meta.message_id; return; }
      cb = std::move(it->second);
      msg_handlers_.erase(it);
    }
    async_channel_->push([cb, response] { cb(response); });
  }

  std::unique_ptr<Route> route_;
  std::shared_ptr<AsynExec::channel_t> async_channel_;
  MessageClass<msgcls_handler_t> message_class_;
  SpinLock handlers_lock_;
  std::map<int, msgrsp_handler_t> msg_handlers_;
  std::atomic<int> msg_id_counter_{0};
  int client_id_ = -3;
  bool down_ = false;
};

// Process-wide transfer; first use reads listen_addr / async_exec_num / listen_thread_num,
// binds, starts the handler pool and the receiver threads. ref: transfer.h:270-298
template <typename Route>
inline Transfer<Route>& global_transfer() {
  static Transfer<Route>* transfer = [] {
    auto* t = new Transfer<Route>();
    std::string listen_addr = global_config().get_config("listen_addr", "").to_string();
    int async_exec_num = global_config().get_config("async_exec_num", "4").to_int32();
    int listen_thread_num = global_config().get_config("listen_thread_num", "1").to_int32();
    if (listen_addr.empty()) t->listen();
    else t->listen(listen_addr);
    t->init_async_channel(async_exec_num);
    t->set_thread_num(listen_thread_num);
    t->service_start();
    return t;
  }();
  return *transfer;
}

}  // namespace ssb
