// async_exec.h — thread pool over a closable channel, plus the run-on-n-threads helper.
// ref: /root/reference/src/core/AsynExec.h:38-144.
// AsynExec(n).open() spawns n threads looping `channel.pop(task); task()` and returns a
// shared_ptr<channel> whose deleter closes the channel and joins the threads.
#pragma once
#include <functional>
#include <memory>
#include <thread>
#include <vector>
#include "channel.h"
#include "utils/barrier.h"
namespace ssb {

class AsynExec {
 public:
  typedef std::function<void()> task_t;
  typedef BasicChannel<task_t> channel_t;

  explicit AsynExec(int thread_num = 1) : thread_num_(thread_num) {}
  int thread_num() const { return thread_num_; }

  std::shared_ptr<channel_t> open() {
    auto threads = std::make_shared<std::vector<std::thread>>();
    channel_t* ch = new channel_t();
    for (int i = 0; i < thread_num_; ++i)
      threads->emplace_back([ch] {
        task_t task;
        while (ch->pop(task)) task();
      });
    return std::shared_ptr<channel_t>(ch, [threads](channel_t* c) {
      c->close();
      for (auto& t : *threads)
        if (t.joinable()) t.join();
      delete c;
    });
  }

 private:
  int thread_num_;
};

// Run `task` on n threads of `channel` and block until all n finish. ref: AsynExec.h:123-144
inline void async_exec(int n, const std::function<void(int)>& task, std::shared_ptr<AsynExec::channel_t> channel) {
  CounterBarrier done(n);
  for (int i = 0; i < n; ++i)
    channel->push([i, &task, &done] {
      task(i);
      done.count_down();
    });
  done.block();
}

}  // namespace ssb
