// corpus_loader.h — background text-corpus reader for the GPU workers: parses lines of integer
// word ids (the data format of the reference's tools, src/tools/gen-word2vec-data.py) into
// fixed-size token batches, one reader thread filling a bounded queue so the host never stalls
// the device.  The reference's workers read their shard with scan_file_by_line on the training
// threads themselves (utils/file.h:14-33); here parsing overlaps GPU work and the consumer copies
// each batch straight into pinned memory.
#pragma once
#include <cstdio>
#include <thread>
#include "utils/all.h"
namespace ssb {

class CorpusLoader {
 public:
  // sentences are padded/cut to `sent_len` tokens; pad_id marks padding (SSB_EMPTY_KEY = skipped by dedup)
  // split_n > 1: this reader keeps only the non-empty lines whose index % split_n == split_i — the
  // disjoint per-worker data split that Hadoop streaming gives the reference's workers
  // (tools/hadoop-worker.sh:8-21), without having to cut the file up first.
  CorpusLoader(const std::string& path, int batch_tokens, int sent_len, uint32_t pad_id, int epochs, int queue_depth = 4,
               int split_n = 1, int split_i = 0)
      : path_(path), batch_(batch_tokens), sent_len_(sent_len), pad_(pad_id), epochs_(epochs),
        split_n_(split_n < 1 ? 1 : split_n), split_i_(split_i), queue_(queue_depth) {
    CHECK_GT(batch_, 0);
    CHECK(split_i_ >= 0 && split_i_ < split_n_) << "split index " << split_i_ << " outside [0, " << split_n_ << ")";
    CHECK_GT(sent_len_, 0);
    CHECK_EQ(batch_ % sent_len_, 0) << "batch_tokens must be a multiple of sent_len";
    thread_ = std::thread([this] { run(); });
  }
  ~CorpusLoader() {
    stop_.store(true);
    std::vector<uint32_t> drain;
    while (!done_.load() && queue_.size() > 0) queue_.wait_and_pop(drain);
    // unblock a producer stuck on a full queue, then join
    while (!done_.load()) { if (queue_.size() > 0) queue_.wait_and_pop(drain); else std::this_thread::yield(); }
    if (thread_.joinable()) thread_.join();
  }
  // Copies the next batch into out[batch_tokens]; returns the number of real (non-pad) tokens, or -1 at the end.
  long next(uint32_t* out) {
    std::vector<uint32_t> b;
    queue_.wait_and_pop(b);
    if (b.empty()) { queue_.push(std::vector<uint32_t>()); return -1; }  // keep the end marker for later calls
    long real = (long)b.back();
    b.pop_back();
    std::memcpy(out, b.data(), sizeof(uint32_t) * (size_t)batch_);
    return real;
  }
  int batch_tokens() const { return batch_; }

 private:
  void run() {
    std::vector<uint32_t> cur;
    cur.reserve(batch_ + 1);
    long real = 0;
    for (int ep = 0; ep < epochs_ && !stop_.load(); ++ep) {
      FILE* f = std::fopen(path_.c_str(), "r");
      if (!f) { LOG(ERROR) << "cannot open corpus " << path_; break; }
      LineFileReader reader;
      long line_no = 0;
      while (!stop_.load()) {
        char* line = reader.getline(f);
        if (!line) break;
        int n = 0;
        char* p = line;
        if (split_n_ > 1) {
          // index non-empty lines only, so every reader sees the same numbering
          char* q = p;
          while (*q == ' ' || *q == '\t' || *q == '\r' || *q == '\n') ++q;
          if (*q == '\0') continue;
          if (line_no++ % split_n_ != split_i_) continue;
        }
        while (n < sent_len_) {
          char* end = nullptr;
          unsigned long v = std::strtoul(p, &end, 10);
          if (end == p) break;
          cur.push_back((uint32_t)v);
          ++n; ++real;
          p = end;
        }
        if (n == 0) continue;
        for (; n < sent_len_; ++n) cur.push_back(pad_);
        if ((int)cur.size() == batch_) {
          cur.push_back((uint32_t)real);
          queue_.push(std::move(cur));
          cur = std::vector<uint32_t>();
          cur.reserve(batch_ + 1);
          real = 0;
        }
      }
      std::fclose(f);
    }
    if (!cur.empty() && !stop_.load()) {
      cur.resize(batch_, pad_);
      cur.push_back((uint32_t)real);
      queue_.push(std::move(cur));
    }
    queue_.push(std::vector<uint32_t>());  // end marker
    done_.store(true);
  }
  std::string path_;
  int batch_, sent_len_;
  uint32_t pad_;
  int epochs_;
  int split_n_, split_i_;
  queue_with_capacity<std::vector<uint32_t>> queue_;
  std::thread thread_;
  std::atomic<bool> stop_{false}, done_{false};
};

}  // namespace ssb
