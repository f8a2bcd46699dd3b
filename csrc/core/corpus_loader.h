// corpus_loader.h — parallel text-corpus reader for the GPU workers: parses lines of integer word ids (the
// data format of the reference's tools, src/tools/gen-word2vec-data.py) into fixed-size token batches.
//
// The reference's workers read their shard with scan_file_by_line on the training threads themselves
// (utils/file.h:14-33: one FILE*, one mutex, one line at a time). A B200 step eats ~1 G tokens/s, ten times
// what one parsing thread delivers, so here the file is mapped once, the starts of its sentence lines are
// indexed in one memchr pass, and N parser threads fill an ordered ring of batches ahead of the consumer,
// which copies each batch straight into pinned memory while the GPU works on the previous one.
//
// A "sentence line" is a line whose first non-blank character is a digit; anything else (blank lines,
// comments, headers) is skipped and NOT numbered — swiftsnails_b200.train.count_sentences uses the same rule,
// so every rank agrees on how many minibatches a corpus holds. Sentences are cut / padded to `sent_len`.
#pragma once
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <atomic>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>
#include "utils/all.h"
namespace ssb {

// Read-only view of a whole file (mmap; an empty or unmappable file gives an empty view).
class MappedFile {
 public:
  explicit MappedFile(const std::string& path) {
    fd_ = ::open(path.c_str(), O_RDONLY);
    if (fd_ < 0) return;
    struct stat st;
    if (::fstat(fd_, &st) != 0 || st.st_size <= 0) return;
    void* p = ::mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE | MAP_POPULATE, fd_, 0);   // pre-fault: one bulk pass, not a fault per page
    if (p == MAP_FAILED) return;
    data_ = static_cast<const char*>(p);
    size_ = (size_t)st.st_size;
    ::madvise(p, size_, MADV_SEQUENTIAL);
  }
  ~MappedFile() {
    if (data_) ::munmap(const_cast<char*>(data_), size_);
    if (fd_ >= 0) ::close(fd_);
  }
  MappedFile(const MappedFile&) = delete;
  MappedFile& operator=(const MappedFile&) = delete;
  bool opened() const { return fd_ >= 0; }
  const char* data() const { return data_; }
  size_t size() const { return size_; }

 private:
  int fd_ = -1;
  const char* data_ = nullptr;
  size_t size_ = 0;
};

inline bool corpus_is_blank(char c) { return c == ' ' || c == '\t' || c == '\r'; }

// First non-blank character of the line [p, e) is a digit.
inline bool corpus_is_sentence(const char* p, const char* e) {
  while (p < e && corpus_is_blank(*p)) ++p;
  return p < e && *p >= '0' && *p <= '9';
}

// Parses up to max_tokens unsigned integers of the line [p, e) into out; stops at the first thing that is not
// a number. Returns the number of tokens written.
inline int corpus_parse_line(const char* p, const char* e, uint32_t* out, int max_tokens) {
  int n = 0;
  while (n < max_tokens) {
    while (p < e && corpus_is_blank(*p)) ++p;
    if (p >= e || *p < '0' || *p > '9') break;
    uint32_t v = 0;
    while (p < e && *p >= '0' && *p <= '9') { v = v * 10u + (uint32_t)(*p - '0'); ++p; }
    out[n++] = v;
  }
  return n;
}

class CorpusLoader {
 public:
  // sentences are padded/cut to `sent_len` tokens; pad_id marks padding (SSB_EMPTY_KEY = skipped by dedup).
  // split_n > 1: this reader keeps only the sentence lines whose index % split_n == split_i — the disjoint
  // per-worker data split that Hadoop streaming gives the reference's workers (tools/hadoop-worker.sh:8-21),
  // without having to cut the file up first. max_id > 0: token ids >= max_id are replaced by padding and
  // counted (an id >= vocab would alias another word's output vector). threads <= 0: chosen from the host.
  CorpusLoader(const std::string& path, int batch_tokens, int sent_len, uint32_t pad_id, int epochs, int queue_depth = 4,
               int split_n = 1, int split_i = 0, int threads = 0, uint32_t max_id = 0)
      : file_(path), batch_(batch_tokens), sent_len_(sent_len), pad_(pad_id), epochs_(epochs < 1 ? 1 : epochs),
        max_id_(max_id), depth_(queue_depth < 2 ? 2 : queue_depth) {
    CHECK_GT(batch_, 0);
    CHECK_GT(sent_len_, 0);
    CHECK_EQ(batch_ % sent_len_, 0) << "batch_tokens must be a multiple of sent_len";
    if (split_n < 1) split_n = 1;
    CHECK(split_i >= 0 && split_i < split_n) << "split index " << split_i << " outside [0, " << split_n << ")";
    if (!file_.opened()) LOG(ERROR) << "cannot open corpus " << path;
    index_lines(split_n, split_i);
    lines_per_batch_ = batch_ / sent_len_;
    const long long total = (long long)epochs_ * (long long)starts_.size();
    n_batches_ = (total + lines_per_batch_ - 1) / lines_per_batch_;
    if (threads <= 0) {
      const char* env = std::getenv("SSB200_LOADER_THREADS");
      threads = env ? std::atoi(env) : 0;
      if (threads <= 0) threads = (int)std::min<unsigned>(8u, std::max(1u, std::thread::hardware_concurrency() / 2));
    }
    if ((long long)threads > n_batches_) threads = (int)std::max<long long>(1, n_batches_);
    if (depth_ < threads + 2) depth_ = threads + 2;     // every parser thread can be a batch ahead of the consumer
    slots_.resize(depth_);
    for (auto& s : slots_) { s.data.resize(batch_); s.batch = -1; }
    for (int t = 0; t < threads; ++t) workers_.emplace_back([this] { run(); });
  }
  ~CorpusLoader() {
    {
      std::lock_guard<std::mutex> lk(m_);
      stop_ = true;
    }
    cv_free_.notify_all();
    cv_ready_.notify_all();
    for (auto& t : workers_) if (t.joinable()) t.join();
  }
  // Copies the next batch into out[batch_tokens]; returns the number of real (non-pad) tokens, or -1 at the end.
  long next(uint32_t* out) {
    std::unique_lock<std::mutex> lk(m_);
    if (consumed_ >= n_batches_) return -1;
    Slot& s = slots_[(size_t)(consumed_ % depth_)];
    cv_ready_.wait(lk, [&] { return s.batch == consumed_ || stop_; });
    if (s.batch != consumed_) return -1;
    lk.unlock();
    std::memcpy(out, s.data.data(), sizeof(uint32_t) * (size_t)batch_);   // the slot is ours until consumed_ moves on
    const long real = s.real;
    lk.lock();
    ++consumed_;
    lk.unlock();
    cv_free_.notify_all();
    return real;
  }
  int batch_tokens() const { return batch_; }
  long long sentences() const { return (long long)starts_.size(); }       // sentence lines of this reader's share, per pass
  long long batches() const { return n_batches_; }
  long long bad_tokens() const { return bad_tokens_.load(); }             // ids >= max_id replaced by padding so far

 private:
  struct Slot {
    std::vector<uint32_t> data;
    long real = 0;
    long long batch = -1;
  };
  void index_lines(int split_n, int split_i) {
    const char* p = file_.data();
    const char* end = p ? p + file_.size() : nullptr;
    long long line_no = 0;
    while (p && p < end) {
      const char* nl = static_cast<const char*>(std::memchr(p, '\n', (size_t)(end - p)));
      const char* e = nl ? nl : end;
      if (corpus_is_sentence(p, e)) {
        if (line_no % split_n == split_i) { starts_.push_back((size_t)(p - file_.data())); }
        ++line_no;
      }
      p = e + 1;
    }
  }
  void parse_batch(long long j, Slot& s) {
    const char* base = file_.data();
    const char* end = base + file_.size();
    const long long total = (long long)epochs_ * (long long)starts_.size();
    long real = 0, bad = 0;
    uint32_t* out = s.data.data();
    for (int k = 0; k < lines_per_batch_; ++k, out += sent_len_) {
      const long long g = j * lines_per_batch_ + k;
      int n = 0;
      if (g < total) {
        const char* p = base + starts_[(size_t)(g % (long long)starts_.size())];
        const char* nl = static_cast<const char*>(std::memchr(p, '\n', (size_t)(end - p)));
        n = corpus_parse_line(p, nl ? nl : end, out, sent_len_);
        if (max_id_ > 0)
          for (int x = 0; x < n; ++x)
            if (out[x] >= max_id_) { out[x] = pad_; ++bad; }
      }
      real += n;
      for (; n < sent_len_; ++n) out[n] = pad_;
    }
    s.real = real - bad;
    if (bad) bad_tokens_.fetch_add(bad);
  }
  void run() {
    for (;;) {
      long long j;
      Slot* s;
      {
        std::unique_lock<std::mutex> lk(m_);
        j = next_batch_;
        if (j >= n_batches_ || stop_) return;
        ++next_batch_;
        // slot j % depth is free once the consumer has taken batch j - depth
        cv_free_.wait(lk, [&] { return j < consumed_ + depth_ || stop_; });
        if (stop_) return;
        s = &slots_[(size_t)(j % depth_)];
      }
      parse_batch(j, *s);
      {
        std::lock_guard<std::mutex> lk(m_);
        s->batch = j;
      }
      cv_ready_.notify_all();
    }
  }
  MappedFile file_;
  int batch_, sent_len_;
  uint32_t pad_;
  int epochs_;
  uint32_t max_id_;
  int depth_;
  int lines_per_batch_ = 1;
  long long n_batches_ = 0;
  std::vector<size_t> starts_;
  std::vector<Slot> slots_;
  std::vector<std::thread> workers_;
  std::mutex m_;
  std::condition_variable cv_free_, cv_ready_;
  long long next_batch_ = 0, consumed_ = 0;
  bool stop_ = false;
  std::atomic<long long> bad_tokens_{0};
};

// Word frequencies of a corpus (for the unigram^0.75 negative-sampling table: the noise distribution has to
// come from the corpus itself). counts[w] += occurrences of word id w < vocab in the sentence lines; ids
// >= vocab are counted in the return value. The file is cut into `threads` byte ranges at line starts, each
// range is parsed into its own histogram, and the histograms are added up.
inline long long corpus_count_words(const std::string& path, uint32_t vocab, unsigned long long* counts, int threads) {
  MappedFile file(path);
  if (!file.opened()) return -1;
  if (!file.data()) return 0;
  if (threads <= 0) threads = (int)std::min<unsigned>(8u, std::max(1u, std::thread::hardware_concurrency() / 2));
  const char* base = file.data();
  const char* end = base + file.size();
  std::vector<const char*> cut((size_t)threads + 1, end);
  cut[0] = base;
  for (int t = 1; t < threads; ++t) {
    const char* p = base + file.size() / (size_t)threads * (size_t)t;
    const char* nl = static_cast<const char*>(std::memchr(p, '\n', (size_t)(end - p)));
    cut[(size_t)t] = nl ? nl + 1 : end;
    if (cut[(size_t)t] < cut[(size_t)t - 1]) cut[(size_t)t] = cut[(size_t)t - 1];
  }
  std::vector<std::vector<unsigned long long>> hist((size_t)threads);
  std::vector<long long> bad((size_t)threads, 0);
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t) {
    pool.emplace_back([&, t] {
      auto& h = hist[(size_t)t];
      h.assign(vocab, 0ull);
      std::vector<uint32_t> buf(1 << 16);
      const char* p = cut[(size_t)t];
      const char* stop = cut[(size_t)t + 1];
      while (p < stop) {
        const char* nl = static_cast<const char*>(std::memchr(p, '\n', (size_t)(end - p)));
        const char* e = nl ? nl : end;
        if (corpus_is_sentence(p, e)) {
          const char* q = p;
          for (;;) {                                   // a line may hold more tokens than the buffer: parse it in pieces
            const int n = corpus_parse_line(q, e, buf.data(), (int)buf.size());
            for (int x = 0; x < n; ++x) { if (buf[(size_t)x] < vocab) ++h[buf[(size_t)x]]; else ++bad[(size_t)t]; }
            if (n < (int)buf.size()) break;
            // advance q past the n tokens just parsed
            for (int x = 0; x < n; ++x) {
              while (q < e && corpus_is_blank(*q)) ++q;
              while (q < e && *q >= '0' && *q <= '9') ++q;
            }
          }
        }
        p = e + 1;
      }
    });
  }
  for (auto& th : pool) th.join();
  long long total_bad = 0;
  for (int t = 0; t < threads; ++t) {
    total_bad += bad[(size_t)t];
    for (uint32_t w = 0; w < vocab; ++w) counts[w] += hist[(size_t)t][w];
  }
  return total_bad;
}

}  // namespace ssb
