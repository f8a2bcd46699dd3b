// ref_ps.cpp — master / server / worker binaries for the REFERENCE arm of the benchmark.
//
// Everything below `#include "swiftsnails.h"` talks to the UNMODIFIED reference headers copied to
// baseline/_ref/src (FrankSzn/SwiftSnails). The reference's own apps (src/apps/word2vec, …) are
// absent from its snapshot, so the application layer — row type, access methods,
// the word2vec / pull-push algorithms and main() — is written here against the reference's public
// API: BaseAlgorithm, SwiftWorker, SwiftServer, SwiftMaster, PullAccessMethod / PushAccessMethod,
// GlobalParamCache, global_pull_access().pull_with_barrier / global_push_access().push_with_barrier.
// None of swiftsnails_b200's code is on this path.
//
//   ref_ps -role master|server|worker -config <conf> [-data <file>] [-app word2vec|bench]
#include <atomic>
#include <chrono>
#include <cmath>
#include <memory>
#include <mutex>
#include <random>
#include <unordered_map>
#include <unordered_set>
#include "swiftsnails.h"

using namespace swift_snails;
typedef index_t rkey_t;

static int g_dim = 128;

// Row type: float vector with the members the reference requires of Val / Grad
// (BinaryBuffer << >>, ostream <<, reset(), init(bool)) — SURVEY.md §2.12(1).
struct Row {
  std::vector<float> d;
  void init(bool random = false) {
    d.assign(g_dim, 0.f);
    if (random) for (auto& x : d) x = (float)((rand() / (double)RAND_MAX - 0.5) / g_dim);  // vec1.h:223-226
  }
  void reset() { std::fill(d.begin(), d.end(), 0.f); }
  size_t size() const { return d.size(); }
  float& operator[](size_t i) { return d[i]; }
  const float& operator[](size_t i) const { return d[i]; }
};
static BinaryBuffer& operator<<(BinaryBuffer& bb, const Row& r) {
  bb << (unsigned int)r.d.size();
  for (float x : r.d) bb << x;
  return bb;
}
static BinaryBuffer& operator>>(BinaryBuffer& bb, Row& r) {
  unsigned int n = 0;
  bb >> n;
  r.d.resize(n);
  for (unsigned int i = 0; i < n; ++i) bb >> r.d[i];
  return bb;
}
static std::ostream& operator<<(std::ostream& os, const Row& r) {
  os << "Vec:\t";
  for (float x : r.d) os << x << " ";
  return os;
}

// Server-side parameter of a key: the row plus its AdaGrad accumulator. Only the row is serialised, so
// the accumulator never leaves the server (the reference's stock SwiftServer uses ONE type for the table
// entry and the pulled value — core/system/server/init.h:33,49-69 — so "Param != Val" is expressed by
// what the BinaryBuffer operators carry).
struct AdaRow {
  std::vector<float> d, acc;
  void init(bool random = false) {
    d.assign(g_dim, 0.f);
    if (random) for (auto& x : d) x = (float)((rand() / (double)RAND_MAX - 0.5) / g_dim);  // vec1.h:223-226
  }
  size_t size() const { return d.size(); }
  float& operator[](size_t i) { return d[i]; }
  const float& operator[](size_t i) const { return d[i]; }
};
static BinaryBuffer& operator<<(BinaryBuffer& bb, const AdaRow& r) {
  bb << (unsigned int)r.d.size();
  for (float x : r.d) bb << x;
  return bb;
}
static BinaryBuffer& operator>>(BinaryBuffer& bb, AdaRow& r) {
  unsigned int n = 0;
  bb >> n;
  r.d.resize(n);
  for (unsigned int i = 0; i < n; ++i) bb >> r.d[i];
  return bb;
}
static std::ostream& operator<<(std::ostream& os, const AdaRow& r) {
  os << "Vec:\t";
  for (float x : r.d) os << x << " ";
  if (!r.acc.empty()) { os << "\tAcc:\t"; for (float x : r.acc) os << x << " "; }
  return os;
}

typedef SparseTable<rkey_t, AdaRow> table_t;
static float g_lr = 0.05f;
static bool g_adagrad = true;
static float g_eps = 1e-8f;

class RowPull : public PullAccessMethod<table_t, AdaRow, AdaRow> {
 public:
  void init_param(const rkey_t&, AdaRow& param) override { param.init(true); }
  void get_pull_value(const rkey_t&, const AdaRow& param, AdaRow& val) override { val.d = param.d; }
};
// The user-supplied server update (sparse_access_method.h:36-46): AdaGrad by default — the same rule the
// GPU arm applies in its push kernel — or plain SGD (config `optimizer: sgd`).
class RowPush : public PushAccessMethod<table_t, Row, AdaRow> {
 public:
  void merge_push_value(const rkey_t&, Row& a, const Row& b) override { for (size_t i = 0; i < a.size(); ++i) a[i] += b[i]; }
  void apply_push_value(const rkey_t&, AdaRow& param, const Row& grad) override {
    const size_t n = std::min(param.size(), grad.size());
    if (g_adagrad) {
      if (param.acc.size() != param.d.size()) param.acc.assign(param.d.size(), 0.f);
      for (size_t i = 0; i < n; ++i) {
        const float g = grad[i];
        param.acc[i] += g * g;
        param.d[i] -= g_lr * g / std::sqrt(param.acc[i] + g_eps);
      }
    } else {
      for (size_t i = 0; i < n; ++i) param.d[i] -= g_lr * grad[i];
    }
  }
};

static double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
static float sigm(float x) { return 1.f / (1.f + std::exp(-x)); }
static float softplus(float x) { return std::max(x, 0.f) + std::log1p(std::exp(-std::fabs(x))); }

typedef GlobalParamCache<rkey_t, AdaRow, Row> cache_t;

// ---- word2vec SGNS against the reference worker API -------------------------------------------
// The SAME objective as the GPU arm (swiftsnails_b200/models/word2vec.py, csrc/kernels/sgns_tc.cu): the
// minibatch's token stream is cut into tiles of 128 consecutive centre tokens; a centre's contexts are
// the tokens within its shrunk window inside its sentence; each tile shares `neg_pool` (16) negatives
// drawn from unigram^0.75, each weighted m_i * K / neg_pool for a centre with m_i contexts (K negatives
// per pair in expectation); gradients are summed over the minibatch and pushed once per key.
// Threads: the reference's own idiom — `async_channel_thread_num` threads run scan_file_by_line over one
// FILE* (utils/file.h:14-33) through async_exec (core/AsynExec.h:123-144); each thread collects
// `minibatch_lines` sentences, then pulls, computes and pushes them with its own GlobalParamCache.
class RefWord2Vec : public BaseAlgorithm<rkey_t, AdaRow, Row, std::vector<rkey_t>> {
 public:
  RefWord2Vec() {
    vocab_ = (unsigned)global_config().get_config("vocab").to_int32();
    window_ = global_config().get_config("window").to_int32();
    negative_ = global_config().get_config("negative").to_int32();
    neg_pool_ = global_config().get_config("neg_pool").to_int32();
    batch_lines_ = global_config().get_config("minibatch_lines").to_int32();
    num_iters_ = global_config().get_config("num_iters").to_int32();
    threads_ = global_config().get_config("async_channel_thread_num").to_int32();
    prepopulate_ = global_config().get_config("prepopulate").to_int32();
  }
  // Create every input and output row before the timed region (the GPU arm does the same with its table: a
  // steady-state model whose rows all exist). The reference creates rows lazily on the first pull
  // (PullAccessAgent::get_pull_value, parameter/sparsetable.h:142-149), so pulling every key once does it.
  void populate_all() {
    const rkey_t total = 2u * vocab_;
    const rkey_t chunk = 50000;
    cache_t cache;
    for (rkey_t k0 = 0; k0 < total; k0 += chunk) {
      std::unordered_set<rkey_t> keys;
      for (rkey_t k = k0; k < std::min(total, k0 + chunk); ++k) keys.insert(k);
      global_pull_access<rkey_t, AdaRow, Row>().pull_with_barrier(keys, cache);
      cache.params().clear();
      cache.grads().clear();
    }
  }
  rcd_t parse_record(const std::string& line) override {
    rcd_t out;
    for (auto& t : split(line, " \t")) out.push_back((rkey_t)std::stoul(t));
    return out;
  }
  // unigram^0.75 table from the corpus itself, as classic word2vec does (not timed)
  void build_unigram_table() {
    std::unordered_map<rkey_t, double> cnt;
    FILE* f = fopen(data_path().c_str(), "r");
    CHECK(f) << "cannot open data";
    LineFileReader reader;
    while (char* l = reader.getline(f)) for (auto w : parse_record(l)) cnt[w] += 1;
    fclose(f);
    double acc = 0;
    for (auto& kv : cnt) { acc += std::pow(kv.second, 0.75); neg_words_.push_back(kv.first); neg_cdf_.push_back(acc); }
  }
  struct ThreadState {
    cache_t cache;
    std::mt19937 rng;
    std::vector<std::string> batch;
    double words = 0, loss = 0, loss_n = 0, pull_s = 0, push_s = 0, pull_keys = 0;
    double loss_first = 0, n_first = 0, loss_last = 0, n_last = 0;   // first / last minibatch of this thread
    int batches = 0;
  };
  void train() override {
    build_unigram_table();
    create_async_channel();
    if (prepopulate_) populate_all();
    std::vector<std::unique_ptr<ThreadState>> states;
    for (int t = 0; t < threads_; ++t) { states.emplace_back(new ThreadState); states.back()->rng.seed(12345 + 977 * t); }
    double t0 = now_s();
    for (int it = 0; it < num_iters_; ++it) {
      FILE* f = fopen(data_path().c_str(), "r");
      CHECK(f) << "cannot open data";
      std::mutex file_mut;
      std::atomic<int> next_state{0};
      AsynExec::task_t task = [&] {
        ThreadState& st = *states[next_state++ % threads_];
        scan_file_by_line(f, file_mut, [&](const std::string& line) {
          st.batch.push_back(line);
          if ((int)st.batch.size() == batch_lines_) { train_batch(st); st.batch.clear(); }
        });
        if (!st.batch.empty()) { train_batch(st); st.batch.clear(); }
      };
      async_exec(threads_, std::move(task), async_channel());
      fclose(f);
    }
    double dt = now_s() - t0;
    ThreadState sum;
    for (auto& s : states) {
      sum.words += s->words; sum.loss += s->loss; sum.loss_n += s->loss_n; sum.pull_s += s->pull_s; sum.push_s += s->push_s;
      sum.pull_keys += s->pull_keys; sum.loss_first += s->loss_first; sum.n_first += s->n_first; sum.loss_last += s->loss_last;
      sum.n_last += s->n_last; sum.batches += s->batches;
    }
    printf("RESULT words %.0f train_s %.6f words_per_s %.1f mean_loss %.6f loss_first %.6f loss_last %.6f batches %d threads %d "
           "pull_s %.6f push_s %.6f pull_keys %.0f\n",
           sum.words, dt, sum.words / dt, sum.loss_n ? sum.loss / sum.loss_n : 0.0, sum.n_first ? sum.loss_first / sum.n_first : 0.0,
           sum.n_last ? sum.loss_last / sum.n_last : 0.0, sum.batches, threads_, sum.pull_s, sum.push_s, sum.pull_keys);
    fflush(stdout);
  }

 private:
  rkey_t sample_negative(std::mt19937& rng) {
    double u = (rng() / 4294967296.0) * neg_cdf_.back();
    size_t i = std::lower_bound(neg_cdf_.begin(), neg_cdf_.end(), u) - neg_cdf_.begin();
    return neg_words_[std::min(i, neg_words_.size() - 1)] + vocab_;
  }
  void train_batch(ThreadState& st);
  unsigned vocab_;
  int window_, negative_, neg_pool_, batch_lines_, num_iters_, threads_, prepopulate_ = 0;
  std::vector<rkey_t> neg_words_;
  std::vector<double> neg_cdf_;
};

void RefWord2Vec::train_batch(ThreadState& st) {
  const int kTile = 128;
  // the minibatch as one token stream, with the sentence bounds of every position
  std::vector<rkey_t> tok;
  std::vector<std::pair<int, int>> bound;   // [lo, hi) of the sentence holding position p
  std::unordered_set<rkey_t> keys;
  for (auto& l : st.batch) {
    rcd_t s = parse_record(l);
    if (s.size() < 2) continue;
    const int lo = (int)tok.size(), hi = lo + (int)s.size();
    for (auto w : s) { keys.insert(w); keys.insert(w + vocab_); tok.push_back(w); bound.emplace_back(lo, hi); }
  }
  if (tok.empty()) return;
  const int n = (int)tok.size(), tiles = (n + kTile - 1) / kTile;
  std::vector<rkey_t> negs((size_t)tiles * neg_pool_);
  for (auto& k : negs) { k = sample_negative(st.rng); keys.insert(k); }
  double t = now_s();
  global_pull_access<rkey_t, AdaRow, Row>().pull_with_barrier(keys, st.cache);
  st.pull_s += now_s() - t;
  st.pull_keys += keys.size();
  auto& params = st.cache.params();
  auto& grads = st.cache.grads();
  std::vector<float> da(g_dim);
  double bl = 0, bn = 0;
  for (int i = 0; i < n; ++i) {
    const int eff = window_ - (int)(st.rng() % (unsigned)window_);
    const int lo = std::max(i - eff, bound[i].first), hi = std::min(i + eff, bound[i].second - 1);
    const int m_i = hi - lo;                        // contexts of this centre (the range minus itself)
    st.words += 1;
    if (m_i <= 0) continue;
    AdaRow& a = params[tok[i]];
    Row& ga = grads[tok[i]];
    if (ga.size() == 0) ga.init();
    std::fill(da.begin(), da.end(), 0.f);
    auto step = [&](rkey_t ok, float label, float weight) {
      AdaRow& o = params[ok];
      Row& go = grads[ok];
      if (go.size() == 0) go.init();
      float z = 0;
      for (int d = 0; d < g_dim; ++d) z += a[d] * o[d];
      const float g = weight * (sigm(z) - label);
      bl += weight * (label > 0 ? softplus(-z) : softplus(z));
      for (int d = 0; d < g_dim; ++d) { da[d] += g * o[d]; go[d] += g * a[d]; }
    };
    for (int j = lo; j <= hi; ++j) {
      if (j == i) continue;
      step(tok[j] + vocab_, 1.f, 1.f);
      bn += 1;
    }
    const float w = (float)m_i * (float)negative_ / (float)neg_pool_;
    const rkey_t* tn = &negs[(size_t)(i / kTile) * neg_pool_];
    for (int k = 0; k < neg_pool_; ++k) step(tn[k], 0.f, w);
    for (int d = 0; d < g_dim; ++d) ga[d] += da[d];
  }
  st.loss += bl; st.loss_n += bn;
  if (st.batches == 0) { st.loss_first = bl; st.n_first = bn; }
  st.loss_last = bl; st.n_last = bn;
  st.batches += 1;
  t = now_s();
  global_push_access<rkey_t, AdaRow, Row>().push_with_barrier(keys, st.cache);
  st.push_s += now_s() - t;
  params.clear();
  grads.clear();
}

// ---- synthetic pull/push benchmark (BASELINE config 1 / sweep) --------------------------------
class RefBench : public BaseAlgorithm<rkey_t, AdaRow, Row, int> {
 public:
  rcd_t parse_record(const std::string&) override { return 0; }
  void train() override {
    int nkeys = global_config().get_config("bench_keys").to_int32();
    int iters = global_config().get_config("num_iters").to_int32();
    unsigned space = (unsigned)global_config().get_config("bench_key_space").to_int32();
    std::mt19937 rng(99 + global_transfer<ServerWorkerRoute>().client_id());
    std::unordered_set<rkey_t> keys;
    while ((int)keys.size() < nkeys) keys.insert(rng() % space);
    cache_t cache;
    global_pull_access<rkey_t, AdaRow, Row>().pull_with_barrier(keys, cache);  // warm-up: creates the rows
    double pull_s = 0, push_s = 0;
    for (int it = 0; it < iters; ++it) {
      double t = now_s();
      global_pull_access<rkey_t, AdaRow, Row>().pull_with_barrier(keys, cache);
      pull_s += now_s() - t;
      for (auto k : keys) { Row& g = cache.grads()[k]; g.d.assign(g_dim, 0.001f); }
      t = now_s();
      global_push_access<rkey_t, AdaRow, Row>().push_with_barrier(keys, cache);
      push_s += now_s() - t;
    }
    printf("BENCH pull_keys_per_s %.1f push_keys_per_s %.1f keys %zu dim %d iters %d\n", keys.size() * (double)iters / pull_s,
           keys.size() * (double)iters / push_s, keys.size(), g_dim, iters);
    fflush(stdout);
  }
};

int main(int argc, char** argv) {
  fms::CMDLine cmd(argc, argv);
  std::string role = cmd.getValue(cmd.registerParameter("role", "master|server|worker"));
  std::string conf = cmd.getValue(cmd.registerParameter("config", "config file"));
  std::string app = cmd.hasParameter("app") ? cmd.getValue("app") : "word2vec";
  global_config().load_conf(conf);
  global_config().parse();
  g_dim = global_config().get_config("dim").to_int32();
  g_lr = global_config().get_config("learning_rate").to_float();
  g_adagrad = global_config().get_config("optimizer").to_string() != "sgd";
  if (role == "master") {
    SwiftMaster master(conf);
    master();
  } else if (role == "server") {
    SwiftServer<rkey_t, AdaRow, Row, RowPull, RowPush> server(conf);
    server();
  } else {
    std::string data = cmd.hasParameter("data") ? cmd.getValue("data") : "";
    if (app == "bench") {
      RefBench alg;
      SwiftWorker<RefBench> worker(conf, alg);
      worker();
    } else {
      RefWord2Vec alg;
      alg.set_data_path(data);
      SwiftWorker<RefWord2Vec> worker(conf, alg);
      worker();
    }
  }
  fflush(stdout);
  return 0;  // static destructors run the reference's own teardown (~Transfer: poison pills, join)
}
