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
#include <chrono>
#include <cmath>
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

typedef SparseTable<rkey_t, Row> table_t;
static float g_lr = 0.05f;

class RowPull : public PullAccessMethod<table_t, Row, Row> {
 public:
  void init_param(const rkey_t&, Row& param) override { param.init(true); }
  void get_pull_value(const rkey_t&, const Row& param, Row& val) override { val = param; }
};
class RowPushSgd : public PushAccessMethod<table_t, Row, Row> {
 public:
  void merge_push_value(const rkey_t&, Row& a, const Row& b) override { for (size_t i = 0; i < a.size(); ++i) a[i] += b[i]; }
  void apply_push_value(const rkey_t&, Row& param, const Row& grad) override {
    for (size_t i = 0; i < param.size() && i < grad.size(); ++i) param[i] -= g_lr * grad[i];
  }
};

static double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
static float sigm(float x) { return 1.f / (1.f + std::exp(-x)); }
static float softplus(float x) { return std::max(x, 0.f) + std::log1p(std::exp(-std::fabs(x))); }

typedef GlobalParamCache<rkey_t, Row, Row> cache_t;

// ---- word2vec SGNS against the reference worker API -------------------------------------------
class RefWord2Vec : public BaseAlgorithm<rkey_t, Row, Row, std::vector<rkey_t>> {
 public:
  RefWord2Vec() {
    vocab_ = (unsigned)global_config().get_config("vocab").to_int32();
    window_ = global_config().get_config("window").to_int32();
    negative_ = global_config().get_config("negative").to_int32();
    batch_lines_ = global_config().get_config("minibatch_lines").to_int32();
    num_iters_ = global_config().get_config("num_iters").to_int32();
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
  rkey_t sample_negative() {
    double u = (rng_() / 4294967296.0) * neg_cdf_.back();
    size_t i = std::lower_bound(neg_cdf_.begin(), neg_cdf_.end(), u) - neg_cdf_.begin();
    return neg_words_[std::min(i, neg_words_.size() - 1)] + vocab_;
  }
  void train() override {
    build_unigram_table();
    double t0 = now_s();
    for (int it = 0; it < num_iters_; ++it) {
      FILE* f = fopen(data_path().c_str(), "r");
      CHECK(f) << "cannot open data";
      LineFileReader reader;
      std::vector<std::string> batch;
      while (true) {
        char* l = reader.getline(f);
        if (l) batch.emplace_back(l);
        if ((!l && !batch.empty()) || (int)batch.size() == batch_lines_) { train_batch(batch); batch.clear(); }
        if (!l) break;
      }
      fclose(f);
    }
    double dt = now_s() - t0;
    printf("RESULT words %.0f train_s %.6f words_per_s %.1f mean_loss %.6f pull_s %.6f push_s %.6f pull_keys %.0f\n", words_, dt,
           words_ / dt, loss_n_ ? loss_ / loss_n_ : 0.0, pull_s_, push_s_, pull_keys_);
    fflush(stdout);
  }

 private:
  void train_batch(std::vector<std::string>& lines) {
    std::vector<rcd_t> sents;
    std::unordered_set<rkey_t> keys;
    size_t n_centres = 0;
    for (auto& l : lines) {
      rcd_t s = parse_record(l);
      if (s.size() < 2) continue;
      for (auto w : s) { keys.insert(w); keys.insert(w + vocab_); }
      n_centres += s.size();
      sents.push_back(std::move(s));
    }
    // `negative_` negatives per centre word, shared by that centre's context pairs (the same
    // sharing idea as the GPU arm, which shares 16 negatives across a 128-centre tile)
    std::vector<rkey_t> negs(n_centres * negative_);
    for (auto& k : negs) { k = sample_negative(); keys.insert(k); }
    double t = now_s();
    global_pull_access<rkey_t, Row, Row>().pull_with_barrier(keys, cache_);
    pull_s_ += now_s() - t;
    pull_keys_ += keys.size();
    auto& params = cache_.params();
    auto& grads = cache_.grads();
    size_t np = 0;
    std::vector<float> da(g_dim);
    for (auto& s : sents) {
      for (size_t i = 0; i < s.size(); ++i) {
        int eff = window_ - (int)(rng_() % (unsigned)window_);
        Row& a = params[s[i]];
        Row& ga = grads[s[i]];
        if (ga.size() == 0) ga.init();
        for (int j = -eff; j <= eff; ++j) {
          if (j == 0 || (long)i + j < 0 || i + j >= s.size()) continue;
          std::fill(da.begin(), da.end(), 0.f);
          auto step = [&](rkey_t ok, float label) {
            Row& o = params[ok];
            Row& go = grads[ok];
            if (go.size() == 0) go.init();
            float z = 0;
            for (int d = 0; d < g_dim; ++d) z += a[d] * o[d];
            float g = sigm(z) - label;
            loss_ += label > 0 ? softplus(-z) : softplus(z);
            for (int d = 0; d < g_dim; ++d) { da[d] += g * o[d]; go[d] += g * a[d]; }
          };
          rkey_t ctx = s[i + j] + vocab_;
          step(ctx, 1.f);
          for (int k = 0; k < negative_; ++k) { rkey_t nk = negs[np * negative_ + k]; if (nk != ctx) step(nk, 0.f); }
          for (int d = 0; d < g_dim; ++d) ga[d] += da[d];
          loss_n_ += 1;
        }
        ++np;
        words_ += 1;
      }
    }
    t = now_s();
    global_push_access<rkey_t, Row, Row>().push_with_barrier(keys, cache_);
    push_s_ += now_s() - t;
    params.clear();
    grads.clear();
  }
  unsigned vocab_;
  int window_, negative_, batch_lines_, num_iters_;
  cache_t cache_;
  std::mt19937 rng_{12345};
  std::vector<rkey_t> neg_words_;
  std::vector<double> neg_cdf_;
  double words_ = 0, loss_ = 0, loss_n_ = 0, pull_s_ = 0, push_s_ = 0, pull_keys_ = 0;
};

// ---- synthetic pull/push benchmark (BASELINE config 1 / sweep) --------------------------------
class RefBench : public BaseAlgorithm<rkey_t, Row, Row, int> {
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
    global_pull_access<rkey_t, Row, Row>().pull_with_barrier(keys, cache);  // warm-up: creates the rows
    double pull_s = 0, push_s = 0;
    for (int it = 0; it < iters; ++it) {
      double t = now_s();
      global_pull_access<rkey_t, Row, Row>().pull_with_barrier(keys, cache);
      pull_s += now_s() - t;
      for (auto k : keys) { Row& g = cache.grads()[k]; g.d.assign(g_dim, 0.001f); }
      t = now_s();
      global_push_access<rkey_t, Row, Row>().push_with_barrier(keys, cache);
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
  if (role == "master") {
    SwiftMaster master(conf);
    master();
  } else if (role == "server") {
    SwiftServer<rkey_t, Row, Row, RowPull, RowPushSgd> server(conf);
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
