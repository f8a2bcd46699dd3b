// algorithms.h — the worker-side applications on the CPU / ZeroMQ path.
// The reference's src/apps/{word2vec,logistic_regression} are absent from its snapshot
// (SURVEY.md §0); these are written fresh against the BaseAlgorithm contract
// (ref: /root/reference/src/core/framework/SwiftWorker.h:19-57) and the CLI/data format implied
// by src/tools (run_worker.sh, gen-word2vec-data.py: one sentence of integer word ids per line).
//   word2vec : skip-gram negative sampling (Mikolov et al.), minibatch of lines per pull/push
//   lr       : sparse logistic regression, libsvm-like lines `label k:v k:v …`
//   bench    : synthetic sparse-key pull/push (BASELINE config 1 and the keys/batch sweep)
#pragma once
#include <random>
#include <unordered_map>
#include <unordered_set>
#include "../core/swiftsnails.h"
namespace ssb {

typedef index_t key_t_;
typedef VecF val_t_;
typedef VecF grad_t_;
typedef GlobalParamCache<key_t_, val_t_, grad_t_> cache_t_;

inline float sigmoidf(float x) { return 1.f / (1.f + std::exp(-x)); }
inline float softplusf(float x) { return std::max(x, 0.f) + std::log1p(std::exp(-std::fabs(x))); }

struct PsIo {
  // pull/push through the cluster, or a local stand-in when local_train=1
  bool local = false;
  int dim = 1;
  float lr = 0.05f, init_scale = 0.01f;
  uint64_t seed = 0x5EED;
  std::unordered_map<key_t_, val_t_> local_store;  // the "server" of single-process mode
  void pull(const std::vector<key_t_>& keys, cache_t_& cache) {
    if (!local) { global_pull_access<key_t_, val_t_, grad_t_>().pull_with_barrier(keys, cache); return; }
    rwlock_write_guard g(cache.rwlock());
    for (auto k : keys) {
      auto& p = local_store[k];
      if (p.size() == 0) { p.init(dim); for (int j = 0; j < dim; ++j) p[j] = row_init_value(k, j, seed, init_scale); }
      cache.params()[k] = p;
      auto& gr = cache.grads()[k];
      gr.init(dim);
    }
  }
  void push(const std::vector<key_t_>& keys, cache_t_& cache) {
    if (!local) { global_push_access<key_t_, val_t_, grad_t_>().push_with_barrier(keys, cache); return; }
    rwlock_write_guard g(cache.rwlock());
    for (auto k : keys) {
      auto* gr = cache.grads().find(k);
      auto it = local_store.find(k);
      if (!gr || it == local_store.end() || (int)gr->size() != dim) continue;
      for (int j = 0; j < dim; ++j) it->second[j] -= lr * (*gr)[j];
      gr->reset();
    }
  }
};

struct TrainStats {
  double loss_sum = 0, loss_cnt = 0, words = 0, pull_keys = 0, push_keys = 0;
  double train_s = 0, pull_s = 0, push_s = 0;
};

// ------------------------------------------------------------------ word2vec
class Word2VecAlg : public BaseAlgorithm<key_t_, val_t_, grad_t_, std::vector<key_t_>> {
 public:
  Word2VecAlg() {
    auto& c = global_config();
    dim_ = c.get_config("dim", "128").to_int32();
    vocab_ = (uint32_t)c.get_config("vocab", "1000000").to_int64();
    window_ = c.get_config("window", "5").to_int32();
    negative_ = c.get_config("negative", "5").to_int32();
    batch_lines_ = c.get_config("minibatch_lines", "64").to_int32();
    num_iters_ = c.get_config("num_iters", "1").to_int32();
    io_.dim = dim_;
    io_.lr = c.get_config("learning_rate", "0.05").to_float();
    io_.init_scale = c.get_config("init_scale", "0").to_float();
    if (io_.init_scale == 0.f) io_.init_scale = 1.0f / dim_;
    rng_.seed(c.get_config("seed", "1").to_int32());
  }
  rcd_t parse_record(const std::string& line) override {
    rcd_t out;
    for (auto& tok : split(line, " \t")) out.push_back((key_t_)std::stoul(tok));
    return out;
  }
  void train() override {
    io_.local = local_train();
    Timer total;
    for (int iter = 0; iter < num_iters_; ++iter) {
      FILE* f = std::fopen(data_path().c_str(), "r");
      PCHECK(f != nullptr) << "cannot open " << data_path();
      std::mutex mut;
      std::vector<std::string> lines;
      scan_file_by_batch(f, mut, (size_t)batch_lines_, [this](std::vector<std::string>& batch) { train_batch(batch); });
      std::fclose(f);
    }
    stats.train_s = total.elapsed();
    LOG(WARNING) << "word2vec: words=" << stats.words << " time=" << stats.train_s << "s words/s=" << stats.words / stats.train_s
                 << " mean_loss=" << (stats.loss_cnt ? stats.loss_sum / stats.loss_cnt : 0.0);
  }
  TrainStats stats;

 private:
  void train_batch(std::vector<std::string>& lines) {
    std::vector<rcd_t> sents;
    std::unordered_set<key_t_> keyset;
    std::vector<std::vector<key_t_>> negs;
    for (auto& l : lines) {
      rcd_t s = parse_record(l);
      if (s.size() < 2) continue;
      for (auto w : s) { keyset.insert(w); keyset.insert(w + vocab_); }
      sents.push_back(std::move(s));
    }
    // negatives are drawn per (centre, context) pair; sample them up front so they can be pulled
    std::uniform_int_distribution<uint32_t> uni(0, vocab_ - 1);
    std::vector<key_t_> neg_pool;
    size_t n_pairs_max = 0;
    for (auto& s : sents) n_pairs_max += s.size() * 2 * window_;
    neg_pool.resize(n_pairs_max * negative_);
    for (auto& k : neg_pool) { k = uni(rng_) + vocab_; keyset.insert(k); }
    std::vector<key_t_> keys(keyset.begin(), keyset.end());
    Timer t;
    io_.pull(keys, cache_);
    stats.pull_s += t.elapsed();
    stats.pull_keys += keys.size();
    auto& params = cache_.params();
    auto& grads = cache_.grads();
    size_t np = 0;
    std::vector<float> da(dim_);
    for (auto& s : sents) {
      for (size_t i = 0; i < s.size(); ++i) {
        int eff = window_ - (int)(rng_() % (uint32_t)window_);
        val_t_& a = params[s[i]];
        grad_t_& ga = grads[s[i]];
        if (ga.size() == 0) ga.init(dim_);
        for (int j = -eff; j <= eff; ++j) {
          if (j == 0 || (long)i + j < 0 || i + j >= s.size()) continue;
          std::fill(da.begin(), da.end(), 0.f);
          auto step = [&](key_t_ out_key, float label) {
            val_t_& o = params[out_key];
            grad_t_& go = grads[out_key];
            if (go.size() == 0) go.init(dim_);
            float z = 0;
            for (int d = 0; d < dim_; ++d) z += a[d] * o[d];
            float g = sigmoidf(z) - label;
            stats.loss_sum += label > 0 ? softplusf(-z) : softplusf(z);
            for (int d = 0; d < dim_; ++d) { da[d] += g * o[d]; go[d] += g * a[d]; }
          };
          key_t_ ctx = s[i + j] + vocab_;
          step(ctx, 1.f);
          for (int k = 0; k < negative_; ++k) {
            key_t_ nk = neg_pool[np * negative_ + k];
            if (nk != ctx) step(nk, 0.f);
          }
          for (int d = 0; d < dim_; ++d) ga[d] += da[d];
          ++np;
          stats.loss_cnt += 1;
        }
        stats.words += 1;
      }
    }
    t.start();
    io_.push(keys, cache_);
    stats.push_s += t.elapsed();
    stats.push_keys += keys.size();
    cache_.clear();
  }
  int dim_, window_, negative_, batch_lines_, num_iters_;
  uint32_t vocab_;
  PsIo io_;
  cache_t_ cache_;
  std::mt19937 rng_;
};

// ------------------------------------------------------------------ logistic regression
struct LrRecord { float label = 0; std::vector<std::pair<key_t_, float>> feats; };

class LogisticRegressionAlg : public BaseAlgorithm<key_t_, val_t_, grad_t_, LrRecord> {
 public:
  LogisticRegressionAlg() {
    auto& c = global_config();
    batch_lines_ = c.get_config("minibatch_lines", "256").to_int32();
    num_iters_ = c.get_config("num_iters", "1").to_int32();
    io_.dim = 1;
    io_.lr = c.get_config("learning_rate", "0.05").to_float();
    io_.init_scale = c.get_config("init_scale", "0.0").to_float();
  }
  // `label key:value key:value …` (value defaults to 1)
  rcd_t parse_record(const std::string& line) override {
    rcd_t r;
    auto toks = split(line, " \t");
    if (toks.empty()) return r;
    r.label = std::stof(toks[0]) > 0 ? 1.f : 0.f;
    for (size_t i = 1; i < toks.size(); ++i) {
      auto kv = key_value_split(toks[i], ":");
      r.feats.emplace_back((key_t_)std::stoul(kv.first), kv.second.empty() ? 1.f : std::stof(kv.second));
    }
    return r;
  }
  void train() override {
    io_.local = local_train();
    Timer total;
    for (int iter = 0; iter < num_iters_; ++iter) {
      FILE* f = std::fopen(data_path().c_str(), "r");
      PCHECK(f != nullptr) << "cannot open " << data_path();
      std::mutex mut;
      scan_file_by_batch(f, mut, (size_t)batch_lines_, [this](std::vector<std::string>& batch) { train_batch(batch); });
      std::fclose(f);
    }
    stats.train_s = total.elapsed();
    LOG(WARNING) << "lr: examples=" << stats.words << " time=" << stats.train_s << "s mean_loss="
                 << (stats.loss_cnt ? stats.loss_sum / stats.loss_cnt : 0.0);
  }
  TrainStats stats;

 private:
  void train_batch(std::vector<std::string>& lines) {
    std::vector<rcd_t> recs;
    std::unordered_set<key_t_> keyset;
    for (auto& l : lines) {
      rcd_t r = parse_record(l);
      if (r.feats.empty()) continue;
      for (auto& f : r.feats) keyset.insert(f.first);
      recs.push_back(std::move(r));
    }
    std::vector<key_t_> keys(keyset.begin(), keyset.end());
    io_.pull(keys, cache_);
    stats.pull_keys += keys.size();
    auto& params = cache_.params();
    auto& grads = cache_.grads();
    for (auto& r : recs) {
      float z = 0;
      for (auto& f : r.feats) z += params[f.first][0] * f.second;
      float g = sigmoidf(z) - r.label;
      stats.loss_sum += r.label > 0 ? softplusf(-z) : softplusf(z);
      stats.loss_cnt += 1;
      stats.words += 1;
      for (auto& f : r.feats) {
        grad_t_& gr = grads[f.first];
        if (gr.size() == 0) gr.init(1);
        gr[0] += g * f.second;
      }
    }
    io_.push(keys, cache_);
    stats.push_keys += keys.size();
    cache_.clear();
  }
  int batch_lines_, num_iters_;
  PsIo io_;
  cache_t_ cache_;
};

// ------------------------------------------------------------------ pull/push benchmark
class PullPushBenchAlg : public BaseAlgorithm<key_t_, val_t_, grad_t_, int> {
 public:
  PullPushBenchAlg() {
    auto& c = global_config();
    dim_ = c.get_config("dim", "64").to_int32();
    keys_per_batch_ = c.get_config("bench_keys", "100000").to_int32();
    iters_ = c.get_config("num_iters", "5").to_int32();
    key_space_ = (uint32_t)c.get_config("bench_key_space", "10000000").to_int64();
    io_.dim = dim_;
  }
  rcd_t parse_record(const std::string&) override { return 0; }
  void train() override {
    io_.local = local_train();
    std::mt19937 rng(1234 + (io_.local ? 0 : ps_transfer().client_id()));
    std::unordered_set<key_t_> ks;
    while ((int)ks.size() < keys_per_batch_) ks.insert(rng() % key_space_);
    std::vector<key_t_> keys(ks.begin(), ks.end());
    io_.pull(keys, cache_);  // warm-up pull creates the rows
    double pull_s = 0, push_s = 0;
    for (int it = 0; it < iters_; ++it) {
      Timer t;
      io_.pull(keys, cache_);
      pull_s += t.elapsed();
      {
        rwlock_write_guard g(cache_.rwlock());
        for (auto k : keys) {
          auto& gr = cache_.grads()[k];
          if ((int)gr.size() != dim_) gr.init(dim_);
          for (int d = 0; d < dim_; ++d) gr[d] = 0.001f;
        }
      }
      t.start();
      io_.push(keys, cache_);
      push_s += t.elapsed();
    }
    stats.pull_keys = (double)keys.size() * iters_;
    stats.push_keys = stats.pull_keys;
    stats.pull_s = pull_s;
    stats.push_s = push_s;
    std::printf("BENCH pull_keys_per_s %.1f push_keys_per_s %.1f keys %zu dim %d iters %d\n",
                stats.pull_keys / pull_s, stats.push_keys / push_s, keys.size(), dim_, iters_);
    std::fflush(stdout);
  }
  TrainStats stats;

 private:
  int dim_, keys_per_batch_, iters_;
  uint32_t key_space_;
  PsIo io_;
  cache_t_ cache_;
};

// ------------------------------------------------------------------ role runners
typedef SparseTable<key_t_, RowParam> table_t_;

inline void run_master() { SwiftMaster master; master(); }

inline void run_server() {
  auto& c = global_config();
  int dim = c.get_config("dim", "128").to_int32();
  std::string update = c.get_config("update", "sgd").to_string();
  float lr = c.get_config("learning_rate", "0.05").to_float();
  float init_scale = c.get_config("init_scale", "0").to_float();
  if (init_scale == 0.f) init_scale = c.has("init_scale") ? 0.f : 1.0f / dim;
  uint64_t seed = (uint64_t)c.get_config("table_seed", "24301").to_int64();
  RowPullMethod<table_t_> pull(dim, update == "adagrad", seed, init_scale);
  if (update == "adagrad") {
    AdaGradPushMethod<table_t_> push(lr, c.get_config("adagrad_eps", "1e-8").to_float());
    SwiftServer<key_t_, RowParam, RowPullMethod<table_t_>, AdaGradPushMethod<table_t_>> server(pull, push);
    server();
  } else {
    SgdPushMethod<table_t_> push(lr);
    SwiftServer<key_t_, RowParam, RowPullMethod<table_t_>, SgdPushMethod<table_t_>> server(pull, push);
    server();
  }
}

inline int run_worker(const std::string& app, const std::string& data) {
  if (app == "word2vec") {
    Word2VecAlg alg;
    alg.set_data_path(data);
    SwiftWorker<Word2VecAlg> worker(alg);
    worker();
    std::printf("RESULT words %.0f train_s %.6f words_per_s %.1f mean_loss %.6f pull_s %.6f push_s %.6f pull_keys %.0f\n",
                alg.stats.words, alg.stats.train_s, alg.stats.words / alg.stats.train_s,
                alg.stats.loss_cnt ? alg.stats.loss_sum / alg.stats.loss_cnt : 0.0, alg.stats.pull_s, alg.stats.push_s,
                alg.stats.pull_keys);
  } else if (app == "lr") {
    LogisticRegressionAlg alg;
    alg.set_data_path(data);
    SwiftWorker<LogisticRegressionAlg> worker(alg);
    worker();
    std::printf("RESULT examples %.0f train_s %.6f mean_loss %.6f\n", alg.stats.words, alg.stats.train_s,
                alg.stats.loss_cnt ? alg.stats.loss_sum / alg.stats.loss_cnt : 0.0);
  } else if (app == "bench") {
    PullPushBenchAlg alg;
    SwiftWorker<PullPushBenchAlg> worker(alg);
    worker();
  } else {
    LOG(ERROR) << "unknown app '" << app << "' (word2vec | lr | bench)";
    return 2;
  }
  std::fflush(stdout);
  return 0;
}

}  // namespace ssb
