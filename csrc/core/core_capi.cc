// core_capi.cc — C ABI of the host core for the Python layer (declared in include/ssb200_core.h).
#include <fstream>
#include "../../include/ssb200_core.h"
#include "../apps/algorithms.h"
#include "corpus_loader.h"

using namespace ssb;

extern "C" {

int ssb_core_version(void) { return 100; }

// ---- config ---------------------------------------------------------------------------------
void ssb_config_clear(void) { global_config().clear(); }
int ssb_config_load(const char* path) {
  std::ifstream probe(path);
  if (!probe.good()) return -1;
  global_config().load_conf(path);
  global_config().parse();
  return 0;
}
void ssb_config_set(const char* key, const char* value) { global_config().set(key, value); }
int ssb_config_parse_line(const char* line) { return global_config().parse_line(line) ? 1 : 0; }
int ssb_config_get(const char* key, char* out, int out_len) {
  if (!global_config().has(key)) return -1;
  const std::string& v = global_config().get_config(key).to_string();
  if ((int)v.size() + 1 > out_len) return -2;
  std::memcpy(out, v.c_str(), v.size() + 1);
  return (int)v.size();
}
int ssb_config_keys(char* out, int out_len) {
  std::string all;
  for (auto& kv : global_config().items()) { all += kv.first; all += '\n'; }
  if ((int)all.size() + 1 > out_len) return -2;
  std::memcpy(out, all.c_str(), all.size() + 1);
  return (int)all.size();
}

// ---- hashing / hash fragments ----------------------------------------------------------------
uint64_t ssb_hash_code(uint64_t x) { return get_hash_code(x); }
int ssb_hashfrag_build(int num_nodes, int frag_num, uint32_t* map_out) {
  if (num_nodes < 1 || frag_num < num_nodes) return -1;
  BasicHashFrag<uint32_t> h;
  h.init(num_nodes, frag_num);
  std::memcpy(map_out, h.map_table().data(), sizeof(uint32_t) * frag_num);
  return 0;
}
void ssb_hashfrag_route(const uint32_t* keys, int64_t n, const uint32_t* map, int frag_num, int32_t* node_out) {
  for (int64_t i = 0; i < n; ++i) node_out[i] = (int32_t)map[get_hash_code((uint64_t)keys[i]) % (uint64_t)frag_num];
}

// ---- binary buffer round trip (used by the wire-format tests) ----------------------------------
int ssb_buffer_selftest(void) {
  BinaryBuffer bb;
  bb << (int32_t)-7 << (uint64_t)1234567890123ull << 3.5f << 2.25 << true << (uint8_t)9;
  for (int i = 0; i < 1000; ++i) bb << (int32_t)i;  // grows past the 1 KiB initial capacity
  int32_t a; uint64_t b; float c; double d; bool e; uint8_t f;
  bb >> a >> b >> c >> d >> e >> f;
  if (a != -7 || b != 1234567890123ull || c != 3.5f || d != 2.25 || !e || f != 9) return 1;
  for (int i = 0; i < 1000; ++i) { int32_t v; bb >> v; if (v != i) return 2; }
  return bb.read_finished() ? 0 : 3;
}

// ---- text parameter dump (the reference's checkpoint format) + loader ----------------------------
// line = "<key>\t<value>\n"; a row of dim > 1 prints like the reference's Vec: "Vec:\t v0 v1 … "
// (ref: parameter/sparsetable.h:49-56, utils/vec1.h:106-112); dim 1 prints the scalar.
int64_t ssb_dump_write_text(const char* path, const uint32_t* keys, const float* rows, int64_t n, int dim, int append) {
  FILE* f = std::fopen(path, append ? "a" : "w");
  if (!f) return -1;
  std::vector<char> buf(1 << 20);
  std::setvbuf(f, buf.data(), _IOFBF, buf.size());
  for (int64_t i = 0; i < n; ++i) {
    const float* r = rows + i * dim;
    if (dim == 1) {
      std::fprintf(f, "%u\t%g\n", keys[i], r[0]);
    } else {
      std::fprintf(f, "%u\tVec:\t", keys[i]);
      for (int j = 0; j < dim; ++j) std::fprintf(f, "%g ", r[j]);
      std::fputc('\n', f);
    }
  }
  std::fclose(f);
  return n;
}
// Param-style line: "<key>\tVec:\t v0 v1 … \tAcc:\t a0 a1 … \n" — the row followed by its optimizer state (AdaGrad
// accumulator), which is what the reference's dump prints when the table's Param carries one (the whole Param goes
// through operator<<, parameter/sparsetable.h:49-56). state == nullptr writes plain rows. digits: significant digits per
// number (6 = the reference's default ostream precision; 9 round-trips a float exactly).
// state_dim: floats of optimizer state per entry (state words x dim; AdaGrad = dim, FTRL = 2 x dim)
int64_t ssb_dump_write_text3(const char* path, const uint32_t* keys, const float* rows, const float* state, int64_t n, int dim,
                             int state_dim, int append, int digits) {
  if (state_dim <= 0) state_dim = dim;
  FILE* f = std::fopen(path, append ? "a" : "w");
  if (!f) return -1;
  std::vector<char> buf(1 << 20);
  std::setvbuf(f, buf.data(), _IOFBF, buf.size());
  if (digits < 1) digits = 6;
  if (digits > 17) digits = 17;
  for (int64_t i = 0; i < n; ++i) {
    const float* r = rows + i * dim;
    if (dim == 1 && !state) {
      std::fprintf(f, "%u\t%.*g\n", keys[i], digits, r[0]);
      continue;
    }
    std::fprintf(f, "%u\tVec:\t", keys[i]);
    for (int j = 0; j < dim; ++j) std::fprintf(f, "%.*g ", digits, r[j]);
    if (state) {
      std::fputs("\tAcc:\t", f);
      const float* a = state + i * state_dim;
      for (int j = 0; j < state_dim; ++j) std::fprintf(f, "%.*g ", digits, a[j]);
    }
    std::fputc('\n', f);
  }
  std::fclose(f);
  return n;
}
int64_t ssb_dump_write_text2(const char* path, const uint32_t* keys, const float* rows, const float* state, int64_t n, int dim,
                             int append, int digits) {
  return ssb_dump_write_text3(path, keys, rows, state, n, dim, dim, append, digits);
}
// Reads either line form. state may be nullptr (an "Acc:" part is then skipped); rows / state shorter than dim are
// zero-padded. *has_state (optional) is set to 1 if any line carried an "Acc:" part. Returns the rows read (<= max_n).
int64_t ssb_dump_read_text3(const char* path, uint32_t* keys, float* rows, float* state, int64_t max_n, int dim, int state_dim,
                            int* has_state) {
  if (state_dim <= 0) state_dim = dim;
  FILE* f = std::fopen(path, "r");
  if (!f) return -1;
  LineFileReader reader;
  int64_t n = 0;
  if (has_state) *has_state = 0;
  auto parse_vec = [](char* p, float* out, int dim) -> char* {
    for (int j = 0; j < dim; ++j) {
      char* end = nullptr;
      const float v = std::strtof(p, &end);
      if (end == p) { if (out) for (; j < dim; ++j) out[j] = 0.f; break; }
      if (out) out[j] = v;
      p = end;
    }
    return p;
  };
  while (n < max_n) {
    char* line = reader.getline(f);
    if (!line) break;
    char* tab = std::strchr(line, '\t');
    if (!tab) continue;
    keys[n] = (uint32_t)std::strtoul(line, nullptr, 10);
    char* p = tab + 1;
    if (std::strncmp(p, "Vec:", 4) == 0) p += 4;
    p = parse_vec(p, rows + n * dim, dim);
    float* st = state ? state + n * state_dim : nullptr;
    char* acc = std::strstr(p, "Acc:");
    if (acc) {
      if (has_state) *has_state = 1;
      parse_vec(acc + 4, st, state_dim);
    } else if (st) {
      for (int j = 0; j < state_dim; ++j) st[j] = 0.f;
    }
    ++n;
  }
  std::fclose(f);
  return n;
}
int64_t ssb_dump_read_text2(const char* path, uint32_t* keys, float* rows, float* state, int64_t max_n, int dim, int* has_state) {
  return ssb_dump_read_text3(path, keys, rows, state, max_n, dim, dim, has_state);
}
// returns the number of rows read (<= max_n); rows shorter than dim are zero-padded
int64_t ssb_dump_read_text(const char* path, uint32_t* keys, float* rows, int64_t max_n, int dim) {
  FILE* f = std::fopen(path, "r");
  if (!f) return -1;
  LineFileReader reader;
  int64_t n = 0;
  while (n < max_n) {
    char* line = reader.getline(f);
    if (!line) break;
    char* tab = std::strchr(line, '\t');
    if (!tab) continue;
    keys[n] = (uint32_t)std::strtoul(line, nullptr, 10);
    char* p = tab + 1;
    if (std::strncmp(p, "Vec:", 4) == 0) p += 4;
    float* r = rows + n * dim;
    for (int j = 0; j < dim; ++j) {
      char* end = nullptr;
      r[j] = std::strtof(p, &end);
      if (end == p) { for (; j < dim; ++j) r[j] = 0.f; break; }
      p = end;
    }
    ++n;
  }
  std::fclose(f);
  return n;
}

// ---- CPU sparse table (oracle for the HBM table; SGD / AdaGrad access methods) -------------------
struct CpuTable {
  table_t_ table;
  RowPullMethod<table_t_> pull;
  SgdPushMethod<table_t_> sgd;
  AdaGradPushMethod<table_t_> adagrad;
  int dim;
  bool use_adagrad;
  CpuTable(int shards, int dim_, bool ada, float lr, float eps, uint64_t seed, float scale)
      : table(shards), pull(dim_, ada, seed, scale), sgd(lr), adagrad(lr, eps), dim(dim_), use_adagrad(ada) {}
};
void* ssb_cpu_table_new(int shard_num, int dim, int adagrad, float lr, float eps, uint64_t seed, float init_scale) {
  return new CpuTable(shard_num, dim, adagrad != 0, lr, eps, seed, init_scale);
}
void ssb_cpu_table_free(void* t) { delete static_cast<CpuTable*>(t); }
void ssb_cpu_table_pull(void* tp, const uint32_t* keys, int64_t n, float* rows_out) {
  auto* t = static_cast<CpuTable*>(tp);
  PullAccessAgent<table_t_, RowPullMethod<table_t_>> agent(t->table, t->pull);
  VecF val;
  for (int64_t i = 0; i < n; ++i) {
    agent.get_pull_value(keys[i], val);
    std::memcpy(rows_out + i * t->dim, val.data(), sizeof(float) * t->dim);
  }
}
int64_t ssb_cpu_table_push(void* tp, const uint32_t* keys, const float* grads, int64_t n) {
  auto* t = static_cast<CpuTable*>(tp);
  int64_t missing = 0;
  VecF g((size_t)t->dim);
  for (int64_t i = 0; i < n; ++i) {
    std::memcpy(g.data(), grads + i * t->dim, sizeof(float) * t->dim);
    bool ok;
    if (t->use_adagrad) {
      PushAccessAgent<table_t_, AdaGradPushMethod<table_t_>> agent(t->table, t->adagrad);
      ok = agent.apply_push_value(keys[i], g);
    } else {
      PushAccessAgent<table_t_, SgdPushMethod<table_t_>> agent(t->table, t->sgd);
      ok = agent.apply_push_value(keys[i], g);
    }
    if (!ok) ++missing;
  }
  return missing;
}
int64_t ssb_cpu_table_size(void* tp) { return (int64_t) static_cast<CpuTable*>(tp)->table.size(); }
int ssb_cpu_table_dump(void* tp, const char* path) {
  std::ofstream out(path);
  if (!out.good()) return -1;
  static_cast<CpuTable*>(tp)->table.output(out);
  return 0;
}

// ---- corpus loader (host-side data pipeline of the GPU workers) ----------------------------------
void* ssb_corpus_open(const char* path, int batch_tokens, int sent_len, uint32_t pad_id, int epochs) {
  FILE* f = std::fopen(path, "r");
  if (!f) return nullptr;
  std::fclose(f);
  if (batch_tokens <= 0 || sent_len <= 0 || batch_tokens % sent_len != 0) return nullptr;   // refuse, do not abort
  return new CorpusLoader(path, batch_tokens, sent_len, pad_id, epochs);
}
void* ssb_corpus_open_split(const char* path, int batch_tokens, int sent_len, uint32_t pad_id, int epochs, int split_n,
                            int split_i) {
  FILE* f = std::fopen(path, "r");
  if (!f) return nullptr;
  std::fclose(f);
  if (split_n < 1 || split_i < 0 || split_i >= split_n) return nullptr;
  if (batch_tokens <= 0 || sent_len <= 0 || batch_tokens % sent_len != 0) return nullptr;   // refuse, do not abort
  return new CorpusLoader(path, batch_tokens, sent_len, pad_id, epochs, 4, split_n, split_i);
}
// Full form: parser threads (<= 0: chosen from the host) and an id bound (ids >= max_id become padding; 0 = off).
void* ssb_corpus_open_ex(const char* path, int batch_tokens, int sent_len, uint32_t pad_id, int epochs, int split_n,
                         int split_i, int threads, uint32_t max_id) {
  FILE* f = std::fopen(path, "r");
  if (!f) return nullptr;
  std::fclose(f);
  if (split_n < 1 || split_i < 0 || split_i >= split_n) return nullptr;
  if (batch_tokens <= 0 || sent_len <= 0 || batch_tokens % sent_len != 0) return nullptr;   // refuse, do not abort
  return new CorpusLoader(path, batch_tokens, sent_len, pad_id, epochs, 4, split_n, split_i, threads, max_id);
}
// out[0] = sentence lines of this reader's share per pass, out[1] = minibatches it will deliver,
// out[2] = token ids >= max_id replaced by padding so far
void ssb_corpus_stats(void* h, long long* out) {
  CorpusLoader* c = static_cast<CorpusLoader*>(h);
  out[0] = c->sentences(); out[1] = c->batches(); out[2] = c->bad_tokens();
}
// counts[vocab] += word frequencies of the corpus; returns the number of ids >= vocab, or -1 if the file cannot be opened
long long ssb_corpus_count_words(const char* path, uint32_t vocab, unsigned long long* counts, int threads) {
  return corpus_count_words(path, vocab, counts, threads);
}
long ssb_corpus_next(void* h, uint32_t* out) { return static_cast<CorpusLoader*>(h)->next(out); }
void ssb_corpus_close(void* h) { delete static_cast<CorpusLoader*>(h); }

// ---- sparse-LR example files: `label key:value key:value …` (value defaults to 1), the format the CPU worker's lr
// app reads (csrc/apps/algorithms.h, LogisticRegressionAlg::parse_record). Parses the lines whose index % split_n ==
// split_i into padded arrays for the GPU step: keys[n][max_feat] (0xFFFFFFFF = empty position), vals[n][max_feat],
// labels[n] (1 if label > 0). keys == NULL: only count. key_mod > 0: keys are taken modulo key_mod (hashed space).
// out_stats[0] = examples, [1] = features dropped because a line had more than max_feat. Returns examples or -1.
long long ssb_lr_parse_file(const char* path, int max_feat, uint32_t* keys, float* vals, float* labels, long long max_examples,
                            unsigned long long key_mod, int split_n, int split_i, long long* out_stats) {
  if (!path || max_feat < 1 || split_n < 1 || split_i < 0 || split_i >= split_n) return -1;
  MappedFile mf(path);
  if (!mf.opened()) return -1;
  const char* p = mf.data();
  if (!p) { if (out_stats) { out_stats[0] = 0; out_stats[1] = 0; } return 0; }     // empty file
  const char* end = p + mf.size();
  long long n = 0, dropped = 0, line_no = 0;
  while (p < end) {
    const char* eol = static_cast<const char*>(std::memchr(p, '\n', (size_t)(end - p)));
    if (!eol) eol = end;
    const char* q = p;
    while (q < eol && (*q == ' ' || *q == '\t' || *q == '\r')) ++q;
    const bool is_example = q < eol && (std::isdigit((unsigned char)*q) || *q == '-' || *q == '+' || *q == '.');
    if (is_example) {
      const bool mine = (line_no % split_n) == split_i;
      ++line_no;
      if (mine) {
        if (keys && n >= max_examples) break;
        char* e = nullptr;
        const float lab = std::strtof(q, &e);
        q = e;
        int f = 0;
        uint32_t* kr = keys ? keys + n * max_feat : nullptr;
        float* vr = vals ? vals + n * max_feat : nullptr;
        while (q < eol) {
          while (q < eol && (*q == ' ' || *q == '\t' || *q == '\r')) ++q;
          if (q >= eol || !std::isdigit((unsigned char)*q)) break;
          unsigned long long k = std::strtoull(q, &e, 10);
          q = e;
          float v = 1.f;
          if (q < eol && *q == ':') { v = std::strtof(q + 1, &e); q = e; }
          if (key_mod) k %= key_mod;
          if (f < max_feat) { if (kr) { kr[f] = (uint32_t)k; vr[f] = v; } ++f; } else ++dropped;
        }
        if (kr) {
          for (; f < max_feat; ++f) { kr[f] = 0xFFFFFFFFu; vr[f] = 0.f; }
          labels[n] = lab > 0 ? 1.f : 0.f;
        }
        ++n;
      }
    }
    p = eol + 1;
  }
  if (out_stats) { out_stats[0] = n; out_stats[1] = dropped; }
  return n;
}

// ---- roles (one role per process; they use the process-wide transfer) ----------------------------
int ssb_run_master(const char* config_path) {
  if (ssb_config_load(config_path) != 0) return -1;
  run_master();
  ps_transfer().shutdown();
  return 0;
}
int ssb_run_server(const char* config_path) {
  if (ssb_config_load(config_path) != 0) return -1;
  run_server();
  ps_transfer().shutdown();
  return 0;
}
int ssb_run_worker(const char* config_path, const char* app, const char* data_path) {
  if (ssb_config_load(config_path) != 0) return -1;
  int rc = run_worker(app, data_path ? data_path : "");
  if (global_config().get_config("local_train", "0").to_int32() == 0) ps_transfer().shutdown();
  return rc;
}


// ---- a GPU rank's control plane: register with the master, route, fragment map, group exchange, shutdown ----------
// ref: system/node_init.h:16-152, master/terminate.h:15-109, server/terminate.h:16-54; see system/group_exchange.h.
static GpuNode* g_node = nullptr;

// master_addr "tcp://ip:port"; listen_addr "" = an ephemeral port on the first non-loopback interface, or e.g.
// "tcp://127.0.0.1:0". Returns the rank (>= 0) or a negative number.
int ssb_node_start(const char* master_addr, const char* listen_addr, int init_timeout_s) {
  if (g_node) return -2;
  if (!master_addr || !*master_addr) return -1;
  global_config().set("master_addr", master_addr);
  if (listen_addr && *listen_addr) global_config().set("listen_addr", listen_addr);
  if (init_timeout_s > 0) global_config().set("init_timeout", std::to_string(init_timeout_s));
  g_node = new GpuNode();
  return g_node->start();
}
int ssb_node_rank(void) { return g_node ? g_node->rank() : -1; }
int ssb_node_world(void) { return g_node ? g_node->world() : -1; }
// The fragment map the master built (NODE_ASKFOR_HASHFRAG). out[0] = nodes, out[1] = fragments; map_out may be
// NULL (size query). Returns the number of fragments, -1 without a node, -2 if cap is too small.
int ssb_node_hashfrag(int* out, uint32_t* map_out, int cap) {
  if (!g_node) return -1;
  auto& hf = shared_hashfrag();
  if (out) { out[0] = hf.num_nodes(); out[1] = hf.num_frags(); }
  if (!map_out) return hf.num_frags();
  if (cap < hf.num_frags()) return -2;
  for (int i = 0; i < hf.num_frags(); ++i) map_out[i] = (uint32_t)hf.map_table()[(size_t)i];
  return hf.num_frags();
}
// All-gather through the master: out receives the ranks' byte strings back to back, lens[r] their lengths.
// Returns the total bytes, -1 without a node, -2 timed out, -3 out too small.
long ssb_node_exchange(const void* buf, long len, void* out, long out_cap, long* lens, int timeout_s) {
  if (!g_node) return -1;
  std::vector<std::string> parts;
  if (!g_node->exchange(std::string(static_cast<const char*>(buf), (size_t)len), parts, 1000L * timeout_s)) return -2;
  long total = 0;
  for (auto& p : parts) total += (long)p.size();
  if (total > out_cap) return -3;
  char* o = static_cast<char*>(out);
  for (size_t r = 0; r < parts.size(); ++r) {
    std::memcpy(o, parts[r].data(), parts[r].size());
    o += parts[r].size();
    if (lens) lens[r] = (long)parts[r].size();
  }
  return total;
}
// WORKER_FINISH_WORK, then wait for SERVER_TOLD_TO_TERMINATE. 0 ok, -2 timed out. Afterwards: dump, ssb_node_dump_done().
int ssb_node_finish(int timeout_s) {
  if (!g_node) return -1;
  return g_node->finish(1000L * timeout_s) ? 0 : -2;
}
void ssb_node_dump_done(void) { if (g_node) g_node->dump_done(); }
void ssb_node_stop(void) {
  if (!g_node) return;
  ps_transfer().shutdown();
  delete g_node;
  g_node = nullptr;
}

}  // extern "C"
