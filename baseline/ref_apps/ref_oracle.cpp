// ref_oracle.cpp — the reference's OWN hash-fragment routing and sparse table as a test oracle (SURVEY.md §4, §8
// "Oracle" column). Compiled against the unmodified headers in baseline/_ref/src with the shim headers; none of
// swiftsnails_b200's code is on this path.
//
//   ref_oracle route <num_nodes> <frag_num> <n_keys> <seed>         -> "<key> <node_id>" per line, then the map
//   ref_oracle sgd   <shard_num> <dim> <lr> <n_keys> <steps> <seed> -> the table's own text dump after SGD pushes
//
// Keys and gradients come from a tiny LCG that the Python test repeats, so both sides see the same numbers.
#include <cstdio>
#include <cstdlib>
#include <unistd.h>
#include <iostream>
#include <string>
#include <vector>
#include "swiftsnails.h"

using namespace swift_snails;
typedef index_t okey_t;

static uint32_t lcg(uint32_t& s) { s = s * 1664525u + 1013904223u; return s; }

// the reference's config is file based (ConfigParser::load_conf + parse, utils/ConfigParser.h:52-57)
static void load_config(const std::string& text) {
  const std::string path = "/tmp/ref_oracle_" + std::to_string((long)getpid()) + ".conf";
  FILE* f = fopen(path.c_str(), "w");
  fputs(text.c_str(), f);
  fclose(f);
  global_config().load_conf(path);
  global_config().parse();
  remove(path.c_str());
}

static int g_dim = 4;
struct ORow {
  std::vector<float> d;
};
static std::ostream& operator<<(std::ostream& os, const ORow& r) {
  os << "Vec:\t";
  for (float x : r.d) os << x << " ";
  return os;
}
typedef SparseTable<okey_t, ORow> otable_t;
static float g_lr = 0.1f;
class OPull : public PullAccessMethod<otable_t, ORow, ORow> {
 public:
  // deterministic init so that the other side can restate it: element j of key k = ((k * 31 + j) % 17 - 8) / 16
  void init_param(const okey_t& key, ORow& p) override {
    p.d.resize(g_dim);
    for (int j = 0; j < g_dim; ++j) p.d[j] = (float)((int)((key * 31u + (unsigned)j) % 17u) - 8) / 16.0f;
  }
  void get_pull_value(const okey_t&, const ORow& p, ORow& v) override { v = p; }
};
class OPush : public PushAccessMethod<otable_t, ORow, ORow> {
 public:
  void merge_push_value(const okey_t&, ORow&, const ORow&) override {}
  void apply_push_value(const okey_t&, ORow& p, const ORow& g) override {
    for (int j = 0; j < g_dim; ++j) p.d[j] -= g_lr * g.d[j];
  }
};

int main(int argc, char** argv) {
  if (argc < 2) return 2;
  const std::string mode = argv[1];
  if (mode == "route" && argc >= 6) {
    const int num_nodes = atoi(argv[2]), frag_num = atoi(argv[3]), n = atoi(argv[4]);
    uint32_t seed = (uint32_t)strtoul(argv[5], nullptr, 10);
    load_config("frag_num: " + std::string(argv[3]) + "\n");   // BasicHashFrag::init reads it (hashfrag.h:33)
    BasicHashFrag<okey_t> hf;
    hf.set_num_nodes(num_nodes);
    hf.init();
    for (int i = 0; i < n; ++i) {
      const okey_t k = lcg(seed);
      printf("%u %d\n", k, hf.to_node_id(k));
    }
    // the wire form of the map (hashfrag.h:55-62): num_nodes, num_frags, map[num_frags]
    BinaryBuffer bb;
    hf.serialize(bb);
    int nn = 0, nf = 0;
    bb >> nn;
    bb >> nf;
    printf("MAP %d %d", nn, nf);
    for (int i = 0; i < nf; ++i) { index_t v; bb >> v; printf(" %u", v); }
    printf("\n");
    (void)frag_num;
    return 0;
  }
  if (mode == "sgd" && argc >= 8) {
    const int shard_num = atoi(argv[2]);
    g_dim = atoi(argv[3]);
    g_lr = (float)atof(argv[4]);
    const int n = atoi(argv[5]), steps = atoi(argv[6]);
    uint32_t seed = (uint32_t)strtoul(argv[7], nullptr, 10);
    load_config("shard_num: " + std::string(argv[2]) + "\n");   // SparseTable's ctor reads it (sparsetable.h:77)
    otable_t table;
    auto pull = make_pull_access<otable_t, OPull>(table);
    auto push = make_push_access<otable_t, OPush>(table);
    std::vector<okey_t> keys(n);
    for (auto& k : keys) k = lcg(seed) % 100000u;
    for (int s = 0; s < steps; ++s) {
      for (int i = 0; i < n; ++i) {
        ORow v;
        pull->get_pull_value(keys[i], v);          // lazy init on first touch (sparsetable.h:142-149)
        ORow g;
        g.d.resize(g_dim);
        for (int j = 0; j < g_dim; ++j) g.d[j] = (float)((int)(lcg(seed) % 2001u) - 1000) / 1000.0f;
        push->apply_push_value(keys[i], g);        // sparsetable.h:181-192
      }
    }
    table.output();                                // "key \t value" per entry, every shard (sparsetable.h:100-104)
    (void)shard_num;
    return 0;
  }
  return 2;
}
