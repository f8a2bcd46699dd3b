// swift_worker — `swift_worker -config worker.conf -data data.txt [-app word2vec|lr|bench]`
// ref: /root/reference/src/tools/run_worker.sh:13 (CLI contract).
#include "algorithms.h"
int main(int argc, char** argv) {
  ssb::CMDLine cmd(argc, argv);
  std::string conf = cmd.registerParameter("config", "path of the config file");
  std::string data = cmd.registerParameter("data", "training data file");
  std::string app = cmd.registerParameter("app", "word2vec | lr | bench (default: config key `app`, else word2vec)");
  if (!cmd.hasParameter(conf)) { cmd.print_help(); return 1; }
  ssb::global_config().load_conf(cmd.getValue(conf));
  ssb::global_config().parse();
  std::string a = cmd.getValue(app, ssb::global_config().get_config("app", "word2vec").to_string());
  int rc = ssb::run_worker(a, cmd.getValue(data, std::string("")));
  if (ssb::global_config().get_config("local_train", "0").to_int32() == 0) ssb::ps_transfer().shutdown();
  return rc;
}
