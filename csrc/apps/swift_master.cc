// swift_master — `swift_master -config master.conf`
// ref: /root/reference/src/tools/run_master.sh:1 (CLI contract).
#include "algorithms.h"
int main(int argc, char** argv) {
  ssb::CMDLine cmd(argc, argv);
  std::string conf = cmd.registerParameter("config", "path of the config file");
  if (!cmd.hasParameter(conf)) { cmd.print_help(); return 1; }
  ssb::global_config().load_conf(cmd.getValue(conf));
  ssb::global_config().parse();
  ssb::run_master();
  ssb::ps_transfer().shutdown();
  return 0;
}
