// swift_server — `swift_server -config server.conf`
// ref: /root/reference/src/tools/run_server.sh:7 (CLI contract). The final parameter dump goes
// to stdout (or `param_dump_path`), as the reference's Hadoop-streaming deployment expects.
#include "algorithms.h"
int main(int argc, char** argv) {
  ssb::CMDLine cmd(argc, argv);
  std::string conf = cmd.registerParameter("config", "path of the config file");
  if (!cmd.hasParameter(conf)) { cmd.print_help(); return 1; }
  ssb::global_config().load_conf(cmd.getValue(conf));
  ssb::global_config().parse();
  ssb::run_server();
  ssb::ps_transfer().shutdown();
  return 0;
}
