// shell.h — run shell commands / pipes. ref: /root/reference/src/utils/shell.h:7-62
// (GlobalShell via popen("set -o pipefail; …")).
#pragma once
#include <cstdio>
#include <memory>
#include <string>
#include "common.h"
namespace ssb {

class GlobalShell {
 public:
  // open a pipe to/from a shell command; closes with guarded_pclose
  static std::shared_ptr<FILE> make_pipe(const std::string& cmd, const char* mode = "r") {
    std::string full = "set -o pipefail; " + cmd;
    std::string wrapped = "bash -c '" + escape(full) + "'";
    FILE* f = guarded_popen(wrapped.c_str(), mode);
    PCHECK(f != nullptr) << "popen failed: " << cmd;
    return std::shared_ptr<FILE>(f, [](FILE* p) { if (p) guarded_pclose(p); });
  }
  // run and return the exit status
  static int execute(const std::string& cmd) {
    std::string full = "bash -c '" + escape("set -o pipefail; " + cmd) + "'";
    FILE* f = guarded_popen(full.c_str(), "r");
    PCHECK(f != nullptr) << "popen failed: " << cmd;
    char buf[4096];
    while (std::fread(buf, 1, sizeof(buf), f) > 0) {}
    return guarded_pclose(f);
  }
  static std::string get_command_output(const std::string& cmd) {
    auto p = make_pipe(cmd, "r");
    std::string out;
    char buf[4096];
    size_t n;
    while ((n = std::fread(buf, 1, sizeof(buf), p.get())) > 0) out.append(buf, n);
    return out;
  }

 private:
  static std::string escape(const std::string& s) {
    std::string o;
    for (char c : s) {
      if (c == '\'') o += "'\\''";
      else o += c;
    }
    return o;
  }
};

}  // namespace ssb
