// cmdline.h — `-name value` / `--name value` argv parser with typed getters and a help
// registry. ref: /root/reference/src/utils/CMDLine.h:29-197 (namespace fms, from libFM).
// Written fresh; the apps use `-config <conf>` and `-data <file>` (tools/run_*.sh).
#pragma once
#include <map>
#include <string>
#include <vector>
#include "common.h"
#include "string_util.h"
namespace ssb {

class CMDLine {
 public:
  CMDLine(int argc, char** argv) {
    program_ = argc > 0 ? argv[0] : "";
    for (int i = 1; i < argc; ++i) {
      std::string a = argv[i];
      if (a.size() > 1 && a[0] == '-' && !is_number(a)) {
        std::string name = a.substr(a[1] == '-' ? 2 : 1);
        std::string value;
        if (i + 1 < argc) {
          std::string nxt = argv[i + 1];
          if (!(nxt.size() > 1 && nxt[0] == '-' && !is_number(nxt))) { value = nxt; ++i; }
        }
        values_[name] = value;
      } else {
        positional_.push_back(a);
      }
    }
  }
  // registers a parameter for the help text and returns its name (libFM idiom)
  std::string registerParameter(const std::string& name, const std::string& help) {
    help_[name] = help;
    return name;
  }
  bool hasParameter(const std::string& name) const { return values_.count(name) > 0; }
  std::string getValue(const std::string& name) const {
    auto it = values_.find(name);
    CHECK(it != values_.end()) << "missing command line parameter -" << name;
    return it->second;
  }
  std::string getValue(const std::string& name, const std::string& dflt) const {
    auto it = values_.find(name);
    return it == values_.end() ? dflt : it->second;
  }
  int getValue(const std::string& name, int dflt) const { return hasParameter(name) ? std::stoi(getValue(name)) : dflt; }
  double getValue(const std::string& name, double dflt) const { return hasParameter(name) ? std::stod(getValue(name)) : dflt; }
  // `;` or `,` separated lists
  std::vector<std::string> getStrValues(const std::string& name) const { return split(getValue(name), ";,"); }
  std::vector<int> getIntValues(const std::string& name) const {
    std::vector<int> out;
    for (auto& s : getStrValues(name)) out.push_back(std::stoi(s));
    return out;
  }
  std::vector<double> getDblValues(const std::string& name) const {
    std::vector<double> out;
    for (auto& s : getStrValues(name)) out.push_back(std::stod(s));
    return out;
  }
  const std::vector<std::string>& positional() const { return positional_; }
  void print_help(std::ostream& os = std::cout) const {
    os << "usage: " << program_ << " [options]\n";
    for (auto& kv : help_) os << "  -" << kv.first << "\t" << kv.second << "\n";
  }
  // true when every given parameter was registered
  bool checkParameters() const {
    bool ok = true;
    for (auto& kv : values_)
      if (!help_.count(kv.first)) { LOG(ERROR) << "unknown parameter -" << kv.first; ok = false; }
    return ok;
  }

 private:
  static bool is_number(const std::string& s) {
    char* end = nullptr;
    std::strtod(s.c_str(), &end);
    return end && *end == '\0';
  }
  std::string program_;
  std::map<std::string, std::string> values_, help_;
  std::vector<std::string> positional_;
};

}  // namespace ssb
