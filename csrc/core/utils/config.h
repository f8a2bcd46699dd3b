// config.h — `key: value` config files with `import`, comments and typed getters.
// ref: /root/reference/src/utils/ConfigParser.h:25-129.
// Same grammar: `# comment`, `import <path>` (recursive, relative to the importing file's
// directory or the cwd), first ':' splits (so tcp://… values work, :106), duplicate keys keep
// the FIRST value (map insert, :118), missing key aborts (:73). register_config (commented
// out in the reference but used by its tests, :60-66) is provided.
#pragma once
#include <fstream>
#include <map>
#include <string>
#include "common.h"
#include "string_util.h"
namespace ssb {

class ConfigParser {
 public:
  struct Item {
    std::string value;
    int32_t to_int32() const { return (int32_t)std::stol(value); }
    int64_t to_int64() const { return (int64_t)std::stoll(value); }
    float to_float() const { return std::stof(value); }
    double to_double() const { return std::stod(value); }
    const std::string& to_string() const { return value; }
    bool to_bool() const { return value == "true" || value == "True" || value == "1"; }
    bool empty() const { return value.empty(); }
  };

  ConfigParser() = default;
  explicit ConfigParser(const std::string& path) { load_conf(path); }

  void load_conf(const std::string& path) { paths_.push_back(path); }
  void parse() {
    for (auto& p : paths_) parse_file(p, 0);
    paths_.clear();
  }
  void clear() { items_.clear(); paths_.clear(); }
  void register_config(const std::string& key, const std::string& value) { items_.insert({key, Item{value}}); }
  void set(const std::string& key, const std::string& value) { items_[key] = Item{value}; }
  bool has(const std::string& key) const { return items_.count(key) > 0; }
  const Item& get_config(const std::string& key) const {
    auto it = items_.find(key);
    CHECK(it != items_.end()) << "config key '" << key << "' is missing";
    return it->second;
  }
  Item get_config(const std::string& key, const std::string& dflt) const {
    auto it = items_.find(key);
    return it == items_.end() ? Item{dflt} : it->second;
  }
  const std::map<std::string, Item>& items() const { return items_; }
  bool parse_line(const std::string& raw, const std::string& dir = "", int depth = 0) {
    std::string line = raw;
    size_t hash = line.find('#');
    if (hash != std::string::npos) line = line.substr(0, hash);
    line = trim(line);
    if (line.empty()) return false;
    if (headswith(line, "import ") || headswith(line, "import\t")) {
      std::string inc = trim(line.substr(6));
      CHECK(depth < 16) << "config import nesting too deep at " << inc;
      std::string full = inc;
      if (!inc.empty() && inc[0] != '/' && !dir.empty()) {
        std::ifstream probe(dir + "/" + inc);
        if (probe.good()) full = dir + "/" + inc;
      }
      parse_file(full, depth + 1);
      return true;
    }
    auto kv = key_value_split(line, ":");
    CHECK(!kv.first.empty()) << "bad config line: " << raw;
    items_.insert({kv.first, Item{kv.second}});
    return true;
  }

 private:
  void parse_file(const std::string& path, int depth) {
    std::ifstream in(path);
    CHECK(in.good()) << "cannot open config file " << path;
    std::string dir;
    size_t slash = path.find_last_of('/');
    if (slash != std::string::npos) dir = path.substr(0, slash);
    std::string line;
    while (std::getline(in, line)) parse_line(line, dir, depth);
  }
  std::map<std::string, Item> items_;
  std::vector<std::string> paths_;
};

inline ConfigParser& global_config() {
  static ConfigParser cfg;
  return cfg;
}

// `wire_compat: reference`: pull requests carry (key, empty value) pairs like the reference's worker
// (global_pull_access.h:84-88) instead of this framework's bare key array. Read once.
inline bool reference_wire() {
  static const bool on = global_config().get_config("wire_compat", "").to_string() == "reference";
  return on;
}

}  // namespace ssb
