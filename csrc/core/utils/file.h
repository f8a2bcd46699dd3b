// file.h — parallel line scanner: many threads share one FILE*, each locks, reads one line,
// unlocks, and processes it. ref: /root/reference/src/utils/file.h:14-33. This is the apps'
// data reader (SwiftWorker loop).
#pragma once
#include <cstdio>
#include <functional>
#include <mutex>
#include <string>
#include "string_util.h"
namespace ssb {

inline void scan_file_by_line(FILE* file, std::mutex& file_mut, const std::function<void(const std::string&)>& handler) {
  LineFileReader reader;
  std::string line;
  while (true) {
    {
      std::lock_guard<std::mutex> lk(file_mut);
      char* res = reader.getline(file);
      if (res == nullptr) return;
      line.assign(res, reader.length());
    }
    handler(line);
  }
}

// Batched variant (not in the reference): grabs up to `batch` lines per lock acquisition.
inline void scan_file_by_batch(FILE* file, std::mutex& file_mut, size_t batch,
                               const std::function<void(std::vector<std::string>&)>& handler) {
  LineFileReader reader;
  std::vector<std::string> lines;
  while (true) {
    lines.clear();
    {
      std::lock_guard<std::mutex> lk(file_mut);
      while (lines.size() < batch) {
        char* res = reader.getline(file);
        if (res == nullptr) break;
        lines.emplace_back(res, reader.length());
      }
    }
    if (lines.empty()) return;
    handler(lines);
  }
}

}  // namespace ssb
