#pragma once
#include <cstdio>
#define RAW_LOG(sev, ...) do { if (::ssb_glog_shim::sev >= ::ssb_glog_shim::min_level()) { std::fprintf(stderr, __VA_ARGS__); std::fputc('\n', stderr); } } while (0)
#define RAW_DLOG(sev, ...) do { } while (0)
