// hash.h — MurmurHash3 fmix64. ref: /root/reference/src/utils/HashFunction.h:16-24.
// Used for key -> fragment (hashfrag) and key -> intra-server shard, bit-exact with the
// reference and with the device version in csrc/kernels/common.cuh.
#pragma once
#include <cstdint>
namespace ssb {
inline uint64_t get_hash_code(uint64_t x) {
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ULL;
  x ^= x >> 33;
  return x;
}
}  // namespace ssb
