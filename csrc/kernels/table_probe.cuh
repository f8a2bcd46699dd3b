// table_probe.cuh — the slot claim protocol of the HBM table, shared by the table kernels (table.cu) and the fused
// sparse-LR step (lr.cu). See the header comment of table.cu for the protocol.
#pragma once
#include "common.cuh"

namespace ssb {

struct Probe {
  int owner;
  uint32_t slot;   // slot within the owner shard, or SSB_INVALID_SLOT
  bool is_new;     // we hold the slot LOCKED and must publish it
};

// Warp-cooperative: all 32 lanes call with the same key.
__device__ __forceinline__ Probe probe_insert(const SsbTableView& v, uint32_t key, bool insert, int lane) {
  Probe p;
  p.owner = owner_of(key, v.frag_map, v.frag_num);
  p.slot = SSB_INVALID_SLOT;
  p.is_new = false;
  if (p.owner < 0 || p.owner >= v.num_shards) { p.owner = 0; return p; }
  uint32_t* kbase = v.keys[p.owner];
  const uint32_t nb = v.capacity / SSB_BUCKET;
  uint32_t b = reduce_range(slot_hash(key), nb);
  uint32_t lock_spins = 0;
  for (uint32_t probes = 0; probes < nb;) {
    uint32_t kv = ld_acquire_sys(kbase + (size_t)b * SSB_BUCKET + lane);
    unsigned m = __ballot_sync(0xffffffffu, kv == key);
    if (m) { p.slot = b * SSB_BUCKET + (__ffs(m) - 1); return p; }
    unsigned locked = __ballot_sync(0xffffffffu, kv == SSB_LOCKED_KEY);
    if (locked) {
      // the holder is a resident warp that is writing the row's initial value and never waits, so this resolves in
      // microseconds; bounded all the same (~0.5 s): a slot that stays LOCKED is reported like a full table
      if (++lock_spins > (1u << 22)) return p;
      __nanosleep(100);
      continue;
    }
    unsigned empties = __ballot_sync(0xffffffffu, kv == SSB_EMPTY_KEY);
    if (empties) {
      if (!insert) return p;
      int e = __ffs(empties) - 1;
      uint32_t old = 0;
      if (lane == e) old = cas_sys(kbase + (size_t)b * SSB_BUCKET + e, SSB_EMPTY_KEY, SSB_LOCKED_KEY);
      old = __shfl_sync(0xffffffffu, old, e);
      if (old == SSB_EMPTY_KEY) { p.slot = b * SSB_BUCKET + e; p.is_new = true; return p; }
      continue;  // lost the race: re-read the bucket
    }
    b = (b + 1 == nb) ? 0 : b + 1;
    ++probes;
  }
  return p;  // table full
}

__device__ __forceinline__ void publish(const SsbTableView& v, const Probe& p, uint32_t key, int lane) {
  __syncwarp();
  if (lane == 0) st_release_sys(v.keys[p.owner] + p.slot, key);
}

}  // namespace ssb
