// dedup.cu — K2: key de-duplication with an inverse index.
// ref: parameter/global_param_cache.h:107-108 (params/grads are hash maps => one entry per
//      unique key), global_push_access.h:80-99 (one (key,grad) per unique key is pushed).
// GPU form: an open-addressing hash *set* in scratch memory. Entries are 64-bit
// (epoch << 32 | key); an entry whose epoch differs from the current call's epoch counts as
// empty, so the scratch never needs clearing between calls.
#include "common.cuh"

namespace ssb {

__global__ void __launch_bounds__(256)
dedup_insert_kernel(const uint32_t* __restrict__ keys, int64_t n, unsigned long long* set,
                    uint32_t* __restrict__ set_ids, uint32_t mask, uint32_t epoch,
                    uint32_t* __restrict__ unique, uint32_t* __restrict__ inverse,
                    uint32_t* __restrict__ n_unique) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t n_round = (n + 31) / 32 * 32;
  const int lane = threadIdx.x & 31;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n_round; i += stride) {
    uint32_t k = (i < n) ? keys[i] : SSB_EMPTY_KEY;
    bool won = false;
    uint32_t h = SSB_INVALID_SLOT;
    if (k != SSB_EMPTY_KEY) {
      const unsigned long long mine = ((unsigned long long)epoch << 32) | k;
      h = slot_hash(k) & mask;
      while (true) {
        unsigned long long e = *((volatile unsigned long long*)(set + h));
        if (e == mine) break;
        if ((uint32_t)(e >> 32) != epoch) {
          unsigned long long old = atomicCAS(set + h, e, mine);
          if (old == e) { won = true; break; }
          if (old == mine) break;
          if ((uint32_t)(old >> 32) != epoch) continue;  // raced with a stale->stale change; retry
        }
        h = (h + 1) & mask;
      }
    }
    // warp-aggregated id allocation for the winners
    unsigned wmask = __ballot_sync(0xffffffffu, won);
    if (wmask) {
      int leader = __ffs(wmask) - 1;
      uint32_t base = 0;
      if (lane == leader) base = atomicAdd(n_unique, (uint32_t)__popc(wmask));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (won) {
        uint32_t id = base + __popc(wmask & ((1u << lane) - 1));
        unique[id] = k;
        set_ids[h] = id;
      }
    }
    if (i < n) inverse[i] = h;  // set slot for now; resolved to the id by the second kernel
  }
}

__global__ void __launch_bounds__(256)
dedup_resolve_kernel(const uint32_t* __restrict__ set_ids, uint32_t* __restrict__ inverse, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint32_t h = inverse[i];
    if (h != SSB_INVALID_SLOT) inverse[i] = set_ids[h];
  }
}

}  // namespace ssb

extern "C" int ssb_dedup(const uint32_t* keys, int64_t n, uint64_t* set, uint32_t* set_ids,
                         uint32_t set_cap, uint32_t epoch, uint32_t* unique, uint32_t* inverse,
                         uint32_t* n_unique, int reset, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (set_cap == 0 || (set_cap & (set_cap - 1)) != 0 || (int64_t)set_cap < 2 * n || epoch == 0)
    return (int)cudaErrorInvalidValue;
  if (reset) {
    cudaError_t e = cudaMemsetAsync(n_unique, 0, sizeof(uint32_t), s);
    if (e != cudaSuccess) return (int)e;
  }
  if (n <= 0) return 0;
  int blocks = ssb::grid_for(n, 256 * 2);
  ssb::dedup_insert_kernel<<<blocks, 256, 0, s>>>(keys, n, (unsigned long long*)set, set_ids,
                                                  set_cap - 1, epoch, unique, inverse, n_unique);
  SSB_LAUNCH_CHECK();
  ssb::dedup_resolve_kernel<<<blocks, 256, 0, s>>>(set_ids, inverse, n);
  SSB_LAUNCH_CHECK();
  return 0;
}
