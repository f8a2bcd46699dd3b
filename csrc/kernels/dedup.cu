// dedup.cu — K2: key de-duplication with an inverse index.
// ref: parameter/global_param_cache.h:107-108 (params/grads are hash maps => one entry per
//      unique key), global_push_access.h:80-99 (one (key,grad) per unique key is pushed).
// GPU form: an open-addressing hash *set* in scratch memory. Entries are 64-bit
// (epoch << 32 | key); an entry whose epoch differs from the current call's epoch counts as
// empty, so the scratch never needs clearing between calls.
#include <cstdio>
#include "common.cuh"

namespace ssb {

// Each block owns kDedupKeysPerBlock consecutive keys (kDedupKPT per thread). New keys get their ids
// from ONE atomicAdd per block: the winners are counted with a block-wide scan first. (One atomic
// per warp — 17 K same-address atomics for a 557 K-key word2vec batch — serialised in a single L2
// slice and made this kernel latency-bound: 39 us at < 10 % of every throughput metric.)
constexpr int kDedupKPT = 2;
constexpr int kDedupKeysPerBlock = 256 * kDedupKPT;

__global__ void __launch_bounds__(256)
dedup_insert_kernel(const uint32_t* __restrict__ keys, int64_t n, unsigned long long* set,
                    uint32_t* __restrict__ set_ids, uint32_t mask, uint32_t epoch,
                    uint32_t* __restrict__ unique, uint32_t* __restrict__ inverse,
                    uint32_t* __restrict__ n_unique) {
  __shared__ uint32_t warp_off[8];
  __shared__ uint32_t block_base;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t i0 = (int64_t)blockIdx.x * kDedupKeysPerBlock + threadIdx.x;
  uint32_t k[kDedupKPT], h[kDedupKPT];
#pragma unroll
  for (int j = 0; j < kDedupKPT; ++j) {
    int64_t i = i0 + j * 256;
    k[j] = (i < n) ? keys[i] : SSB_EMPTY_KEY;
  }
  unsigned won = 0;
#pragma unroll
  for (int j = 0; j < kDedupKPT; ++j) {
    h[j] = SSB_INVALID_SLOT;
    if (k[j] == SSB_EMPTY_KEY) continue;
    const unsigned long long mine = ((unsigned long long)epoch << 32) | k[j];
    uint32_t hh = slot_hash(k[j]) & mask;
    while (true) {
      unsigned long long e = *((volatile unsigned long long*)(set + hh));
      if (e == mine) break;
      if ((uint32_t)(e >> 32) != epoch) {
        unsigned long long old = atomicCAS(set + hh, e, mine);
        if (old == e) { won |= 1u << j; break; }
        if (old == mine) break;
        if ((uint32_t)(old >> 32) != epoch) continue;  // raced with a stale->stale change; retry
      }
      hh = (hh + 1) & mask;
    }
    h[j] = hh;
  }
  // block-wide exclusive scan of the winner counts, then one id-range claim for the whole block
  const uint32_t cnt = (uint32_t)__popc(won);
  uint32_t incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) warp_off[warp] = incl;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) { uint32_t t = warp_off[w]; warp_off[w] = run; run += t; }
    block_base = run ? atomicAdd(n_unique, run) : 0u;
  }
  __syncthreads();
  uint32_t id = block_base + warp_off[warp] + incl - cnt;
#pragma unroll
  for (int j = 0; j < kDedupKPT; ++j) {
    if (won & (1u << j)) {
      unique[id] = k[j];
      set_ids[h[j]] = id;
      ++id;
    }
    int64_t i = i0 + j * 256;
    if (i < n) inverse[i] = h[j];  // set slot for now; resolved to the id by the second kernel
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
  int blocks = (int)((n + ssb::kDedupKeysPerBlock - 1) / ssb::kDedupKeysPerBlock);
  ssb::dedup_insert_kernel<<<blocks, 256, 0, s>>>(keys, n, (unsigned long long*)set, set_ids,
                                                  set_cap - 1, epoch, unique, inverse, n_unique);
  SSB_LAUNCH_CHECK();
  ssb::dedup_resolve_kernel<<<ssb::grid_for(n, 256 * 2), 256, 0, s>>>(set_ids, inverse, n);
  SSB_LAUNCH_CHECK();
  return 0;
}

// -------------------------------------------------------------------------------------------
// ssb_dedup_static — the same de-duplication for callers that must not allocate.
//
// ssb_dedup above needs a scratch hash set that somebody has to allocate on the device. Two kinds
// of caller cannot do that: code embedded in a process whose device memory is entirely owned and
// accounted for by its own arena or pool (a serving process, or a framework whose caching allocator
// must not be bypassed, hands this library raw buffers and forbids any further cudaMalloc so its
// memory budget stays exact), and code on a latency-critical path, where cudaMalloc/cudaFree would
// stall the whole device. For them the scratch lives in the library image itself: a file-scope
// __device__ array of fixed size. Every GPU that loads the library carries its own copy of that
// array (1.5 MiB) from load until the process exits.
//
// Limits and rules:
//   * at most SSB_DEDUP_STATIC_MAX_KEYS (65 536) keys per call; larger batches are refused with
//     cudaErrorInvalidValue and a message on stderr — use ssb_dedup with your own scratch instead;
//   * the array's name is only a host-side handle, not a device address, so the first call on each
//     device (with that device current) asks the runtime for the real address with
//     cudaGetSymbolAddress and remembers it; make that first call outside any stream capture;
//   * nothing here calls cudaMalloc, cudaFree or any allocating library routine: the only runtime
//     calls are the address query, two cudaMemsetAsync on the caller's stream and the launches;
//   * all calls on a device share the one array and it keeps whatever the previous call left in
//     it; each call therefore clears the hash-set part (and the output counter) on the caller's
//     stream before inserting. Two calls that overlap on the GPU — different streams, or different
//     host threads — race on the shared scratch: serialise them.
#include <mutex>

namespace ssb {
constexpr uint32_t kStaticDedupMaxKeys = 65536;
constexpr uint32_t kStaticDedupCap = 2 * kStaticDedupMaxKeys;  // hash-set slots (power of two)
// [0, kStaticDedupCap) 64-bit set entries, then kStaticDedupCap 32-bit ids packed into 64-bit words
static __device__ __align__(256) unsigned long long g_static_dedup_scratch[kStaticDedupCap + kStaticDedupCap / 2];

static std::mutex g_static_dedup_mutex;
static unsigned long long* g_static_dedup_ptr[64] = {nullptr};  // per device ordinal
}  // namespace ssb

extern "C" int ssb_dedup_static(const uint32_t* keys, int64_t n, uint32_t* unique, uint32_t* inverse,
                                uint32_t* n_unique, void* stream) {
  using namespace ssb;
  cudaStream_t s = (cudaStream_t)stream;
  if (n < 0 || n > (int64_t)kStaticDedupMaxKeys) {
    fprintf(stderr, "ssb_dedup_static: n = %lld exceeds the built-in scratch (max %u keys per call); "
                    "use ssb_dedup with caller-provided scratch\n", (long long)n, kStaticDedupMaxKeys);
    return (int)cudaErrorInvalidValue;
  }
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= 64) return (int)cudaErrorInvalidDevice;
  unsigned long long* scratch = nullptr;
  {
    // one-time set-up per device: turn the symbol into this device's address
    std::lock_guard<std::mutex> lk(g_static_dedup_mutex);
    if (g_static_dedup_ptr[dev] == nullptr) {
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_static_dedup_scratch);
      if (e != cudaSuccess) return (int)e;
      g_static_dedup_ptr[dev] = static_cast<unsigned long long*>(p);
    }
    scratch = g_static_dedup_ptr[dev];
  }
  unsigned long long* set = scratch;
  uint32_t* set_ids = reinterpret_cast<uint32_t*>(scratch + kStaticDedupCap);
  // the set must start empty (epoch 0 everywhere) and the counter at zero; set_ids needs no
  // clearing: an id is only read for a slot whose set entry was written by this call
  e = cudaMemsetAsync(set, 0, sizeof(unsigned long long) * kStaticDedupCap, s);
  if (e != cudaSuccess) return (int)e;
  e = cudaMemsetAsync(n_unique, 0, sizeof(uint32_t), s);
  if (e != cudaSuccess) return (int)e;
  if (n == 0) return 0;
  int blocks = (int)((n + kDedupKeysPerBlock - 1) / kDedupKeysPerBlock);
  dedup_insert_kernel<<<blocks, 256, 0, s>>>(keys, n, set, set_ids, kStaticDedupCap - 1, /*epoch=*/1u, unique,
                                             inverse, n_unique);
  SSB_LAUNCH_CHECK();
  dedup_resolve_kernel<<<grid_for(n, 256 * 2), 256, 0, s>>>(set_ids, inverse, n);
  SSB_LAUNCH_CHECK();
  return 0;
}
