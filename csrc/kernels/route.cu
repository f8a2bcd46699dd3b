// route.cu — K1: key hashing, fragment routing and per-owner bucketing.
// ref: utils/HashFunction.h:16-24 (fmix64), parameter/hashfrag.h:48-53 (to_node_id),
//      parameter/global_pull_access.h:58-72 (arrange_local_vals: bucket keys per server).
// The reference walks a std::map per key on the CPU; here one thread hashes a key, block-level
// shared-memory histograms cut the global atomics to num_owners per block.
#include "common.cuh"

namespace ssb {

int g_blocks_per_sm_cap = 0;

__global__ void fmix64_kernel(const uint64_t* __restrict__ in, uint64_t* __restrict__ out, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = fmix64(in[i]);
}

__global__ void __launch_bounds__(256)
hash_route_kernel(const uint32_t* __restrict__ keys, int64_t n, const uint32_t* __restrict__ frag_map,
                  uint32_t frag_num, int num_owners, int32_t* __restrict__ owner,
                  uint32_t* __restrict__ counts) {
  __shared__ uint32_t hist[SSB_MAX_SHARDS];
  if (threadIdx.x < SSB_MAX_SHARDS) hist[threadIdx.x] = 0;
  __syncthreads();
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint32_t k = keys[i];
    int o = (k == SSB_EMPTY_KEY) ? -1 : owner_of(k, frag_map, frag_num);
    owner[i] = o;
    if (o >= 0 && o < num_owners) atomicAdd(&hist[o], 1u);
  }
  __syncthreads();
  if (threadIdx.x < num_owners && hist[threadIdx.x]) atomicAdd(&counts[threadIdx.x], hist[threadIdx.x]);
}

// Scatter keys into per-owner contiguous ranges. Warp-aggregated cursor claims: lanes with the
// same owner elect a leader that does one atomicAdd for the group.
__global__ void __launch_bounds__(256)
bucket_scatter_kernel(const uint32_t* __restrict__ keys, const int32_t* __restrict__ owner, int64_t n,
                      const uint32_t* __restrict__ offsets, uint32_t* __restrict__ cursor,
                      uint32_t* __restrict__ bucketed, uint32_t* __restrict__ perm) {
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t n_round = (n + 31) / 32 * 32;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n_round; i += stride) {
    int o = (i < n) ? owner[i] : -1;
    unsigned active = __ballot_sync(0xffffffffu, o >= 0);
    if (o >= 0) {
      unsigned peers = __match_any_sync(active, o);
      int leader = __ffs(peers) - 1;
      int rank = __popc(peers & ((1u << (threadIdx.x & 31)) - 1));
      uint32_t base = 0;
      if ((threadIdx.x & 31) == leader) base = atomicAdd(&cursor[o], (uint32_t)__popc(peers));
      base = __shfl_sync(peers, base, leader);
      uint32_t dst = offsets[o] + base + rank;
      bucketed[dst] = keys[i];
      if (perm) perm[dst] = (uint32_t)i;
    }
  }
}

__global__ void fill_u32_kernel(uint32_t* p, uint32_t v, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    p[i] = v;
}

}  // namespace ssb

extern "C" {

int ssb_fmix64(const uint64_t* in, uint64_t* out, int64_t n, void* stream) {
  if (n <= 0) return 0;
  ssb::fmix64_kernel<<<ssb::grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(in, out, n);
  SSB_LAUNCH_CHECK();
  return 0;
}

int ssb_hash_route(const uint32_t* keys, int64_t n, const uint32_t* frag_map, uint32_t frag_num,
                   int32_t num_owners, int32_t* owner, uint32_t* counts, void* stream) {
  if (n <= 0) return 0;
  if (num_owners > SSB_MAX_SHARDS || frag_num == 0) return (int)cudaErrorInvalidValue;
  ssb::hash_route_kernel<<<ssb::grid_for(n, 256 * 8), 256, 0, (cudaStream_t)stream>>>(
      keys, n, frag_map, frag_num, num_owners, owner, counts);
  SSB_LAUNCH_CHECK();
  return 0;
}

int ssb_bucket_scatter(const uint32_t* keys, const int32_t* owner, int64_t n, const uint32_t* offsets,
                       uint32_t* cursor, uint32_t* bucketed, uint32_t* perm, void* stream) {
  if (n <= 0) return 0;
  ssb::bucket_scatter_kernel<<<ssb::grid_for(n, 256 * 4), 256, 0, (cudaStream_t)stream>>>(
      keys, owner, n, offsets, cursor, bucketed, perm);
  SSB_LAUNCH_CHECK();
  return 0;
}

int ssb_fill_u32(uint32_t* p, uint32_t v, int64_t n, void* stream) {
  if (n <= 0) return 0;
  ssb::fill_u32_kernel<<<ssb::grid_for(n, 256 * 4), 256, 0, (cudaStream_t)stream>>>(p, v, n);
  SSB_LAUNCH_CHECK();
  return 0;
}

// 0 removes the cap. Host-side setting, applies to launches issued afterwards.
void ssb_set_blocks_per_sm_cap(int blocks_per_sm) { ssb::g_blocks_per_sm_cap = blocks_per_sm < 0 ? 0 : blocks_per_sm; }

const char* ssb_error_string(int err) { return cudaGetErrorString((cudaError_t)err); }
int ssb_version(void) { return 100; }

}  // extern "C"
