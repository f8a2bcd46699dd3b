// common.cuh — shared device helpers for the sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/ssb200.h"
#include "../../include/ssb200_rule.cuh"

#define SSB_LAUNCH_CHECK()                                 \
  do {                                                     \
    cudaError_t e__ = cudaGetLastError();                  \
    if (e__ != cudaSuccess) return (int)e__;               \
  } while (0)

namespace ssb {

constexpr int kWarp = 32;
constexpr int kNumSMs = 148;  // B200: 2 dies x 74

// MurmurHash3 fmix64 finaliser. ref: utils/HashFunction.h:16-24 — used for key -> fragment.
__host__ __device__ __forceinline__ uint64_t fmix64(uint64_t x) {
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ULL;
  x ^= x >> 33;
  return x;
}

// Second, independent hash for the intra-shard bucket (the reference reuses fmix64 % shard_num,
// sparsetable.h:115; an independent mix avoids correlating bucket choice with the owner choice).
__host__ __device__ __forceinline__ uint32_t slot_hash(uint32_t key) {
  uint32_t h = key * 0x9E3779B1u;
  h ^= h >> 16;
  h *= 0x85ebca6bu;
  h ^= h >> 13;
  h *= 0xc2b2ae35u;
  h ^= h >> 16;
  return h;
}

// Lemire range reduction: uniform h in [0,2^32) -> [0,n)
__host__ __device__ __forceinline__ uint32_t reduce_range(uint32_t h, uint32_t n) {
  return (uint32_t)(((uint64_t)h * (uint64_t)n) >> 32);
}

__device__ __forceinline__ int owner_of(uint32_t key, const uint32_t* __restrict__ frag_map,
                                        uint32_t frag_num) {
  uint64_t h = fmix64((uint64_t)key);
  // 64-bit modulo keeps bit-exact parity with hashfrag.h:48-53; power-of-two fast path.
  uint32_t f = ((frag_num & (frag_num - 1)) == 0) ? (uint32_t)(h & (uint64_t)(frag_num - 1))
                                                  : (uint32_t)(h % (uint64_t)frag_num);
  return (int)__ldg(frag_map + f) - 1;
}

// Deterministic lazy row init: element j of key's row. Stateless so any GPU computes the same
// value (ref: PullAccessMethod::init_param, sparse_access_method.h:21; vec1.h:223-226 randInit).
__host__ __device__ __forceinline__ float init_value(uint32_t key, uint32_t j, uint64_t seed,
                                                     float scale) {
  uint64_t h = fmix64(((uint64_t)key << 32 | (uint64_t)j) ^ seed);
  float u = (float)(uint32_t)(h >> 40) * (1.0f / 16777216.0f);  // [0,1)
  return (u - 0.5f) * scale;
}

// Lazy init of element j of key's row under the table's init kind (SSB_INIT_*).
__device__ __forceinline__ float init_of(const SsbTableView& v, uint32_t key, uint32_t j) {
  switch (v.init_kind) {
    case SSB_INIT_ZERO: return 0.f;
    case SSB_INIT_CONSTANT: return v.init_scale;
    case SSB_INIT_UNIFORM_POS: return (init_value(key, j, v.seed, 1.0f) + 0.5f) * v.init_scale;
    default: return init_value(key, j, v.seed, v.init_scale);
  }
}

// ---- memory-model helpers (system scope: the other side may be a peer GPU over NVLink) ----
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t cas_sys(uint32_t* p, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.relaxed.sys.global.cas.b32 %0, [%1], %2, %3;"
               : "=r"(old)
               : "l"(p), "r"(cmp), "r"(val)
               : "memory");
  return old;
}
// ld_row_f4/f1, red_add_f4/f1_sys, atom_add_f4/f1_sys: include/ssb200_rule.cuh (public: user update rules use them)
// Local (gpu-scope) vector reduction — gradient accumulation into the worker's own buffer.
__device__ __forceinline__ void red_add_f4_gpu(float* p, float4 v) {
  asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ float fast_sigmoid(float x) {
  return __fdividef(1.0f, 1.0f + __expf(-x));
}
// -log(sigmoid(x)) = softplus(-x), numerically stable
__device__ __forceinline__ float neg_log_sigmoid(float x) {
  return fmaxf(-x, 0.0f) + __logf(1.0f + __expf(-fabsf(x)));
}

// Philox4x32-10 counter RNG (K9). Stateless: (seed, counter) -> 4 x uint32.
struct Philox {
  static constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
  __host__ __device__ static inline void round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
#ifdef __CUDA_ARCH__
    uint32_t hi0 = __umulhi(M0, c[0]), hi1 = __umulhi(M1, c[2]);
#else
    uint32_t hi0 = (uint32_t)(((uint64_t)M0 * c[0]) >> 32), hi1 = (uint32_t)(((uint64_t)M1 * c[2]) >> 32);
#endif
    uint32_t lo0 = M0 * c[0], lo1 = M1 * c[2];
    uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
  }
  __host__ __device__ static inline void gen(uint64_t seed, uint64_t ctr, uint32_t (&out)[4]) {
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    out[0] = (uint32_t)ctr; out[1] = (uint32_t)(ctr >> 32); out[2] = 0x5353425fu; out[3] = 0x42323030u;
#pragma unroll
    for (int i = 0; i < 10; ++i) { round(out, k0, k1); k0 += W0; k1 += W1; }
  }
};

// Optional cap on resident blocks per SM for the (grid-stride) table / dedup / sampling kernels.
// In the pipelined worker these kernels run next to the persistent tensor-core kernel of the
// previous step; a grid that fills every SM would make that kernel's CTAs wait for them to drain.
extern int g_blocks_per_sm_cap;   // 0 = no cap (defined in route.cu)

inline int grid_for(int64_t work_items, int per_block, int max_blocks = kNumSMs * 16) {
  int64_t b = (work_items + per_block - 1) / per_block;
  if (b < 1) b = 1;
  if (g_blocks_per_sm_cap > 0 && max_blocks > kNumSMs * g_blocks_per_sm_cap) max_blocks = kNumSMs * g_blocks_per_sm_cap;
  if (b > max_blocks) b = max_blocks;
  return (int)b;
}

}  // namespace ssb
