// table.cu — K3/K4/K7/K8/K10/K11: the HBM-resident sharded sparse table.
//
// ref: parameter/sparsetable.h:69-121 (SparseTable), :123-160 (PullAccessAgent: lazy
//      init_param on miss), :162-198 (PushAccessAgent::apply_push_value, Hogwild),
//      parameter/global_pull_access.h:40-120 + system/server/init.h:49-69 (pull RPC),
//      parameter/global_push_access.h:36-149 + system/server/init.h:116-132 (push RPC).
//
// Design (not a translation): a shard is three flat arrays in HBM — keys[cap], rows[cap][D],
// state[cap][D] — mapped into every worker process (own memory or a peer's over NVLink).
// A pull is ONE kernel: hash the key to its owner shard, probe the owner's 128-byte key bucket
// with one warp-wide load (local or remote), claim+initialise the row if it is missing, and
// gather the row straight into the worker's cache — no request/response messages, no host.
// A push is ONE kernel: vector `red`/`atom` of the update straight into the owner's row.
//
// Slot claim protocol (readers never observe a half-initialised row):
//   CAS(key slot: EMPTY -> LOCKED) ; write init row ; st.release.sys(key slot = key).
// Probers that see LOCKED in their bucket re-read until it resolves (the holder is a resident
// warp that never waits, so this cannot deadlock). No deletion, no rehash: capacity is fixed,
// which is also what makes caching (shard, slot) on the worker legal.
#include <cuda_bf16.h>
#include <algorithm>
#include <cstdlib>
#include "common.cuh"
#include "table_probe.cuh"

namespace ssb {

__device__ __forceinline__ uint2 pack_bf16x4(float4 v) {
  __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
  uint2 pk;
  pk.x = *reinterpret_cast<uint32_t*>(&lo);
  pk.y = *reinterpret_cast<uint32_t*>(&hi);
  return pk;
}
__device__ __forceinline__ uint2 ld_shadow_u2(const void* p) {
  uint2 r;
  asm volatile("ld.relaxed.sys.global.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ uint32_t effective_n(int64_t n, const uint32_t* n_dev) {
  uint32_t ne = (uint32_t)n;
  if (n_dev) { uint32_t d = *n_dev; if (d < ne) ne = d; }
  return ne;
}

// ---- pull ---------------------------------------------------------------------------------
// A warp works on kPullK keys at a time so that kPullK bucket probes, then kPullK row reads, are
// in flight together (one key at a time leaves HBM / NVLink latency exposed twice per key).
// Fast path: the first bucket is probed with relaxed system-scope loads and, on a hit, the row
// is read through an address that depends on the loaded key word; rows are only ever written
// before their key is published (st.release.sys) and are read with strong .sys loads that bypass
// L1, so the dependent read observes the initialised row. Anything else — miss in the first
// bucket, LOCKED slot, insert — takes the acquire-based probe_insert() slow path.
constexpr int kPullK = 4;

// Direct-mapped worker-side slot cache: 2^bits entries of (key << 32 | global slot); empty = all ones
// (0xFFFFFFFF is not a legal key). Fibonacci hashing spreads dense word ids without collisions.
__device__ __forceinline__ uint32_t slot_cache_index(uint32_t key, int bits) {
  return (key * 0x9E3779B1u) >> (32 - bits);
}

__device__ __forceinline__ void pull_emit(const SsbTableView& v, uint32_t key, bool valid, const Probe& p, uint32_t w,
                                          int lane, float* __restrict__ rows_out, __nv_bfloat16* __restrict__ rows_bf16,
                                          float* __restrict__ grads_zero, uint32_t* __restrict__ gslot_out,
                                          uint32_t* __restrict__ stats, bool have_pre, float4 preloaded) {
  const int D = v.dim;
  const bool have = p.slot != SSB_INVALID_SLOT;
  float* row = have ? v.rows[p.owner] + (size_t)p.slot * D : nullptr;
  __nv_bfloat16* shadow = (have && v.shadow[p.owner]) ? reinterpret_cast<__nv_bfloat16*>(v.shadow[p.owner]) + (size_t)p.slot * D : nullptr;
  // bf16-only pull of an existing row: read the 2-byte mirror instead of the fp32 row
  if (have && !p.is_new && shadow && !rows_out && rows_bf16 && (D & 3) == 0) {
    __nv_bfloat16* oh = rows_bf16 + (size_t)w * D;
    float* gz0 = grads_zero ? grads_zero + (size_t)w * D : nullptr;
    for (int j = lane * 4; j < D; j += 128) {
      uint2 pk = (have_pre && j < 128) ? make_uint2(__float_as_uint(preloaded.x), __float_as_uint(preloaded.y)) : ld_shadow_u2(shadow + j);
      *reinterpret_cast<uint2*>(oh + j) = pk;
      if (gz0) *reinterpret_cast<float4*>(gz0 + j) = make_float4(0, 0, 0, 0);
    }
    if (lane == 0) gslot_out[w] = (uint32_t)p.owner * v.capacity + p.slot;
    return;
  }
  float* out = rows_out ? rows_out + (size_t)w * D : nullptr;
  __nv_bfloat16* outh = rows_bf16 ? rows_bf16 + (size_t)w * D : nullptr;
  float* gz = grads_zero ? grads_zero + (size_t)w * D : nullptr;
  const bool want = out || outh;
  if ((D & 3) == 0) {
    for (int j = lane * 4; j < D; j += 128) {
      float4 val;
      if (have && !p.is_new) {
        val = (have_pre && j < 128) ? preloaded : (want ? ld_row_f4(row + j) : make_float4(0, 0, 0, 0));
      } else if (valid) {
        val.x = init_of(v, key, j);
        val.y = init_of(v, key, j + 1);
        val.z = init_of(v, key, j + 2);
        val.w = init_of(v, key, j + 3);
        if (p.is_new) {
          *reinterpret_cast<float4*>(row + j) = val;
          if (shadow) *reinterpret_cast<uint2*>(shadow + j) = pack_bf16x4(val);
        }
      } else {
        val = make_float4(0, 0, 0, 0);
      }
      if (out) *reinterpret_cast<float4*>(out + j) = val;
      if (outh) *reinterpret_cast<uint2*>(outh + j) = pack_bf16x4(val);
      if (gz) *reinterpret_cast<float4*>(gz + j) = make_float4(0, 0, 0, 0);
    }
  } else {
    for (int j = lane; j < D; j += 32) {
      float val;
      if (have && !p.is_new) {
        val = want ? ld_row_f1(row + j) : 0.f;
      } else if (valid) {
        val = init_of(v, key, j);
        if (p.is_new) { row[j] = val; if (shadow) shadow[j] = __float2bfloat16_rn(val); }
      } else {
        val = 0.f;
      }
      if (out) out[j] = val;
      if (outh) outh[j] = __float2bfloat16_rn(val);
      if (gz) gz[j] = 0.f;
    }
  }
  if (p.is_new) publish(v, p, key, lane);
  if (lane == 0) {
    gslot_out[w] = have ? (uint32_t)p.owner * v.capacity + p.slot : SSB_INVALID_SLOT;
    if (stats) {
      if (p.is_new) atomicAdd(stats + 0, 1u);
      if (valid && !have) atomicAdd(stats + 1, 1u);
    }
  }
}

struct PullArgs {
  uint32_t ne;
  uint32_t* gslot_out;
  float* rows_out;
  __nv_bfloat16* rows_bf16;
  float* grads_zero;
  int insert;
  uint32_t* stats;
  unsigned long long* slot_cache;
  int cache_bits;       // negative: dense key space, entry index = key & mask (no two keys share an entry)
};

__device__ __forceinline__ uint32_t cache_index(uint32_t key, int cache_bits) {
  return cache_bits < 0 ? (key & ((1u << -cache_bits) - 1u)) : slot_cache_index(key, cache_bits);
}

// One group of kPullK keys [w0, w0 + kPullK): cslot[k] = global slot from the worker's slot cache or
// SSB_INVALID_SLOT. All of key / cslot are warp-uniform.
__device__ __forceinline__ void pull_group(const SsbTableView& v, const PullArgs& a, uint32_t w0, const uint32_t (&key)[kPullK],
                                           const uint32_t (&cslot)[kPullK], int lane) {
  const int D = v.dim;
  const bool want = a.rows_out || a.rows_bf16;
  const bool fast_rows = want && (D & 3) == 0;
  const bool from_shadow = !a.rows_out && a.rows_bf16 && v.shadow[0] != nullptr && (D & 3) == 0;
  const uint32_t nb = v.capacity / SSB_BUCKET;
  uint32_t kv[kPullK], bucket[kPullK];   // bucket[k] holds the SLOT (within the owner) for cache hits
  int owner[kPullK];
  bool valid[kPullK];
  unsigned cached = 0;                   // bit k: key k was resolved by the slot cache
  // phase 1: independent first-bucket probes for the keys the cache did not resolve
#pragma unroll
  for (int k = 0; k < kPullK; ++k) {
    valid[k] = key[k] < SSB_LOCKED_KEY && w0 + k < a.ne;
    kv[k] = SSB_EMPTY_KEY;
    if (valid[k] && cslot[k] != SSB_INVALID_SLOT) {
      cached |= 1u << k;
      owner[k] = (int)(cslot[k] / v.capacity);
      bucket[k] = cslot[k] - (uint32_t)owner[k] * v.capacity;
      continue;
    }
    owner[k] = valid[k] ? owner_of(key[k], v.frag_map, v.frag_num) : 0;
    if (owner[k] < 0 || owner[k] >= v.num_shards) { owner[k] = 0; valid[k] = valid[k] && false; }
    bucket[k] = reduce_range(slot_hash(key[k]), nb);
    if (valid[k]) kv[k] = ld_relaxed_sys(v.keys[owner[k]] + (size_t)bucket[k] * SSB_BUCKET + lane);
  }
  // phase 2: resolve hits, issue their row reads together
  Probe p[kPullK];
  float4 pre[kPullK];
#pragma unroll
  for (int k = 0; k < kPullK; ++k) pre[k] = make_float4(0.f, 0.f, 0.f, 0.f);
  bool fast[kPullK];
#pragma unroll
  for (int k = 0; k < kPullK; ++k) {
    p[k].owner = owner[k]; p[k].slot = SSB_INVALID_SLOT; p[k].is_new = false;
    const bool hit = (cached >> k) & 1u;
    unsigned m = __ballot_sync(0xffffffffu, valid[k] && kv[k] == key[k]);
    fast[k] = hit || m != 0;
    if (fast[k]) {
      p[k].slot = hit ? bucket[k] : bucket[k] * SSB_BUCKET + (__ffs(m) - 1);
      if (from_shadow) {
        if (lane * 4 < D) {
          uint2 pk = ld_shadow_u2(reinterpret_cast<const __nv_bfloat16*>(v.shadow[owner[k]]) + (size_t)p[k].slot * D + lane * 4);
          pre[k].x = __uint_as_float(pk.x); pre[k].y = __uint_as_float(pk.y);
        }
      } else if (fast_rows && lane * 4 < D) {
        pre[k] = ld_row_f4(v.rows[owner[k]] + (size_t)p[k].slot * D + lane * 4);
      }
    }
  }
  // phase 3: slow path for the rest (miss in first bucket, locked slot, insert), then emit
#pragma unroll
  for (int k = 0; k < kPullK; ++k) {
    const uint32_t w = w0 + k;
    if (w >= a.ne) break;
    if (!fast[k] && valid[k]) p[k] = probe_insert(v, key[k], a.insert != 0, lane);
    pull_emit(v, key[k], valid[k], p[k], w, lane, a.rows_out, a.rows_bf16, a.grads_zero, a.gslot_out, a.stats,
              fast[k] && fast_rows, pre[k]);
    // remember the slot (after pull_emit: a row this warp created has been published by now)
    if (a.slot_cache && lane == 0 && valid[k] && !((cached >> k) & 1u) && p[k].slot != SSB_INVALID_SLOT)
      a.slot_cache[cache_index(key[k], a.cache_bits)] =
          ((unsigned long long)key[k] << 32) | ((uint32_t)p[k].owner * v.capacity + p[k].slot);
  }
}

// kCached = false: every key probes its owner's bucket (K probes, then K row reads, in flight together).
// kCached = true: a warp takes 32 keys at a time; each LANE looks one key up in the worker's own
// key -> (shard, slot) cache (local memory, one dependent access per 32 keys instead of one per key),
// then the warp walks the 32 keys in groups of K: hits go straight to their row reads — for a remote
// owner ONE NVLink round trip per key instead of two — misses take the probing path and are remembered.
// Slots never move and are cached only once published, so an entry stays true.
template <bool kCached>
__global__ void __launch_bounds__(256, 4)
pull_kernel(const __grid_constant__ SsbTableView v, const uint32_t* __restrict__ keys, int64_t n,
            const uint32_t* __restrict__ n_dev, const PullArgs args) {
  const int lane = threadIdx.x & 31;
  PullArgs a = args;
  a.ne = effective_n(n, n_dev);
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  const uint32_t warp_id = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (!kCached) {
    a.slot_cache = nullptr;
    for (uint32_t w0 = warp_id * kPullK; w0 < a.ne; w0 += warps * kPullK) {
      uint32_t key[kPullK], cslot[kPullK];
#pragma unroll
      for (int k = 0; k < kPullK; ++k) {
        key[k] = (w0 + k < a.ne) ? __ldg(keys + w0 + k) : SSB_EMPTY_KEY;
        cslot[k] = SSB_INVALID_SLOT;
      }
      pull_group(v, a, w0, key, cslot, lane);
    }
  } else {
    const uint32_t total_slots = v.capacity * (uint32_t)v.num_shards;
    for (uint32_t c0 = warp_id * 32; c0 < a.ne; c0 += warps * 32) {
      const uint32_t my_key = (c0 + lane < a.ne) ? __ldg(keys + c0 + lane) : SSB_EMPTY_KEY;
      uint32_t my_slot = SSB_INVALID_SLOT;
      if (my_key < SSB_LOCKED_KEY) {
        const unsigned long long ce = *((const volatile unsigned long long*)(a.slot_cache + cache_index(my_key, a.cache_bits)));
        if ((uint32_t)(ce >> 32) == my_key && (uint32_t)ce < total_slots) my_slot = (uint32_t)ce;
      }
#pragma unroll 1
      for (int g = 0; g < 32; g += kPullK) {
        if (c0 + g >= a.ne) break;
        uint32_t key[kPullK], cslot[kPullK];
#pragma unroll
        for (int k = 0; k < kPullK; ++k) {
          key[k] = __shfl_sync(0xffffffffu, my_key, g + k);
          cslot[k] = __shfl_sync(0xffffffffu, my_slot, g + k);
        }
        pull_group(v, a, c0 + g, key, cslot, lane);
      }
    }
  }
}

// ---- pull, bulk-copy form (default for bf16-mirror pulls; SSB200_PULL_BULK=0/1 forces) ------------
// The kernel above keeps every row it is waiting for in registers (a float4 per lane per key), and its
// per-key state caps a warp at kPullK = 4 rows in flight: 128 rows per SM at 64 registers, which is what
// bounds it (its time scales with 1 / resident warps: 3.7 TB/s of 6.5). Here a LANE resolves one key
// (slot cache, else one pass over the key's first bucket with four 16-byte loads ... eight in total) and
// starts ONE bulk copy (cp.async.bulk, the TMA engine's 1-D form) of its row into the warp's slice of
// shared memory; the 32 copies of a warp complete on the warp's mbarrier. No register is held while a
// row is in flight: 32 rows per warp, 384 per SM. Keys the fast path cannot resolve (not in the first
// bucket, locked, to be inserted, illegal) go through probe_insert / pull_emit exactly as above while
// the copies fly. Rows must be 16-byte multiples of at most 512 bytes (D % 4 == 0, D <= 128).
constexpr int kBulkWarps = 4;

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bulk_mbar_init(unsigned long long* bar) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_addr(bar)) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void bulk_mbar_expect(unsigned long long* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_mbar_wait(unsigned long long* bar, uint32_t parity) {   // bounded: traps instead of hanging
  uint32_t done = 0;
  for (uint32_t it = 0; it < (1u << 26); ++it) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(smem_addr(bar)), "r"(parity) : "memory");
    if (done) return;
  }
  __trap();
}
__device__ __forceinline__ void bulk_copy_g2s(void* dst_smem, const void* src, uint32_t bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_addr(dst_smem)), "l"(src), "r"(bytes), "r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ uint4 ld_relaxed_sys_v4(const uint4* p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}

template <bool kCached>
__global__ void __launch_bounds__(kBulkWarps * 32)
pull_bulk_kernel(const __grid_constant__ SsbTableView v, const uint32_t* __restrict__ keys, int64_t n,
                 const uint32_t* __restrict__ n_dev, const PullArgs args) {
  extern __shared__ __align__(128) uint8_t bulk_smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  PullArgs a = args;
  a.ne = effective_n(n, n_dev);
  if (!kCached) a.slot_cache = nullptr;
  const int D = v.dim;
  const bool from_shadow = !a.rows_out && a.rows_bf16 && v.shadow[0] != nullptr;
  const uint32_t row_bytes = (uint32_t)D * (from_shadow ? 2u : 4u);
  uint8_t* my_rows = bulk_smem + (size_t)warp * 32 * row_bytes;
  unsigned long long* bar = reinterpret_cast<unsigned long long*>(bulk_smem + (size_t)kBulkWarps * 32 * row_bytes) + warp;
  if (lane == 0) bulk_mbar_init(bar);
  __syncwarp();
  uint32_t phase = 0;
  const uint32_t nb = v.capacity / SSB_BUCKET;
  const uint32_t total_slots = v.capacity * (uint32_t)v.num_shards;
  const uint32_t warps = gridDim.x * kBulkWarps;
  for (uint32_t c0 = (blockIdx.x * kBulkWarps + warp) * 32; c0 < a.ne; c0 += warps * 32) {
    const uint32_t w = c0 + lane;
    const uint32_t key = (w < a.ne) ? __ldg(keys + w) : SSB_EMPTY_KEY;
    uint32_t gs = SSB_INVALID_SLOT;                 // global slot of MY key once the fast path has it
    if (key < SSB_LOCKED_KEY) {
      if (kCached) {
        const unsigned long long ce = *((const volatile unsigned long long*)(a.slot_cache + cache_index(key, a.cache_bits)));
        if ((uint32_t)(ce >> 32) == key && (uint32_t)ce < total_slots) gs = (uint32_t)ce;
      }
      if (gs == SSB_INVALID_SLOT) {
        const int o = owner_of(key, v.frag_map, v.frag_num);
        if (o >= 0 && o < v.num_shards) {
          const uint32_t b = reduce_range(slot_hash(key), nb);
          const uint4* bp = reinterpret_cast<const uint4*>(v.keys[o] + (size_t)b * SSB_BUCKET);
          int idx = -1;
#pragma unroll
          for (int q = 0; q < SSB_BUCKET / 4; ++q) {
            // relaxed, as in pull_kernel's fast path (an acquire costs an L1 invalidate per load): a row is
            // written before its key is published with st.release.sys, and the copy below is issued on an
            // address that depends on the key word read here and is served from L2 / HBM, never from L1
            const uint4 kq = ld_relaxed_sys_v4(bp + q);
            if (kq.x == key) idx = q * 4;
            if (kq.y == key) idx = q * 4 + 1;
            if (kq.z == key) idx = q * 4 + 2;
            if (kq.w == key) idx = q * 4 + 3;
          }
          if (idx >= 0) {
            gs = (uint32_t)o * v.capacity + b * SSB_BUCKET + (uint32_t)idx;
            if (kCached) a.slot_cache[cache_index(key, a.cache_bits)] = ((unsigned long long)key << 32) | gs;
          }
        }
      }
    }
    const unsigned fast = __ballot_sync(0xffffffffu, gs != SSB_INVALID_SLOT);
    if (fast) {
      if (lane == 0) bulk_mbar_expect(bar, (uint32_t)__popc(fast) * row_bytes);
      __syncwarp();
      if (gs != SSB_INVALID_SLOT) {
        const uint32_t o = gs / v.capacity, sl = gs - o * v.capacity;
        const uint8_t* src = from_shadow ? reinterpret_cast<const uint8_t*>(v.shadow[o]) + (size_t)sl * row_bytes
                                         : reinterpret_cast<const uint8_t*>(v.rows[o]) + (size_t)sl * row_bytes;
        bulk_copy_g2s(my_rows + (size_t)lane * row_bytes, src, row_bytes, bar);
      }
    }
    // everything the fast path did not resolve: the probing path, one key at a time, while the copies fly
    unsigned todo = __ballot_sync(0xffffffffu, w < a.ne && gs == SSB_INVALID_SLOT);
    while (todo) {
      const int src_lane = __ffs(todo) - 1;
      todo &= todo - 1;
      const uint32_t k = __shfl_sync(0xffffffffu, key, src_lane);
      const bool vld = k < SSB_LOCKED_KEY;
      Probe p;
      p.owner = 0; p.slot = SSB_INVALID_SLOT; p.is_new = false;
      if (vld) p = probe_insert(v, k, a.insert != 0, lane);
      pull_emit(v, k, vld, p, c0 + src_lane, lane, a.rows_out, a.rows_bf16, a.grads_zero, a.gslot_out, a.stats, false,
                make_float4(0.f, 0.f, 0.f, 0.f));
      if (kCached && lane == 0 && vld && p.slot != SSB_INVALID_SLOT)
        a.slot_cache[cache_index(k, a.cache_bits)] = ((unsigned long long)k << 32) | ((uint32_t)p.owner * v.capacity + p.slot);
    }
    if (fast) {
      bulk_mbar_wait(bar, phase);
      phase ^= 1;
      unsigned m = fast;
      while (m) {
        const int r = __ffs(m) - 1;
        m &= m - 1;
        const uint32_t g = __shfl_sync(0xffffffffu, gs, r);
        const size_t wr = (size_t)(c0 + r);
        const uint8_t* sr = my_rows + (size_t)r * row_bytes;
        for (int j = lane * 4; j < D; j += 128) {
          if (from_shadow) {
            *reinterpret_cast<uint2*>(a.rows_bf16 + wr * D + j) = *reinterpret_cast<const uint2*>(sr + (size_t)j * 2);
          } else {
            const float4 val = *reinterpret_cast<const float4*>(sr + (size_t)j * 4);
            if (a.rows_out) *reinterpret_cast<float4*>(a.rows_out + wr * D + j) = val;
            if (a.rows_bf16) *reinterpret_cast<uint2*>(a.rows_bf16 + wr * D + j) = pack_bf16x4(val);
          }
          if (a.grads_zero) *reinterpret_cast<float4*>(a.grads_zero + wr * D + j) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (lane == 0) a.gslot_out[wr] = g;
      }
      __syncwarp();   // every lane has read its share of the slice before the next chunk's copies land in it
    }
  }
}

__global__ void __launch_bounds__(256)
gather_slots_kernel(const __grid_constant__ SsbTableView v, const uint32_t* __restrict__ gslot, int64_t n,
                    const uint32_t* __restrict__ n_dev, float* __restrict__ rows_out) {
  const int lane = threadIdx.x & 31;
  const uint32_t ne = effective_n(n, n_dev);
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  const int D = v.dim;
  for (uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < ne; w += warps) {
    uint32_t g = __ldg(gslot + w);
    float* out = rows_out + (size_t)w * D;
    if (g == SSB_INVALID_SLOT) {
      for (int j = lane; j < D; j += 32) out[j] = 0.f;
      continue;
    }
    uint32_t o = g / v.capacity, s = g - o * v.capacity;
    const float* row = v.rows[o] + (size_t)s * D;
    if ((D & 3) == 0) {
      for (int j = lane * 4; j < D; j += 128) *reinterpret_cast<float4*>(out + j) = ld_row_f4(row + j);
    } else {
      for (int j = lane; j < D; j += 32) out[j] = ld_row_f1(row + j);
    }
  }
}

// ---- push ---------------------------------------------------------------------------------
// The push kernels proper — rule_push_kernel<Rule, kShadow> (a worker's fp32 gradient rows into the owners' rows),
// rule_push_bf16_kernel<Rule> (owner-side apply of bf16 inbox segments) and rule_mirror_refresh_kernel — live in
// the public header include/ssb200_rule.cuh together with the update rules (Sgd, AdaGrad, Ftrl), so that a user's
// rule is instantiated into the same kernels without editing this file. Below: the kernels that SHIP gradient rows
// to their owners; their "rows I own myself" branch applies the rule through the header's rule_apply_row.
struct ShipOnly {};                 // Rule argument of the send kernels: ship remote rows, leave the local rows alone
template <class Rule> struct applies_rule { static constexpr bool value = true; };
template <> struct applies_rule<ShipOnly> { static constexpr bool value = false; };

// Cross-GPU hand-off flags for the inbox push (replaces a collective as the barrier): after its
// shipment kernel a rank stores the step number into flags[owner][rank] on every owner; an owner
// waits until all of its flags reached the step before applying its inbox. The waiter only spins
// on ITS OWN memory, every rank runs on its own GPU, and the spin is bounded (trap, not hang).
__global__ void flags_signal_kernel(const __grid_constant__ SsbInboxView ib, int num_shards, int self, uint32_t value) {
  __threadfence_system();
  if ((int)threadIdx.x < num_shards) st_release_sys(ib.flags[threadIdx.x] + self, value);
}
// Private-region inboxes: every sender owns a sub-region of each owner's inbox, so space is claimed
// with LOCAL atomics on `cursor[owner]` and the row counts are published once per step with plain
// peer stores (no remote atomic ever serialises on an owner's counter). value != 0 also raises the
// hand-off flags.
__global__ void inbox_publish_kernel(const __grid_constant__ SsbInboxView ib, int num_shards, int self,
                                     uint32_t* __restrict__ cursor, uint32_t value) {
  const int o = threadIdx.x;
  if (o < num_shards) {
    const uint32_t c = cursor[o];
    cursor[o] = 0;
    if (o != self) *((volatile uint32_t*)ib.count[o]) = c < ib.capacity ? c : ib.capacity;
  }
  __threadfence_system();
  __syncthreads();
  if (value != 0 && o < num_shards) st_release_sys(ib.flags[o] + self, value);
}
// spin_limit iterations of ~0.2 us each. err == nullptr: a wait that runs out traps (the launch fails); otherwise
// err[0] is set to 1 and the kernel returns, so that a straggling rank (a host stall, a backup being written) costs
// the others an error the host can read and act on instead of a sticky CUDA context error.
__global__ void flags_wait_kernel(const uint32_t* __restrict__ my_flags, int num_shards, uint32_t value, uint32_t spin_limit,
                                  uint32_t* __restrict__ err) {
  if ((int)threadIdx.x < num_shards) {
    uint32_t it = 0;
    while ((int32_t)(ld_acquire_sys(my_flags + threadIdx.x) - value) < 0) {
      __nanosleep(200);
      if (++it > spin_limit) {
        if (err) { atomicExch(err, 1u); break; }
        __trap();
      }
    }
  }
  __syncthreads();
  __threadfence_system();
}

// ---- inbox push: one-way gradient shipment, owner applies the update --------------------------------
// The AdaGrad update is not a pure reduction (it reads the accumulator), so pushing it with remote
// `atom` moves every row three times across NVLink (operand out, old value back, `red` out). Here a
// worker ships (slot, grad row) ONCE into the owner's inbox with plain peer stores; after a barrier
// the owner applies its inbox with the ordinary push kernel on local memory (full HBM speed).
// Rows the worker owns itself skip the inbox. Space in an owner's inbox is claimed with one remote
// atomicAdd per (block, owner). ref: the reference's push is exactly "send (key, grad)* to the owner,
// the owner applies" (global_push_access.h:123-149, server/init.h:116-132) — this is its NVLink form.
constexpr int kSendRows = 64;   // rows per block

template <class Rule, bool kBf16>
__global__ void __launch_bounds__(256)
push_send_kernel(const __grid_constant__ SsbTableView v, const __grid_constant__ SsbInboxView ib,
                 const uint32_t* __restrict__ gslot, const float* __restrict__ grads, int64_t n,
                 const uint32_t* __restrict__ n_dev, int self_shard, const __grid_constant__ rule::Hyper h,
                 uint32_t* __restrict__ overflow, uint32_t* __restrict__ cursor) {
  __shared__ uint32_t cnt[SSB_MAX_SHARDS], base[SSB_MAX_SHARDS];
  __shared__ uint32_t row_slot[kSendRows];
  __shared__ uint16_t row_owner[kSendRows], row_rank[kSendRows];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t ne = effective_n(n, n_dev);
  const int D = v.dim;
  for (uint32_t r0 = blockIdx.x * kSendRows; r0 < ne; r0 += gridDim.x * kSendRows) {
    if (threadIdx.x < SSB_MAX_SHARDS) cnt[threadIdx.x] = 0;
    __syncthreads();
    if (threadIdx.x < kSendRows) {
      const uint32_t r = r0 + threadIdx.x;
      uint32_t g = (r < ne) ? __ldg(gslot + r) : SSB_INVALID_SLOT;
      uint16_t o = 0xFFFF, rk = 0;
      uint32_t sl = 0;
      if (g != SSB_INVALID_SLOT) {
        o = (uint16_t)(g / v.capacity);
        sl = g - (uint32_t)o * v.capacity;
        if ((int)o != self_shard) rk = (uint16_t)atomicAdd(&cnt[o], 1u);
      }
      row_owner[threadIdx.x] = o; row_rank[threadIdx.x] = rk; row_slot[threadIdx.x] = sl;
    }
    __syncthreads();
    if (threadIdx.x < v.num_shards && cnt[threadIdx.x])   // claim once per (block, owner): local cursor (private regions) or remote counter
      base[threadIdx.x] = cursor ? atomicAdd(cursor + threadIdx.x, cnt[threadIdx.x]) : atomicAdd(ib.count[threadIdx.x], cnt[threadIdx.x]);
    __syncthreads();
    for (int k = warp; k < kSendRows; k += 8) {
      const uint32_t r = r0 + k;
      const int o = row_owner[k];
      if (o == 0xFFFF) continue;
      const float* gr = grads + (size_t)r * D;
      if (o == self_shard) {
        // local row: apply directly (this worker's gradient buffer holds a slot once: single writer, mirror in step)
        if constexpr (applies_rule<Rule>::value)
          rule_apply_row<Rule, true>(v, (uint32_t)o * v.capacity + row_slot[k], lane, h,
                                     [gr](int j) { return *reinterpret_cast<const float4*>(gr + j); }, nullptr);
      } else {
        const uint32_t idx = base[o] + row_rank[k];
        if (idx >= ib.capacity) { if (lane == 0) atomicAdd(overflow, 1u); continue; }
        if (lane == 0) ib.slots[o][idx] = row_slot[k];
        if (kBf16) {
          __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(ib.rows[o]) + (size_t)idx * D;
          for (int j = lane * 4; j < D; j += 128) {
            const float4 gv = *reinterpret_cast<const float4*>(gr + j);
            __nv_bfloat162 lo = __floats2bfloat162_rn(gv.x, gv.y), hi = __floats2bfloat162_rn(gv.z, gv.w);
            uint2 pk;
            pk.x = *reinterpret_cast<uint32_t*>(&lo); pk.y = *reinterpret_cast<uint32_t*>(&hi);
            *reinterpret_cast<uint2*>(dst + j) = pk;
          }
        } else {
          float* dst = ib.rows[o] + (size_t)idx * D;
          for (int j = lane * 4; j < D; j += 128) *reinterpret_cast<float4*>(dst + j) = *reinterpret_cast<const float4*>(gr + j);
        }
      }
    }
    __syncthreads();
  }
}

// Ship-only form of the send kernel (the own rows are applied by a separate launch next to it, cluster.Inbox): no
// block-wide phases at all. A warp takes 32 gradient rows (one per lane), groups its lanes by owner
// (__match_any_sync) so that each owner's cursor gets ONE local atomic per warp, then moves the remote rows four at
// a time: four independent 512-byte reads in flight per warp, then four converted 256-byte peer stores.
template <bool kBf16>
__global__ void __launch_bounds__(256)
push_ship_kernel(const __grid_constant__ SsbTableView v, const __grid_constant__ SsbInboxView ib,
                 const uint32_t* __restrict__ gslot, const float* __restrict__ grads, int64_t n,
                 const uint32_t* __restrict__ n_dev, int self_shard, uint32_t* __restrict__ overflow,
                 uint32_t* __restrict__ cursor) {
  const int lane = threadIdx.x & 31;
  const uint32_t ne = effective_n(n, n_dev);
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  const int D = v.dim;
  for (uint32_t w0 = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 32; w0 < ne; w0 += warps * 32) {
    const uint32_t r = w0 + lane;
    const uint32_t g = r < ne ? __ldg(gslot + r) : SSB_INVALID_SLOT;
    int o = -1;
    uint32_t sl = 0;
    if (g != SSB_INVALID_SLOT) { o = (int)(g / v.capacity); sl = g - (uint32_t)o * v.capacity; }
    const bool remote = o >= 0 && o != self_shard;
    const unsigned rm = __ballot_sync(0xffffffffu, remote);
    if (!rm) continue;
    // one cursor claim per (warp, owner)
    const unsigned peers = __match_any_sync(0xffffffffu, remote ? o : -1);
    const int leader = __ffs(peers) - 1;
    uint32_t base = 0;
    if (remote && lane == leader) base = cursor ? atomicAdd(cursor + o, (uint32_t)__popc(peers)) : atomicAdd(ib.count[o], (uint32_t)__popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    uint32_t idx = base + __popc(peers & ((1u << lane) - 1));
    if (remote && idx >= ib.capacity) { atomicAdd(overflow, 1u); idx = 0xFFFFFFFFu; }
    if (remote && idx != 0xFFFFFFFFu) ib.slots[o][idx] = sl;
    unsigned todo = __ballot_sync(0xffffffffu, remote && idx != 0xFFFFFFFFu);
    while (todo) {
      int src[4];
      float4 gv[4];
      int cnt = 0;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        src[k] = todo ? __ffs(todo) - 1 : -1;
        if (todo) { todo &= todo - 1; ++cnt; }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (k < cnt)
          for (int j = lane * 4; j < D && j < lane * 4 + 4; j += 4)       // D = 128: one float4 per lane
            gv[k] = *reinterpret_cast<const float4*>(grads + (size_t)(w0 + src[k]) * D + j);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (k >= cnt) continue;
        const int ok = __shfl_sync(0xffffffffu, o, src[k]);
        const uint32_t ik = __shfl_sync(0xffffffffu, idx, src[k]);
        if (D == 128) {
          if (kBf16) *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(ib.rows[ok]) + (size_t)ik * D + lane * 4) = pack_bf16x4(gv[k]);
          else *reinterpret_cast<float4*>(ib.rows[ok] + (size_t)ik * D + lane * 4) = gv[k];
        } else {
          const float* gr = grads + (size_t)(w0 + src[k]) * D;
          for (int j = lane * 4; j < D; j += 128) {
            const float4 x = *reinterpret_cast<const float4*>(gr + j);
            if (kBf16) *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(ib.rows[ok]) + (size_t)ik * D + j) = pack_bf16x4(x);
            else *reinterpret_cast<float4*>(ib.rows[ok] + (size_t)ik * D + j) = x;
          }
        }
      }
    }
  }
}

// ---- chunked inbox push: ship and apply overlap -----------------------------------------------------
// The inbox push above hands a whole shipment over at once (send -> publish counts -> wait for every sender ->
// apply), so NVLink shipping and the owner's HBM update never overlap. Here a sender's sub-region is cut into
// chunks of kChunkRows rows; blocks of the send kernel count the rows they have written into each chunk
// (local atomics) and the block that completes a chunk raises its ready flag on the owner (st.release.sys);
// the last block publishes the total. The owner's apply kernel — on another stream, next to its own send
// kernel — takes (sender, chunk) items as they become ready. Flags carry the step number, so nothing is
// cleared on the owner. ref: global_push_access.h:123-149 + server/init.h:116-132 ("send (key, grad)*, the
// server applies"), carried by NVLink stores and pipelined chunk by chunk.
constexpr uint32_t kChunkRows = 256;

struct SendChunkState {          // sender-local device memory, zero before the first call
  uint32_t* cursor;              // [SSB_MAX_SHARDS] rows claimed per owner this step
  uint32_t* done;                // [SSB_MAX_SHARDS][max_chunks] rows written per (owner, chunk) this step
  uint32_t* finished;            // [1] blocks of the send kernel that are done
};

template <class Rule>
__global__ void __launch_bounds__(256)
push_send_chunks_kernel(const __grid_constant__ SsbTableView v, const __grid_constant__ SsbChunkInboxView ib,
                        const uint32_t* __restrict__ gslot, const float* __restrict__ grads, int64_t n,
                        const uint32_t* __restrict__ n_dev, int self_shard, const __grid_constant__ rule::Hyper h, uint32_t tag,
                        uint32_t* __restrict__ overflow, SendChunkState st) {
  __shared__ uint32_t cnt[SSB_MAX_SHARDS], base[SSB_MAX_SHARDS];
  __shared__ uint32_t row_slot[kSendRows];
  __shared__ uint16_t row_owner[kSendRows], row_rank[kSendRows];
  __shared__ uint32_t is_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t ne = effective_n(n, n_dev);
  const int D = v.dim;
  for (uint32_t r0 = blockIdx.x * kSendRows; r0 < ne; r0 += gridDim.x * kSendRows) {
    if (threadIdx.x < SSB_MAX_SHARDS) cnt[threadIdx.x] = 0;
    __syncthreads();
    if (threadIdx.x < kSendRows) {
      const uint32_t r = r0 + threadIdx.x;
      uint32_t g = (r < ne) ? __ldg(gslot + r) : SSB_INVALID_SLOT;
      uint16_t o = 0xFFFF, rk = 0;
      uint32_t sl = 0;
      if (g != SSB_INVALID_SLOT) {
        o = (uint16_t)(g / v.capacity);
        sl = g - (uint32_t)o * v.capacity;
        if ((int)o != self_shard) rk = (uint16_t)atomicAdd(&cnt[o], 1u);
      }
      row_owner[threadIdx.x] = o; row_rank[threadIdx.x] = rk; row_slot[threadIdx.x] = sl;
    }
    __syncthreads();
    if (threadIdx.x < v.num_shards && cnt[threadIdx.x]) base[threadIdx.x] = atomicAdd(st.cursor + threadIdx.x, cnt[threadIdx.x]);
    __syncthreads();
    for (int k = warp; k < kSendRows; k += 8) {
      const uint32_t r = r0 + k;
      const int o = row_owner[k];
      if (o == 0xFFFF) continue;
      const float* gr = grads + (size_t)r * D;
      if (o == self_shard) {
        // local row: apply directly (this worker's gradient buffer holds a slot once: single writer, mirror in step)
        if constexpr (applies_rule<Rule>::value)
          rule_apply_row<Rule, true>(v, (uint32_t)o * v.capacity + row_slot[k], lane, h,
                                     [gr](int j) { return *reinterpret_cast<const float4*>(gr + j); }, nullptr);
      } else {
        const uint32_t idx = base[o] + row_rank[k];
        if (idx >= ib.seg_rows) { if (lane == 0) atomicAdd(overflow, 1u); continue; }
        if (lane == 0) ib.slots[o][idx] = row_slot[k];
        __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(ib.rows[o]) + (size_t)idx * D;
        for (int j = lane * 4; j < D; j += 128) *reinterpret_cast<uint2*>(dst + j) = pack_bf16x4(*reinterpret_cast<const float4*>(gr + j));
      }
    }
    __threadfence_system();     // this thread's peer stores are performed before the block reports them
    __syncthreads();
    if (threadIdx.x < v.num_shards && cnt[threadIdx.x] && (int)threadIdx.x != self_shard) {
      const uint32_t o = threadIdx.x;
      const uint32_t lo = min(base[o], ib.seg_rows), hi = min(base[o] + cnt[o], ib.seg_rows);   // rows past the region were dropped
      for (uint32_t c = lo / kChunkRows; c * kChunkRows < hi; ++c) {
        const uint32_t a = max(lo, c * kChunkRows), b = min(hi, (c + 1) * kChunkRows);
        if (b <= a) continue;
        const uint32_t old = atomicAdd(st.done + (size_t)o * ib.max_chunks + c, b - a);
        if (old + (b - a) == kChunkRows) {
          __threadfence_system();   // order the other blocks' rows (seen through the counter) before the flag
          st_release_sys(ib.ready[o] + c, tag);
        }
      }
    }
    __syncthreads();
  }
  // last block out: publish the totals (which also cover each region's last, partial chunk), reset the step state
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    is_last = (atomicAdd(st.finished, 1u) == gridDim.x - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (is_last) {
    __threadfence_system();
    __shared__ uint32_t tot[SSB_MAX_SHARDS];
    if (threadIdx.x < SSB_MAX_SHARDS) {
      uint32_t t = 0;
      if ((int)threadIdx.x < v.num_shards && (int)threadIdx.x != self_shard) {
        t = min(*((volatile uint32_t*)(st.cursor + threadIdx.x)), ib.seg_rows);
        *((volatile uint32_t*)ib.total[threadIdx.x]) = t;
        st_release_sys(ib.total[threadIdx.x] + 1, tag);
      }
      tot[threadIdx.x] = t;
      st.cursor[threadIdx.x] = 0;
    }
    __syncthreads();
    for (int o = 0; o < v.num_shards; ++o) {
      const uint32_t nc = (tot[o] + kChunkRows - 1) / kChunkRows;
      for (uint32_t c = threadIdx.x; c < nc; c += blockDim.x) st.done[(size_t)o * ib.max_chunks + c] = 0;
    }
    if (threadIdx.x == 0) *st.finished = 0;
  }
}

// Owner side: items (chunk c, sender s), chunk-major so that the first chunks of every sender come first. A block
// waits (bounded) until its item is known — the chunk's ready flag carries this step's tag, or the sender's total
// does (then the chunk holds min(C, total - c*C) rows, possibly none) — and applies it to the local shard.
// A wait that runs out sets err[0] and gives up on the item (reported by the host) instead of trapping.
template <class Rule>
__global__ void __launch_bounds__(256)
push_apply_chunks_kernel(const __grid_constant__ SsbTableView v, const uint32_t* __restrict__ slots,
                         const __nv_bfloat16* __restrict__ rows, const uint32_t* __restrict__ ready,
                         const uint32_t* __restrict__ total, uint32_t seg_rows, uint32_t max_chunks, int world, int self,
                         uint32_t tag, const __grid_constant__ rule::Hyper h, uint32_t spin_limit, uint32_t* __restrict__ err) {
  __shared__ uint32_t n_rows_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int D = v.dim;
  const uint32_t items = max_chunks * (uint32_t)world;
  for (uint32_t item = blockIdx.x; item < items; item += gridDim.x) {
    const uint32_t c = item / (uint32_t)world;
    const int s = (int)(item - c * (uint32_t)world);
    if (s == self) continue;
    if (threadIdx.x == 0) {
      uint32_t n = 0, it = 0;
      for (;;) {
        if (ld_acquire_sys(ready + (size_t)s * max_chunks + c) == tag) { n = kChunkRows; break; }
        if (ld_acquire_sys(total + 2 * s + 1) == tag) {
          const uint32_t t = *((const volatile uint32_t*)(total + 2 * s));
          n = t > c * kChunkRows ? min(kChunkRows, t - c * kChunkRows) : 0u;
          break;
        }
        __nanosleep(100);
        if (++it > spin_limit) { atomicExch(err, 1u); break; }
      }
      n_rows_s = n;
    }
    __syncthreads();
    const uint32_t nr = n_rows_s;
    const size_t seg0 = (size_t)s * seg_rows + (size_t)c * kChunkRows;
    for (uint32_t r = warp; r < nr; r += 8) {
      const uint32_t sl = __ldg(slots + seg0 + r);
      if (sl >= v.capacity) continue;
      const __nv_bfloat16* gr = rows + (seg0 + r) * D;
      // several senders' chunks can hold the same row: no mirror writes here, the launcher refreshes it afterwards
      rule_apply_row<Rule, false>(v, sl, lane, h, [gr](int j) {
        const uint2 pk = *reinterpret_cast<const uint2*>(gr + j);
        const float2 a01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&pk.x));
        const float2 a23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&pk.y));
        return make_float4(a01.x, a01.y, a23.x, a23.y);
      }, nullptr);
    }
    __syncthreads();
  }
}

// ---- bulk load (checkpoint resume; new capability, the reference never reads a dump back) ----
__global__ void __launch_bounds__(256)
table_load_kernel(const __grid_constant__ SsbTableView v, const uint32_t* __restrict__ keys,
                  const float* __restrict__ rows, const float* __restrict__ state, int64_t n,
                  uint32_t* __restrict__ stats) {
  const int lane = threadIdx.x & 31;
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  const int D = v.dim;
  for (uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < (uint32_t)n; w += warps) {
    const uint32_t key = __ldg(keys + w);
    if (key >= SSB_LOCKED_KEY) continue;
    Probe p = probe_insert(v, key, true, lane);
    if (p.slot == SSB_INVALID_SLOT) { if (lane == 0 && stats) atomicAdd(stats + 1, 1u); continue; }
    float* row = v.rows[p.owner] + (size_t)p.slot * D;
    for (int j = lane; j < D; j += 32) row[j] = rows[(size_t)w * D + j];
    if (v.shadow[p.owner]) {
      __nv_bfloat16* sh = reinterpret_cast<__nv_bfloat16*>(v.shadow[p.owner]) + (size_t)p.slot * D;
      for (int j = lane; j < D; j += 32) sh[j] = __float2bfloat16_rn(rows[(size_t)w * D + j]);
    }
    if (state && v.state[p.owner]) {
      const int SD = rule_state_stride(v);          // state_words * D floats per entry
      float* st = v.state[p.owner] + (size_t)p.slot * SD;
      for (int j = lane; j < SD; j += 32) st[j] = state[(size_t)w * SD + j];
    }
    if (p.is_new) { publish(v, p, key, lane); if (lane == 0 && stats) atomicAdd(stats + 0, 1u); }
  }
}

// ---- dump: compact the live slots of one shard (K11) ----------------------------------------
// ref: sparsetable.h:49-56,100-104 (text dump key \t value), system/server/init.h:138-149.
__global__ void __launch_bounds__(256)
dump_compact_kernel(const uint32_t* __restrict__ keys, const float* __restrict__ rows, uint32_t capacity,
                    int D, uint32_t* __restrict__ out_keys, float* __restrict__ out_rows,
                    uint32_t* __restrict__ n_live, uint32_t max_out) {
  const int lane = threadIdx.x & 31;
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  const uint32_t nb = capacity / SSB_BUCKET;
  for (uint32_t b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < nb; b += warps) {
    uint32_t k = keys[(size_t)b * SSB_BUCKET + lane];
    unsigned live = __ballot_sync(0xffffffffu, k < SSB_LOCKED_KEY);
    if (!live) continue;
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(n_live, (uint32_t)__popc(live));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (k < SSB_LOCKED_KEY) {
      uint32_t dst = base + __popc(live & ((1u << lane) - 1));
      if (dst < max_out) out_keys[dst] = k;
    }
    if (out_rows) {
      unsigned m = live;
      while (m) {
        int src = __ffs(m) - 1;
        m &= m - 1;
        uint32_t dst = base + __popc(live & ((1u << src) - 1));
        if (dst >= max_out) break;
        const float* r = rows + ((size_t)b * SSB_BUCKET + src) * D;
        float* o = out_rows + (size_t)dst * D;
        for (int j = lane; j < D; j += 32) o[j] = r[j];
      }
    }
  }
}

// Chunked form for checkpoints: compacts the live slots of the bucket range [bucket_lo, bucket_hi) of one shard,
// copying the row AND (when the shard has one) the optimizer-state row of a slot to the same output index, so a
// shard of any size streams through bounded host / device buffers with rows and state aligned.
__global__ void __launch_bounds__(256)
dump_compact_range_kernel(const uint32_t* __restrict__ keys, const float* __restrict__ rows, const float* __restrict__ state,
                          uint32_t bucket_lo, uint32_t bucket_hi, int D, int SD, uint32_t* __restrict__ out_keys,
                          float* __restrict__ out_rows, float* __restrict__ out_state, uint32_t* __restrict__ n_live,
                          uint32_t max_out) {
  const int lane = threadIdx.x & 31;
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  for (uint32_t b = bucket_lo + ((blockIdx.x * blockDim.x + threadIdx.x) >> 5); b < bucket_hi; b += warps) {
    const uint32_t k = keys[(size_t)b * SSB_BUCKET + lane];
    const unsigned live = __ballot_sync(0xffffffffu, k < SSB_LOCKED_KEY);
    if (!live) continue;
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(n_live, (uint32_t)__popc(live));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (k < SSB_LOCKED_KEY) {
      const uint32_t dst = base + __popc(live & ((1u << lane) - 1));
      if (dst < max_out) out_keys[dst] = k;
    }
    unsigned m = live;
    while (m) {
      const int src = __ffs(m) - 1;
      m &= m - 1;
      const uint32_t dst = base + __popc(live & ((1u << src) - 1));
      if (dst >= max_out) break;
      const size_t slot = (size_t)b * SSB_BUCKET + src;
      for (int j = lane; j < D; j += 32) out_rows[(size_t)dst * D + j] = rows[slot * D + j];
      if (out_state)
        for (int j = lane; j < SD; j += 32) out_state[(size_t)dst * SD + j] = state[slot * SD + j];
    }
  }
}

inline int row_grid(int64_t n) {
  // warp per row, 8 warps per block; cap at a few waves of 148 SMs x 8 resident blocks
  return grid_for(n, 8, kNumSMs * 8);
}

}  // namespace ssb

extern "C" {

static int check_view(const SsbTableView* v) {
  if (!v || v->num_shards < 1 || v->num_shards > SSB_MAX_SHARDS || v->dim < 1 ||
      v->capacity == 0 || v->capacity % SSB_BUCKET != 0 || v->frag_num == 0 || !v->frag_map)
    return (int)cudaErrorInvalidValue;
  if ((uint64_t)v->capacity * (uint64_t)v->num_shards >= 0xFFFFFFFFull) return (int)cudaErrorInvalidValue;
  return 0;
}

int ssb_pull_cached(const SsbTableView* view, const uint32_t* keys, int64_t n, const uint32_t* n_dev,
                    uint32_t* gslot, float* rows_out, void* rows_bf16, float* grads_zero, int insert,
                    uint32_t* stats, uint64_t* slot_cache, int32_t cache_bits, void* stream) {
  if (int e = check_view(view)) return e;
  if (slot_cache && ((cache_bits < 0 ? -cache_bits : cache_bits) < 4 || (cache_bits < 0 ? -cache_bits : cache_bits) > 31))
    return (int)cudaErrorInvalidValue;
  if (n <= 0) return 0;
  ssb::PullArgs a{0u, gslot, rows_out, (__nv_bfloat16*)rows_bf16, grads_zero, insert, stats, (unsigned long long*)slot_cache, cache_bits};
  // Which pull kernel: the bulk-copy form (rows travel by cp.async.bulk into shared memory, no registers held while
  // they are in flight) is the default for bf16-mirror pulls — the training step's pull. Alone it is no faster than
  // the register form (168 vs 161 us at 2 GPUs), but next to the tensor-core kernel of the previous step, whose
  // reductions keep the load/store units busy, the whole pipelined step is: 0.474 vs 0.507 ms on one GPU, 0.577 vs
  // 0.590 on two. SSB200_PULL_BULK=0 / 1 forces one form for every pull.
  static const int bulk_env = [] { const char* e = getenv("SSB200_PULL_BULK"); return e ? atoi(e) : -1; }();
  const bool from_shadow = !rows_out && rows_bf16 && view->shadow[0] != nullptr;
  const int use_bulk = bulk_env >= 0 ? bulk_env : (from_shadow ? 1 : 0);
  const uint32_t row_bytes = (uint32_t)view->dim * (from_shadow ? 2u : 4u);
  if (use_bulk && (rows_out || rows_bf16) && (view->dim & 3) == 0 && row_bytes % 16 == 0 && row_bytes <= 512) {
    const size_t smem = (size_t)ssb::kBulkWarps * 32 * row_bytes + 64;
    static bool attr_set_dev[64] = {false};   // the attribute is per device
    int dev_id = 0;
    if (cudaGetDevice(&dev_id) != cudaSuccess || dev_id < 0 || dev_id >= 64) return (int)cudaErrorInvalidDevice;
    bool& attr_set = attr_set_dev[dev_id];
    if (!attr_set) {
      cudaError_t e1 = cudaFuncSetAttribute(ssb::pull_bulk_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, ssb::kBulkWarps * 32 * 512 + 64);
      cudaError_t e2 = cudaFuncSetAttribute(ssb::pull_bulk_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, ssb::kBulkWarps * 32 * 512 + 64);
      if (e1 != cudaSuccess) return (int)e1;
      if (e2 != cudaSuccess) return (int)e2;
      attr_set = true;
    }
    const int blocks = ssb::grid_for(n, ssb::kBulkWarps * 32, ssb::kNumSMs * 3);
    if (slot_cache) ssb::pull_bulk_kernel<true><<<blocks, ssb::kBulkWarps * 32, smem, (cudaStream_t)stream>>>(*view, keys, n, n_dev, a);
    else ssb::pull_bulk_kernel<false><<<blocks, ssb::kBulkWarps * 32, smem, (cudaStream_t)stream>>>(*view, keys, n, n_dev, a);
  } else if (slot_cache) {
    ssb::pull_kernel<true><<<ssb::grid_for(n, 32, ssb::kNumSMs * 8), 256, 0, (cudaStream_t)stream>>>(*view, keys, n, n_dev, a);
  } else {
    ssb::pull_kernel<false><<<ssb::row_grid(n), 256, 0, (cudaStream_t)stream>>>(*view, keys, n, n_dev, a);
  }
  SSB_LAUNCH_CHECK();
  return 0;
}

int ssb_pull(const SsbTableView* view, const uint32_t* keys, int64_t n, const uint32_t* n_dev,
             uint32_t* gslot, float* rows_out, void* rows_bf16, float* grads_zero, int insert,
             uint32_t* stats, void* stream) {
  return ssb_pull_cached(view, keys, n, n_dev, gslot, rows_out, rows_bf16, grads_zero, insert, stats, nullptr, 0, stream);
}

int ssb_gather_slots(const SsbTableView* view, const uint32_t* gslot, int64_t n, const uint32_t* n_dev,
                     float* rows_out, void* stream) {
  if (int e = check_view(view)) return e;
  if (n <= 0) return 0;
  ssb::gather_slots_kernel<<<ssb::row_grid(n), 256, 0, (cudaStream_t)stream>>>(*view, gslot, n, n_dev, rows_out);
  SSB_LAUNCH_CHECK();
  return 0;
}

static int map_rule_err(int e) { return e < 0 ? (int)cudaErrorInvalidValue : e; }

int ssb_push_sgd(const SsbTableView* view, const uint32_t* gslot, float* grads, int64_t n,
                 const uint32_t* n_dev, float lr, int zero_grads, void* stream) {
  if (int e = check_view(view)) return e;
  const ssb::rule::Hyper h{lr, 0.f, 0.f, 0.f, 0.f, 0u};
  return map_rule_err(ssb::launch_rule_push<ssb::rule::Sgd>(view, gslot, grads, n, n_dev, &h, zero_grads, -1, stream, ssb::row_grid(n)));
}

int ssb_push_adagrad(const SsbTableView* view, const uint32_t* gslot, float* grads, int64_t n,
                     const uint32_t* n_dev, float lr, float eps, int zero_grads, void* stream) {
  if (int e = check_view(view)) return e;
  const ssb::rule::Hyper h{lr, eps, 0.f, 0.f, 0.f, 0u};
  return map_rule_err(ssb::launch_rule_push<ssb::rule::AdaGrad>(view, gslot, grads, n, n_dev, &h, zero_grads, -1, stream, ssb::row_grid(n)));
}

int ssb_push_send2(const SsbTableView* view, const SsbInboxView* inbox, const uint32_t* gslot, const float* grads,
                   int64_t n, const uint32_t* n_dev, int32_t self_shard, float lr, float eps, int32_t adagrad,
                   int32_t bf16_rows, uint32_t* overflow, uint32_t* cursor, void* stream);

int ssb_push_send(const SsbTableView* view, const SsbInboxView* inbox, const uint32_t* gslot, const float* grads,
                  int64_t n, const uint32_t* n_dev, int32_t self_shard, float lr, float eps, int32_t adagrad,
                  int32_t bf16_rows, uint32_t* overflow, void* stream) {
  return ssb_push_send2(view, inbox, gslot, grads, n, n_dev, self_shard, lr, eps, adagrad, bf16_rows, overflow, nullptr, stream);
}

int ssb_push_send2(const SsbTableView* view, const SsbInboxView* inbox, const uint32_t* gslot, const float* grads,
                   int64_t n, const uint32_t* n_dev, int32_t self_shard, float lr, float eps, int32_t adagrad,
                   int32_t bf16_rows, uint32_t* overflow, uint32_t* cursor, void* stream) {
  if (int e = check_view(view)) return e;
  if (!inbox || inbox->capacity == 0 || (view->dim & 3) != 0) return (int)cudaErrorInvalidValue;
  if (n <= 0) return 0;
  int blocks = ssb::grid_for(n, ssb::kSendRows, ssb::kNumSMs * 8);
  // `adagrad` is a rule id: 0 SGD, 1 AdaGrad, -1 ship only (the caller applies its own rows with any rule of
  // include/ssb200_rule.cuh: ssb_rule_<name>_push(..., only_shard = self_shard))
  const ssb::rule::Hyper h{lr, eps, 0.f, 0.f, 0.f, 0u};
  if (adagrad == 1 && ssb::rule_check(view, 1)) return (int)cudaErrorInvalidValue;
#define SSB_SEND(R, B) ssb::push_send_kernel<R, B><<<blocks, 256, 0, (cudaStream_t)stream>>>(*view, *inbox, gslot, grads, n, n_dev, self_shard, h, overflow, cursor)
  if (adagrad == 1) { if (bf16_rows) SSB_SEND(ssb::rule::AdaGrad, true); else SSB_SEND(ssb::rule::AdaGrad, false); }
  else if (adagrad == 0) { if (bf16_rows) SSB_SEND(ssb::rule::Sgd, true); else SSB_SEND(ssb::rule::Sgd, false); }
  else if (adagrad == -1) {
    const int sb = ssb::grid_for(n, 256, ssb::kNumSMs * 8);        // 8 warps x 32 rows per block trip
    if (bf16_rows) ssb::push_ship_kernel<true><<<sb, 256, 0, (cudaStream_t)stream>>>(*view, *inbox, gslot, grads, n, n_dev, self_shard, overflow, cursor);
    else ssb::push_ship_kernel<false><<<sb, 256, 0, (cudaStream_t)stream>>>(*view, *inbox, gslot, grads, n, n_dev, self_shard, overflow, cursor);
  }
  else return (int)cudaErrorInvalidValue;
#undef SSB_SEND
  SSB_LAUNCH_CHECK();
  return 0;
}

int ssb_push_send_chunks(const SsbTableView* view, const SsbChunkInboxView* inbox, const uint32_t* gslot, const float* grads,
                         int64_t n, const uint32_t* n_dev, int32_t self_shard, float lr, float eps, int32_t adagrad, uint32_t tag,
                         uint32_t* overflow, uint32_t* cursor, uint32_t* done, uint32_t* finished, void* stream) {
  if (int e = check_view(view)) return e;
  if (!inbox || inbox->seg_rows == 0 || inbox->seg_rows % ssb::kChunkRows != 0 ||
      inbox->max_chunks != inbox->seg_rows / ssb::kChunkRows || (view->dim & 3) != 0 || !overflow || !cursor || !done ||
      !finished || tag == 0)
    return (int)cudaErrorInvalidValue;
  if (n < 0) n = 0;
  // the kernel always runs (also with n == 0): its last block publishes this step's totals to every owner
  int blocks = ssb::grid_for(n > 0 ? n : 1, ssb::kSendRows, ssb::kNumSMs * 8);
  ssb::SendChunkState st{cursor, done, finished};
  const ssb::rule::Hyper h{lr, eps, 0.f, 0.f, 0.f, 0u};
  if (adagrad == 1 && ssb::rule_check(view, 1)) return (int)cudaErrorInvalidValue;
  if (adagrad == 1)
    ssb::push_send_chunks_kernel<ssb::rule::AdaGrad><<<blocks, 256, 0, (cudaStream_t)stream>>>(*view, *inbox, gslot, grads, n, n_dev, self_shard, h, tag, overflow, st);
  else if (adagrad == 0)
    ssb::push_send_chunks_kernel<ssb::rule::Sgd><<<blocks, 256, 0, (cudaStream_t)stream>>>(*view, *inbox, gslot, grads, n, n_dev, self_shard, h, tag, overflow, st);
  else
    ssb::push_send_chunks_kernel<ssb::ShipOnly><<<blocks, 256, 0, (cudaStream_t)stream>>>(*view, *inbox, gslot, grads, n, n_dev, self_shard, h, tag, overflow, st);
  SSB_LAUNCH_CHECK();
  return 0;
}

int ssb_push_apply_chunks(const SsbTableView* local_view, const uint32_t* slots, const void* rows_bf16, const uint32_t* ready,
                          const uint32_t* total, uint32_t seg_rows, int32_t world, int32_t self, uint32_t tag, float lr, float eps,
                          int32_t adagrad, uint32_t spin_limit, uint32_t* err, void* stream) {
  if (int e = check_view(local_view)) return e;
  if (local_view->num_shards != 1 || (local_view->dim & 3) != 0 || seg_rows == 0 || seg_rows % ssb::kChunkRows != 0 || world < 1 ||
      world > SSB_MAX_SHARDS || self < 0 || self >= world || !slots || !rows_bf16 || !ready || !total || !err || tag == 0)
    return (int)cudaErrorInvalidValue;
  if (adagrad && !local_view->state[0]) return (int)cudaErrorInvalidValue;
  if (world == 1) return 0;
  if (spin_limit == 0) spin_limit = 1u << 26;          // x 100 ns: several seconds
  const uint32_t max_chunks = seg_rows / ssb::kChunkRows;
  // one block per SM at most: the waiting blocks of this kernel must never fill the SMs the send kernel needs
  const int blocks = (int)std::min<uint64_t>((uint64_t)max_chunks * (uint64_t)world, (uint64_t)ssb::kNumSMs);
  const ssb::rule::Hyper h{lr, eps, 0.f, 0.f, 0.f, 0u};
  if (adagrad)
    ssb::push_apply_chunks_kernel<ssb::rule::AdaGrad><<<blocks, 256, 0, (cudaStream_t)stream>>>(*local_view, slots, (const __nv_bfloat16*)rows_bf16, ready, total, seg_rows, max_chunks, world, self, tag, h, spin_limit, err);
  else
    ssb::push_apply_chunks_kernel<ssb::rule::Sgd><<<blocks, 256, 0, (cudaStream_t)stream>>>(*local_view, slots, (const __nv_bfloat16*)rows_bf16, ready, total, seg_rows, max_chunks, world, self, tag, h, spin_limit, err);
  SSB_LAUNCH_CHECK();
  if (local_view->shadow[0]) {     // rows listed by several senders had several writers: set the mirror from the final rows
    dim3 grid((unsigned)std::max(1, ssb::kNumSMs * 4 / world), (unsigned)world);
    ssb::rule_mirror_refresh_kernel<0><<<grid, 256, 0, (cudaStream_t)stream>>>(*local_view, slots, seg_rows, total, seg_rows, 2);
    SSB_LAUNCH_CHECK();
  }
  return 0;
}

int ssb_push_bf16_seg(const SsbTableView* view, const uint32_t* gslot, const void* grads_bf16, int64_t seg_rows,
                      int32_t num_segs, const uint32_t* counts, float lr, float eps, int32_t adagrad, void* stream) {
  if (int e = check_view(view)) return e;
  if ((view->dim & 3) != 0 || num_segs < 1 || num_segs > SSB_MAX_SHARDS || (num_segs > 1 && !counts)) return (int)cudaErrorInvalidValue;
  if (seg_rows <= 0) return 0;
  const int gx = ssb::grid_for(seg_rows, 8, ssb::kNumSMs * 8 / (num_segs > 1 ? 2 : 1));
  const ssb::rule::Hyper h{lr, eps, 0.f, 0.f, 0.f, 0u};
  if (adagrad)
    return map_rule_err(ssb::launch_rule_push_bf16_seg<ssb::rule::AdaGrad>(view, gslot, grads_bf16, seg_rows, counts, num_segs, seg_rows, &h, stream, gx));
  return map_rule_err(ssb::launch_rule_push_bf16_seg<ssb::rule::Sgd>(view, gslot, grads_bf16, seg_rows, counts, num_segs, seg_rows, &h, stream, gx));
}

int ssb_push_bf16(const SsbTableView* view, const uint32_t* gslot, const void* grads_bf16, int64_t n, const uint32_t* n_dev,
                  float lr, float eps, int32_t adagrad, void* stream) {
  return ssb_push_bf16_seg(view, gslot, grads_bf16, n, 1, n_dev, lr, eps, adagrad, stream);
}

int ssb_inbox_publish(const SsbInboxView* inbox, int32_t num_shards, int32_t self, uint32_t* cursor, uint32_t value, void* stream) {
  if (!inbox || !cursor || num_shards < 1 || num_shards > SSB_MAX_SHARDS) return (int)cudaErrorInvalidValue;
  ssb::inbox_publish_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*inbox, num_shards, self, cursor, value);
  SSB_LAUNCH_CHECK();
  return 0;
}

int ssb_flags_signal(const SsbInboxView* inbox, int32_t num_shards, int32_t self, uint32_t value, void* stream) {
  if (!inbox || num_shards < 1 || num_shards > SSB_MAX_SHARDS) return (int)cudaErrorInvalidValue;
  ssb::flags_signal_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(*inbox, num_shards, self, value);
  SSB_LAUNCH_CHECK();
  return 0;
}

int ssb_flags_wait2(const uint32_t* my_flags, int32_t num_shards, uint32_t value, uint32_t spin_limit, uint32_t* err, void* stream) {
  if (num_shards < 1 || num_shards > SSB_MAX_SHARDS) return (int)cudaErrorInvalidValue;
  if (spin_limit == 0) {
    // default ~27 s (2^27 x 0.2 us); SSB200_WAIT_LIMIT overrides (iterations)
    static const uint32_t env_limit = [] { const char* e = getenv("SSB200_WAIT_LIMIT"); return e ? (uint32_t)strtoul(e, nullptr, 10) : 0u; }();
    spin_limit = env_limit ? env_limit : (1u << 27);
  }
  ssb::flags_wait_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(my_flags, num_shards, value, spin_limit, err);
  SSB_LAUNCH_CHECK();
  return 0;
}

int ssb_flags_wait(const uint32_t* my_flags, int32_t num_shards, uint32_t value, void* stream) {
  return ssb_flags_wait2(my_flags, num_shards, value, 0, nullptr, stream);
}

int ssb_table_load(const SsbTableView* view, const uint32_t* keys, const float* rows, const float* state,
                   int64_t n, uint32_t* stats, void* stream) {
  if (int e = check_view(view)) return e;
  if (n <= 0) return 0;
  ssb::table_load_kernel<<<ssb::row_grid(n), 256, 0, (cudaStream_t)stream>>>(*view, keys, rows, state, n, stats);
  SSB_LAUNCH_CHECK();
  return 0;
}

int ssb_dump_compact_range(const uint32_t* keys, const float* rows, const float* state, uint32_t capacity, int32_t dim,
                           uint32_t slot_lo, uint32_t slot_hi, uint32_t* out_keys, float* out_rows, float* out_state,
                           uint32_t* n_live, uint32_t max_out, int32_t state_dim, void* stream) {
  if (state_dim <= 0) state_dim = dim;       // floats of optimizer state per entry (state_words * dim)
  if (capacity == 0 || capacity % SSB_BUCKET != 0 || slot_lo % SSB_BUCKET != 0 || slot_hi % SSB_BUCKET != 0 ||
      slot_lo > slot_hi || slot_hi > capacity || dim < 1 || !keys || !rows || !out_keys || !out_rows || !n_live ||
      (out_state && !state))
    return (int)cudaErrorInvalidValue;
  cudaError_t e = cudaMemsetAsync(n_live, 0, sizeof(uint32_t), (cudaStream_t)stream);
  if (e != cudaSuccess) return (int)e;
  if (slot_lo == slot_hi) return 0;
  const uint32_t b_lo = slot_lo / SSB_BUCKET, b_hi = slot_hi / SSB_BUCKET;
  ssb::dump_compact_range_kernel<<<ssb::row_grid(b_hi - b_lo), 256, 0, (cudaStream_t)stream>>>(
      keys, rows, state, b_lo, b_hi, dim, state_dim, out_keys, out_rows, out_state, n_live, max_out);
  SSB_LAUNCH_CHECK();
  return 0;
}

int ssb_dump_compact(const uint32_t* keys, const float* rows, uint32_t capacity, int32_t dim,
                     uint32_t* out_keys, float* out_rows, uint32_t* n_live, uint32_t max_out,
                     void* stream) {
  if (capacity == 0 || capacity % SSB_BUCKET != 0) return (int)cudaErrorInvalidValue;
  cudaError_t e = cudaMemsetAsync(n_live, 0, sizeof(uint32_t), (cudaStream_t)stream);
  if (e != cudaSuccess) return (int)e;
  ssb::dump_compact_kernel<<<ssb::row_grid(capacity / SSB_BUCKET), 256, 0, (cudaStream_t)stream>>>(
      keys, rows, capacity, dim, out_keys, out_rows, n_live, max_out);
  SSB_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
