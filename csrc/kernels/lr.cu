// lr.cu — K6: sparse logistic regression forward/backward on the worker's pulled cache.
// The reference's logistic_regression app is absent from the snapshot; the
// loop shape is the BaseAlgorithm contract (core/framework/SwiftWorker.h:19-57): pull feature
// weights, per-example dot + sigmoid + log-loss, accumulate per-feature gradients, push
// (AdaGrad applied on the owner shard, parameter/sparse_access_method.h:33-46).
// One warp per example: lanes stride over the example's features (Criteo shape: 39).
#include <algorithm>
#include "common.cuh"

namespace ssb {

__global__ void __launch_bounds__(256)
lr_fwd_bwd_kernel(const float* __restrict__ params, float* __restrict__ grads,
                  const uint32_t* __restrict__ inv, const float* __restrict__ x,
                  const float* __restrict__ labels, int64_t n_ex, int n_feat, float* __restrict__ loss_out) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  float loss_acc = 0.f;
  for (int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; e < n_ex; e += warps) {
    const uint32_t* ie = inv + e * n_feat;
    const float* xe = x ? x + e * n_feat : nullptr;
    float z = 0.f;
    for (int f = lane; f < n_feat; f += 32) {
      uint32_t r = ie[f];
      if (r != SSB_INVALID_SLOT) z += __ldg(params + r) * (xe ? xe[f] : 1.0f);
    }
    z = warp_sum(z);
    const float y = labels[e];
    const float pr = fast_sigmoid(z);
    // log-loss: y*softplus(-z) + (1-y)*softplus(z)
    loss_acc += y * neg_log_sigmoid(z) + (1.f - y) * neg_log_sigmoid(-z);
    const float g = pr - y;
    for (int f = lane; f < n_feat; f += 32) {
      uint32_t r = ie[f];
      if (r != SSB_INVALID_SLOT) atomicAdd(grads + r, g * (xe ? xe[f] : 1.0f));
    }
  }
  __shared__ float sloss[8];
  if (lane == 0) sloss[threadIdx.x >> 5] = loss_acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += sloss[w];
    if (t != 0.f) atomicAdd(loss_out, t);
    if (blockIdx.x == 0) atomicAdd(loss_out + 1, (float)n_ex);
  }
}

}  // namespace ssb

extern "C" int ssb_lr_fwd_bwd(const float* params, float* grads, const uint32_t* inv, const float* x,
                              const float* labels, int64_t n_ex, int32_t n_feat, float* loss_out, void* stream) {
  if (n_ex <= 0) return 0;
  if (n_feat < 1) return (int)cudaErrorInvalidValue;
  ssb::lr_fwd_bwd_kernel<<<ssb::grid_for(n_ex, 8, ssb::kNumSMs * 8), 256, 0, (cudaStream_t)stream>>>(
      params, grads, inv, x, labels, n_ex, n_feat, loss_out);
  SSB_LAUNCH_CHECK();
  return 0;
}

// ================================================================================================================
// K6 fused: one sparse-LR training step as ONE persistent cooperative kernel.
//
// ref: the loop shape is the reference worker's (core/framework/SwiftWorker.h:19-57: collect the minibatch's keys ->
// pull -> compute -> push; the server applies the update rule per key once per minibatch, parameter/
// sparse_access_method.h:33-46). The reference does this with a request/response round trip per server; the unfused
// path here (dedup -> pull -> lr_fwd_bwd -> push) with four launches. This kernel runs four phases separated by
// grid-wide barriers and dedups through a worker-side ENTRY per feature id:
//
//   A  resolve   lane = example, loop over its features. One 16-byte entry per feature id {tag, idx, slot}: `slot`
//                is the feature's global table slot (found once, by the ordinary probe / insert protocol); `tag` is
//                the last step that touched the feature and `idx` its position in that step's compact arrays. A
//                lane that finds a stale tag allocates an index and installs (step, idx) with one 64-bit CAS; the
//                loser of a race takes the winner's idx (its own allocation stays a hole). The idx is stashed per
//                instance ([feature][example], coalesced). One random 16-byte read per instance, mostly L2 hits.
//   B  gather    one lane per allocated idx: read the weight from the owner shard (own HBM or a peer's over NVLink —
//                once per unique key and rank) into wc[idx].
//   C  compute   lane = example: dot over wc[stash], sigmoid, log-loss, g = p - y; g * x is added to a block-local
//                shared-memory table of 64-bit fixed-point sums (two native 32-bit smem atomics; a feature the whole
//                warp shares — the numeric fields — is warp-reduced first), spilling to gc[idx]; the block flushes
//                its table at the end of the phase. The compact wc / gc arrays (4 bytes per touched key) stay in L2.
//   D  push      one lane per allocated idx: take gc[idx] (and zero it), apply the update rule (Rule::apply1 of
//                include/ssb200_rule.cuh: SGD, AdaGrad, FTRL) to the owner's weight and state.
//
// Tried and dropped: staging a chunk's feature values through shared memory with coalesced loads (40 KB per block at
// 39 features). The strided reads it removes were L1 hits; the shared memory it takes away from L1 made the resolve
// phase slower (201 -> 253 us) and the whole step 0.63 -> 0.73 ms.
// Worker-side memory, allocated once and sized for 180 GB of HBM: 16 bytes per feature id (4 GB at 2^28). Without a
// dense feature space (entry_count == capacity * shards) the entry index is the global slot and every instance
// probes the table (slower, general). Chunks of 32 examples are handed out dynamically, so warps finish together.
// The grid is launched cooperatively (all blocks co-resident): the barriers cannot deadlock, and they are bounded.
// ================================================================================================================
#include "table_probe.cuh"

namespace ssb {
namespace lrf {

struct Args {
  SsbTableView v;
  const uint32_t* keys;      // [n_ex][n_feat] (or [n_feat][n_ex] when feature_major); SSB_EMPTY_KEY = no feature in this position
  const float* x;            // same layout as keys, or nullptr (all ones)
  int feature_major;         // 1: keys / x are stored feature by feature, so that the 32 examples of a warp are contiguous
  const float* labels;       // [n_ex] in {0, 1}
  int64_t n_ex;
  int n_feat;
  uint4* ent;                // [entry_count] {tag, idx, slot, -}; zero tag, slot = SSB_INVALID_SLOT initially
  uint64_t dense_keys;       // > 0: entries are indexed by key (< dense_keys); 0: by global slot
  uint32_t* stash;           // [n_feat][n_ex] idx of every instance
  uint32_t* touched;         // [max_touched] idx -> global slot (SSB_INVALID_SLOT = hole)
  float* wc;                 // [max_touched] weights of this step
  float* gc;                 // [max_touched] gradient sums of this step (zero between steps)
  uint32_t max_touched;
  uint32_t* counters;        // [0] allocated idx, [1] grid barrier, [2] idx overflow, [3] keys without a slot,
                             // [4] / [5] chunk tickets of phases A / C, [8..12] globaltimer stamps (block 0)
  uint32_t step;             // != 0, different from every earlier step's
  float* loss_out;           // [2]: sum of log-loss, examples
  rule::Hyper h;
};

constexpr int kThreads = 256;
constexpr int kAggBits = 11, kAggSize = 1 << kAggBits;      // block-local gradient table: 2048 x (4 + 4 + 4) bytes (1024: compute phase 248 vs 217 us; 4096: no further gain)
constexpr float kFixScale = 68719476736.0f;                  // 2^36: |block sum| < 2^27 fits int64 with room
constexpr uint32_t kAggEmpty = 0xFFFFFFFFu;

__device__ __forceinline__ uint32_t ld_acquire_gpu_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_entry(const uint4* p) {          // strong load: other warps install tags concurrently
  uint4 r;
  asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ float ld_strong_f32(const float* p) {
  float r;
  asm volatile("ld.relaxed.gpu.global.f32 %0, [%1];" : "=f"(r) : "l"(p) : "memory");
  return r;
}

// All blocks are co-resident (cooperative launch). `phase` counts the barriers passed so far.
__device__ __forceinline__ void grid_barrier(uint32_t* bar, uint32_t& phase) {
  __syncthreads();
  if (threadIdx.x == 0) {
    ++phase;
    const uint32_t target = phase * gridDim.x;
    __threadfence();
    atomicAdd(bar, 1u);
    uint32_t spins = 0;
    while (ld_acquire_gpu_u32(bar) < target) {
      __nanosleep(40);
      if (++spins > (1u << 26)) __trap();     // seconds: a block that never arrives is a bug, not a wait
    }
    __threadfence();
  }
  __syncthreads();
}

__device__ __forceinline__ int64_t next_chunk(uint32_t* ticket, int lane) {
  uint32_t c = 0;
  if (lane == 0) c = atomicAdd(ticket, 1u);
  return (int64_t)__shfl_sync(0xffffffffu, c, 0);
}

// The slow path of phase A for the lanes in `need`: find (or insert) the key's table slot, one key at a time with
// the warp-cooperative probe. Returns this lane's global slot (SSB_INVALID_SLOT if the table is full).
__device__ __forceinline__ uint32_t probe_lanes(const Args& a, uint32_t key, unsigned need, int lane) {
  uint32_t gs = SSB_INVALID_SLOT;
  while (need) {
    const int src = __ffs(need) - 1;
    need &= need - 1;
    const uint32_t k = __shfl_sync(0xffffffffu, key, src);
    Probe p = probe_insert(a.v, k, true, lane);
    if (p.slot != SSB_INVALID_SLOT && p.is_new) {
      if (lane == 0) {
        const float w0 = init_of(a.v, k, 0);                      // dim 1; the state is zero since allocation
        a.v.rows[p.owner][p.slot] = w0;
        if (a.v.shadow[p.owner]) reinterpret_cast<__nv_bfloat16*>(a.v.shadow[p.owner])[p.slot] = __float2bfloat16_rn(w0);
      }
      publish(a.v, p, k, lane);
    }
    const uint32_t g = (p.slot == SSB_INVALID_SLOT) ? SSB_INVALID_SLOT : (uint32_t)p.owner * a.v.capacity + p.slot;
    if (lane == src) {
      gs = g;
      if (g == SSB_INVALID_SLOT) atomicAdd(a.counters + 3, 1u);
    }
  }
  return gs;
}

constexpr int kGroup = 4;       // features handled together in phase A: kGroup independent entry reads in flight per lane
constexpr int kReserve = 64;    // compact indices a warp reserves at a time (<= 64: retired with two stores per lane)

template <class Rule>
__global__ void __launch_bounds__(kThreads)
lr_fused_step_kernel(const __grid_constant__ Args a) {
  __shared__ uint32_t agg_key[kAggSize];
  __shared__ uint32_t agg_lo[kAggSize], agg_hi[kAggSize];     // 64-bit fixed-point sums as two words: native 32-bit ATOMS.ADD
                                                              // (a 64-bit or float shared atomic compiles to a CAS spin loop)
  __shared__ float sloss[kThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int F = a.n_feat;
  const int64_t n_ex = a.n_ex;
  const int64_t chunks = (n_ex + 31) / 32;
  uint32_t phase = 0;
  auto stamp = [&](int i) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      uint32_t t;
      asm volatile("mov.u32 %0, %%globaltimer_lo;" : "=r"(t));
      a.counters[8 + i] = t;
    }
  };
  stamp(0);
  for (int i = threadIdx.x; i < kAggSize; i += kThreads) { agg_key[i] = kAggEmpty; agg_lo[i] = 0u; agg_hi[i] = 0u; }

  // ---- A: resolve + compact index -----------------------------------------------------------------------------
  // Compact indices come from a warp-private reservation (kReserve at a time from the one global counter: one
  // global atomic per kReserve new keys instead of one per group); what a warp leaves unused becomes holes.
  uint32_t res_base = 0, res_left = 0;          // warp-uniform
  const int n_groups = (F + kGroup - 1) / kGroup;
  for (int64_t c = next_chunk(a.counters + 4, lane); c < chunks; c = next_chunk(a.counters + 4, lane)) {
    const int64_t e = c * 32 + lane;
    const bool live = e < n_ex;
    // element (e, f): example-major [e][f] (lanes 4 * n_feat bytes apart: a sector per key) or feature-major [f][e]
    // (32 consecutive keys per load)
    const size_t ks_e = a.feature_major ? 1 : (size_t)F, ks_f = a.feature_major ? (size_t)n_ex : 1;
    const uint32_t* ke = a.keys + (live ? (size_t)e : 0) * ks_e;
    // warps start at different feature groups, so the features every example has (the numeric fields) are not hit
    // by all warps at the same moment: the first warp installs the tag, the others then see it with a plain load
    const int rot = (int)(c % n_groups);
#pragma unroll 1
    for (int gi = 0; gi < n_groups; ++gi) {
      const int f0 = ((gi + rot) % n_groups) * kGroup;
      uint32_t key[kGroup], ei[kGroup];
      uint4 E[kGroup];
      bool valid[kGroup];
#pragma unroll
      for (int g = 0; g < kGroup; ++g) {
        key[g] = (live && f0 + g < F) ? __ldg(ke + (size_t)(f0 + g) * ks_f) : SSB_EMPTY_KEY;
        valid[g] = key[g] < SSB_LOCKED_KEY;
      }
      if (a.dense_keys) {
#pragma unroll
        for (int g = 0; g < kGroup; ++g) {
          if (valid[g] && key[g] >= a.dense_keys) { valid[g] = false; atomicAdd(a.counters + 3, 1u); }   // outside the declared space
          ei[g] = key[g];
        }
      } else {
#pragma unroll
        for (int g = 0; g < kGroup; ++g) {
          ei[g] = probe_lanes(a, key[g], __ballot_sync(0xffffffffu, valid[g]), lane);
          valid[g] = valid[g] && ei[g] != SSB_INVALID_SLOT;
        }
      }
#pragma unroll
      for (int g = 0; g < kGroup; ++g) E[g] = valid[g] ? ld_entry(a.ent + ei[g]) : make_uint4(a.step, SSB_INVALID_SLOT, 0u, 0u);
#pragma unroll
      for (int g = 0; g < kGroup; ++g) {
        uint32_t gs = a.dense_keys ? E[g].z : ei[g];
        uint32_t idx = E[g].y;
        if (a.dense_keys) {            // first sight of this feature id: find / insert its table slot, remember it
          const unsigned need = __ballot_sync(0xffffffffu, valid[g] && gs == SSB_INVALID_SLOT);
          if (need) {
            const uint32_t got = probe_lanes(a, key[g], need, lane);
            if (valid[g] && gs == SSB_INVALID_SLOT) {
              gs = got;
              if (gs != SSB_INVALID_SLOT) reinterpret_cast<uint32_t*>(a.ent + ei[g])[2] = gs; else valid[g] = false;
            }
          }
        }
        const bool stale = valid[g] && E[g].x != a.step;
        const unsigned sm = __ballot_sync(0xffffffffu, stale);
        if (sm) {
          // one lane per distinct entry among the stale lanes does the install; its peers take its result
          unsigned peers = 0;
          if (stale) peers = __match_any_sync(sm, ei[g]);
          const bool leader = stale && (__ffs(peers) - 1 == lane);
          const unsigned lm = __ballot_sync(0xffffffffu, leader);
          const uint32_t cnt = (uint32_t)__popc(lm);
          if (cnt > res_left) {                      // warp-uniform: retire the rest of the reservation, take a new one
            if ((uint32_t)lane < res_left && res_base + lane < a.max_touched) a.touched[res_base + lane] = SSB_INVALID_SLOT;
            if (res_left > 32 && (uint32_t)lane + 32 < res_left && res_base + lane + 32 < a.max_touched)
              a.touched[res_base + lane + 32] = SSB_INVALID_SLOT;
            uint32_t nb = 0;
            if (lane == 0) nb = atomicAdd(a.counters + 0, (uint32_t)kReserve);
            res_base = __shfl_sync(0xffffffffu, nb, 0);
            res_left = kReserve;
          }
          if (leader) {
            const uint32_t my = res_base + __popc(lm & ((1u << lane) - 1));
            if (my < a.max_touched) {
              const unsigned long long expect = ((unsigned long long)E[g].y << 32) | E[g].x;
              const unsigned long long want = ((unsigned long long)my << 32) | a.step;
              const unsigned long long old = atomicCAS(reinterpret_cast<unsigned long long*>(a.ent + ei[g]), expect, want);
              if (old == expect) { idx = my; a.touched[my] = gs; }
              else { idx = (uint32_t)(old >> 32); a.touched[my] = SSB_INVALID_SLOT; }     // another warp was first: a hole
            } else {
              atomicAdd(a.counters + 2, 1u);
              idx = SSB_INVALID_SLOT;
            }
          }
          res_base += cnt;
          res_left -= cnt;
          if (stale) idx = __shfl_sync(sm, idx, __ffs(peers) - 1);
        }
        if (live && f0 + g < F) a.stash[(size_t)(f0 + g) * n_ex + e] = valid[g] ? idx : SSB_INVALID_SLOT;
      }
    }
  }
  if ((uint32_t)lane < res_left && res_base + lane < a.max_touched) a.touched[res_base + lane] = SSB_INVALID_SLOT;
  if (res_left > 32 && (uint32_t)lane + 32 < res_left && res_base + lane + 32 < a.max_touched)
    a.touched[res_base + lane + 32] = SSB_INVALID_SLOT;
  grid_barrier(a.counters + 1, phase);
  stamp(1);

  // ---- B: one weight read per touched key ----------------------------------------------------------------------
  const uint32_t n_idx = min(*reinterpret_cast<volatile uint32_t*>(a.counters + 0), a.max_touched);
  for (uint32_t i = blockIdx.x * kThreads + threadIdx.x; i < n_idx; i += gridDim.x * kThreads) {
    const uint32_t gs = a.touched[i];
    if (gs == SSB_INVALID_SLOT) continue;
    const uint32_t o = gs / a.v.capacity, s = gs - o * a.v.capacity;
    a.wc[i] = ld_row_f1(a.v.rows[o] + s);
  }
  grid_barrier(a.counters + 1, phase);
  stamp(2);

  // ---- C: forward / backward -----------------------------------------------------------------------------------
  float loss_acc = 0.f;
  for (int64_t c = next_chunk(a.counters + 5, lane); c < chunks; c = next_chunk(a.counters + 5, lane)) {
    const int64_t e = c * 32 + lane;
    const bool live = e < n_ex;
    const size_t xs_e = a.feature_major ? 1 : (size_t)F, xs_f = a.feature_major ? (size_t)n_ex : 1;
    const float* xe = a.x ? a.x + (live ? (size_t)e : 0) * xs_e : nullptr;
    float z = 0.f;
#pragma unroll 8
    for (int f = 0; f < F; ++f) {
      const uint32_t idx = live ? a.stash[(size_t)f * n_ex + e] : SSB_INVALID_SLOT;
      if (idx != SSB_INVALID_SLOT) z += ld_strong_f32(a.wc + idx) * (xe ? __ldg(xe + (size_t)f * xs_f) : 1.0f);
    }
    float g = 0.f;
    if (live) {
      const float y = a.labels[e];
      g = fast_sigmoid(z) - y;
      loss_acc += y * neg_log_sigmoid(z) + (1.f - y) * neg_log_sigmoid(-z);
    }
#pragma unroll 2
    for (int f = 0; f < F; ++f) {
      const uint32_t idx = live ? a.stash[(size_t)f * n_ex + e] : SSB_INVALID_SLOT;
      float cgrad = (idx != SSB_INVALID_SLOT) ? g * (xe ? __ldg(xe + (size_t)f * xs_f) : 1.0f) : 0.f;
      uint32_t tgt = idx;
      // a feature the whole warp shares (a numeric field): one add for the warp
      const uint32_t idx0 = __shfl_sync(0xffffffffu, idx, 0);
      if (__all_sync(0xffffffffu, idx == idx0)) {
        cgrad = warp_sum(cgrad);
        if (lane != 0) tgt = SSB_INVALID_SLOT;
      }
      if (tgt == SSB_INVALID_SLOT || cgrad == 0.f) continue;
      const uint32_t hsh = (tgt * 0x9E3779B1u) >> (32 - kAggBits);
      uint32_t cur = agg_key[hsh];
      if (cur == kAggEmpty) cur = atomicCAS(&agg_key[hsh], kAggEmpty, tgt);
      if (cur == kAggEmpty || cur == tgt) {
        const unsigned long long dv = (unsigned long long)__float2ll_rn(cgrad * kFixScale);
        const uint32_t dl = (uint32_t)dv, dh = (uint32_t)(dv >> 32);
        const uint32_t old = atomicAdd(&agg_lo[hsh], dl);
        const uint32_t up = dh + ((uint32_t)(old + dl) < old ? 1u : 0u);        // carry out of the low word
        if (up) atomicAdd(&agg_hi[hsh], up);
      } else {
        atomicAdd(a.gc + tgt, cgrad);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kAggSize; i += kThreads) {
    const uint32_t k = agg_key[i];
    if (k != kAggEmpty) {
      const long long fx = (long long)(((unsigned long long)agg_hi[i] << 32) | (unsigned long long)agg_lo[i]);
      const float vsum = (float)((double)fx * (1.0 / (double)kFixScale));
      if (vsum != 0.f) atomicAdd(a.gc + k, vsum);
    }
  }
  loss_acc = warp_sum(loss_acc);
  if (lane == 0) sloss[warp] = loss_acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < kThreads / 32; ++w) t += sloss[w];
    if (t != 0.f) atomicAdd(a.loss_out, t);
    if (blockIdx.x == 0) atomicAdd(a.loss_out + 1, (float)n_ex);
  }
  grid_barrier(a.counters + 1, phase);
  stamp(3);

  // ---- D: push, once per touched key ---------------------------------------------------------------------------
  const int SW = (a.v.state_words > 0 ? a.v.state_words : 1);
  for (uint32_t i = blockIdx.x * kThreads + threadIdx.x; i < n_idx; i += gridDim.x * kThreads) {
    const uint32_t gs = a.touched[i];
    if (gs == SSB_INVALID_SLOT) continue;
    const float g = ld_strong_f32(a.gc + i);
    if (g == 0.f) continue;
    a.gc[i] = 0.f;
    const uint32_t o = gs / a.v.capacity, s = gs - o * a.v.capacity;
    float* row = a.v.rows[o] + s;
    float* st = (Rule::kStateWords > 0) ? a.v.state[o] + (size_t)s * SW : nullptr;
    const float d = Rule::apply1(row, st, 1, g, a.h);
    if (a.v.shadow[o]) {
      const float o1 = atom_add_f1_sys(row, d);
      reinterpret_cast<__nv_bfloat16*>(a.v.shadow[o])[s] = __float2bfloat16_rn(o1 + d);
    } else {
      red_add_f1_sys(row, d);
    }
  }
  stamp(4);        // block 0's own end: the other blocks finish within the same few microseconds
}

template <class Rule>
int launch(const Args& a, int blocks_per_sm, cudaStream_t stream) {
  static int max_blocks = 0;
  if (max_blocks == 0) {
    int per_sm = 0, dev = 0, sms = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, lr_fused_step_kernel<Rule>, kThreads, 0) != cudaSuccess) return -1;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    max_blocks = per_sm * sms;
    if (max_blocks <= 0) return -1;
  }
  int blocks = max_blocks;
  if (blocks_per_sm > 0) blocks = std::min(max_blocks, blocks_per_sm * kNumSMs);
  const int64_t chunks = (a.n_ex + 31) / 32;
  blocks = (int)std::max<int64_t>(1, std::min<int64_t>(blocks, (chunks + kThreads / 32 - 1) / (kThreads / 32)));
  cudaError_t e = cudaMemsetAsync(a.counters, 0, 8 * sizeof(uint32_t), stream);      // [8..12] are overwritten by the kernel
  if (e != cudaSuccess) return (int)e;
  void* params[] = {(void*)&a};
  e = cudaLaunchCooperativeKernel((const void*)lr_fused_step_kernel<Rule>, dim3(blocks), dim3(kThreads), params, 0, stream);
  return (int)e;
}

}  // namespace lrf
}  // namespace ssb

// rule_id: 0 SGD, 1 AdaGrad, 2 FTRL (include/ssb200_rule.cuh). counters: 16 x uint32 device scratch ([0..7] zeroed here):
// [0] allocated compact indices, [2] index overflow (must stay 0), [3] keys without a slot (table full or key outside
// the declared feature space; must stay 0).
extern "C" int ssb_lr_fused_step(const SsbTableView* view, const uint32_t* keys, const float* x, const float* labels, int64_t n_ex,
                                 int32_t n_feat, void* entries, uint64_t dense_keys, uint32_t* stash, uint32_t* touched,
                                 float* wc, float* gc, uint32_t max_touched, uint32_t* counters, uint32_t step, float* loss_out,
                                 int32_t rule_id, const void* hyper, int32_t blocks_per_sm, void* stream) {
  // blocks_per_sm bit 30: keys / x are feature-major ([n_feat][n_ex])
  const int feature_major = (blocks_per_sm >> 30) & 1;
  blocks_per_sm &= 0x3FFFFFFF;
  if (!view || view->dim != 1 || !keys || !labels || n_feat < 1 || !entries || !stash || !touched || !wc || !gc || !counters ||
      !loss_out || !hyper || step == 0 || max_touched == 0)
    return (int)cudaErrorInvalidValue;
  if (n_ex <= 0) return 0;
  ssb::lrf::Args a;
  a.v = *view; a.keys = keys; a.x = x; a.labels = labels; a.n_ex = n_ex; a.n_feat = n_feat; a.feature_major = feature_major;
  a.ent = reinterpret_cast<uint4*>(entries); a.dense_keys = dense_keys; a.stash = stash; a.touched = touched; a.wc = wc; a.gc = gc;
  a.max_touched = max_touched; a.counters = counters; a.step = step; a.loss_out = loss_out;
  a.h = *static_cast<const ssb::rule::Hyper*>(hyper);
  cudaStream_t s = (cudaStream_t)stream;
  switch (rule_id) {
    case 0: return ssb::lrf::launch<ssb::rule::Sgd>(a, blocks_per_sm, s);
    case 1: if (ssb::rule_check(view, 1)) return (int)cudaErrorInvalidValue; return ssb::lrf::launch<ssb::rule::AdaGrad>(a, blocks_per_sm, s);
    case 2: if (ssb::rule_check(view, 2)) return (int)cudaErrorInvalidValue; return ssb::lrf::launch<ssb::rule::Ftrl>(a, blocks_per_sm, s);
    default: return (int)cudaErrorInvalidValue;
  }
}
