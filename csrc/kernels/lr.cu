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
// pull -> compute -> push, the server applies the update rule per key once per minibatch, parameter/
// sparse_access_method.h:33-46). The reference does this with a request/response round trip per server; the unfused
// path here (dedup -> pull -> lr_fwd_bwd -> push) with four launches and a sort-free hash dedup. This kernel does it
// in four phases separated by grid-wide barriers, with the TABLE ITSELF as the dedup domain:
//
//   A  resolve   lane = example, loop over its features: key -> global slot through a dense key->slot map (4 bytes
//                per feature id, filled on first sight by the ordinary probe / insert protocol); the first lane to
//                touch a slot in this step (tag exchange, pre-checked with a plain load) appends it to the step's
//                touched list; the slot is stashed per instance ([feature][example], coalesced).
//   B  gather    one lane per TOUCHED slot: read the weight from the owner shard (own HBM or a peer's over NVLink,
//                once per unique key and rank) into the worker-side entry of that slot.
//   C  compute   lane = example: dot over the stashed slots, sigmoid, log-loss, g = p - y; the per-feature gradient
//                g * x goes to a block-local shared-memory table of 64-bit fixed-point sums (native integer smem
//                atomics; features shared by a whole warp — the numeric fields — are warp-reduced first), spilling
//                to the entry's accumulator in global memory; the block flushes its table at the end of the phase.
//                Hot features therefore cost one global atomic per block, not one per example.
//   D  push      one lane per touched slot: take the accumulated gradient (and zero it), apply the update rule
//                (Rule::apply1 of include/ssb200_rule.cuh: SGD, AdaGrad, FTRL) to the owner's weight and state.
//
// Worker-side memory (allocated once, sized for 180 GB of HBM): slot_of[2^feature_bits] (1 GB at 2^28) and one
// 16-byte entry {tag, w, gacc} per global slot (2 GB at 2^27 slots).
// The grid is launched cooperatively (all blocks co-resident), so the barriers cannot deadlock; they are bounded.
// ================================================================================================================
#include "table_probe.cuh"

namespace ssb {
namespace lrf {

struct __align__(16) Entry {
  uint32_t tag;    // step number of the last step that touched this slot
  float w;         // the weight as pulled in this step
  float gacc;      // this worker's gradient sum for the step (zero between steps)
  uint32_t pad;
};

struct Args {
  SsbTableView v;
  const uint32_t* keys;      // [n_ex][n_feat]; SSB_EMPTY_KEY = no feature in this position
  const float* x;            // [n_ex][n_feat] or nullptr (all ones)
  const float* labels;       // [n_ex] in {0, 1}
  int64_t n_ex;
  int n_feat;
  uint32_t* slot_of;         // [slot_of_size] key -> global slot, SSB_INVALID_SLOT = not seen yet; nullptr = always probe
  uint64_t slot_of_size;
  Entry* ent;                // [num_shards * capacity]
  uint32_t* inst_slot;       // [n_feat][n_ex] stash
  uint32_t* touched;         // [max_touched]
  uint32_t max_touched;
  uint32_t* counters;        // [0] touched count, [1] grid barrier, [2] touched overflow, [3] table full / unresolved
  uint32_t step;             // != 0, different from the previous step's
  float* loss_out;           // [2]: sum of log-loss, examples
  rule::Hyper h;
};

constexpr int kThreads = 256;
constexpr int kAggBits = 11, kAggSize = 1 << kAggBits;      // block-local gradient table: 2048 x (4 + 8) bytes
constexpr float kFixScale = 68719476736.0f;                  // 2^36: |block sum| < 2^27 fits int64 with room
constexpr uint32_t kAggEmpty = 0xFFFFFFFFu;

__device__ __forceinline__ uint32_t ld_acquire_gpu_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
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

// key -> global slot for one lane's key; lanes that miss the map go through the warp-cooperative probe one by one.
__device__ __forceinline__ uint32_t resolve(const Args& a, uint32_t key, bool valid, int lane) {
  uint32_t gs = SSB_INVALID_SLOT;
  const bool mapped = valid && a.slot_of && key < a.slot_of_size;
  // entries only ever change from INVALID to their final value (other warps of this launch may be writing them)
  if (mapped) gs = *reinterpret_cast<volatile const uint32_t*>(a.slot_of + key);
  unsigned need = __ballot_sync(0xffffffffu, valid && gs == SSB_INVALID_SLOT);
  while (need) {
    const int src = __ffs(need) - 1;
    need &= need - 1;
    const uint32_t k = __shfl_sync(0xffffffffu, key, src);
    Probe p = probe_insert(a.v, k, true, lane);
    if (p.slot != SSB_INVALID_SLOT && p.is_new) {
      if (lane == 0) {
        a.v.rows[p.owner][p.slot] = init_of(a.v, k, 0);      // dim 1; the state is zero since allocation
        if (a.v.shadow[p.owner]) reinterpret_cast<__nv_bfloat16*>(a.v.shadow[p.owner])[p.slot] = __float2bfloat16_rn(init_of(a.v, k, 0));
      }
      publish(a.v, p, k, lane);
    }
    const uint32_t g = (p.slot == SSB_INVALID_SLOT) ? SSB_INVALID_SLOT : (uint32_t)p.owner * a.v.capacity + p.slot;
    if (lane == src) {
      gs = g;
      if (g == SSB_INVALID_SLOT) atomicAdd(a.counters + 3, 1u);
      else if (mapped) a.slot_of[key] = g;
    }
  }
  return gs;
}

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
  const int64_t gwarp = (int64_t)blockIdx.x * (kThreads / 32) + warp, nwarps = (int64_t)gridDim.x * (kThreads / 32);
  uint32_t phase = 0;
  for (int i = threadIdx.x; i < kAggSize; i += kThreads) { agg_key[i] = kAggEmpty; agg_lo[i] = 0u; agg_hi[i] = 0u; }

  // ---- A: resolve + touched list ------------------------------------------------------------------------------
  for (int64_t c = gwarp; c < chunks; c += nwarps) {
    const int64_t e = c * 32 + lane;
    const bool live = e < n_ex;
    const uint32_t* ke = a.keys + (live ? e : 0) * F;
#pragma unroll 1
    for (int f = 0; f < F; ++f) {
      const uint32_t key = live ? __ldg(ke + f) : SSB_EMPTY_KEY;
      const bool valid = live && key < SSB_LOCKED_KEY;
      const uint32_t gs = resolve(a, key, valid, lane);
      bool first = false;
      if (gs != SSB_INVALID_SLOT) {
        Entry* en = a.ent + gs;
        if (*reinterpret_cast<volatile uint32_t*>(&en->tag) != a.step) first = atomicExch(&en->tag, a.step) != a.step;
      }
      const unsigned fm = __ballot_sync(0xffffffffu, first);
      if (fm) {
        uint32_t base = 0;
        if (lane == __ffs(fm) - 1) base = atomicAdd(a.counters + 0, (uint32_t)__popc(fm));
        base = __shfl_sync(0xffffffffu, base, __ffs(fm) - 1);
        if (first) {
          const uint32_t pos = base + __popc(fm & ((1u << lane) - 1));
          if (pos < a.max_touched) a.touched[pos] = gs; else atomicAdd(a.counters + 2, 1u);
        }
      }
      if (live) a.inst_slot[(size_t)f * n_ex + e] = gs;
    }
  }
  grid_barrier(a.counters + 1, phase);

  // ---- B: one weight read per touched slot ---------------------------------------------------------------------
  const uint32_t n_touched = min(*reinterpret_cast<volatile uint32_t*>(a.counters + 0), a.max_touched);
  for (uint32_t i = blockIdx.x * kThreads + threadIdx.x; i < n_touched; i += gridDim.x * kThreads) {
    const uint32_t gs = a.touched[i];
    const uint32_t o = gs / a.v.capacity, s = gs - o * a.v.capacity;
    a.ent[gs].w = ld_row_f1(a.v.rows[o] + s);
  }
  grid_barrier(a.counters + 1, phase);

  // ---- C: forward / backward -----------------------------------------------------------------------------------
  float loss_acc = 0.f;
  for (int64_t c = gwarp; c < chunks; c += nwarps) {
    const int64_t e = c * 32 + lane;
    const bool live = e < n_ex;
    const float* xe = a.x ? a.x + (live ? e : 0) * F : nullptr;
    float z = 0.f;
#pragma unroll 4
    for (int f = 0; f < F; ++f) {
      const uint32_t gs = live ? a.inst_slot[(size_t)f * n_ex + e] : SSB_INVALID_SLOT;
      if (gs != SSB_INVALID_SLOT) z += *reinterpret_cast<volatile float*>(&a.ent[gs].w) * (xe ? __ldg(xe + f) : 1.0f);
    }
    float g = 0.f;
    if (live) {
      const float y = a.labels[e];
      g = fast_sigmoid(z) - y;
      loss_acc += y * neg_log_sigmoid(z) + (1.f - y) * neg_log_sigmoid(-z);
    }
#pragma unroll 1
    for (int f = 0; f < F; ++f) {
      const uint32_t gs = live ? a.inst_slot[(size_t)f * n_ex + e] : SSB_INVALID_SLOT;
      float cgrad = (gs != SSB_INVALID_SLOT) ? g * (xe ? __ldg(xe + f) : 1.0f) : 0.f;
      uint32_t tgt = gs;
      // a feature the whole warp shares (a numeric field): one add for the warp
      const uint32_t gs0 = __shfl_sync(0xffffffffu, gs, 0);
      if (__all_sync(0xffffffffu, gs == gs0)) {
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
      } else
        atomicAdd(&a.ent[tgt].gacc, cgrad);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kAggSize; i += kThreads) {
    const uint32_t k = agg_key[i];
    if (k != kAggEmpty) {
      const long long fx = (long long)(((unsigned long long)agg_hi[i] << 32) | (unsigned long long)agg_lo[i]);
      const float vsum = (float)((double)fx * (1.0 / (double)kFixScale));
      if (vsum != 0.f) atomicAdd(&a.ent[k].gacc, vsum);
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

  // ---- D: push, once per touched slot --------------------------------------------------------------------------
  const int SW = (a.v.state_words > 0 ? a.v.state_words : 1);
  for (uint32_t i = blockIdx.x * kThreads + threadIdx.x; i < n_touched; i += gridDim.x * kThreads) {
    const uint32_t gs = a.touched[i];
    Entry* en = a.ent + gs;
    const float g = *reinterpret_cast<volatile float*>(&en->gacc);
    if (g == 0.f) continue;
    en->gacc = 0.f;
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
  cudaError_t e = cudaMemsetAsync(a.counters, 0, 4 * sizeof(uint32_t), stream);
  if (e != cudaSuccess) return (int)e;
  void* params[] = {(void*)&a};
  e = cudaLaunchCooperativeKernel((const void*)lr_fused_step_kernel<Rule>, dim3(blocks), dim3(kThreads), params, 0, stream);
  return (int)e;
}

}  // namespace lrf
}  // namespace ssb

// rule_id: 0 SGD, 1 AdaGrad, 2 FTRL (include/ssb200_rule.cuh). counters: 4 x uint32 device scratch (zeroed here):
// [0] touched slots, [2] touched-list overflow (must stay 0), [3] keys that found no slot (table full; must stay 0).
extern "C" int ssb_lr_fused_step(const SsbTableView* view, const uint32_t* keys, const float* x, const float* labels, int64_t n_ex,
                                 int32_t n_feat, uint32_t* slot_of, uint64_t slot_of_size, void* entries, uint32_t* inst_slot,
                                 uint32_t* touched, uint32_t max_touched, uint32_t* counters, uint32_t step, float* loss_out,
                                 int32_t rule_id, const void* hyper, int32_t blocks_per_sm, void* stream) {
  if (!view || view->dim != 1 || !keys || !labels || n_feat < 1 || !entries || !inst_slot || !touched || !counters || !loss_out ||
      !hyper || step == 0 || max_touched == 0)
    return (int)cudaErrorInvalidValue;
  if (n_ex <= 0) return 0;
  ssb::lrf::Args a;
  a.v = *view; a.keys = keys; a.x = x; a.labels = labels; a.n_ex = n_ex; a.n_feat = n_feat; a.slot_of = slot_of;
  a.slot_of_size = slot_of ? slot_of_size : 0; a.ent = reinterpret_cast<ssb::lrf::Entry*>(entries); a.inst_slot = inst_slot;
  a.touched = touched; a.max_touched = max_touched; a.counters = counters; a.step = step; a.loss_out = loss_out; a.h = *static_cast<const ssb::rule::Hyper*>(hyper);
  cudaStream_t s = (cudaStream_t)stream;
  switch (rule_id) {
    case 0: return ssb::lrf::launch<ssb::rule::Sgd>(a, blocks_per_sm, s);
    case 1: if (ssb::rule_check(view, 1)) return (int)cudaErrorInvalidValue; return ssb::lrf::launch<ssb::rule::AdaGrad>(a, blocks_per_sm, s);
    case 2: if (ssb::rule_check(view, 2)) return (int)cudaErrorInvalidValue; return ssb::lrf::launch<ssb::rule::Ftrl>(a, blocks_per_sm, s);
    default: return (int)cudaErrorInvalidValue;
  }
}
