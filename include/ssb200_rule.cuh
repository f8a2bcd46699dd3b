// ssb200_rule.cuh — the device-side update rule: the seam where a user's optimizer enters the push path.
//
// ref: core/parameter/sparse_access_method.h:21-46 — the reference's user supplies a PullAccessMethod
//      (init_param / get_pull_value) and a PushAccessMethod (apply_push_value) as C++ classes that the server
//      calls once per key on the host. Here the rule is a stateless struct of __device__ functions that the push
//      kernels are instantiated with; everything else (slot decode, atomics on rows that other GPUs update at the
//      same time, the bf16 row mirror, bf16 gradients that crossed NVLink) stays in the kernels of this header.
//
// Writing a rule (see Sgd / AdaGrad / Ftrl below for three complete ones):
//
//   struct MyRule {
//     static constexpr int  kStateWords  = 1;      // floats of optimizer state per weight (0, 1, 2 ...)
//     static constexpr bool kSkipZeroGrad = true;  // a zero gradient changes nothing (true for almost every rule)
//     static constexpr int  kInit = SSB_INIT_UNIFORM;   // how a missing row is created by the pull kernels
//     // Four consecutive weights of one row. `w` points at the weights, `st` at state word 0 of the same four
//     // weights, word k is at st + k * D. Other warps (other senders' segments, other GPUs) may update the same
//     // row at the same moment (Hogwild, as in the reference's server): touch state with the atomics below.
//     // Return the INCREMENT to add to the weights; the kernel adds it atomically and keeps the mirror in step.
//     static __device__ float4 apply4(const float* w, float* st, int D, float4 g, const Hyper& h);
//     static __device__ float  apply1(const float* w, float* st, int D, float g, const Hyper& h);   // D % 4 != 0
//   };
//   SSB200_EXPORT_RULE(myrule, MyRule)       // in a .cu of your own: extern "C" ssb_rule_myrule_* entry points
//
// `swiftsnails_b200.ops.rules.Rule.from_source("myrule", path)` compiles such a file for sm_100a and loads it;
// `framework.PushMethod("myrule", lr=..)` / `ShardedTable.push_rule` then use it by name. No file of the framework
// is edited. The table must have been created with `with_state >= kStateWords`.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include "ssb200.h"

namespace ssb {

// ---- system-scope accesses to rows / state (the other writer may be a peer GPU over NVLink) --------------------
// Streaming 16-byte load that bypasses L1 allocation (rows are mutated by other SMs / GPUs).
__device__ __forceinline__ float4 ld_row_f4(const float* p) {
  float4 r;
  asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p)
               : "memory");
  return r;
}
__device__ __forceinline__ float ld_row_f1(const float* p) {
  float r;
  asm volatile("ld.relaxed.sys.global.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p) : "memory");
  return r;
}
// Vector reduction into (possibly remote) memory: REDG.E.ADD.F32x4 .SYS
__device__ __forceinline__ void red_add_f4_sys(float* p, float4 v) {
  asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void red_add_f1_sys(float* p, float v) {
  asm volatile("red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
// The same, returning the value before the add.
__device__ __forceinline__ float4 atom_add_f4_sys(float* p, float4 v) {
  float4 r;
  asm volatile("atom.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4], {%5,%6,%7,%8};"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
  return r;
}
__device__ __forceinline__ float atom_add_f1_sys(float* p, float v) {
  float r;
  asm volatile("atom.relaxed.sys.global.add.f32 %0, [%1], %2;" : "=f"(r) : "l"(p), "f"(v) : "memory");
  return r;
}

namespace rule {

// Hyper-parameters, passed to the kernel by value. lr / eps have their usual meaning; a, b, c are the rule's own
// (Ftrl: a = l1, b = l2; a rule with momentum would put its coefficient there).
struct Hyper {
  float lr, eps, a, b, c;
  uint32_t step;        // the caller's step counter (rules with bias correction)
};

// w -= lr * g.  ref: the word2vec app's SGD push (apps/word2vec/word2vec.h, `param -= lr * grad`).
struct Sgd {
  static constexpr int kStateWords = 0;
  static constexpr bool kSkipZeroGrad = true;
  static constexpr int kInit = SSB_INIT_UNIFORM;
  static __device__ __forceinline__ float4 apply4(const float*, float*, int, float4 g, const Hyper& h) {
    return make_float4(-h.lr * g.x, -h.lr * g.y, -h.lr * g.z, -h.lr * g.w);
  }
  static __device__ __forceinline__ float apply1(const float*, float*, int, float g, const Hyper& h) { return -h.lr * g; }
};

// acc += g^2 ; w -= lr * g / sqrt(acc + eps).  One vector atomic claims "my" accumulator value, so concurrent
// updates of one weight each see a distinct, growing accumulator (no lost g^2).
struct AdaGrad {
  static constexpr int kStateWords = 1;
  static constexpr bool kSkipZeroGrad = true;
  static constexpr int kInit = SSB_INIT_UNIFORM;
  static __device__ __forceinline__ float4 apply4(const float*, float* st, int, float4 g, const Hyper& h) {
    const float4 g2 = make_float4(g.x * g.x, g.y * g.y, g.z * g.z, g.w * g.w);
    const float4 a = atom_add_f4_sys(st, g2);
    return make_float4(-h.lr * g.x * rsqrtf(a.x + g2.x + h.eps), -h.lr * g.y * rsqrtf(a.y + g2.y + h.eps),
                       -h.lr * g.z * rsqrtf(a.z + g2.z + h.eps), -h.lr * g.w * rsqrtf(a.w + g2.w + h.eps));
  }
  static __device__ __forceinline__ float apply1(const float*, float* st, int, float g, const Hyper& h) {
    const float g2 = g * g;
    const float a = atom_add_f1_sys(st, g2);
    return -h.lr * g * rsqrtf(a + g2 + h.eps);
  }
};

// FTRL-proximal (McMahan et al. 2013), the usual optimizer of sparse logistic regression on a parameter server.
// Per weight: n += g^2 ; sigma = (sqrt(n) - sqrt(n_old)) / alpha ; z += g - sigma * w ;
//             w = 0 if |z| <= l1 else -(z - sign(z) * l1) / ((beta + sqrt(n)) / alpha + l2).
// State word 0 = z, word 1 = n; both only ever receive ADDS, so they go in with atomics and no update is lost when
// several senders hit one weight; the weight read for sigma * w may be one update old (Hogwild, as on the
// reference's server). alpha = h.lr, beta = h.eps, l1 = h.a, l2 = h.b. Rows start at zero (kInit).
struct Ftrl {
  static constexpr int kStateWords = 2;
  static constexpr bool kSkipZeroGrad = true;
  static constexpr int kInit = SSB_INIT_ZERO;
  static __device__ __forceinline__ float weight(float z, float n, const Hyper& h) {
    const float az = fabsf(z);
    if (az <= h.a) return 0.f;
    return -(z - copysignf(h.a, z)) / ((h.eps + sqrtf(n)) / h.lr + h.b);
  }
  static __device__ __forceinline__ float apply1(const float* w, float* st, int D, float g, const Hyper& h) {
    const float g2 = g * g;
    const float n_old = atom_add_f1_sys(st + D, g2);
    const float w_cur = ld_row_f1(w);
    const float sigma = (sqrtf(n_old + g2) - sqrtf(n_old)) / h.lr;
    const float dz = g - sigma * w_cur;
    const float z_old = atom_add_f1_sys(st, dz);
    return weight(z_old + dz, n_old + g2, h) - w_cur;
  }
  static __device__ __forceinline__ float4 apply4(const float* w, float* st, int D, float4 g, const Hyper& h) {
    const float4 g2 = make_float4(g.x * g.x, g.y * g.y, g.z * g.z, g.w * g.w);
    const float4 n_old = atom_add_f4_sys(st + D, g2);
    const float4 w_cur = ld_row_f4(w);
    const float inv = 1.f / h.lr;
    const float4 dz = make_float4(g.x - (sqrtf(n_old.x + g2.x) - sqrtf(n_old.x)) * inv * w_cur.x,
                                  g.y - (sqrtf(n_old.y + g2.y) - sqrtf(n_old.y)) * inv * w_cur.y,
                                  g.z - (sqrtf(n_old.z + g2.z) - sqrtf(n_old.z)) * inv * w_cur.z,
                                  g.w - (sqrtf(n_old.w + g2.w) - sqrtf(n_old.w)) * inv * w_cur.w);
    const float4 z_old = atom_add_f4_sys(st, dz);
    return make_float4(weight(z_old.x + dz.x, n_old.x + g2.x, h) - w_cur.x, weight(z_old.y + dz.y, n_old.y + g2.y, h) - w_cur.y,
                       weight(z_old.z + dz.z, n_old.z + g2.z, h) - w_cur.z, weight(z_old.w + dz.w, n_old.w + g2.w, h) - w_cur.w);
  }
};

}  // namespace rule

// ---- the kernels a rule is instantiated into ---------------------------------------------------------------------
__device__ __forceinline__ uint2 rule_pack_bf16x4(float4 v) {
  __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
  uint2 pk;
  pk.x = *reinterpret_cast<uint32_t*>(&lo);
  pk.y = *reinterpret_cast<uint32_t*>(&hi);
  return pk;
}
__device__ __forceinline__ uint32_t rule_effective_n(int64_t n, const uint32_t* n_dev) {
  uint32_t ne = (uint32_t)n;
  if (n_dev) { uint32_t d = *n_dev; if (d < ne) ne = d; }
  return ne;
}
// floats of state per row as the table was allocated (0 in old views means 1)
__device__ __forceinline__ int rule_state_stride(const SsbTableView& v) { return (v.state_words > 0 ? v.state_words : 1) * v.dim; }

// One row, one warp: apply the rule to every 4-weight group with a non-zero gradient.
// kMirror: this launch is the ONLY writer of the row (a worker's dedup'd gradient buffer holds a slot once), so the
// bf16 mirror can be written from the value the atomic returned. Launches with several writers per row (inbox
// segments) pass kMirror = false and run rule_mirror_refresh_kernel afterwards.
template <class Rule, bool kMirror, class GradLoad>
__device__ __forceinline__ void rule_apply_row(const SsbTableView& v, uint32_t g, int lane, const rule::Hyper& h, GradLoad load4,
                                               float* zero_fp32_grads) {
  const int D = v.dim;
  const uint32_t o = g / v.capacity, s = g - o * v.capacity;
  float* row = v.rows[o] + (size_t)s * D;
  float* st = (Rule::kStateWords > 0) ? v.state[o] + (size_t)s * rule_state_stride(v) : nullptr;
  __nv_bfloat16* sh = (kMirror && v.shadow[o]) ? reinterpret_cast<__nv_bfloat16*>(v.shadow[o]) + (size_t)s * D : nullptr;
  for (int j = lane * 4; j < D; j += 128) {
    const float4 gv = load4(j);
    if (Rule::kSkipZeroGrad && gv.x == 0.f && gv.y == 0.f && gv.z == 0.f && gv.w == 0.f) continue;
    const float4 d = Rule::apply4(row + j, st ? st + j : nullptr, D, gv, h);
    if (kMirror && sh) {
      const float4 o4 = atom_add_f4_sys(row + j, d);
      *reinterpret_cast<uint2*>(sh + j) = rule_pack_bf16x4(make_float4(o4.x + d.x, o4.y + d.y, o4.z + d.z, o4.w + d.w));
    } else {
      red_add_f4_sys(row + j, d);
    }
    if (zero_fp32_grads) *reinterpret_cast<float4*>(zero_fp32_grads + j) = make_float4(0, 0, 0, 0);
  }
}

// Push of a worker's fp32 gradient rows straight into the owners' rows (local HBM or a peer's over NVLink).
// only_shard >= 0: apply only the rows that shard owns (the worker's own rows when the rest goes through the inbox).
// 32 registers = 8 blocks per SM for Sgd / AdaGrad: rules that wait on an `atom` per group hide it with resident warps
// (at 40 registers / 6 blocks the AdaGrad push ran 20 % slower). Rules with two or more state words get 48 registers.
template <class Rule> constexpr int rule_min_blocks() { return Rule::kStateWords >= 2 ? 5 : 8; }
template <class Rule, bool kShadow>
__global__ void __launch_bounds__(256, rule_min_blocks<Rule>())
rule_push_kernel(const __grid_constant__ SsbTableView v, const uint32_t* __restrict__ gslot, float* grads, int64_t n,
                 const uint32_t* __restrict__ n_dev, const __grid_constant__ rule::Hyper h, int zero_grads, int only_shard) {
  const int lane = threadIdx.x & 31;
  const uint32_t ne = rule_effective_n(n, n_dev);
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  const int D = v.dim;
  for (uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < ne; w += warps) {
    const uint32_t g = __ldg(gslot + w);
    if (g == SSB_INVALID_SLOT) continue;
    if (only_shard >= 0 && (int)(g / v.capacity) != only_shard) continue;
    float* gr = grads + (size_t)w * D;
    if ((D & 3) == 0) {
      rule_apply_row<Rule, kShadow>(v, g, lane, h, [gr](int j) { return *reinterpret_cast<const float4*>(gr + j); },
                                    zero_grads ? gr : nullptr);
    } else {
      const uint32_t o = g / v.capacity, s = g - o * v.capacity;
      float* row = v.rows[o] + (size_t)s * D;
      float* st = (Rule::kStateWords > 0) ? v.state[o] + (size_t)s * rule_state_stride(v) : nullptr;
      __nv_bfloat16* sh = (kShadow && v.shadow[o]) ? reinterpret_cast<__nv_bfloat16*>(v.shadow[o]) + (size_t)s * D : nullptr;
      for (int j = lane; j < D; j += 32) {
        const float gv = gr[j];
        if (Rule::kSkipZeroGrad && gv == 0.f) continue;
        const float d = Rule::apply1(row + j, st ? st + j : nullptr, D, gv, h);
        if (kShadow && sh) { const float o1 = atom_add_f1_sys(row + j, d); sh[j] = __float2bfloat16_rn(o1 + d); }
        else red_add_f1_sys(row + j, d);
        if (zero_grads) gr[j] = 0.f;
      }
    }
  }
}

// Which rows of a multi-writer launch really had a second writer. `seen[slot]` holds 2 * tag once a launch with that
// tag has touched the slot and 2 * tag + 1 once it has touched it AGAIN (atomicMax: tags grow from launch to launch).
// Rows that stay at 2 * tag had ONE writer, so the mirror that writer stored from its atomic's return value is final;
// only the rows at 2 * tag + 1 (the hot words that every sender carries) need the trailing refresh, which walks the
// slot lists again but touches the 768 bytes of a row only when its mark says so. No list, no shared counter.
// One-shard views only (slot == global slot).
struct MirrorTrack {
  uint32_t* seen;        // [capacity], zero-initialised
  uint32_t tag;          // != 0, larger for every later launch, < 2^31
};

// Owner-side apply of inbox segments whose gradient rows travelled as bf16 (half the NVLink bytes); the rule runs in
// fp32 on fp32 rows / state. blockIdx.y = segment (one per sender): segment y holds rows
// [y * seg_stride, y * seg_stride + n_dev[y]). Hot rows arrive from every sender, so a row can have several writers
// in this launch. kTrack = false: the mirror is NOT written here, rule_mirror_refresh_kernel rewrites it for every
// row afterwards. kTrack = true: every writer stores the mirror from its atomic's return value and second writers
// are recorded (MirrorTrack), so that the refresh only revisits rows that need it. D % 4 == 0.
template <class Rule, bool kTrack>
__global__ void __launch_bounds__(256)
rule_push_bf16_kernel(const __grid_constant__ SsbTableView v, const uint32_t* __restrict__ gslot,
                      const __nv_bfloat16* __restrict__ grads, int64_t n, const uint32_t* __restrict__ n_dev,
                      const __grid_constant__ rule::Hyper h, int64_t seg_stride, const __grid_constant__ MirrorTrack tr) {
  if (n_dev) n_dev += blockIdx.y;
  gslot += (size_t)blockIdx.y * seg_stride;
  grads += (size_t)blockIdx.y * seg_stride * v.dim;
  const int lane = threadIdx.x & 31;
  const uint32_t ne = rule_effective_n(n, n_dev);
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  const int D = v.dim;
  for (uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < ne; w += warps) {
    const uint32_t g = __ldg(gslot + w);
    if (g == SSB_INVALID_SLOT) continue;
    if (kTrack) {
      if (g >= v.capacity) continue;                                 // one-shard view: slot == global slot
      if (lane == 0 && atomicMax(tr.seen + g, 2u * tr.tag) >= 2u * tr.tag)     // not the first writer of this launch
        atomicMax(tr.seen + g, 2u * tr.tag + 1u);
    }
    const __nv_bfloat16* gr = grads + (size_t)w * D;
    rule_apply_row<Rule, kTrack>(v, g, lane, h, [gr](int j) {
      const uint2 pk = *reinterpret_cast<const uint2*>(gr + j);
      const float2 a01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&pk.x));
      const float2 a23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&pk.y));
      return make_float4(a01.x, a01.y, a23.x, a23.y);
    }, nullptr);
  }
}

// mirror[slot] = bf16(row[slot]) for every slot of the given segments. Runs in stream order AFTER a launch that had
// several writers per row: every writer of that launch is done, so whichever warp refreshes a row (a row listed by
// three senders is refreshed three times) stores the same, final value — the mirror is bit-exactly bf16(row).
// The rows were just written, so the reads are L2 hits.
__device__ __forceinline__ void rule_refresh_row(const SsbTableView& v, uint32_t g, int lane) {
  const int D = v.dim;
  const uint32_t o = g / v.capacity, s = g - o * v.capacity;
  if (o >= (uint32_t)v.num_shards || !v.shadow[o]) return;
  const float* row = v.rows[o] + (size_t)s * D;
  __nv_bfloat16* sh = reinterpret_cast<__nv_bfloat16*>(v.shadow[o]) + (size_t)s * D;
  if ((D & 3) == 0) {
    for (int j = lane * 4; j < D; j += 128) *reinterpret_cast<uint2*>(sh + j) = rule_pack_bf16x4(ld_row_f4(row + j));
  } else {
    for (int j = lane; j < D; j += 32) sh[j] = __float2bfloat16_rn(ld_row_f1(row + j));
  }
}

// The refresh after a TRACKED apply: walks the same slot lists, rewrites the mirror of the rows marked "several writers".
template <int kUnused = 0>
__global__ void __launch_bounds__(256)
rule_mirror_refresh_tracked_kernel(const __grid_constant__ SsbTableView v, const uint32_t* __restrict__ gslot, int64_t n,
                                   const uint32_t* __restrict__ n_dev, int64_t seg_stride, const __grid_constant__ MirrorTrack tr) {
  const int lane = threadIdx.x & 31;
  if (n_dev) n_dev += blockIdx.y;
  gslot += (size_t)blockIdx.y * seg_stride;
  const uint32_t ne = rule_effective_n(n, n_dev);
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  const uint32_t mark = 2u * tr.tag + 1u;
  // a warp looks at 32 slots at once (one per lane) and then refreshes the marked ones together
  for (uint32_t w0 = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 32; w0 < ne; w0 += warps * 32) {
    const uint32_t wl = w0 + lane;
    const uint32_t gl = wl < ne ? __ldg(gslot + wl) : SSB_INVALID_SLOT;
    const bool hit = gl != SSB_INVALID_SLOT && gl < v.capacity && *reinterpret_cast<volatile const uint32_t*>(tr.seen + gl) == mark;
    unsigned todo = __ballot_sync(0xffffffffu, hit);
    while (todo) {
      const int src = __ffs(todo) - 1;
      todo &= todo - 1;
      rule_refresh_row(v, __shfl_sync(0xffffffffu, gl, src), lane);
    }
  }
}

template <int kUnused = 0>
__global__ void __launch_bounds__(256)
rule_mirror_refresh_kernel(const __grid_constant__ SsbTableView v, const uint32_t* __restrict__ gslot, int64_t n,
                           const uint32_t* __restrict__ n_dev, int64_t seg_stride, int n_dev_stride) {
  if (n_dev) n_dev += (size_t)blockIdx.y * n_dev_stride;
  gslot += (size_t)blockIdx.y * seg_stride;
  const int lane = threadIdx.x & 31;
  const uint32_t ne = rule_effective_n(n, n_dev);
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  const int D = v.dim;
  (void)D;
  for (uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; w < ne; w += warps) {
    const uint32_t g = __ldg(gslot + w);
    if (g != SSB_INVALID_SLOT) rule_refresh_row(v, g, lane);
  }
}

// ---- owner PULLS its gradient rows: apply straight from the senders' outboxes -------------------------------------
// Measured on eight B200s: kernel STORES to seven peers at once move ~360 GB/s per GPU, kernel LOADS from seven peers
// ~620 GB/s. So instead of every sender writing rows into the owners' inboxes, a sender compacts what it has for each
// owner into ITS OWN memory (an outbox: local stores at HBM speed), raises the hand-off flags, and the owner's apply
// kernel reads the rows over NVLink while it applies them — the shipment and the apply are one kernel, and the
// rows cross the fabric as loads. src.rows[y] / src.slots[y] / src.count[y] point at sender y's sub-region FOR THIS
// owner (peer memory); blockIdx.y = sender. A warp reads 32 slots with one coalesced load and keeps the next row's
// 256 bytes in flight while it applies the current one. One-shard view (the owner's own shard); D % 4 == 0.
__device__ __forceinline__ uint2 ld_peer_u2(const void* p) {
  uint2 r;
  asm volatile("ld.relaxed.sys.global.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ uint32_t ld_peer_u32(const uint32_t* p) {
  uint32_t r;
  asm volatile("ld.relaxed.sys.global.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ float4 rule_unpack_bf16x4(uint2 pk) {
  const float2 a01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&pk.x));
  const float2 a23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&pk.y));
  return make_float4(a01.x, a01.y, a23.x, a23.y);
}

template <class Rule>
__global__ void __launch_bounds__(256)
rule_apply_outbox_kernel(const __grid_constant__ SsbTableView v, const __grid_constant__ SsbInboxView src, int self,
                         const __grid_constant__ rule::Hyper h) {
  const int y = blockIdx.y;
  if (y == self || !src.rows[y]) return;
  const int lane = threadIdx.x & 31;
  const int D = v.dim;
  const uint32_t* slots = src.slots[y];
  const __nv_bfloat16* rows = reinterpret_cast<const __nv_bfloat16*>(src.rows[y]);
  uint32_t ne = ld_peer_u32(src.count[y]);
  if (ne > src.capacity) ne = src.capacity;
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  for (uint32_t w0 = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 32; w0 < ne; w0 += warps * 32) {
    const uint32_t nr = min(32u, ne - w0);
    const uint32_t sl = (uint32_t)lane < nr ? ld_peer_u32(slots + w0 + lane) : SSB_INVALID_SLOT;
    if (D == 128) {
      // rows in groups of kAhead: the next group's loads (kAhead x 256 bytes per warp) are in flight over NVLink while
      // the current group is applied (each row's update waits on an atomic's round trip to local HBM)
      constexpr uint32_t kAhead = 4;
      uint2 cur[kAhead], nxt[kAhead];
#pragma unroll
      for (uint32_t i = 0; i < kAhead; ++i) cur[i] = i < nr ? ld_peer_u2(rows + (size_t)(w0 + i) * D + lane * 4) : make_uint2(0u, 0u);
      for (uint32_t k0 = 0; k0 < nr; k0 += kAhead) {
#pragma unroll
        for (uint32_t i = 0; i < kAhead; ++i)
          nxt[i] = (k0 + kAhead + i < nr) ? ld_peer_u2(rows + (size_t)(w0 + k0 + kAhead + i) * D + lane * 4) : make_uint2(0u, 0u);
#pragma unroll
        for (uint32_t i = 0; i < kAhead; ++i) {
          if (k0 + i >= nr) break;
          const uint32_t g = __shfl_sync(0xffffffffu, sl, (int)(k0 + i));
          if (g < v.capacity) {
            const float4 gv = rule_unpack_bf16x4(cur[i]);
            rule_apply_row<Rule, false>(v, g, lane, h, [gv](int) { return gv; }, nullptr);
          }
        }
#pragma unroll
        for (uint32_t i = 0; i < kAhead; ++i) cur[i] = nxt[i];
      }
    } else {
      for (uint32_t k = 0; k < nr; ++k) {
        const uint32_t g = __shfl_sync(0xffffffffu, sl, (int)k);
        if (g >= v.capacity) continue;
        const __nv_bfloat16* gr = rows + (size_t)(w0 + k) * D;
        rule_apply_row<Rule, false>(v, g, lane, h, [gr](int j) { return rule_unpack_bf16x4(ld_peer_u2(gr + j)); }, nullptr);
      }
    }
  }
}

// mirror[slot] = bf16(row[slot]) for every slot the senders listed for this owner (after rule_apply_outbox_kernel).
template <int kUnused = 0>
__global__ void __launch_bounds__(256)
rule_mirror_refresh_outbox_kernel(const __grid_constant__ SsbTableView v, const __grid_constant__ SsbInboxView src, int self) {
  const int y = blockIdx.y;
  if (y == self || !src.rows[y]) return;
  const int lane = threadIdx.x & 31;
  uint32_t ne = ld_peer_u32(src.count[y]);
  if (ne > src.capacity) ne = src.capacity;
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  for (uint32_t w0 = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 32; w0 < ne; w0 += warps * 32) {
    const uint32_t nr = min(32u, ne - w0);
    const uint32_t sl = (uint32_t)lane < nr ? ld_peer_u32(src.slots[y] + w0 + lane) : SSB_INVALID_SLOT;
    for (uint32_t k = 0; k < nr; ++k) {
      const uint32_t g = __shfl_sync(0xffffffffu, sl, (int)k);
      if (g < v.capacity) rule_refresh_row(v, g, lane);
    }
  }
}

// ---- host-side launchers -----------------------------------------------------------------------------------------
inline int rule_grid(int64_t rows, int max_blocks_per_sm = 8) {
  int64_t b = (rows + 7) / 8;                  // 8 warps per block, one row per warp per trip
  const int64_t cap = 148LL * max_blocks_per_sm;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}
inline bool rule_view_has_shadow(const SsbTableView* v) {
  for (int i = 0; i < v->num_shards; ++i) if (v->shadow[i]) return true;
  return false;
}
inline int rule_check(const SsbTableView* v, int state_words) {
  if (!v || v->dim <= 0 || v->num_shards <= 0 || v->num_shards > SSB_MAX_SHARDS) return -1;
  if (state_words > 0) {
    if ((v->state_words > 0 ? v->state_words : 1) < state_words) return -3;      // table allocated with too little state
    for (int i = 0; i < v->num_shards; ++i) if (!v->state[i]) return -3;
  }
  return 0;
}

template <class Rule>
int launch_rule_push(const SsbTableView* view, const uint32_t* gslot, float* grads, int64_t n, const uint32_t* n_dev,
                     const rule::Hyper* hyper, int zero_grads, int only_shard, void* stream, int grid = 0) {
  if (int e = rule_check(view, Rule::kStateWords)) return e;
  if (!hyper) return -1;
  if (n <= 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  if (grid <= 0) grid = rule_grid(n);
  if (rule_view_has_shadow(view))
    rule_push_kernel<Rule, true><<<grid, 256, 0, s>>>(*view, gslot, grads, n, n_dev, *hyper, zero_grads, only_shard);
  else
    rule_push_kernel<Rule, false><<<grid, 256, 0, s>>>(*view, gslot, grads, n, n_dev, *hyper, zero_grads, only_shard);
  return (int)cudaGetLastError();
}

// Segmented bf16 apply (+ mirror refresh when the shards keep a mirror). segments = 1, seg_stride = 0: one flat list.
template <class Rule>
int launch_rule_push_bf16_seg(const SsbTableView* view, const uint32_t* gslot, const void* grads_bf16, int64_t seg_rows,
                              const uint32_t* n_dev, int segments, int64_t seg_stride, const rule::Hyper* hyper, void* stream,
                              int gx = 0, const MirrorTrack* track = nullptr) {
  if (int e = rule_check(view, Rule::kStateWords)) return e;
  if (!hyper || (view->dim & 3) || segments <= 0 || (segments > 1 && !n_dev)) return -1;
  if (seg_rows <= 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  if (gx <= 0) gx = rule_grid(seg_rows) / segments;
  if (gx < 1) gx = 1;
  dim3 grid((unsigned)gx, (unsigned)segments);
  const __nv_bfloat16* gb = reinterpret_cast<const __nv_bfloat16*>(grads_bf16);
  const bool shadow = rule_view_has_shadow(view);
  if (shadow && track && track->seen && track->tag != 0 && view->num_shards == 1) {
    rule_push_bf16_kernel<Rule, true><<<grid, 256, 0, s>>>(*view, gslot, gb, seg_rows, n_dev, *hyper, seg_stride, *track);
    rule_mirror_refresh_tracked_kernel<0><<<grid, 256, 0, s>>>(*view, gslot, seg_rows, n_dev, seg_stride, *track);
    return (int)cudaGetLastError();
  }
  rule_push_bf16_kernel<Rule, false><<<grid, 256, 0, s>>>(*view, gslot, gb, seg_rows, n_dev, *hyper, seg_stride, MirrorTrack{});
  if (shadow) rule_mirror_refresh_kernel<0><<<grid, 256, 0, s>>>(*view, gslot, seg_rows, n_dev, seg_stride, 1);
  return (int)cudaGetLastError();
}

// Apply (and refresh the mirror) straight from the senders' outboxes. src: per sender y the sub-region it filled for
// this owner; senders = world size; max_rows = rows a sub-region can hold (sizes the grid).
template <class Rule>
int launch_rule_apply_outbox(const SsbTableView* view, const SsbInboxView* src, int self, int senders, int64_t max_rows,
                             const rule::Hyper* hyper, void* stream) {
  if (int e = rule_check(view, Rule::kStateWords)) return e;
  if (!hyper || !src || (view->dim & 3) || view->num_shards != 1 || senders < 1 || senders > SSB_MAX_SHARDS) return -1;
  if (senders == 1) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  int gx = (int)((max_rows + 255) / 256);                 // a block trip covers 8 warps x 32 rows
  const int cap = 148 * 8 / (senders > 2 ? 2 : 1);
  if (gx > cap) gx = cap;
  if (gx < 1) gx = 1;
  dim3 grid((unsigned)gx, (unsigned)senders);
  rule_apply_outbox_kernel<Rule><<<grid, 256, 0, s>>>(*view, *src, self, *hyper);
  if (rule_view_has_shadow(view)) rule_mirror_refresh_outbox_kernel<0><<<grid, 256, 0, s>>>(*view, *src, self);
  return (int)cudaGetLastError();
}

}  // namespace ssb

// extern "C" entry points of a rule, found by name from Python (ops/rules.py):
//   ssb_rule_<name>_info(int out[4])            -> {kStateWords, kInit, 0, 0}
//   ssb_rule_<name>_push(...)                   -> fp32 gradient rows, local or direct-remote
//   ssb_rule_<name>_push_bf16_seg(...)          -> owner-side apply of bf16 inbox segments
//   ssb_rule_<name>_push_bf16_seg_tracked(...)  -> the same with second-writer tracking (MirrorTrack)
//   ssb_rule_<name>_apply_outbox(...)           -> owner-side apply that READS the rows from the senders' outboxes
#define SSB200_EXPORT_RULE(name, Rule)                                                                                      \
  extern "C" int ssb_rule_##name##_info(int32_t* out) {                                                                     \
    out[0] = Rule::kStateWords; out[1] = Rule::kInit; out[2] = 0; out[3] = 0;                                              \
    return 0;                                                                                                               \
  }                                                                                                                         \
  extern "C" int ssb_rule_##name##_push(const SsbTableView* view, const uint32_t* gslot, float* grads, int64_t n,           \
                                        const uint32_t* n_dev, const ssb::rule::Hyper* hyper, int32_t zero_grads,          \
                                        int32_t only_shard, void* stream) {                                                \
    return ssb::launch_rule_push<Rule>(view, gslot, grads, n, n_dev, hyper, zero_grads, only_shard, stream);                \
  }                                                                                                                         \
  extern "C" int ssb_rule_##name##_push_bf16_seg(const SsbTableView* view, const uint32_t* gslot, const void* grads_bf16,   \
                                                 int64_t seg_rows, const uint32_t* n_dev, int32_t segments,                \
                                                 int64_t seg_stride, const ssb::rule::Hyper* hyper, void* stream) {        \
    return ssb::launch_rule_push_bf16_seg<Rule>(view, gslot, grads_bf16, seg_rows, n_dev, segments, seg_stride, hyper,      \
                                                stream);                                                                   \
  }                                                                                                                         \
  extern "C" int ssb_rule_##name##_apply_outbox(const SsbTableView* view, const SsbInboxView* src, int32_t self,            \
                                                int32_t senders, int64_t max_rows, const ssb::rule::Hyper* hyper,          \
                                                void* stream) {                                                            \
    return ssb::launch_rule_apply_outbox<Rule>(view, src, self, senders, max_rows, hyper, stream);                          \
  }                                                                                                                         \
  extern "C" int ssb_rule_##name##_push_bf16_seg_tracked(const SsbTableView* view, const uint32_t* gslot,                   \
                                                         const void* grads_bf16, int64_t seg_rows, const uint32_t* n_dev,   \
                                                         int32_t segments, int64_t seg_stride,                              \
                                                         const ssb::rule::Hyper* hyper, uint32_t* seen, uint32_t tag,       \
                                                         void* stream) {                                                    \
    const ssb::MirrorTrack tr{seen, tag};                                                                                   \
    return ssb::launch_rule_push_bf16_seg<Rule>(view, gslot, grads_bf16, seg_rows, n_dev, segments, seg_stride, hyper,      \
                                                stream, 0, &tr);                                                            \
  }
