// tc_common.cuh — tcgen05 / TMEM / TMA / mbarrier wrappers and the tile geometry shared by the
// tensor-core SGNS kernels (sgns_tc.cu: pipelined, the default; sgns_tc_serial.cu: the round-1
// serial form kept for A/B measurements).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cstring>
#include "common.cuh"

namespace ssb {
namespace tc {

constexpr int kD = 128;          // embedding dim
constexpr int kM = 128;          // centre tokens per tile
constexpr int kHalo = 8;         // max window
constexpr int kCtx = kM + 2 * kHalo;   // 144 context rows
constexpr int kNeg = 16;
constexpr int kN = kCtx + kNeg;  // 160

// bf16 operands: a "column block" is 64 elements = 128 bytes wide.
constexpr uint32_t kInBlk = kM * 128;   // bytes of one [128 rows x 128 B] column block of In / G
constexpr uint32_t kOnBlk = kN * 128;   // bytes of one [160 rows x 128 B] column block of ON
constexpr int kInBlks = kD / 64;        // 2
constexpr uint32_t kInBytes = kInBlks * kInBlk;   // 32768
constexpr uint32_t kOnBytes = kInBlks * kOnBlk;   // 40960

// ---- PTX wrappers -----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// One try_wait with a suspend-time hint: the thread sleeps in hardware until the phase completes or about
// `hint_ns` have passed (no issue slots are spent meanwhile; the wake-up on completion is immediate).
__device__ __forceinline__ bool mbar_try(unsigned long long* bar, uint32_t parity, uint32_t hint_ns) {
  uint32_t done;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(done)
      : "r"(smem_u32(bar)), "r"(parity), "r"(hint_ns)
      : "memory");
  return done != 0;
}
// Bounded wait: a lost arrival traps (reported as a launch failure) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
#pragma unroll 1
  for (uint32_t it = 0; it < (1u << 22); ++it)
    if (mbar_try(bar, parity, 1000u)) return;
  __trap();
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(unsigned long long* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// TMEM -> registers, 32 lanes x 32 bit, N consecutive columns per thread. The *_issue forms only
// start the load; tmem_wait_ld() must come before the registers are read.
__device__ __forceinline__ void tmem_ld16_issue(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld8_issue(uint32_t taddr, uint32_t (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// The wait, tied to the registers it guards: the compiler sees them as (re)defined here, so no use of
// them can be scheduled above the wait even when other work sits between the load and the wait.
__device__ __forceinline__ void tmem_wait_ld16(uint32_t (&r)[16]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                 "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
               :
               : "memory");
}
__device__ __forceinline__ void tmem_wait_ld8(uint32_t (&r)[8]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7])
               :
               : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  tmem_ld16_issue(taddr, r);
  tmem_wait_ld16(r);
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tma_gather4(uint32_t dst_smem, const CUtensorMap* map, unsigned long long* bar,
                                            int col, uint32_t r0, uint32_t r1, uint32_t r2, uint32_t r3) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cta.global.tile::gather4.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
      ::"r"(dst_smem), "l"((uint64_t)map), "r"(smem_u32(bar)), "r"(col), "r"(r0), "r"(r1), "r"(r2), "r"(r3)
      : "memory");
}

// ---- descriptors ------------------------------------------------------------------------------
// Shared-memory matrix descriptor, 128-byte swizzle (layout_type 2), version 1 (Blackwell).
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// Instruction descriptor: D=f32, A=B=bf16, dense.
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, int a_mn_major, int b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// byte offset of element (row r, bf16 column c) inside an [R rows] matrix stored as column blocks
// of 64 bf16, each block [R x 128 B] with the 128-byte swizzle (16-byte chunk index ^= row % 8).
__device__ __forceinline__ uint32_t swz_off(uint32_t r, uint32_t c, uint32_t blk_bytes) {
  uint32_t cb = c >> 6, ch = (c & 63) >> 3, e = c & 7;
  return cb * blk_bytes + r * 128 + ((ch ^ (r & 7)) << 4) + (e << 1);
}

// Direct ("fused pull") mode: the TMA gathers operand rows straight from the peer-mapped bf16 mirror of
// the sharded table — local HBM or a peer's over NVLink — instead of from the worker's cache. All
// shards sit side by side in one virtual range (parallel/peer.py), so ONE tensor map with a 256-byte
// row pitch describes the whole distributed table: row = shard * shard_rows + mirror_row_off + slot.
struct DirectMap {
  const uint32_t* gslot;      // cache row -> global slot (shard * capacity + slot); nullptr = cache mode
  uint32_t capacity;          // slots per shard
  uint32_t shard_rows;        // 256-byte rows per shard segment
  uint32_t mirror_row_off;    // first mirror row inside a segment
};

// one exp per logit: t = 1 + e^-x ; sigmoid = 1/t ; -log sigmoid(x) = log t ; softplus(x) = x + log t.
// Three MUFU ops (ex2 / rcp / lg2, approximate forms: ~2 ulp, far inside the bf16 rounding G gets anyway).
__device__ __forceinline__ void sig_logt(float x, float& sig, float& logt) {
  x = fminf(fmaxf(x, -30.f), 30.f);
  float e, l;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-1.4426950408889634f * x));
  const float t = 1.0f + e;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(sig) : "f"(t));
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(t));
  logt = 0.6931471805599453f * l;
}

struct TcArgs {
  const __nv_bfloat16* params;
  float* grads;
  const uint32_t* inv_in;
  const uint32_t* inv_out;
  int n_tok, sent_len, window;
  uint64_t seed, offset;
  const uint32_t* neg_idx;
  float neg_k;
  float* loss_out;
  int mode;
  float* debug;
  DirectMap dm;
};

}  // namespace tc
}  // namespace ssb
