// sgns_tc.cu — K5: the word2vec SGNS step on 5th-generation tensor cores (tcgen05 / TMEM / TMA).
//
// The reference's worker computes, per (centre, context, negatives) sample, dot products on
// `Vec` (utils/vec1.h:97-104) + sigmoid, and accumulates into cache.grads() between
// pull_with_barrier and push_with_barrier (core/framework/SwiftWorker.h:88-113, SURVEY §3.4
// hot loop I).  Its app is absent from the snapshot, so the math is standard SGNS written fresh.
//
// B200 formulation (not a translation): a tile is 128 CONSECUTIVE centre tokens. Their contexts
// all lie in the 144-token window [t0-8, t0+136), so the tile's positive logits are a band of
// one dense GEMM and the whole forward + backward is three tensor-core GEMMs:
//
//   S    [128 x 160] = In[128 x D] . ON[160 x D]^T        ON = 144 context rows ++ 16 negatives
//   G    [128 x 160] = dL/dS  (band mask, window shrink, sigmoid — epilogue on TMEM)
//   dIn  [128 x D  ] = G . ON                              (A = G K-major,   B = ON MN-major)
//   dON^T[D   x 160] = In^T . G                            (A = In MN-major, B = G  MN-major)
//
// Rows are gathered from the worker's bf16 cache (written by the pull kernel next to the fp32
// table) by row index straight into 128B-swizzled shared memory (TMA tile::gather4, or a plain
// vector-load path used for bring-up and as fallback), feed `tcgen05.mma.kind::f16` (bf16 in,
// fp32 accumulate), accumulators live in TMEM (448 of 512 columns), and gradients leave in fp32
// through `red` into the cache's grad buffer: 288 row reads + 288 row reductions per 128 tokens
// instead of ~1500 each for per-pair processing.  bf16 (not tf32) because every operand is
// used both K-major and MN-major, and 32-bit MN-major operands need a different swizzle
// (SWIZZLE_128B_BASE32B) than their K-major use — 16-bit operands take plain SWIZZLE_128B both ways.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cstring>
#include "common.cuh"

namespace ssb {
namespace tc {

constexpr int kD = 128;          // embedding dim (floats)
constexpr int kM = 128;          // centre tokens per tile
constexpr int kHalo = 8;         // max window
constexpr int kCtx = kM + 2 * kHalo;   // 144 context rows
constexpr int kNeg = 16;
constexpr int kN = kCtx + kNeg;  // 160
constexpr int kThreads = 256;

// bf16 operands: a "column block" is 64 elements = 128 bytes wide.
constexpr uint32_t kInBlk = kM * 128;   // bytes of one [128 rows x 128 B] column block of In / G
constexpr uint32_t kOnBlk = kN * 128;   // bytes of one [160 rows x 128 B] column block of ON
constexpr int kInBlks = kD / 64;        // 2
constexpr int kGBlks = (kN + 63) / 64;  // 3 (the last block is half used)
constexpr uint32_t kOffIn = 0;
constexpr uint32_t kOffOn = kOffIn + kInBlks * kInBlk;    // 32768
constexpr uint32_t kOffG = kOffOn + kInBlks * kOnBlk;     // 73728
constexpr uint32_t kOffMisc = kOffG + kGBlks * kInBlk;    // 122880
constexpr uint32_t kSmemBytes = kOffMisc + 2048;

constexpr uint32_t kTmemS = 0, kTmemDIn = 160, kTmemDOn = 288, kTmemCols = 512;

struct Misc {
  uint32_t in_row[kM];
  uint32_t on_row[kN];
  uint32_t col_used[kN];
  unsigned long long tma_bar;
  unsigned long long mma_bar;
  uint32_t tmem_base;
  float loss[8];
  float pairs[8];
};
static_assert(sizeof(Misc) <= 2048, "misc region too large");

// ---- PTX wrappers -----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// Bounded wait: a lost arrival traps (reported as a launch failure) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
  uint32_t done = 0;
  for (uint32_t it = 0; it < (1u << 22); ++it) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, 1000000;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (done) return;
  }
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
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tma_gather4(uint32_t dst_smem, const CUtensorMap* map, unsigned long long* bar,
                                            int col, uint32_t r0, uint32_t r1, uint32_t r2, uint32_t r3) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes"
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

// ---- the kernel -------------------------------------------------------------------------------
// mode 0: rows loaded with vector loads + swizzled st.shared; mode 1: TMA tile::gather4.
__global__ void __launch_bounds__(kThreads, 1)
sgns_block_tc_kernel(const __nv_bfloat16* __restrict__ params, float* __restrict__ grads,
                     const uint32_t* __restrict__ inv_in, const uint32_t* __restrict__ inv_out,
                     int64_t n_tok, int sent_len, int window, uint64_t seed, uint64_t offset,
                     const uint32_t* __restrict__ neg_idx, float neg_k, float* __restrict__ loss_out,
                     const __grid_constant__ CUtensorMap tmap, int mode, float* __restrict__ debug) {
  extern __shared__ __align__(1024) uint8_t smem[];
  Misc* misc = reinterpret_cast<Misc*>(smem + kOffMisc);
  const uint32_t sbase = smem_u32(smem);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t n_tiles = (n_tok + kM - 1) / kM;

  if (tid == 0) {
    mbar_init(&misc->tma_bar, 1);
    mbar_init(&misc->mma_bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(&misc->tmem_base, kTmemCols);
  // G starts all-zero; each tile rewrites only the band windows and the negative columns
  for (uint32_t i = tid; i < kGBlks * kInBlk / 16; i += kThreads)
    reinterpret_cast<float4*>(smem + kOffG)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = misc->tmem_base;

  uint32_t tma_phase = 0, mma_phase = 0;
  float loss_acc = 0.f, pair_acc = 0.f;

  for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t t0 = tile * kM;
    // ---- 1. row indices -------------------------------------------------------------------
    if (tid < kM) {
      int64_t g = t0 + tid;
      misc->in_row[tid] = (g < n_tok) ? inv_in[g] : SSB_INVALID_SLOT;
    }
    if (tid < kN) {
      uint32_t r;
      if (tid < kCtx) {
        int64_t g = t0 - kHalo + tid;
        r = (g >= 0 && g < n_tok) ? inv_out[g] : SSB_INVALID_SLOT;
      } else {
        r = neg_idx[tile * kNeg + (tid - kCtx)];
      }
      misc->on_row[tid] = r;
      misc->col_used[tid] = 0;
    }
    __syncthreads();

    // ---- 2. gather rows into swizzled smem -------------------------------------------------
    if (mode == 1) {
      if (tid == 0) {
        mbar_expect_tx(&misc->tma_bar, (kM + kN) * kD * 2);
        for (int r = 0; r < kM; r += 4) {
          uint32_t i0 = misc->in_row[r], i1 = misc->in_row[r + 1], i2 = misc->in_row[r + 2], i3 = misc->in_row[r + 3];
          i0 = i0 == SSB_INVALID_SLOT ? 0 : i0; i1 = i1 == SSB_INVALID_SLOT ? 0 : i1;
          i2 = i2 == SSB_INVALID_SLOT ? 0 : i2; i3 = i3 == SSB_INVALID_SLOT ? 0 : i3;
#pragma unroll
          for (int cb = 0; cb < kInBlks; ++cb)
            tma_gather4(sbase + kOffIn + cb * kInBlk + r * 128, &tmap, &misc->tma_bar, cb * 64, i0, i1, i2, i3);
        }
        for (int r = 0; r < kN; r += 4) {
          uint32_t i0 = misc->on_row[r], i1 = misc->on_row[r + 1], i2 = misc->on_row[r + 2], i3 = misc->on_row[r + 3];
          i0 = i0 == SSB_INVALID_SLOT ? 0 : i0; i1 = i1 == SSB_INVALID_SLOT ? 0 : i1;
          i2 = i2 == SSB_INVALID_SLOT ? 0 : i2; i3 = i3 == SSB_INVALID_SLOT ? 0 : i3;
#pragma unroll
          for (int cb = 0; cb < kInBlks; ++cb)
            tma_gather4(sbase + kOffOn + cb * kOnBlk + r * 128, &tmap, &misc->tma_bar, cb * 64, i0, i1, i2, i3);
        }
      }
      mbar_wait(&misc->tma_bar, tma_phase);
      tma_phase ^= 1;
    } else {
      // half a warp per row: 16 lanes x 16 B = 256 B of bf16
      for (int r = warp * 2 + (lane >> 4); r < kM + kN; r += kThreads / 16) {
        const bool is_in = r < kM;
        const uint32_t rr = is_in ? r : r - kM;
        const int l16 = lane & 15;
        uint32_t row = is_in ? misc->in_row[rr] : misc->on_row[rr];
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (row != SSB_INVALID_SLOT) v = *reinterpret_cast<const uint4*>(params + (size_t)row * kD + l16 * 8);
        uint32_t off = (is_in ? kOffIn : kOffOn) + swz_off(rr, l16 * 8, is_in ? kInBlk : kOnBlk);
        *reinterpret_cast<uint4*>(smem + off) = v;
      }
      fence_proxy_async();
    }
    __syncthreads();

    // ---- 3. S = In . ON^T ------------------------------------------------------------------
    if (tid == 0) {
      tc_fence_after();
      constexpr uint32_t idesc = make_idesc(kM, kN, 0, 0);
#pragma unroll
      for (int s = 0; s < kD / 16; ++s) {
        uint64_t a = make_desc(sbase + kOffIn + (s >> 2) * kInBlk + (s & 3) * 32, 16, 1024);
        uint64_t b = make_desc(sbase + kOffOn + (s >> 2) * kOnBlk + (s & 3) * 32, 16, 1024);
        umma_bf16(tmem + kTmemS, a, b, idesc, s > 0);
      }
      umma_commit(&misc->mma_bar);
    }
    mbar_wait(&misc->mma_bar, mma_phase);
    mma_phase ^= 1;
    tc_fence_after();

    // ---- 4. epilogue 1: logits -> loss and dL/dS (band + negatives) --------------------------
    if (warp < 4) {
      const int i = tid;                       // tile row = TMEM lane
      const int64_t gi = t0 + i;
      const uint32_t my_in = misc->in_row[i];
      const bool row_ok = my_in != SSB_INVALID_SLOT;
      int eff = 0;
      if (row_ok) {
        uint32_t r4[4];
        Philox::gen(seed, offset + (uint64_t)gi, r4);
        eff = window - (int)(r4[0] % (uint32_t)window);
      }
      const int64_t sent = gi / sent_len;
      const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
      float m_i = 0.f;
      // band window: columns [32*warp, 32*warp + 48)
#pragma unroll
      for (int blk = 0; blk < 3; ++blk) {
        float v[16];
        const int c0 = warp * 32 + blk * 16;
        tmem_ld16(lane_base + kTmemS + c0, v);
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int c = c0 + j;
          const int rel = c - (i + kHalo);
          const int64_t tok = t0 - kHalo + c;
          bool ok = row_ok && rel != 0 && rel >= -eff && rel <= eff && tok >= 0 && tok < n_tok &&
                    (tok / sent_len) == sent && misc->on_row[c] != SSB_INVALID_SLOT;
          float g = 0.f;
          if (ok) {
            g = fast_sigmoid(v[j]) - 1.0f;
            loss_acc += neg_log_sigmoid(v[j]);
            m_i += 1.f;
            misc->col_used[c] = 1;
          }
          *reinterpret_cast<__nv_bfloat16*>(smem + kOffG + swz_off(i, c, kInBlk)) = __float2bfloat16_rn(g);
        }
      }
      pair_acc += m_i;
      // negatives: weight m_i * K / 16 each (K negatives per pair in expectation)
      {
        float v[16];
        tmem_ld16(lane_base + kTmemS + kCtx, v);
        const float w = m_i * neg_k * (1.0f / kNeg);
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const uint32_t nr = misc->on_row[kCtx + j];
          float g = 0.f;
          if (m_i > 0.f && nr != SSB_INVALID_SLOT) {
            g = w * fast_sigmoid(v[j]);
            loss_acc += w * neg_log_sigmoid(-v[j]);
            misc->col_used[kCtx + j] = 1;
          }
          *reinterpret_cast<__nv_bfloat16*>(smem + kOffG + swz_off(i, kCtx + j, kInBlk)) = __float2bfloat16_rn(g);
        }
      }
      fence_proxy_async();
      tc_fence_before();
    }
    __syncthreads();
    if (debug && tile == 0)
      for (uint32_t i = tid; i < kGBlks * kInBlk / 2; i += kThreads)
        debug[i] = __bfloat162float(reinterpret_cast<__nv_bfloat16*>(smem + kOffG)[i]);

    // ---- 5. dIn = G . ON ;  dON^T = In^T . G ------------------------------------------------
    if (tid == 0) {
      tc_fence_after();
      constexpr uint32_t idesc2 = make_idesc(kM, kD, 0, 1);
#pragma unroll
      for (int s = 0; s < kN / 16; ++s) {     // K = 160 context/negative rows, 16 per step
        uint64_t a = make_desc(sbase + kOffG + (s >> 2) * kInBlk + (s & 3) * 32, 16, 1024);
        uint64_t b = make_desc(sbase + kOffOn + s * 2048, kOnBlk, 1024);
        umma_bf16(tmem + kTmemDIn, a, b, idesc2, s > 0);
      }
      constexpr uint32_t idesc3 = make_idesc(kD, kN, 1, 1);
#pragma unroll
      for (int s = 0; s < kM / 16; ++s) {     // K = 128 centre tokens, 16 per step
        uint64_t a = make_desc(sbase + kOffIn + s * 2048, kInBlk, 1024);
        uint64_t b = make_desc(sbase + kOffG + s * 2048, kInBlk, 1024);
        umma_bf16(tmem + kTmemDOn, a, b, idesc3, s > 0);
      }
      umma_commit(&misc->mma_bar);
    }
    mbar_wait(&misc->mma_bar, mma_phase);
    mma_phase ^= 1;
    tc_fence_after();

    // ---- 6. epilogue 2: gradients out through vector red -------------------------------------
    if (warp < 4) {
      const int i = tid;
      const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
      const uint32_t my_in = misc->in_row[i];
      if (debug && tile == 0) {
        float* d1 = debug + kGBlks * kInBlk / 2;
        for (int blk = 0; blk < 28; ++blk) {   // all 448 used columns: S | dIn | dON^T
          float v[16];
          tmem_ld16(lane_base + blk * 16, v);
          for (int j = 0; j < 16; ++j) d1[(size_t)i * 448 + blk * 16 + j] = v[j];
        }
      }
      // dIn row i -> grads[in_row[i]]
#pragma unroll
      for (int blk = 0; blk < kD / 16; ++blk) {
        float v[16];
        tmem_ld16(lane_base + kTmemDIn + blk * 16, v);
        if (my_in != SSB_INVALID_SLOT) {
          float* dst = grads + (size_t)my_in * kD + blk * 16;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float4 x = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
            if (x.x != 0.f || x.y != 0.f || x.z != 0.f || x.w != 0.f) red_add_f4_gpu(dst + 4 * q, x);
          }
        }
      }
      // dON^T: lane = d, columns = context/negative row j -> grads[on_row[j]][d]
#pragma unroll
      for (int blk = 0; blk < kN / 16; ++blk) {
        float v[16];
        tmem_ld16(lane_base + kTmemDOn + blk * 16, v);
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int c = blk * 16 + j;
          const uint32_t row = misc->on_row[c];
          if (misc->col_used[c] && row != SSB_INVALID_SLOT) atomicAdd(grads + (size_t)row * kD + i, v[j]);
        }
      }
      tc_fence_before();
    }
    __syncthreads();
  }

  // ---- teardown -------------------------------------------------------------------------------
  if (warp < 4) {
    loss_acc = warp_sum(loss_acc);
    pair_acc = warp_sum(pair_acc);
    if (lane == 0) { misc->loss[warp] = loss_acc; misc->pairs[warp] = pair_acc; }
  }
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    float l = misc->loss[0] + misc->loss[1] + misc->loss[2] + misc->loss[3];
    float p = misc->pairs[0] + misc->pairs[1] + misc->pairs[2] + misc->pairs[3];
    if (l != 0.f) atomicAdd(loss_out, l);
    if (p != 0.f) atomicAdd(loss_out + 1, p);
  }
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, kTmemCols);
  }
}

// ---- host: tensor map over the cache ------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

static int make_row_tmap(CUtensorMap* map, const void* base, uint64_t n_rows, uint32_t box_rows) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return (int)cudaErrorSymbolNotFound;
  cuuint64_t dims[2] = {(cuuint64_t)kD, (cuuint64_t)n_rows};
  cuuint64_t strides[1] = {(cuuint64_t)kD * 2};
  cuuint32_t box[2] = {64, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : 100000 + (int)r;
}

}  // namespace tc
}  // namespace ssb

extern "C" {

// Block-formulation SGNS step on tensor cores. params: bf16 worker cache [n_rows, 128];
// grads: fp32 [n_rows, 128].
// inv_in/inv_out [n_tok]: cache row of each token's input / output vector; neg_idx
// [ceil(n_tok/128) * 16]: cache rows of each tile's negatives. window <= 8.
// mode: 0 = vector-load gather, 1 = TMA tile::gather4.
int ssb_sgns_block_tc(const void* params, float* grads, int64_t n_rows, int32_t dim, const uint32_t* inv_in,
                      const uint32_t* inv_out, int64_t n_tok, int32_t sent_len, int32_t window, uint64_t seed,
                      uint64_t offset, const uint32_t* neg_idx, float neg_k, float* loss_out, int32_t mode,
                      int32_t box_rows, float* debug, void* stream) {
  using namespace ssb::tc;
  if (n_tok <= 0) return 0;
  if (dim != kD || window < 1 || window > kHalo || sent_len < 1) return (int)cudaErrorInvalidValue;
  CUtensorMap tmap;
  memset(&tmap, 0, sizeof(tmap));
  if (mode == 1) {
    int e = make_row_tmap(&tmap, params, (uint64_t)n_rows, box_rows > 0 ? (uint32_t)box_rows : 1u);
    if (e) return e;
  }
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(sgns_block_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (e != cudaSuccess) return (int)e;
    attr_set = true;
  }
  int64_t tiles = (n_tok + kM - 1) / kM;
  int blocks = (int)(tiles < ssb::kNumSMs ? tiles : ssb::kNumSMs);
  sgns_block_tc_kernel<<<blocks, kThreads, kSmemBytes, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)params, grads, inv_in, inv_out, n_tok, sent_len, window, seed, offset, neg_idx, neg_k,
      loss_out, tmap, mode, debug);
  SSB_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
