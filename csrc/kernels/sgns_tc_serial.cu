// sgns_tc_serial.cu — the round-1 form of the tensor-core SGNS tile kernel: one serial chain per tile
// (S-MMA -> epilogue 1 -> gradient MMAs -> 3-pass staged epilogue 2) run by all 16 consumer warps.
// Superseded by the pipelined kernel in sgns_tc.cu; kept selectable (mode bit 8) so the two can be
// timed against each other on the same inputs (benchmarks/tc_kernel_bench.py). Same maths, same
// oracle tests. See sgns_tc.cu for the formulation and the reference citations.
#include "tc_common.cuh"

namespace ssb {
namespace tc {
namespace serial {

constexpr int kConsumerWarps = 16;
constexpr int kConsumers = kConsumerWarps * 32;   // 512 epilogue / MMA-issue threads
constexpr int kThreads = kConsumers + 32;         // + 1 producer warp (TMA)
constexpr int kGBlks = 4;               // 160 columns used; padded to 256 so the second dON MMA stays in bounds
constexpr uint32_t kOffIn = 0;                               // 2 stages
constexpr uint32_t kOffOn = kOffIn + 2 * kInBytes;           // 65536, 2 stages
constexpr uint32_t kOffG = kOffOn + 2 * kOnBytes;            // 147456
constexpr uint32_t kOffMisc = kOffG + kGBlks * kInBlk;       // 212992
constexpr uint32_t kSmemBytes = kOffMisc + 8192;

// TMEM columns: S (160) | dIn (128) | dON rows 0..127 (128); dON rows 128..255 aliases S.
constexpr uint32_t kTmemS = 0, kTmemDIn = 160, kTmemDOnA = 288, kTmemDOnB = 0, kTmemCols = 512;

struct Misc {
  uint32_t in_row[2][kM];
  uint32_t on_row[2][kN];
  uint32_t in_tab[2][kM];       // direct mode: row of the (peer-mapped) bf16 table mirror to gather from
  uint32_t on_tab[2][kN];
  uint32_t out_rows[kM + kN];   // gradient destination rows of the current tile (INVALID = skip)
  uint8_t lead_in[2][kM];       // first position in the tile holding the same row (tile-local dedup)
  uint8_t lead_on[2][kN];
  unsigned long long full_bar[2];
  unsigned long long empty_bar[2];
  unsigned long long lead_bar[2];
  uint32_t ctx_has_invalid[2];  // a context row inside the batch is INVALID (padding key): count m_i the slow way
  unsigned long long mma_bar;
  uint32_t tmem_base;
  float loss[kConsumerWarps];
  float pairs[kConsumerWarps];
};
static_assert(sizeof(Misc) <= 8192, "misc region too large");

// ---- the kernel -------------------------------------------------------------------------------
// Warp roles: warps 0..15 = consumers (thread 0 issues the MMAs, all 512 threads run the TMEM
// epilogues: warp w owns TMEM lanes 32*(w%4).. and column part w/4), warp 16 = producer
// (row indices + TMA gather4 of the NEXT tile while the current one is computed; 2 stages).
// mode 0: producer copies rows with vector loads + swizzled st.shared; mode 1: TMA tile::gather4.
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory"); }

// kDiag (mode bit 4, experimental): epilogue 1 walks each row's context band by diagonal instead of by
// column block — see the comment at the epilogue.
template <bool kDiag>
__global__ void __maxnreg__(64)
sgns_block_tc_serial_kernel(const __nv_bfloat16* __restrict__ params, float* __restrict__ grads,
                     const uint32_t* __restrict__ inv_in, const uint32_t* __restrict__ inv_out,
                     int n_tok, int sent_len, int window, uint64_t seed, uint64_t offset,
                     const uint32_t* __restrict__ neg_idx, float neg_k, float* __restrict__ loss_out,
                     const __grid_constant__ CUtensorMap tmap, int mode, float* __restrict__ debug, const DirectMap dm) {
  extern __shared__ __align__(1024) uint8_t smem[];
  Misc* misc = reinterpret_cast<Misc*>(smem + kOffMisc);
  const uint32_t sbase = smem_u32(smem);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n_tiles = (n_tok + kM - 1) / kM;

  if (tid == 0) {
    mbar_init(&misc->full_bar[0], 1);
    mbar_init(&misc->full_bar[1], 1);
    mbar_init(&misc->empty_bar[0], 1);
    mbar_init(&misc->empty_bar[1], 1);
    mbar_init(&misc->lead_bar[0], 1);
    mbar_init(&misc->lead_bar[1], 1);
    mbar_init(&misc->mma_bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(&misc->tmem_base, kTmemCols);
  // G starts all-zero; each tile rewrites only the band windows and the negative columns
  for (uint32_t i = tid; i < kGBlks * kInBlk / 16; i += kThreads)
    reinterpret_cast<float4*>(smem + kOffG)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = misc->tmem_base;

  if (warp == kConsumerWarps) {
    // =========================== producer warp ===========================================
    // Per tile: (wait for the stage) -> publish the row indices -> start the gather (TMA) ->
    // prefetch the NEXT tile's indices into registers -> tile-local dedup -> signal lead_bar.
    // The dedup and the index prefetch run while the TMA engine fills the stage.
    uint32_t ri[4], ro[5], ti[4], to[5];
    auto tab_row = [&](uint32_t cache_row) -> uint32_t {
      if (dm.gslot == nullptr || cache_row == SSB_INVALID_SLOT) return cache_row;
      const uint32_t g = __ldg(dm.gslot + cache_row);
      if (g == SSB_INVALID_SLOT) return SSB_INVALID_SLOT;
      const uint32_t sh = g / dm.capacity;
      return sh * dm.shard_rows + dm.mirror_row_off + (g - sh * dm.capacity);
    };
    auto load_idx = [&](int tile) {
      const int t0 = tile * kM;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int g = t0 + lane + 32 * k;
        ri[k] = (tile < n_tiles && g < n_tok) ? __ldg(inv_in + g) : SSB_INVALID_SLOT;
      }
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        const int r = lane + 32 * k;
        uint32_t v = SSB_INVALID_SLOT;
        if (tile < n_tiles) {
          if (r < kCtx) {
            const int g = t0 - kHalo + r;
            if (g >= 0 && g < n_tok) v = __ldg(inv_out + g);
          } else {
            v = __ldg(neg_idx + (size_t)tile * kNeg + (r - kCtx));
          }
        }
        ro[k] = v;
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) ti[k] = tab_row(ri[k]);
#pragma unroll
      for (int k = 0; k < 5; ++k) to[k] = tab_row(ro[k]);
    };
    load_idx(blockIdx.x);
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const int b = it & 1;
      const int t0 = tile * kM;
      if (it >= 2) mbar_wait(&misc->empty_bar[b], ((it >> 1) - 1) & 1);   // stage freed by the MMAs of tile it-2
      bool bad = false;
      if (dm.gslot != nullptr) {
        // Masked-out positions still feed the MMAs (weight 0 in G), so they must gather FINITE data:
        // 0 * NaN = NaN. In the cache, row 0 is always a real row; in the peer-mapped range row 0 is
        // the key array of shard 0 (0xFFFFFFFF = bf16 NaN). Substitute a valid row of this tile.
        uint32_t fb = SSB_INVALID_SLOT;
#pragma unroll
        for (int k = 0; k < 4; ++k) if (ti[k] != SSB_INVALID_SLOT) fb = ti[k];
#pragma unroll
        for (int k = 0; k < 5; ++k) if (to[k] != SSB_INVALID_SLOT) fb = to[k];
        const unsigned m = __ballot_sync(0xffffffffu, fb != SSB_INVALID_SLOT);
        fb = __shfl_sync(0xffffffffu, fb, m ? __ffs(m) - 1 : 0);
        if (m == 0) fb = dm.mirror_row_off;
#pragma unroll
        for (int k = 0; k < 4; ++k) if (ti[k] == SSB_INVALID_SLOT) ti[k] = fb;
#pragma unroll
        for (int k = 0; k < 5; ++k) if (to[k] == SSB_INVALID_SLOT) to[k] = fb;
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) { misc->in_row[b][lane + 32 * k] = ri[k]; misc->in_tab[b][lane + 32 * k] = ti[k]; }
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        const int r = lane + 32 * k;
        misc->on_row[b][r] = ro[k];
        misc->on_tab[b][r] = to[k];
        const int g = t0 - kHalo + r;
        bad = bad || (r < kCtx && g >= 0 && g < n_tok && ro[k] == SSB_INVALID_SLOT);
      }
      const unsigned any_bad = __ballot_sync(0xffffffffu, bad);
      if (lane == 0) misc->ctx_has_invalid[b] = any_bad ? 1u : 0u;
      __syncwarp();
      const uint32_t in_base = kOffIn + b * kInBytes, on_base = kOffOn + b * kOnBytes;
      if (mode & 1) {
        if (lane == 0) mbar_expect_tx(&misc->full_bar[b], (kM + kN) * kD * 2);
        __syncwarp();
        // 72 groups of 4 rows (32 In + 40 ON), each 2 column blocks; spread over the lanes
        for (int g = lane; g < (kM + kN) / 4; g += 32) {
          const bool is_in = g < kM / 4;
          const int r = is_in ? g * 4 : (g - kM / 4) * 4;
          const uint32_t* rows = is_in ? &misc->in_tab[b][r] : &misc->on_tab[b][r];   // == cache rows unless direct
          uint32_t i0 = rows[0], i1 = rows[1], i2 = rows[2], i3 = rows[3];
          i0 = i0 == SSB_INVALID_SLOT ? 0 : i0; i1 = i1 == SSB_INVALID_SLOT ? 0 : i1;
          i2 = i2 == SSB_INVALID_SLOT ? 0 : i2; i3 = i3 == SSB_INVALID_SLOT ? 0 : i3;
          const uint32_t dst = sbase + (is_in ? in_base : on_base) + r * 128;
          const uint32_t blk = is_in ? kInBlk : kOnBlk;
          tma_gather4(dst, &tmap, &misc->full_bar[b], 0, i0, i1, i2, i3);
          tma_gather4(dst + blk, &tmap, &misc->full_bar[b], 64, i0, i1, i2, i3);
        }
      } else {
        // half a warp per row: 16 lanes x 16 B = 256 B of bf16
        for (int r = lane >> 4; r < kM + kN; r += 2) {
          const bool is_in = r < kM;
          const uint32_t rr = is_in ? r : r - kM;
          const int l16 = lane & 15;
          uint32_t row = is_in ? misc->in_tab[b][rr] : misc->on_tab[b][rr];
          uint4 v = make_uint4(0u, 0u, 0u, 0u);
          if (row != SSB_INVALID_SLOT) v = *reinterpret_cast<const uint4*>(params + (size_t)row * kD + l16 * 8);
          uint32_t off = (is_in ? in_base : on_base) + swz_off(rr, l16 * 8, is_in ? kInBlk : kOnBlk);
          *reinterpret_cast<uint4*>(smem + off) = v;
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(&misc->full_bar[b]);
      }
      // tile-local dedup: lead[p] = first position with the same row id. Duplicate positions fold
      // their dL/dS entries into the leader's row/column of G, so every distinct row gets ONE
      // gradient reduction per tile (hot words would otherwise serialise on the same address).
      {
        int ld[5];
#pragma unroll
        for (int k = 0; k < 4; ++k) ld[k] = lane + 32 * k;
        for (int j = 0; j < kM; ++j) {
          const uint32_t x = misc->in_row[b][j];
#pragma unroll
          for (int k = 0; k < 4; ++k) if (ri[k] == x) ld[k] = min(ld[k], j);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) misc->lead_in[b][lane + 32 * k] = (uint8_t)ld[k];
#pragma unroll
        for (int k = 0; k < 5; ++k) ld[k] = lane + 32 * k;
        for (int j = 0; j < kN; ++j) {
          const uint32_t x = misc->on_row[b][j];
#pragma unroll
          for (int k = 0; k < 5; ++k) if (ro[k] == x) ld[k] = min(ld[k], j);
        }
#pragma unroll
        for (int k = 0; k < 5; ++k) misc->lead_on[b][lane + 32 * k] = (uint8_t)ld[k];
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&misc->lead_bar[b]);
      load_idx(tile + gridDim.x);   // next tile's indices, in flight while the consumers work
    }
  } else {
    // =========================== consumer warps ==========================================
    const int q = warp & 3, part = warp >> 2;       // TMEM lane quarter, column part
    const int i = q * 32 + lane;                    // tile row = TMEM lane
    const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);
    uint32_t mma_phase = 0, it = 0;
    float loss_acc = 0.f, pair_acc = 0.f;

    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const int b = it & 1;
      const int t0 = tile * kM;
      const uint32_t in_s = sbase + kOffIn + b * kInBytes, on_s = sbase + kOffOn + b * kOnBytes;
      mbar_wait(&misc->full_bar[b], (it >> 1) & 1);

      // ---- S = In . ON^T --------------------------------------------------------------------
      if (tid == 0) {
        tc_fence_after();
        constexpr uint32_t idesc = make_idesc(kM, kN, 0, 0);
#pragma unroll
        for (int s = 0; s < kD / 16; ++s) {
          uint64_t a = make_desc(in_s + (s >> 2) * kInBlk + (s & 3) * 32, 16, 1024);
          uint64_t bd = make_desc(on_s + (s >> 2) * kOnBlk + (s & 3) * 32, 16, 1024);
          umma_bf16(tmem + kTmemS, a, bd, idesc, s > 0);
        }
        umma_commit(&misc->mma_bar);
      }
      mbar_wait(&misc->mma_bar, mma_phase);
      mma_phase ^= 1;
      tc_fence_after();

      // Gradient destination rows for epilogue 2, copied out of the stage's index arrays now (the
      // producer may refill them as soon as the gradient MMAs retire). Non-leaders are skipped:
      // their contribution has been folded into the leader's row / column of G.
      mbar_wait(&misc->lead_bar[b], (it >> 1) & 1);
      if (tid < kM) {
        misc->out_rows[tid] = (misc->lead_in[b][tid] == tid) ? misc->in_row[b][tid] : SSB_INVALID_SLOT;
      } else if (tid < kM + kN) {
        const int c = tid - kM;
        misc->out_rows[tid] = (misc->lead_on[b][c] == c) ? misc->on_row[b][c] : SSB_INVALID_SLOT;
      }

      // ---- epilogue 1: logits -> loss and dL/dS. Column part p owns band block p (16 of the 48
      //      window columns; part 3 has none) and negatives [4p, 4p+4) -------------------------
      {
        const int gi = t0 + i;
        const uint32_t my_in = misc->in_row[b][i];
        const bool row_ok = my_in != SSB_INVALID_SLOT;
        const uint32_t li = misc->lead_in[b][i];
        int eff = 0;
        if (row_ok) {
          uint32_t r4[4];
          Philox::gen(seed, offset + (uint64_t)gi, r4);
          eff = window - (int)(r4[0] % (uint32_t)window);
        }
        const int s_lo = (gi / sent_len) * sent_len;
        const int tok_lo = max(s_lo, 0), tok_hi = min(s_lo + sent_len, n_tok);
        // column c <-> token t0 - 8 + c ; centre i <-> column i + 8. Valid context columns of row i
        // form one range [c_lo, c_hi] minus the centre itself (plus, rarely, INVALID rows inside it).
        const int self_c = i + kHalo;
        int c_lo = 1, c_hi = 0;
        if (row_ok) {
          c_lo = max(gi - eff, tok_lo) - (t0 - kHalo);
          c_hi = min(gi + eff, tok_hi - 1) - (t0 - kHalo);
        }
        const bool chk_rows = misc->ctx_has_invalid[b] != 0;
        auto col_ok = [&](int c) -> bool {
          return c >= c_lo && c <= c_hi && c != self_c && (!chk_rows || misc->on_row[b][c] != SSB_INVALID_SLOT);
        };
        // G[lead(i)][lead(c)] += g   (bf16x2 shared-memory reduction; G was zeroed by the previous tile)
        auto g_add = [&](int c, float g) {
          const uint32_t lc = misc->lead_on[b][c];
          __nv_bfloat162 pk = (lc & 1) ? __floats2bfloat162_rn(0.f, g) : __floats2bfloat162_rn(g, 0.f);
          const uint32_t addr = sbase + kOffG + swz_off(li, lc & ~1u, kInBlk);
          asm volatile("red.shared::cta.add.noftz.bf16x2 [%0], %1;" ::"r"(addr), "r"(*reinterpret_cast<uint32_t*>(&pk)) : "memory");
        };
        if constexpr (!kDiag) {
          if (part < 3) {
            float v[16];
            const int c0 = q * 32 + part * 16;          // band window of this lane quarter: [32q, 32q+48)
            tmem_ld16(lane_base + kTmemS + c0, v);
  #pragma unroll
            for (int j = 0; j < 16; ++j) {
              if (col_ok(c0 + j)) {
                float sg, lt;
                sig_logt(v[j], sg, lt);
                g_add(c0 + j, sg - 1.0f);
                loss_acc += lt;
                pair_acc += 1.f;
              }
            }
          }
          {
            // negatives: weight m_i * K / 16 each, m_i = number of valid contexts of centre i. The 16
            // negatives are split 4/2/4/6 over the column parts to even out the band work (the
            // middle band block holds most of a row's valid contexts, part 3 has no band block).
            float m_i = 0.f;
            if (row_ok) {
              if (chk_rows) {
  #pragma unroll
                for (int c = 0; c <= 2 * kHalo; ++c) m_i += col_ok(i + c) ? 1.f : 0.f;
              } else {
                m_i = (float)max(c_hi - c_lo, 0);
              }
            }
            const int n_lo = (part == 0) ? 0 : (part == 1) ? 4 : (part == 2) ? 6 : 10;
            const int n_hi = (part == 0) ? 4 : (part == 1) ? 6 : (part == 2) ? 10 : 16;
            float v[16];
            tmem_ld16(lane_base + kTmemS + kCtx, v);
            const float w = m_i * neg_k * (1.0f / kNeg);
            if (m_i > 0.f) {
  #pragma unroll
              for (int j = 0; j < 16; ++j) {
                if (j >= n_lo && j < n_hi && misc->on_row[b][kCtx + j] != SSB_INVALID_SLOT) {
                  float sg, lt;
                  sig_logt(v[j], sg, lt);
                  g_add(kCtx + j, w * sg);
                  loss_acc += w * (v[j] + lt);
                }
              }
            }
          }
} else {
          // Diagonal form. In the column-block form above a warp runs the sigmoid path for each of its 16
          // columns with only the ~1/3 of its lanes whose band contains that column. Here the band warps
          // hold the whole 48-column window of their lane quarter (indexed per lane, so it lives in local
          // memory) and step through the OFFSETS a = 1..window on one side of the centre: lane i looks at
          // column i + 8 -/+ a, so every lane whose window reaches that far works in the same iteration.
          // part 0: left side, part 1: right side, parts 2 and 3: eight negatives each.
          if (part < 2) {
            float v[48];
            const int cq = q * 32;
            {
              float t16[16];
#pragma unroll
              for (int blk = 0; blk < 3; ++blk) {
                tmem_ld16(lane_base + kTmemS + cq + blk * 16, t16);
#pragma unroll
                for (int j = 0; j < 16; ++j) v[blk * 16 + j] = t16[j];
              }
            }
            const int sgn = (part == 0) ? -1 : 1;
            for (int a = 1; a <= window; ++a) {
              const int c = self_c + sgn * a;            // in [cq, cq + 48) for every lane: lane + 8 -/+ a, a <= 8
              if (col_ok(c)) {
                float sg, lt;
                sig_logt(v[c - cq], sg, lt);
                g_add(c, sg - 1.0f);
                loss_acc += lt;
                pair_acc += 1.f;
              }
            }
          } else {
            float m_i = 0.f;
            if (row_ok) {
              if (chk_rows) {
#pragma unroll
                for (int c = 0; c <= 2 * kHalo; ++c) m_i += col_ok(i + c) ? 1.f : 0.f;
              } else {
                m_i = (float)max(c_hi - c_lo, 0);
              }
            }
            const int n_lo = (part == 2) ? 0 : 8, n_hi = n_lo + 8;
            float v[16];
            tmem_ld16(lane_base + kTmemS + kCtx, v);
            const float w = m_i * neg_k * (1.0f / kNeg);
            if (m_i > 0.f) {
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                if (j >= n_lo && j < n_hi && misc->on_row[b][kCtx + j] != SSB_INVALID_SLOT) {
                  float sg, lt;
                  sig_logt(v[j], sg, lt);
                  g_add(kCtx + j, w * sg);
                  loss_acc += w * (v[j] + lt);
                }
              }
            }
          }
        }
        fence_proxy_async();
        tc_fence_before();
      }
      consumer_sync();
      if (debug && tile == 0)
        for (uint32_t k = tid; k < kGBlks * kInBlk / 2; k += kConsumers)
          debug[k] = __bfloat162float(reinterpret_cast<__nv_bfloat16*>(smem + kOffG)[k]);

      // ---- dIn = G . ON ;  dON = G^T . In (rows 0..127 and 128..255) ---------------------------
      if (tid == 0) {
        tc_fence_after();
        constexpr uint32_t idesc2 = make_idesc(kM, kD, 0, 1);
#pragma unroll
        for (int s = 0; s < kN / 16; ++s) {     // K = 160 context/negative rows, 16 per step
          uint64_t a = make_desc(sbase + kOffG + (s >> 2) * kInBlk + (s & 3) * 32, 16, 1024);
          uint64_t bd = make_desc(on_s + s * 2048, kOnBlk, 1024);
          umma_bf16(tmem + kTmemDIn, a, bd, idesc2, s > 0);
        }
        constexpr uint32_t idesc3 = make_idesc(128, kD, 1, 1);
#pragma unroll
        for (int s = 0; s < kM / 16; ++s) {     // K = 128 centre tokens, 16 per step
          uint64_t bd = make_desc(in_s + s * 2048, kInBlk, 1024);
          uint64_t a0 = make_desc(sbase + kOffG + s * 2048, kInBlk, 1024);
          uint64_t a1 = make_desc(sbase + kOffG + 2 * kInBlk + s * 2048, kInBlk, 1024);
          umma_bf16(tmem + kTmemDOnA, a0, bd, idesc3, s > 0);
          umma_bf16(tmem + kTmemDOnB, a1, bd, idesc3, s > 0);
        }
        umma_commit(&misc->empty_bar[b]);   // the stage's In / ON are free once these MMAs retire
        umma_commit(&misc->mma_bar);
      }
      mbar_wait(&misc->mma_bar, mma_phase);
      mma_phase ^= 1;
      tc_fence_after();

      if (debug && tile == 0 && part == 0) {
        float* d1 = debug + kGBlks * kInBlk / 2;
        for (int blk = 0; blk < 26; ++blk) {   // 416 columns: dONb | pad | dIn | dONa
          float vv[16];
          tmem_ld16(lane_base + blk * 16, vv);
          for (int j = 0; j < 16; ++j) d1[(size_t)i * 416 + blk * 16 + j] = vv[j];
        }
      }

      // ---- epilogue 2: TMEM -> fp32 staging in shared memory (the G region is dead now) -> one
      //      coalesced 512-byte vector red per gradient row. TMEM holds a row per lane; staging
      //      transposes that into a row per warp so 32 lanes reduce 32 x 16 contiguous bytes.
      {
        float* stage = reinterpret_cast<float*>(smem + kOffG);      // [128 rows][128 fp32], 16 B chunks swizzled by row
        const uint32_t tcol[3] = {kTmemDIn, kTmemDOnA, kTmemDOnB};
        const int nrows[3] = {kM, kM, kN - kM};
        const int row_base[3] = {0, kM, 2 * kM};                    // into misc->out_rows
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          if (k < 2 || q == 0) {                                    // dON rows 128..159 live in lanes 0..31 only
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              float v[16];
              tmem_ld16(lane_base + tcol[k] + part * 32 + h * 16, v);
#pragma unroll
              for (int x4 = 0; x4 < 4; ++x4) {
                const int chunk = part * 8 + h * 4 + x4;
                *reinterpret_cast<float4*>(stage + i * kD + ((chunk ^ (i & 7)) << 2)) =
                    make_float4(v[4 * x4], v[4 * x4 + 1], v[4 * x4 + 2], v[4 * x4 + 3]);
              }
            }
          }
          tc_fence_before();
          consumer_sync();
          for (int r = warp; r < nrows[k]; r += kConsumerWarps) {
            const uint32_t row = misc->out_rows[row_base[k] + r];
            if (row == SSB_INVALID_SLOT) continue;
            const float4 x = *reinterpret_cast<const float4*>(stage + r * kD + ((lane ^ (r & 7)) << 2));
            if ((x.x != 0.f || x.y != 0.f || x.z != 0.f || x.w != 0.f) && !(mode & 2))
              red_add_f4_gpu(grads + (size_t)row * kD + lane * 4, x);
          }
          consumer_sync();
        }
        // leave G zeroed for the next tile (columns 0..191; block 3 is never read back)
        for (uint32_t z = tid; z < 3 * kInBlk / 16; z += kConsumers)
          reinterpret_cast<float4*>(smem + kOffG)[z] = make_float4(0.f, 0.f, 0.f, 0.f);
        fence_proxy_async();
      }
      consumer_sync();
    }
    loss_acc = warp_sum(loss_acc);
    pair_acc = warp_sum(pair_acc);
    if (lane == 0) { misc->loss[warp] = loss_acc; misc->pairs[warp] = pair_acc; }
  }

  // ---- teardown -------------------------------------------------------------------------------
  tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    float l = 0.f, p = 0.f;
    for (int w = 0; w < kConsumerWarps; ++w) { l += misc->loss[w]; p += misc->pairs[w]; }
    if (l != 0.f) atomicAdd(loss_out, l);
    if (p != 0.f) atomicAdd(loss_out + 1, p);
  }
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, kTmemCols);
  }
}

}  // namespace serial

int launch_block_tc_serial(const TcArgs& a, const CUtensorMap& tmap, void* stream) {
  using namespace serial;
  static bool attr_set[64] = {false};   // the attribute is per device: a process may drive several
  int dev_id = 0;
  if (cudaGetDevice(&dev_id) != cudaSuccess || dev_id < 0 || dev_id >= 64) return (int)cudaErrorInvalidDevice;
  if (!attr_set[dev_id]) {
    cudaError_t e = cudaFuncSetAttribute(sgns_block_tc_serial_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(sgns_block_tc_serial_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (e != cudaSuccess) return (int)e;
    attr_set[dev_id] = true;
  }
  int64_t tiles = ((int64_t)a.n_tok + kM - 1) / kM;
  int blocks = (int)(tiles < ssb::kNumSMs ? tiles : ssb::kNumSMs);
  if (a.mode & 4)
    sgns_block_tc_serial_kernel<true><<<blocks, kThreads, kSmemBytes, (cudaStream_t)stream>>>(
        a.params, a.grads, a.inv_in, a.inv_out, a.n_tok, a.sent_len, a.window, a.seed, a.offset, a.neg_idx, a.neg_k,
        a.loss_out, tmap, a.mode, a.debug, a.dm);
  else
    sgns_block_tc_serial_kernel<false><<<blocks, kThreads, kSmemBytes, (cudaStream_t)stream>>>(
        a.params, a.grads, a.inv_in, a.inv_out, a.n_tok, a.sent_len, a.window, a.seed, a.offset, a.neg_idx, a.neg_k,
        a.loss_out, tmap, a.mode, a.debug, a.dm);
  SSB_LAUNCH_CHECK();
  return 0;
}

}  // namespace tc
}  // namespace ssb
