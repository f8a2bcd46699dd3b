// sgns.cu — word2vec skip-gram with negative sampling on the worker's pulled cache.
//
// The reference's word2vec app is absent from the snapshot: its loop is
// "pull keys -> dot products on Vec (utils/vec1.h:97-104) + sigmoid -> accumulate
// cache.grads()[k] -> push". The math here is the standard SGNS objective (Mikolov et al.)
// written fresh:   L = -log s(a.p) - w * sum_n log s(-a.neg_n),  w = K / n_neg
// with one set of n_neg negatives shared by the 128 (centre, context) pairs of a tile.
//
// Two kernels: pair generation (dynamic window, compacted) and the CUDA-core tile kernel.
// The tcgen05/TMEM variant of the tile kernel lives in sgns_tc.cu; this one is the
// always-correct baseline it is checked against.
#include "common.cuh"

namespace ssb {

// ---- pair generation ----------------------------------------------------------------------
// One thread per token (the centre). Window shrink b ~ U[0, window) as in classic word2vec, so
// the effective half-window is window - b. Pairs never cross a sentence boundary.
__global__ void __launch_bounds__(256)
sgns_make_pairs_kernel(const uint32_t* __restrict__ inv_in, const uint32_t* __restrict__ inv_out,
                       int64_t n_tok, int sent_len, int window, uint64_t seed, uint64_t offset,
                       uint32_t* __restrict__ pairs_in, uint32_t* __restrict__ pairs_out,
                       uint32_t* __restrict__ n_pairs) {
  const int lane = threadIdx.x & 31;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t n_round = (n_tok + 31) / 32 * 32;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n_round; i += stride) {
    int cnt = 0, lo = 0, hi = 0;
    uint32_t ci = SSB_INVALID_SLOT;
    if (i < n_tok) {
      ci = inv_in[i];
      uint32_t r[4];
      Philox::gen(seed, offset + (uint64_t)i, r);
      int eff = window - (int)(r[0] % (uint32_t)window);
      int pos = (int)(i % sent_len);
      lo = -min(eff, pos);
      hi = min(eff, sent_len - 1 - pos);
      if (i + hi >= n_tok) hi = (int)(n_tok - 1 - i);
      if (ci != SSB_INVALID_SLOT)
        for (int j = lo; j <= hi; ++j)
          if (j != 0 && inv_out[i + j] != SSB_INVALID_SLOT) ++cnt;
    }
    // warp exclusive scan of cnt
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    int total = __shfl_sync(0xffffffffu, incl, 31);
    uint32_t base = 0;
    if (lane == 31 && total > 0) base = atomicAdd(n_pairs, (uint32_t)total);
    base = __shfl_sync(0xffffffffu, base, 31);
    if (cnt > 0) {
      uint32_t dst = base + (uint32_t)(incl - cnt);
      for (int j = lo; j <= hi; ++j) {
        if (j == 0) continue;
        uint32_t co = inv_out[i + j];
        if (co == SSB_INVALID_SLOT) continue;
        pairs_in[dst] = ci;
        pairs_out[dst] = co;
        ++dst;
      }
    }
  }
}

// ---- CUDA-core tile kernel -----------------------------------------------------------------
// Block = 8 warps = one tile of 128 pairs at a time (persistent, grid-stride over tiles).
// Lane l holds float4 slices {l, l+32, ...} of each row; the tile's negatives sit in smem.
constexpr int kTilePairs = 128;

template <int NNEG, int VPL>
__global__ void __launch_bounds__(256)
sgns_tiles_kernel(const float* __restrict__ params, float* __restrict__ grads, int D,
                  const uint32_t* __restrict__ pairs_in, const uint32_t* __restrict__ pairs_out,
                  const uint32_t* __restrict__ n_pairs_dev, int64_t max_pairs,
                  const uint32_t* __restrict__ neg_idx, float neg_weight, float* __restrict__ loss_out) {
  extern __shared__ float smem[];
  float* sneg = smem;                  // [NNEG][D]
  float* sdneg = smem + NNEG * D;      // [NNEG][D]
  __shared__ uint32_t sneg_row[NNEG];
  __shared__ float sloss[8];

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t n_pairs = *n_pairs_dev;
  if ((int64_t)n_pairs > max_pairs) n_pairs = (uint32_t)max_pairs;
  const uint32_t n_tiles = (n_pairs + kTilePairs - 1) / kTilePairs;
  const int nvec = D >> 2;
  float loss_acc = 0.f;

  for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    __syncthreads();
    if (threadIdx.x < NNEG) sneg_row[threadIdx.x] = neg_idx[(size_t)tile * NNEG + threadIdx.x];
    __syncthreads();
    for (int idx = threadIdx.x; idx < NNEG * nvec; idx += blockDim.x) {
      int n = idx / nvec, j = idx - n * nvec;
      uint32_t r = sneg_row[n];
      float4 v = (r == SSB_INVALID_SLOT) ? make_float4(0, 0, 0, 0)
                                         : *reinterpret_cast<const float4*>(params + (size_t)r * D + 4 * j);
      reinterpret_cast<float4*>(sneg)[idx] = v;
      reinterpret_cast<float4*>(sdneg)[idx] = make_float4(0, 0, 0, 0);
    }
    __syncthreads();

    float4 dneg[NNEG][VPL];
#pragma unroll
    for (int n = 0; n < NNEG; ++n)
#pragma unroll
      for (int v = 0; v < VPL; ++v) dneg[n][v] = make_float4(0, 0, 0, 0);

    const uint32_t p0 = tile * kTilePairs + warp * (kTilePairs / 8);
    for (int q = 0; q < kTilePairs / 8; ++q) {
      const uint32_t pi = p0 + q;
      if (pi >= n_pairs) break;
      const uint32_t ri = pairs_in[pi], ro = pairs_out[pi];
      float4 a[VPL], p[VPL];
      float s[NNEG + 1];
#pragma unroll
      for (int n = 0; n <= NNEG; ++n) s[n] = 0.f;
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
        int j = lane + 32 * v;
        if (j < nvec) {
          a[v] = *reinterpret_cast<const float4*>(params + (size_t)ri * D + 4 * j);
          p[v] = *reinterpret_cast<const float4*>(params + (size_t)ro * D + 4 * j);
          s[NNEG] += a[v].x * p[v].x + a[v].y * p[v].y + a[v].z * p[v].z + a[v].w * p[v].w;
#pragma unroll
          for (int n = 0; n < NNEG; ++n) {
            float4 ng = reinterpret_cast<const float4*>(sneg)[n * nvec + j];
            s[n] += a[v].x * ng.x + a[v].y * ng.y + a[v].z * ng.z + a[v].w * ng.w;
          }
        } else {
          a[v] = make_float4(0, 0, 0, 0);
          p[v] = make_float4(0, 0, 0, 0);
        }
      }
#pragma unroll
      for (int n = 0; n <= NNEG; ++n) s[n] = warp_sum(s[n]);

      const float gpos = fast_sigmoid(s[NNEG]) - 1.0f;
      float l = neg_log_sigmoid(s[NNEG]);
      float g[NNEG];
#pragma unroll
      for (int n = 0; n < NNEG; ++n) {
        const uint32_t nr = sneg_row[n];
        const bool skip = (nr == ro) || (nr == SSB_INVALID_SLOT);  // word2vec skips negative == target
        g[n] = skip ? 0.f : neg_weight * fast_sigmoid(s[n]);
        l += skip ? 0.f : neg_weight * neg_log_sigmoid(-s[n]);
      }
      loss_acc += l;  // identical on all lanes; lane 0's copy is used
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
        int j = lane + 32 * v;
        if (j < nvec) {
          float4 da = make_float4(gpos * p[v].x, gpos * p[v].y, gpos * p[v].z, gpos * p[v].w);
#pragma unroll
          for (int n = 0; n < NNEG; ++n) {
            float4 ng = reinterpret_cast<const float4*>(sneg)[n * nvec + j];
            da.x += g[n] * ng.x; da.y += g[n] * ng.y; da.z += g[n] * ng.z; da.w += g[n] * ng.w;
            dneg[n][v].x += g[n] * a[v].x; dneg[n][v].y += g[n] * a[v].y;
            dneg[n][v].z += g[n] * a[v].z; dneg[n][v].w += g[n] * a[v].w;
          }
          red_add_f4_gpu(grads + (size_t)ri * D + 4 * j, da);
          red_add_f4_gpu(grads + (size_t)ro * D + 4 * j,
                         make_float4(gpos * a[v].x, gpos * a[v].y, gpos * a[v].z, gpos * a[v].w));
        }
      }
    }
    // reduce the per-warp negative gradients through shared memory, then one red per element
#pragma unroll
    for (int n = 0; n < NNEG; ++n)
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
        int j = lane + 32 * v;
        if (j < nvec) {
          float* d = sdneg + (size_t)(n * nvec + j) * 4;
          atomicAdd(d + 0, dneg[n][v].x); atomicAdd(d + 1, dneg[n][v].y);
          atomicAdd(d + 2, dneg[n][v].z); atomicAdd(d + 3, dneg[n][v].w);
        }
      }
    __syncthreads();
    for (int idx = threadIdx.x; idx < NNEG * nvec; idx += blockDim.x) {
      int n = idx / nvec, j = idx - n * nvec;
      uint32_t r = sneg_row[n];
      if (r != SSB_INVALID_SLOT)
        red_add_f4_gpu(grads + (size_t)r * D + 4 * j, reinterpret_cast<const float4*>(sdneg)[idx]);
    }
  }
  // loss reduction: lane 0 of each warp -> block -> one atomic
  if (lane == 0) sloss[warp] = loss_acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += sloss[w];
    if (t != 0.f) atomicAdd(loss_out, t);
    if (blockIdx.x == 0) atomicAdd(loss_out + 1, (float)n_pairs);
  }
}

template <int NNEG, int VPL>
int launch_tiles(const float* params, float* grads, int D, const uint32_t* pairs_in, const uint32_t* pairs_out,
                 const uint32_t* n_pairs, int64_t max_pairs, const uint32_t* neg_idx, float neg_weight,
                 float* loss_out, cudaStream_t s) {
  size_t smem = (size_t)2 * NNEG * D * sizeof(float);
  auto kern = sgns_tiles_kernel<NNEG, VPL>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  int64_t tiles = (max_pairs + kTilePairs - 1) / kTilePairs;
  int blocks = (int)(tiles < kNumSMs * 4 ? (tiles < 1 ? 1 : tiles) : kNumSMs * 4);
  kern<<<blocks, 256, smem, s>>>(params, grads, D, pairs_in, pairs_out, n_pairs, max_pairs, neg_idx,
                                 neg_weight, loss_out);
  SSB_LAUNCH_CHECK();
  return 0;
}

}  // namespace ssb

extern "C" {

int ssb_sgns_make_pairs(const uint32_t* inv_in, const uint32_t* inv_out, int64_t n_tok, int32_t sent_len,
                        int32_t window, uint64_t seed, uint64_t offset, uint32_t* pairs_in,
                        uint32_t* pairs_out, uint32_t* n_pairs, void* stream) {
  if (n_tok <= 0) return 0;
  if (sent_len < 1 || window < 1) return (int)cudaErrorInvalidValue;
  cudaError_t e = cudaMemsetAsync(n_pairs, 0, sizeof(uint32_t), (cudaStream_t)stream);
  if (e != cudaSuccess) return (int)e;
  ssb::sgns_make_pairs_kernel<<<ssb::grid_for(n_tok, 256), 256, 0, (cudaStream_t)stream>>>(
      inv_in, inv_out, n_tok, sent_len, window, seed, offset, pairs_in, pairs_out, n_pairs);
  SSB_LAUNCH_CHECK();
  return 0;
}

int ssb_sgns_tiles(const float* params, float* grads, int32_t dim, const uint32_t* pairs_in,
                   const uint32_t* pairs_out, const uint32_t* n_pairs, int64_t max_pairs,
                   const uint32_t* neg_idx, int32_t n_neg, float neg_weight, float* loss_out, void* stream) {
  if (max_pairs <= 0) return 0;
  if (dim % 4 != 0 || dim > 256 || dim < 4) return (int)cudaErrorInvalidValue;
  cudaStream_t s = (cudaStream_t)stream;
#define SSB_DISPATCH(NN)                                                                                   \
  case NN:                                                                                                 \
    return dim <= 128 ? ssb::launch_tiles<NN, 1>(params, grads, dim, pairs_in, pairs_out, n_pairs,         \
                                                 max_pairs, neg_idx, neg_weight, loss_out, s)              \
                      : ssb::launch_tiles<NN, 2>(params, grads, dim, pairs_in, pairs_out, n_pairs,         \
                                                 max_pairs, neg_idx, neg_weight, loss_out, s);
  switch (n_neg) {
    SSB_DISPATCH(4)
    SSB_DISPATCH(8)
    SSB_DISPATCH(16)
    default:
      return (int)cudaErrorInvalidValue;
  }
#undef SSB_DISPATCH
}

}  // extern "C"
