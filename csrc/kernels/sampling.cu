// sampling.cu — K9: negative sampling with Walker's alias method and a Philox counter RNG.
// The reference apps (absent from the snapshot) would draw negatives from the
// unigram^0.75 table of classic word2vec on the CPU; here one thread draws 2 samples from one
// Philox4x32-10 block, O(1) per sample regardless of vocabulary size.
#include "common.cuh"

namespace ssb {

__global__ void __launch_bounds__(256)
neg_sample_kernel(const float* __restrict__ prob, const uint32_t* __restrict__ alias, uint32_t vocab,
                  uint64_t seed, uint64_t offset, uint32_t key_offset, uint32_t* __restrict__ out, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t n2 = (n + 1) / 2;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < n2; t += stride) {
    uint32_t r[4];
    Philox::gen(seed, offset + (uint64_t)t, r);
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      int64_t i = 2 * t + s;
      if (i >= n) break;
      uint32_t col = reduce_range(r[2 * s], vocab);
      float u = (float)(r[2 * s + 1] >> 8) * (1.0f / 16777216.0f);
      uint32_t pick = (u < __ldg(prob + col)) ? col : __ldg(alias + col);
      out[i] = pick + key_offset;
    }
  }
}

// The step's key list in one pass: keys[0..B) = the tokens (input vectors), keys[B..2B) = token + vocab (output
// vectors; padding stays padding). Replaces three eager framework ops (copy_, add, masked_fill_) per step.
__global__ void __launch_bounds__(256)
build_keys_kernel(const uint32_t* __restrict__ tokens, int64_t n, uint32_t vocab, uint32_t* __restrict__ keys) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t t = __ldg(tokens + i);
    keys[i] = t;
    keys[n + i] = (t >= SSB_LOCKED_KEY) ? t : t + vocab;
  }
}

}  // namespace ssb

extern "C" int ssb_build_keys(const uint32_t* tokens, int64_t n, uint32_t vocab, uint32_t* keys, void* stream) {
  if (n <= 0) return 0;
  if (!tokens || !keys) return (int)cudaErrorInvalidValue;
  ssb::build_keys_kernel<<<ssb::grid_for(n, 256 * 4), 256, 0, (cudaStream_t)stream>>>(tokens, n, vocab, keys);
  SSB_LAUNCH_CHECK();
  return 0;
}

extern "C" int ssb_neg_sample(const float* prob, const uint32_t* alias, uint32_t vocab, uint64_t seed,
                              uint64_t offset, uint32_t key_offset, uint32_t* out, int64_t n, void* stream) {
  if (n <= 0) return 0;
  if (vocab == 0) return (int)cudaErrorInvalidValue;
  ssb::neg_sample_kernel<<<ssb::grid_for((n + 1) / 2, 256), 256, 0, (cudaStream_t)stream>>>(
      prob, alias, vocab, seed, offset, key_offset, out, n);
  SSB_LAUNCH_CHECK();
  return 0;
}
