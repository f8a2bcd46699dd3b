// p2p.cu — K12: peer copy probe (the NVLink roofline denominator) — a plain grid of 16-byte
// loads and stores where either side may be a peer mapping. ref: transfer/transfer.h as a whole
// (the reference's transport is ZeroMQ over TCP; this measures what replaces it).
#include "common.cuh"

namespace ssb {
__global__ void __launch_bounds__(512)
p2p_copy_kernel(const int4* __restrict__ src, int4* __restrict__ dst, int64_t n16) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  // 4 independent 16-byte requests in flight per thread
  for (; i + 3 * stride < n16; i += 4 * stride) {
    int4 a = src[i], b = src[i + stride], c = src[i + 2 * stride], d = src[i + 3 * stride];
    dst[i] = a; dst[i + stride] = b; dst[i + 2 * stride] = c; dst[i + 3 * stride] = d;
  }
  for (; i < n16; i += stride) dst[i] = src[i];
}
}  // namespace ssb

extern "C" int ssb_p2p_copy(const void* src, void* dst, int64_t n_bytes, int32_t blocks, void* stream) {
  if (n_bytes <= 0) return 0;
  if (n_bytes % 16 != 0) return (int)cudaErrorInvalidValue;
  if (blocks <= 0) blocks = ssb::kNumSMs * 4;
  ssb::p2p_copy_kernel<<<blocks, 512, 0, (cudaStream_t)stream>>>((const int4*)src, (int4*)dst, n_bytes / 16);
  SSB_LAUNCH_CHECK();
  return 0;
}
