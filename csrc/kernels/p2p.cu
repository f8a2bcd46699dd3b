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

// Reduction-throughput probe: how fast can all SMs add fp32 rows into a [n_rows x 128] buffer at
// random rows? pattern 0: one `red.v4.f32` per warp covers a whole 512-byte row (32 lanes x 16 B);
// 1: one scalar `red.f32` per warp covers a 128-byte quarter of a row (four per row — the form the
// tensor-core SGNS kernel's epilogue 2 uses); 2: one `red.v4.f32` per warp covers 128 bytes of four
// different rows; 3: plain 16-byte stores of a whole row (no read-modify-write), for reference.
// Every warp touches `iters` random rows; bytes per launch = warps * iters * 512.
__global__ void __launch_bounds__(512)
red_probe_kernel(float* __restrict__ buf, uint32_t n_rows, int iters, int pattern) {
  const int lane = threadIdx.x & 31;
  const uint32_t gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  uint32_t st = gw * 0x9E3779B1u + 12345u;
  for (int it = 0; it < iters; ++it) {
    st = st * 1664525u + 1013904223u;
    const uint32_t row = reduce_range(slot_hash(st), n_rows);
    float* r = buf + (size_t)row * 128;
    const float4 v = make_float4(1.f, 2.f, 3.f, 4.f);
    if (pattern == 0) {
      red_add_f4_gpu(r + lane * 4, v);
    } else if (pattern == 1) {
#pragma unroll
      for (int qd = 0; qd < 4; ++qd)
        asm volatile("red.relaxed.gpu.global.add.f32 [%0], %1;" ::"l"(r + qd * 32 + lane), "f"(1.f) : "memory");
    } else if (pattern == 2) {
#pragma unroll
      for (int qd = 0; qd < 4; ++qd) {
        const uint32_t row2 = reduce_range(slot_hash(st + 77u * (uint32_t)(lane & 3) + (uint32_t)qd), n_rows);
        red_add_f4_gpu(buf + (size_t)row2 * 128 + qd * 32 + (lane >> 2) * 4, v);
      }
    } else {
      *reinterpret_cast<float4*>(r + lane * 4) = v;
    }
  }
}
}  // namespace ssb

extern "C" int ssb_red_probe(float* buf, uint32_t n_rows, int32_t iters, int32_t pattern, int32_t blocks, void* stream) {
  if (!buf || n_rows == 0 || iters <= 0) return (int)cudaErrorInvalidValue;
  if (blocks <= 0) blocks = ssb::kNumSMs;
  ssb::red_probe_kernel<<<blocks, 512, 0, (cudaStream_t)stream>>>(buf, n_rows, iters, pattern);
  SSB_LAUNCH_CHECK();
  return 0;
}

extern "C" int ssb_p2p_copy(const void* src, void* dst, int64_t n_bytes, int32_t blocks, void* stream) {
  if (n_bytes <= 0) return 0;
  if (n_bytes % 16 != 0) return (int)cudaErrorInvalidValue;
  if (blocks <= 0) blocks = ssb::kNumSMs * 4;
  ssb::p2p_copy_kernel<<<blocks, 512, 0, (cudaStream_t)stream>>>((const int4*)src, (int4*)dst, n_bytes / 16);
  SSB_LAUNCH_CHECK();
  return 0;
}
