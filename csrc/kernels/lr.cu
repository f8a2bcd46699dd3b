// lr.cu — K6: sparse logistic regression forward/backward on the worker's pulled cache.
// The reference's logistic_regression app is absent from the snapshot; the
// loop shape is the BaseAlgorithm contract (core/framework/SwiftWorker.h:19-57): pull feature
// weights, per-example dot + sigmoid + log-loss, accumulate per-feature gradients, push
// (AdaGrad applied on the owner shard, parameter/sparse_access_method.h:33-46).
// One warp per example: lanes stride over the example's features (Criteo shape: 39).
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
