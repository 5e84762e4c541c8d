// a11: build_optimize (model/base.py:94-99): staircase exponential decay + tf.train.AdamOptimizer as
// driven by optimize_loss, fused over ONE flat parameter buffer (all trainables of the graph live in
// one allocation, so the data-parallel gradient all-reduce and this update are each a single call).
//   lr   = lr0 * decay_rate ^ floor(global_step / decay_steps)          (global_step BEFORE the step)
//   lr_t = lr * sqrt(1 - beta2^t) / (1 - beta1^t),  t = global_step + 1
//   m = b1 m + (1-b1) g ; v = b2 v + (1-b2) g^2 ; theta -= lr_t * m / (sqrt(v) + eps)   (SURVEY A.6:
//   epsilon is outside the bias correction, unlike torch.optim.Adam)
// global_step lives on the device so the whole train step can sit in one CUDA graph.
#include <cuda_bf16.h>
#include "common.cuh"
#include "indnet_internal.h"
#include "../../include/indnet.h"

namespace indnet {

__global__ void adam_prep_kernel(long long* global_step, float* hyper, float lr0, int decay_steps, float decay_rate,
                                 float beta1, float beta2, int advance) {
  const long long gs = *global_step;
  const float lr = lr0 * powf(decay_rate, floorf((float)gs / (float)decay_steps));
  const double t = (double)(gs + 1);
  const float lr_t = (float)((double)lr * sqrt(1.0 - pow((double)beta2, t)) / (1.0 - pow((double)beta1, t)));
  hyper[0] = lr_t;
  hyper[1] = lr;
  if (advance) *global_step = gs + 1;
}

__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, long long n, const float* __restrict__ hyper, float beta1,
                            float beta2, float eps, float gscale, __nv_bfloat16* __restrict__ shadow,
                            long long* __restrict__ advance, const int* __restrict__ enable) {
  if (enable && *enable == 0) return;          // a deferred slice with nothing pending (first step, or already flushed)
  const float lr_t = hyper[0];
  if (advance && blockIdx.x == 0 && threadIdx.x == 0) *advance += 1;    // the step's last slice closes the step
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float gi = g[i] * gscale;
    const float mi = beta1 * m[i] + (1.0f - beta1) * gi;
    const float vi = beta2 * v[i] + (1.0f - beta2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    const float pi = p[i] - lr_t * mi / (sqrtf(vi) + eps);
    p[i] = pi;
    if (shadow) shadow[i] = __float2bfloat16(pi);   // bf16 copy read by the tensor-core mode kernels
  }
}

}  // namespace indnet

using namespace indnet;

extern "C" int indnet_adam_step(indnet_stream st_, float* params, const float* grads, float* m, float* v, long long n,
                                long long* global_step, float* hyper, float lr0, int decay_steps, float decay_rate,
                                float beta1, float beta2, float eps, float grad_scale, void* shadow_bf16) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(n >= 0 && decay_steps > 0, "adam_step: bad sizes");
  IN_REQUIRE(global_step && hyper, "adam_step: null step/hyper pointer");
  adam_prep_kernel<<<1, 1, 0, st>>>(global_step, hyper, lr0, decay_steps, decay_rate, beta1, beta2, 1);
  IN_LAUNCH_CHECK("adam_prep_kernel");
  if (n == 0) return INDNET_OK;
  IN_REQUIRE(params && grads && m && v, "adam_step: null pointer");
  int blocks = ceil_div(n, 256 * 4);
  if (blocks > 148 * 8) blocks = 148 * 8;
  adam_kernel<<<blocks, 256, 0, st>>>(params, grads, m, v, n, hyper, beta1, beta2, eps, grad_scale,
                                           (__nv_bfloat16*)shadow_bf16, nullptr, nullptr);
  IN_LAUNCH_CHECK("adam_kernel");
  return INDNET_OK;
}

// The same update in two halves, so that a slice of the flat buffer whose gradient is final early in the
// backward pass (M, 96 % of the parameters) can be updated on a side stream underneath the rest of the pass:
// prep(advance = 0) -> apply(slice) ... apply(rest, advance_step = &global_step).  prep is idempotent for a given
// global_step, and the dropout streams of the step keep reading the un-advanced counter until the end.
extern "C" int indnet_adam_prep(indnet_stream st_, long long* global_step, float* hyper, float lr0, int decay_steps,
                                float decay_rate, float beta1, float beta2, int advance) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(decay_steps > 0 && global_step && hyper, "adam_prep: bad arguments");
  adam_prep_kernel<<<1, 1, 0, st>>>(global_step, hyper, lr0, decay_steps, decay_rate, beta1, beta2, advance ? 1 : 0);
  IN_LAUNCH_CHECK("adam_prep_kernel");
  return INDNET_OK;
}

extern "C" int indnet_adam_apply(indnet_stream st_, float* params, const float* grads, float* m, float* v, long long n,
                                 const float* hyper, float beta1, float beta2, float eps, float grad_scale,
                                 void* shadow_bf16, long long* advance_step, const int* enable) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(n >= 0 && hyper, "adam_apply: bad arguments");
  IN_REQUIRE(n > 0 || !advance_step, "adam_apply: an empty slice cannot close the step");
  if (n == 0) return INDNET_OK;
  IN_REQUIRE(params && grads && m && v, "adam_apply: null pointer");
  int blocks = ceil_div(n, 256 * 4);
  if (blocks > 148 * 8) blocks = 148 * 8;
  adam_kernel<<<blocks, 256, 0, st>>>(params, grads, m, v, n, hyper, beta1, beta2, eps, grad_scale,
                                           (__nv_bfloat16*)shadow_bf16, advance_step, enable);
  IN_LAUNCH_CHECK("adam_kernel");
  return INDNET_OK;
}
