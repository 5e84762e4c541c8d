// a4 (FP32 parity mode): structured self-attention pooling, model/layer_module.py:49-98.
//   proj  = tanh(h * Wa1 + ba1)            (GEMM with bias+tanh epilogue, sgemm.cu)
//   score = proj * w_a2                     (:76)
//   alpha = softmax_t(score + (1-m)*(-2^31))  (:84-87; mask binarised as in model/graph.py:52)
//   e     = sum_t alpha_t h_t               (:90-96)
// One CTA per sentence; the score dot products are warp reductions, the softmax runs in one warp,
// the weighted sum is one thread per feature with coalesced reads of h.
#include <cuda_bf16.h>
#include "common.cuh"
#include "indnet_internal.h"
#include "../../include/indnet.h"

namespace indnet {

constexpr int ATT_NT = 256;

__device__ __forceinline__ float ldf(const float* p) { return *p; }
__device__ __forceinline__ float ldf(const __nv_bfloat16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ void stf(float* p, float v) { *p = v; }
__device__ __forceinline__ void stf(__nv_bfloat16* p, float v) { *p = __float2bfloat16(v); }

// Sentence x of a batch laid out as `group` rows per episode lands in one of two dense outputs:
// the first `split` rows of every episode (the support set, model/graph.py:57-58) go to `a`, the
// rest (the query set, :60-62) to `b` - the reference's two tf.slice copies become addressing.
template <typename T>
__device__ __forceinline__ T* split_row(T* a, T* b, int x, int group, int split, int H) {
  const int g = x / group, r = x % group;
  return (r < split) ? a + ((long long)g * split + r) * H
                     : b + ((long long)g * (group - split) + (r - split)) * H;
}

template <typename TI>
__global__ void __launch_bounds__(ATT_NT)
attn_pool_fwd_kernel(const TI* __restrict__ hout, const TI* __restrict__ proj,
                     const int* __restrict__ mask, const float* __restrict__ w2,
                     float* __restrict__ enc_a, float* __restrict__ enc_b, float* __restrict__ alphas,
                     int L, int H, int group, int split) {
  extern __shared__ float sc[];   // [L]
  const int x = blockIdx.x;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32, nw = ATT_NT / 32;
  const TI* pj = proj + (long long)x * L * H;
  for (int t = warp; t < L; t += nw) {
    float s = 0.f;
    for (int h = lane; h < H; h += 32) s = fmaf(ldf(pj + (long long)t * H + h), w2[h], s);
    s = warp_sum(s);
    if (lane == 0) {
      const float m = mask[(long long)x * L + t] > 0 ? 1.0f : 0.0f;
      sc[t] = s + (1.0f - m) * (-2147483648.0f);
    }
  }
  __syncthreads();
  if (warp == 0) {
    float mx = -INFINITY;
    for (int t = lane; t < L; t += 32) mx = fmaxf(mx, sc[t]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int t = lane; t < L; t += 32) {
      float e = expf(sc[t] - mx);
      sc[t] = e;
      sum += e;
    }
    sum = warp_sum(sum);
    const float inv = 1.0f / sum;
    for (int t = lane; t < L; t += 32) {
      float a = sc[t] * inv;
      sc[t] = a;
      alphas[(long long)x * L + t] = a;
    }
  }
  __syncthreads();
  const TI* hp = hout + (long long)x * L * H;
  float* enc = split_row(enc_a, enc_b, x, group, split, H);
  for (int h = threadIdx.x; h < H; h += ATT_NT) {
    float acc = 0.f;
    for (int t = 0; t < L; ++t) acc = fmaf(sc[t], ldf(hp + (long long)t * H + h), acc);
    enc[h] = acc;
  }
}

// Backward of the pooling part.  Given denc [X,H] (alphas itself carries no loss gradient):
//   dalpha_t = denc . h_t ; dscore_t = alpha_t (dalpha_t - sum_s alpha_s dalpha_s)
//   dhout[t] = alpha_t * denc        (the Dense path adds dpre * Wa1^T afterwards, beta = 1)
//   dpre[t]  = dscore_t * w2 * (1 - proj_t^2)
//   dw2_part[x] = sum_t dscore_t * proj_t   (summed over sentences by colsum)
template <typename TI>
__global__ void __launch_bounds__(ATT_NT)
attn_pool_bwd_kernel(const float* __restrict__ denc_a, const float* __restrict__ denc_b,
                     const TI* __restrict__ hout,
                     const TI* proj /* may alias dpre */, const float* __restrict__ alphas,
                     const float* __restrict__ w2, float* __restrict__ dhout, TI* dpre,
                     float* __restrict__ dw2_part, float* __restrict__ dsum_part, int L, int H, int group,
                     int split) {
  extern __shared__ float sm[];   // dsc[L], al[L]
  float* dsc = sm;
  float* al = sm + L;
  const int x = blockIdx.x;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32, nw = ATT_NT / 32;
  const TI* hp = hout + (long long)x * L * H;
  const TI* pj = proj + (long long)x * L * H;
  const float* de = split_row(denc_a, denc_b, x, group, split, H);
  for (int t = warp; t < L; t += nw) {
    float s = 0.f;
    for (int h = lane; h < H; h += 32) s = fmaf(de[h], ldf(hp + (long long)t * H + h), s);
    s = warp_sum(s);
    if (lane == 0) { dsc[t] = s; al[t] = alphas[(long long)x * L + t]; }
  }
  __syncthreads();
  if (warp == 0) {
    float dot = 0.f;
    for (int t = lane; t < L; t += 32) dot = fmaf(al[t], dsc[t], dot);
    dot = warp_sum(dot);
    for (int t = lane; t < L; t += 32) dsc[t] = al[t] * (dsc[t] - dot);
  }
  __syncthreads();
  for (int h = threadIdx.x; h < H; h += ATT_NT) {
    const float g = de[h], w = w2[h];
    float dw = 0.f, dsum = 0.f;
    for (int t = 0; t < L; ++t) {
      const long long o = ((long long)x * L + t) * H + h;
      const float p = ldf(pj + (long long)t * H + h);
      const float dp = dsc[t] * w * (1.0f - p * p);
      if (dhout) dhout[o] = al[t] * g;
      stf(dpre + o, dp);
      dw = fmaf(dsc[t], p, dw);
      dsum += dp;
    }
    dw2_part[(long long)x * H + h] = dw;
    if (dsum_part) dsum_part[(long long)x * H + h] = dsum;   // Dense-bias gradient partial, summed in fp32
  }
}

}  // namespace indnet

using namespace indnet;

extern "C" int indnet_attn_pool_fwd(indnet_stream st_, const float* hout, const float* proj, const int* mask,
                                    const float* w_a2, float* enc_a, float* enc_b, float* alphas, int X, int L,
                                    int H, int group, int split) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(X >= 0 && L > 0 && H > 0, "attn_pool_fwd: bad sizes X=%d L=%d H=%d", X, L, H);
  if (X == 0) return INDNET_OK;
  IN_REQUIRE(hout && proj && mask && w_a2 && enc_a && alphas, "attn_pool_fwd: null pointer");
  IN_REQUIRE(group > 0 && split >= 0 && split <= group && X % group == 0, "attn_pool_fwd: bad group/split %d/%d", group, split);
  IN_REQUIRE(split == group || enc_b, "attn_pool_fwd: second output missing");
  IN_REQUIRE(L <= 8192, "attn_pool_fwd: sequence_length %d too long", L);
  attn_pool_fwd_kernel<float><<<X, ATT_NT, sizeof(float) * L, st>>>(hout, proj, mask, w_a2, enc_a, enc_b, alphas, L, H, group, split);
  IN_LAUNCH_CHECK("attn_pool_fwd_kernel");
  return INDNET_OK;
}

extern "C" int indnet_attn_pool_bwd(indnet_stream st_, const float* denc_a, const float* denc_b, const float* hout,
                                    const float* proj, const float* alphas, const float* w_a2, float* dhout,
                                    float* dpre, float* dw2_part, int X, int L, int H, int group, int split) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(X >= 0 && L > 0 && H > 0, "attn_pool_bwd: bad sizes X=%d L=%d H=%d", X, L, H);
  if (X == 0) return INDNET_OK;
  IN_REQUIRE(denc_a && hout && proj && alphas && w_a2 && dhout && dpre && dw2_part, "attn_pool_bwd: null pointer");
  IN_REQUIRE(group > 0 && split >= 0 && split <= group && X % group == 0, "attn_pool_bwd: bad group/split %d/%d", group, split);
  IN_REQUIRE(split == group || denc_b, "attn_pool_bwd: second gradient input missing");
  IN_REQUIRE(L <= 4096, "attn_pool_bwd: sequence_length %d too long", L);
  attn_pool_bwd_kernel<float><<<X, ATT_NT, sizeof(float) * 2 * L, st>>>(denc_a, denc_b, hout, proj, alphas, w_a2, dhout, dpre, dw2_part, nullptr, L, H, group, split);
  IN_LAUNCH_CHECK("attn_pool_bwd_kernel");
  return INDNET_OK;
}

// bf16 tensor-core mode: same kernels reading the bf16 copies of h and proj (and writing dpre in bf16,
// the operand format of the Dense-gradient GEMMs), which halves their HBM traffic.
extern "C" int indnet_attn_pool_fwd_bf16(indnet_stream st_, const void* hout_bf, const void* proj_bf, const int* mask,
                                         const float* w_a2, float* enc_a, float* enc_b, float* alphas, int X, int L,
                                         int H, int group, int split) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(X >= 0 && L > 0 && H > 0 && L <= 8192, "attn_pool_fwd_bf16: bad sizes X=%d L=%d H=%d", X, L, H);
  if (X == 0) return INDNET_OK;
  IN_REQUIRE(hout_bf && proj_bf && mask && w_a2 && enc_a && alphas, "attn_pool_fwd_bf16: null pointer");
  IN_REQUIRE(group > 0 && split >= 0 && split <= group && X % group == 0, "attn_pool_fwd_bf16: bad group/split %d/%d", group, split);
  IN_REQUIRE(split == group || enc_b, "attn_pool_fwd_bf16: second output missing");
  attn_pool_fwd_kernel<__nv_bfloat16><<<X, ATT_NT, sizeof(float) * L, st>>>(
      (const __nv_bfloat16*)hout_bf, (const __nv_bfloat16*)proj_bf, mask, w_a2, enc_a, enc_b, alphas, L, H, group, split);
  IN_LAUNCH_CHECK("attn_pool_fwd_kernel<bf16>");
  return INDNET_OK;
}

extern "C" int indnet_attn_pool_bwd_bf16(indnet_stream st_, const float* denc_a, const float* denc_b,
                                         const void* hout_bf, const void* proj_bf, const float* alphas,
                                         const float* w_a2, float* dhout, void* dpre_bf, float* dw2_part,
                                         float* dsum_part, int X, int L, int H, int group, int split) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(X >= 0 && L > 0 && H > 0 && L <= 4096, "attn_pool_bwd_bf16: bad sizes X=%d L=%d H=%d", X, L, H);
  if (X == 0) return INDNET_OK;
  IN_REQUIRE(denc_a && hout_bf && proj_bf && alphas && w_a2 && dpre_bf && dw2_part, "attn_pool_bwd_bf16: null pointer");
  IN_REQUIRE(group > 0 && split >= 0 && split <= group && X % group == 0, "attn_pool_bwd_bf16: bad group/split %d/%d", group, split);
  IN_REQUIRE(split == group || denc_b, "attn_pool_bwd_bf16: second gradient input missing");
  attn_pool_bwd_kernel<__nv_bfloat16><<<X, ATT_NT, sizeof(float) * 2 * L, st>>>(
      denc_a, denc_b, (const __nv_bfloat16*)hout_bf, (const __nv_bfloat16*)proj_bf, alphas, w_a2, dhout,
      (__nv_bfloat16*)dpre_bf, dw2_part, dsum_part, L, H, group, split);
  IN_LAUNCH_CHECK("attn_pool_bwd_kernel<bf16>");
  return INDNET_OK;
}
