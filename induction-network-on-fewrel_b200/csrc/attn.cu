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


// ---- bf16 tensor-core mode, H a multiple of 8: the same math with 16-byte accesses.  A warp reads a whole
// row of 256 features with one load instruction (8 bf16 per lane); the sums over time run as 8 groups of
// time steps in parallel (thread = (time group, 8-feature chunk)) and meet in shared memory.
__device__ __forceinline__ void unpack8(const uint4& q, float* f) {
  f[0] = __uint_as_float(q.x << 16); f[1] = __uint_as_float(q.x & 0xFFFF0000u);
  f[2] = __uint_as_float(q.y << 16); f[3] = __uint_as_float(q.y & 0xFFFF0000u);
  f[4] = __uint_as_float(q.z << 16); f[5] = __uint_as_float(q.z & 0xFFFF0000u);
  f[6] = __uint_as_float(q.w << 16); f[7] = __uint_as_float(q.w & 0xFFFF0000u);
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
constexpr int ATT_TG = ATT_NT / 32;      // time groups

__global__ void __launch_bounds__(ATT_NT)
attn_pool_fwd_v8_kernel(const __nv_bfloat16* __restrict__ hout, const __nv_bfloat16* __restrict__ proj,
                        const int* __restrict__ mask, const float* __restrict__ w2, float* __restrict__ enc_a,
                        float* __restrict__ enc_b, __nv_bfloat16* enc_a_bf, __nv_bfloat16* enc_b_bf,
                        float* __restrict__ alphas, int L, int H, int group, int split) {
  extern __shared__ float sm[];   // sc[L] | red[ATT_TG][H]
  float* sc = sm;
  float* red = sm + L;
  const int x = blockIdx.x;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int chunks = H / 8;
  const uint4* pj = reinterpret_cast<const uint4*>(proj + (long long)x * L * H);
  const uint4* hp = reinterpret_cast<const uint4*>(hout + (long long)x * L * H);
  for (int t = warp; t < L; t += ATT_TG) {
    float s = 0.f;
    for (int c = lane; c < chunks; c += 32) {
      float f[8];
      unpack8(pj[(long long)t * chunks + c], f);
      const float* w = w2 + c * 8;          // parameter views are only 4-byte aligned
#pragma unroll
      for (int j = 0; j < 8; ++j) s = fmaf(f[j], w[j], s);
    }
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
  for (int c = lane; c < chunks; c += 32) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int t = warp; t < L; t += ATT_TG) {
      float f[8];
      unpack8(hp[(long long)t * chunks + c], f);
      const float a = sc[t];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = fmaf(a, f[j], acc[j]);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) red[warp * H + c * 8 + j] = acc[j];
  }
  __syncthreads();
  float* enc = split_row(enc_a, enc_b, x, group, split, H);
  __nv_bfloat16* enc_bf = enc_a_bf ? split_row(enc_a_bf, enc_b_bf, x, group, split, H) : nullptr;
  for (int h = threadIdx.x; h < H; h += ATT_NT) {
    float acc = 0.f;
#pragma unroll
    for (int g = 0; g < ATT_TG; ++g) acc += red[g * H + h];
    enc[h] = acc;
    if (enc_bf) enc_bf[h] = __float2bfloat16(acc);      // operand copy for the bf16 GEMMs that follow
  }
}

__global__ void __launch_bounds__(ATT_NT)
attn_pool_bwd_v8_kernel(const float* __restrict__ denc_a, const float* __restrict__ denc_b,
                        const __nv_bfloat16* __restrict__ hout, const __nv_bfloat16* proj /* may alias dpre */,
                        const float* __restrict__ alphas, const float* __restrict__ w2, float* __restrict__ dhout,
                        __nv_bfloat16* dpre, float* __restrict__ dw2_part, float* __restrict__ dsum_part,
                        float* __restrict__ dw2_sum, float* __restrict__ dsum_sum, int L, int H, int group, int split) {
  extern __shared__ float sm[];   // dsc[L] | al[L] | red_w[ATT_TG][H] | red_s[ATT_TG][H]
  float* dsc = sm;
  float* al = sm + L;
  float* red_w = sm + 2 * L;
  float* red_s = red_w + ATT_TG * H;
  const int x = blockIdx.x;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int chunks = H / 8;
  const uint4* hp = reinterpret_cast<const uint4*>(hout + (long long)x * L * H);
  const uint4* pj = reinterpret_cast<const uint4*>(proj + (long long)x * L * H);
  uint4* dp4 = reinterpret_cast<uint4*>(dpre + (long long)x * L * H);
  const float* de = split_row(denc_a, denc_b, x, group, split, H);
  for (int t = warp; t < L; t += ATT_TG) {
    float s = 0.f;
    for (int c = lane; c < chunks; c += 32) {
      float f[8];
      unpack8(hp[(long long)t * chunks + c], f);
      const float4 da = *reinterpret_cast<const float4*>(de + c * 8), db = *reinterpret_cast<const float4*>(de + c * 8 + 4);
      s += f[0] * da.x + f[1] * da.y + f[2] * da.z + f[3] * da.w + f[4] * db.x + f[5] * db.y + f[6] * db.z + f[7] * db.w;
    }
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
  for (int c = lane; c < chunks; c += 32) {
    float w[8], g[8], dw[8], ds[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { w[j] = w2[c * 8 + j]; g[j] = de[c * 8 + j]; dw[j] = 0.f; ds[j] = 0.f; }
    for (int t = warp; t < L; t += ATT_TG) {
      float pr[8], dp[8];
      unpack8(pj[(long long)t * chunks + c], pr);
      const float d = dsc[t];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        dp[j] = d * w[j] * (1.0f - pr[j] * pr[j]);
        dw[j] = fmaf(d, pr[j], dw[j]);
        ds[j] += dp[j];
      }
      dp4[(long long)t * chunks + c] = make_uint4(pack2(dp[0], dp[1]), pack2(dp[2], dp[3]), pack2(dp[4], dp[5]), pack2(dp[6], dp[7]));
      if (dhout) {
        const float a = al[t];
        float4* o = reinterpret_cast<float4*>(dhout + ((long long)x * L + t) * H + c * 8);
        o[0] = make_float4(a * g[0], a * g[1], a * g[2], a * g[3]);
        o[1] = make_float4(a * g[4], a * g[5], a * g[6], a * g[7]);
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) { red_w[warp * H + c * 8 + j] = dw[j]; red_s[warp * H + c * 8 + j] = ds[j]; }
  }
  __syncthreads();
  for (int h = threadIdx.x; h < H; h += ATT_NT) {
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int g = 0; g < ATT_TG; ++g) { a += red_w[g * H + h]; b += red_s[g * H + h]; }
    if (dw2_part) dw2_part[(long long)x * H + h] = a;
    if (dsum_part) dsum_part[(long long)x * H + h] = b;   // Dense-bias gradient partial, summed in fp32
    if (dw2_sum) { atomicAdd(dw2_sum + h, a); atomicAdd(dsum_sum + h, b); }   // or summed right here
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
                                         const float* w_a2, float* enc_a, float* enc_b, void* enc_a_bf, void* enc_b_bf,
                                         float* alphas, int X, int L, int H, int group, int split) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(X >= 0 && L > 0 && H > 0 && L <= 8192, "attn_pool_fwd_bf16: bad sizes X=%d L=%d H=%d", X, L, H);
  IN_REQUIRE(!enc_a_bf || split == group || enc_b_bf, "attn_pool_fwd_bf16: second bf16 output missing");
  if (X == 0) return INDNET_OK;
  IN_REQUIRE(hout_bf && proj_bf && mask && w_a2 && enc_a && alphas, "attn_pool_fwd_bf16: null pointer");
  IN_REQUIRE(group > 0 && split >= 0 && split <= group && X % group == 0, "attn_pool_fwd_bf16: bad group/split %d/%d", group, split);
  IN_REQUIRE(split == group || enc_b, "attn_pool_fwd_bf16: second output missing");
  const bool v8 = H % 8 == 0 && (((uintptr_t)hout_bf | (uintptr_t)proj_bf) & 15) == 0 &&
                  sizeof(float) * (L + ATT_TG * H) <= 48 * 1024;
  if (v8) {
    attn_pool_fwd_v8_kernel<<<X, ATT_NT, sizeof(float) * (L + ATT_TG * H), st>>>(
        (const __nv_bfloat16*)hout_bf, (const __nv_bfloat16*)proj_bf, mask, w_a2, enc_a, enc_b,
        (__nv_bfloat16*)enc_a_bf, (__nv_bfloat16*)enc_b_bf, alphas, L, H, group, split);
    IN_LAUNCH_CHECK("attn_pool_fwd_v8_kernel");
    return INDNET_OK;
  }
  attn_pool_fwd_kernel<__nv_bfloat16><<<X, ATT_NT, sizeof(float) * L, st>>>(
      (const __nv_bfloat16*)hout_bf, (const __nv_bfloat16*)proj_bf, mask, w_a2, enc_a, enc_b, alphas, L, H, group, split);
  IN_LAUNCH_CHECK("attn_pool_fwd_kernel<bf16>");
  if (enc_a_bf) {                  // generic path: the copies are a separate cast
    const long long na = (long long)(X / group) * split * H, nb = (long long)(X / group) * (group - split) * H;
    int rc = indnet_cast_bf16(st_, enc_a, enc_a_bf, na);
    if (rc) return rc;
    if (nb > 0) { rc = indnet_cast_bf16(st_, enc_b, enc_b_bf, nb); if (rc) return rc; }
  }
  return INDNET_OK;
}

extern "C" int indnet_attn_pool_bwd_bf16(indnet_stream st_, const float* denc_a, const float* denc_b,
                                         const void* hout_bf, const void* proj_bf, const float* alphas,
                                         const float* w_a2, float* dhout, void* dpre_bf, float* dw2_part,
                                         float* dsum_part, float* dw2_sum, float* dsum_sum, int X, int L, int H,
                                         int group, int split) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(X >= 0 && L > 0 && H > 0 && L <= 4096, "attn_pool_bwd_bf16: bad sizes X=%d L=%d H=%d", X, L, H);
  if (X == 0) return INDNET_OK;
  IN_REQUIRE((dw2_sum == nullptr) == (dsum_sum == nullptr), "attn_pool_bwd_bf16: pass both summed outputs or neither");
  if (dw2_sum) {
    IN_CUDA(cudaMemsetAsync(dw2_sum, 0, sizeof(float) * H, st));
    IN_CUDA(cudaMemsetAsync(dsum_sum, 0, sizeof(float) * H, st));
  }
  IN_REQUIRE(denc_a && hout_bf && proj_bf && alphas && w_a2 && dpre_bf && (dw2_part || dw2_sum), "attn_pool_bwd_bf16: null pointer");
  IN_REQUIRE(group > 0 && split >= 0 && split <= group && X % group == 0, "attn_pool_bwd_bf16: bad group/split %d/%d", group, split);
  IN_REQUIRE(split == group || denc_b, "attn_pool_bwd_bf16: second gradient input missing");
  const size_t v8_smem = sizeof(float) * (2 * L + 2 * ATT_TG * H);
  const bool v8 = H % 8 == 0 && v8_smem <= 48 * 1024 &&
                  (((uintptr_t)hout_bf | (uintptr_t)proj_bf | (uintptr_t)dpre_bf | (uintptr_t)denc_a |
                    (uintptr_t)denc_b | (uintptr_t)dhout) & 15) == 0;
  if (v8) {
    attn_pool_bwd_v8_kernel<<<X, ATT_NT, v8_smem, st>>>(denc_a, denc_b, (const __nv_bfloat16*)hout_bf,
                                                        (const __nv_bfloat16*)proj_bf, alphas, w_a2, dhout,
                                                        (__nv_bfloat16*)dpre_bf, dw2_part, dsum_part, dw2_sum, dsum_sum,
                                                        L, H, group, split);
    IN_LAUNCH_CHECK("attn_pool_bwd_v8_kernel");
    return INDNET_OK;
  }
  IN_REQUIRE(dw2_sum == nullptr, "attn_pool_bwd_bf16: summed outputs need H %% 8 == 0 and 16-byte aligned buffers");
  attn_pool_bwd_kernel<__nv_bfloat16><<<X, ATT_NT, sizeof(float) * 2 * L, st>>>(
      denc_a, denc_b, (const __nv_bfloat16*)hout_bf, (const __nv_bfloat16*)proj_bf, alphas, w_a2, dhout,
      (__nv_bfloat16*)dpre_bf, dw2_part, dsum_part, L, H, group, split);
  IN_LAUNCH_CHECK("attn_pool_bwd_kernel<bf16>");
  return INDNET_OK;
}
