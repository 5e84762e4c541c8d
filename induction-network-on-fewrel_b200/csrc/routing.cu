// a6/a7: the induction module, model/layer_module.py:101-162, one warp per (episode, class).
//   e_hat = e_s * W_s is loop invariant (the reference rebuilds it in each unrolled iteration, :132)
//   and is produced once by the GEMM path; this kernel runs the routing iterations:
//   d = softmax_K(b) (:127) ; c_hat = sum_k d_k e_hat_k (:136) ; c = squash(c_hat) (:139,153-162) ;
//   b += e_hat . c (:143-147).  The coupling logits live one-per-lane (K <= 32) and every reduction is
//   a warp shuffle.  The backward recomputes the iterations and back-propagates through all of them
//   (TF autodiff has no stop_gradient on this path).
#include <cuda_bf16.h>
#include "common.cuh"
#include "indnet_internal.h"
#include "../../include/indnet.h"

namespace indnet {

constexpr int RT_WARPS = 4;
constexpr float SQ_EPS = 1e-9f;

__device__ __forceinline__ float squash_factor(float s) { return s / (1.0f + s) / sqrtf(s + SQ_EPS); }
// d/ds [ s / (1+s) / sqrt(s+eps) ] written without 1/s
__device__ __forceinline__ float squash_factor_grad(float s) {
  const float r = sqrtf(s + SQ_EPS);
  return 1.0f / ((1.0f + s) * (1.0f + s) * r) - 0.5f * s / ((1.0f + s) * (s + SQ_EPS) * r);
}

__global__ void __launch_bounds__(RT_WARPS * 32)
routing_fwd_kernel(const float* __restrict__ ehat, const float* __restrict__ b_in, float* __restrict__ cvec,
                   __nv_bfloat16* __restrict__ cvec_bf, int BC, int K, int H, int iters) {
  extern __shared__ float sm[];
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int w = blockIdx.x * RT_WARPS + warp;
  if (w >= BC) return;   // whole warp leaves together; no block-level barrier below
  float* cv = sm + warp * (H + 32);
  float* dk = cv + H;
  const float* __restrict__ e = ehat + (long long)w * K * H;
  float b = (lane < K) ? (b_in ? b_in[(long long)w * K + lane] : 0.f) : -INFINITY;
  for (int it = 0; it < iters; ++it) {
    const float mx = warp_max(b);
    const float ex = (lane < K) ? expf(b - mx) : 0.f;
    const float sum = warp_sum(ex);
    dk[lane] = ex / sum;
    __syncwarp();
    float n2 = 0.f;
    for (int h = lane; h < H; h += 32) {
      float acc = 0.f;
      for (int k = 0; k < K; ++k) acc = fmaf(dk[k], e[(long long)k * H + h], acc);
      cv[h] = acc;
      n2 = fmaf(acc, acc, n2);
    }
    n2 = warp_sum(n2);
    const float f = squash_factor(n2);
    for (int h = lane; h < H; h += 32) cv[h] *= f;
    __syncwarp();
    if (it + 1 < iters) {
      for (int k = 0; k < K; ++k) {
        float dot = 0.f;
        for (int h = lane; h < H; h += 32) dot = fmaf(e[(long long)k * H + h], cv[h], dot);
        dot = warp_sum(dot);
        if (lane == k) b += dot;
      }
    }
    __syncwarp();
  }
  for (int h = lane; h < H; h += 32) {
    cvec[(long long)w * H + h] = cv[h];
    if (cvec_bf) cvec_bf[(long long)w * H + h] = __float2bfloat16(cv[h]);   // operand copy for the bf16 GEMMs
  }
}

__global__ void __launch_bounds__(RT_WARPS * 32)
routing_bwd_kernel(const float* __restrict__ ehat, const float* __restrict__ b_in,
                   const float* __restrict__ dcvec, float* __restrict__ dehat, __nv_bfloat16* __restrict__ dehat_bf,
                   int BC, int K, int H, int iters) {
  extern __shared__ float sm[];
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int w = blockIdx.x * RT_WARPS + warp;
  if (w >= BC) return;
  // per-warp layout: chat[iters][H], cs[iters][H], gc[H], dch[H], d[iters][32], dbs[32], sn[iters rounded to 32]
  const int per_warp = (2 * iters + 2) * H + (iters + 2) * 32;
  float* chat = sm + (long long)warp * per_warp;
  float* cs = chat + iters * H;
  float* gc = cs + iters * H;
  float* dch = gc + H;
  float* dsm = dch + H;
  float* dbs = dsm + iters * 32;
  float* sn = dbs + 32;
  const float* __restrict__ e = ehat + (long long)w * K * H;
  float* __restrict__ de = dehat + (long long)w * K * H;

  // ---- recompute the forward iterations, keeping d, c_hat, c and |c_hat|^2 of each
  float b = (lane < K) ? (b_in ? b_in[(long long)w * K + lane] : 0.f) : -INFINITY;
  for (int it = 0; it < iters; ++it) {
    const float mx = warp_max(b);
    const float ex = (lane < K) ? expf(b - mx) : 0.f;
    const float sum = warp_sum(ex);
    dsm[it * 32 + lane] = ex / sum;
    __syncwarp();
    float n2 = 0.f;
    for (int h = lane; h < H; h += 32) {
      float acc = 0.f;
      for (int k = 0; k < K; ++k) acc = fmaf(dsm[it * 32 + k], e[(long long)k * H + h], acc);
      chat[it * H + h] = acc;
      n2 = fmaf(acc, acc, n2);
    }
    n2 = warp_sum(n2);
    if (lane == 0) sn[it] = n2;
    const float f = squash_factor(n2);
    for (int h = lane; h < H; h += 32) cs[it * H + h] = chat[it * H + h] * f;
    __syncwarp();
    if (it + 1 < iters) {
      for (int k = 0; k < K; ++k) {
        float dot = 0.f;
        for (int h = lane; h < H; h += 32) dot = fmaf(e[(long long)k * H + h], cs[it * H + h], dot);
        dot = warp_sum(dot);
        if (lane == k) b += dot;
      }
    }
    __syncwarp();
  }

  // ---- reverse sweep
  float dbn = 0.f;           // gradient w.r.t. b^{r+1}_lane
  dbs[lane] = 0.f;
  __syncwarp();
  for (int r = iters - 1; r >= 0; --r) {
    const bool last = (r == iters - 1);
    float vg = 0.f;
    for (int h = lane; h < H; h += 32) {
      float g = last ? dcvec[(long long)w * H + h] : 0.f;
      if (!last)
        for (int k = 0; k < K; ++k) g = fmaf(dbs[k], e[(long long)k * H + h], g);
      gc[h] = g;
      vg = fmaf(chat[r * H + h], g, vg);
    }
    vg = warp_sum(vg);
    const float s = sn[r];
    const float f = squash_factor(s);
    const float coef = 2.0f * squash_factor_grad(s) * vg;
    for (int h = lane; h < H; h += 32) dch[h] = f * gc[h] + chat[r * H + h] * coef;
    __syncwarp();
    float dd = 0.f;
    for (int k = 0; k < K; ++k) {
      float dot = 0.f;
      for (int h = lane; h < H; h += 32) dot = fmaf(dch[h], e[(long long)k * H + h], dot);
      dot = warp_sum(dot);
      if (lane == k) dd = dot;
    }
    const float d = (lane < K) ? dsm[r * 32 + lane] : 0.f;
    const float dsum = warp_sum(d * dd);
    const float db = (lane < K) ? (d * (dd - dsum) + dbn) : 0.f;
    for (int k = 0; k < K; ++k) {
      const float dkk = dsm[r * 32 + k];
      const float dbk = dbs[k];
      for (int h = lane; h < H; h += 32) {
        const float val = dkk * dch[h] + dbk * cs[r * H + h];
        const long long o = (long long)k * H + h;
        const float tot = last ? val : de[o] + val;
        de[o] = tot;
        if (dehat_bf && r == 0) dehat_bf[(long long)w * K * H + o] = __float2bfloat16(tot);   // final value
      }
    }
    __syncwarp();
    dbn = db;
    dbs[lane] = db;
    __syncwarp();
  }
}

// ---- fast path (H a multiple of 128, at most 4 iterations): still one warp per (episode, class), but e_hat is
// staged ONCE into the warp's slice of shared memory with 16-byte loads and everything else lives in registers
// (lane owns the features {128 j + 4 lane + i}); the generic kernels above re-read e_hat from L2 inside every
// loop, which made them pure load latency (30-60 us for a few KB of data).
constexpr int RT_MAX_IT = 4;

template <int NV4>
__device__ __forceinline__ void rt_stage(const float* __restrict__ e, float4* se, int K, int lane) {
  const float4* e4 = reinterpret_cast<const float4*>(e);
  for (int i = lane; i < K * NV4 * 32; i += 32) se[i] = e4[i];
  __syncwarp();
}
// dots[k] = sum_h e[k][h] * v[h] for all k; lane k ends up holding dots[k] (others 0)
template <int NV4>
__device__ __forceinline__ float rt_dots(const float4* se, const float4* v, int K, int lane) {
  float mine = 0.f;
#pragma unroll 4
  for (int k = 0; k < K; ++k) {
    float dot = 0.f;
#pragma unroll
    for (int j = 0; j < NV4; ++j) {
      const float4 a = se[(k * NV4 + j) * 32 + lane];
      dot += a.x * v[j].x + a.y * v[j].y + a.z * v[j].z + a.w * v[j].w;
    }
    dot = warp_sum(dot);
    if (lane == k) mine = dot;
  }
  return mine;
}
template <int NV4>
__device__ __forceinline__ float rt_norm2(const float4* v) {
  float n2 = 0.f;
#pragma unroll
  for (int j = 0; j < NV4; ++j) n2 += v[j].x * v[j].x + v[j].y * v[j].y + v[j].z * v[j].z + v[j].w * v[j].w;
  return warp_sum(n2);
}
// acc[h] = sum_k coef_k e[k][h], coef_k held by lane k
template <int NV4>
__device__ __forceinline__ void rt_combine(const float4* se, float coef, float4* acc, int K, int lane) {
#pragma unroll
  for (int j = 0; j < NV4; ++j) acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int k = 0; k < K; ++k) {
    const float ck = __shfl_sync(0xffffffffu, coef, k);
#pragma unroll
    for (int j = 0; j < NV4; ++j) {
      const float4 a = se[(k * NV4 + j) * 32 + lane];
      acc[j].x = fmaf(ck, a.x, acc[j].x); acc[j].y = fmaf(ck, a.y, acc[j].y);
      acc[j].z = fmaf(ck, a.z, acc[j].z); acc[j].w = fmaf(ck, a.w, acc[j].w);
    }
  }
}
__device__ __forceinline__ float rt_softmax(float b, int K, int lane) {
  const float mx = warp_max(b);
  const float ex = (lane < K) ? expf(b - mx) : 0.f;
  return ex / warp_sum(ex);
}

template <int NV4>
__global__ void __launch_bounds__(RT_WARPS * 32)
routing_fwd_fast_kernel(const float* __restrict__ ehat, const float* __restrict__ b_in, float* __restrict__ cvec,
                        __nv_bfloat16* __restrict__ cvec_bf, int BC, int K, int iters, int warps) {
  extern __shared__ float4 sm4[];
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int w = blockIdx.x * warps + warp;
  if (warp >= warps || w >= BC) return;
  constexpr int H = NV4 * 128;
  float4* se = sm4 + (size_t)warp * K * NV4 * 32;
  rt_stage<NV4>(ehat + (long long)w * K * H, se, K, lane);
  float b = (lane < K) ? (b_in ? b_in[(long long)w * K + lane] : 0.f) : -INFINITY;
  float4 cv[NV4];
  for (int it = 0; it < iters; ++it) {
    const float d = rt_softmax(b, K, lane);
    rt_combine<NV4>(se, d, cv, K, lane);
    const float f = squash_factor(rt_norm2<NV4>(cv));
#pragma unroll
    for (int j = 0; j < NV4; ++j) { cv[j].x *= f; cv[j].y *= f; cv[j].z *= f; cv[j].w *= f; }
    if (it + 1 < iters) b += rt_dots<NV4>(se, cv, K, lane);        // lanes >= K add 0 to -inf
  }
#pragma unroll
  for (int j = 0; j < NV4; ++j) {
    const long long o = (long long)w * H + j * 128 + lane * 4;
    *reinterpret_cast<float4*>(cvec + o) = cv[j];
    if (cvec_bf) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(cv[j].x, cv[j].y), hi = __floats2bfloat162_rn(cv[j].z, cv[j].w);
      *reinterpret_cast<uint2*>(cvec_bf + o) = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
    }
  }
}

// The backward keeps every iteration's c_hat, c, d and |c_hat|^2 in the warp's shared-memory slice and runs both
// sweeps as real loops: fully unrolled over the iterations with everything in registers it was ~100 KB of straight-line
// code executed once by a handful of warps, i.e. bound by instruction fetch (ncu: 64 % of the stall samples "no
// instruction", 21 us for 20 classes).
template <int NV4>
__global__ void __launch_bounds__(RT_WARPS * 32)
routing_bwd_fast_kernel(const float* __restrict__ ehat, const float* __restrict__ b_in, const float* __restrict__ dcvec,
                        float* __restrict__ dehat, __nv_bfloat16* __restrict__ dehat_bf, int BC, int K, int iters,
                        int warps) {
  extern __shared__ float4 sm4[];
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int w = blockIdx.x * warps + warp;
  if (warp >= warps || w >= BC) return;
  constexpr int H = NV4 * 128;
  const size_t slice = (size_t)2 * K * NV4 * 32 + (size_t)iters * (2 * NV4 * 32 + 16);     // in float4
  float4* se = sm4 + (size_t)warp * slice;                 // e_hat, then its gradient accumulator
  float4* sde = se + (size_t)K * NV4 * 32;
  float4* schat = sde + (size_t)K * NV4 * 32;              // [iters][NV4][32]
  float4* scs = schat + (size_t)iters * NV4 * 32;          // [iters][NV4][32]
  float* sdsm = reinterpret_cast<float*>(scs + (size_t)iters * NV4 * 32);   // [iters][32] coupling coefficients
  float* ssn = sdsm + iters * 32;                          // [iters] |c_hat|^2
  rt_stage<NV4>(ehat + (long long)w * K * H, se, K, lane);
  // ---- recompute the forward iterations
  float b = (lane < K) ? (b_in ? b_in[(long long)w * K + lane] : 0.f) : -INFINITY;
#pragma unroll 1
  for (int it = 0; it < iters; ++it) {
    const float d = rt_softmax(b, K, lane);
    float4 chat[NV4], cs[NV4];
    rt_combine<NV4>(se, d, chat, K, lane);
    const float sn = rt_norm2<NV4>(chat);
    const float f = squash_factor(sn);
#pragma unroll
    for (int j = 0; j < NV4; ++j) {
      cs[j] = make_float4(chat[j].x * f, chat[j].y * f, chat[j].z * f, chat[j].w * f);
      schat[(it * NV4 + j) * 32 + lane] = chat[j];
      scs[(it * NV4 + j) * 32 + lane] = cs[j];
    }
    sdsm[it * 32 + lane] = d;
    if (lane == 0) ssn[it] = sn;
    if (it + 1 < iters) b += rt_dots<NV4>(se, cs, K, lane);
  }
  __syncwarp();
  // ---- reverse sweep (autodiff through every iteration: no stop_gradient in the reference)
  float dbs = 0.f;           // gradient w.r.t. b^{r+1}_lane
#pragma unroll 1
  for (int r = iters - 1; r >= 0; --r) {
    const bool last = (r == iters - 1);
    float4 g[NV4], dch[NV4], chat[NV4], cs[NV4];
#pragma unroll
    for (int j = 0; j < NV4; ++j) { chat[j] = schat[(r * NV4 + j) * 32 + lane]; cs[j] = scs[(r * NV4 + j) * 32 + lane]; }
    if (last) {
#pragma unroll
      for (int j = 0; j < NV4; ++j) g[j] = *reinterpret_cast<const float4*>(dcvec + (long long)w * H + j * 128 + lane * 4);
    } else {
      rt_combine<NV4>(se, dbs, g, K, lane);
    }
    float vg = 0.f;
#pragma unroll
    for (int j = 0; j < NV4; ++j)
      vg += chat[j].x * g[j].x + chat[j].y * g[j].y + chat[j].z * g[j].z + chat[j].w * g[j].w;
    vg = warp_sum(vg);
    const float sn = ssn[r];
    const float f = squash_factor(sn);
    const float coef = 2.0f * squash_factor_grad(sn) * vg;
#pragma unroll
    for (int j = 0; j < NV4; ++j)
      dch[j] = make_float4(f * g[j].x + chat[j].x * coef, f * g[j].y + chat[j].y * coef,
                           f * g[j].z + chat[j].z * coef, f * g[j].w + chat[j].w * coef);
    const float dd = rt_dots<NV4>(se, dch, K, lane);
    const float d = (lane < K) ? sdsm[r * 32 + lane] : 0.f;
    const float dsum = warp_sum(d * dd);
    const float db = (lane < K) ? (d * (dd - dsum) + dbs) : 0.f;
#pragma unroll 1
    for (int k = 0; k < K; ++k) {
      const float dkk = __shfl_sync(0xffffffffu, d, k), dbk = __shfl_sync(0xffffffffu, dbs, k);
#pragma unroll
      for (int j = 0; j < NV4; ++j) {
        float4 val = make_float4(dkk * dch[j].x + dbk * cs[j].x, dkk * dch[j].y + dbk * cs[j].y,
                                 dkk * dch[j].z + dbk * cs[j].z, dkk * dch[j].w + dbk * cs[j].w);
        float4* dst = &sde[(k * NV4 + j) * 32 + lane];
        if (!last) { const float4 o = *dst; val.x += o.x; val.y += o.y; val.z += o.z; val.w += o.w; }
        *dst = val;
      }
    }
    dbs = db;
  }
  __syncwarp();
  float4* de4 = reinterpret_cast<float4*>(dehat + (long long)w * K * H);
  for (int i = lane; i < K * NV4 * 32; i += 32) {
    const float4 v = sde[i];
    de4[i] = v;
    if (dehat_bf) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
      reinterpret_cast<uint2*>(dehat_bf + (long long)w * K * H)[i] =
          make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
    }
  }
}

// warps per CTA (<= RT_WARPS) whose shared-memory slices fit; 0 = the fast path does not apply
static size_t rt_fast_slice(int K, int H, int iters, int slices) {      // bytes of shared memory per warp
  return (size_t)slices * K * H * sizeof(float) + (slices == 2 ? (size_t)iters * (2 * H * sizeof(float) + 256) : 0);
}
static int rt_fast_warps(int K, int H, int iters, int slices, const void* a, const void* b, const void* c, const void* d) {
  if (H % 128 != 0 || H > 512 || iters > RT_MAX_IT) return 0;
  if ((((uintptr_t)a | (uintptr_t)b | (uintptr_t)c) & 15) != 0 || ((uintptr_t)d & 7) != 0) return 0;
  const size_t per_warp = rt_fast_slice(K, H, iters, slices);
  int w = (int)((200 * 1024) / per_warp);
  return w > RT_WARPS ? RT_WARPS : w;
}

__global__ void squash_fwd_kernel(const float* __restrict__ v, float* __restrict__ out, int rows, int H) {
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) / 32, lane = threadIdx.x % 32;
  if (w >= rows) return;
  float n2 = 0.f;
  for (int h = lane; h < H; h += 32) { float a = v[(long long)w * H + h]; n2 = fmaf(a, a, n2); }
  n2 = warp_sum(n2);
  const float f = squash_factor(n2);
  for (int h = lane; h < H; h += 32) out[(long long)w * H + h] = v[(long long)w * H + h] * f;
}

__global__ void squash_bwd_kernel(const float* __restrict__ v, const float* __restrict__ dout,
                                  float* __restrict__ dv, int rows, int H) {
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) / 32, lane = threadIdx.x % 32;
  if (w >= rows) return;
  float n2 = 0.f, vg = 0.f;
  for (int h = lane; h < H; h += 32) {
    float a = v[(long long)w * H + h];
    n2 = fmaf(a, a, n2);
    vg = fmaf(a, dout[(long long)w * H + h], vg);
  }
  n2 = warp_sum(n2);
  vg = warp_sum(vg);
  const float f = squash_factor(n2), coef = 2.0f * squash_factor_grad(n2) * vg;
  for (int h = lane; h < H; h += 32)
    dv[(long long)w * H + h] = f * dout[(long long)w * H + h] + v[(long long)w * H + h] * coef;
}

}  // namespace indnet

using namespace indnet;

extern "C" int indnet_routing_fwd(indnet_stream st_, const float* ehat, const float* b_in, float* cvec, void* cvec_bf,
                                  int BC, int K, int H, int iters) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(BC >= 0 && K > 0 && H > 0 && iters > 0, "routing_fwd: bad sizes BC=%d K=%d H=%d iters=%d", BC, K, H, iters);
  IN_REQUIRE(K <= 32, "routing_fwd: K=%d support instances per class exceeds the warp-per-class limit of 32", K);
  if (BC == 0) return INDNET_OK;
  IN_REQUIRE(ehat && cvec, "routing_fwd: null pointer");
  if (const int fw = rt_fast_warps(K, H, iters, 1, ehat, cvec, nullptr, cvec_bf)) {
    const size_t fsm = (size_t)fw * K * H * sizeof(float);
#define RT_FWD_FAST(NV4_)                                                                                          \
    do {                                                                                                           \
      IN_CUDA(cudaFuncSetAttribute(routing_fwd_fast_kernel<NV4_>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)); \
      routing_fwd_fast_kernel<NV4_><<<ceil_div(BC, fw), RT_WARPS * 32, fsm, st>>>(ehat, b_in, cvec, (__nv_bfloat16*)cvec_bf, \
                                                                                 BC, K, iters, fw);                \
    } while (0)
    if (H == 128) RT_FWD_FAST(1); else if (H == 256) RT_FWD_FAST(2); else if (H == 384) RT_FWD_FAST(3); else RT_FWD_FAST(4);
#undef RT_FWD_FAST
    IN_LAUNCH_CHECK("routing_fwd_fast_kernel");
    return INDNET_OK;
  }
  size_t smem = sizeof(float) * RT_WARPS * (H + 32);
  IN_REQUIRE(smem <= 48 * 1024, "routing_fwd: H=%d too large", H);
  routing_fwd_kernel<<<ceil_div(BC, RT_WARPS), RT_WARPS * 32, smem, st>>>(ehat, b_in, cvec, (__nv_bfloat16*)cvec_bf, BC, K, H, iters);
  IN_LAUNCH_CHECK("routing_fwd_kernel");
  return INDNET_OK;
}

extern "C" int indnet_routing_bwd(indnet_stream st_, const float* ehat, const float* b_in, const float* dcvec,
                                  float* dehat, void* dehat_bf, int BC, int K, int H, int iters) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(BC >= 0 && K > 0 && H > 0 && iters > 0, "routing_bwd: bad sizes BC=%d K=%d H=%d iters=%d", BC, K, H, iters);
  IN_REQUIRE(K <= 32, "routing_bwd: K=%d support instances per class exceeds the warp-per-class limit of 32", K);
  IN_REQUIRE(iters <= 30, "routing_bwd: too many routing iterations (%d)", iters);
  if (BC == 0) return INDNET_OK;
  IN_REQUIRE(ehat && dcvec && dehat, "routing_bwd: null pointer");
  if (const int fw = rt_fast_warps(K, H, iters, 2, ehat, dcvec, dehat, dehat_bf)) {
    const size_t fsm = (size_t)fw * rt_fast_slice(K, H, iters, 2);
#define RT_BWD_FAST(NV4_)                                                                                          \
    do {                                                                                                           \
      IN_CUDA(cudaFuncSetAttribute(routing_bwd_fast_kernel<NV4_>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)); \
      routing_bwd_fast_kernel<NV4_><<<ceil_div(BC, fw), RT_WARPS * 32, fsm, st>>>(ehat, b_in, dcvec, dehat,           \
                                                                                 (__nv_bfloat16*)dehat_bf, BC, K, iters, fw); \
    } while (0)
    if (H == 128) RT_BWD_FAST(1); else if (H == 256) RT_BWD_FAST(2); else if (H == 384) RT_BWD_FAST(3); else RT_BWD_FAST(4);
#undef RT_BWD_FAST
    IN_LAUNCH_CHECK("routing_bwd_fast_kernel");
    return INDNET_OK;
  }
  size_t smem = sizeof(float) * RT_WARPS * ((size_t)(2 * iters + 2) * H + (iters + 2) * 32);
  IN_REQUIRE(smem <= 227 * 1024, "routing_bwd: H=%d x iters=%d does not fit shared memory", H, iters);
  if (smem > 48 * 1024)
    IN_CUDA(cudaFuncSetAttribute(routing_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  routing_bwd_kernel<<<ceil_div(BC, RT_WARPS), RT_WARPS * 32, smem, st>>>(ehat, b_in, dcvec, dehat, (__nv_bfloat16*)dehat_bf, BC, K, H, iters);
  IN_LAUNCH_CHECK("routing_bwd_kernel");
  return INDNET_OK;
}

extern "C" int indnet_squash_fwd(indnet_stream st_, const float* v, float* out, int rows, int H) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(rows >= 0 && H > 0, "squash_fwd: bad sizes");
  if (rows == 0) return INDNET_OK;
  IN_REQUIRE(v && out, "squash_fwd: null pointer");
  squash_fwd_kernel<<<ceil_div((long long)rows * 32, 128), 128, 0, st>>>(v, out, rows, H);
  IN_LAUNCH_CHECK("squash_fwd_kernel");
  return INDNET_OK;
}

extern "C" int indnet_squash_bwd(indnet_stream st_, const float* v, const float* dout, float* dv, int rows, int H) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(rows >= 0 && H > 0, "squash_bwd: bad sizes");
  if (rows == 0) return INDNET_OK;
  IN_REQUIRE(v && dout && dv, "squash_bwd: null pointer");
  squash_bwd_kernel<<<ceil_div((long long)rows * 32, 128), 128, 0, st>>>(v, dout, dv, rows, H);
  IN_LAUNCH_CHECK("squash_bwd_kernel");
  return INDNET_OK;
}
