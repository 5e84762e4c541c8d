// a2: Base.get_embedding (model/base.py:76-86): three table lookups concatenated on the feature
// axis, x[t, :] = [word_table[word[t]] (E) ; pos1_table[pos1[t]] (P) ; pos2_table[pos2[t]] (P)],
// and its backward (scatter-add of dx into the trainable tables).
#include <cuda_bf16.h>
#include "common.cuh"
#include "indnet_internal.h"
#include "../../include/indnet.h"

namespace indnet {

__global__ void embed_fwd_kernel(const int* __restrict__ word, const int* __restrict__ pos1,
                                 const int* __restrict__ pos2, const float* __restrict__ wt,
                                 const float* __restrict__ p1t, const float* __restrict__ p2t,
                                 float* __restrict__ x, long long T, int E, int P) {
  const int D = E + 2 * P;
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= T * D) return;
  long long t = i / D;
  int d = (int)(i % D);
  float v;
  if (d < E) v = wt[(long long)word[t] * E + d];
  else if (d < E + P) v = p1t[(long long)pos1[t] * P + (d - E)];
  else v = p2t[(long long)pos2[t] * P + (d - E - P)];
  x[i] = v;
}

__global__ void embed_fwd_bf16_kernel(const int* __restrict__ word, const int* __restrict__ pos1,
                                      const int* __restrict__ pos2, const float* __restrict__ wt,
                                      const float* __restrict__ p1t, const float* __restrict__ p2t,
                                      __nv_bfloat16* __restrict__ xb, long long T, int E, int P) {
  // two adjacent columns per thread: one 4-byte store, half the threads; a warp covers a whole 64-column row
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= T * 32) return;
  long long t = i >> 5;
  const int d0 = (int)(i & 31) * 2;
  const long long w = word[t], q1 = pos1[t], q2 = pos2[t];
  float v[2];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int d = d0 + j;
    float x = 0.f;
    if (d < E) x = wt[w * E + d];
    else if (d < E + P) x = p1t[q1 * P + (d - E)];
    else if (d < E + 2 * P) x = p2t[q2 * P + (d - E - P)];
    else if (E + 2 * P + 2 <= 64 && d < E + 2 * P + 2) x = 1.0f;      // the two bias columns (see indnet.h)
    v[j] = x;
  }
  reinterpret_cast<__nv_bfloat162*>(xb)[i] = __floats2bfloat162_rn(v[0], v[1]);
}

// Position tables are tiny ([2L, P]) and every token hits them: accumulate in shared memory per
// block, then one atomic per table entry per block.
__global__ void embed_pos_bwd_kernel(const int* __restrict__ pos1, const int* __restrict__ pos2,
                                     const float* __restrict__ dx, const float* __restrict__ dx_b, int ld,
                                     float* __restrict__ dp1, float* __restrict__ dp2, long long T, int E, int P,
                                     int pos_rows, long long tokens_per_block) {
  extern __shared__ float sm[];  // [2][pos_rows*P]
  const int n = pos_rows * P;
  for (int i = threadIdx.x; i < 2 * n; i += blockDim.x) sm[i] = 0.f;
  __syncthreads();
  const int D = ld;                    // row stride of dx; E = first position column
  long long t0 = blockIdx.x * tokens_per_block;
  long long t1 = min(T, t0 + tokens_per_block);
  long long work = (t1 - t0) * (2 * P);
  for (long long w = threadIdx.x; w < work; w += blockDim.x) {
    long long t = t0 + w / (2 * P);
    int j = (int)(w % (2 * P));
    float g = dx[t * D + E + j];
    if (dx_b) g += dx_b[t * D + E + j];
    if (j < P) atomicAdd(&sm[pos1[t] * P + j], g);
    else atomicAdd(&sm[n + pos2[t] * P + (j - P)], g);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    float a = sm[i], b = sm[n + i];
    if (a != 0.f) atomicAdd(dp1 + i, a);
    if (b != 0.f) atomicAdd(dp2 + i, b);
  }
}

// Same scatter for the split input gradient of the fused BPTT path: dx_a / dx_b rows of 16 floats (64 B), the 2P
// position columns first.  One thread per token: three 16-byte loads per direction, then 2P shared-memory adds.
__global__ void embed_pos_bwd16_kernel(const int* __restrict__ pos1, const int* __restrict__ pos2,
                                       const float* __restrict__ dx_a, const float* __restrict__ dx_b,
                                       float* __restrict__ dp1, float* __restrict__ dp2, long long T, int P,
                                       int pos_rows, long long tokens_per_block) {
  extern __shared__ float sm[];  // [2][pos_rows*P]
  const int n = pos_rows * P;
  for (int i = threadIdx.x; i < 2 * n; i += blockDim.x) sm[i] = 0.f;
  __syncthreads();
  const long long t0 = blockIdx.x * tokens_per_block, t1 = min(T, t0 + tokens_per_block);
  for (long long t = t0 + threadIdx.x; t < t1; t += blockDim.x) {
    float g[12];
    const float4* a = reinterpret_cast<const float4*>(dx_a + t * 16);
    const float4* b = dx_b ? reinterpret_cast<const float4*>(dx_b + t * 16) : nullptr;
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      float4 v = a[q];
      if (b) { const float4 w = b[q]; v.x += w.x; v.y += w.y; v.z += w.z; v.w += w.w; }
      g[4 * q] = v.x; g[4 * q + 1] = v.y; g[4 * q + 2] = v.z; g[4 * q + 3] = v.w;
    }
    const int r1 = pos1[t] * P, r2 = n + pos2[t] * P;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      if (j < P) { atomicAdd(&sm[r1 + j], g[j]); atomicAdd(&sm[r2 + j], g[P + j]); }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    float a = sm[i], b = sm[n + i];
    if (a != 0.f) atomicAdd(dp1 + i, a);
    if (b != 0.f) atomicAdd(dp2 + i, b);
  }
}

__global__ void embed_word_bwd_kernel(const int* __restrict__ word, const float* __restrict__ dx,
                                      float* __restrict__ dwt, long long T, int E, int D) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= T * E) return;
  long long t = i / E;
  int d = (int)(i % E);
  atomicAdd(dwt + (long long)word[t] * E + d, dx[t * D + d]);
}

}  // namespace indnet

using namespace indnet;

extern "C" int indnet_embed_fwd(indnet_stream st_, const int* word, const int* pos1, const int* pos2,
                                const float* word_table, const float* pos1_table, const float* pos2_table,
                                float* x, long long T, int E, int P) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(T >= 0 && E > 0 && P > 0, "embed_fwd: bad sizes T=%lld E=%d P=%d", T, E, P);
  if (T == 0) return INDNET_OK;
  IN_REQUIRE(word && pos1 && pos2 && word_table && pos1_table && pos2_table && x, "embed_fwd: null pointer");
  long long n = T * (E + 2 * P);
  embed_fwd_kernel<<<ceil_div(n, 256), 256, 0, st>>>(word, pos1, pos2, word_table, pos1_table, pos2_table, x, T, E, P);
  IN_LAUNCH_CHECK("embed_fwd_kernel");
  return INDNET_OK;
}

extern "C" int indnet_embed_bwd(indnet_stream st_, const int* word, const int* pos1, const int* pos2,
                                const float* dx, float* dword_table, float* dpos1, float* dpos2,
                                long long T, int E, int P, int pos_rows) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(T >= 0 && E > 0 && P > 0 && pos_rows > 0, "embed_bwd: bad sizes");
  if (T == 0) return INDNET_OK;
  IN_REQUIRE(pos1 && pos2 && dx && dpos1 && dpos2, "embed_bwd: null pointer");
  size_t smem = sizeof(float) * 2 * pos_rows * P;
  IN_REQUIRE(smem <= 48 * 1024, "embed_bwd: position tables too large for the shared-memory accumulator (%zu B)", smem);
  long long tpb = 512;        // ~150 CTAs at 77k tokens; each flushes its shared-memory tables once
  int blocks = ceil_div(T, tpb);
  embed_pos_bwd_kernel<<<blocks, 256, smem, st>>>(pos1, pos2, dx, nullptr, E + 2 * P, dpos1, dpos2, T, E, P, pos_rows, tpb);
  IN_LAUNCH_CHECK("embed_pos_bwd_kernel");
  if (dword_table) {
    IN_REQUIRE(word != nullptr, "embed_bwd: word ids needed for the trainable word table");
    embed_word_bwd_kernel<<<ceil_div(T * E, 256), 256, 0, st>>>(word, dx, dword_table, T, E, E + 2 * P);
    IN_LAUNCH_CHECK("embed_word_bwd_kernel");
  }
  return INDNET_OK;
}

extern "C" int indnet_embed_pos_bwd2(indnet_stream st_, const int* pos1, const int* pos2, const float* dx_a,
                                     const float* dx_b, int ld, int col0, float* dpos1, float* dpos2, long long T,
                                     int P, int pos_rows) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(T >= 0 && P > 0 && pos_rows > 0 && ld >= col0 + 2 * P && col0 >= 0, "embed_pos_bwd2: bad sizes");
  if (T == 0) return INDNET_OK;
  IN_REQUIRE(pos1 && pos2 && dx_a && dpos1 && dpos2, "embed_pos_bwd2: null pointer");
  size_t smem = sizeof(float) * 2 * pos_rows * P;
  IN_REQUIRE(smem <= 48 * 1024, "embed_pos_bwd2: position tables too large for the shared-memory accumulator (%zu B)", smem);
  long long tpb = 512;
  int blocks = ceil_div(T, tpb);
  if (ld == 16 && col0 == 0 && P <= 6 && (((uintptr_t)dx_a | (uintptr_t)dx_b) & 15) == 0) {
    embed_pos_bwd16_kernel<<<blocks, 256, smem, st>>>(pos1, pos2, dx_a, dx_b, dpos1, dpos2, T, P, pos_rows, tpb);
    IN_LAUNCH_CHECK("embed_pos_bwd16_kernel");
    return INDNET_OK;
  }
  embed_pos_bwd_kernel<<<blocks, 256, smem, st>>>(pos1, pos2, dx_a, dx_b, ld, dpos1, dpos2, T, col0, P, pos_rows, tpb);
  IN_LAUNCH_CHECK("embed_pos_bwd_kernel");
  return INDNET_OK;
}

extern "C" int indnet_embed_fwd_bf16(indnet_stream st_, const int* word, const int* pos1, const int* pos2,
                                     const float* word_table, const float* pos1_table, const float* pos2_table,
                                     void* xb, long long T, int E, int P) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(T >= 0 && E > 0 && P > 0 && E + 2 * P <= 64, "embed_fwd_bf16: need E + 2P <= 64 (E=%d P=%d)", E, P);
  if (T == 0) return INDNET_OK;
  IN_REQUIRE(word && pos1 && pos2 && word_table && pos1_table && pos2_table && xb, "embed_fwd_bf16: null pointer");
  embed_fwd_bf16_kernel<<<ceil_div(T * 32, 256), 256, 0, st>>>(word, pos1, pos2, word_table, pos1_table, pos2_table,
                                                               (__nv_bfloat16*)xb, T, E, P);
  IN_LAUNCH_CHECK("embed_fwd_bf16_kernel");
  return INDNET_OK;
}
