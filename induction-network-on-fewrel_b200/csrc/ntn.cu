// a8-a10 (FP32 parity mode): the relation module head, model/layer_module.py:35-47, fused with
// build_loss (model/graph.py:92-94) and build_predict/accuracy (model/base.py:101-107).
//
// The bilinear part (c_i^T M_k e_q for all slices) runs as two batched GEMMs on the internal layout
// M' = M.transpose(0, 2, 1) ([H_i][S][H_j]): T'[b,c][k][j] = cv * M', then
// V[b][q][c*S + k] = qe_b * T'_b^T.  This file takes V (pre-ReLU) and applies
//   v = relu(V) ; logits[q, c'] = sum_{c,k} v[q, c*S+k] * w[(k*N + c), c'] + b[c'] ; probs = sigmoid
// - w keeps the reference's slice-major row order k*N + c (concat(axis=0) at :35, transpose at :37) -
// then the per-query squared error against the one-hot label and the first-max argmax hit.
#include <cuda_bf16.h>
#include "common.cuh"
#include "indnet_internal.h"
#include "../../include/indnet.h"

namespace indnet {

constexpr int NTN_NT = 128;

// tf.nn.relu propagates NaN (fmaxf would silently turn it into 0 and hide an upstream fault)
__device__ __forceinline__ float relu_nan(float v) { return v < 0.f ? 0.f : v; }

__global__ void __launch_bounds__(NTN_NT)
ntn_head_fwd_kernel(const float* __restrict__ V, const float* __restrict__ w, const float* __restrict__ bias,
                    const int* __restrict__ labels, float* __restrict__ probs, float* __restrict__ loss_part,
                    int* __restrict__ correct, int N, int S, float* __restrict__ out2, float* __restrict__ running,
                    unsigned int* __restrict__ ticket, float* __restrict__ dlogit, float* __restrict__ dV,
                    __nv_bfloat16* __restrict__ dV_bf, int ld_bf, float scale) {
  extern __shared__ float pr[];   // [N] probs, then [N] dlogit (training)
  __shared__ bool last;
  __shared__ float sl[NTN_NT / 32], sa[NTN_NT / 32];
  const int row = blockIdx.x;
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32, nw = NTN_NT / 32;
  const int NS = N * S;
  const float* v = V + (long long)row * NS;
  // thread t owns the V columns t, t + 128, ...: for each it reads the N-wide row of w once (contiguous) and
  // keeps up to 8 class partial sums in registers - all loads of a thread are independent, one memory round
  // trip instead of a dependent chain of NS / 32 of them per class
  __shared__ float part[NTN_NT / 32][8];
  for (int c0 = 0; c0 < N; c0 += 8) {
    const int nc = min(8, N - c0);
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    for (int r = threadIdx.x; r < NS; r += NTN_NT) {
      const float x = relu_nan(v[r]);
      const int c = r / S, k = r - c * S;
      const float* wr = w + (long long)(k * N + c) * N + c0;
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (i < nc) acc[i] = fmaf(x, wr[i], acc[i]);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = warp_sum(acc[i]);
    if (lane == 0) {
#pragma unroll
      for (int i = 0; i < 8; ++i) part[warp][i] = acc[i];
    }
    __syncthreads();
    if (threadIdx.x < nc) {
      float t = 0.f;
      for (int q = 0; q < nw; ++q) t += part[q][threadIdx.x];
      const float p = sigmoid_acc(t + bias[c0 + threadIdx.x]);
      pr[c0 + threadIdx.x] = p;
      probs[(long long)row * N + c0 + threadIdx.x] = p;
    }
    __syncthreads();
  }
  __syncthreads();
  if (dV) {
    // training: the backward of the head is local to the row (fused MSE gradient, sigmoid, FC, ReLU mask - exactly
    // ntn_head_bwd_kernel), so it is emitted here and the backward pass starts one launch further down the chain
    float* dl = pr + N;
    if (threadIdx.x < N) {
      const int c = threadIdx.x;
      const float p = pr[c];
      const float g = scale * 2.0f * (p - ((c == labels[row]) ? 1.0f : 0.0f)) * p * (1.0f - p);
      dl[c] = g;
      dlogit[(long long)row * N + c] = g;
    }
    __syncthreads();
    for (int r = threadIdx.x; r < NS; r += NTN_NT) {
      float g = 0.f;
      if (v[r] > 0.f) {
        const int c = r / S, k = r % S;
        const float* wr = w + (long long)(k * N + c) * N;
        for (int cp = 0; cp < N; ++cp) g = fmaf(wr[cp], dl[cp], g);
      }
      dV[(long long)row * NS + r] = g;
      if (dV_bf) dV_bf[(long long)row * ld_bf + r] = __float2bfloat16(g);
    }
    if (dV_bf) for (int r = NS + threadIdx.x; r < ld_bf; r += NTN_NT) dV_bf[(long long)row * ld_bf + r] = __float2bfloat16(0.f);
  }
  if (threadIdx.x == 0) {
    const int y = labels[row];
    float l = 0.f, best = pr[0];
    int arg = 0;
    for (int c = 0; c < N; ++c) {
      const float d = pr[c] - (c == y ? 1.0f : 0.0f);
      l = fmaf(d, d, l);
      if (pr[c] > best) { best = pr[c]; arg = c; }   // strict > : first maximal index wins (tf.argmax)
    }
    loss_part[row] = l;
    correct[row] = (arg == y) ? 1 : 0;
    if (ticket) {
      // the block that finishes last folds the per-row partials - in row order, whatever the block schedule,
      // so the sum is deterministic - exactly as loss_acc_reduce_kernel would in a launch of its own
      __threadfence();
      last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
  }
  if (!ticket) return;
  __syncthreads();
  if (!last) return;
  __threadfence();
  const int rows = gridDim.x;
  float l = 0.f, a = 0.f;
  for (int i = threadIdx.x; i < rows; i += NTN_NT) { l += __ldcg(loss_part + i); a += (float)__ldcg(correct + i); }
  l = warp_sum(l); a = warp_sum(a);
  if (lane == 0) { sl[warp] = l; sa[warp] = a; }
  __syncthreads();
  if (threadIdx.x == 0) {
    l = 0.f; a = 0.f;
    for (int i = 0; i < NTN_NT / 32; ++i) { l += sl[i]; a += sa[i]; }
    const float loss = l / ((float)rows * (float)N), acc = a / (float)rows;
    out2[0] = loss; out2[1] = acc;
    if (running) { running[0] += loss; running[1] += acc; running[2] += 1.0f; }
    *ticket = 0u;                   // ready for the next launch
  }
}

// loss = sum(loss_part) / (rows*N)  (tf.losses.mean_squared_error == mean);
// accuracy = mean(correct).  One block, fixed summation order => deterministic.
__global__ void loss_acc_reduce_kernel(const float* __restrict__ loss_part, const int* __restrict__ correct,
                                       float* __restrict__ out2, float* __restrict__ running, int rows, int N) {
  __shared__ float sl[32];
  __shared__ float sa[32];
  float l = 0.f, a = 0.f;
  for (int i = threadIdx.x; i < rows; i += blockDim.x) { l += loss_part[i]; a += (float)correct[i]; }
  l = warp_sum(l); a = warp_sum(a);
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  if (lane == 0) { sl[warp] = l; sa[warp] = a; }
  __syncthreads();
  if (warp == 0) {
    const int nw = blockDim.x / 32;
    l = lane < nw ? sl[lane] : 0.f;
    a = lane < nw ? sa[lane] : 0.f;
    l = warp_sum(l); a = warp_sum(a);
    if (lane == 0) {
      const float loss = l / ((float)rows * (float)N), acc = a / (float)rows;
      out2[0] = loss; out2[1] = acc;
      // running sums so a training loop can read its means every n steps instead of synchronising each step
      if (running) { running[0] += loss; running[1] += acc; running[2] += 1.0f; }
    }
  }
}

// dlogit[row, c'] = up * p (1 - p) with up = scale * 2 (p - y) (fused MSE, model/graph.py:92-94) or an
// explicit upstream gradient dprobs;  dV[row, r] = (V > 0) * sum_c' w[(k*N+c), c'] dlogit[c']
__global__ void __launch_bounds__(NTN_NT)
ntn_head_bwd_kernel(const float* __restrict__ V, const float* __restrict__ w, const float* __restrict__ probs,
                    const int* __restrict__ labels, const float* __restrict__ dprobs,
                    float* __restrict__ dlogit, float* __restrict__ dV, __nv_bfloat16* __restrict__ dV_bf, int ld_bf,
                    int N, int S, float scale) {
  extern __shared__ float dl[];   // [N]
  const int row = blockIdx.x;
  const int NS = N * S;
  if (threadIdx.x < N) {
    const int c = threadIdx.x;
    const float p = probs[(long long)row * N + c];
    float up;
    if (dprobs) up = dprobs[(long long)row * N + c];
    else up = scale * 2.0f * (p - ((c == labels[row]) ? 1.0f : 0.0f));
    const float g = up * p * (1.0f - p);
    dl[c] = g;
    dlogit[(long long)row * N + c] = g;
  }
  __syncthreads();
  const float* v = V + (long long)row * NS;
  for (int r = threadIdx.x; r < NS; r += NTN_NT) {
    float g = 0.f;
    if (v[r] > 0.f) {
      const int c = r / S, k = r % S;
      const float* wr = w + (long long)(k * N + c) * N;
      for (int cp = 0; cp < N; ++cp) g = fmaf(wr[cp], dl[cp], g);
    }
    dV[(long long)row * NS + r] = g;
    if (dV_bf) dV_bf[(long long)row * ld_bf + r] = __float2bfloat16(g);
  }
  // bf16 rows are padded to a 16-byte multiple for TMA: zero the tail
  if (dV_bf) for (int r = NS + threadIdx.x; r < ld_bf; r += NTN_NT) dV_bf[(long long)row * ld_bf + r] = __float2bfloat16(0.f);
}

// dw[(k*N+c), c'] += sum_rows relu(V[row, c*S+k]) * dlogit[row, c'] over a chunk of rows.
__global__ void ntn_dw_kernel(const float* __restrict__ V, const float* __restrict__ dlogit,
                              float* __restrict__ dw, int rows, int N, int S, int rows_per_block) {
  const int NS = N * S;
  const int id = blockIdx.x * blockDim.x + threadIdx.x;   // r * N + c'
  if (id >= NS * N) return;
  const int r = id / N, cp = id % N;
  const int r0 = blockIdx.y * rows_per_block, r1 = min(rows, r0 + rows_per_block);
  float acc = 0.f;
  for (int row = r0; row < r1; ++row)
    acc = fmaf(relu_nan(V[(long long)row * NS + r]), dlogit[(long long)row * N + cp], acc);
  const int c = r / S, k = r % S;
  atomicAdd(dw + (long long)(k * N + c) * N + cp, acc);
}

// [d0][d1][d2] -> [d0][d2][d1]
__global__ void permute021_kernel(const float* __restrict__ in, float* __restrict__ out, int d1, int d2) {
  __shared__ float tile[32][33];
  const long long base = (long long)blockIdx.z * d1 * d2;
  int c = blockIdx.x * 32 + threadIdx.x;
  for (int j = threadIdx.y; j < 32; j += 8) {
    int r = blockIdx.y * 32 + j;
    if (r < d1 && c < d2) tile[j][threadIdx.x] = in[base + (long long)r * d2 + c];
  }
  __syncthreads();
  int r2 = blockIdx.y * 32 + threadIdx.x;
  for (int j = threadIdx.y; j < 32; j += 8) {
    int c2 = blockIdx.x * 32 + j;
    if (r2 < d1 && c2 < d2) out[base + (long long)c2 * d1 + r2] = tile[threadIdx.x][j];
  }
}

}  // namespace indnet

using namespace indnet;

extern "C" int indnet_ntn_head_fwd(indnet_stream st_, const float* V, const float* w, const float* b, const int* labels,
                                   float* probs, float* loss_part, int* correct, int rows, int N, int S, float* out2,
                                   float* running, unsigned int* ticket, float* dlogit, float* dV, void* dV_bf, int ld_bf,
                                   float scale) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(rows >= 0 && N > 0 && S > 0, "ntn_head_fwd: bad sizes rows=%d N=%d S=%d", rows, N, S);
  if (rows == 0) return INDNET_OK;
  IN_REQUIRE(V && w && b && labels && probs && loss_part && correct, "ntn_head_fwd: null pointer");
  IN_REQUIRE(!out2 == !ticket, "ntn_head_fwd: out2 and ticket go together");
  IN_REQUIRE(!dV == !dlogit && (dV || !dV_bf), "ntn_head_fwd: dlogit and dV go together (dV_bf optional)");
  IN_REQUIRE(!dV || N <= NTN_NT, "ntn_head_fwd: the fused backward needs N <= %d classes", NTN_NT);
  IN_REQUIRE(!dV_bf || ld_bf >= N * S, "ntn_head_fwd: bf16 row stride %d shorter than N*S", ld_bf);
  ntn_head_fwd_kernel<<<rows, NTN_NT, sizeof(float) * 2 * N, st>>>(V, w, b, labels, probs, loss_part, correct, N, S, out2,
                                                                   running, ticket, dlogit, dV, (__nv_bfloat16*)dV_bf, ld_bf,
                                                                   scale);
  IN_LAUNCH_CHECK("ntn_head_fwd_kernel");
  return INDNET_OK;
}

extern "C" int indnet_loss_acc_reduce(indnet_stream st_, const float* loss_part, const int* correct, float* out2,
                                      float* running, int rows, int N) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(rows > 0 && N > 0, "loss_acc_reduce: bad sizes");
  IN_REQUIRE(loss_part && correct && out2, "loss_acc_reduce: null pointer");
  loss_acc_reduce_kernel<<<1, 1024, 0, st>>>(loss_part, correct, out2, running, rows, N);
  IN_LAUNCH_CHECK("loss_acc_reduce_kernel");
  return INDNET_OK;
}

extern "C" int indnet_ntn_head_bwd(indnet_stream st_, const float* V, const float* w, const float* probs,
                                   const int* labels, const float* dprobs, float* dlogit, float* dV, void* dV_bf,
                                   int ld_bf, float* dw, float* db, int rows, int N, int S, float scale) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(rows >= 0 && N > 0 && S > 0, "ntn_head_bwd: bad sizes rows=%d N=%d S=%d", rows, N, S);
  IN_REQUIRE(N <= NTN_NT, "ntn_head_bwd: N=%d classes exceeds %d", N, NTN_NT);
  IN_REQUIRE(V && w && probs && (labels || dprobs) && dlogit && dV && !dw == !db, "ntn_head_bwd: null pointer");
  IN_REQUIRE(!dV_bf || ld_bf >= N * S, "ntn_head_bwd: bf16 row stride %d shorter than N*S", ld_bf);
  if (rows > 0) {
    ntn_head_bwd_kernel<<<rows, NTN_NT, sizeof(float) * N, st>>>(V, w, probs, labels, dprobs, dlogit, dV, (__nv_bfloat16*)dV_bf, ld_bf, N, S, scale);
    IN_LAUNCH_CHECK("ntn_head_bwd_kernel");
  }
  return dw ? indnet_ntn_head_dw(st_, V, dlogit, dw, db, rows, N, S) : INDNET_OK;
}

extern "C" int indnet_ntn_head_dw(indnet_stream st_, const float* V, const float* dlogit, float* dw, float* db, int rows,
                                  int N, int S) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(rows >= 0 && N > 0 && S > 0, "ntn_head_dw: bad sizes rows=%d N=%d S=%d", rows, N, S);
  IN_REQUIRE(V && dlogit && dw && db, "ntn_head_dw: null pointer");
  IN_CUDA(cudaMemsetAsync(dw, 0, sizeof(float) * (size_t)N * S * N, st));
  if (rows == 0) { IN_CUDA(cudaMemsetAsync(db, 0, sizeof(float) * N, st)); return INDNET_OK; }
  int rpb = 64;
  int gy = ceil_div(rows, rpb);
  while (gy > 65535) { rpb *= 2; gy = ceil_div(rows, rpb); }
  ntn_dw_kernel<<<dim3(ceil_div((long long)N * S * N, 128), gy), 128, 0, st>>>(V, dlogit, dw, rows, N, S, rpb);
  IN_LAUNCH_CHECK("ntn_dw_kernel");
  return colsum(st, dlogit, rows, N, N, db, false);
}

extern "C" int indnet_permute021(indnet_stream st_, const float* in, float* out, int d0, int d1, int d2) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(d0 >= 0 && d1 >= 0 && d2 >= 0, "permute021: bad sizes");
  if (d0 == 0 || d1 == 0 || d2 == 0) return INDNET_OK;
  IN_REQUIRE(in && out && in != out, "permute021: null or aliased pointers");
  IN_REQUIRE(d0 <= 65535, "permute021: leading dim too large");
  permute021_kernel<<<dim3(ceil_div(d2, 32), ceil_div(d1, 32), d0), dim3(32, 8), 0, st>>>(in, out, d1, d2);
  IN_LAUNCH_CHECK("permute021_kernel");
  return INDNET_OK;
}
