// FP32-mode GEMM building block: C = alpha * op(A) * op(B) + beta * C (+ bias, + activation),
// row-major with leading dimensions, two-level strided batching and optional split-K.
// This is the "FP32 parity mode" contraction (plain FFMA, no tensor cores, no fast-math) used for
// every dense op of the path when 1e-5 agreement with the reference graph is requested
// (model/graph.py:44 input projection, model/layer_module.py:65 Dense, :132 routing transform,
// :24-30 NTN slices) and for all weight-gradient contractions.  The bf16 tensor-core mode replaces
// the large ones with tcgen05 kernels (tc_*.cu).
#include "common.cuh"
#include "indnet_internal.h"

namespace indnet {

constexpr int BM = 128, BN = 128, BK = 16, TM = 8, TN = 8, NT = 256;

struct GemmP {
  const float* A; const float* B; float* C; const float* bias;
  int M, N, K, lda, ldb, ldc;
  float alpha, beta;
  int act;            // 0 none, 1 tanh, 2 relu
  int splitk, kchunk; // K range per z-slice
  int inner;          // two-level batch: z -> (zo = z / inner, zi = z % inner)
  long long sAo, sAi, sBo, sBi, sCo, sCi;
};

template <bool TA, bool TB>
__global__ void __launch_bounds__(NT) sgemm_kernel(GemmP p) {
  __shared__ __align__(16) float As[BK][BM + 4];
  __shared__ __align__(16) float Bs[BK][BN + 4];

  const int z = blockIdx.z;
  const int batch = z / p.splitk, ks = z % p.splitk;
  const int zo = batch / p.inner, zi = batch % p.inner;
  const float* __restrict__ A = p.A + zo * p.sAo + zi * p.sAi;
  const float* __restrict__ B = p.B + zo * p.sBo + zi * p.sBi;
  float* __restrict__ C = p.C + zo * p.sCo + zi * p.sCi;

  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int k_begin = ks * p.kchunk;
  const int k_end = min(p.K, k_begin + p.kchunk);
  const int tid = threadIdx.x;
  const int tx = tid % 16, ty = tid / 16;

  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  for (int k0 = k_begin; k0 < k_end; k0 += BK) {
    // ---- stage A tile (BM x BK) as As[k][m]
#pragma unroll
    for (int i = 0; i < (BM * BK) / NT; ++i) {
      int idx = tid + i * NT;
      int m, k;
      if (TA) { k = idx / BM; m = idx % BM; } else { m = idx / BK; k = idx % BK; }
      int gm = m0 + m, gk = k0 + k;
      float v = 0.f;
      if (gm < p.M && gk < k_end) v = TA ? A[(long long)gk * p.lda + gm] : A[(long long)gm * p.lda + gk];
      As[k][m] = v;
    }
    // ---- stage B tile (BK x BN) as Bs[k][n]
#pragma unroll
    for (int i = 0; i < (BN * BK) / NT; ++i) {
      int idx = tid + i * NT;
      int n, k;
      if (TB) { n = idx / BK; k = idx % BK; } else { k = idx / BN; n = idx % BN; }
      int gn = n0 + n, gk = k0 + k;
      float v = 0.f;
      if (gn < p.N && gk < k_end) v = TB ? B[(long long)gn * p.ldb + gk] : B[(long long)gk * p.ldb + gn];
      Bs[k][n] = v;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float a[TM], b[TN];
      float4 a0 = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      float4 a1 = *reinterpret_cast<const float4*>(&As[k][ty * 4 + 64]);
      float4 b0 = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      float4 b1 = *reinterpret_cast<const float4*>(&Bs[k][tx * 4 + 64]);
      a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
      b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

  // ---- epilogue
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    int gm = m0 + ty * 4 + (i & 3) + (i >> 2) * 64;
    if (gm >= p.M) continue;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      int gn = n0 + tx * 4 + (j & 3) + (j >> 2) * 64;
      if (gn >= p.N) continue;
      float* c = C + (long long)gm * p.ldc + gn;
      float v = p.alpha * acc[i][j];
      if (p.splitk > 1) {
        // C was pre-scaled by beta (or zeroed) by the launcher; bias joins on slice 0
        if (p.bias && ks == 0) v += p.bias[gn];
        atomicAdd(c, v);
      } else {
        if (p.beta != 0.f) v += p.beta * (*c);
        if (p.bias) v += p.bias[gn];
        if (p.act == 1) v = tanhf(v);
        else if (p.act == 2) v = fmaxf(v, 0.f);
        *c = v;
      }
    }
  }
}

__global__ void scale_kernel(float* C, int M, int N, int ldc, float beta, long long n_total, int inner,
                             long long sCo, long long sCi, int batch) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n_total) return;
  long long per = (long long)M * N;
  int b = (int)(i / per);
  long long r = i % per;
  int m = (int)(r / N), n = (int)(r % N);
  float* c = C + (b / inner) * sCo + (b % inner) * sCi + (long long)m * ldc + n;
  *c = (beta == 0.f) ? 0.f : beta * (*c);
}

int sgemm(cudaStream_t st, bool tA, bool tB, int M, int N, int K, float alpha, const float* A, int lda,
          const float* B, int ldb, float beta, float* C, int ldc, const float* bias, int act, int batch,
          int inner, long long sAo, long long sAi, long long sBo, long long sBi, long long sCo, long long sCi,
          int splitk) {
  IN_REQUIRE(M >= 0 && N >= 0 && K >= 0 && batch >= 0, "sgemm: negative dimension");
  if (M == 0 || N == 0 || batch == 0) return INDNET_OK;
  IN_REQUIRE(A && B && C, "sgemm: null operand");
  IN_REQUIRE(inner >= 1, "sgemm: inner batch count must be >= 1");
  const int gx = ceil_div(N, BN), gy = ceil_div(M, BM);
  if (splitk <= 0) {  // auto: aim at >= 2 waves of 148 SMs when K is long
    splitk = 1;
    long long ctas = (long long)gx * gy * batch;
    if (ctas < 148 && K >= 1024 && act == 0) {
      splitk = (int)((296 + ctas - 1) / ctas);
      int maxs = K / 256;
      if (splitk > maxs) splitk = maxs;
      if (splitk < 1) splitk = 1;
    }
  }
  IN_REQUIRE(!(splitk > 1 && act != 0), "sgemm: split-K cannot apply an activation");
  GemmP p;
  p.A = A; p.B = B; p.C = C; p.bias = bias;
  p.M = M; p.N = N; p.K = K; p.lda = lda; p.ldb = ldb; p.ldc = ldc;
  p.alpha = alpha; p.beta = beta; p.act = act;
  p.splitk = splitk;
  p.kchunk = ((ceil_div(K, splitk) + BK - 1) / BK) * BK;
  if (p.kchunk == 0) p.kchunk = BK;
  p.inner = inner; p.sAo = sAo; p.sAi = sAi; p.sBo = sBo; p.sBi = sBi; p.sCo = sCo; p.sCi = sCi;
  if (splitk > 1 && beta != 1.f) {
    long long n_total = (long long)M * N * batch;
    scale_kernel<<<ceil_div(n_total, 256), 256, 0, st>>>(C, M, N, ldc, beta, n_total, inner, sCo, sCi, batch);
    IN_LAUNCH_CHECK("scale_kernel");
  }
  long long gz = (long long)batch * splitk;
  IN_REQUIRE(gz <= 65535, "sgemm: batch*splitk %lld exceeds grid.z limit", gz);
  dim3 grid(gx, gy, (unsigned)gz);
  if (tA && tB) sgemm_kernel<true, true><<<grid, NT, 0, st>>>(p);
  else if (tA) sgemm_kernel<true, false><<<grid, NT, 0, st>>>(p);
  else if (tB) sgemm_kernel<false, true><<<grid, NT, 0, st>>>(p);
  else sgemm_kernel<false, false><<<grid, NT, 0, st>>>(p);
  IN_LAUNCH_CHECK("sgemm_kernel");
  return INDNET_OK;
}

// out[c] (+)= sum_r A[r, c]; deterministic per block, atomics across row chunks.
__global__ void colsum_kernel(const float* __restrict__ A, int rows, int cols, int lda, float* out, int rows_per_block) {
  __shared__ float red[8][33];
  int tx = threadIdx.x % 32, ty = threadIdx.x / 32;
  int c = blockIdx.x * 32 + tx;
  int r0 = blockIdx.y * rows_per_block;
  int r1 = min(rows, r0 + rows_per_block);
  float s = 0.f;
  if (c < cols)
    for (int r = r0 + ty; r < r1; r += 8) s += A[(long long)r * lda + c];
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && c < cols) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][tx];
    atomicAdd(out + c, t);
  }
}

int colsum(cudaStream_t st, const float* A, int rows, int cols, int lda, float* out, bool accumulate) {
  if (cols == 0) return INDNET_OK;
  if (!accumulate) IN_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * cols, st));
  if (rows == 0) return INDNET_OK;
  int gx = ceil_div(cols, 32);
  int rpb = 256;
  int gy = ceil_div(rows, rpb);
  while (gy > 65535) { rpb *= 2; gy = ceil_div(rows, rpb); }
  colsum_kernel<<<dim3(gx, gy), 256, 0, st>>>(A, rows, cols, lda, out, rpb);
  IN_LAUNCH_CHECK("colsum_kernel");
  return INDNET_OK;
}

}  // namespace indnet
