// Internal (C++) declarations shared between the .cu files of libindnet.  The public C-ABI is in
// include/indnet.h.
#pragma once
#include <cuda_runtime.h>

namespace indnet {

int sgemm(cudaStream_t st, bool tA, bool tB, int M, int N, int K, float alpha, const float* A, int lda,
          const float* B, int ldb, float beta, float* C, int ldc, const float* bias, int act, int batch,
          int inner, long long sAo, long long sAi, long long sBo, long long sBi, long long sCo, long long sCi,
          int splitk);

int colsum(cudaStream_t st, const float* A, int rows, int cols, int lda, float* out, bool accumulate);

}  // namespace indnet
