// C-ABI plumbing: error reporting and the exported wrappers of the GEMM / column-sum building blocks.
#include <cstdarg>
#include <cstring>
#include "common.cuh"
#include "indnet_internal.h"
#include "../../include/indnet.h"

namespace indnet {
static thread_local char g_err[1024] = "";
int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
static unsigned long long g_launches = 0;
void count_launches(int n) { __atomic_add_fetch(&g_launches, (unsigned long long)n, __ATOMIC_RELAXED); }
}  // namespace indnet

using namespace indnet;

extern "C" long long indnet_launch_count(void) { return (long long)__atomic_load_n(&indnet::g_launches, __ATOMIC_RELAXED); }
extern "C" const char* indnet_last_error(void) { return g_err; }
extern "C" int indnet_abi_version(void) { return INDNET_ABI_VERSION; }

extern "C" int indnet_sgemm(indnet_stream st, int transA, int transB, int M, int N, int K, float alpha, const float* A,
                            int lda, const float* B, int ldb, float beta, float* C, int ldc, const float* bias, int act,
                            int batch, int inner, long long sAo, long long sAi, long long sBo, long long sBi,
                            long long sCo, long long sCi, int splitk) {
  return sgemm((cudaStream_t)st, transA != 0, transB != 0, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, bias, act,
               batch, inner, sAo, sAi, sBo, sBi, sCo, sCi, splitk);
}

extern "C" int indnet_colsum(indnet_stream st, const float* A, int rows, int cols, int lda, float* out, int accumulate) {
  IN_REQUIRE(rows >= 0 && cols >= 0, "colsum: bad sizes");
  IN_REQUIRE(cols == 0 || (A && out) || rows == 0, "colsum: null pointer");
  return colsum((cudaStream_t)st, A, rows, cols, lda, out, accumulate != 0);
}

extern "C" int indnet_memset(indnet_stream st, void* dst, int byte_value, long long bytes) {
  IN_REQUIRE(bytes >= 0, "memset: negative size");
  if (bytes == 0) return INDNET_OK;
  IN_REQUIRE(dst != nullptr, "memset: null pointer");
  IN_CUDA(cudaMemsetAsync(dst, byte_value, (size_t)bytes, (cudaStream_t)st));
  return INDNET_OK;
}
