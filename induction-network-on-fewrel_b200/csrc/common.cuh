// Shared helpers for the libindnet kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

#define INDNET_OK 0
#define INDNET_ERR_ARG 1
#define INDNET_ERR_CUDA 2
#define INDNET_ERR_UNSUPPORTED 3

namespace indnet {

int set_error(int code, const char* fmt, ...);
void count_launches(int n);   // feeds indnet_launch_count(): kernels this library has launched

#define IN_REQUIRE(cond, ...)                                              \
  do {                                                                     \
    if (!(cond)) return ::indnet::set_error(INDNET_ERR_ARG, __VA_ARGS__);  \
  } while (0)

#define IN_CUDA(expr)                                                                        \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess)                                                                   \
      return ::indnet::set_error(INDNET_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(_e)); \
  } while (0)

#define IN_LAUNCH_CHECK(name)                                                                 \
  do {                                                                                        \
    ::indnet::count_launches(1);                                                              \
    cudaError_t _e = cudaGetLastError();                                                      \
    if (_e != cudaSuccess)                                                                    \
      return ::indnet::set_error(INDNET_ERR_CUDA, "launch of %s failed: %s", name, cudaGetErrorString(_e)); \
  } while (0)

static inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ float sigmoid_acc(float x) { return 1.0f / (1.0f + expf(-x)); }

// Counter-based uniform in [0,1) used for LSTM output dropout; the same integer mix is restated in
// Python (induction_b200.engine.dropout_keep_mask) so a device mask can be replayed in the oracle.
__host__ __device__ __forceinline__ uint32_t mix32(uint64_t seed, uint64_t idx) {
  uint64_t z = seed + 0x9E3779B97F4A7C15ull * (idx + 1);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z = z ^ (z >> 31);
  return (uint32_t)(z >> 32);
}
__host__ __device__ __forceinline__ bool dropout_keep(uint64_t seed, uint64_t idx, float keep_prob) {
  // keep iff u < keep_prob with u = top 24 bits / 2^24
  return (float)(mix32(seed, idx) >> 8) * (1.0f / 16777216.0f) < keep_prob;
}

}  // namespace indnet
