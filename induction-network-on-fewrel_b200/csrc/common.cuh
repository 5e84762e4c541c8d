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

// Counter-based Bernoulli stream for the LSTM output dropout (model/graph.py:36-38); the same integer mix is
// restated in Python (induction_b200.engine.dropout_keep_mask) so a device mask can be replayed in the oracle.
// key = one 64-bit mix of (seed + step * 1000003) per thread; per element a 32-bit murmur-style finaliser
// (two multiplies) - the per-element cost matters inside the persistent recurrence kernels.
__host__ __device__ __forceinline__ uint32_t dropout_key(uint64_t seed) {
  uint64_t z = seed + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z = z ^ (z >> 31);
  return (uint32_t)(z >> 32);
}
__host__ __device__ __forceinline__ uint32_t dropout_hash(uint32_t key, uint64_t idx) {
  uint32_t h = (uint32_t)idx * 0x9E3779B1u + key + (uint32_t)(idx >> 32) * 0x7FEB352Du;
  h ^= h >> 16; h *= 0x85EBCA6Bu;
  h ^= h >> 13; h *= 0xC2B2AE35u;
  h ^= h >> 16;
  return h;
}
// One hash serves TWO sentences (2x, 2x+1) of the same (time, feature): `idx_even` is the flat [X][L][F] index of
// the even sentence's element, the even sentence takes the low 16 bits, the odd one the high 16 bits;
// keep iff that field < keep_prob * 2^16.
__host__ __device__ __forceinline__ uint32_t dropout_threshold(float keep_prob) { return (uint32_t)(keep_prob * 65536.0f); }
__host__ __device__ __forceinline__ bool dropout_keep_field(uint32_t hash, bool odd, uint32_t thr) {
  return (odd ? (hash >> 16) : (hash & 0xFFFFu)) < thr;
}
// element (sentence x, flat index idx) with `row` = L * F elements per sentence
__host__ __device__ __forceinline__ bool dropout_keep(uint64_t seed, uint64_t idx, int x, uint64_t row, float keep_prob) {
  const bool odd = (x & 1) != 0;
  return dropout_keep_field(dropout_hash(dropout_key(seed), odd ? idx - row : idx), odd, dropout_threshold(keep_prob));
}

}  // namespace indnet
