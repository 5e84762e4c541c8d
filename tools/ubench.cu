// Micro-latency probe for the warp-level primitives the sampler walk is built from (run on the B200 box:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ubench tools/ubench.cu && /tmp/ubench)
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(long long* out, int iters, int seed) {
  const int lane = threadIdx.x & 31;
  const unsigned below = (1u << lane) - 1u;
  unsigned x = seed + lane;
  long long t0, t1;
  // 1. dependent POPC chain
  t0 = clock64();
  for (int i = 0; i < iters; ++i) x = __popc(x | 0x10001u) + x;
  t1 = clock64(); if (lane == 0) out[0] = t1 - t0;
  // 2. dependent ballot chain
  unsigned b = x;
  t0 = clock64();
  for (int i = 0; i < iters; ++i) b = __ballot_sync(0xffffffffu, ((b >> lane) & 1u) ^ (lane & 1));
  t1 = clock64(); if (lane == 0) out[1] = t1 - t0;
  // 3. dependent clz chain
  t0 = clock64();
  for (int i = 0; i < iters; ++i) x = (0xFFFFFFFFu >> __clz(x | 1u)) + x;
  t1 = clock64(); if (lane == 0) out[2] = t1 - t0;
  // 4. the sampler's flat inner step: c = popc(bal & below); bal = ballot(c <= t)
  int t = (x & 31);
  t0 = clock64();
  for (int i = 0; i < iters; ++i) { int c = __popc(b & below); b = __ballot_sync(0xffffffffu, c <= t + (i & 1)); }
  t1 = clock64(); if (lane == 0) out[3] = t1 - t0;
  // 5. same with the data-dependent exit test kept (never taken)
  t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    int c = __popc(b & below);
    unsigned b2 = __ballot_sync(0xffffffffu, c <= t + (i & 1));
    if (b2 == 0x12345u) break;
    b = b2;
  }
  t1 = clock64(); if (lane == 0) out[4] = t1 - t0;
  // 6. dependent IADD chain (reference)
  t0 = clock64();
  for (int i = 0; i < iters; ++i) x = x * 3 + 1;
  t1 = clock64(); if (lane == 0) out[5] = t1 - t0;
  // 7. shuffle chain
  t0 = clock64();
  for (int i = 0; i < iters; ++i) x = __shfl_up_sync(0xffffffffu, x, 1) + 1;
  t1 = clock64(); if (lane == 0) out[6] = t1 - t0;
  // 8. smem store -> __syncthreads -> load chain (block of blockDim threads)
  __shared__ unsigned s[8];
  t0 = clock64();
  for (int i = 0; i < iters; ++i) { if (lane == 0) s[threadIdx.x >> 5] = x; __syncthreads(); x += s[(threadIdx.x >> 5) ^ 1]; __syncthreads(); }
  t1 = clock64(); if (lane == 0) out[7] = t1 - t0;
  if (x == 12345 && b == 77) out[9] = 1;
}
int main() {
  long long* d; cudaMalloc(&d, 80); long long h[10];
  for (int threads : {32, 256}) {
    k<<<1, threads>>>(d, 4096, 7); cudaDeviceSynchronize();
    k<<<1, threads>>>(d, 4096, 7); cudaDeviceSynchronize();
    cudaMemcpy(h, d, 80, cudaMemcpyDeviceToHost);
    printf("threads=%d cycles/iter: popc %.1f ballot %.1f clz-mask %.1f popc+ballot %.1f popc+ballot+exit %.1f imad %.1f shfl %.1f sts+2bar+lds %.1f\n",
           threads, h[0]/4096.0, h[1]/4096.0, h[2]/4096.0, h[3]/4096.0, h[4]/4096.0, h[5]/4096.0, h[6]/4096.0, h[7]/4096.0);
  }
  return 0;
}
