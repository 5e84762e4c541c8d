// a4, bf16 tensor-core mode: structured self-attention pooling (model/layer_module.py:49-98) as ONE kernel.
//   proj  = tanh(h * Wa1 + ba1)        tcgen05 GEMM, 128 token rows x N <= 256 columns per pass, fp32 in TMEM
//   score = proj . w_a2                 reduced in registers while the accumulator is drained (lane = token row)
//   alpha = softmax_t(score + (1 - m) * (-2^31))    per sentence, through 1 KB of shared memory
//   e     = sum_t alpha_t h_t           h re-read from L2 (the CTA's own 128 rows, just loaded by TMA)
// A CTA owns floor(128 / L) WHOLE sentences (3 at L = 40, 1 at L = 128), so the softmax and the weighted sum never
// leave it; proj [T, H] - 39 MB per step at BASELINE configs[1] - is written (bf16) only when the backward pass
// asks for it and never read back in the forward.  Warp-specialised like tc_gemm.cu: warp 0 TMA producer, warp 1
// TMEM allocator + MMA issuer, warps 2-5 epilogue (one TMEM lane quarter each).  H = 512 runs as two passes over
// the same accumulator columns.
#include "common.cuh"
#include "indnet_internal.h"
#include "tc_common.cuh"
#include "../../include/indnet.h"

namespace indnet {

using namespace tc;

constexpr int AF_BM = 128, AF_BK = 64, AF_STAGES = 2, AF_THREADS = 192;
constexpr int AF_A_BYTES = AF_BM * AF_BK * 2;                 // 16 KB

int make_tmap_bf16(CUtensorMap* tm, const void* base, long long inner, long long outer, long long ld,
                   long long batch, long long batch_stride, int box_inner, int box_outer);

struct AttnFusedP {
  const __nv_bfloat16* h;       // [X*L][H]
  const float* bias; const float* w2; const int* mask;
  __nv_bfloat16* proj;          // optional [X*L][H]
  float* enc_a; float* enc_b; __nv_bfloat16* enc_a_bf; __nv_bfloat16* enc_b_bf; float* alphas;
  int X, L, H, NP, group, split, spt;       // NP = accumulator columns per pass, spt = sentences per CTA
};

__device__ __forceinline__ float af_tanh(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(AF_THREADS, 2)
attn_fused_fwd_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, AttnFusedP p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int B_BYTES = p.NP * AF_BK * 2;                 // NP / 64 boxes of 8 KB
  uint8_t* sA = smem;
  uint8_t* sB = smem + AF_STAGES * AF_A_BYTES;
  float* sBias = reinterpret_cast<float*>(sB + AF_STAGES * B_BYTES);      // [H]
  float* sW2 = sBias + p.H;                                               // [H]
  float* sSc = sW2 + p.H;                                                 // [128] scores, then exp(score - max)
  float* sAl = sSc + AF_BM;                                               // [128] alphas
  uint64_t* full = reinterpret_cast<uint64_t*>(sAl + AF_BM);
  uint64_t* empty = full + AF_STAGES;
  uint64_t* tmem_full = empty + AF_STAGES;
  uint64_t* tmem_empty = tmem_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int L = p.L, H = p.H, NP = p.NP;
  const int rows_tile = p.spt * L;                      // token rows this CTA owns (<= 128)
  const long long row0 = (long long)blockIdx.x * rows_tile;
  const int KB = H / AF_BK, NPASS = H / NP;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int i = 0; i < AF_STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(tmem_full, 1);
    mbar_init(tmem_empty, 128);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 256);
  for (int i = threadIdx.x; i < H; i += AF_THREADS) { sBias[i] = p.bias[i]; sW2[i] = p.w2[i]; }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t taddr = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int it = 0;
      for (int pass = 0; pass < NPASS; ++pass) {
        for (int kb = 0; kb < KB; ++kb, ++it) {
          const int s = it % AF_STAGES, ph = (it / AF_STAGES) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          mbar_expect_tx(&full[s], AF_A_BYTES + B_BYTES);
          tma_load_3d(sA + s * AF_A_BYTES, &tmA, &full[s], kb * AF_BK, (int)row0, 0);       // 64 k x 128 token rows
          for (int j = 0; j < NP / 64; ++j)                                                 // Wa1 [K][N]: 64 n x 64 k boxes
            tma_load_3d(sB + s * B_BYTES + j * 8192, &tmB, &full[s], pass * NP + j * 64, kb * AF_BK, 0);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = idesc_bf16(AF_BM, NP, false, true);
      int it = 0;
      for (int pass = 0; pass < NPASS; ++pass) {
        if (pass > 0) { mbar_wait(tmem_empty, (pass - 1) & 1); tc_fence_after(); }     // accumulator drained
        for (int kb = 0; kb < KB; ++kb, ++it) {
          const int s = it % AF_STAGES, ph = (it / AF_STAGES) & 1;
          mbar_wait(&full[s], ph);
          tc_fence_after();
          const uint32_t a_base = smem_u32(sA + s * AF_A_BYTES), b_base = smem_u32(sB + s * B_BYTES);
#pragma unroll
          for (int k = 0; k < AF_BK / 16; ++k) {
            const uint64_t ad = smem_desc_sw128(a_base + k * 32, 0, 1024);                    // K-major A
            const uint64_t bd = smem_desc_sw128(b_base + k * (16 * 128), AF_BK * 128, 1024);  // MN-major B, 64-wide N atoms 8 KB apart
            umma_bf16(taddr, ad, bd, idesc, (kb > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty[s]);
        }
        umma_commit(tmem_full);
      }
    }
  } else {
    // ---- epilogue warps: lane = token row of the tile
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const long long gr = row0 + row;
    const bool valid = row < rows_tile && gr < (long long)p.X * L;
    float score = 0.f;
    for (int pass = 0; pass < NPASS; ++pass) {
      mbar_wait(tmem_full, pass & 1);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < NP / 32; ++c) {
        float v[32];
        tmem_ld32(taddr + ((uint32_t)(q * 32) << 16) + c * 32, v);
        tmem_ld_wait();
        const int n0 = pass * NP + c * 32;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          v[j] = af_tanh(v[j] + sBias[n0 + j]);
          score = fmaf(v[j], sW2[n0 + j], score);
        }
        if (p.proj && valid) {
          uint4* dst = reinterpret_cast<uint4*>(p.proj + gr * H + n0);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            __nv_bfloat162 a = __floats2bfloat162_rn(v[8 * j], v[8 * j + 1]), b = __floats2bfloat162_rn(v[8 * j + 2], v[8 * j + 3]);
            __nv_bfloat162 cc = __floats2bfloat162_rn(v[8 * j + 4], v[8 * j + 5]), d = __floats2bfloat162_rn(v[8 * j + 6], v[8 * j + 7]);
            dst[j] = make_uint4(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b),
                                *reinterpret_cast<uint32_t*>(&cc), *reinterpret_cast<uint32_t*>(&d));
          }
        }
      }
      tc_fence_before();
      mbar_arrive(tmem_empty);
    }
    // ---- masked softmax over each sentence's L rows (layer_module.py:84-87)
    float sc = -INFINITY;
    if (valid) {
      const float m = p.mask[gr] > 0 ? 1.0f : 0.0f;
      sc = score + (1.0f - m) * (-2147483648.0f);
    }
    sSc[row] = sc;
    asm volatile("bar.sync 1, 128;" ::: "memory");
    const int sbase = (row / L) * L;          // first row of this row's sentence
    float e = 0.f;
    if (valid) {
      float mx = -INFINITY;
      for (int t = 0; t < L; ++t) mx = fmaxf(mx, sSc[sbase + t]);
      e = expf(sc - mx);
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    sSc[row] = e;
    asm volatile("bar.sync 1, 128;" ::: "memory");
    float al = 0.f;
    if (valid) {
      float sum = 0.f;
      for (int t = 0; t < L; ++t) sum += sSc[sbase + t];
      al = e / sum;
      p.alphas[gr] = al;
    }
    sAl[row] = al;
    asm volatile("bar.sync 1, 128;" ::: "memory");
    // ---- e = sum_t alpha_t h_t.  16-byte loads of h (re-read from L2: this CTA's rows were just loaded), every load
    // independent: thread = (time group tg of 4, 8-feature chunk); the four partial sums meet in shared memory
    // (the operand ring is free by now).
    const int tid = threadIdx.x - 64;          // 0..127
    const int x_first = (int)(row0 / L);
    const int chunks = H / 8;                  // 16-byte chunks per row
    float* red = reinterpret_cast<float*>(smem);            // [4][H] partial sums
    for (int s = 0; s < p.spt; ++s) {
      const int x = x_first + s;
      if (x >= p.X) break;
      const int g = x / p.group, r = x - g * p.group;
      const long long orow = (r < p.split) ? ((long long)g * p.split + r) : ((long long)g * (p.group - p.split) + (r - p.split));
      float* enc = (r < p.split) ? p.enc_a + orow * H : p.enc_b + orow * H;
      __nv_bfloat16* enc_bf = p.enc_a_bf ? ((r < p.split) ? p.enc_a_bf + orow * H : p.enc_b_bf + orow * H) : nullptr;
      const uint4* hp = reinterpret_cast<const uint4*>(p.h + ((long long)x * L) * H);
      for (int c0 = 0; c0 < chunks; c0 += 32) {
        const int c = c0 + (tid & 31), tg = tid >> 5;
        float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (c < chunks) {
#pragma unroll 4
          for (int t = tg; t < L; t += 4) {
            const uint4 w = hp[(long long)t * chunks + c];
            const float a = sAl[s * L + t];
            acc[0] = fmaf(a, __uint_as_float(w.x << 16), acc[0]); acc[1] = fmaf(a, __uint_as_float(w.x & 0xFFFF0000u), acc[1]);
            acc[2] = fmaf(a, __uint_as_float(w.y << 16), acc[2]); acc[3] = fmaf(a, __uint_as_float(w.y & 0xFFFF0000u), acc[3]);
            acc[4] = fmaf(a, __uint_as_float(w.z << 16), acc[4]); acc[5] = fmaf(a, __uint_as_float(w.z & 0xFFFF0000u), acc[5]);
            acc[6] = fmaf(a, __uint_as_float(w.w << 16), acc[6]); acc[7] = fmaf(a, __uint_as_float(w.w & 0xFFFF0000u), acc[7]);
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) red[tg * H + c * 8 + j] = acc[j];
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      for (int n = tid * 2; n < H; n += 256) {
        const float a0 = red[n] + red[H + n] + red[2 * H + n] + red[3 * H + n];
        const float a1 = red[n + 1] + red[H + n + 1] + red[2 * H + n + 1] + red[3 * H + n + 1];
        *reinterpret_cast<float2*>(enc + n) = make_float2(a0, a1);
        if (enc_bf) *reinterpret_cast<__nv_bfloat162*>(enc_bf + n) = __floats2bfloat162_rn(a0, a1);
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
    }
  }
  __syncthreads();
  if (warp == 1) tmem_dealloc(taddr, 256);
}

}  // namespace indnet

using namespace indnet;

extern "C" int indnet_attn_fused_fwd(indnet_stream st_, const void* hout_bf, const void* wa1_bf, const float* ba1,
                                     const float* w_a2, const int* mask, void* proj_bf, float* enc_a, float* enc_b,
                                     void* enc_a_bf, void* enc_b_bf, float* alphas, int X, int L, int H, int group,
                                     int split) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(X >= 0 && L > 0 && H > 0, "attn_fused_fwd: bad sizes X=%d L=%d H=%d", X, L, H);
  IN_REQUIRE(L <= AF_BM, "attn_fused_fwd: sequence_length %d exceeds the %d-row tile (use the unfused path)", L, AF_BM);
  IN_REQUIRE(H % 64 == 0 && (H <= 256 || H % 256 == 0) && H <= 1024, "attn_fused_fwd: H=%d must be a multiple of 64 up to 256, or 512 / 768 / 1024", H);
  if (X == 0) return INDNET_OK;
  IN_REQUIRE(hout_bf && wa1_bf && ba1 && w_a2 && mask && enc_a && alphas, "attn_fused_fwd: null pointer");
  IN_REQUIRE(group > 0 && split >= 0 && split <= group && X % group == 0, "attn_fused_fwd: bad group/split %d/%d", group, split);
  IN_REQUIRE(split == group || enc_b, "attn_fused_fwd: second output missing");
  IN_REQUIRE(!enc_a_bf || split == group || enc_b_bf, "attn_fused_fwd: second bf16 output missing");
  IN_REQUIRE((((uintptr_t)proj_bf | (uintptr_t)enc_a | (uintptr_t)enc_b) & 15) == 0, "attn_fused_fwd: outputs must be 16-byte aligned");
  IN_REQUIRE((long long)X * L < 2147483647LL, "attn_fused_fwd: too many tokens");
  AttnFusedP p;
  p.h = (const __nv_bfloat16*)hout_bf; p.bias = ba1; p.w2 = w_a2; p.mask = mask; p.proj = (__nv_bfloat16*)proj_bf;
  p.enc_a = enc_a; p.enc_b = enc_b; p.enc_a_bf = (__nv_bfloat16*)enc_a_bf; p.enc_b_bf = (__nv_bfloat16*)enc_b_bf;
  p.alphas = alphas; p.X = X; p.L = L; p.H = H; p.NP = H <= 256 ? H : 256; p.group = group; p.split = split;
  p.spt = AF_BM / L;
  CUtensorMap tmA, tmB;
  int rc = make_tmap_bf16(&tmA, hout_bf, H, (long long)X * L, H, 1, 0, AF_BK, AF_BM);       // [T][H]: box 64 k x 128 rows
  if (rc) return rc;
  rc = make_tmap_bf16(&tmB, wa1_bf, H, H, H, 1, 0, 64, AF_BK);                              // [K][N]: box 64 n x 64 k
  if (rc) return rc;
  const int smem = AF_STAGES * (AF_A_BYTES + p.NP * AF_BK * 2) + (2 * H + 2 * AF_BM) * 4 + 128 + 1024;
  IN_CUDA(cudaFuncSetAttribute(attn_fused_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  attn_fused_fwd_kernel<<<ceil_div(X, p.spt), AF_THREADS, smem, st>>>(tmA, tmB, p);
  IN_LAUNCH_CHECK("attn_fused_fwd_kernel");
  return INDNET_OK;
}
