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
#include <cstdlib>
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
  extern __shared__ __align__(1024) uint8_t smem[];       // declared alignment keeps the shared address space visible (LDS / STS)
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


#ifdef INDNET_ATTN_TIMELINE
__device__ long long g_attnf_dbg[16 * 8];
#define AF_MARK(k) do { if (blockIdx.x == 0 && threadIdx.x == 64 && it < 16) g_attnf_dbg[it * 8 + (k)] = clock64(); } while (0)
#else
#define AF_MARK(k) do { } while (0)
#endif

// ---- H <= 256, persistent: one CTA per SM walks the tiles.  Wa1 (H x H bf16, <= 128 KB) is loaded ONCE per CTA and
// stays in shared memory; the accumulator is double buffered in tensor memory (2 x 256 columns), so tile i+1's
// A load and MMAs run under tile i's epilogue; eight epilogue warps (two per TMEM lane quarter, half of the columns
// each) drain it.  The A buffer is released as soon as the tile's MMAs have completed, so the weighted sum reads h
// back from L2 (where the TMA load just put it).
constexpr int AP_THREADS = 64 + 256;
template <int SPT, int NK>        // SPT sentences per tile of NK = ceil(L / 8) rows per time group, known at compile time; 0: generic
__global__ void __launch_bounds__(AP_THREADS, 1)
attn_fused_fwd_pers_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, AttnFusedP p,
                           int ntiles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int L = p.L, H = p.H;
  const int KB = H / AF_BK;
  const int W_KB_BYTES = H * AF_BK * 2;                 // one K block of Wa1: H / 64 boxes of 8 KB
  uint8_t* sW = smem;                                   // [KB][W_KB_BYTES]
  uint8_t* sA = smem + KB * W_KB_BYTES;                 // [KB][16 KB]
  float* sBias = reinterpret_cast<float*>(sA + KB * AF_A_BYTES);
  float* sW2 = sBias + H;
  float* sScP = sW2 + H;                                // [2][128] score halves
  float* sSc = sScP + 2 * AF_BM;                        // [128]
  float* sAl = sSc + AF_BM;                             // [128]
  uint8_t* sStage = reinterpret_cast<uint8_t*>(sAl + AF_BM);     // [8 warps][32 rows][64 B]: proj chunk, row-major
  uint64_t* w_full = reinterpret_cast<uint64_t*>(sStage + 8 * 2048);
  uint64_t* a_full = w_full + 1;
  uint64_t* a_empty = a_full + 1;
  uint64_t* acc_full = a_empty + 1;                     // [2]
  uint64_t* acc_empty = acc_full + 2;                   // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rows_tile = p.spt * L;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    mbar_init(w_full, 1); mbar_init(a_full, 1); mbar_init(a_empty, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 256); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  for (int i = threadIdx.x; i < H; i += AP_THREADS) { sBias[i] = p.bias[i]; sW2[i] = p.w2[i]; }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t taddr = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(w_full, KB * W_KB_BYTES);
      for (int kb = 0; kb < KB; ++kb)
        for (int j = 0; j < H / 64; ++j)                                         // Wa1 [K][N]: 64 n x 64 k boxes
          tma_load_3d(sW + kb * W_KB_BYTES + j * 8192, &tmB, w_full, j * 64, kb * AF_BK, 0);
      int it = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        if (it > 0) mbar_wait(a_empty, (it - 1) & 1);                            // the previous tile's MMAs have read sA
        mbar_expect_tx(a_full, KB * AF_A_BYTES);
        for (int kb = 0; kb < KB; ++kb)
          tma_load_3d(sA + kb * AF_A_BYTES, &tmA, a_full, kb * AF_BK, tile * rows_tile, 0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = idesc_bf16(AF_BM, H, false, true);
      mbar_wait(w_full, 0);
      int it = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int b = it & 1;
        if (it >= 2) mbar_wait(&acc_empty[b], ((it >> 1) - 1) & 1);              // that accumulator has been drained
        mbar_wait(a_full, it & 1);
        tc_fence_after();
        for (int kb = 0; kb < KB; ++kb) {
          const uint32_t a_base = smem_u32(sA + kb * AF_A_BYTES), b_base = smem_u32(sW + kb * W_KB_BYTES);
#pragma unroll
          for (int k = 0; k < AF_BK / 16; ++k) {
            const uint64_t ad = smem_desc_sw128(a_base + k * 32, 0, 1024);                    // K-major A
            const uint64_t bd = smem_desc_sw128(b_base + k * (16 * 128), AF_BK * 128, 1024);  // MN-major B, 64-wide N atoms 8 KB apart
            umma_bf16(taddr + b * 256, ad, bd, idesc, (kb > 0 || k > 0) ? 1u : 0u);
          }
        }
        umma_commit(a_empty);
        umma_commit(&acc_full[b]);
      }
    }
  } else {
    // ---- epilogue warps: lane = token row of the tile, two warps per row (a half of the columns each)
    const int e = warp - 2;                    // 0..7
    const int q = warp & 3;                    // the TMEM lane quarter this warp may touch
    const int half = e >> 2;
    const int row = q * 32 + lane;
    const int NH = H / 2;                      // columns per half (a multiple of 32)
    const int et = threadIdx.x - 64;           // 0..255
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const int b = it & 1;
      const long long row0 = (long long)tile * rows_tile;
      const long long gr = row0 + row;
      const bool valid = row < rows_tile && gr < (long long)p.X * L;
      AF_MARK(0);
      // the mask of this row is needed right after the drain: fetch it now (a DRAM round trip otherwise in the chain)
      int mrow = 1;
      if (half == 0 && valid) mrow = __ldg(p.mask + gr);
      mbar_wait(&acc_full[b], (it >> 1) & 1);
      tc_fence_after();
      AF_MARK(1);
      float score = 0.f;
      // two register buffers: the TMEM load of chunk c + 1 is in flight while chunk c is processed
      const uint32_t tbase = taddr + ((uint32_t)(q * 32) << 16) + b * 256 + half * NH;
      const int nch = NH / 32;
      float va[32], vb[32];
      tmem_ld32(tbase, va);
      auto chunk = [&](float* v, int c) {
        const int n0 = half * NH + c * 32;
        const float4* b4 = reinterpret_cast<const float4*>(sBias + n0);
        const float4* w4 = reinterpret_cast<const float4*>(sW2 + n0);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 bb = b4[j], ww = w4[j];
          v[4 * j] = af_tanh(v[4 * j] + bb.x); v[4 * j + 1] = af_tanh(v[4 * j + 1] + bb.y);
          v[4 * j + 2] = af_tanh(v[4 * j + 2] + bb.z); v[4 * j + 3] = af_tanh(v[4 * j + 3] + bb.w);
          score = fmaf(v[4 * j], ww.x, score); score = fmaf(v[4 * j + 1], ww.y, score);
          score = fmaf(v[4 * j + 2], ww.z, score); score = fmaf(v[4 * j + 3], ww.w, score);
        }
        if (p.proj) {
          // lane = row would make every 16-byte store instruction touch 32 different 128-byte lines (the LSU works per
          // line): the warp's 32 x 32 chunk goes through 2 KB of shared memory and leaves as 64-byte row segments,
          // four lanes per row, eight lines per instruction
          uint8_t* stg = sStage + e * 2048;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            __nv_bfloat162 a = __floats2bfloat162_rn(v[8 * j], v[8 * j + 1]), bb = __floats2bfloat162_rn(v[8 * j + 2], v[8 * j + 3]);
            __nv_bfloat162 cc = __floats2bfloat162_rn(v[8 * j + 4], v[8 * j + 5]), d = __floats2bfloat162_rn(v[8 * j + 6], v[8 * j + 7]);
            *reinterpret_cast<uint4*>(stg + lane * 64 + ((j ^ ((lane >> 1) & 3)) << 4)) =
                make_uint4(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&bb),
                           *reinterpret_cast<uint32_t*>(&cc), *reinterpret_cast<uint32_t*>(&d));
          }
          __syncwarp();
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int rr = i * 8 + (lane >> 2), cj = lane & 3;          // row of the warp's 32, 16-byte piece of its 64 bytes
            const uint4 w = *reinterpret_cast<const uint4*>(stg + rr * 64 + ((cj ^ ((rr >> 1) & 3)) << 4));
            const int trow = q * 32 + rr;
            const long long grr = row0 + trow;
            if (trow < rows_tile && grr < (long long)p.X * L)
              *reinterpret_cast<uint4*>(p.proj + grr * H + n0 + cj * 8) = w;
          }
          __syncwarp();
        }
      };
#pragma unroll 1
      for (int c = 0; c < nch; c += 2) {
        tmem_ld_wait();
        if (c + 1 < nch) tmem_ld32(tbase + (c + 1) * 32, vb);
        chunk(va, c);
        if (c + 1 < nch) {
          tmem_ld_wait();
          if (c + 2 < nch) tmem_ld32(tbase + (c + 2) * 32, va);
          chunk(vb, c + 1);
        }
      }
      tc_fence_before();
      mbar_arrive(&acc_empty[b]);              // the issuer may overwrite this accumulator (tile it + 2)
      sScP[half * AF_BM + row] = score;
      AF_MARK(2);
      asm volatile("bar.sync 1, 256;" ::: "memory");
      AF_MARK(3);
      // ---- masked softmax over each sentence's L rows (layer_module.py:84-87); one thread per row
      float sc = -INFINITY, ex = 0.f;
      const int sbase = (row / L) * L;
      if (half == 0) {
        if (valid) {
          const float m = mrow > 0 ? 1.0f : 0.0f;
          sc = sScP[row] + sScP[AF_BM + row] + (1.0f - m) * (-2147483648.0f);
        }
        sSc[row] = sc;
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (half == 0 && valid) {
        float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
        int t = 0;
        for (; t + 4 <= L; t += 4) {
          m0 = fmaxf(m0, sSc[sbase + t]); m1 = fmaxf(m1, sSc[sbase + t + 1]);
          m2 = fmaxf(m2, sSc[sbase + t + 2]); m3 = fmaxf(m3, sSc[sbase + t + 3]);
        }
        for (; t < L; ++t) m0 = fmaxf(m0, sSc[sbase + t]);
        ex = expf(sc - fmaxf(fmaxf(m0, m1), fmaxf(m2, m3)));
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (half == 0) sSc[row] = ex;
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (half == 0) {
        float al = 0.f;
        if (valid) {
          float sum = 0.f;                     // same order as the reference sum over t (fp32, sequential)
          for (int t = 0; t < L; ++t) sum += sSc[sbase + t];
          al = ex / sum;
          p.alphas[gr] = al;
        }
        sAl[row] = al;
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      AF_MARK(4);
      // ---- e = sum_t alpha_t h_t.  lane = (16-byte chunk c of 4, time group tg of 8): the eight time groups of a chunk sit
      // in one warp and meet through shuffles.  Every thread first requests ALL of its rows of the tile (rows tg, tg + 8,
      // ... of up to 128: one L2 round trip for the whole tile, h was just put there by the TMA load), then adds each
      // row into the accumulator of the sentence it belongs to (predicated, at most 4 sentences per tile here; tiles
      // of more sentences take the loop below).
      const int x_first = (int)(row0 / L);
      const int chunks = H / 8;                // <= 32
      const int tg = lane & 7, c = e * 4 + (lane >> 3);
      if (SPT > 0) {
        // every load of the thread for the whole tile is requested first (one L2 round trip), then one accumulator
        // set per sentence, eight time groups meeting through shuffles
        uint4 w[SPT > 0 ? SPT * NK : 1];
#pragma unroll
        for (int s2 = 0; s2 < SPT; ++s2) {
          const bool live = x_first + s2 < p.X && c < chunks;
          const uint4* hp = reinterpret_cast<const uint4*>(p.h + (row0 + (long long)s2 * L) * H);
#pragma unroll
          for (int k = 0; k < NK; ++k) {
            const int t = tg + 8 * k;
            w[s2 * NK + k] = (live && t < L) ? __ldcg(hp + (long long)t * chunks + c) : make_uint4(0u, 0u, 0u, 0u);
          }
        }
#pragma unroll
        for (int s2 = 0; s2 < SPT; ++s2) {
          if (x_first + s2 >= p.X) break;
          float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int k = 0; k < NK; ++k) {
            const int t = tg + 8 * k;
            const float a = t < L ? sAl[s2 * L + t] : 0.f;
            const uint4 ww = w[s2 * NK + k];
            acc[0] = fmaf(a, __uint_as_float(ww.x << 16), acc[0]); acc[1] = fmaf(a, __uint_as_float(ww.x & 0xFFFF0000u), acc[1]);
            acc[2] = fmaf(a, __uint_as_float(ww.y << 16), acc[2]); acc[3] = fmaf(a, __uint_as_float(ww.y & 0xFFFF0000u), acc[3]);
            acc[4] = fmaf(a, __uint_as_float(ww.z << 16), acc[4]); acc[5] = fmaf(a, __uint_as_float(ww.z & 0xFFFF0000u), acc[5]);
            acc[6] = fmaf(a, __uint_as_float(ww.w << 16), acc[6]); acc[7] = fmaf(a, __uint_as_float(ww.w & 0xFFFF0000u), acc[7]);
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 1);
            acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 2);
            acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 4);
          }
          if (tg == 0 && c < chunks) {
            const int x = x_first + s2;
            const int g = x / p.group, r = x - g * p.group;
            const long long orow = (r < p.split) ? ((long long)g * p.split + r) : ((long long)g * (p.group - p.split) + (r - p.split));
            float* enc = ((r < p.split) ? p.enc_a + orow * H : p.enc_b + orow * H) + c * 8;
            *reinterpret_cast<float4*>(enc) = make_float4(acc[0], acc[1], acc[2], acc[3]);
            *reinterpret_cast<float4*>(enc + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
            if (p.enc_a_bf) {
              __nv_bfloat16* eb = ((r < p.split) ? p.enc_a_bf + orow * H : p.enc_b_bf + orow * H) + c * 8;
              __nv_bfloat162 a = __floats2bfloat162_rn(acc[0], acc[1]), bb = __floats2bfloat162_rn(acc[2], acc[3]);
              __nv_bfloat162 cc = __floats2bfloat162_rn(acc[4], acc[5]), d = __floats2bfloat162_rn(acc[6], acc[7]);
              *reinterpret_cast<uint4*>(eb) = make_uint4(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&bb),
                                                         *reinterpret_cast<uint32_t*>(&cc), *reinterpret_cast<uint32_t*>(&d));
            }
          }
        }
      } else {
      for (int s = 0; s < p.spt; ++s) {
        const int x = x_first + s;
        if (x >= p.X) break;
        float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (c < chunks) {
          const uint4* hp = reinterpret_cast<const uint4*>(p.h + ((long long)x * L) * H);
#pragma unroll 4
          for (int t = tg; t < L; t += 8) {
            const uint4 w = hp[(long long)t * chunks + c];
            const float a = sAl[s * L + t];
            acc[0] = fmaf(a, __uint_as_float(w.x << 16), acc[0]); acc[1] = fmaf(a, __uint_as_float(w.x & 0xFFFF0000u), acc[1]);
            acc[2] = fmaf(a, __uint_as_float(w.y << 16), acc[2]); acc[3] = fmaf(a, __uint_as_float(w.y & 0xFFFF0000u), acc[3]);
            acc[4] = fmaf(a, __uint_as_float(w.z << 16), acc[4]); acc[5] = fmaf(a, __uint_as_float(w.z & 0xFFFF0000u), acc[5]);
            acc[6] = fmaf(a, __uint_as_float(w.w << 16), acc[6]); acc[7] = fmaf(a, __uint_as_float(w.w & 0xFFFF0000u), acc[7]);
          }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 1);
          acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 2);
          acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 4);
        }
        if (tg == 0 && c < chunks) {
          const int g = x / p.group, r = x - g * p.group;
          const long long orow = (r < p.split) ? ((long long)g * p.split + r) : ((long long)g * (p.group - p.split) + (r - p.split));
          float* enc = ((r < p.split) ? p.enc_a + orow * H : p.enc_b + orow * H) + c * 8;
          *reinterpret_cast<float4*>(enc) = make_float4(acc[0], acc[1], acc[2], acc[3]);
          *reinterpret_cast<float4*>(enc + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
          if (p.enc_a_bf) {
            __nv_bfloat16* eb = ((r < p.split) ? p.enc_a_bf + orow * H : p.enc_b_bf + orow * H) + c * 8;
            __nv_bfloat162 a = __floats2bfloat162_rn(acc[0], acc[1]), bb = __floats2bfloat162_rn(acc[2], acc[3]);
            __nv_bfloat162 cc = __floats2bfloat162_rn(acc[4], acc[5]), d = __floats2bfloat162_rn(acc[6], acc[7]);
            *reinterpret_cast<uint4*>(eb) = make_uint4(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&bb),
                                                       *reinterpret_cast<uint32_t*>(&cc), *reinterpret_cast<uint32_t*>(&d));
          }
        }
      }
      }
      AF_MARK(5);
      asm volatile("bar.sync 1, 256;" ::: "memory");      // sAl / sSc are rewritten by the next tile
      AF_MARK(6);
    }
  }
  __syncthreads();
  if (warp == 1) tmem_dealloc(taddr, 512);
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
  if (H <= 256 && H % 64 == 0 && !getenv("INDNET_ATTN_FWD_RING")) {
    // persistent variant: Wa1 resident, double-buffered accumulator, eight epilogue warps
    p.NP = H;
    const int ntiles = ceil_div(X, p.spt);
    const int smem = (H / AF_BK) * (H * AF_BK * 2 + AF_A_BYTES) + (2 * H + 4 * AF_BM) * 4 + 8 * 2048 + 128;
    const int grid = ntiles < 148 ? ntiles : 148;
    if (p.spt == 3 && (L + 7) / 8 == 5) {           // max_length 40 (the reference's default and every training config)
      IN_CUDA(cudaFuncSetAttribute(attn_fused_fwd_pers_kernel<3, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      attn_fused_fwd_pers_kernel<3, 5><<<grid, AP_THREADS, smem, st>>>(tmA, tmB, p, ntiles);
    } else {
      IN_CUDA(cudaFuncSetAttribute(attn_fused_fwd_pers_kernel<0, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      attn_fused_fwd_pers_kernel<0, 0><<<grid, AP_THREADS, smem, st>>>(tmA, tmB, p, ntiles);
    }
    IN_LAUNCH_CHECK("attn_fused_fwd_pers_kernel");
    return INDNET_OK;
  }
  const int smem = AF_STAGES * (AF_A_BYTES + p.NP * AF_BK * 2) + (2 * H + 2 * AF_BM) * 4 + 128;
  IN_CUDA(cudaFuncSetAttribute(attn_fused_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  attn_fused_fwd_kernel<<<ceil_div(X, p.spt), AF_THREADS, smem, st>>>(tmA, tmB, p);
  IN_LAUNCH_CHECK("attn_fused_fwd_kernel");
  return INDNET_OK;
}

namespace indnet {

// ------------------------------------------------------------------------------------------------
// Backward of the attention block up to d(loss)/d(h), one kernel (H a multiple of 64, <= 256):
//   dalpha_t = denc . h_t ; dscore_t = alpha_t (dalpha_t - sum_s alpha_s dalpha_s)          (softmax + weighted sum)
//   dpre[t]  = dscore_t * w2 * (1 - proj_t^2) ; dw2 += dscore_t proj_t ; dba += dpre[t]     (score, tanh, bias)
//   dh[t]    = alpha_t * denc + dpre[t] * Wa1^T                                              (pooling path + Dense path)
// A CTA owns floor(128 / L) whole sentences.  h and proj tiles arrive by TMA (swizzled K-major tiles); dpre is
// written IN PLACE over the proj tile, which makes it the A operand of the tcgen05 GEMM dpre * Wa1^T (B = the bf16
// Dense kernel itself: [in][out] row-major is a K-major operand for this product), streamed through the shared
// memory the h tile no longer needs.  dpre also goes to HBM once (bf16) for the Dense-kernel gradient GEMM.
constexpr int AB_THREADS = 512;         // 16 row groups: the CUDA-core phases are latency bound, more warps hide it
constexpr int AB_RG = AB_THREADS / 32;
#ifdef INDNET_ATTN_TIMELINE
__device__ long long g_attn_dbg[16 * 8];
#define AB_MARK(k) do { if (blockIdx.x == 0 && threadIdx.x == 0 && it < 16) g_attn_dbg[it * 8 + (k)] = clock64(); } while (0)
#else
#define AB_MARK(k) do { } while (0)
#endif
constexpr int AB_TILE = AF_BM * 256 * 2;          // 64 KB: one [128 x 256] bf16 tile as 4 K-blocks of 16 KB
constexpr int AB_W_BYTES = 256 * 128;             // one K-block of Wa1: [<= 256 rows (in)][64 k (out)]

struct AttnFusedBwdP {
  const float* denc_a; const float* denc_b; const float* alphas; const float* w2;
  const __nv_bfloat16* h;
  __nv_bfloat16* dpre; __nv_bfloat16* dh;
  float* dw2; float* dba;
  int X, L, H, group, split, spt;
};

__device__ __forceinline__ void unpack8f(const uint4& q, float* f) {
  f[0] = __uint_as_float(q.x << 16); f[1] = __uint_as_float(q.x & 0xFFFF0000u);
  f[2] = __uint_as_float(q.y << 16); f[3] = __uint_as_float(q.y & 0xFFFF0000u);
  f[4] = __uint_as_float(q.z << 16); f[5] = __uint_as_float(q.z & 0xFFFF0000u);
  f[6] = __uint_as_float(q.w << 16); f[7] = __uint_as_float(q.w & 0xFFFF0000u);
}
__device__ __forceinline__ uint32_t pack2f(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// Persistent: one CTA per SM loops over tiles.  Shared memory (about 200 KB):
//   sW   the whole bf16 Dense kernel, resident (B operand, 4 K-blocks x [256 rows (in)][64 k (out)] swizzled) - loaded
//        once per CTA instead of once per tile, and the 16 MMAs of a tile then issue back to back
//   sP   proj tile -> dpre tile (A operand), 4 K-blocks x [128 rows][128 B] swizzled; the NEXT tile's proj is
//        fetched as soon as this tile's MMAs have completed, under the epilogue
//   sDenc [spt][H], sCol [2][H] (dw2 / dba column sums of the whole CTA), sAl [128], sDa [128]
__global__ void __launch_bounds__(AB_THREADS, 1)
attn_fused_bwd_kernel(const __grid_constant__ CUtensorMap tmP, const __grid_constant__ CUtensorMap tmW,
                      const __grid_constant__ CUtensorMap tmDpre, const __grid_constant__ CUtensorMap tmDh, AttnFusedBwdP p,
                      int ntiles) {
  extern __shared__ __align__(1024) uint8_t smem[];       // declared alignment keeps the shared address space visible
  uint8_t* sW = smem;
  uint8_t* sP = smem + 4 * AB_W_BYTES;
  float* sCol = reinterpret_cast<float*>(smem + 4 * AB_W_BYTES + AB_TILE);      // [2][256]
  float* sDenc = sCol + 2 * 256;
  const int L = p.L, H = p.H;
  const int rows_tile = p.spt * L;
  float* sAl = sDenc + 2 * (size_t)p.spt * H;   // [2][128] alpha          (sDenc is [2][spt][H]: this tile / next tile)
  float* sDa = sAl + 2 * AF_BM;                 // [128] dalpha, then dscore
  int* sSent = reinterpret_cast<int*>(sDa + AF_BM);     // [128] row -> sentence of the tile (r / L)
  float* sPart = reinterpret_cast<float*>(sSent + AF_BM);      // [128][33] partial dalpha sums (row, chunk)
  uint64_t* in_full = reinterpret_cast<uint64_t*>(sPart + 33 * AF_BM);      // 16896 bytes: keeps 8-byte alignment
  uint64_t* w_full = in_full + 1;
  uint64_t* acc_full = w_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long T = (long long)p.X * L;
  const int KB = H / 64;
  auto load_proj = [&](int tile) {
    mbar_expect_tx(in_full, (uint32_t)KB * AF_A_BYTES);
    for (int kb = 0; kb < KB; ++kb) tma_load_3d(sP + kb * AF_A_BYTES, &tmP, in_full, kb * 64, tile * rows_tile, 0);
  };
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmP); tma_prefetch_desc(&tmW);
    mbar_init(in_full, 1); mbar_init(w_full, 1); mbar_init(acc_full, 1);
    fence_barrier_init();
    mbar_expect_tx(w_full, (uint32_t)KB * H * 128);
    for (int kb = 0; kb < KB; ++kb) tma_load_3d(sW + kb * AB_W_BYTES, &tmW, w_full, kb * 64, 0, 0);   // Wa1 [n = in][k = out]
    if ((int)blockIdx.x < ntiles) load_proj(blockIdx.x);
  }
  if (warp == 1) tmem_alloc(tmem_slot, 256);
  for (int i = threadIdx.x; i < 2 * 256; i += AB_THREADS) sCol[i] = 0.f;
  for (int r = threadIdx.x; r < AF_BM; r += AB_THREADS) sSent[r] = r / L;
  // thread = (16-byte chunk c of a row: 8 features, row group g); a warp covers one whole row per step
  const int c = threadIdx.x & 31, g = threadIdx.x >> 5;
  const bool cvalid = c < H / 8;
  // operands that come straight from L2 / HBM are fetched one tile AHEAD (h chunks into registers right after phase A
  // has consumed the previous ones; denc rows and alphas into the other half of their double buffers)
  // phase A operands: coalesced (a warp reads one whole row per instruction: lane = 16-byte chunk, rows g + 16 i)
  constexpr int AB_RPT = AF_BM / AB_RG;                    // rows per thread
  uint4 hv[AB_RPT];
  auto fetch_h = [&](int tile) {
    const long long row0 = (long long)tile * rows_tile;
#pragma unroll
    for (int i = 0; i < AB_RPT; ++i) {
      const int r = g + AB_RG * i;
      const long long gr = row0 + r;
      hv[i] = (cvalid && r < rows_tile && gr < T) ? *reinterpret_cast<const uint4*>(p.h + gr * H + c * 8) : make_uint4(0, 0, 0, 0);
    }
  };
  auto fetch_small = [&](int tile, int buf) {
    // asynchronous global -> shared copies (no thread waits for the data here): warp s copies sentence s's upstream
    // gradient row (H floats, 16-byte pieces), the next warp the tile's attention weights
    const int x_first = tile * p.spt;
    float* dD = sDenc + (size_t)buf * p.spt * H;
    for (int s = warp; s <= p.spt; s += AB_RG) {
      if (s < p.spt) {
        const int x = x_first + s;
        if (x < p.X) {
          const int gg = x / p.group, r = x - gg * p.group;
          const float* src = (r < p.split) ? p.denc_a + ((long long)gg * p.split + r) * H
                                           : p.denc_b + ((long long)gg * (p.group - p.split) + (r - p.split)) * H;
          for (int n = lane * 4; n < H; n += 128)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(&dD[s * H + n])), "l"(src + n) : "memory");
        } else {
          for (int n = lane; n < H; n += 32) dD[s * H + n] = 0.f;
        }
      } else {
        const long long row0 = (long long)tile * rows_tile;
        for (int r = lane; r < AF_BM; r += 32) {
          const long long gr = row0 + r;
          if (r < rows_tile && gr < T)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(&sAl[buf * AF_BM + r])), "l"(p.alphas + gr) : "memory");
          else sAl[buf * AF_BM + r] = 0.f;
        }
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  if ((int)blockIdx.x < ntiles) { fetch_h(blockIdx.x); fetch_small(blockIdx.x, 0); }
  float w2v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) w2v[j] = cvalid ? p.w2[c * 8 + j] : 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t taddr = *tmem_slot;
  float dw[8], db[8];              // this thread's share of the column sums, over all its tiles
#pragma unroll
  for (int j = 0; j < 8; ++j) { dw[j] = 0.f; db[j] = 0.f; }

  int it = 0;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
    const long long row0 = (long long)tile * rows_tile;
    const int buf = it & 1;
    const float* tDenc = sDenc + (size_t)buf * p.spt * H;
    const float* tAl = sAl + buf * AF_BM;
    AB_MARK(0);
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();               // this tile's denc / alpha (fetched one tile ago) are in shared memory
    AB_MARK(1);
    // ---- phase A: dalpha_r = denc[sentence(r)] . h_r.  Each thread forms the partial sum of its 8 features for its
    // rows and parks it in shared memory ([row][33]); 128 threads then add up the 32 partials of their row - no
    // shuffle reductions on the path
#pragma unroll
    for (int i = 0; i < AB_RPT; ++i) {
      const int r = g + AB_RG * i;
      float dot = 0.f;
      if (cvalid && r < rows_tile) {
        float hf[8];
        unpack8f(hv[i], hf);
        const float4* de4 = reinterpret_cast<const float4*>(tDenc + sSent[r] * H + c * 8);
        const float4 d0 = de4[0], d1 = de4[1];
        dot = hf[0] * d0.x + hf[1] * d0.y + hf[2] * d0.z + hf[3] * d0.w + hf[4] * d1.x + hf[5] * d1.y + hf[6] * d1.z + hf[7] * d1.w;
      }
      sPart[r * 33 + c] = dot;
    }
    AB_MARK(2);
    // the next tile's operands: in flight under everything that follows
    if (tile + (int)gridDim.x < ntiles) { fetch_h(tile + gridDim.x); fetch_small(tile + gridDim.x, buf ^ 1); }
    __syncthreads();
    // dscore_r = alpha_r (dalpha_r - sum_s alpha_s dalpha_s): warp s reduces sentence s with shuffles
    for (int sw = warp; sw < p.spt; sw += AB_RG) {
      const int sb = sw * L;
      float da[4] = {0.f, 0.f, 0.f, 0.f}, al[4] = {0.f, 0.f, 0.f, 0.f}, part = 0.f;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int t = lane + 32 * k;
        if (t < L) {
          const int r = sb + t;
          float v = 0.f;
#pragma unroll 8
          for (int q4 = 0; q4 < 32; ++q4) v += sPart[r * 33 + q4];
          da[k] = v; al[k] = tAl[r];
          part = fmaf(al[k], v, part);
        }
      }
      const float dsum = warp_sum(part);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int t = lane + 32 * k;
        if (t < L) sDa[sb + t] = al[k] * (da[k] - dsum);
      }
    }
    for (int r = rows_tile + threadIdx.x; r < AF_BM; r += AB_THREADS) sDa[r] = 0.f;
    __syncthreads();
    AB_MARK(3);
    // ---- phase B: dpre in place over proj (bf16), its copy to HBM, and the column sums dw2 / dba
    mbar_wait(in_full, it & 1);
    AB_MARK(4);
    if (cvalid) {
      const uint32_t tile_off = (uint32_t)(c >> 3) * AF_A_BYTES, cc = (uint32_t)(c & 7);
#pragma unroll
      for (int i = 0; i < AF_BM / AB_RG; ++i) {
        const int r = g + AB_RG * i;
        uint4* slot = reinterpret_cast<uint4*>(sP + tile_off + r * 128 + ((cc ^ (uint32_t)(r & 7)) << 4));
        float pr[8], dp[8];
        unpack8f(*slot, pr);
        const float ds = sDa[r];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          dp[j] = ds * w2v[j] * (1.0f - pr[j] * pr[j]);
          dw[j] = fmaf(ds, pr[j], dw[j]);
          db[j] += dp[j];
        }
        const uint4 o = make_uint4(pack2f(dp[0], dp[1]), pack2f(dp[2], dp[3]), pack2f(dp[4], dp[5]), pack2f(dp[6], dp[7]));
        *slot = o;
      }
    }
    AB_MARK(5);
    fence_proxy_async();             // dpre tile -> visible to the MMA's operand reads
    __syncthreads();
    AB_MARK(6);
    // ---- dh_acc[128, H] = dpre[128, H] * Wa1^T: one thread issues 4 x KB MMAs back to back; the same tile goes to
    // HBM (bf16, for the Dense-kernel gradient GEMM) by TMA stores of exactly this tile's rows
    if (warp == 1 && lane == 0) {
      tc_fence_after();
      if (it == 0) mbar_wait(w_full, 0);
      tc_fence_after();
      const uint32_t idesc = idesc_bf16(AF_BM, H, false, false);
      for (int kb = 0; kb < KB; ++kb) {
        const uint32_t a_base = smem_u32(sP + kb * AF_A_BYTES), b_base = smem_u32(sW + kb * AB_W_BYTES);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16(taddr, smem_desc_sw128(a_base + k * 32, 0, 1024), smem_desc_sw128(b_base + k * 32, 0, 1024), idesc,
                    (kb > 0 || k > 0) ? 1u : 0u);
      }
      umma_commit(acc_full);
      for (int kb = 0; kb < KB; ++kb) tma_store_3d(&tmDpre, smem_u32(sP + kb * AF_A_BYTES), kb * 64, (int)row0, 0);
      bulk_commit();
      bulk_wait_read<0>();           // sP has been read by the stores (the MMAs are tracked by acc_full)
    }
    // ---- epilogue: dh = acc + alpha_r * denc[sentence(r)] -> bf16 into the (now free) A tile in the swizzled layout,
    // then TMA stores: global stores with lane = row would touch 32 lines per instruction
    {
      const int q = warp & 3, half = warp >> 2;          // column group of this warp (AB_RG / 4 groups)
      const int r = q * 32 + lane;
      const bool valid = r < rows_tile && row0 + r < T;
      mbar_wait(acc_full, it & 1);
      tc_fence_after();
      AB_MARK(7);
      __syncthreads();               // warp 1 lane 0 has also seen its dpre stores read the tile
      const float al = tAl[r];
      const float* de = tDenc + (valid ? sSent[r] : 0) * H;
      const int nchunks = H / 32;
      for (int ch = half; ch < nchunks; ch += AB_RG / 4) {
        float v[32];
        tmem_ld32(taddr + ((uint32_t)(q * 32) << 16) + ch * 32, v);
        tmem_ld_wait();
        uint8_t* rowp = sP + (ch >> 1) * AF_A_BYTES + r * 128;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float o[8];
          const float4 d0 = *reinterpret_cast<const float4*>(de + ch * 32 + j * 8), d1 = *reinterpret_cast<const float4*>(de + ch * 32 + j * 8 + 4);
          o[0] = fmaf(al, d0.x, v[j * 8 + 0]); o[1] = fmaf(al, d0.y, v[j * 8 + 1]); o[2] = fmaf(al, d0.z, v[j * 8 + 2]); o[3] = fmaf(al, d0.w, v[j * 8 + 3]);
          o[4] = fmaf(al, d1.x, v[j * 8 + 4]); o[5] = fmaf(al, d1.y, v[j * 8 + 5]); o[6] = fmaf(al, d1.z, v[j * 8 + 6]); o[7] = fmaf(al, d1.w, v[j * 8 + 7]);
          const uint32_t chunk16 = (uint32_t)((ch & 1) * 4 + j);
          *reinterpret_cast<uint4*>(rowp + ((chunk16 ^ (uint32_t)(r & 7)) << 4)) =
              make_uint4(pack2f(o[0], o[1]), pack2f(o[2], o[3]), pack2f(o[4], o[5]), pack2f(o[6], o[7]));
        }
      }
      tc_fence_before();
      fence_proxy_async();
      __syncthreads();
      if (threadIdx.x == 0) {
        for (int kb = 0; kb < KB; ++kb) tma_store_3d(&tmDh, smem_u32(sP + kb * AF_A_BYTES), kb * 64, (int)row0, 0);
        bulk_commit();
        bulk_wait_read<0>();
        // the tile is free again: fetch the next tile's proj (it lands under the next tile's phase A)
        if (tile + (int)gridDim.x < ntiles) load_proj(tile + gridDim.x);
      }
    }
  }
  if (threadIdx.x == 0) bulk_wait_all<0>();
  __syncthreads();
  // ---- the CTA's column sums: shared-memory reduction over the row groups, one global atomic per column
  if (cvalid) {
#pragma unroll
    for (int j = 0; j < 8; ++j) { atomicAdd(&sCol[c * 8 + j], dw[j]); atomicAdd(&sCol[256 + c * 8 + j], db[j]); }
  }
  __syncthreads();
  if (threadIdx.x < H) {
    atomicAdd(p.dw2 + threadIdx.x, sCol[threadIdx.x]);
    atomicAdd(p.dba + threadIdx.x, sCol[256 + threadIdx.x]);
  }
  if (warp == 1) { tc_fence_after(); tmem_dealloc(taddr, 256); }
}

}  // namespace indnet

#ifdef INDNET_ATTN_TIMELINE
extern "C" int indnet_debug_attn_fwd(long long* out) { return (int)cudaMemcpyFromSymbol(out, indnet::g_attnf_dbg, sizeof(long long) * 128); }
extern "C" int indnet_debug_attn(long long* out) { return (int)cudaMemcpyFromSymbol(out, indnet::g_attn_dbg, sizeof(long long) * 128); }
#endif
extern "C" int indnet_attn_fused_bwd(indnet_stream st_, const float* denc_a, const float* denc_b, const void* hout_bf,
                                     const void* proj_bf, const float* alphas, const float* w_a2, const void* wa1_bf,
                                     void* dpre_bf, void* dhout_bf, float* dw2_sum, float* dba_sum, int X, int L, int H,
                                     int group, int split, int accumulate) {
  using namespace indnet;
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(X >= 0 && L > 0 && H > 0, "attn_fused_bwd: bad sizes X=%d L=%d H=%d", X, L, H);
  IN_REQUIRE(L <= AF_BM && H % 64 == 0 && H <= 256, "attn_fused_bwd: needs L <= 128 and H a multiple of 64 up to 256 (L=%d H=%d)", L, H);
  IN_REQUIRE(denc_a && hout_bf && proj_bf && alphas && w_a2 && wa1_bf && dpre_bf && dhout_bf && dw2_sum && dba_sum,
             "attn_fused_bwd: null pointer");
  IN_REQUIRE(group > 0 && split >= 0 && split <= group && X % group == 0, "attn_fused_bwd: bad group/split %d/%d", group, split);
  IN_REQUIRE(split == group || denc_b, "attn_fused_bwd: second gradient input missing");
  IN_REQUIRE((((uintptr_t)dpre_bf | (uintptr_t)dhout_bf) & 15) == 0, "attn_fused_bwd: outputs must be 16-byte aligned");
  if (!accumulate) {
    IN_CUDA(cudaMemsetAsync(dw2_sum, 0, sizeof(float) * H, st));
    IN_CUDA(cudaMemsetAsync(dba_sum, 0, sizeof(float) * H, st));
  }
  if (X == 0) return INDNET_OK;
  AttnFusedBwdP p;
  p.denc_a = denc_a; p.denc_b = denc_b; p.alphas = alphas; p.w2 = w_a2;
  p.h = (const __nv_bfloat16*)hout_bf;
  p.dpre = (__nv_bfloat16*)dpre_bf; p.dh = (__nv_bfloat16*)dhout_bf; p.dw2 = dw2_sum; p.dba = dba_sum;
  p.X = X; p.L = L; p.H = H; p.group = group; p.split = split;
  p.spt = AF_BM / L < 4 ? AF_BM / L : 4;      // sentences per tile: their upstream-gradient rows are double buffered in shared memory
  IN_REQUIRE(((uintptr_t)hout_bf & 15) == 0, "attn_fused_bwd: hout_bf must be 16-byte aligned");
  CUtensorMap tmP, tmW;
  int rc = make_tmap_bf16(&tmP, proj_bf, H, (long long)X * L, H, 1, 0, 64, AF_BM);
  if (rc) return rc;
  rc = make_tmap_bf16(&tmW, wa1_bf, H, H, H, 1, 0, 64, H);              // [n = in][k = out]: box 64 k x H rows (n)
  if (rc) return rc;
  CUtensorMap tmDpre, tmDh;
  rc = make_tmap_bf16(&tmDpre, dpre_bf, H, (long long)X * L, H, 1, 0, 64, p.spt * L);       // box = exactly this tile's rows
  if (rc) return rc;
  rc = make_tmap_bf16(&tmDh, dhout_bf, H, (long long)X * L, H, 1, 0, 64, p.spt * L);
  if (rc) return rc;
  const int smem = AB_TILE + 4 * AB_W_BYTES + (2 * 256 + 2 * p.spt * H + (4 + 33) * AF_BM + 2) * 4 + 128;
  IN_REQUIRE(smem <= 227 * 1024, "attn_fused_bwd: L=%d H=%d does not fit shared memory", L, H);
  IN_CUDA(cudaFuncSetAttribute(attn_fused_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int ntiles = ceil_div(X, p.spt);
  attn_fused_bwd_kernel<<<ntiles < 148 ? ntiles : 148, AB_THREADS, smem, st>>>(tmP, tmW, tmDpre, tmDh, p, ntiles);
  IN_LAUNCH_CHECK("attn_fused_bwd_kernel");
  return INDNET_OK;
}
