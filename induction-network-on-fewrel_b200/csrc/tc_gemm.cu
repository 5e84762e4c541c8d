// bf16 tensor-core mode GEMM: C[M,N] = A * B (+ bias, tanh/relu, += C) with bf16 operands staged by
// TMA into 128-byte-swizzled shared memory, tcgen05.mma (M=128, N=128, K=16 per instruction)
// accumulating fp32 in TMEM, and a tcgen05.ld epilogue.  Warp-specialised: warp 0 = TMA producer,
// warp 1 = TMEM allocator + single-thread MMA issuer, warps 2-5 = epilogue (one TMEM lane quarter
// each).  4-stage smem ring, mbarrier full/empty pairs, tcgen05.commit releases stages.
//
// Operand layouts (no transposes anywhere in the pipeline - every contraction of the path maps onto
// one of these):
//   A "K-major"  : stored [M][K] row-major          A "MN-major" : stored [K][M] row-major
//   B "K-major"  : stored [N][K] row-major (B^T)    B "MN-major" : stored [K][N] row-major
// Batched (grid.z) through 3-D tensor maps, or split-K (grid.z) with fp32 atomic accumulation.
#include <cstring>
#include "common.cuh"
#include "indnet_internal.h"
#include "tc_common.cuh"
#include "../../include/indnet.h"

namespace indnet {

using namespace tc;

constexpr int G_BM = 128, G_BN = 128, G_BK = 64, G_STAGES = 3;   // 3 stages: 2 CTAs per SM
constexpr int G_A_BYTES = G_BM * G_BK * 2, G_B_BYTES = G_BN * G_BK * 2;   // 16 KB each
constexpr int G_SMEM = G_STAGES * (G_A_BYTES + G_B_BYTES) + 256 + 1024;   // + barriers + alignment slack
constexpr int G_THREADS = 192;

__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct TcGemmRank1 { const float* rs; const float* rv_a; const float* rv_b; int rv_rows, group, split; };

struct TcGemmP {
  float* C; __nv_bfloat16* Cb; const float* bias;
  // optional rank-1 term of the epilogue: C[m][n] += rs[m] * v(m / rv_rows)[n], v(x) = row x of the pair of
  // matrices (rv_a, rv_b) in the split layout of the attention kernels (group / split), N wide
  const float* rs; const float* rv_a; const float* rv_b; int rv_rows, rv_group, rv_split;
  int M, N, K, ldc;
  long long sC;        // batch stride of C (elements)
  int act, accumulate, splitk, kb_per_split, batch;
  int vec;        // 1: outputs / bias / rank-1 vectors are 16-byte aligned -> row-path epilogue
};

// NS = 1: plain bf16 operands.  NS = 3: FP32-accurate mode - every fp32 operand arrives as THREE bf16 matrices
// (head, middle, tail: x = h + m + l to 24 mantissa bits, indnet_cast_bf16_split3) and every K step issues the six
// products whose weight is at least 2^-16 of the leading one (hh, hm, mh, hl, lh, mm) into the same fp32
// accumulator: the contraction of the reference's fp32 graph (model/graph.py:47 dtype=tf.float32) on the tensor
// cores instead of FFMA.  tanh in the epilogue is then the accurate one.
struct TcMaps { CUtensorMap a[3]; CUtensorMap b[3]; };

template <bool A_MN, bool B_MN, int NS>
__global__ void __launch_bounds__(G_THREADS, NS == 3 ? 1 : 2)
tc_gemm_kernel(const __grid_constant__ TcMaps tm, TcGemmP p) {
  constexpr int STAGES = NS == 3 ? 2 : G_STAGES;
  constexpr int A_ST = NS * G_A_BYTES, B_ST = NS * G_B_BYTES;       // one stage: NS tiles of each operand
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * A_ST;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * (A_ST + B_ST));
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
  // NS == 3 only: the head*head products of one 64-wide K block go to one of two "main" accumulators, which the
  // epilogue warps drain into registers (round-to-nearest adds) while the other one fills; see the issuer below
  uint64_t* acc_full = tmem_full + 2;
  uint64_t* acc_empty = acc_full + 2;
  constexpr uint32_t TM_COLS = NS == 3 ? 512 : G_BN;                // main[0], main[1], small (, unused)
  constexpr uint32_t TM_SMALL = 2 * G_BN;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.y * G_BM, n0 = blockIdx.x * G_BN;
  const int z = blockIdx.z;
  const int batch = z / p.splitk, ks = z % p.splitk;
  const int total_kb = (p.K + G_BK - 1) / G_BK;
  const int kb0 = ks * p.kb_per_split;
  const int num_kb = min(total_kb, kb0 + p.kb_per_split) - kb0;

  if (warp == 0 && lane == 0) {
    for (int t = 0; t < NS; ++t) { tma_prefetch_desc(&tm.a[t]); tma_prefetch_desc(&tm.b[t]); }
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(tmem_full, 1);
    if (NS == 3) for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 128); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t taddr = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % STAGES, ph = (kb / STAGES) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        mbar_expect_tx(&full[s], A_ST + B_ST);
        const int k = (kb0 + kb) * G_BK;
#pragma unroll
        for (int t = 0; t < NS; ++t) {
          uint8_t* a = sA + s * A_ST + t * G_A_BYTES;
          uint8_t* b = sB + s * B_ST + t * G_B_BYTES;
          if (A_MN) {   // stored [K][M]: two boxes of 64 (M) x 64 (K rows), one per 128-byte M atom
            tma_load_3d(a, &tm.a[t], &full[s], m0, k, batch);
            tma_load_3d(a + G_A_BYTES / 2, &tm.a[t], &full[s], m0 + 64, k, batch);
          } else {      // stored [M][K]: one box of 64 (K) x 128 (M rows)
            tma_load_3d(a, &tm.a[t], &full[s], k, m0, batch);
          }
          if (B_MN) {
            tma_load_3d(b, &tm.b[t], &full[s], n0, k, batch);
            tma_load_3d(b + G_B_BYTES / 2, &tm.b[t], &full[s], n0 + 64, k, batch);
          } else {
            tma_load_3d(b, &tm.b[t], &full[s], k, n0, batch);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = idesc_bf16(G_BM, G_BN, A_MN, B_MN);
      // K-major: 8-row groups 1024 B apart, k-step = 32 B inside the 128 B row.
      // MN-major: 8-k-row groups 1024 B apart (SBO), 64-wide MN atoms G_BK*128 B apart (LBO), k-step = 16 rows.
      constexpr uint32_t a_lbo = A_MN ? G_BK * 128 : 0, b_lbo = B_MN ? G_BK * 128 : 0;
      constexpr uint32_t a_kstep = A_MN ? 16 * 128 : 32, b_kstep = B_MN ? 16 * 128 : 32;
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % STAGES, ph = (kb / STAGES) & 1;
        mbar_wait(&full[s], ph);
        if (NS == 3 && kb >= 2) mbar_wait(&acc_empty[kb & 1], ((kb >> 1) - 1) & 1);
        tc_fence_after();
        const uint32_t a_base = smem_u32(sA + s * A_ST), b_base = smem_u32(sB + s * B_ST);
#pragma unroll
        for (int k = 0; k < G_BK / 16; ++k) {
          if (NS == 1) {
            const uint64_t ad = smem_desc_sw128(a_base + k * a_kstep, a_lbo, 1024);
            const uint64_t bd = smem_desc_sw128(b_base + k * b_kstep, b_lbo, 1024);
            umma_bf16(taddr, ad, bd, idesc, (kb > 0 || k > 0) ? 1u : 0u);
          } else {
            // The tensor core adds each MMA's result to the TMEM accumulator with truncation, so a long chain
            // into one large accumulator drifts towards zero (measured 2.6e-5 of max|C| at K = 4096, against
            // 8e-7 for SGEMM).  Hence: the five cross terms of weight 2^-8 .. 2^-16 - (m,m) (l,h) (h,l) (m,h)
            // (h,m), smallest first - go to their own accumulator, 2^8 smaller than the result, for the whole
            // K range, and the head*head term of each K block goes to a fresh "main" accumulator that the
            // epilogue warps add into registers.  Operand t sits t tiles into the stage.
            constexpr int TA_[5] = {1, 2, 0, 1, 0}, TB_[5] = {1, 0, 2, 0, 1};
#pragma unroll
            for (int q = 0; q < 5; ++q) {
              const uint64_t ad = smem_desc_sw128(a_base + TA_[q] * G_A_BYTES + k * a_kstep, a_lbo, 1024);
              const uint64_t bd = smem_desc_sw128(b_base + TB_[q] * G_B_BYTES + k * b_kstep, b_lbo, 1024);
              umma_bf16(taddr + TM_SMALL, ad, bd, idesc, (kb > 0 || k > 0 || q > 0) ? 1u : 0u);
            }
            const uint64_t ad = smem_desc_sw128(a_base + k * a_kstep, a_lbo, 1024);
            const uint64_t bd = smem_desc_sw128(b_base + k * b_kstep, b_lbo, 1024);
            umma_bf16(taddr + (uint32_t)(kb & 1) * G_BN, ad, bd, idesc, k > 0 ? 1u : 0u);
          }
        }
        umma_commit(&empty[s]);
        if (NS == 3) umma_commit(&acc_full[kb & 1]);
      }
      umma_commit(tmem_full);
    }
  } else {
    // epilogue: warp w may only touch TMEM lanes [32*(w%4), 32*(w%4)+32)
    const int q = warp & 3;
    const int row = m0 + q * 32 + lane;
    if (NS == 3) {
      const uint32_t lane_base = taddr + ((uint32_t)(q * 32) << 16);
      float acc[G_BN];
#pragma unroll
      for (int j = 0; j < G_BN; ++j) acc[j] = 0.f;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&acc_full[kb & 1], (kb >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < G_BN / 32; ++c) {
          float v[32];
          tmem_ld32(lane_base + (uint32_t)(kb & 1) * G_BN + c * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) acc[c * 32 + j] += v[j];
        }
        tc_fence_before();
        mbar_arrive(&acc_empty[kb & 1]);
      }
      mbar_wait(tmem_full, 0);
      tc_fence_after();
      // total = sum of the K blocks + the small terms, parked in main[0] for the common epilogue below
#pragma unroll
      for (int c = 0; c < G_BN / 16; ++c) {
        float v[16];
        tmem_ld16(lane_base + TM_SMALL + c * 16, v);
        tmem_ld_wait();
        uint32_t r[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) r[j] = __float_as_uint(acc[c * 16 + j] + v[j]);
        tmem_st16(lane_base + c * 16, r);
      }
      tmem_st_wait();
    } else {
      mbar_wait(tmem_full, 0);
      tc_fence_after();
    }
    // All loads of this tile have been consumed once tmem_full fires, so the operand ring is free:
    // reuse it as a per-warp 32x33 fp32 transpose tile so that every global store instruction writes
    // one contiguous 128-byte row segment (lane = column) instead of 32 scattered rows.
    float* tile = reinterpret_cast<float*>(smem) + (warp - 2) * (32 * 33);
    const int row0 = m0 + q * 32;
    float* Cbase = p.C ? p.C + batch * p.sC : nullptr;
    __nv_bfloat16* Cbbase = p.Cb ? p.Cb + batch * p.sC : nullptr;
#pragma unroll 1
    for (int c = 0; c < G_BN / 32; ++c) {
      float v[32];
      tmem_ld32(taddr + ((uint32_t)(q * 32) << 16) + c * 32, v);
      tmem_ld_wait();
      const int nc = n0 + c * 32;
      if (p.vec && nc + 32 <= p.N) {
        // Row path (everything 16-byte aligned, full 32-column chunk): the lane keeps its own row, i.e. 32
        // consecutive output columns, and moves them with 16-byte loads / stores - no shared-memory
        // transpose, 4 (bf16) or 8 (fp32) store instructions per chunk instead of 32.
        if (row < p.M) {
          const long long o = (long long)row * p.ldc + nc;
          if (p.bias) {
            const float4* b4 = reinterpret_cast<const float4*>(p.bias + nc);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 b = __ldg(b4 + j);
              v[4 * j] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
            }
          }
          if (p.accumulate) {
            const float4* c4 = reinterpret_cast<const float4*>(Cbase + o);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 b = c4[j];
              v[4 * j] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
            }
          } else if (p.rs) {
            const float rs = p.rs[row];
            const int xr = row / p.rv_rows;
            const int g = xr / p.rv_group, r = xr - g * p.rv_group;
            const float* vp = (r < p.rv_split) ? p.rv_a + ((long long)g * p.rv_split + r) * p.N
                                               : p.rv_b + ((long long)g * (p.rv_group - p.rv_split) + (r - p.rv_split)) * p.N;
            const float4* v4 = reinterpret_cast<const float4*>(vp + nc);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 b = __ldg(v4 + j);
              v[4 * j] = fmaf(rs, b.x, v[4 * j]); v[4 * j + 1] = fmaf(rs, b.y, v[4 * j + 1]);
              v[4 * j + 2] = fmaf(rs, b.z, v[4 * j + 2]); v[4 * j + 3] = fmaf(rs, b.w, v[4 * j + 3]);
            }
          }
          if (p.act == 1) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = NS == 3 ? tanhf(v[j]) : tanh_approx(v[j]);
          } else if (p.act == 2) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
          }
          if (Cbase) {
            float4* c4 = reinterpret_cast<float4*>(Cbase + o);
#pragma unroll
            for (int j = 0; j < 8; ++j) c4[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
          }
          if (Cbbase) {
            uint4* c4 = reinterpret_cast<uint4*>(Cbbase + o);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              __nv_bfloat162 a = __floats2bfloat162_rn(v[8 * j], v[8 * j + 1]), b = __floats2bfloat162_rn(v[8 * j + 2], v[8 * j + 3]);
              __nv_bfloat162 cc = __floats2bfloat162_rn(v[8 * j + 4], v[8 * j + 5]), d = __floats2bfloat162_rn(v[8 * j + 6], v[8 * j + 7]);
              c4[j] = make_uint4(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b),
                                 *reinterpret_cast<uint32_t*>(&cc), *reinterpret_cast<uint32_t*>(&d));
            }
          }
        }
        continue;
      }
#pragma unroll
      for (int j = 0; j < 32; ++j) tile[lane * 33 + j] = v[j];
      if (p.rs) tile[lane * 33 + 32] = (row0 + lane < p.M) ? p.rs[row0 + lane] : 0.f;   // padding column: rs of row `lane`
      __syncwarp();
      const int n = n0 + c * 32 + lane;
      if (n < p.N) {
        const float bv = (p.bias && (p.splitk == 1 || ks == 0)) ? p.bias[n] : 0.f;
        const int rmax = min(32, p.M - row0);
        const long long o0 = (long long)row0 * p.ldc + n;
        if (p.splitk > 1) {
          for (int rr = 0; rr < rmax; ++rr) atomicAdd(Cbase + o0 + (long long)rr * p.ldc, tile[rr * 33 + lane] + bv);
        } else {
          // "+= C": fetch the whole column strip first so the 32 loads are in flight together
          // (interleaved with the stores they would each wait a full memory round trip)
          float old[32];
          if (p.accumulate) {
#pragma unroll
            for (int rr = 0; rr < 32; ++rr) old[rr] = (rr < rmax) ? Cbase[o0 + (long long)rr * p.ldc] : 0.f;
          } else if (p.rs) {
            // rank-1 term: rs of the strip's rows sits in the tile's padding column; a 32-row strip touches
            // at most a few vectors, each fetched once; the row -> vector index advances incrementally
            int xr = row0 / p.rv_rows, rem = row0 - xr * p.rv_rows;
            float vv = 0.f;
            bool fresh = true;
#pragma unroll
            for (int rr = 0; rr < 32; ++rr) {
              if (rr < rmax && fresh) {
                const int g = xr / p.rv_group, r = xr - g * p.rv_group;
                const float* vp = (r < p.rv_split) ? p.rv_a + ((long long)g * p.rv_split + r) * p.N
                                                   : p.rv_b + ((long long)g * (p.rv_group - p.rv_split) + (r - p.rv_split)) * p.N;
                vv = vp[n];
              }
              old[rr] = tile[rr * 33 + 32] * vv;
              fresh = (++rem == p.rv_rows);
              if (fresh) { rem = 0; ++xr; }
            }
          }
#pragma unroll
          for (int rr = 0; rr < 32; ++rr) {
            if (rr < rmax) {
              float x = tile[rr * 33 + lane] + bv;
              if (p.accumulate || p.rs) x += old[rr];
              if (p.act == 1) x = NS == 3 ? tanhf(x) : tanh_approx(x);   // bf16 mode: MUFU.TANH is well inside the 2e-2 budget
              else if (p.act == 2) x = fmaxf(x, 0.f);
              const long long o = o0 + (long long)rr * p.ldc;
              if (Cbase) Cbase[o] = x;
              if (Cbbase) Cbbase[o] = __float2bfloat16(x);
            }
          }
        }
      }
      __syncwarp();
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) tmem_dealloc(taddr, TM_COLS);
}

// ---------------------------------------------------------------- stacked weight-gradient GEMM
// dW = [A1 | A2]^T * B for the LSTM kernels: A1 = x [K][64] and A2 = h_prev [K][128] (both stored token-major,
// i.e. MN-major A), B = dz [K][N].  One CTA owns BOTH M tiles of the stacked operand [x(64); h(128); 0(64)]
// for its N tile and K slice, so every dz tile is fetched once and multiplied twice (two accumulators, 256
// TMEM columns).  Split-K over the tokens, fp32 atomics into the TF kernel gradient [D+u][N]: stacked row r
// lands in row r (r < D, the x features), is dropped (D <= r < 64, padding), or in row D + r - 64 (h units).
constexpr int D_STAGES = 4;
constexpr int D_A_BYTES = 4 * 64 * G_BK * 2;      // four 64-wide M atoms of 8 KB: x, h[0:64], h[64:128], zeros
constexpr int D_SMEM = D_STAGES * (D_A_BYTES + G_B_BYTES) + 256 + 1024;

struct TcDualP { float* C; float* dbias; int D, N, K, ldc, kb_per_split; };

__global__ void __launch_bounds__(G_THREADS, 1)
tc_gemm_dual_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmH,
                    const __grid_constant__ CUtensorMap tmB, TcDualP p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + D_STAGES * D_A_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + D_STAGES * (D_A_BYTES + G_B_BYTES));
  uint64_t* empty = full + D_STAGES;
  uint64_t* tmem_full = empty + D_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * G_BN, ks = blockIdx.z;
  const int total_kb = (p.K + G_BK - 1) / G_BK;
  const int kb0 = ks * p.kb_per_split;
  const int num_kb = min(total_kb, kb0 + p.kb_per_split) - kb0;

  // the fourth atom of every stage is the zero padding of the stacked operand: written once, never loaded
  for (int s = 0; s < D_STAGES; ++s)
    for (int i = threadIdx.x; i < 8192 / 16; i += G_THREADS)
      reinterpret_cast<uint4*>(sA + s * D_A_BYTES + 3 * 8192)[i] = make_uint4(0, 0, 0, 0);
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmH);
    tma_prefetch_desc(&tmB);
    for (int i = 0; i < D_STAGES; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 2 * G_BN);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t taddr = *tmem_slot;


  if (warp == 0) {
    if (lane == 0) {
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % D_STAGES, ph = (kb / D_STAGES) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        mbar_expect_tx(&full[s], 3 * 8192 + G_B_BYTES);
        const int k = (kb0 + kb) * G_BK;
        uint8_t* a = sA + s * D_A_BYTES;
        uint8_t* b = sB + s * G_B_BYTES;
        tma_load_3d(a, &tmX, &full[s], 0, k, 0);                 // x: 64 features x 64 tokens
        tma_load_3d(a + 8192, &tmH, &full[s], 0, k, 0);          // h_prev units 0..63
        tma_load_3d(a + 16384, &tmH, &full[s], 64, k, 0);        // h_prev units 64..127
        tma_load_3d(b, &tmB, &full[s], n0, k, 0);
        tma_load_3d(b + G_B_BYTES / 2, &tmB, &full[s], n0 + 64, k, 0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = idesc_bf16(G_BM, G_BN, true, true);
      constexpr uint32_t lbo = G_BK * 128, kstep = 16 * 128;     // MN-major operands, as in tc_gemm_kernel
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % D_STAGES, ph = (kb / D_STAGES) & 1;
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint32_t a_base = smem_u32(sA + s * D_A_BYTES), b_base = smem_u32(sB + s * G_B_BYTES);
#pragma unroll
        for (int k = 0; k < G_BK / 16; ++k) {
          const uint64_t bd = smem_desc_sw128(b_base + k * kstep, lbo, 1024);
          const uint64_t a0 = smem_desc_sw128(a_base + k * kstep, lbo, 1024);            // atoms x | h[0:64]
          const uint64_t a1 = smem_desc_sw128(a_base + 16384 + k * kstep, lbo, 1024);    // atoms h[64:128] | 0
          umma_bf16(taddr, a0, bd, idesc, (kb > 0 || k > 0) ? 1u : 0u);
          umma_bf16(taddr + G_BN, a1, bd, idesc, (kb > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&empty[s]);
      }
      umma_commit(tmem_full);
    }
  } else {
    // epilogue: warp w may only touch TMEM lanes [32*(w%4), 32*(w%4)+32); lane = stacked row within the M tile
    const int q = warp & 3;
    const int r = q * 32 + lane;
    mbar_wait(tmem_full, 0);
    tc_fence_after();
#pragma unroll 1
    for (int tile = 0; tile < 2; ++tile) {
      const int sr = tile * G_BM + r;                                      // stacked row
      const int dest = sr < 64 ? (sr < p.D ? sr : -1) : (sr < 64 + 128 ? p.D + sr - 64 : -1);
      // x's column D is 1.0: stacked row D is the column sum of dz = the LSTM bias gradient
      float* bias_row = (p.dbias && sr == p.D && p.D < 64) ? p.dbias : nullptr;
#pragma unroll 1
      for (int c = 0; c < G_BN / 32; ++c) {
        float v[32];
        tmem_ld32(taddr + tile * G_BN + ((uint32_t)(q * 32) << 16) + c * 32, v);
        tmem_ld_wait();
        if (dest >= 0 || bias_row) {
          // N is a multiple of the tile width here, rows are 16-byte aligned: 16-byte vector reductions
          float4* out = reinterpret_cast<float4*>((bias_row ? bias_row : p.C + (long long)dest * p.ldc) + n0 + c * 32);
#pragma unroll
          for (int j = 0; j < 8; ++j) atomicAdd(out + j, make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]));
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) tmem_dealloc(taddr, 2 * G_BN);
}

// ---------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

// bf16 tensor [batch][outer][inner] (inner contiguous, row stride ld elements), box = box_inner x box_outer x 1
int make_tmap_bf16(CUtensorMap* tm, const void* base, long long inner, long long outer, long long ld,
                   long long batch, long long batch_stride, int box_inner, int box_outer) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return set_error(INDNET_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  IN_REQUIRE((ld * 2) % 16 == 0, "tensor map: row stride %lld elements is not a multiple of 16 bytes", ld);
  IN_REQUIRE(((uintptr_t)base & 15) == 0, "tensor map: base pointer not 16-byte aligned");
  if (batch < 1) batch = 1;
  if (batch_stride <= 0) batch_stride = outer * ld;
  IN_REQUIRE((batch_stride * 2) % 16 == 0, "tensor map: batch stride is not a multiple of 16 bytes");
  cuuint64_t dims[3] = {(cuuint64_t)inner, (cuuint64_t)outer, (cuuint64_t)batch};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)batch_stride * 2};
  cuuint32_t box[3] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(INDNET_ERR_CUDA, "cuTensorMapEncodeTiled failed with %d", (int)r);
  return INDNET_OK;
}

// General tiled map: rank 3 or 4, element type bf16 (elem_bytes 2) or uint32 (4), innermost dimension contiguous;
// strides_bytes[i] = byte stride of dimension i + 1; 128-byte swizzle or none; out-of-range reads give zeros.
int make_tmap_nd(CUtensorMap* tm, const void* base, int elem_bytes, int rank, const unsigned long long* dims,
                 const unsigned long long* strides_bytes, const unsigned* box, bool swizzle128) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return set_error(INDNET_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  IN_REQUIRE(rank >= 1 && rank <= 5 && (elem_bytes == 2 || elem_bytes == 4), "tensor map: bad rank / element size");
  IN_REQUIRE(((uintptr_t)base & 15) == 0, "tensor map: base pointer not 16-byte aligned");
  cuuint64_t d[5], sb[4];
  cuuint32_t b[5], es[5];
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) {
    IN_REQUIRE(strides_bytes[i] % 16 == 0, "tensor map: stride %d is not a multiple of 16 bytes", i);
    sb[i] = strides_bytes[i];
  }
  CUresult r = fn(tm, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_UINT32, (cuuint32_t)rank,
                  const_cast<void*>(base), d, sb, b, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(INDNET_ERR_CUDA, "cuTensorMapEncodeTiled (rank %d) failed with %d", rank, (int)r);
  return INDNET_OK;
}

struct TcSplit { const void* a[2]; const void* b[2]; };      // middle and tail parts of A and B (FP32-accurate mode)

int tc_gemm(cudaStream_t st, bool a_mn, bool b_mn, int M, int N, int K, const void* A, long long lda,
            const void* B, long long ldb, float* C, void* Cb, int ldc, const float* bias, int act, int accumulate,
            int batch, long long sA, long long sB, long long sC, int splitk, const TcGemmRank1* r1 = nullptr,
            const TcSplit* split = nullptr) {
  IN_REQUIRE(M > 0 && N > 0 && K > 0 && batch > 0, "tc_gemm: bad sizes M=%d N=%d K=%d batch=%d", M, N, K, batch);
  IN_REQUIRE(A && B && (C || Cb), "tc_gemm: null operand");
  const int total_kb = ceil_div(K, G_BK);
  const int gx = ceil_div(N, G_BN), gy = ceil_div(M, G_BM);
  if (splitk <= 0) {
    splitk = 1;
    long long ctas = (long long)gx * gy * batch;
    if (batch == 1 && ctas < 148 && total_kb >= 16 && act == 0 && !Cb) {
      splitk = (int)((296 + ctas - 1) / ctas);
      if (splitk > total_kb / 4) splitk = total_kb / 4;
      if (splitk < 1) splitk = 1;
    }
  }
  IN_REQUIRE(splitk == 1 || (batch == 1 && act == 0 && !Cb && C), "tc_gemm: split-K needs fp32 output, no activation, no batch");
  TcGemmP p;
  p.C = C; p.Cb = (__nv_bfloat16*)Cb; p.bias = bias;
  p.rs = nullptr; p.rv_a = p.rv_b = nullptr; p.rv_rows = p.rv_group = 1; p.rv_split = 0;
  if (r1) {
    IN_REQUIRE(splitk == 1 && batch == 1 && !accumulate, "tc_gemm: the rank-1 epilogue needs splitk = batch = 1 and no accumulate");
    IN_REQUIRE(r1->rs && r1->rv_a && r1->rv_rows > 0 && r1->group > 0 && r1->split >= 0 && r1->split <= r1->group &&
               (r1->split == r1->group || r1->rv_b), "tc_gemm: bad rank-1 epilogue arguments");
    p.rs = r1->rs; p.rv_a = r1->rv_a; p.rv_b = r1->rv_b; p.rv_rows = r1->rv_rows; p.rv_group = r1->group; p.rv_split = r1->split;
  }
  p.M = M; p.N = N; p.K = K; p.ldc = ldc; p.sC = sC;
  p.act = act; p.accumulate = accumulate; p.splitk = splitk; p.batch = batch;
  p.kb_per_split = ceil_div(total_kb, splitk);
  splitk = ceil_div(total_kb, p.kb_per_split);      // no empty K slices
  p.splitk = splitk;
  {
    auto al16 = [](const void* q) { return ((uintptr_t)q & 15) == 0; };
    bool v = splitk == 1;
    if (C) v = v && al16(C) && ldc % 4 == 0 && (batch == 1 || sC % 4 == 0);
    if (Cb) v = v && al16(Cb) && ldc % 8 == 0 && (batch == 1 || sC % 8 == 0);
    if (bias) v = v && al16(bias);
    if (p.rs) v = v && al16(p.rv_a) && (!p.rv_b || al16(p.rv_b)) && N % 4 == 0;
    p.vec = v ? 1 : 0;
  }
  if (splitk > 1 && !accumulate) {
    IN_REQUIRE(ldc == N, "tc_gemm: split-K zero-fill expects a dense C");
    IN_CUDA(cudaMemsetAsync(C, 0, sizeof(float) * (size_t)M * N, st));
  }
  TcMaps tm;
  memset(&tm, 0, sizeof(tm));
  const int ns = split ? 3 : 1;
  for (int t = 0; t < ns; ++t) {
    const void* At = t == 0 ? A : split->a[t - 1];
    const void* Bt = t == 0 ? B : split->b[t - 1];
    IN_REQUIRE(At && Bt, "tc_gemm: missing operand part %d", t);
    int rc;
    if (a_mn) rc = make_tmap_bf16(&tm.a[t], At, M, K, lda, batch, sA, 64, G_BK);
    else rc = make_tmap_bf16(&tm.a[t], At, K, M, lda, batch, sA, G_BK, G_BM);
    if (rc) return rc;
    if (b_mn) rc = make_tmap_bf16(&tm.b[t], Bt, N, K, ldb, batch, sB, 64, G_BK);
    else rc = make_tmap_bf16(&tm.b[t], Bt, K, N, ldb, batch, sB, G_BK, G_BN);
    if (rc) return rc;
  }
  dim3 grid(gx, gy, batch * splitk);
  IN_REQUIRE(grid.z <= 65535, "tc_gemm: batch*splitk too large");
  constexpr int G_SMEM3 = 2 * 3 * (G_A_BYTES + G_B_BYTES) + 256 + 1024;
#define LAUNCH_TC2(AM, BM_, NS_, SM_)                                                                           \
  do {                                                                                                          \
    /* per device and cheap: set on every launch (legal during stream capture) */                                 \
    IN_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<AM, BM_, NS_>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_)); \
    tc_gemm_kernel<AM, BM_, NS_><<<grid, G_THREADS, SM_, st>>>(tm, p);                                          \
  } while (0)
#define LAUNCH_TC(AM, BM_) do { if (ns == 3) LAUNCH_TC2(AM, BM_, 3, G_SMEM3); else LAUNCH_TC2(AM, BM_, 1, G_SMEM); } while (0)
  if (a_mn && b_mn) LAUNCH_TC(true, true);
  else if (a_mn) LAUNCH_TC(true, false);
  else if (b_mn) LAUNCH_TC(false, true);
  else LAUNCH_TC(false, false);
#undef LAUNCH_TC
#undef LAUNCH_TC2
  IN_LAUNCH_CHECK("tc_gemm_kernel");
  return INDNET_OK;
}

__global__ void cast_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, long long n) {
  long long i = (blockIdx.x * (long long)blockDim.x + threadIdx.x) * 4;
  if (i + 3 < n) {
    const float4 v = *reinterpret_cast<const float4*>(in + i);
    __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
    uint2 o;
    o.x = *reinterpret_cast<uint32_t*>(&a);
    o.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(out + i) = o;
  } else {
    for (; i < n; ++i) out[i] = __float2bfloat16(in[i]);
  }
}

// out[r][0..ldo) (bf16) = in[r][0..cols) (fp32), zero in the padding columns [cols, ldo)
__global__ void cast_bf16_rows_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, long long rows,
                                      int cols, int ldo) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= rows * ldo) return;
  const long long r = i / ldo;
  const int c = (int)(i % ldo);
  out[i] = __float2bfloat16(c < cols ? in[r * cols + c] : 0.f);
}

// x = head + middle + tail, each a bf16 (24 mantissa bits in all): rows [rows][cols] with input stride ld_in are
// written with the padded stride ld_out (a multiple of 8 elements for TMA), zeros in the padding
__global__ void cast_split3_kernel(const float* __restrict__ in, long long ld_in, long long rows, int cols, int ld_out,
                                   __nv_bfloat16* __restrict__ hi, __nv_bfloat16* __restrict__ mid, __nv_bfloat16* __restrict__ lo) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= rows * ld_out) return;
  const long long r = i / ld_out;
  const int c = (int)(i - r * ld_out);
  const float x = c < cols ? in[r * ld_in + c] : 0.f;
  const __nv_bfloat16 h = __float2bfloat16(x);
  const float r1 = x - __bfloat162float(h);
  const __nv_bfloat16 m = __float2bfloat16(r1);
  const float r2 = r1 - __bfloat162float(m);
  hi[i] = h; mid[i] = m; lo[i] = __float2bfloat16(r2);
}

}  // namespace indnet

using namespace indnet;

extern "C" int indnet_cast_bf16_split3(indnet_stream st, const float* in, long long ld_in, long long rows, int cols,
                                       int ld_out, void* hi, void* mid, void* lo) {
  IN_REQUIRE(rows >= 0 && cols >= 0 && ld_out >= cols && ld_in >= cols, "cast_bf16_split3: bad sizes");
  if (rows == 0 || ld_out == 0) return INDNET_OK;
  IN_REQUIRE(in && hi && mid && lo, "cast_bf16_split3: null pointer");
  cast_split3_kernel<<<ceil_div(rows * ld_out, 256), 256, 0, (cudaStream_t)st>>>(in, ld_in, rows, cols, ld_out,
                                                                                (__nv_bfloat16*)hi, (__nv_bfloat16*)mid,
                                                                                (__nv_bfloat16*)lo);
  IN_LAUNCH_CHECK("cast_split3_kernel");
  return INDNET_OK;
}

extern "C" int indnet_tc_gemm3(indnet_stream st, int a_mn, int b_mn, int M, int N, int K, const void* Ah, const void* Am,
                               const void* Al, long long lda, const void* Bh, const void* Bm, const void* Bl,
                               long long ldb, float* C, int ldc, const float* bias, int act, int accumulate, int batch,
                               long long sA, long long sB, long long sC, int splitk) {
  TcSplit sp{{Am, Al}, {Bm, Bl}};
  return tc_gemm((cudaStream_t)st, a_mn != 0, b_mn != 0, M, N, K, Ah, lda, Bh, ldb, C, nullptr, ldc, bias, act, accumulate,
                 batch, sA, sB, sC, splitk, nullptr, &sp);
}

extern "C" int indnet_tc_gemm(indnet_stream st, int a_mn, int b_mn, int M, int N, int K, const void* A, long long lda,
                              const void* B, long long ldb, float* C, void* Cb, int ldc, const float* bias, int act,
                              int accumulate, int batch, long long sA, long long sB, long long sC, int splitk) {
  return tc_gemm((cudaStream_t)st, a_mn != 0, b_mn != 0, M, N, K, A, lda, B, ldb, C, Cb, ldc, bias, act, accumulate,
                 batch, sA, sB, sC, splitk);
}

extern "C" int indnet_cast_bf16(indnet_stream st, const float* in, void* out, long long n) {
  IN_REQUIRE(n >= 0, "cast_bf16: negative size");
  if (n == 0) return INDNET_OK;
  IN_REQUIRE(in && out, "cast_bf16: null pointer");
  IN_REQUIRE((((uintptr_t)in) & 15) == 0 && (((uintptr_t)out) & 7) == 0, "cast_bf16: unaligned pointers");
  cast_bf16_kernel<<<ceil_div(ceil_div(n, 4), 256), 256, 0, (cudaStream_t)st>>>(in, (__nv_bfloat16*)out, n);
  IN_LAUNCH_CHECK("cast_bf16_kernel");
  return INDNET_OK;
}

extern "C" int indnet_cast_bf16_rows(indnet_stream st, const float* in, void* out, long long rows, int cols, int ldo) {
  IN_REQUIRE(rows >= 0 && cols >= 0 && ldo >= cols, "cast_bf16_rows: bad sizes");
  if (rows == 0 || ldo == 0) return INDNET_OK;
  IN_REQUIRE(in && out, "cast_bf16_rows: null pointer");
  cast_bf16_rows_kernel<<<ceil_div(rows * ldo, 256), 256, 0, (cudaStream_t)st>>>(in, (__nv_bfloat16*)out, rows, cols, ldo);
  IN_LAUNCH_CHECK("cast_bf16_rows_kernel");
  return INDNET_OK;
}

extern "C" int indnet_tc_gemm_rank1(indnet_stream st, int a_mn, int b_mn, int M, int N, int K, const void* A,
                                    long long lda, const void* B, long long ldb, float* C, void* Cb, int ldc,
                                    const float* rs, const float* rv_a, const float* rv_b, int rv_rows, int group,
                                    int split) {
  TcGemmRank1 r1{rs, rv_a, rv_b, rv_rows, group, split};
  return tc_gemm((cudaStream_t)st, a_mn != 0, b_mn != 0, M, N, K, A, lda, B, ldb, C, Cb, ldc, nullptr, 0, 0, 1, 0, 0, 0,
                 1, &r1);
}

extern "C" int indnet_tc_gemm_lstm_dw(indnet_stream st_, const void* x_bf, const void* hprev_bf, const void* dz_bf,
                                      float* dkernel, float* dbias, long long T, int D, int u, int splitk,
                                      int accumulate) {
  cudaStream_t st = (cudaStream_t)st_;
  const int N = 4 * u;
  IN_REQUIRE(T > 0 && D > 0 && D <= 64 && u == 128, "tc_gemm_lstm_dw: needs E + 10 <= 64 and hidden_size 128 (D=%d u=%d)", D, u);
  IN_REQUIRE(x_bf && hprev_bf && dz_bf && dkernel, "tc_gemm_lstm_dw: null pointer");
  IN_REQUIRE(((uintptr_t)dkernel & 15) == 0, "tc_gemm_lstm_dw: gradient buffer not 16-byte aligned");
  IN_REQUIRE(T <= 2147483647LL, "tc_gemm_lstm_dw: too many tokens");
  const int total_kb = ceil_div((int)T, G_BK);
  const int gx = N / G_BN;
  if (splitk <= 0) splitk = 148 / gx;                        // one CTA per SM (160 KB of operand ring each)
  if (splitk > total_kb) splitk = total_kb;
  if (splitk < 1) splitk = 1;
  TcDualP p;
  IN_REQUIRE(!dbias || (((uintptr_t)dbias & 15) == 0 && D < 64), "tc_gemm_lstm_dw: dbias needs 16-byte alignment and D < 64");
  p.C = dkernel; p.dbias = dbias; p.D = D; p.N = N; p.K = (int)T; p.ldc = N;
  if (dbias && !accumulate) IN_CUDA(cudaMemsetAsync(dbias, 0, sizeof(float) * N, st));
  p.kb_per_split = ceil_div(total_kb, splitk);
  splitk = ceil_div(total_kb, p.kb_per_split);
  if (!accumulate) IN_CUDA(cudaMemsetAsync(dkernel, 0, sizeof(float) * (size_t)(D + u) * N, st));
  CUtensorMap tmX, tmH, tmB;
  int rc = make_tmap_bf16(&tmX, x_bf, 64, T, 64, 1, 0, 64, G_BK);           // [T][64]: box 64 features x 64 tokens
  if (rc) return rc;
  rc = make_tmap_bf16(&tmH, hprev_bf, u, T, u, 1, 0, 64, G_BK);             // [T][128]: boxes of 64 units
  if (rc) return rc;
  rc = make_tmap_bf16(&tmB, dz_bf, N, T, N, 1, 0, 64, G_BK);
  if (rc) return rc;
  // the attribute is per device and cheap to set: do it on every launch (legal during stream capture)
  IN_CUDA(cudaFuncSetAttribute(tc_gemm_dual_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, D_SMEM));
  tc_gemm_dual_kernel<<<dim3(gx, 1, splitk), G_THREADS, D_SMEM, st>>>(tmX, tmH, tmB, p);
  IN_LAUNCH_CHECK("tc_gemm_dual_kernel");
  return INDNET_OK;
}
