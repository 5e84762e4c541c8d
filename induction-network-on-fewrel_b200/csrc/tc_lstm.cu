// a3, bf16 tensor-core mode: the BiLSTM of model/graph.py:34-49 and its BPTT as persistent tcgen05 kernels
// whose gate threads never touch global memory.
//
// Forward, swap-AB.  Per step t and direction, for a tile of 32 sentences handled as two halves of 16:
//     Z^T[4u, 16] = [W_x^T | W_h^T][4u, 64+u] * [x_t ; h_{t-1}][64+u, 16]
// * MMA-M = gate units: four M=128 tiles, one per gate (i, j, f, o - the TF kernel column order), so TMEM
//   lane m of every tile is unit m and a thread reads all four gates of its unit from its own lane; MMA-N =
//   the 16 sentences of a half.
// * BOTH weight operands live in tensor memory for all L steps (W_h^T 256 columns, W_x^T 128 columns, staged
//   once with tcgen05.st): A-from-TMEM MMAs run at the math rate (8 cycles at N = 16), reading them from
//   shared memory costs 4 KB per instruction.  The remaining 128 columns are the accumulators (4 gates x 32).
// * The LSTM bias (with the +1.0 forget bias) is folded into the MMA: x carries 1.0 in columns D
//   and D+1 (indnet_embed_fwd_bf16) and W_x^T rows D / D+1 hold the bias as a bf16 head + bf16 remainder.
//   The i, f, o rows of the packed weights are pre-scaled by 0.5 (exact in bf16), so
//   sigmoid(z) = 0.5 + 0.5 * tanh(z / 2) costs one MUFU and one FMA.
// * The two halves are phase shifted: while the gate threads work on half A (tcgen05.ld, 5 MUFU per element,
//   bf16 stores to shared memory) the tensor pipe runs half B's MMAs and vice versa, so a step costs
//   max(gate math, MMA) instead of their sum.
// * Everything the gate threads produce goes to shared memory: h_t straight into the swizzled B-operand tile
//   of the next step; gates / c / dropped h into staging tiles.  ONE elected thread moves tiles: x_t in by
//   TMA loads (4-deep ring), h_t / h_{t-1} / c_t / gates out by TMA tile stores (bulk async groups; rows past
//   X are clipped by the tensor map).  All tiles are double buffered against the stores' reads.
// Warps 0-15: gate threads (TMEM lane quarter = warp % 4, sentence group = warp / 4); warp 16: the issuer.
//
// BPTT.  dh_rec^T[u, 16] = W_h[u, 4u] * dz^T[4u, 16] per half; W_h (rows D.. of the bf16 TF kernel) staged once
// into tensor memory.  Saved gates / c / upstream dh come in by TMA loads (2 stages per half), dz goes out by
// TMA stores from the very tile that is the MMA's B operand.  The position-column input gradient
// dx^T[f, 16] = W_x[f, 4u] * dz^T rides behind the recurrent MMAs (second A operand: 6 K-blocks in tensor
// memory, 2 in shared memory).  The bias gradients are NOT formed here: x's column D is 1.0, so row D of the
// stacked weight-gradient GEMM (indnet_tc_gemm_lstm_dw) is the column sum of dz.
#include <cstring>
#include "common.cuh"
#include "indnet_internal.h"
#include "tc_common.cuh"
#include "../../include/indnet.h"

namespace indnet {

using namespace tc;

constexpr int LT_U = 128;            // units per direction this kernel is specialised for
constexpr int LT_KX = 64;            // padded input width (D + 2 bias columns <= 64)
constexpr int LT_K = LT_KX + LT_U;   // 192
constexpr int LT_NS = 32;            // sentences per CTA
constexpr int LT_NH = 16;            // sentences per half (MMA N)
constexpr int LT_GATE_THREADS = 512;
constexpr int LT_THREADS = LT_GATE_THREADS + 32;
constexpr int LT_XBUFS = 4;          // x tiles in flight
// tensor-memory columns (forward)
constexpr uint32_t LT_TM_WH = 0, LT_TM_WX = 256, LT_TM_ACC = 384;
// shared-memory tiles (forward)
constexpr int LT_X_BYTES = LT_NS * 128;                 // x_t: 32 rows x 64 bf16
constexpr int LT_H_BYTES = 2 * LT_NH * 128;             // h of a half: 2 K-blocks x 16 rows x 128 B (swizzled)
constexpr int LT_G_BYTES = LT_NH * LT_U * 8;            // gates of a half: [16][128] x (i, j, f, o) bf16
constexpr int LT_C_BYTES = LT_NH * LT_U * 2;            // c of a half: [16][128] bf16
constexpr int LT_OFF_X = 0;
constexpr int LT_OFF_H = LT_OFF_X + LT_XBUFS * LT_X_BYTES;      // [half][buf]
constexpr int LT_OFF_O = LT_OFF_H + 4 * LT_H_BYTES;            // dropped h (DROP only), [half][buf]
constexpr int LT_OFF_G = LT_OFF_O + 4 * LT_H_BYTES;            // [half][buf] (SAVE only)
constexpr int LT_OFF_C = LT_OFF_G + 4 * LT_G_BYTES;            // [half][buf] (SAVE only)
constexpr int LT_OFF_BAR = LT_OFF_C + 4 * LT_C_BYTES;
constexpr int LT_SMEM = LT_OFF_BAR + 256 + 1024;

int make_tmap_bf16(CUtensorMap* tm, const void* base, long long inner, long long outer, long long ld,
                   long long batch, long long batch_stride, int box_inner, int box_outer);
int make_tmap_nd(CUtensorMap* tm, const void* base, int elem_bytes, int rank, const unsigned long long* dims,
                 const unsigned long long* strides_bytes, const unsigned* box, bool swizzle128);

// WT[dir][g*128 + m][k] (bf16, k-major, 192 wide), the A operand of the forward MMAs:
//   k < D            W_x[k][g*128+m]
//   k = D, D + 1     bias (+1.0 for the forget gate) as bf16 head and bf16 remainder (x carries 1.0 there)
//   D + 2 <= k < 64  0
//   k >= 64          W_h[k-64][g*128+m]
// all scaled by 0.5 for g = i, f, o (the kernel evaluates sigmoid(z) as 0.5 + 0.5 tanh(z/2)).
// kernel = TF BasicLSTMCell kernel [D+u][4u] (W_x rows then W_h rows).
__global__ void lstm_pack_weights_kernel(const float* __restrict__ k_fw, const float* __restrict__ k_bw,
                                         const float* __restrict__ b_fw, const float* __restrict__ b_bw,
                                         __nv_bfloat16* __restrict__ WT, int D) {
  const int dir = blockIdx.y;
  const float* __restrict__ kern = dir ? k_bw : k_fw;
  const float* __restrict__ bias = dir ? b_bw : b_fw;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;      // over 4u * K with the unit index fastest
  if (idx >= 4 * LT_U * LT_K) return;
  const int col = idx % (4 * LT_U), k = idx / (4 * LT_U);
  const int g = col / LT_U;
  const float sc = (g == 1) ? 1.0f : 0.5f;
  float v = 0.f;
  if (k < D) v = sc * kern[(long long)k * (4 * LT_U) + col];
  else if (k >= LT_KX) v = sc * kern[(long long)(D + k - LT_KX) * (4 * LT_U) + col];
  else if (k == D || k == D + 1) {
    const float b = sc * (bias[col] + (g == 2 ? 1.0f : 0.0f));
    const float head = __bfloat162float(__float2bfloat16(b));
    v = (k == D) ? head : b - head;
  }
  WT[((long long)dir * 4 * LT_U + col) * LT_K + k] = __float2bfloat16(v);
}

// Optional per-step timeline of CTA (0,0) (build with INDNET_NVCC_DEFINES=-DINDNET_LSTM_TIMELINE, read with
// tools/lstm_dbg.py); compiled out of the product build.
#ifdef INDNET_LSTM_TIMELINE
__device__ long long g_lstm_dbg[64 * 8];
__device__ long long g_lstm_dbg_b[64 * 8];
#define LT_MARK(cond, s, k) do { if ((cond) && blockIdx.x == 0 && blockIdx.y == 0 && (s) < 64) g_lstm_dbg[(s) * 8 + (k)] = clock64(); } while (0)
#define LB_MARK(cond, s, k) do { if ((cond) && blockIdx.x == 0 && blockIdx.y == 0 && (s) < 64) g_lstm_dbg_b[(s) * 8 + (k)] = clock64(); } while (0)
#else
#define LT_MARK(cond, s, k) do { } while (0)
#define LB_MARK(cond, s, k) do { } while (0)
#endif
__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ void sts_bf16(uint32_t saddr, __nv_bfloat16 v) {
  asm volatile("st.shared.b16 [%0], %1;" ::"r"(saddr), "h"(*reinterpret_cast<unsigned short*>(&v)) : "memory");
}
__device__ __forceinline__ void sts_u64(uint32_t saddr, uint32_t lo, uint32_t hi) {
  asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(saddr), "r"(lo), "r"(hi) : "memory");
}
__device__ __forceinline__ uint2 lds_u64(uint32_t saddr) {
  uint2 v;
  asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(saddr) : "memory");
  return v;
}
__device__ __forceinline__ float lds_bf16(uint32_t saddr) {
  unsigned short v;
  asm volatile("ld.shared.b16 %0, [%1];" : "=h"(v) : "r"(saddr) : "memory");
  return __uint_as_float((uint32_t)v << 16);
}
__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); }

struct LstmTcP {
  int X, L;
  float keep_prob; unsigned long long seed; const long long* seed_step;
  const __nv_bfloat16* wt;      // packed weights WT[dir][4u][192] (lstm_pack_weights_kernel)
};

// DROP: DropoutWrapper(output_keep_prob) on the emitted h (the recurrent h stays un-dropped).
// SAVE: also emit what BPTT needs (gates, c, h_{t-1}); off for inference.
// NHALF: 2 = a CTA owns 32 sentences as two phase-shifted halves (full chip); 1 = 16 sentences, no overlap but half
// the gate work per step - the host picks it when that still fits the chip in one wave (small batches are pure
// per-step latency).
template <bool DROP, bool SAVE, int NHALF>
__global__ void __launch_bounds__(LT_THREADS, 1)
tc_lstm_fwd_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmHout,
                   const __grid_constant__ CUtensorMap tmHprev, const __grid_constant__ CUtensorMap tmC,
                   const __grid_constant__ CUtensorMap tmG, LstmTcP p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* x_full = reinterpret_cast<uint64_t*>(smem + LT_OFF_BAR);     // [4]
  uint64_t* mma_done = x_full + LT_XBUFS;                                 // [2] per half
  uint64_t* h_ready = mma_done + 2;                                       // [2] per half
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(h_ready + 2);
  const uint32_t sX_s = smem_u32(smem + LT_OFF_X), sH_s = smem_u32(smem + LT_OFF_H), sO_s = smem_u32(smem + LT_OFF_O);
  const uint32_t sG_s = smem_u32(smem + LT_OFF_G), sC_s = smem_u32(smem + LT_OFF_C);

  const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;
  const bool gate_warp = warp < LT_GATE_THREADS / 32;
  const int dir = blockIdx.y;
  constexpr int NS = LT_NH * NHALF;
  const int x0 = blockIdx.x * NS;
  const int L = p.L;

  // h_{-1} = 0: the B operand of step 0 (buffer 1 of each half); the other tiles are written before they are read
  for (int i = threadIdx.x; i < 4 * LT_H_BYTES / 16; i += LT_THREADS)
    reinterpret_cast<uint4*>(smem + LT_OFF_H)[i] = make_uint4(0, 0, 0, 0);
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmHout);
    if (SAVE) { tma_prefetch_desc(&tmHprev); tma_prefetch_desc(&tmC); tma_prefetch_desc(&tmG); }
    for (int b = 0; b < LT_XBUFS; ++b) mbar_init(&x_full[b], 1);
    for (int h = 0; h < 2; ++h) { mbar_init(&mma_done[h], 1); mbar_init(&h_ready[h], LT_GATE_THREADS); }
    fence_barrier_init();
  }
  __syncwarp();
  if (warp == 0) tmem_alloc(tmem_slot, 512);            // W_h^T 256 | W_x^T 128 | accumulators 4 gates x (16 + 16)
  fence_proxy_async();          // the zero-fill of sH must be visible to the tensor core's operand reads
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t taddr = *tmem_slot;
  if (gate_warp) {
    // stage the resident A operands: warp w -> gate w / 4, lane quarter w % 4; thread = one row (unit) of
    // [W_x^T | W_h^T]: 64 + 128 k values = 32 + 64 packed columns
    const int g = warp >> 2, row = (warp & 3) * 32 + lane;
    const uint4* src = reinterpret_cast<const uint4*>(p.wt + ((long long)(dir * 4 + g) * LT_U + row) * LT_K);
    const uint32_t lane_sel = (uint32_t)((warp & 3) * 32) << 16;
#pragma unroll
    for (int c = 0; c < 6; ++c) {
      uint32_t r[16];
#pragma unroll
      for (int v = 0; v < 4; ++v) {
        const uint4 q = src[c * 4 + v];
        r[v * 4 + 0] = q.x; r[v * 4 + 1] = q.y; r[v * 4 + 2] = q.z; r[v * 4 + 3] = q.w;
      }
      const uint32_t dst = (c < 2) ? (taddr + LT_TM_WX + g * 32 + c * 16) : (taddr + LT_TM_WH + g * 64 + (c - 2) * 16);
      tmem_st16(dst + lane_sel, r);
    }
    tmem_st_wait();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (!gate_warp) {
    // ---------------------------------------------------------------- issuer warp
    constexpr uint32_t idesc = idesc_bf16(128, LT_NH, false, false);
    auto tn = [&](int s) { return dir ? (L - 1 - s) : s; };
    auto load_x = [&](int s) {
      mbar_expect_tx(&x_full[s % LT_XBUFS], NS * 128);
      tma_load_3d(smem + LT_OFF_X + (s % LT_XBUFS) * LT_X_BYTES, &tmX, &x_full[s % LT_XBUFS], 0, tn(s), x0);
    };
    if (elect_one()) {
      for (int s = 0; s < LT_XBUFS && s < L; ++s) load_x(s);
      if (SAVE) {             // h_{t-1} of the first step is zero
        for (int h = 0; h < NHALF; ++h)
          for (int kb = 0; kb < 2; ++kb)
            tma_store_4d(&tmHprev, sH_s + (h * 2 + 1) * LT_H_BYTES + kb * (LT_NH * 128), kb * 64, tn(0), x0 + h * LT_NH, dir);
        bulk_commit();
      }
    }
    __syncwarp();
    for (int s = 0; s <= L; ++s) {
#pragma unroll 1
      for (int h = 0; h < NHALF; ++h) {
        if (s > 0) {
          mbar_wait_wd(&h_ready[h], (s - 1) & 1);       // gate threads published step s-1 of this half, drained its accumulators
          tc_fence_after();
        }
        if (s < L && h == 0) mbar_wait_wd(&x_full[s % LT_XBUFS], (s / LT_XBUFS) & 1);
        if (elect_one()) {
          // tiles of step s-2 (same buffers as step s) must have been read by their stores before the gate
          // threads may overwrite them, i.e. before this step's MMAs are committed
          // (the prologue's zero tile of half 0 has only one younger group at s = 1: wait for everything there)
          if (NHALF == 1 || (s == 1 && h == 0)) { if (s > 0) bulk_wait_read<0>(); }
          else if (s > 0) bulk_wait_read<1>();
          if (s < L) {
            const uint32_t xb16 = (sX_s + (s % LT_XBUFS) * LT_X_BYTES + h * (LT_NH * 128)) >> 4;
            const uint32_t hb16 = (sH_s + (h * 2 + ((s + 1) & 1)) * LT_H_BYTES) >> 4;       // h_{s-1}: buffer (s-1) & 1
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              const uint32_t acc = taddr + LT_TM_ACC + g * 32 + h * LT_NH;
#pragma unroll
              for (int k4 = 0; k4 < 4; ++k4)
                umma_bf16_ts(acc, taddr + LT_TM_WX + g * 32 + k4 * 8, smem_desc_sw128_k(xb16, k4 * 32), idesc, k4 > 0 ? 1u : 0u);
#pragma unroll
              for (int kk = 0; kk < 8; ++kk)
                umma_bf16_ts(acc, taddr + LT_TM_WH + g * 64 + kk * 8,
                             smem_desc_sw128_k(hb16, (kk >> 2) * (LT_NH * 128) + (kk & 3) * 32), idesc, 1u);
            }
            umma_commit(&mma_done[h]);
          }
          if (s > 0) {
            // tile stores of step s-1 (buffers (s-1) & 1)
            const int b = (s - 1) & 1, t = tn(s - 1), xs = x0 + h * LT_NH;
            const uint32_t hsrc = sH_s + (h * 2 + b) * LT_H_BYTES;
            const uint32_t osrc = DROP ? sO_s + (h * 2 + b) * LT_H_BYTES : hsrc;
#pragma unroll
            for (int kb = 0; kb < 2; ++kb) {
              tma_store_3d(&tmHout, osrc + kb * (LT_NH * 128), dir * LT_U + kb * 64, t, xs);
              if (SAVE && s < L) tma_store_4d(&tmHprev, hsrc + kb * (LT_NH * 128), kb * 64, tn(s), xs, dir);
            }
            if (SAVE) {
              tma_store_4d(&tmC, sC_s + (h * 2 + b) * LT_C_BYTES, 0, t, xs, dir);
              tma_store_4d(&tmG, sG_s + (h * 2 + b) * LT_G_BYTES, 0, t, xs, dir);
            }
            bulk_commit();
            // both halves of step s-1 have consumed x_{s-1}: its slot takes x_{s-1+XBUFS}
            if (h == NHALF - 1 && s - 1 + LT_XBUFS < L) load_x(s - 1 + LT_XBUFS);
          }
        }
        __syncwarp();
      }
    }
    if (elect_one()) bulk_wait_all<0>();
    __syncwarp();
  } else {
    // ---------------------------------------------------------------- gate warps
    // thread = (unit m = TMEM lane, sentence group sg): 4 sentences of half A, then 4 of half B, per step
    const int lq = warp & 3, sg = warp >> 2;
    const int m = lq * 32 + lane;
    const uint32_t acc0 = taddr + LT_TM_ACC + ((uint32_t)(lq * 32) << 16) + sg * 4;
    // element (row r, unit m) of a swizzled h tile: K-block m / 64, 16-byte chunk (m % 64) / 8 XOR (r % 8)
    const uint32_t hbase = (uint32_t)(m >> 6) * (LT_NH * 128) + (uint32_t)(m & 7) * 2;
    const uint32_t hchunk = (uint32_t)(m & 63) >> 3;
    uint32_t hoff[4];
#pragma unroll
    for (int qq = 0; qq < 4; ++qq) {
      const uint32_t r = (uint32_t)(sg * 4 + qq);
      hoff[qq] = hbase + r * 128 + ((hchunk ^ (r & 7)) << 4);
    }
    const uint32_t goff = (uint32_t)(sg * 4) * (LT_U * 8) + (uint32_t)m * 8;     // + qq * 1024
    const uint32_t coff = (uint32_t)(sg * 4) * (LT_U * 2) + (uint32_t)m * 2;     // + qq * 256
    uint32_t dkey = 0, dthr = 0, didx = 0;
    float inv_keep = 1.f;
    if (DROP) {
      unsigned long long seed = p.seed;
      if (p.seed_step) seed += (unsigned long long)(*p.seed_step) * 1000003ull;
      dkey = dropout_key(seed);
      dthr = dropout_threshold(p.keep_prob);
      inv_keep = 1.0f / p.keep_prob;
      // flat index of (sentence x0 + sg*4, time tn(0), feature dir*u + m) in [X][L][2u]
      didx = ((uint32_t)(x0 + sg * 4) * (uint32_t)L + (uint32_t)(dir ? L - 1 : 0)) * (2 * LT_U) + dir * LT_U + m;
    }
    const uint32_t drow = (uint32_t)L * (2 * LT_U);              // one sentence further
    const uint32_t dstep = dir ? (uint32_t)(-(2 * LT_U)) : (uint32_t)(2 * LT_U);   // one step further
    float c[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) c[q] = 0.f;

    for (int s = 0; s < L; ++s) {
      const int b = s & 1;
#pragma unroll
      for (int h = 0; h < NHALF; ++h) {
        mbar_wait_wd(&mma_done[h], s & 1);
        tc_fence_after();
        LT_MARK(threadIdx.x == 0, s, h * 4 + 0);
        float zi[4], zj[4], zf[4], zo[4];
        tmem_ld4(acc0 + h * LT_NH + 0 * 32, zi);
        tmem_ld4(acc0 + h * LT_NH + 1 * 32, zj);
        tmem_ld4(acc0 + h * LT_NH + 2 * 32, zf);
        tmem_ld4(acc0 + h * LT_NH + 3 * 32, zo);
        tmem_ld_wait();
        LT_MARK(threadIdx.x == 0, s, h * 4 + 1);
        const uint32_t tile = (uint32_t)(h * 2 + b);
        const uint32_t hdst = sH_s + tile * LT_H_BYTES, odst = sO_s + tile * LT_H_BYTES;
        const uint32_t gdst = sG_s + tile * LT_G_BYTES + goff, cdst = sC_s + tile * LT_C_BYTES + coff;
        uint32_t dhash = 0;
#pragma unroll
        for (int qq = 0; qq < 4; ++qq) {
          // the accumulators already hold z/2 (i, f, o) resp. z (j) with the bias added
          const float gi = fmaf(tanh_fast(zi[qq]), 0.5f, 0.5f);
          const float gj = tanh_fast(zj[qq]);
          const float gf = fmaf(tanh_fast(zf[qq]), 0.5f, 0.5f);
          const float go = fmaf(tanh_fast(zo[qq]), 0.5f, 0.5f);
          const float cn = fmaf(c[h * 4 + qq], gf, gi * gj);
          const float hv = tanh_fast(cn) * go;
          c[h * 4 + qq] = cn;
          sts_bf16(hdst + hoff[qq], __float2bfloat16(hv));           // next step's B operand (and h / h_prev stores)
          if (SAVE) {
            sts_u64(gdst + qq * (LT_U * 8), pack_bf16x2(gi, gj), pack_bf16x2(gf, go));
            sts_bf16(cdst + qq * (LT_U * 2), __float2bfloat16(cn));
          }
          if (DROP) {
            if ((qq & 1) == 0) dhash = dropout_hash(dkey, didx + (uint32_t)(h * LT_NH + qq) * drow);   // serves qq and qq + 1
            const bool keep = dropout_keep_field(dhash, qq & 1, dthr);
            sts_bf16(odst + hoff[qq], __float2bfloat16(keep ? hv * inv_keep : 0.f));
          }
        }
        LT_MARK(threadIdx.x == 0, s, h * 4 + 2);
        fence_proxy_async();      // shared-memory tiles -> visible to the MMA's operand reads and the TMA stores
        tc_fence_before();        // TMEM loads ordered before the next MMAs overwrite the accumulators
        mbar_arrive(&h_ready[h]);
        LT_MARK(threadIdx.x == 0, s, h * 4 + 3);
      }
      if (DROP) didx += dstep;
    }
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc(taddr, 512);
}

// ------------------------------------------------------------------------------------------------
// BPTT, bf16 tensor-core mode (see the file header).  Shared memory: per (half, stage) one input tile set
// {gates [16][128] x 8 B, dh [16][128] bf16, c_prev [16][128] bf16}, the c tile of the first step, the dz
// tile (B operand, [8 K-blocks][32 rows][128 B] swizzled, rows 0-15 / 16-31 = the halves, double buffered) and,
// with DX, two K-blocks of W_x.
constexpr int LB_KB = 4 * LT_U / 64;                            // 8 K-blocks of dz
constexpr int LB_IN_G = LT_NH * LT_U * 8, LB_IN_D = LT_NH * LT_U * 2, LB_IN_C = LT_NH * LT_U * 2;
constexpr int LB_IN_BYTES = LB_IN_G + LB_IN_D + LB_IN_C;       // 24576
constexpr int LB_Z_BYTES = LB_KB * LT_NS * 128;                 // 32768: dz of BOTH halves, [8 K-blocks][32 rows][128 B]
constexpr int LB_WX_TKB = 6;                                    // dx path: K-blocks of the W_x rows held in tensor memory
constexpr int LB_WX_BYTES = (LB_KB - LB_WX_TKB) * 128 * 128;    // ... the other two stay in shared memory (32 KB)
constexpr int LB_OFF_IN = 0;                                    // [half][stage]
constexpr int LB_OFF_C0 = LB_OFF_IN + 4 * LB_IN_BYTES;         // [half]
constexpr int LB_OFF_Z = LB_OFF_C0 + 2 * LB_IN_C;              // [buf]
constexpr int LB_OFF_WX = LB_OFF_Z + 2 * LB_Z_BYTES;
constexpr int LB_DXT_BYTES = LT_NS * 16 * 4;                    // dx of a step: [32 sentences][16 features] fp32
constexpr int LB_SMEM_NODX = LB_OFF_WX + 256 + 1024;
constexpr int LB_SMEM_DX = LB_OFF_WX + LB_WX_BYTES + 2 * LB_DXT_BYTES + 256 + 1024;
static_assert(LB_OFF_Z % 1024 == 0 && LB_OFF_WX % 1024 == 0, "swizzled tiles need 1024-byte alignment");
constexpr int LB_DX_LD = 16;                                    // dx2 row stride in floats (64-byte rows)
constexpr uint32_t LB_TM_WH = 0, LB_TM_ACC = 256, LB_TM_ACCX = 288, LB_TM_WX = 320;

struct LstmTcBwdP {
  const __nv_bfloat16* wh_fw; const __nv_bfloat16* wh_bw;     // W_h = rows D.. of the bf16 TF kernels, [u][4u]
  float* dx2; int dxn;                  // optional: dx2[dir][tok][16], the first dxn (<= 16) input-gradient features
  const __nv_bfloat16* wx_fw; const __nv_bfloat16* wx_bw; int wx_rows;   // rows dx_first.. of the TF kernels, valid rows
  int X, L;
  float keep_prob; unsigned long long seed; const long long* seed_step;
};

template <bool DROP, bool DX, int NHALF>
__global__ void __launch_bounds__(LT_THREADS, 1)
tc_lstm_bwd_kernel(const __grid_constant__ CUtensorMap tmG, const __grid_constant__ CUtensorMap tmC,
                   const __grid_constant__ CUtensorMap tmDh, const __grid_constant__ CUtensorMap tmDz,
                   const __grid_constant__ CUtensorMap tmXf, const __grid_constant__ CUtensorMap tmXb,
                   const __grid_constant__ CUtensorMap tmDx, LstmTcBwdP p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sWx = smem + LB_OFF_WX;                      // DX only: [2 kb][128 rows][128 B]
  uint8_t* sDx = sWx + LB_WX_BYTES;                     // DX only: [2][32][16] fp32 staging of the dx rows
  uint64_t* in_full = reinterpret_cast<uint64_t*>(smem + LB_OFF_WX + (DX ? LB_WX_BYTES + 2 * LB_DXT_BYTES : 0));   // [half][stage]
  uint64_t* c0_full = in_full + 4;                      // [half]
  uint64_t* z_ready = c0_full + 2;                      // [half]
  uint64_t* mma_done = z_ready + 2;                     // [half]
  uint64_t* dx_done = mma_done + 2;                     // one per step (both halves)
  uint64_t* wx_full = dx_done + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wx_full + 1);
  const uint32_t sIn_s = smem_u32(smem + LB_OFF_IN), sC0_s = smem_u32(smem + LB_OFF_C0), sZ_s = smem_u32(smem + LB_OFF_Z);

  const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;
  const bool gate_warp = warp < LT_GATE_THREADS / 32;
  const int dir = blockIdx.y;
  constexpr int NS = LT_NH * NHALF;
  const int x0 = blockIdx.x * NS;
  const int L = p.L;
  auto tb = [&](int step) { return dir ? step : (L - 1 - step); };     // BPTT walks the processing order backwards

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmG); tma_prefetch_desc(&tmC); tma_prefetch_desc(&tmDh); tma_prefetch_desc(&tmDz);
    for (int i = 0; i < 4; ++i) mbar_init(&in_full[i], 1);
    for (int h = 0; h < 2; ++h) {
      mbar_init(&c0_full[h], 1);
      mbar_init(&z_ready[h], LT_GATE_THREADS);
      mbar_init(&mma_done[h], 1);
    }
    mbar_init(dx_done, 1);
    mbar_init(wx_full, 1);
    fence_barrier_init();
    if (DX) {
      const CUtensorMap* tmX = dir ? &tmXb : &tmXf;
      tma_prefetch_desc(tmX);
      mbar_expect_tx(wx_full, LB_WX_BYTES);
      for (int kb = LB_WX_TKB; kb < LB_KB; ++kb) tma_load_3d(sWx + (kb - LB_WX_TKB) * 16384, tmX, wx_full, kb * 64, 0, 0);
    }
  }
  __syncwarp();
  if (warp == 0) tmem_alloc(tmem_slot, 512);            // W_h 256 | dh_rec 32 | dx 32 | W_x rows 192
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t taddr = *tmem_slot;
  if (gate_warp) {
    // stage W_h: warp w writes lane quarter w % 4, k range [(w / 4) * 128, +128) = 64 packed columns
    const int kq = warp >> 2, row = (warp & 3) * 32 + lane;
    const uint32_t lane_sel = (uint32_t)((warp & 3) * 32) << 16;
    const uint4* src = reinterpret_cast<const uint4*>((dir ? p.wh_bw : p.wh_fw) + (long long)row * (4 * LT_U) + kq * LT_U);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      uint32_t r[16];
#pragma unroll
      for (int v = 0; v < 4; ++v) {
        const uint4 q = src[c * 4 + v];
        r[v * 4 + 0] = q.x; r[v * 4 + 1] = q.y; r[v * 4 + 2] = q.z; r[v * 4 + 3] = q.w;
      }
      tmem_st16(taddr + LB_TM_WH + lane_sel + kq * (LT_U / 2) + c * 16, r);
    }
    if (DX) {
      // W_x rows, K-blocks 0..5: k range [(w / 4) * 96, +96) = 48 packed columns
      const bool have = row < p.wx_rows;
      const uint4* xs = reinterpret_cast<const uint4*>((dir ? p.wx_bw : p.wx_fw) + (long long)(have ? row : 0) * (4 * LT_U) + kq * 96);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        uint32_t r[16];
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          const uint4 q = have ? xs[c * 4 + v] : make_uint4(0, 0, 0, 0);
          r[v * 4 + 0] = q.x; r[v * 4 + 1] = q.y; r[v * 4 + 2] = q.z; r[v * 4 + 3] = q.w;
        }
        tmem_st16(taddr + LB_TM_WX + lane_sel + kq * 48 + c * 16, r);
      }
    }
    tmem_st_wait();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (!gate_warp) {
    // ---------------------------------------------------------------- issuer warp
    constexpr uint32_t idesc = idesc_bf16(128, LT_NH, false, false);
    const uint32_t wx16 = smem_u32(sWx) >> 4;
    auto load_in = [&](int step, int h) {      // inputs of processing step `step`, half h -> stage step & 1
      uint64_t* bar = &in_full[h * 2 + (step & 1)];
      uint8_t* dst = smem + LB_OFF_IN + (h * 2 + (step & 1)) * LB_IN_BYTES;
      const int xs = x0 + h * LT_NH, t = tb(step);
      const bool more = step + 1 < L;
      mbar_expect_tx(bar, LB_IN_G + LB_IN_D + (more ? LB_IN_C : 0));
      tma_load_4d(dst, &tmG, bar, 0, t, xs, dir);
      tma_load_3d(dst + LB_IN_G, &tmDh, bar, dir * LT_U, t, xs);
      if (more) tma_load_4d(dst + LB_IN_G + LB_IN_D, &tmC, bar, 0, tb(step + 1), xs, dir);
    };
    if (elect_one()) {
      for (int h = 0; h < NHALF; ++h) {
        mbar_expect_tx(&c0_full[h], LB_IN_C);
        tma_load_4d(smem + LB_OFF_C0 + h * LB_IN_C, &tmC, &c0_full[h], 0, tb(0), x0 + h * LT_NH, dir);
      }
      for (int step = 0; step < 2 && step < L; ++step)
        for (int h = 0; h < NHALF; ++h) load_in(step, h);
    }
    __syncwarp();
    if (DX) mbar_wait_wd(wx_full, 0);
    // dx of a finished step: the 10 position features sit in TMEM lanes 0..9, which this warp (16 % 4 == 0) can
    // read; they leave through a small staging tile and one TMA store, so no gate thread touches global memory
    auto drain_dx = [&](int step_done) {
      float v[32];
      mbar_wait_wd(dx_done, step_done & 1);
      tc_fence_after();
      tmem_ld32(taddr + LB_TM_ACCX, v);
      tmem_ld_wait();
      float* tile = reinterpret_cast<float*>(sDx + (step_done & 1) * LB_DXT_BYTES);
      if (lane < LB_DX_LD) {
#pragma unroll
        for (int q = 0; q < NS; ++q) tile[q * LB_DX_LD + lane] = v[q];
      }
      fence_proxy_async();
      tc_fence_before();
      __syncwarp();
      if (elect_one()) {
        tma_store_4d(&tmDx, smem_u32(tile), 0, tb(step_done), x0, dir);
        bulk_commit();
      }
      __syncwarp();
    };
    for (int step = 0; step < L; ++step) {
#pragma unroll 1
      for (int h = 0; h < NHALF; ++h) {
        mbar_wait_wd(&z_ready[h], step & 1);      // dz(step) of this half is in its tile; its input stage is free
        tc_fence_after();
        if (elect_one()) {
          // the gate threads start on dz(step+1) once the commit below fires; that buffer was last stored by the
          // group of step-1 (committed a whole half step ago): its reads must have finished first
          if (h == 0 && step > 0) bulk_wait_read<0>();
          const uint32_t zsrc = sZ_s + (uint32_t)(step & 1) * LB_Z_BYTES;
          const uint32_t zb16 = (zsrc + h * (LT_NH * 128)) >> 4;        // this half's 16 rows of every K-block
          if (step + 1 < L) {                     // the last step's dh_rec has no consumer
#pragma unroll
            for (int kk = 0; kk < 4 * LB_KB; ++kk)
              umma_bf16_ts(taddr + LB_TM_ACC + h * LT_NH, taddr + LB_TM_WH + kk * 8,
                           smem_desc_sw128_k(zb16, (kk >> 2) * (LT_NS * 128) + (kk & 3) * 32), idesc, kk > 0 ? 1u : 0u);
            umma_commit(&mma_done[h]);
          }
        }
        __syncwarp();
        if (h == NHALF - 1) {
          // dx of the previous step has had a whole step to finish: drain it (this warp reads TMEM lanes 0..31)
          // right before this step's dx MMAs reuse the accumulator - the issuer never waits on the tensor pipe
          if (DX && step > 0) drain_dx(step - 1);
          if (elect_one()) {
            const uint32_t zsrc = sZ_s + (uint32_t)(step & 1) * LB_Z_BYTES;
            if (DX) {
              // once per step for both halves (N = 32): the two shared-memory K-blocks cost their 4 KB operand
              // read per instruction whatever N is
              constexpr uint32_t idesc32 = idesc_bf16(128, NS, false, false);
              const uint32_t zf16 = zsrc >> 4;
#pragma unroll
              for (int kk = 0; kk < 4 * LB_KB; ++kk) {
                const uint64_t bd = smem_desc_sw128_k(zf16, (kk >> 2) * (LT_NS * 128) + (kk & 3) * 32);
                if ((kk >> 2) < LB_WX_TKB)
                  umma_bf16_ts(taddr + LB_TM_ACCX, taddr + LB_TM_WX + kk * 8, bd, idesc32, kk > 0 ? 1u : 0u);
                else
                  umma_bf16(taddr + LB_TM_ACCX, smem_desc_sw128_k(wx16, ((kk >> 2) - LB_WX_TKB) * 16384 + (kk & 3) * 32),
                            bd, idesc32, 1u);
              }
              umma_commit(dx_done);
            }
            // dz of this step to HBM for the weight-gradient GEMMs: the operand tile itself is the source
            const int t = tb(step);
#pragma unroll
            for (int kb = 0; kb < LB_KB; ++kb) tma_store_4d(&tmDz, zsrc + kb * (LT_NS * 128), kb * 64, t, x0, dir);
            bulk_commit();
          }
          __syncwarp();
        }
        if (elect_one()) {
          if (step + 2 < L) load_in(step + 2, h);
        }
        __syncwarp();
      }
    }
    if (DX) drain_dx(L - 1);
    if (elect_one()) bulk_wait_all<0>();
    __syncwarp();
  } else {
    // ---------------------------------------------------------------- gate-gradient warps
    const int lq = warp & 3, sg = warp >> 2;
    const int m = lq * 32 + lane;
    const uint32_t lane_sel = (uint32_t)(lq * 32) << 16;
    const uint32_t acc0 = taddr + LB_TM_ACC + lane_sel + sg * 4;
    // dz element (row r, column g*128 + m) of the swizzled tile: K-block g*2 + m/64
    const uint32_t zbase = (uint32_t)(m >> 6) * (LT_NS * 128) + (uint32_t)(m & 7) * 2;
    const uint32_t zchunk = (uint32_t)(m & 63) >> 3;
    uint32_t zoff[4];
#pragma unroll
    for (int qq = 0; qq < 4; ++qq) {
      const uint32_t r = (uint32_t)(sg * 4 + qq);
      zoff[qq] = zbase + r * 128 + ((zchunk ^ (r & 7)) << 4);
    }
    const uint32_t goff = (uint32_t)(sg * 4) * (LT_U * 8) + (uint32_t)m * 8;     // + qq * 1024
    const uint32_t coff = (uint32_t)(sg * 4) * (LT_U * 2) + (uint32_t)m * 2;     // + qq * 256
    uint32_t dkey = 0, dthr = 0, didx = 0;
    float inv_keep = 1.f;
    if (DROP) {
      unsigned long long seed = p.seed;
      if (p.seed_step) seed += (unsigned long long)(*p.seed_step) * 1000003ull;
      dkey = dropout_key(seed);
      dthr = dropout_threshold(p.keep_prob);
      inv_keep = 1.0f / p.keep_prob;
      didx = ((uint32_t)(x0 + sg * 4) * (uint32_t)L + (uint32_t)(dir ? 0 : L - 1)) * (2 * LT_U) + dir * LT_U + m;
    }
    const uint32_t drow = (uint32_t)L * (2 * LT_U);
    const uint32_t dstep = dir ? (uint32_t)(2 * LT_U) : (uint32_t)(-(2 * LT_U));
    float dcr[8], cc[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) { dcr[q] = 0.f; cc[q] = 0.f; }

    for (int step = 0; step < L; ++step) {
      const bool more = step + 1 < L;
#pragma unroll
      for (int h = 0; h < NHALF; ++h) {
        const uint32_t in_s = sIn_s + (uint32_t)(h * 2 + (step & 1)) * LB_IN_BYTES;
        mbar_wait_wd(&in_full[h * 2 + (step & 1)], (step >> 1) & 1);
        LB_MARK(threadIdx.x == 0, step, h * 4 + 0);
        if (step == 0) {
          mbar_wait_wd(&c0_full[h], 0);
#pragma unroll
          for (int qq = 0; qq < 4; ++qq) cc[h * 4 + qq] = lds_bf16(sC0_s + h * LB_IN_C + coff + qq * (LT_U * 2));
        }
        uint2 gw[4];
        float dh[4], cp[4];
        uint32_t dhash = 0;
#pragma unroll
        for (int qq = 0; qq < 4; ++qq) {
          gw[qq] = lds_u64(in_s + goff + qq * (LT_U * 8));
          dh[qq] = lds_bf16(in_s + LB_IN_G + coff + qq * (LT_U * 2));
          cp[qq] = more ? lds_bf16(in_s + LB_IN_G + LB_IN_D + coff + qq * (LT_U * 2)) : 0.f;
          if (DROP) {
            if ((qq & 1) == 0) dhash = dropout_hash(dkey, didx + (uint32_t)(h * LT_NH + qq) * drow);   // serves qq and qq + 1
            dh[qq] = dropout_keep_field(dhash, qq & 1, dthr) ? dh[qq] * inv_keep : 0.f;
          }
        }
        if (step > 0) {
          float dhr[4];
          mbar_wait_wd(&mma_done[h], (step - 1) & 1);
          tc_fence_after();
          tmem_ld4(acc0 + h * LT_NH, dhr);
          tmem_ld_wait();
#pragma unroll
          for (int qq = 0; qq < 4; ++qq) dh[qq] += dhr[qq];      // the recurrent part is added after the dropout mask
        }
        LB_MARK(threadIdx.x == 0, step, h * 4 + 1);
        const uint32_t zdst = sZ_s + (uint32_t)(step & 1) * LB_Z_BYTES + h * (LT_NH * 128);
#pragma unroll
        for (int qq = 0; qq < 4; ++qq) {
          const int q = h * 4 + qq;
          const float gi = bf16_lo(gw[qq].x), gj = bf16_hi(gw[qq].x), gf = bf16_lo(gw[qq].y), go = bf16_hi(gw[qq].y);
          const float tc_ = tanh_fast(cc[q]);
          const float dob = dh[qq] * tc_ * go * (1.f - go);
          const float dc = dcr[q] + dh[qq] * go * (1.f - tc_ * tc_);
          const float di = dc * gj * gi * (1.f - gi);
          const float dj = dc * gi * (1.f - gj * gj);
          const float df = dc * cp[qq] * gf * (1.f - gf);
          dcr[q] = dc * gf;
          cc[q] = cp[qq];                        // c_{t_prev} is the cell state of the next step processed
          const uint32_t zp = zdst + zoff[qq];
          sts_bf16(zp, __float2bfloat16(di));
          sts_bf16(zp + 2 * (LT_NS * 128), __float2bfloat16(dj));
          sts_bf16(zp + 4 * (LT_NS * 128), __float2bfloat16(df));
          sts_bf16(zp + 6 * (LT_NS * 128), __float2bfloat16(dob));
        }
        LB_MARK(threadIdx.x == 0, step, h * 4 + 2);
        fence_proxy_async();
        tc_fence_before();
        mbar_arrive(&z_ready[h]);
        LB_MARK(threadIdx.x == 0, step, h * 4 + 3);
      }
      if (DROP) didx += dstep;
    }
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc(taddr, 512);
}

// out[c] = sum_r A[r][c] for a bf16 matrix (bias gradient from the bf16 dz stream)
__global__ void colsum_bf16_kernel(const __nv_bfloat16* __restrict__ A, long long rows, int cols, float* out,
                                   int rows_per_block) {
  __shared__ float red[8][33];
  int tx = threadIdx.x % 32, ty = threadIdx.x / 32;
  int c = blockIdx.x * 32 + tx;
  long long r0 = (long long)blockIdx.y * rows_per_block;
  long long r1 = min(rows, r0 + rows_per_block);
  float s = 0.f;
  if (c < cols)
    for (long long r = r0 + ty; r < r1; r += 8) s += __bfloat162float(A[r * cols + c]);
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && c < cols) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][tx];
    atomicAdd(out + c, t);
  }
}

}  // namespace indnet

using namespace indnet;

#ifdef INDNET_LSTM_TIMELINE
extern "C" int indnet_debug_lstm(long long* out) {
  return (int)cudaMemcpyFromSymbol(out, g_lstm_dbg, sizeof(long long) * 512);
}
extern "C" int indnet_debug_lstm_bwd(long long* out) {
  return (int)cudaMemcpyFromSymbol(out, g_lstm_dbg_b, sizeof(long long) * 512);
}
#endif

namespace {
typedef unsigned long long u64;
// [2][X][L][inner] with `eb`-byte elements as a 4-D map (inner, L, X, 2), box (box_inner, 1, 16, 1)
int map_dir_tok(CUtensorMap* tm, const void* base, int eb, int inner, int X, int L, int box_inner, bool sw,
                int rows = LT_NH) {
  const u64 dims[4] = {(u64)inner, (u64)L, (u64)X, 2};
  const u64 strides[3] = {(u64)inner * eb, (u64)L * inner * eb, (u64)X * L * inner * eb};
  const unsigned box[4] = {(unsigned)box_inner, 1, (unsigned)rows, 1};
  return make_tmap_nd(tm, base, eb, 4, dims, strides, box, sw);
}
// [X][L][inner] bf16 as a 3-D map (inner, L, X), box (box_inner, 1, rows)
int map_tok(CUtensorMap* tm, const void* base, int inner, int X, int L, int box_inner, int rows, bool sw) {
  const u64 dims[3] = {(u64)inner, (u64)L, (u64)X};
  const u64 strides[2] = {(u64)inner * 2, (u64)L * inner * 2};
  const unsigned box[3] = {(unsigned)box_inner, 1, (unsigned)rows};
  return make_tmap_nd(tm, base, 2, 3, dims, strides, box, sw);
}
}  // namespace

// The weight packing of indnet_lstm_fwd_tc as a call of its own (it depends on the weights only, so a caller can run it
// beside the embedding gather and pass NULL weights to indnet_lstm_fwd_tc).
extern "C" int indnet_lstm_pack_tc(indnet_stream st_, const float* kernel_fw, const float* bias_fw, const float* kernel_bw,
                                   const float* bias_bw, void* wt_scratch, int D, int u) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(u == LT_U && D > 0 && D + 2 <= LT_KX, "lstm_pack_tc: needs hidden_size 128 and E + 12 <= 64 (u=%d D=%d)", u, D);
  IN_REQUIRE(kernel_fw && bias_fw && kernel_bw && bias_bw && wt_scratch, "lstm_pack_tc: null pointer");
  lstm_pack_weights_kernel<<<dim3(ceil_div(4 * LT_U * LT_K, 256), 2), 256, 0, st>>>(kernel_fw, kernel_bw, bias_fw, bias_bw,
                                                                                    (__nv_bfloat16*)wt_scratch, D);
  IN_LAUNCH_CHECK("lstm_pack_weights_kernel");
  return INDNET_OK;
}

extern "C" int indnet_lstm_fwd_tc(indnet_stream st_, const void* x_bf, const float* kernel_fw, const float* bias_fw,
                                  const float* kernel_bw, const float* bias_bw, void* wt_scratch, float* hout,
                                  void* hout_bf, void* gates_bf, void* cstate_bf, void* hprev_bf, int X, int L, int D,
                                  int u, float keep_prob, unsigned long long seed, const long long* seed_step) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(X >= 0 && L > 0, "lstm_fwd_tc: bad sizes X=%d L=%d", X, L);
  IN_REQUIRE(u == LT_U && D > 0 && D + 2 <= LT_KX,
             "lstm_fwd_tc: specialised for hidden_size 128 and E+10 <= 62 (got u=%d D=%d)", u, D);
  IN_REQUIRE(keep_prob > 0.f && keep_prob <= 1.f, "lstm_fwd_tc: keep_prob must be in (0, 1], got %f", keep_prob);
  IN_REQUIRE(hout == nullptr, "lstm_fwd_tc: the fp32 copy of the output is no longer produced - pass hout = NULL");
  if (X == 0) return INDNET_OK;
  const bool packed = !kernel_fw && !bias_fw && !kernel_bw && !bias_bw;     // wt_scratch filled by indnet_lstm_pack_tc
  IN_REQUIRE(x_bf && (packed || (kernel_fw && bias_fw && kernel_bw && bias_bw)) && wt_scratch && hout_bf, "lstm_fwd_tc: null pointer");
  const bool save = gates_bf != nullptr;
  IN_REQUIRE(save == (cstate_bf != nullptr) && save == (hprev_bf != nullptr),
             "lstm_fwd_tc: pass all of gates_bf / cstate_bf / hprev_bf (training) or none (inference)");
  IN_REQUIRE((long long)X * L * (2 * LT_U) < 4294967296LL, "lstm_fwd_tc: X*L*2u exceeds 32-bit element indexing");
  if (!packed) {
    lstm_pack_weights_kernel<<<dim3(ceil_div(4 * LT_U * LT_K, 256), 2), 256, 0, st>>>(kernel_fw, kernel_bw, bias_fw, bias_bw,
                                                                                      (__nv_bfloat16*)wt_scratch, D);
    IN_LAUNCH_CHECK("lstm_pack_weights_kernel");
  }
  CUtensorMap tmX, tmHout, tmHprev, tmC, tmG;
  // one wave of 16-sentence CTAs when the batch is small enough, else 32-sentence CTAs with phase-shifted halves
  const int nhalf = (2 * ceil_div(X, LT_NH) <= 148) ? 1 : 2;
  const int ns = LT_NH * nhalf;
  int rc = map_tok(&tmX, x_bf, LT_KX, X, L, LT_KX, ns, true);
  if (rc) return rc;
  rc = map_tok(&tmHout, hout_bf, 2 * LT_U, X, L, 64, LT_NH, true);
  if (rc) return rc;
  tmHprev = tmC = tmG = tmHout;
  if (save) {
    rc = map_dir_tok(&tmHprev, hprev_bf, 2, LT_U, X, L, 64, true);
    if (rc) return rc;
    rc = map_dir_tok(&tmC, cstate_bf, 2, LT_U, X, L, LT_U, false);
    if (rc) return rc;
    rc = map_dir_tok(&tmG, gates_bf, 4, 2 * LT_U, X, L, 2 * LT_U, false);      // (i, j, f, o) bf16 = two 32-bit words per unit
    if (rc) return rc;
  }
  LstmTcP p;
  p.X = X; p.L = L; p.keep_prob = keep_prob; p.seed = seed; p.seed_step = seed_step;
  p.wt = (const __nv_bfloat16*)wt_scratch;
  dim3 grid(ceil_div(X, ns), 2);
  const bool drop = keep_prob < 1.0f;
#define LT_FWD_LAUNCH2(DROP_, SAVE_, NH_)                                                                       \
  do {                                                                                                          \
    /* per device and cheap: set on every launch (legal during stream capture) */                                 \
    IN_CUDA(cudaFuncSetAttribute(tc_lstm_fwd_kernel<DROP_, SAVE_, NH_>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                 LT_SMEM));                                                                     \
    tc_lstm_fwd_kernel<DROP_, SAVE_, NH_><<<grid, LT_THREADS, LT_SMEM, st>>>(tmX, tmHout, tmHprev, tmC, tmG, p);  \
  } while (0)
#define LT_FWD_LAUNCH(DROP_, SAVE_) do { if (nhalf == 1) LT_FWD_LAUNCH2(DROP_, SAVE_, 1); else LT_FWD_LAUNCH2(DROP_, SAVE_, 2); } while (0)
  if (drop && save) LT_FWD_LAUNCH(true, true);
  else if (drop) LT_FWD_LAUNCH(true, false);
  else if (save) LT_FWD_LAUNCH(false, true);
  else LT_FWD_LAUNCH(false, false);
#undef LT_FWD_LAUNCH2
#undef LT_FWD_LAUNCH
  IN_LAUNCH_CHECK("tc_lstm_fwd_kernel");
  return INDNET_OK;
}

extern "C" int indnet_lstm_bwd_tc(indnet_stream st_, const void* dhout_bf, const void* gates_bf, const void* cstate_bf,
                                  const void* kernel_fw_bf, const void* kernel_bw_bf, void* dz_bf, float* dx2,
                                  int dx_first, int X, int L, int D, int u, float keep_prob, unsigned long long seed,
                                  const long long* seed_step) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(X >= 0 && L > 0, "lstm_bwd_tc: bad sizes X=%d L=%d", X, L);
  IN_REQUIRE(u == LT_U && D > 0 && D <= LT_KX, "lstm_bwd_tc: specialised for hidden_size 128 and E+10 <= 64 (got u=%d D=%d)", u, D);
  IN_REQUIRE(keep_prob > 0.f && keep_prob <= 1.f, "lstm_bwd_tc: keep_prob must be in (0, 1], got %f", keep_prob);
  if (X == 0) return INDNET_OK;
  IN_REQUIRE(dhout_bf && gates_bf && cstate_bf && kernel_fw_bf && kernel_bw_bf && dz_bf, "lstm_bwd_tc: null pointer");
  IN_REQUIRE((long long)X * L * (2 * LT_U) < 4294967296LL, "lstm_bwd_tc: X*L*2u exceeds 32-bit element indexing");
  LstmTcBwdP p;
  // W_h = rows D.. of the bf16 TF kernel [D+u][4u]; D*4u*2 bytes is a multiple of 16 for every D
  p.wh_fw = (const __nv_bfloat16*)kernel_fw_bf + (long long)D * 4 * LT_U;
  p.wh_bw = (const __nv_bfloat16*)kernel_bw_bf + (long long)D * 4 * LT_U;
  IN_REQUIRE((((uintptr_t)p.wh_fw | (uintptr_t)p.wh_bw) & 15) == 0, "lstm_bwd_tc: bf16 kernels must be 16-byte aligned");
  p.X = X; p.L = L; p.keep_prob = keep_prob; p.seed = seed; p.seed_step = seed_step;
  p.dx2 = dx2; p.dxn = 0;
  p.wx_fw = p.wx_bw = nullptr; p.wx_rows = 0;
  CUtensorMap tmG, tmC, tmDh, tmDz, tmXf, tmXb;
  int rc = map_dir_tok(&tmG, gates_bf, 4, 2 * LT_U, X, L, 2 * LT_U, false);
  if (rc) return rc;
  rc = map_dir_tok(&tmC, cstate_bf, 2, LT_U, X, L, LT_U, false);
  if (rc) return rc;
  rc = map_tok(&tmDh, dhout_bf, 2 * LT_U, X, L, LT_U, LT_NH, false);
  if (rc) return rc;
  const int nhalf = (2 * ceil_div(X, LT_NH) <= 148) ? 1 : 2;
  const int ns = LT_NH * nhalf;
  rc = map_dir_tok(&tmDz, dz_bf, 2, 4 * LT_U, X, L, 64, true, ns);
  if (rc) return rc;
  CUtensorMap tmDx;
  tmXf = tmXb = tmDx = tmDh;
  if (dx2) {
    rc = map_dir_tok(&tmDx, dx2, 4, LB_DX_LD, X, L, LB_DX_LD, false, ns);
    if (rc) return rc;
    IN_REQUIRE(dx_first >= 0 && dx_first < D && D - dx_first <= LB_DX_LD,
               "lstm_bwd_tc: the fused input gradient covers at most %d features (got first=%d of D=%d)", LB_DX_LD, dx_first, D);
    p.dxn = D - dx_first;
    // rows dx_first.. of the TF kernel [D+u][4u] as the A operand: box 64 (k) x 128 rows (rows past the matrix: zero fill)
    const __nv_bfloat16* xf = (const __nv_bfloat16*)kernel_fw_bf + (long long)dx_first * 4 * LT_U;
    const __nv_bfloat16* xb = (const __nv_bfloat16*)kernel_bw_bf + (long long)dx_first * 4 * LT_U;
    p.wx_fw = xf; p.wx_bw = xb; p.wx_rows = D + LT_U - dx_first;
    rc = make_tmap_bf16(&tmXf, xf, 4 * LT_U, D + LT_U - dx_first, 4 * LT_U, 1, 0, 64, 128);
    if (rc) return rc;
    rc = make_tmap_bf16(&tmXb, xb, 4 * LT_U, D + LT_U - dx_first, 4 * LT_U, 1, 0, 64, 128);
    if (rc) return rc;
  }
  dim3 grid(ceil_div(X, ns), 2);
  const bool drop = keep_prob < 1.0f;
#define LB_LAUNCH2(DROP_, DX_, NH_)                                                                             \
  do {                                                                                                          \
    const int smem = DX_ ? LB_SMEM_DX : LB_SMEM_NODX;                                                           \
    IN_CUDA(cudaFuncSetAttribute(tc_lstm_bwd_kernel<DROP_, DX_, NH_>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); \
    tc_lstm_bwd_kernel<DROP_, DX_, NH_><<<grid, LT_THREADS, smem, st>>>(tmG, tmC, tmDh, tmDz, tmXf, tmXb, tmDx, p);  \
  } while (0)
#define LB_LAUNCH(DROP_, DX_) do { if (nhalf == 1) LB_LAUNCH2(DROP_, DX_, 1); else LB_LAUNCH2(DROP_, DX_, 2); } while (0)
  if (drop && dx2) LB_LAUNCH(true, true);
  else if (drop) LB_LAUNCH(true, false);
  else if (dx2) LB_LAUNCH(false, true);
  else LB_LAUNCH(false, false);
#undef LB_LAUNCH
#undef LB_LAUNCH2
  IN_LAUNCH_CHECK("tc_lstm_bwd_kernel");
  return INDNET_OK;
}

extern "C" int indnet_colsum_bf16(indnet_stream st_, const void* A, long long rows, int cols, float* out) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(rows >= 0 && cols >= 0, "colsum_bf16: bad sizes");
  if (cols == 0) return INDNET_OK;
  IN_REQUIRE(out != nullptr, "colsum_bf16: null output");
  IN_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * cols, st));
  if (rows == 0) return INDNET_OK;
  IN_REQUIRE(A != nullptr, "colsum_bf16: null input");
  int rpb = 256;
  long long gy = (rows + rpb - 1) / rpb;
  while (gy > 65535) { rpb *= 2; gy = (rows + rpb - 1) / rpb; }
  colsum_bf16_kernel<<<dim3(ceil_div(cols, 32), (unsigned)gy), 256, 0, st>>>((const __nv_bfloat16*)A, rows, cols, out, rpb);
  IN_LAUNCH_CHECK("colsum_bf16_kernel");
  return INDNET_OK;
}
