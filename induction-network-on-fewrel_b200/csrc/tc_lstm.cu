// a3, bf16 tensor-core mode: the BiLSTM of model/graph.py:34-49 as ONE persistent tcgen05 kernel with
// the input projection fused in.
//
// Swap-AB formulation.  Per step t and direction, for a tile of NS sentences:
//     Z^T[4u, NS] = [W_x^T | W_h^T][4u, 64+u] * [x_t ; h_{t-1}][64+u, NS]
// * MMA-M = gate units: four M=128 tiles, one per gate (i, j, f, o - the TF kernel column order), so
//   TMEM lane m of every tile is unit m and one thread reads all four gates of its unit from its own
//   lane; MMA-N = sentences (NS = 32), so a step's accumulators are 4 x NS TMEM columns.
// * The weights are loaded ONCE and stay resident for all L steps: W_x^T (64 KB) by TMA into
//   128-byte-swizzled shared memory, W_h^T - the operand of the serial, recurrent MMAs - into tensor
//   memory (A-from-TMEM MMAs run at the math rate, not at the shared-memory read rate).  x_t tiles
//   stream in by TMA (three buffers); h_{t-1} is written by the gate threads straight into the swizzled
//   B-operand layout.  Two accumulator buffers let the x part of step s+1 be issued while the gate
//   threads still read step s.
// * Gate math (forget bias 1.0, no length masking - SURVEY A.1/A.2) runs on the accumulators right
//   after tcgen05.ld; c stays in registers for the whole sequence.
// Warps 0-15: gate threads (TMEM lane quarter = warp % 4, sentence quarter = warp / 4); warp 16 issues the
// TMA loads and the MMAs through one elected lane.
#include <cstring>
#include "common.cuh"
#include "indnet_internal.h"
#include "tc_common.cuh"
#include "../../include/indnet.h"

namespace indnet {

using namespace tc;

constexpr int LT_U = 128;            // units per direction this kernel is specialised for
constexpr int LT_KX = 64;            // padded input width (D <= 64)
constexpr int LT_K = LT_KX + LT_U;   // 192
constexpr int LT_KB = LT_K / 64;     // 3 K-blocks of 64 (one 128-byte swizzle row each)
constexpr int LT_NS = 32;            // sentences per CTA (MMA N)
constexpr int LT_THREADS = 512;
constexpr int LT_SPT = LT_NS / (LT_THREADS / 128);   // sentences per gate thread
static_assert(LT_SPT == 8 || LT_SPT == 16, "gate threads use 8- or 16-column TMEM loads");
__device__ __forceinline__ void tmem_ld_spt(uint32_t taddr, float* v) {
  if constexpr (LT_SPT == 16) tmem_ld16(taddr, v); else tmem_ld8(taddr, v);
}
constexpr int LT_W_BYTES = 4 * 128 * 128;                    // W_x only (4 gate tiles x 64 k): 65536; W_h lives in TMEM
constexpr int LT_XB_BYTES = LT_NS * 128;                     // one x tile
constexpr int LT_HB_BYTES = 2 * LT_NS * 128;                 // h tile: 2 K-blocks
constexpr int LT_XBUFS = 3;                                 // x tiles in flight (see issue_x / issue_h)
constexpr int LT_SMEM = LT_W_BYTES + LT_XBUFS * LT_XB_BYTES + LT_HB_BYTES + 256 + 1024;

int make_tmap_bf16(CUtensorMap* tm, const void* base, long long inner, long long outer, long long ld,
                   long long batch, long long batch_stride, int box_inner, int box_outer);

// WT[dir][g*128 + m][k] (bf16, k-major, 192 wide): k < D -> W_x[k][g*128+m]; D <= k < 64 -> 0;
// k >= 64 -> W_h[k-64][g*128+m].  kernel = TF BasicLSTMCell kernel [D+u][4u] (W_x rows then W_h rows).
__global__ void lstm_pack_weights_kernel(const float* __restrict__ k_fw, const float* __restrict__ k_bw,
                                         __nv_bfloat16* __restrict__ WT, int D) {
  const int dir = blockIdx.y;
  const float* __restrict__ kern = dir ? k_bw : k_fw;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;      // over 4u * K with the unit index fastest
  if (idx >= 4 * LT_U * LT_K) return;
  const int col = idx % (4 * LT_U), k = idx / (4 * LT_U);
  float v = 0.f;
  if (k < D) v = kern[(long long)k * (4 * LT_U) + col];
  else if (k >= LT_KX) v = kern[(long long)(D + k - LT_KX) * (4 * LT_U) + col];
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
__device__ __forceinline__ float sigmoid_fast(float x) { return fmaf(tanh_fast(0.5f * x), 0.5f, 0.5f); }

struct LstmTcP {
  const float* bias_fw; const float* bias_bw;
  float* hout; __nv_bfloat16* hout_bf;
  __nv_bfloat16* gates_bf;      // [2][X][L][u][4]  (i, j, f, o of a unit adjacent: one 8-byte store / load)
  __nv_bfloat16* cstate;        // [2][X][L][u]
  __nv_bfloat16* hprev_bf;      // [2][X][L][u]
  const __nv_bfloat16* wt;      // packed weights WT[dir][4u][192] (lstm_pack_weights_kernel)
  int X, L;
  float keep_prob; unsigned long long seed; const long long* seed_step;
};

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ void sts_bf16(uint32_t saddr, __nv_bfloat16 v) {
  asm volatile("st.shared.b16 [%0], %1;" ::"r"(saddr), "h"(*reinterpret_cast<unsigned short*>(&v)) : "memory");
}
__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); }

// Warps 0-15 are gate warps (TMEM lane quarter = warp % 4, sentence quarter = warp / 4); warp 16 issues the
// TMA loads and the MMAs (one elected lane, operands in uniform registers).  DROP / HOUT32 are
// compile-time so the per-element code carries no uniform branches.
constexpr int LT_GATE_THREADS = LT_THREADS;
constexpr int LT_FWD_THREADS = LT_THREADS + 32;

template <bool DROP, bool HOUT32>
__global__ void __launch_bounds__(LT_FWD_THREADS, 1)
tc_lstm_fwd_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX, LstmTcP p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sW = smem;                                   // W_x^T: [4 gates][128 rows][128 B]
  uint8_t* sX = smem + LT_W_BYTES;                      // [3][NS rows][128 B]
  uint8_t* sH = sX + LT_XBUFS * LT_XB_BYTES;            // [2 kb][NS rows][128 B]
  uint64_t* w_full = reinterpret_cast<uint64_t*>(sH + LT_HB_BYTES);
  uint64_t* x_full = w_full + 1;                        // [3]
  uint64_t* mma_done = x_full + LT_XBUFS;
  uint64_t* h_ready = mma_done + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(h_ready + 1);

  const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;
  const bool gate_warp = warp < LT_GATE_THREADS / 32;
  const int dir = blockIdx.y;
  const int x0 = blockIdx.x * LT_NS;
  const int L = p.L;

  for (int i = threadIdx.x; i < LT_HB_BYTES / 16; i += LT_FWD_THREADS) reinterpret_cast<uint4*>(sH)[i] = make_uint4(0, 0, 0, 0);   // h_{-1} = 0
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmX);
    mbar_init(w_full, 1);
    for (int b = 0; b < LT_XBUFS; ++b) mbar_init(&x_full[b], 1);
    mbar_init(mma_done, 1);
    mbar_init(h_ready, LT_GATE_THREADS);
    fence_barrier_init();
  }
  __syncwarp();
  if (warp == 0) tmem_alloc(tmem_slot, 512);            // two accumulator buffers (2 x 4 gates x NS columns) + W_h^T (256)
  fence_proxy_async();          // the zero-fill of sH must be visible to the tensor core's operand reads
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t taddr = *tmem_slot;
  constexpr uint32_t idesc = idesc_bf16(128, LT_NS, false, false);
  const uint32_t w16 = smem_u32(sW) >> 4, x16 = smem_u32(sX) >> 4, h16 = smem_u32(sH) >> 4;
  const uint32_t wh_t = taddr + 8 * LT_NS;                 // W_h^T: gate g at columns g*64, lane = unit
  if (gate_warp) {
    // The recurrent weights are the A operand of every step's serial MMAs.  Read from shared memory
    // they cost 4 KB per 128x32x16 MMA and the tensor pipe runs at that read rate; resident in tensor
    // memory the MMA is bounded by its math.  Warp w stages gate w / 4, lane quarter w % 4: thread =
    // one row (unit) of W_h^T, 128 k values = 64 packed columns.
    const int g = warp >> 2, row = (warp & 3) * 32 + lane;
    const uint4* src = reinterpret_cast<const uint4*>(p.wt + ((long long)(dir * 4 + g) * LT_U + row) * LT_K + LT_KX);
    const uint32_t dst = wh_t + ((uint32_t)((warp & 3) * 32) << 16) + g * (LT_U / 2);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      uint32_t r[16];
#pragma unroll
      for (int v = 0; v < 4; ++v) {
        const uint4 q = src[c * 4 + v];
        r[v * 4 + 0] = q.x; r[v * 4 + 1] = q.y; r[v * 4 + 2] = q.z; r[v * 4 + 3] = q.w;
      }
      tmem_st16(dst + c * 16, r);
    }
    tmem_st_wait();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (!gate_warp) {
    // ---- issuer warp.  The x part of a step does not depend on the recurrence, so it is issued a step
    // early into the other accumulator buffer (buffer = s & 1) and only the h part sits on the serial
    // path between two gate phases.  x tiles: three buffers, tile of step s in buffer s % 3.
    auto load_x = [&](int s) {
      if (elect_one()) {
        const int tn = dir ? (L - 1 - s) : s;
        mbar_expect_tx(&x_full[s % LT_XBUFS], LT_XB_BYTES);
        tma_load_3d(sX + (s % LT_XBUFS) * LT_XB_BYTES, &tmX, &x_full[s % LT_XBUFS], 0, tn, x0);
      }
      __syncwarp();
    };
    auto issue_x = [&](int s) {
      mbar_wait(&x_full[s % LT_XBUFS], (s / LT_XBUFS) & 1);
      tc_fence_after();
      const uint32_t b16 = x16 + (s % LT_XBUFS) * (LT_XB_BYTES >> 4);
      const uint32_t acc = taddr + (s & 1) * (4 * LT_NS);
      if (elect_one()) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4) {
            const uint64_t ad = smem_desc_sw128_k(w16, g * 16384 + k4 * 32);
            const uint64_t bd = smem_desc_sw128_k(b16, k4 * 32);
            umma_bf16(acc + g * LT_NS, ad, bd, idesc, k4 > 0 ? 1u : 0u);
          }
        }
      }
      __syncwarp();
    };
    auto issue_h = [&](int s) {
      const uint32_t acc = taddr + (s & 1) * (4 * LT_NS);
      if (elect_one()) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
#pragma unroll
          for (int kb = 0; kb < 2; ++kb) {
            const uint32_t b16 = h16 + kb * ((LT_NS * 128) >> 4);
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4) {
              // A = W_h^T of gate g from tensor memory: 8 columns (16 k) per MMA
              const uint64_t bd = smem_desc_sw128_k(b16, k4 * 32);
              umma_bf16_ts(acc + g * LT_NS, wh_t + g * (LT_U / 2) + kb * 32 + k4 * 8, bd, idesc, 1u);
            }
          }
        }
        umma_commit(mma_done);
      }
      __syncwarp();
    };
    if (elect_one()) {
      mbar_expect_tx(w_full, LT_W_BYTES);                 // resident W_x^T: 4 boxes of 64 (k) x 128 (rows)
      for (int g = 0; g < 4; ++g) tma_load_3d(sW + g * 16384, &tmW, w_full, 0, g * 128, dir);
    }
    __syncwarp();
    for (int s = 0; s < LT_XBUFS && s < L; ++s) load_x(s);
    mbar_wait(w_full, 0);
    issue_x(0);
    issue_h(0);
    if (L > 1) issue_x(1);
    for (int s = 0; s + 1 < L; ++s) {
      mbar_wait(h_ready, s & 1);        // every gate thread has published h_s and drained accumulator buffer s & 1
      tc_fence_after();
      LT_MARK(lane == 0, s, 5);
      issue_h(s + 1);
      LT_MARK(lane == 0, s, 6);
      if (s + 2 < L) issue_x(s + 2);
      if (s + 3 < L) load_x(s + 3);     // into the buffer of x_s: its MMAs completed before h_s could exist
      LT_MARK(lane == 0, s, 7);
    }
  } else {
    // ---- gate warps: thread = (unit m = TMEM lane, sentence quarter hf), LT_SPT sentences each
    unsigned long long seed = p.seed;
    if (DROP && p.seed_step) seed += (unsigned long long)(*p.seed_step) * 1000003ull;
    const int m = (warp & 3) * 32 + lane;
    const int hf = warp >> 2;
    const float* __restrict__ bias = dir ? p.bias_bw : p.bias_fw;
    const float bi = bias[m], bj = bias[LT_U + m], bf = bias[2 * LT_U + m] + 1.0f, bo = bias[3 * LT_U + m];
    const uint32_t lane_addr0 = taddr + ((uint32_t)((warp & 3) * 32) << 16) + hf * LT_SPT;
    const float inv_keep = 1.0f / p.keep_prob;
    float c[LT_SPT], hp[LT_SPT];
#pragma unroll
    for (int q = 0; q < LT_SPT; ++q) { c[q] = 0.f; hp[q] = 0.f; }
    const int hkb = m >> 6, hk = m & 63;                     // destination inside the swizzled h tile
    const uint32_t hdst = smem_u32(sH) + hkb * (LT_NS * 128) + (hk & 7) * 2;
    const uint32_t hchunk = (uint32_t)(hk >> 3);
    const int xs0 = x0 + hf * LT_SPT;                        // first sentence of this thread
    uint32_t valid = 0;
#pragma unroll
    for (int q = 0; q < LT_SPT; ++q) valid |= (xs0 + q < p.X) ? (1u << q) : 0u;
    // Running byte pointers of sentence q = 0 at the current time step; sentence q adds q * (a
    // per-array stride), one wide multiply-add per access.
    const int t0 = dir ? (L - 1) : 0;
    const long long e0 = ((long long)(dir * p.X + xs0) * L + t0) * LT_U + m;            // [dir][X][L][u] element
    const long long o0 = ((long long)xs0 * L + t0) * (2 * LT_U) + dir * LT_U + m;       // [X][L][2u] element
    const long long estep = dir ? -(long long)LT_U : (long long)LT_U;
    const long long ostep = dir ? -(long long)(2 * LT_U) : (long long)(2 * LT_U);
    const uint32_t qe = (uint32_t)L * LT_U, qo = (uint32_t)L * 2 * LT_U;                // elements between sentences
    char* g_ptr = reinterpret_cast<char*>(p.gates_bf) + e0 * 8;
    char* c_ptr = reinterpret_cast<char*>(p.cstate) + e0 * 2;
    char* hp_ptr = reinterpret_cast<char*>(p.hprev_bf) + e0 * 2;
    char* ob_ptr = reinterpret_cast<char*>(p.hout_bf) + o0 * 2;
    char* of_ptr = reinterpret_cast<char*>(p.hout) + o0 * 4;
    long long oi = o0;

    for (int s = 0; s < L; ++s) {
      LT_MARK(threadIdx.x == 0, s, 0);
      mbar_wait(mma_done, s & 1);
      __syncwarp();
      tc_fence_after();
      LT_MARK(threadIdx.x == 0, s, 1);
      const uint32_t lane_addr = lane_addr0 + (s & 1) * (4 * LT_NS);
      // two halves of 4 sentences: halves the live accumulator registers of the gate math
#pragma unroll
      for (int half = 0; half < LT_SPT / 4; ++half) {
        float zi[4], zj[4], zf[4], zo[4];
        tmem_ld4(lane_addr + 0 * LT_NS + half * 4, zi);
        tmem_ld4(lane_addr + 1 * LT_NS + half * 4, zj);
        tmem_ld4(lane_addr + 2 * LT_NS + half * 4, zf);
        tmem_ld4(lane_addr + 3 * LT_NS + half * 4, zo);
        tmem_ld_wait();
#pragma unroll
        for (int qq = 0; qq < 4; ++qq) {
          const int q = half * 4 + qq;
          const int sl = hf * LT_SPT + q;
          const float gi = sigmoid_fast(zi[qq] + bi);
          const float gj = tanh_fast(zj[qq] + bj);
          const float gf = sigmoid_fast(zf[qq] + bf);
          const float go = sigmoid_fast(zo[qq] + bo);
          const float cn = fmaf(c[q], gf, gi * gj);
          const float h = tanh_fast(cn) * go;
          c[q] = cn;
          // next step's B operand: row = sentence, 16-byte chunk XOR (row & 7)
          sts_bf16(hdst + sl * 128 + ((hchunk ^ (uint32_t)(sl & 7)) << 4), __float2bfloat16(h));
          if (valid & (1u << q)) {
            const unsigned long long eq = (unsigned long long)qe * (uint32_t)q, oq = (unsigned long long)qo * (uint32_t)q;
            *reinterpret_cast<uint2*>(g_ptr + eq * 8) = make_uint2(pack_bf16x2(gi, gj), pack_bf16x2(gf, go));
            *reinterpret_cast<__nv_bfloat16*>(c_ptr + eq * 2) = __float2bfloat16(cn);
            *reinterpret_cast<__nv_bfloat16*>(hp_ptr + eq * 2) = __float2bfloat16(hp[q]);
            float ho = h;
            if (DROP) ho = dropout_keep(seed, (unsigned long long)(oi + (long long)oq), p.keep_prob) ? h * inv_keep : 0.f;
            if (HOUT32) *reinterpret_cast<float*>(of_ptr + oq * 4) = ho;
            *reinterpret_cast<__nv_bfloat16*>(ob_ptr + oq * 2) = __float2bfloat16(ho);
          }
          hp[q] = h;
        }
      }
      fence_proxy_async();      // h stores -> visible to the MMA's shared-memory reads
      tc_fence_before();        // TMEM loads ordered before the next step's MMAs overwrite the columns
      mbar_arrive(h_ready);
      LT_MARK(threadIdx.x == 0, s, 4);
      g_ptr += estep * 8; c_ptr += estep * 2; hp_ptr += estep * 2; ob_ptr += ostep * 2; of_ptr += ostep * 4; oi += ostep;
    }
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc(taddr, 512);
}

// ------------------------------------------------------------------------------------------------
// BPTT, bf16 tensor-core mode.  Per step (reverse processing order) and tile of NS sentences:
//     dh_rec^T[u, NS] = W_h[u, 4u] * dz^T[4u, NS]
// A = W_h exactly as the TF kernel stores it (rows D.. of [D+u][4u], K-major over the 4u gate columns),
// staged ONCE into tensor memory (lane = hidden unit, 256 packed columns), so the step's 32 MMAs run at
// the math rate.  B = dz of the step, written by the gate-gradient threads (bf16, swizzled shared
// memory).  dh_rec comes back through TMEM (NS columns); dc_rec stays in registers.  dz is also streamed
// to HBM in bf16 for the weight-gradient GEMMs, and its column sums (bias gradients) are formed here.
// Warps 0-15: gate-gradient threads; warp 16: MMA issuer.
constexpr int LB_KB = 4 * LT_U / 64;                           // 8 K-blocks
constexpr int LB_Z_BYTES = LB_KB * LT_NS * 128;                // 32768
constexpr int LB_WX_TKB = 6;                                   // dx path: K-blocks of the W_x rows held in tensor memory
constexpr int LB_WX_BYTES = (LB_KB - LB_WX_TKB) * 128 * 128;   // ... the remaining two stay in shared memory (32 KB)
constexpr int LB_SMEM = LB_Z_BYTES + 256 + 1024;
constexpr int LB_SMEM_DX = 2 * LB_Z_BYTES + LB_WX_BYTES + 256 + 1024;   // dz double-buffered: dx MMAs outlive the step
constexpr int LB_DX_LD = 16;                                   // dx2 row stride in floats (64-byte rows)
constexpr int LB_THREADS = LT_THREADS + 32;

struct LstmTcBwdP {
  const __nv_bfloat16* dhout; const __nv_bfloat16* gates_bf; const __nv_bfloat16* cstate; __nv_bfloat16* dz_bf;
  const __nv_bfloat16* wh_fw; const __nv_bfloat16* wh_bw;     // W_h = rows D.. of the bf16 TF kernels, [u][4u]
  float* dbias_fw; float* dbias_bw;     // optional: bias gradients (column sums of dz), accumulated with atomics
  float* dx2; int dxn;                  // optional: dx2[dir][tok][16], the first dxn (<= 16) input-gradient features
  const __nv_bfloat16* wx_fw; const __nv_bfloat16* wx_bw; int wx_rows;   // rows dx_first.. of the TF kernels, valid rows
  int X, L;
  float keep_prob; unsigned long long seed; const long long* seed_step;
};

// DX: also form the input gradient dx^T[f, NS] = W_x[f, 4u] * dz^T of the step (rows n0.. of the TF kernel as a
// second A operand with its own accumulator and barrier): the MMAs run behind the recurrent ones on an
// otherwise idle tensor pipe and replace a GEMM launch that re-read all of dz.  The operand fills what is left
// of tensor memory (192 columns = 6 of its 8 K-blocks); the last two K-blocks come from shared memory (32 KB),
// which keeps the shared-memory bandwidth for the gate threads.
template <bool DROP, bool DX>
__global__ void __launch_bounds__(LB_THREADS, 1)
tc_lstm_bwd_kernel(const __grid_constant__ CUtensorMap tmXf, const __grid_constant__ CUtensorMap tmXb, LstmTcBwdP p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  // dz tile [8 kb][NS rows][128 B]; with DX two of them (step s uses buffer s & 1): the dx MMAs of step s
  // still read dz(s) while the gate threads already write dz(s+1)
  uint8_t* sZ = smem;
  constexpr int ZBUFS = DX ? 2 : 1;
  uint8_t* sWx = smem + ZBUFS * LB_Z_BYTES;             // DX only: [8 kb][128 rows][128 B]
  uint64_t* mma_done = reinterpret_cast<uint64_t*>(sZ + ZBUFS * LB_Z_BYTES + (DX ? LB_WX_BYTES : 0));
  uint64_t* z_ready = mma_done + 1;
  uint64_t* dx_done = mma_done + 2;
  uint64_t* wx_full = mma_done + 3;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mma_done + 4);

  const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;
  const bool gate_warp = warp < LT_THREADS / 32;
  const int dir = blockIdx.y;
  const int x0 = blockIdx.x * LT_NS;
  const int L = p.L;

  if (threadIdx.x == 0) {
    mbar_init(mma_done, 1);
    mbar_init(z_ready, LT_THREADS);
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
  if (warp == 0) tmem_alloc(tmem_slot, 512);            // W_h: 256 columns, dh_rec accumulator: NS columns, dx accumulator: NS
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t taddr = *tmem_slot;
  const uint32_t wh_t = taddr, acc_t = taddr + 2 * LT_U, accx_t = acc_t + LT_NS, wx_t = accx_t + LT_NS;
  constexpr uint32_t idesc = idesc_bf16(128, LT_NS, false, false);
  const uint32_t z16 = smem_u32(sZ) >> 4;
  if (gate_warp) {
    // stage W_h: warp w writes lane quarter w % 4, k range [(w / 4) * 128, +128) = 64 packed columns
    const int kq = warp >> 2, row = (warp & 3) * 32 + lane;
    const uint4* src = reinterpret_cast<const uint4*>((dir ? p.wh_bw : p.wh_fw) + (long long)row * (4 * LT_U) + kq * LT_U);
    const uint32_t dst = wh_t + ((uint32_t)((warp & 3) * 32) << 16) + kq * (LT_U / 2);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      uint32_t r[16];
#pragma unroll
      for (int v = 0; v < 4; ++v) {
        const uint4 q = src[c * 4 + v];
        r[v * 4 + 0] = q.x; r[v * 4 + 1] = q.y; r[v * 4 + 2] = q.z; r[v * 4 + 3] = q.w;
      }
      tmem_st16(dst + c * 16, r);
    }
    if (DX) {
      // W_x rows, K-blocks 0..5: warp w writes lane quarter w % 4, k range [(w / 4) * 96, +96) = 48 packed columns
      const bool have = row < p.wx_rows;
      const uint4* xs = reinterpret_cast<const uint4*>((dir ? p.wx_bw : p.wx_fw) + (long long)(have ? row : 0) * (4 * LT_U) + kq * 96);
      const uint32_t xd = wx_t + ((uint32_t)((warp & 3) * 32) << 16) + kq * 48;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        uint32_t r[16];
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          const uint4 q = have ? xs[c * 4 + v] : make_uint4(0, 0, 0, 0);
          r[v * 4 + 0] = q.x; r[v * 4 + 1] = q.y; r[v * 4 + 2] = q.z; r[v * 4 + 3] = q.w;
        }
        tmem_st16(xd + c * 16, r);
      }
    }
    tmem_st_wait();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (!gate_warp) {
    const uint32_t wx16 = smem_u32(sWx) >> 4;
    if (DX) mbar_wait(wx_full, 0);
    for (int step = 0; step < (DX ? L : L - 1); ++step) {
      mbar_wait(z_ready, step & 1);
      tc_fence_after();
      LB_MARK(lane == 0, step, 5);
      const uint32_t zb16 = z16 + (DX ? (uint32_t)(step & 1) * (LB_Z_BYTES >> 4) : 0u);
      if (elect_one()) {
        if (step + 1 < L) {                             // the last step's dh_rec has no consumer
#pragma unroll
          for (int kb = 0; kb < LB_KB; ++kb) {
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4) {
              const uint64_t bd = smem_desc_sw128_k(zb16, kb * (LT_NS * 128) + k4 * 32);
              umma_bf16_ts(acc_t, wh_t + kb * 32 + k4 * 8, bd, idesc, (kb > 0 || k4 > 0) ? 1u : 0u);
            }
          }
          umma_commit(mma_done);
        }
      }
      __syncwarp();
      LB_MARK(lane == 0, step, 6);
      if (DX) {
        // Off the recurrence's critical path, own barrier - and issued only once the recurrent MMAs are done
        // and the gate warps have had time to pull dh_rec out of tensor memory: shared-memory-operand MMAs
        // keep the tensor pipe busy for ~1300 cycles and TMEM loads issued meanwhile wait behind them.
        if (step + 1 < L) { mbar_wait(mma_done, step & 1); __nanosleep(256); }
      }
      if (elect_one()) {
        if (DX) {
#pragma unroll
          for (int kb = 0; kb < LB_KB; ++kb) {
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4) {
              const uint64_t bd = smem_desc_sw128_k(zb16, kb * (LT_NS * 128) + k4 * 32);
              if (kb < LB_WX_TKB) {
                umma_bf16_ts(accx_t, wx_t + kb * 32 + k4 * 8, bd, idesc, (kb > 0 || k4 > 0) ? 1u : 0u);
              } else {
                const uint64_t ad = smem_desc_sw128_k(wx16, (kb - LB_WX_TKB) * 16384 + k4 * 32);
                umma_bf16(accx_t, ad, bd, idesc, 1u);
              }
            }
          }
          umma_commit(dx_done);
        }
      }
      __syncwarp();
    }
  } else {
    unsigned long long seed = p.seed;
    if (DROP && p.seed_step) seed += (unsigned long long)(*p.seed_step) * 1000003ull;
    const int m = (warp & 3) * 32 + lane;
    const int hf = warp >> 2;
    const uint32_t lane_addr = acc_t + ((uint32_t)((warp & 3) * 32) << 16) + hf * LT_SPT;
    const float inv_keep = 1.0f / p.keep_prob;
    // dz element (sentence row sl, k = g*128 + m) inside the swizzled tile: K-block g*2 + m/64
    const uint32_t zdst = smem_u32(sZ) + (m >> 6) * (LT_NS * 128) + (m & 7) * 2;
    const uint32_t zchunk = (uint32_t)((m & 63) >> 3);
    const int xs0 = x0 + hf * LT_SPT;
    // out-of-range sentences of the last tile read sentence X-1 (any valid address) and are zeroed via dh
    uint32_t valid = 0;
#pragma unroll
    for (int q = 0; q < LT_SPT; ++q) valid |= (xs0 + q < p.X) ? (1u << q) : 0u;
    const int xs_c = min(xs0, p.X - 1);                  // clamped first sentence (whole thread out of range)
    const uint32_t qmax = (uint32_t)max(0, min(LT_SPT - 1, p.X - 1 - xs_c));   // clamp for q offsets
    // running byte pointers of (clamped) sentence q = 0 at the current time step; sentence q adds q * stride
    const int t0 = dir ? 0 : (L - 1);                    // time index of the first step processed here
    const long long e0 = ((long long)(dir * p.X + xs_c) * L + t0) * LT_U + m;
    const long long o0 = ((long long)xs_c * L + t0) * (2 * LT_U) + dir * LT_U + m;
    const long long estep = dir ? (long long)LT_U : -(long long)LT_U;           // BPTT walks time backwards
    const long long ostep = dir ? (long long)(2 * LT_U) : -(long long)(2 * LT_U);
    const uint32_t qe = (uint32_t)L * LT_U, qo = (uint32_t)L * 2 * LT_U;
    const char* g_ptr = reinterpret_cast<const char*>(p.gates_bf) + e0 * 8;
    const char* c_ptr = reinterpret_cast<const char*>(p.cstate) + e0 * 2;
    const char* do_ptr = reinterpret_cast<const char*>(p.dhout) + o0 * 2;
    char* z_ptr = reinterpret_cast<char*>(p.dz_bf) + (e0 - m) * 8 + m * 2;       // [tok][4u] bf16, column m
    // dx2[dir][tok][16]: this thread's feature = its TMEM lane m (only lanes < dxn store), sentence q adds L rows
    const bool dx_lane = DX && m < p.dxn;
    char* x_ptr = DX ? reinterpret_cast<char*>(p.dx2) + ((e0 - m) / LT_U) * (LB_DX_LD * 4) + m * 4 : nullptr;
    const uint32_t lanex_addr = accx_t + ((uint32_t)((warp & 3) * 32) << 16) + hf * LT_SPT;
    auto store_dx = [&](int step_done) {                 // dx of processing step `step_done` (x_ptr still points at it)
      if ((warp & 3) * 32 < p.dxn) {                     // warp-uniform: this lane quarter holds needed features
        float dxv[LT_SPT];
        mbar_wait(dx_done, step_done & 1);
        __syncwarp();
        tc_fence_after();
        tmem_ld_spt(lanex_addr, dxv);
        tmem_ld_wait();
        if (dx_lane) {
#pragma unroll
          for (int q = 0; q < LT_SPT; ++q)
            if ((valid >> q) & 1u) *reinterpret_cast<float*>(x_ptr + (unsigned long long)((uint32_t)L * (uint32_t)q) * (LB_DX_LD * 4)) = dxv[q];
        }
      }
    };
    long long oi = o0;
    float sb_i = 0.f, sb_j = 0.f, sb_f = 0.f, sb_o = 0.f;     // this unit's bias-gradient partial sums
    float dcr[LT_SPT], cc[LT_SPT];
#pragma unroll
    for (int q = 0; q < LT_SPT; ++q) {
      dcr[q] = 0.f;
      const unsigned long long eq = (unsigned long long)qe * min((uint32_t)q, qmax);
      cc[q] = __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(c_ptr + eq * 2));
    }

    for (int step = 0; step < L; ++step) {
      LB_MARK(threadIdx.x == 0, step, 0);
      // ---- this step's saved activations: every load is issued before anything waits on them
      uint2 gw[LT_SPT];
      float dh[LT_SPT], cp[LT_SPT];
#pragma unroll
      for (int q = 0; q < LT_SPT; ++q) {
        const uint32_t qq = min((uint32_t)q, qmax);
        const unsigned long long eq = (unsigned long long)qe * qq, oq = (unsigned long long)qo * qq;
        gw[q] = *reinterpret_cast<const uint2*>(g_ptr + eq * 8);
        dh[q] = __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(do_ptr + oq * 2));
        cp[q] = (step + 1 < L) ? __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(c_ptr + estep * 2 + eq * 2)) : 0.f;
        if (DROP) dh[q] = dropout_keep(seed, (unsigned long long)(oi + (long long)oq), p.keep_prob) ? dh[q] * inv_keep : 0.f;
      }
      if (step > 0) {
        float dhr[LT_SPT];
        mbar_wait(mma_done, (step - 1) & 1);
        __syncwarp();
        tc_fence_after();
        LB_MARK(threadIdx.x == 0, step, 1);
        tmem_ld_spt(lane_addr, dhr);
        tmem_ld_wait();
        LB_MARK(threadIdx.x == 0, step, 2);
#pragma unroll
        for (int q = 0; q < LT_SPT; ++q) dh[q] += dhr[q];      // the recurrent part is added after the dropout mask
      }
#pragma unroll
      for (int q = 0; q < LT_SPT; ++q) {
        const int sl = hf * LT_SPT + q;
        const bool ok = (valid >> q) & 1u;
        const float dhq = ok ? dh[q] : 0.f;
        const float gi = bf16_lo(gw[q].x), gj = bf16_hi(gw[q].x), gf = bf16_lo(gw[q].y), go = bf16_hi(gw[q].y);
        const float tc_ = tanh_fast(cc[q]);
        const float dob = dhq * tc_ * go * (1.f - go);
        const float dc = dcr[q] + dhq * go * (1.f - tc_ * tc_);
        const float di = dc * gj * gi * (1.f - gi);
        const float dj = dc * gi * (1.f - gj * gj);
        const float df = dc * cp[q] * gf * (1.f - gf);
        dcr[q] = dc * gf;
        cc[q] = cp[q];                        // c_{t_prev} is the cell state of the next step processed
        sb_i += di; sb_j += dj; sb_f += df; sb_o += dob;
        const __nv_bfloat16 bi = __float2bfloat16(di), bj = __float2bfloat16(dj), bf = __float2bfloat16(df),
                            bo = __float2bfloat16(dob);
        const uint32_t zp = zdst + (DX ? (uint32_t)(step & 1) * LB_Z_BYTES : 0u) + sl * 128 + ((zchunk ^ (uint32_t)(sl & 7)) << 4);
        sts_bf16(zp, bi);
        sts_bf16(zp + 2 * (LT_NS * 128), bj);
        sts_bf16(zp + 4 * (LT_NS * 128), bf);
        sts_bf16(zp + 6 * (LT_NS * 128), bo);
        if (ok) {
          __nv_bfloat16* gz = reinterpret_cast<__nv_bfloat16*>(z_ptr + (unsigned long long)qe * (uint32_t)q * 8);
          gz[0] = bi; gz[LT_U] = bj; gz[2 * LT_U] = bf; gz[3 * LT_U] = bo;
        }
      }
      LB_MARK(threadIdx.x == 0, step, 3);
      if (DX && step > 0) store_dx(step - 1);           // long complete by now; must precede this step's arrival
      fence_proxy_async();
      tc_fence_before();
      mbar_arrive(z_ready);
      LB_MARK(threadIdx.x == 0, step, 4);
      g_ptr += estep * 8; c_ptr += estep * 2; do_ptr += ostep * 2; z_ptr += estep * 8; oi += ostep;
      if (DX && step > 0) x_ptr += (estep / LT_U) * (LB_DX_LD * 4);
    }
    if (DX) store_dx(L - 1);
    if (p.dbias_fw) {
      float* db = dir ? p.dbias_bw : p.dbias_fw;
      atomicAdd(db + m, sb_i); atomicAdd(db + LT_U + m, sb_j); atomicAdd(db + 2 * LT_U + m, sb_f); atomicAdd(db + 3 * LT_U + m, sb_o);
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
extern "C" int indnet_lstm_fwd_tc(indnet_stream st_, const void* x_bf, const float* kernel_fw, const float* bias_fw,
                                  const float* kernel_bw, const float* bias_bw, void* wt_scratch, float* hout,
                                  void* hout_bf, void* gates_bf, void* cstate_bf, void* hprev_bf, int X, int L, int D,
                                  int u, float keep_prob, unsigned long long seed, const long long* seed_step) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(X >= 0 && L > 0, "lstm_fwd_tc: bad sizes X=%d L=%d", X, L);
  IN_REQUIRE(u == LT_U && D > 0 && D <= LT_KX, "lstm_fwd_tc: specialised for hidden_size 128 and E+10 <= 64 (got u=%d D=%d)", u, D);
  IN_REQUIRE(keep_prob > 0.f && keep_prob <= 1.f, "lstm_fwd_tc: keep_prob must be in (0, 1], got %f", keep_prob);
  if (X == 0) return INDNET_OK;
  IN_REQUIRE(x_bf && kernel_fw && bias_fw && kernel_bw && bias_bw && wt_scratch && cstate_bf && hout_bf && gates_bf &&
             hprev_bf, "lstm_fwd_tc: null pointer");
  IN_REQUIRE((long long)2 * X * L < 2147483647LL / 1, "lstm_fwd_tc: token count exceeds 32-bit indexing");
  lstm_pack_weights_kernel<<<dim3(ceil_div(4 * LT_U * LT_K, 256), 2), 256, 0, st>>>(kernel_fw, kernel_bw,
                                                                                    (__nv_bfloat16*)wt_scratch, D);
  IN_LAUNCH_CHECK("lstm_pack_weights_kernel");
  CUtensorMap tmW, tmX;
  // WT as [dir][4u rows][192]: box 64 (k) x 128 (rows)
  int rc = make_tmap_bf16(&tmW, wt_scratch, LT_K, 4 * LT_U, LT_K, 2, (long long)4 * LT_U * LT_K, 64, 128);
  if (rc) return rc;
  // x as [X][L][64]: box 64 x 1 (step) x NS (sentences) -> NS rows of 128 B
  {
    typedef CUresult (*Fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                           const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                           CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fp = nullptr;
    cudaDriverEntryPointQueryResult q;
    IN_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
    IN_REQUIRE(fp && q == cudaDriverEntryPointSuccess, "lstm_fwd_tc: cuTensorMapEncodeTiled unavailable");
    cuuint64_t dims[3] = {(cuuint64_t)LT_KX, (cuuint64_t)L, (cuuint64_t)X};
    cuuint64_t strides[2] = {(cuuint64_t)LT_KX * 2, (cuuint64_t)L * LT_KX * 2};
    cuuint32_t box[3] = {(cuuint32_t)LT_KX, 1, (cuuint32_t)LT_NS};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = ((Fn)fp)(&tmX, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(x_bf), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    IN_REQUIRE(r == CUDA_SUCCESS, "lstm_fwd_tc: cuTensorMapEncodeTiled(x) failed with %d", (int)r);
  }
  LstmTcP p;
  p.bias_fw = bias_fw; p.bias_bw = bias_bw;
  p.hout = hout; p.hout_bf = (__nv_bfloat16*)hout_bf;
  p.gates_bf = (__nv_bfloat16*)gates_bf;
  p.cstate = (__nv_bfloat16*)cstate_bf; p.hprev_bf = (__nv_bfloat16*)hprev_bf;
  p.wt = (const __nv_bfloat16*)wt_scratch;
  p.X = X; p.L = L; p.keep_prob = keep_prob; p.seed = seed; p.seed_step = seed_step;
  dim3 grid(ceil_div(X, LT_NS), 2);
  const bool drop = keep_prob < 1.0f, h32 = hout != nullptr;
#define LT_FWD_LAUNCH(DROP_, H32_)                                                                              \
  do {                                                                                                          \
    /* per device and cheap: set on every launch (legal during stream capture) */                                 \
    IN_CUDA(cudaFuncSetAttribute(tc_lstm_fwd_kernel<DROP_, H32_>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                   LT_SMEM));                                                                   \
    tc_lstm_fwd_kernel<DROP_, H32_><<<grid, LT_FWD_THREADS, LT_SMEM, st>>>(tmW, tmX, p);                          \
  } while (0)
  if (drop && h32) LT_FWD_LAUNCH(true, true);
  else if (drop) LT_FWD_LAUNCH(true, false);
  else if (h32) LT_FWD_LAUNCH(false, true);
  else LT_FWD_LAUNCH(false, false);
#undef LT_FWD_LAUNCH
  IN_LAUNCH_CHECK("tc_lstm_fwd_kernel");
  return INDNET_OK;
}

extern "C" int indnet_lstm_bwd_tc(indnet_stream st_, const void* dhout_bf, const void* gates_bf, const void* cstate_bf,
                                  const void* kernel_fw_bf, const void* kernel_bw_bf, void* dz_bf, float* dbias_fw,
                                  float* dbias_bw, float* dx2, int dx_first, int X, int L, int D, int u,
                                  float keep_prob, unsigned long long seed, const long long* seed_step) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(X >= 0 && L > 0, "lstm_bwd_tc: bad sizes X=%d L=%d", X, L);
  IN_REQUIRE(u == LT_U && D > 0 && D <= LT_KX, "lstm_bwd_tc: specialised for hidden_size 128 and E+10 <= 64 (got u=%d D=%d)", u, D);
  IN_REQUIRE(keep_prob > 0.f && keep_prob <= 1.f, "lstm_bwd_tc: keep_prob must be in (0, 1], got %f", keep_prob);
  if (X == 0) return INDNET_OK;
  IN_REQUIRE(dhout_bf && gates_bf && cstate_bf && kernel_fw_bf && kernel_bw_bf && dz_bf, "lstm_bwd_tc: null pointer");
  LstmTcBwdP p;
  // W_h = rows D.. of the bf16 TF kernel [D+u][4u]; D*4u*2 bytes is a multiple of 16 for every D
  p.wh_fw = (const __nv_bfloat16*)kernel_fw_bf + (long long)D * 4 * LT_U;
  p.wh_bw = (const __nv_bfloat16*)kernel_bw_bf + (long long)D * 4 * LT_U;
  IN_REQUIRE((((uintptr_t)p.wh_fw | (uintptr_t)p.wh_bw) & 15) == 0, "lstm_bwd_tc: bf16 kernels must be 16-byte aligned");
  p.dhout = (const __nv_bfloat16*)dhout_bf; p.gates_bf = (const __nv_bfloat16*)gates_bf;
  p.cstate = (const __nv_bfloat16*)cstate_bf; p.dz_bf = (__nv_bfloat16*)dz_bf;
  p.X = X; p.L = L; p.keep_prob = keep_prob; p.seed = seed; p.seed_step = seed_step;
  IN_REQUIRE((dbias_fw == nullptr) == (dbias_bw == nullptr), "lstm_bwd_tc: pass both bias gradients or neither");
  p.dbias_fw = dbias_fw; p.dbias_bw = dbias_bw;
  if (dbias_fw) {
    IN_CUDA(cudaMemsetAsync(dbias_fw, 0, sizeof(float) * 4 * LT_U, st));
    IN_CUDA(cudaMemsetAsync(dbias_bw, 0, sizeof(float) * 4 * LT_U, st));
  }
  p.dx2 = dx2; p.dxn = 0;
  p.wx_fw = p.wx_bw = nullptr; p.wx_rows = 0;
  CUtensorMap tmXf, tmXb;
  memset(&tmXf, 0, sizeof(tmXf));
  memset(&tmXb, 0, sizeof(tmXb));
  if (dx2) {
    IN_REQUIRE(dx_first >= 0 && dx_first < D && D - dx_first <= LB_DX_LD,
               "lstm_bwd_tc: the fused input gradient covers at most %d features (got first=%d of D=%d)", LB_DX_LD, dx_first, D);
    p.dxn = D - dx_first;
    // rows dx_first.. of the TF kernel [D+u][4u] as the A operand: box 64 (k) x 128 rows (rows past the matrix: zero fill)
    const __nv_bfloat16* xf = (const __nv_bfloat16*)kernel_fw_bf + (long long)dx_first * 4 * LT_U;
    const __nv_bfloat16* xb = (const __nv_bfloat16*)kernel_bw_bf + (long long)dx_first * 4 * LT_U;
    p.wx_fw = xf; p.wx_bw = xb; p.wx_rows = D + LT_U - dx_first;
    int rc = make_tmap_bf16(&tmXf, xf, 4 * LT_U, D + LT_U - dx_first, 4 * LT_U, 1, 0, 64, 128);
    if (rc) return rc;
    rc = make_tmap_bf16(&tmXb, xb, 4 * LT_U, D + LT_U - dx_first, 4 * LT_U, 1, 0, 64, 128);
    if (rc) return rc;
  }
  dim3 grid(ceil_div(X, LT_NS), 2);
  const bool drop = keep_prob < 1.0f;
#define LB_LAUNCH(DROP_, DX_)                                                                                   \
  do {                                                                                                          \
    const int smem = DX_ ? LB_SMEM_DX : LB_SMEM;                                                                \
    /* per device and cheap: set on every launch (legal during stream capture) */                                 \
    IN_CUDA(cudaFuncSetAttribute(tc_lstm_bwd_kernel<DROP_, DX_>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); \
    tc_lstm_bwd_kernel<DROP_, DX_><<<grid, LB_THREADS, smem, st>>>(tmXf, tmXb, p);                                \
  } while (0)
  if (drop && dx2) LB_LAUNCH(true, true);
  else if (drop) LB_LAUNCH(true, false);
  else if (dx2) LB_LAUNCH(false, true);
  else LB_LAUNCH(false, false);
#undef LB_LAUNCH
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
