// a3 at hidden_size 256 (BASELINE configs[4]: the encoder+induction inference sweep, model/base.py:16 takes the
// width as a free parameter): the BiLSTM forward as a persistent tcgen05 kernel on a CLUSTER of four CTAs.
//
// W_h^T of one direction is [4u, u] = 1024 x 256 bf16 = 512 KB - twice one SM's tensor memory - so the gate rows
// are split over four SMs: CTA r of the cluster owns units [64r, 64r + 64) with all four gates, as two M = 128
// tiles with rows {i (64 units); j (64 units)} and {f (64 units); o (64 units)}.  Per CTA that is 256 rows x
// (64 + 256) k = 160 KB, resident in tensor memory for all L steps (W_h 256 columns + W_x 64 columns, bias and
// the 0.5 sigmoid pre-scale folded in exactly as in tc_lstm.cu); the accumulators take 2 tiles x 2 halves x NH.
//   Z^T[256 rows, NH] = [W_x^T | W_h^T][256, 64 + 256] * [x_t ; h_{t-1}][64 + 256, NH]      per half, per step
// Every CTA needs the whole h_{t-1} (256 units) as its B operand and produces 64 units of h_t: after the gate
// math a CTA writes its 64 units for the NH sentences as ONE swizzled K-block tile in its own shared memory and
// a sender warp pushes that tile into K-block r of the h tile of all four CTAs with cp.async.bulk over
// distributed shared memory, completing bytes on the destination's mbarrier - no cluster-wide barrier inside
// the loop; the recurrence itself orders the buffers (a CTA cannot produce h_{t+1} before it has consumed every
// CTA's h_t).  The same tile leaves for HBM by a TMA store.
// A unit's i / f rows sit in TMEM lanes l and its j / o rows in lanes 64 + l, which different warps own: the
// j / o warps apply their activations and hand (tanh j, sigmoid o) to the i / f warps through shared memory
// (one named barrier per warp pair), which then update c (fp32 registers) and emit h.
// Two halves of NH sentences are phase shifted as in tc_lstm.cu.  Forward / inference only (no BPTT streams).
// Warps 0-15: gate warps; warp 16: TMA loads + MMA issue; warp 17: h exchange + stores.
#include <cstring>
#include "common.cuh"
#include "indnet_internal.h"
#include "tc_common.cuh"
#include "../../include/indnet.h"

namespace indnet {

using namespace tc;

constexpr int L2_U = 256, L2_KX = 64, L2_K = L2_KX + L2_U;     // 320
constexpr int L2_CL = 4;                 // CTAs per cluster
constexpr int L2_UC = L2_U / L2_CL;      // 64 units per CTA
constexpr int L2_GATE_THREADS = 512, L2_THREADS = L2_GATE_THREADS + 64;
constexpr int L2_XBUFS = 4;
constexpr uint32_t L2_TM_WH = 0, L2_TM_WX = 256, L2_TM_ACC = 320;

int make_tmap_nd(CUtensorMap* tm, const void* base, int elem_bytes, int rank, const unsigned long long* dims,
                 const unsigned long long* strides_bytes, const unsigned* box, bool swizzle128);

// WT[dir][r][tile][row][k] (bf16, k-major, 320 wide): tile 0 rows = {i, j}, tile 1 rows = {f, o} of the 64 units
// of cluster rank r; k < D: W_x, k = D / D+1: bias head / remainder (+1.0 forget bias), k >= 64: W_h[k - 64];
// i, f, o rows scaled by 0.5.  kernel = TF BasicLSTMCell kernel [D+u][4u], bias [4u].
__global__ void lstm256_pack_weights_kernel(const float* __restrict__ k_fw, const float* __restrict__ k_bw,
                                            const float* __restrict__ b_fw, const float* __restrict__ b_bw,
                                            __nv_bfloat16* __restrict__ WT, int D) {
  const int dir = blockIdx.y;
  const float* __restrict__ kern = dir ? k_bw : k_fw;
  const float* __restrict__ bias = dir ? b_bw : b_fw;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;      // over (4u rows) x K with the row fastest
  if (idx >= 4 * L2_U * L2_K) return;
  const int prow = idx % (4 * L2_U), k = idx / (4 * L2_U);     // packed row = ((r * 2 + tile) * 128 + row)
  const int r = prow / 256, tile = (prow / 128) & 1, row = prow % 128;
  const int g = tile * 2 + (row >= 64 ? 1 : 0);               // tile 0: i, j; tile 1: f, o
  const int unit = r * L2_UC + (row & 63);
  const int col = g * L2_U + unit;                            // column of the TF kernel
  const float sc = (g == 1) ? 1.0f : 0.5f;
  float v = 0.f;
  if (k < D) v = sc * kern[(long long)k * (4 * L2_U) + col];
  else if (k >= L2_KX) v = sc * kern[(long long)(D + k - L2_KX) * (4 * L2_U) + col];
  else if (k == D || k == D + 1) {
    const float b = sc * (bias[col] + (g == 2 ? 1.0f : 0.0f));
    const float head = __bfloat162float(__float2bfloat16(b));
    v = (k == D) ? head : b - head;
  }
  WT[((long long)dir * 4 * L2_U + prow) * L2_K + k] = __float2bfloat16(v);
}

__device__ __forceinline__ float tanh_fast2(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t saddr, uint32_t rank) {      // shared::cta -> shared::cluster of `rank`
  uint32_t d;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(d) : "r"(saddr), "r"(rank));
  return d;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared memory of this CTA -> shared memory of a CTA of the cluster; completes `bytes` on that CTA's mbarrier
__device__ __forceinline__ void dsmem_copy(uint32_t dst_cluster, uint32_t src_cta, uint32_t bytes, uint32_t bar_cluster) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst_cluster), "r"(src_cta), "r"(bytes), "r"(bar_cluster) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void sts_bf16_256(uint32_t saddr, __nv_bfloat16 v) {
  asm volatile("st.shared.b16 [%0], %1;" ::"r"(saddr), "h"(*reinterpret_cast<unsigned short*>(&v)) : "memory");
}

struct Lstm256P {
  int X, L;
  const __nv_bfloat16* wt;      // packed weights WT[dir][r][tile][128][320]
};

template <int N> __device__ __forceinline__ void tmem_ld_n(uint32_t taddr, float* v) {
  if constexpr (N == 8) tmem_ld8(taddr, v); else tmem_ld4(taddr, v);
}

// NH = sentences per half (MMA N): 32 for throughput, 16 when the batch is small
template <int NH>
struct L2Smem {
  static constexpr int X_BYTES = 2 * NH * 128;                 // x_t of both halves
  static constexpr int H_BYTES = 4 * NH * 128;                 // h of a half: 4 K-blocks (one per CTA of the cluster)
  static constexpr int O_BYTES = NH * 128;                     // this CTA's K-block of h_t
  static constexpr int E_BYTES = L2_UC * NH * 8;               // (tanh j, sigmoid o) hand-over, [sentence][unit] float2
  static constexpr int OFF_X = 0;
  static constexpr int OFF_H = OFF_X + L2_XBUFS * X_BYTES;     // [half][buf]
  static constexpr int OFF_O = OFF_H + 4 * H_BYTES;            // [half][buf]
  static constexpr int OFF_E = OFF_O + 4 * O_BYTES;            // [half]
  static constexpr int OFF_BAR = OFF_E + 2 * E_BYTES;
  static constexpr int BYTES = OFF_BAR + 256 + 1024;
};

template <int NH>
__global__ void __cluster_dims__(L2_CL, 1, 1) __launch_bounds__(L2_THREADS, 1)
tc_lstm256_fwd_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmHout, Lstm256P p) {
  using S = L2Smem<NH>;
  constexpr int SPT = NH / 4;            // sentences per gate thread and half
  extern __shared__ uint8_t smem_raw[];
  // every CTA of the cluster must use the same offsets: the dynamic segment starts at the same shared address in each
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* x_full = reinterpret_cast<uint64_t*>(smem + S::OFF_BAR);     // [4]
  uint64_t* mma_done = x_full + L2_XBUFS;                                  // [half]
  uint64_t* out_ready = mma_done + 2;                                      // [half]: the 256 i/f threads wrote this CTA's h K-block
  uint64_t* h_full = out_ready + 2;                                        // [half][step parity]: all four K-blocks of h landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(h_full + 4);
  const uint32_t sX_s = smem_u32(smem + S::OFF_X), sH_s = smem_u32(smem + S::OFF_H), sO_s = smem_u32(smem + S::OFF_O);
  const uint32_t sE_s = smem_u32(smem + S::OFF_E);

  const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;
  const int dir = blockIdx.y;
  const uint32_t rank = cluster_rank();
  const int x0 = (blockIdx.x / L2_CL) * (2 * NH);
  const int L = p.L;
  constexpr uint32_t H_STEP_BYTES = L2_CL * S::O_BYTES;          // bytes that complete one h_full phase

  for (int i = threadIdx.x; i < 4 * S::H_BYTES / 16; i += L2_THREADS)
    reinterpret_cast<uint4*>(smem + S::OFF_H)[i] = make_uint4(0, 0, 0, 0);     // h_{-1} = 0
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmHout);
    for (int b = 0; b < L2_XBUFS; ++b) mbar_init(&x_full[b], 1);
    for (int h = 0; h < 2; ++h) {
      mbar_init(&mma_done[h], 1);
      mbar_init(&out_ready[h], L2_GATE_THREADS / 2);
      mbar_init(&h_full[h * 2 + 0], 1);
      mbar_init(&h_full[h * 2 + 1], 1);
    }
    fence_barrier_init();
    // arm the phases of h_0 and h_1 before any CTA of the cluster can send (cluster barrier below)
    for (int h = 0; h < 2; ++h) {
      if (L >= 2) mbar_expect_tx(&h_full[h * 2 + 0], H_STEP_BYTES);
      if (L >= 3) mbar_expect_tx(&h_full[h * 2 + 1], H_STEP_BYTES);
    }
  }
  __syncwarp();
  if (warp == 0) tmem_alloc(tmem_slot, 512);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t taddr = *tmem_slot;
  if (warp < L2_GATE_THREADS / 32) {
    // stage this CTA's 256 weight rows: thread = row (lane quarter warp % 4); the 20 chunks of 32 k values of the
    // two tiles are split over the four warps of a quarter
    const int part = warp >> 2, row = (warp & 3) * 32 + lane;
    const uint32_t lane_sel = (uint32_t)((warp & 3) * 32) << 16;
#pragma unroll
    for (int cc = 0; cc < 5; ++cc) {
      const int c = part * 5 + cc, tile = c / 10, ck = c % 10;
      const uint4* src = reinterpret_cast<const uint4*>(
          p.wt + (((long long)dir * 4 + rank) * 256 + tile * 128 + row) * L2_K + ck * 32);
      uint32_t r[16];
#pragma unroll
      for (int v = 0; v < 4; ++v) {
        const uint4 q = src[v];
        r[v * 4 + 0] = q.x; r[v * 4 + 1] = q.y; r[v * 4 + 2] = q.z; r[v * 4 + 3] = q.w;
      }
      const uint32_t dst = (ck < 2) ? (taddr + L2_TM_WX + tile * 32 + ck * 16) : (taddr + L2_TM_WH + tile * 128 + (ck - 2) * 16);
      tmem_st16(dst + lane_sel, r);
    }
    tmem_st_wait();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  cluster_sync_all();            // every CTA's barriers are initialised and armed before anyone sends

  auto tn = [&](int s) { return dir ? (L - 1 - s) : s; };
  if (warp == 16) {
    // ---------------------------------------------------------------- TMA loads + MMA issue
    constexpr uint32_t idesc = idesc_bf16(128, NH, false, false);
    auto load_x = [&](int s) {
      mbar_expect_tx(&x_full[s % L2_XBUFS], S::X_BYTES);
      tma_load_3d(smem + S::OFF_X + (s % L2_XBUFS) * S::X_BYTES, &tmX, &x_full[s % L2_XBUFS], 0, tn(s), x0);
    };
    if (elect_one()) for (int s = 0; s < L2_XBUFS && s < L; ++s) load_x(s);
    __syncwarp();
    for (int s = 0; s < L; ++s) {
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        if (s > 0) {
          mbar_wait_wd(&h_full[h * 2 + ((s - 1) & 1)], ((s - 1) >> 1) & 1);      // h_{s-1}: all four K-blocks are here
          // its phase is consumed: arm the same barrier for h_{s+1} (nobody can send that before it has our h_s)
          if (s + 1 <= L - 2 && elect_one()) mbar_expect_tx(&h_full[h * 2 + ((s + 1) & 1)], H_STEP_BYTES);
          __syncwarp();
        }
        if (h == 0) mbar_wait_wd(&x_full[s % L2_XBUFS], (s / L2_XBUFS) & 1);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t xb16 = (sX_s + (s % L2_XBUFS) * S::X_BYTES + h * (NH * 128)) >> 4;
          const uint32_t hb16 = (sH_s + (h * 2 + ((s + 1) & 1)) * S::H_BYTES) >> 4;          // h_{s-1}: buffer (s-1) & 1
#pragma unroll
          for (int t = 0; t < 2; ++t) {
            const uint32_t acc = taddr + L2_TM_ACC + t * (2 * NH) + h * NH;
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4)
              umma_bf16_ts(acc, taddr + L2_TM_WX + t * 32 + k4 * 8, smem_desc_sw128_k(xb16, k4 * 32), idesc, k4 > 0 ? 1u : 0u);
#pragma unroll
            for (int kk = 0; kk < 16; ++kk)
              umma_bf16_ts(acc, taddr + L2_TM_WH + t * 128 + kk * 8,
                           smem_desc_sw128_k(hb16, (kk >> 2) * (NH * 128) + (kk & 3) * 32), idesc, 1u);
          }
          umma_commit(&mma_done[h]);
          // x_{s-1} has been consumed by both halves (their MMAs completed before h_{s-1} could exist)
          if (h == 1 && s >= 1 && s - 1 + L2_XBUFS < L) load_x(s - 1 + L2_XBUFS);
        }
        __syncwarp();
      }
    }
  } else if (warp == 17) {
    // ---------------------------------------------------------------- h exchange + output stores
    for (int s = 0; s < L; ++s) {
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        mbar_wait_wd(&out_ready[h], s & 1);
        if (elect_one()) {
          bulk_wait_read<1>();             // the store that read this half's other buffer two half steps ago
          const uint32_t src = sO_s + (h * 2 + (s & 1)) * S::O_BYTES;
          if (s + 1 < L) {                 // the last h has no consumer
            const uint32_t dst = sH_s + (h * 2 + (s & 1)) * S::H_BYTES + rank * S::O_BYTES;
            const uint32_t bar = smem_u32(&h_full[h * 2 + (s & 1)]);
#pragma unroll
            for (uint32_t q = 0; q < L2_CL; ++q) dsmem_copy(map_to_cta(dst, q), src, S::O_BYTES, map_to_cta(bar, q));
          }
          tma_store_3d(&tmHout, src, dir * L2_U + rank * L2_UC, tn(s), x0 + h * NH);
          bulk_commit();
        }
        __syncwarp();
      }
    }
    if (elect_one()) bulk_wait_all<0>();
    __syncwarp();
  } else {
    // ---------------------------------------------------------------- gate warps
    const int lq = warp & 3, sg = warp >> 2;
    const bool owner = lq < 2;                       // lanes 0..63: i and f rows; lanes 64..127: j and o rows
    const int ul = (lq & 1) * 32 + lane;             // unit within this CTA's 64
    const uint32_t acc0 = taddr + L2_TM_ACC + ((uint32_t)(lq * 32) << 16) + sg * SPT;
    const int pair_bar = 1 + (lq & 1) * 4 + sg;      // named barrier shared by the i/f warp and the j/o warp of a unit group
    const uint32_t eoff = (uint32_t)(sg * SPT) * (L2_UC * 8) + (uint32_t)ul * 8;        // + q * 512
    const uint32_t ochunk = (uint32_t)ul >> 3, obyte = (uint32_t)(ul & 7) * 2;
    float c[2 * SPT];
#pragma unroll
    for (int q = 0; q < 2 * SPT; ++q) c[q] = 0.f;
    for (int s = 0; s < L; ++s) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        mbar_wait_wd(&mma_done[h], s & 1);
        tc_fence_after();
        float z0[SPT], z1[SPT];
        tmem_ld_n<SPT>(acc0 + 0 * (2 * NH) + h * NH, z0);      // owner: i, helper: j
        tmem_ld_n<SPT>(acc0 + 1 * (2 * NH) + h * NH, z1);      // owner: f, helper: o
        tmem_ld_wait();
        tc_fence_before();
        const uint32_t ebase = sE_s + h * S::E_BYTES + eoff;
        if (!owner) {
#pragma unroll
          for (int q = 0; q < SPT; ++q) {
            const float gj = tanh_fast2(z0[q]);
            const float go = fmaf(tanh_fast2(z1[q]), 0.5f, 0.5f);
            asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(ebase + q * (L2_UC * 8)), "f"(gj), "f"(go) : "memory");
          }
          named_bar_sync(pair_bar, 64);
        } else {
          float gi[SPT], gf[SPT];
#pragma unroll
          for (int q = 0; q < SPT; ++q) {
            gi[q] = fmaf(tanh_fast2(z0[q]), 0.5f, 0.5f);
            gf[q] = fmaf(tanh_fast2(z1[q]), 0.5f, 0.5f);
          }
          named_bar_sync(pair_bar, 64);
          const uint32_t odst = sO_s + (h * 2 + (s & 1)) * S::O_BYTES + obyte;
#pragma unroll
          for (int q = 0; q < SPT; ++q) {
            float gj, go;
            asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(gj), "=f"(go) : "r"(ebase + q * (L2_UC * 8)) : "memory");
            const float cn = fmaf(c[h * SPT + q], gf[q], gi[q] * gj);
            const float hv = tanh_fast2(cn) * go;
            c[h * SPT + q] = cn;
            const uint32_t row = (uint32_t)(sg * SPT + q);
            sts_bf16_256(odst + row * 128 + ((ochunk ^ (row & 7)) << 4), __float2bfloat16(hv));
          }
          fence_proxy_async();      // this CTA's h K-block -> visible to the bulk copies and the TMA store
          mbar_arrive(&out_ready[h]);
        }
      }
    }
  }
  // nobody leaves while a peer may still write into its shared memory or wait on its barriers
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 0) tmem_dealloc(taddr, 512);
}

}  // namespace indnet

using namespace indnet;

extern "C" int indnet_lstm256_fwd_tc(indnet_stream st_, const void* x_bf, const float* kernel_fw, const float* bias_fw,
                                     const float* kernel_bw, const float* bias_bw, void* wt_scratch, void* hout_bf,
                                     int X, int L, int D, int u) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(X >= 0 && L > 0, "lstm256_fwd_tc: bad sizes X=%d L=%d", X, L);
  IN_REQUIRE(u == L2_U && D > 0 && D + 2 <= L2_KX, "lstm256_fwd_tc: specialised for hidden_size 256 and E+10 <= 62 (got u=%d D=%d)", u, D);
  if (X == 0) return INDNET_OK;
  IN_REQUIRE(x_bf && kernel_fw && bias_fw && kernel_bw && bias_bw && wt_scratch && hout_bf, "lstm256_fwd_tc: null pointer");
  lstm256_pack_weights_kernel<<<dim3(ceil_div(4 * L2_U * L2_K, 256), 2), 256, 0, st>>>(kernel_fw, kernel_bw, bias_fw, bias_bw,
                                                                                       (__nv_bfloat16*)wt_scratch, D);
  IN_LAUNCH_CHECK("lstm256_pack_weights_kernel");
  // 64-sentence tiles (two halves of 32) for throughput; 32-sentence tiles when that still gives every cluster work
  const bool big = ceil_div(X, 64) * 2 >= 36;
  const int nh = big ? 32 : 16;
  typedef unsigned long long u64;
  CUtensorMap tmX, tmHout;
  {
    const u64 dims[3] = {(u64)L2_KX, (u64)L, (u64)X};
    const u64 strides[2] = {(u64)L2_KX * 2, (u64)L * L2_KX * 2};
    const unsigned box[3] = {(unsigned)L2_KX, 1, (unsigned)(2 * nh)};
    int rc = make_tmap_nd(&tmX, x_bf, 2, 3, dims, strides, box, true);
    if (rc) return rc;
  }
  {
    const u64 dims[3] = {(u64)(2 * L2_U), (u64)L, (u64)X};
    const u64 strides[2] = {(u64)(2 * L2_U) * 2, (u64)L * (2 * L2_U) * 2};
    const unsigned box[3] = {64, 1, (unsigned)nh};
    int rc = make_tmap_nd(&tmHout, hout_bf, 2, 3, dims, strides, box, true);
    if (rc) return rc;
  }
  Lstm256P p;
  p.X = X; p.L = L; p.wt = (const __nv_bfloat16*)wt_scratch;
  dim3 grid(L2_CL * ceil_div(X, 2 * nh), 2);
  if (big) {
    IN_CUDA(cudaFuncSetAttribute(tc_lstm256_fwd_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, L2Smem<32>::BYTES));
    tc_lstm256_fwd_kernel<32><<<grid, L2_THREADS, L2Smem<32>::BYTES, st>>>(tmX, tmHout, p);
  } else {
    IN_CUDA(cudaFuncSetAttribute(tc_lstm256_fwd_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, L2Smem<16>::BYTES));
    tc_lstm256_fwd_kernel<16><<<grid, L2_THREADS, L2Smem<16>::BYTES, st>>>(tmX, tmHout, p);
  }
  IN_LAUNCH_CHECK("tc_lstm256_fwd_kernel");
  return INDNET_OK;
}
