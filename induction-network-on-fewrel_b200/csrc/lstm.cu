// a3 (FP32 parity mode): the BiLSTM recurrence of model/graph.py:34-49 as persistent kernels.
//
// The input projection x_t * W_x + bias for all timesteps is done beforehand as one GEMM
// (the first D rows of the TF kernel [D+u, 4u] are W_x, the last u rows are W_h), so
// each CTA here owns a tile of SB sentences of one direction for ALL L steps: h and c stay in
// shared memory, the gate math (i, j, f, o order, forget bias 1.0) is fused, and no length masking
// is applied (the reference passes no sequence_length - padding steps are processed, the backward
// direction starts on the padding).  Activations needed by BPTT are saved; the backward kernel walks
// the steps in reverse and emits dz (gradient w.r.t. the pre-activation gates) for the weight GEMMs.
#include "common.cuh"
#include "indnet_internal.h"
#include "../../include/indnet.h"

namespace indnet {

constexpr int SB = 16;   // sentences per CTA
constexpr int ST = 8;    // sentences per thread (register blocking)
constexpr int LSTM_NT = 256;

__global__ void __launch_bounds__(LSTM_NT)
lstm_fwd_kernel(const float* __restrict__ xproj, const float* __restrict__ Wh_fw,
                const float* __restrict__ Wh_bw, float* __restrict__ hout, float* __restrict__ gates,
                float* __restrict__ cstate, float* __restrict__ hprev, int X, int L, int u,
                float keep_prob, unsigned long long seed, const long long* __restrict__ seed_step) {
  if (seed_step) seed += (unsigned long long)(*seed_step) * 1000003ull;
  extern __shared__ __align__(16) float sm[];
  float* hs = sm;                  // [2][u][SB]
  float* cs = sm + 2 * u * SB;     // [u][SB]
  const int dir = blockIdx.y;
  const int x0 = blockIdx.x * SB;
  const float* __restrict__ Wh = dir ? Wh_bw : Wh_fw;
  const int tid = threadIdx.x;
  const int U4 = 4 * u;
  for (int i = tid; i < 3 * u * SB; i += LSTM_NT) sm[i] = 0.f;
  __syncthreads();
  const int items = u * (SB / ST);
  const float inv_keep = 1.0f / keep_prob;

  for (int step = 0; step < L; ++step) {
    const int t = dir ? (L - 1 - step) : step;
    const float* hcur = hs + (step & 1) * u * SB;
    float* hnxt = hs + ((step & 1) ^ 1) * u * SB;
    for (int item = tid; item < items; item += LSTM_NT) {
      const int unit = item % u, sg = item / u;
      float acc[4][ST];
#pragma unroll
      for (int s = 0; s < ST; ++s) {
        const int x = x0 + sg * ST + s;
        if (x < X) {
          const float* xp = xproj + (((long long)dir * X + x) * L + t) * U4 + unit;
#pragma unroll
          for (int g = 0; g < 4; ++g) acc[g][s] = xp[g * u];
        } else {
#pragma unroll
          for (int g = 0; g < 4; ++g) acc[g][s] = 0.f;
        }
      }
      const float* wp = Wh + unit;
      const float* hp = hcur + sg * ST;
#pragma unroll 4
      for (int k = 0; k < u; ++k) {
        const float w0 = wp[0], w1 = wp[u], w2 = wp[2 * u], w3 = wp[3 * u];
        const float4 ha = *reinterpret_cast<const float4*>(hp);
        const float4 hb = *reinterpret_cast<const float4*>(hp + 4);
        const float hv[ST] = {ha.x, ha.y, ha.z, ha.w, hb.x, hb.y, hb.z, hb.w};
#pragma unroll
        for (int s = 0; s < ST; ++s) {
          acc[0][s] = fmaf(hv[s], w0, acc[0][s]);
          acc[1][s] = fmaf(hv[s], w1, acc[1][s]);
          acc[2][s] = fmaf(hv[s], w2, acc[2][s]);
          acc[3][s] = fmaf(hv[s], w3, acc[3][s]);
        }
        wp += U4;
        hp += SB;
      }
#pragma unroll
      for (int s = 0; s < ST; ++s) {
        const int sl = sg * ST + s;
        const int x = x0 + sl;
        const float gi = sigmoid_acc(acc[0][s]);
        const float gj = tanhf(acc[1][s]);
        const float gf = sigmoid_acc(acc[2][s] + 1.0f);
        const float go = sigmoid_acc(acc[3][s]);
        const float c = cs[unit * SB + sl] * gf + gi * gj;
        const float h = tanhf(c) * go;
        cs[unit * SB + sl] = c;
        hnxt[unit * SB + sl] = h;
        if (x < X) {
          const long long tok = ((long long)dir * X + x) * L + t;
          float* gp = gates + tok * U4 + unit;
          gp[0] = gi; gp[u] = gj; gp[2 * u] = gf; gp[3 * u] = go;
          cstate[tok * u + unit] = c;
          hprev[tok * u + unit] = hcur[unit * SB + sl];
          const long long oi = ((long long)x * L + t) * (2 * u) + dir * u + unit;
          float ho = h;
          if (keep_prob < 1.0f) ho = dropout_keep(seed, (unsigned long long)oi, x, (unsigned long long)L * 2 * u, keep_prob) ? h * inv_keep : 0.f;
          hout[oi] = ho;
        }
      }
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(LSTM_NT)
lstm_bwd_kernel(const float* __restrict__ dhout, const float* __restrict__ gates,
                const float* __restrict__ cstate, const float* __restrict__ WhT_all,
                float* __restrict__ dz, int X, int L, int u, float keep_prob, unsigned long long seed,
                const long long* __restrict__ seed_step) {
  if (seed_step) seed += (unsigned long long)(*seed_step) * 1000003ull;
  extern __shared__ __align__(16) float sm[];
  const int U4 = 4 * u;
  float* dzs = sm;                    // [4u][SB]
  float* dhr = sm + U4 * SB;          // [u][SB]
  float* dcr = dhr + u * SB;          // [u][SB]
  const int dir = blockIdx.y;
  const int x0 = blockIdx.x * SB;
  const float* __restrict__ WhT = WhT_all + (long long)dir * U4 * u;   // [4u][u]
  const int tid = threadIdx.x;
  for (int i = tid; i < 6 * u * SB; i += LSTM_NT) sm[i] = 0.f;
  __syncthreads();
  const float inv_keep = 1.0f / keep_prob;
  const int items2 = u * (SB / ST);

  for (int step = 0; step < L; ++step) {
    const int pstep = L - 1 - step;                 // forward processing index of this step
    const int t = dir ? (L - 1 - pstep) : pstep;
    const int tprev = dir ? (t + 1) : (t - 1);      // time index processed just before t (if pstep>0)
    // ---- phase 1: elementwise gate gradients
    for (int item = tid; item < u * SB; item += LSTM_NT) {
      const int unit = item % u, sl = item / u;
      const int x = x0 + sl;
      float di = 0.f, dj = 0.f, df = 0.f, dob = 0.f;
      if (x < X) {
        const long long tok = ((long long)dir * X + x) * L + t;
        const long long oi = ((long long)x * L + t) * (2 * u) + dir * u + unit;
        float dh = dhout[oi];
        if (keep_prob < 1.0f) dh = dropout_keep(seed, (unsigned long long)oi, x, (unsigned long long)L * 2 * u, keep_prob) ? dh * inv_keep : 0.f;
        dh += dhr[unit * SB + sl];
        const float* gp = gates + tok * U4 + unit;
        const float gi = gp[0], gj = gp[u], gf = gp[2 * u], go = gp[3 * u];
        const float c = cstate[tok * u + unit];
        const float cprev = (pstep > 0) ? cstate[(((long long)dir * X + x) * L + tprev) * u + unit] : 0.f;
        const float tc = tanhf(c);
        dob = dh * tc * go * (1.f - go);
        const float dc = dcr[unit * SB + sl] + dh * go * (1.f - tc * tc);
        di = dc * gj * gi * (1.f - gi);
        dj = dc * gi * (1.f - gj * gj);
        df = dc * cprev * gf * (1.f - gf);
        dcr[unit * SB + sl] = dc * gf;
        float* zp = dz + tok * U4 + unit;
        zp[0] = di; zp[u] = dj; zp[2 * u] = df; zp[3 * u] = dob;
      }
      dzs[(unit)*SB + sl] = di;
      dzs[(u + unit) * SB + sl] = dj;
      dzs[(2 * u + unit) * SB + sl] = df;
      dzs[(3 * u + unit) * SB + sl] = dob;
    }
    __syncthreads();
    // ---- phase 2: dh_rec[s, m] = sum_n dz[s, n] * Wh[m, n]
    if (pstep > 0) {
      for (int item = tid; item < items2; item += LSTM_NT) {
        const int m = item % u, sg = item / u;
        float acc[ST];
#pragma unroll
        for (int s = 0; s < ST; ++s) acc[s] = 0.f;
        const float* wp = WhT + m;
        const float* zp = dzs + sg * ST;
#pragma unroll 4
        for (int n = 0; n < U4; ++n) {
          const float w = *wp;
          const float4 za = *reinterpret_cast<const float4*>(zp);
          const float4 zb = *reinterpret_cast<const float4*>(zp + 4);
          acc[0] = fmaf(za.x, w, acc[0]); acc[1] = fmaf(za.y, w, acc[1]);
          acc[2] = fmaf(za.z, w, acc[2]); acc[3] = fmaf(za.w, w, acc[3]);
          acc[4] = fmaf(zb.x, w, acc[4]); acc[5] = fmaf(zb.y, w, acc[5]);
          acc[6] = fmaf(zb.z, w, acc[6]); acc[7] = fmaf(zb.w, w, acc[7]);
          wp += u;
          zp += SB;
        }
#pragma unroll
        for (int s = 0; s < ST; ++s) dhr[m * SB + sg * ST + s] = acc[s];
      }
    }
    __syncthreads();
  }
}

__global__ void transpose_kernel(const float* __restrict__ in, float* __restrict__ out, int rows, int cols) {
  __shared__ float tile[32][33];
  int c = blockIdx.x * 32 + threadIdx.x;
  for (int j = threadIdx.y; j < 32; j += 8) {
    int r = blockIdx.y * 32 + j;
    if (r < rows && c < cols) tile[j][threadIdx.x] = in[(long long)r * cols + c];
  }
  __syncthreads();
  int r2 = blockIdx.y * 32 + threadIdx.x;
  for (int j = threadIdx.y; j < 32; j += 8) {
    int c2 = blockIdx.x * 32 + j;
    if (r2 < rows && c2 < cols) out[(long long)c2 * rows + r2] = tile[threadIdx.x][j];
  }
}

}  // namespace indnet

using namespace indnet;

extern "C" int indnet_lstm_fwd(indnet_stream st_, const float* xproj, const float* Wh_fw, const float* Wh_bw,
                               float* hout, float* gates, float* cstate, float* hprev, int X, int L, int u,
                               float keep_prob, unsigned long long seed, const long long* seed_step) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(X >= 0 && L > 0 && u > 0, "lstm_fwd: bad sizes X=%d L=%d u=%d", X, L, u);
  IN_REQUIRE(keep_prob > 0.f && keep_prob <= 1.f, "lstm_fwd: keep_prob must be in (0, 1], got %f", keep_prob);
  if (X == 0) return INDNET_OK;
  IN_REQUIRE(xproj && Wh_fw && Wh_bw && hout && gates && cstate && hprev, "lstm_fwd: null pointer");
  size_t smem = sizeof(float) * 3 * u * SB;
  IN_REQUIRE(smem <= 227 * 1024, "lstm_fwd: hidden size %d too large for the FP32 recurrence tile", u);
  if (smem > 48 * 1024)
    IN_CUDA(cudaFuncSetAttribute(lstm_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid(ceil_div(X, SB), 2);
  lstm_fwd_kernel<<<grid, LSTM_NT, smem, st>>>(xproj, Wh_fw, Wh_bw, hout, gates, cstate, hprev, X, L, u, keep_prob, seed, seed_step);
  IN_LAUNCH_CHECK("lstm_fwd_kernel");
  return INDNET_OK;
}

extern "C" int indnet_lstm_bwd(indnet_stream st_, const float* dhout, const float* gates, const float* cstate,
                               const float* Wh_fw, const float* Wh_bw, float* whT_scratch, float* dz, int X, int L,
                               int u, float keep_prob, unsigned long long seed, const long long* seed_step) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(X >= 0 && L > 0 && u > 0, "lstm_bwd: bad sizes X=%d L=%d u=%d", X, L, u);
  IN_REQUIRE(keep_prob > 0.f && keep_prob <= 1.f, "lstm_bwd: keep_prob must be in (0, 1], got %f", keep_prob);
  if (X == 0) return INDNET_OK;
  IN_REQUIRE(dhout && gates && cstate && Wh_fw && Wh_bw && whT_scratch && dz, "lstm_bwd: null pointer");
  size_t smem = sizeof(float) * 6 * u * SB;
  IN_REQUIRE(smem <= 227 * 1024, "lstm_bwd: hidden size %d too large for the FP32 recurrence tile", u);
  dim3 tg(ceil_div(4 * u, 32), ceil_div(u, 32));
  transpose_kernel<<<tg, dim3(32, 8), 0, st>>>(Wh_fw, whT_scratch, u, 4 * u);
  IN_LAUNCH_CHECK("transpose_kernel");
  transpose_kernel<<<tg, dim3(32, 8), 0, st>>>(Wh_bw, whT_scratch + (long long)4 * u * u, u, 4 * u);
  IN_LAUNCH_CHECK("transpose_kernel");
  if (smem > 48 * 1024)
    IN_CUDA(cudaFuncSetAttribute(lstm_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid(ceil_div(X, SB), 2);
  lstm_bwd_kernel<<<grid, LSTM_NT, smem, st>>>(dhout, gates, cstate, whT_scratch, dz, X, L, u, keep_prob, seed, seed_step);
  IN_LAUNCH_CHECK("lstm_bwd_kernel");
  return INDNET_OK;
}

// ------------------------------------------------------------------------------------------------
// indnet_lstm_bwd_static: the BPTT entry point for callers that are not allowed to allocate.
//
// indnet_lstm_bwd needs a [2][4u][u] float scratch for the transposed recurrent weights and makes
// the caller provide it.  Some callers cannot: code running inside a host framework whose device
// allocator they do not own (a TF / JAX custom-op or a plugin callback that only receives the input
// and output buffers), real-time loops that forbid cudaMalloc because it synchronises the device,
// and processes whose whole memory budget was handed to another pool at start-up.  For them the
// scratch lives in this library's image: a fixed file-scope __device__ array sized for u <= 128
// (512 KiB).  Every GPU that loads the library carries those 512 KiB until the process exits.
//
// Rules that follow from the scratch being one shared array per device:
//  * u > INDNET_STATIC_MAX_U is rejected with an error - the array is not grown.
//  * Calls that may overlap on the same GPU (different streams, different host threads) race on it;
//    serialise them.  The array keeps whatever the last call wrote; nothing in it has to start at
//    zero because each call rewrites the [2][4u][u] region it uses before reading it.
//  * The device address of the array is looked up once per device, on the first call made with
//    that device current (cudaGetSymbolAddress; the array's name in host code is only a host-side
//    handle, not a device pointer).  Make that first call before starting any stream capture.
//    A first call made while the stream is capturing is refused with a clear error instead of
//    breaking the capture.  The cached address belongs to the device's current context: after
//    cudaDeviceReset() the library has to be reloaded.
//  * No cudaMalloc / cudaFree / allocating library routine is called here.
#define INDNET_STATIC_MAX_U 128
#define INDNET_STATIC_MAX_DEVICES 64

static __device__ __align__(256) float g_whT_static[2 * 4 * INDNET_STATIC_MAX_U * INDNET_STATIC_MAX_U];
static float* g_whT_static_addr[INDNET_STATIC_MAX_DEVICES] = {nullptr};

static int static_scratch_for_current_device(float** out, cudaStream_t st) {
  int dev = -1;
  IN_CUDA(cudaGetDevice(&dev));
  IN_REQUIRE(dev >= 0 && dev < INDNET_STATIC_MAX_DEVICES, "lstm_bwd_static: device ordinal %d out of range", dev);
  float* p = __atomic_load_n(&g_whT_static_addr[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    IN_CUDA(cudaStreamIsCapturing(st, &cap));
    IN_REQUIRE(cap == cudaStreamCaptureStatusNone,
               "lstm_bwd_static: the first call on a device looks up the scratch array's address and must be made "
               "outside stream capture");
    IN_CUDA(cudaGetSymbolAddress((void**)&p, g_whT_static));
    IN_REQUIRE(p != nullptr, "lstm_bwd_static: cudaGetSymbolAddress returned a null address");
    __atomic_store_n(&g_whT_static_addr[dev], p, __ATOMIC_RELEASE);   // same value from any thread
  }
  *out = p;
  return INDNET_OK;
}

extern "C" int indnet_lstm_bwd_static(indnet_stream st_, const float* dhout, const float* gates, const float* cstate,
                                      const float* Wh_fw, const float* Wh_bw, float* dz, int X, int L, int u,
                                      float keep_prob, unsigned long long seed, const long long* seed_step) {
  cudaStream_t st = (cudaStream_t)st_;
  IN_REQUIRE(X >= 0 && L > 0 && u > 0, "lstm_bwd_static: bad sizes X=%d L=%d u=%d", X, L, u);
  IN_REQUIRE(u <= INDNET_STATIC_MAX_U,
             "lstm_bwd_static: hidden size u=%d exceeds the built-in scratch (max u=%d); use indnet_lstm_bwd with "
             "caller-provided scratch", u, INDNET_STATIC_MAX_U);
  IN_REQUIRE(keep_prob > 0.f && keep_prob <= 1.f, "lstm_bwd_static: keep_prob must be in (0, 1], got %f", keep_prob);
  if (X == 0) return INDNET_OK;
  IN_REQUIRE(dhout && gates && cstate && Wh_fw && Wh_bw && dz, "lstm_bwd_static: null pointer");
  float* whT = nullptr;
  int rc = static_scratch_for_current_device(&whT, st);
  if (rc != INDNET_OK) return rc;
  size_t smem = sizeof(float) * 6 * u * SB;
  dim3 tg(ceil_div(4 * u, 32), ceil_div(u, 32));
  transpose_kernel<<<tg, dim3(32, 8), 0, st>>>(Wh_fw, whT, u, 4 * u);
  IN_LAUNCH_CHECK("transpose_kernel");
  transpose_kernel<<<tg, dim3(32, 8), 0, st>>>(Wh_bw, whT + (long long)4 * u * u, u, 4 * u);
  IN_LAUNCH_CHECK("transpose_kernel");
  if (smem > 48 * 1024)
    IN_CUDA(cudaFuncSetAttribute(lstm_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid(ceil_div(X, SB), 2);
  lstm_bwd_kernel<<<grid, LSTM_NT, smem, st>>>(dhout, gates, cstate, whT, dz, X, L, u, keep_prob, seed, seed_step);
  IN_LAUNCH_CHECK("lstm_bwd_kernel");
  return INDNET_OK;
}
