// a1: JSONFileDataLoader.next_one_tf (util/data_loader.py:259-309) on the device, bit-exact.
//
// The reference consumes two independent Mersenne-Twister streams per episode:
//   * CPython `random`:  random.sample(rel2scope.keys(), N)                       (:260)
//   * NumPy legacy global RandomState: np.random.choice(range(lo,hi), K+Q, False)   (:267, one per class)
//     == lo + permutation(hi-lo)[:K+Q], then np.random.permutation(N*Q)            (:296)
// Both are rejection samplers, so the stream position after an episode is data dependent and the
// draws cannot be partitioned ahead of time.  Kernel 1 (one CTA) walks the streams:
//   - MT19937 blocks of 624 words are regenerated cooperatively (3-phase parallel twist);
//   - the Python draws (few) are taken by one thread; the pool / set algorithms of random.sample are
//     restated literally;
//   - the NumPy Fisher-Yates draws (~1000 per class) are consumed up to 256 at a time by the CTA:
//     lane p's index is i - (#accepted before p), its mask follows from that index, and "accepted iff
//     masked draw <= index" is a triangular system solved by a ballot/prefix-count fixed-point
//     iteration (exact, and it crosses the mask changes inside a batch); accepted (i, j) pairs go to
//     a j-list.
// Kernel 2 turns each j-list into the first K+Q entries of its permutation WITHOUT simulating the
// swaps: the value that ends at position p is found by tracing p backwards through the swap list
// (identity start), one thread per output.  Kernel 3 assembles support||permuted-query row ids,
// labels, and gathers the four [instances, L] int32 tables (:268-271, :286-307).
// Stream states use the (key[624], pos) layout of random.getstate() / np.random.get_state(), so the
// host can hand its streams in and take them back.
#include "common.cuh"
#include "indnet_internal.h"
#include "../../include/indnet.h"

namespace indnet {

constexpr int MT_N = 624, MT_M = 397;
constexpr int SMP_NT = 256;

__device__ __forceinline__ uint32_t mt_temper(uint32_t y) {
  y ^= y >> 11;
  y ^= (y << 7) & 0x9D2C5680u;
  y ^= (y << 15) & 0xEFC60000u;
  y ^= y >> 18;
  return y;
}
__device__ __forceinline__ uint32_t mt_mix(uint32_t a, uint32_t b, uint32_t c) {
  const uint32_t y = (a & 0x80000000u) | (b & 0x7FFFFFFFu);
  return c ^ (y >> 1) ^ ((y & 1u) ? 0x9908B0DFu : 0u);
}
// Cooperative in-place twist of key[624] by the whole CTA.  new[i] needs old[i], old[i+1] and
// (old|new)[(i+397)%624]: phases [0,227) | [227,454) | [454,624) keep every read well defined.
__device__ void mt_twist_block(uint32_t* key) {
  const int t = threadIdx.x;
  __syncthreads();
  uint32_t r = 0;
  if (t < 227) r = mt_mix(key[t], key[t + 1], key[t + MT_M]);
  __syncthreads();
  if (t < 227) key[t] = r;
  __syncthreads();
  if (t < 227) r = mt_mix(key[t + 227], key[t + 228], key[t]);
  __syncthreads();
  if (t < 227) key[t + 227] = r;
  __syncthreads();
  if (t < 170) {
    const int i = t + 454;
    r = mt_mix(key[i], key[(i + 1) % MT_N], key[i - 227]);   // i == 623 reads new key[0]
  }
  __syncthreads();
  if (t < 170) key[t + 454] = r;
  __syncthreads();
}
// Sequential twist by one thread (Python stream: a handful of draws per episode).
__device__ void mt_twist_serial(uint32_t* key) {
  for (int i = 0; i < MT_N; ++i) key[i] = mt_mix(key[i], key[(i + 1) % MT_N], key[(i + MT_M) % MT_N]);
}
struct SerialMT {
  uint32_t* key; int pos;
  __device__ uint32_t next() {
    if (pos >= MT_N) { mt_twist_serial(key); pos = 0; }
    return mt_temper(key[pos++]);
  }
  // CPython Random._randbelow_with_getrandbits
  __device__ int randbelow(int n) {
    const int k = 32 - __clz(n);
    uint32_t r = next() >> (32 - k);
    while (r >= (uint32_t)n) r = next() >> (32 - k);
    return (int)r;
  }
};

// NumPy random_interval mask: smallest 2^k - 1 >= mx (mx >= 1): all ones below the top set bit.
__device__ __forceinline__ uint32_t mask_for(uint32_t mx) { return 0xFFFFFFFFu >> __clz(mx); }

// status codes written to *status
enum { SMP_OK = 0, SMP_SCOPE_TOO_SMALL = 1, SMP_SCOPE_TOO_LARGE = 2 };

__global__ void __launch_bounds__(SMP_NT)
sampler_walk_kernel(uint32_t* __restrict__ py_state, uint32_t* __restrict__ np_state,
                    const int* __restrict__ scopes, int R, int B, int N, int K, int Q, int n_max,
                    int* __restrict__ classes, int* __restrict__ jlist, int* __restrict__ status) {
  __shared__ uint32_t pkey[MT_N];
  __shared__ uint32_t nkey[MT_N];
  __shared__ int pool[1024];
  __shared__ uint32_t seen[32];
  const int tid = threadIdx.x, lane = tid & 31;
  for (int i = tid; i < MT_N; i += SMP_NT) { pkey[i] = py_state[i]; nkey[i] = np_state[i]; }
  __syncthreads();

  // ---- phase A: the Python stream, random.sample(range(R), N) per episode (thread 0)
  if (tid == 0) {
    SerialMT mt{pkey, (int)py_state[MT_N]};
    int setsize = 21;
    if (N > 5) {   // setsize += 4 ** ceil(log(N*3, 4))
      int p = 1;
      while (p < 3 * N) p *= 4;
      setsize += p;
    }
    for (int b = 0; b < B; ++b) {
      int* out = classes + (long long)b * N;
      if (R <= setsize) {
        for (int i = 0; i < R; ++i) pool[i] = i;
        for (int i = 0; i < N; ++i) {
          const int j = mt.randbelow(R - i);
          out[i] = pool[j];
          pool[j] = pool[R - i - 1];
        }
      } else {
        for (int i = 0; i < 32; ++i) seen[i] = 0u;
        for (int i = 0; i < N; ++i) {
          int j = mt.randbelow(R);
          while (seen[j >> 5] & (1u << (j & 31))) j = mt.randbelow(R);
          seen[j >> 5] |= 1u << (j & 31);
          out[i] = j;
        }
      }
    }
    for (int i = 0; i < MT_N; ++i) py_state[i] = pkey[i];
    py_state[MT_N] = (uint32_t)mt.pos;
  }
  __syncthreads();

  // ---- phase B: the NumPy stream.  jobs: for each episode, N class permutations then the query one.
  // The whole CTA consumes up to 256 draws per round.  Lane p (stream order) needs c_p = number of
  // draws accepted before it: then its Fisher-Yates index is i_p = i - c_p, its mask is mask(i_p),
  // and it is accepted iff (y_p & mask(i_p)) <= i_p.  That is a triangular system (lane p depends
  // only on lanes before it), so iterating "recount, re-decide" from any start reaches the exact
  // sequential answer; it settles in a few rounds because only draws near a threshold ever flip.
  // Every thread keeps the same (job, i, cur) in registers - all updates derive from shared counts.
  __shared__ uint32_t ntemp[MT_N];
  __shared__ __align__(16) int wcount[SMP_NT / 32];
  __shared__ int s_pos;
  const int warp = tid >> 5;
  const uint32_t below = (1u << lane) - 1u;
  const int jobs = B * (N + 1);
  int cur = (int)np_state[MT_N];
  for (int k = tid; k < MT_N; k += SMP_NT) ntemp[k] = mt_temper(nkey[k]);
  __syncthreads();
  int job = 0, i = 0, flag = SMP_OK;
  bool fresh = true;
  while (true) {
    if (fresh) {
      if (job >= jobs) break;
      const int b = job / (N + 1), ci = job % (N + 1);
      int n;
      if (ci < N) {
        const int c = classes[(long long)b * N + ci];
        n = scopes[2 * c + 1] - scopes[2 * c];
        if (n < K + Q) { flag = SMP_SCOPE_TOO_SMALL; break; }
        if (n > n_max) { flag = SMP_SCOPE_TOO_LARGE; break; }
      } else {
        n = N * Q;
      }
      i = n - 1;
      fresh = false;
      if (i <= 0) { ++job; fresh = true; continue; }   // permutation(1): no draws
    }
    if (cur >= MT_N) {
      mt_twist_block(nkey);
      for (int k = tid; k < MT_N; k += SMP_NT) ntemp[k] = mt_temper(nkey[k]);
      __syncthreads();
      cur = 0;
    }
    const int idx = cur + tid;
    const bool valid = idx < MT_N;
    const uint32_t y = valid ? ntemp[idx] : 0u;
    // Two-level fixed point.  Outer: each warp publishes how many of its 32 draws it accepts and
    // reads its carry-in (accepted by the warps before it).  Inner: with that carry the warp settles
    // its own 32 lanes exactly by ballot iteration (no block barrier).  Repeat until no warp's count
    // moves.  First guess: the expected number of earlier accepts at the batch's acceptance rate.
    const uint32_t m0 = mask_for((uint32_t)i);
    int c = (int)(((long long)tid * (i + 1)) >> (32 - __clz(m0)));
    uint32_t v = 0;
    bool acc_me = false;
    {
      const int ip = i - c;
      if (valid && ip >= 1) { v = y & mask_for((uint32_t)ip); acc_me = v <= (uint32_t)ip; }
    }
    uint32_t bal = __ballot_sync(0xffffffffu, acc_me);
    int total = 0;
    while (true) {
      const int published = __popc(bal);
      if (lane == 0) wcount[warp] = published;
      __syncthreads();
      const int4 w0 = *reinterpret_cast<const int4*>(&wcount[0]);
      const int4 w1 = *reinterpret_cast<const int4*>(&wcount[4]);
      const int wc[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
      total = ((wc[0] + wc[1]) + (wc[2] + wc[3])) + ((wc[4] + wc[5]) + (wc[6] + wc[7]));
      int prefix = 0;
#pragma unroll
      for (int w = 0; w < SMP_NT / 32; ++w) prefix += (w < warp) ? wc[w] : 0;
      while (true) {      // warp-local: exact for this carry-in
        c = prefix + __popc(bal & below);
        const int ip = i - c;
        bool again = false;
        if (valid && ip >= 1) { v = y & mask_for((uint32_t)ip); again = v <= (uint32_t)ip; }
        const uint32_t bal2 = __ballot_sync(0xffffffffu, again);
        acc_me = again;
        if (bal2 == bal) break;
        bal = bal2;
      }
      if (!__syncthreads_or(__popc(bal) != published)) break;
    }
    int consumed = min(SMP_NT, MT_N - cur);
    if (total == i) {                 // the permutation completes inside this round
      if (acc_me && c == total - 1) s_pos = tid;
      __syncthreads();
      consumed = s_pos + 1;
    }
    if (acc_me) jlist[(long long)job * n_max + (i - c)] = (int)v;
    i -= total;
    cur += consumed;
    if (i <= 0) { ++job; fresh = true; }
  }
  __syncthreads();
  for (int k = tid; k < MT_N; k += SMP_NT) np_state[k] = nkey[k];
  if (tid == 0) { np_state[MT_N] = (uint32_t)cur; *status = flag; }
}

// Value that ends at position p after the Fisher-Yates swaps (i = n-1..1, swap a[i], a[j_i]) applied
// to the identity: walk the swaps backwards (i ascending) moving the tracked position.
__global__ void sampler_trace_kernel(const int* __restrict__ scopes, const int* __restrict__ classes,
                                     const int* __restrict__ jlist, int B, int N, int K, int Q, int n_max,
                                     int pick_max, int* __restrict__ picks) {
  const long long id = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const int jobs = B * (N + 1);
  if (id >= (long long)jobs * pick_max) return;
  const int job = (int)(id / pick_max), p = (int)(id % pick_max);
  const int b = job / (N + 1), ci = job % (N + 1);
  int n, want;
  if (ci < N) {
    const int c = classes[(long long)b * N + ci];
    n = scopes[2 * c + 1] - scopes[2 * c];
    want = K + Q;
  } else {
    n = N * Q;
    want = N * Q;
  }
  if (p >= want || p >= n) return;
  const int* jl = jlist + (long long)job * n_max;
  int pos = p;
  for (int i = 1; i < n; ++i) {
    const int j = jl[i];
    if (pos == i) pos = j;
    else if (pos == j) pos = i;
  }
  picks[(long long)job * pick_max + p] = pos;
}

// rows[b][r]: r < N*K -> support (class-major); else query slot p = r - N*K holding pre-permutation
// query perm_q[p] (class = perm_q[p] / Q).  Then gather the four tables.
__global__ void sampler_gather_kernel(const int* __restrict__ scopes, const int* __restrict__ classes,
                                      const int* __restrict__ picks, int pick_max, int B, int N, int K, int Q,
                                      int L, const int* __restrict__ t_word, const int* __restrict__ t_pos1,
                                      const int* __restrict__ t_pos2, const int* __restrict__ t_mask,
                                      int* __restrict__ rows, int* __restrict__ labels, int* __restrict__ o_word,
                                      int* __restrict__ o_pos1, int* __restrict__ o_pos2, int* __restrict__ o_mask) {
  const int n_s = N * (K + Q), NK = N * K;
  const int b = blockIdx.x / n_s, r = blockIdx.x % n_s;
  int ci, slot;
  if (r < NK) { ci = r / K; slot = r % K; }
  else {
    const int pq = picks[((long long)b * (N + 1) + N) * pick_max + (r - NK)];
    ci = pq / Q; slot = K + pq % Q;
    if (threadIdx.x == 0) labels[(long long)b * (N * Q) + (r - NK)] = ci;
  }
  const int c = classes[(long long)b * N + ci];
  const int row = scopes[2 * c] + picks[((long long)b * (N + 1) + ci) * pick_max + slot];
  if (threadIdx.x == 0) rows[(long long)b * n_s + r] = row;
  const long long src = (long long)row * L, dst = (long long)blockIdx.x * L;
  for (int t = threadIdx.x; t < L; t += blockDim.x) {
    o_word[dst + t] = t_word[src + t];
    o_pos1[dst + t] = t_pos1[src + t];
    o_pos2[dst + t] = t_pos2[src + t];
    o_mask[dst + t] = t_mask[src + t];
  }
}

}  // namespace indnet

using namespace indnet;

extern "C" long long indnet_sampler_scratch_bytes(int B, int N, int K, int Q, int n_max) {
  if (B < 0 || N <= 0 || K < 0 || Q < 0 || n_max <= 0) return -1;
  const long long jobs = (long long)B * (N + 1);
  const long long pick_max = (K + Q > N * Q) ? (K + Q) : (N * Q);
  const long long stride = (n_max > N * Q) ? n_max : (N * Q);   // the query permutation has N*Q entries
  return 4 * ((long long)B * N + jobs * stride + jobs * pick_max + 64);
}

static int sample_walk(cudaStream_t st, unsigned int* py_state, unsigned int* np_state, const int* scopes, int R,
                       int B, int N, int K, int Q, int n_max, void* scratch, int* status) {
  IN_REQUIRE(B >= 0 && N > 0 && K > 0 && Q > 0 && R > 0 && n_max > 0, "sample_episodes: bad sizes");
  IN_REQUIRE(N <= R, "sample_episodes: Sample larger than population (N=%d classes of %d relations)", N, R);
  IN_REQUIRE(R <= 1024, "sample_episodes: at most 1024 relations supported, got %d", R);
  if (B == 0) return INDNET_OK;
  IN_REQUIRE(py_state && np_state && scopes && scratch && status, "sample_episodes: null pointer");
  const long long jobs = (long long)B * (N + 1);
  const int pick_max = (K + Q > N * Q) ? (K + Q) : (N * Q);
  if (n_max < N * Q) n_max = N * Q;   // j-list stride must also hold the N*Q-entry query permutation
  int* classes = (int*)scratch;
  int* jlist = classes + (((long long)B * N + 15) / 16) * 16;
  int* picks = jlist + jobs * n_max;
  sampler_walk_kernel<<<1, SMP_NT, 0, st>>>(py_state, np_state, scopes, R, B, N, K, Q, n_max, classes, jlist, status);
  IN_LAUNCH_CHECK("sampler_walk_kernel");
  sampler_trace_kernel<<<ceil_div(jobs * pick_max, 128), 128, 0, st>>>(scopes, classes, jlist, B, N, K, Q, n_max,
                                                                       pick_max, picks);
  IN_LAUNCH_CHECK("sampler_trace_kernel");
  return INDNET_OK;
}

static int sample_gather(cudaStream_t st, const int* scopes, const int* t_word, const int* t_pos1, const int* t_pos2,
                         const int* t_mask, int L, int B, int N, int K, int Q, int n_max, const void* scratch,
                         int* rows, int* labels, int* o_word, int* o_pos1, int* o_pos2, int* o_mask) {
  IN_REQUIRE(B >= 0 && N > 0 && K > 0 && Q > 0 && L > 0 && n_max > 0, "sample_gather: bad sizes");
  if (B == 0) return INDNET_OK;
  IN_REQUIRE(scopes && t_word && t_pos1 && t_pos2 && t_mask && scratch && rows && labels && o_word && o_pos1 &&
             o_pos2 && o_mask, "sample_gather: null pointer");
  const long long jobs = (long long)B * (N + 1);
  const int pick_max = (K + Q > N * Q) ? (K + Q) : (N * Q);
  if (n_max < N * Q) n_max = N * Q;
  const int* classes = (const int*)scratch;
  const int* jlist = classes + (((long long)B * N + 15) / 16) * 16;
  const int* picks = jlist + jobs * n_max;
  sampler_gather_kernel<<<B * N * (K + Q), 64, 0, st>>>(scopes, classes, picks, pick_max, B, N, K, Q, L, t_word,
                                                       t_pos1, t_pos2, t_mask, rows, labels, o_word, o_pos1, o_pos2,
                                                       o_mask);
  IN_LAUNCH_CHECK("sampler_gather_kernel");
  return INDNET_OK;
}

extern "C" int indnet_sample_episodes(indnet_stream st_, unsigned int* py_state, unsigned int* np_state,
                                      const int* scopes, int R, const int* t_word, const int* t_pos1,
                                      const int* t_pos2, const int* t_mask, int L, int B, int N, int K, int Q,
                                      int n_max, void* scratch, int* rows, int* labels, int* o_word, int* o_pos1,
                                      int* o_pos2, int* o_mask, int* status) {
  cudaStream_t st = (cudaStream_t)st_;
  int rc = sample_walk(st, py_state, np_state, scopes, R, B, N, K, Q, n_max, scratch, status);
  if (rc != INDNET_OK) return rc;
  return sample_gather(st, scopes, t_word, t_pos1, t_pos2, t_mask, L, B, N, K, Q, n_max, scratch, rows, labels,
                       o_word, o_pos1, o_pos2, o_mask);
}

extern "C" int indnet_sample_walk(indnet_stream st, unsigned int* py_state, unsigned int* np_state, const int* scopes,
                                  int R, int B, int N, int K, int Q, int n_max, void* scratch, int* status) {
  return sample_walk((cudaStream_t)st, py_state, np_state, scopes, R, B, N, K, Q, n_max, scratch, status);
}

extern "C" int indnet_sample_gather(indnet_stream st, const int* scopes, const int* t_word, const int* t_pos1,
                                    const int* t_pos2, const int* t_mask, int L, int B, int N, int K, int Q, int n_max,
                                    const void* scratch, int* rows, int* labels, int* o_word, int* o_pos1,
                                    int* o_pos2, int* o_mask) {
  return sample_gather((cudaStream_t)st, scopes, t_word, t_pos1, t_pos2, t_mask, L, B, N, K, Q, n_max, scratch, rows,
                       labels, o_word, o_pos1, o_pos2, o_mask);
}
