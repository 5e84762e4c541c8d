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

__device__ __forceinline__ bool warp_id_is_zero(int tid) { return tid < 32; }

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
  // Warp 0 walks the stream in rows of 32 draws.  With c = draws accepted so far in the row before
  // lane l, lane l's Fisher-Yates index is i - c, its mask follows from that index, and it is accepted
  // iff (y & mask) <= index.  That is a triangular system over the lanes, settled exactly by ballot
  // iteration from the expected-rate guess (it normally confirms on the second ballot, and it crosses
  // the mask changes inside the row); the carry into the next row is then exact, so there is no
  // iteration across rows.  Meanwhile the whole CTA keeps a ring of untempered MT19937 blocks ahead of
  // the consumer (3-phase parallel twist); the state handed back is the block the consumer stopped in.
  constexpr int RING = 4;      // power of two
  __shared__ uint32_t ring[RING][MT_N];
  __shared__ int s_blk, s_off, s_done;
  const uint32_t below = (1u << lane) - 1u;
  const int jobs = B * (N + 1);
  for (int k = tid; k < MT_N; k += SMP_NT) ring[0][k] = nkey[k];
  __syncthreads();
  int gen = 1;                               // blocks generated so far (block b lives in ring[b % RING])
  int blk = 0, off = (int)np_state[MT_N];    // next draw to consume = block blk, offset off (off may be 624)
  int done = 0, flag = SMP_OK;
  // warp-0 job state
  int job = 0, i = 0;
  bool fresh = true;
  int* jrow = jlist;
  // permutation length of job j (>= 0), or -(error code); jobs past the end report 0
  auto job_size = [&](int j) -> int {
    if (j >= jobs) return 0;
    const int b = j / (N + 1), ci = j % (N + 1);
    if (ci == N) return N * Q;
    const int c = classes[(long long)b * N + ci];
    const int n = scopes[2 * c + 1] - scopes[2 * c];
    if (n < K + Q) return -SMP_SCOPE_TOO_SMALL;
    if (n > n_max) return -SMP_SCOPE_TOO_LARGE;
    return n;
  };
  int n_next = job_size(0);
  const long long clk0 = clock64();
  int n_pass = 0, n_round = 0;
  while (!done) {
    // ---- keep at least one full block (plus a row) generated beyond the consumer
    while ((gen - 1 - blk) * MT_N - off < MT_N + 32 && gen - blk < RING) {
      const uint32_t* o = ring[(gen - 1) % RING];
      uint32_t* w = ring[gen % RING];
      if (tid < 227) w[tid] = mt_mix(o[tid], o[tid + 1], o[tid + MT_M]);
      __syncthreads();
      if (tid < 227) w[tid + 227] = mt_mix(o[tid + 227], o[tid + 228], w[tid]);
      __syncthreads();
      if (tid < 170) {
        const int k = tid + 454;
        w[k] = mt_mix(o[k], (k + 1 < MT_N) ? o[k + 1] : w[0], w[k - 227]);
      }
      __syncthreads();
      ++gen;
    }
    if (warp_id_is_zero(tid)) {
      // Groups of GR rows (32 draws each) are consumed while a whole group is available in the ring.
      // Everything that does not depend on the carry (ring reads, tempering) is done for the whole
      // group up front so those instruction streams overlap; only the accept chain is serial.
      constexpr int GR = 4;
      while (true) {
        if (fresh) {
          if (job >= jobs) { done = 1; break; }
          i = n_next - 1;
          if (n_next < 0) { flag = -n_next; done = 1; break; }
          jrow = jlist + (long long)job * n_max;
          n_next = job_size(job + 1);          // prefetch: its global-load latency hides behind this job's rows
          fresh = false;
          if (i <= 0) { ++job; fresh = true; continue; }   // permutation(1): no draws
        }
        if ((gen - blk) * MT_N - off < 32 * GR) break;     // need more blocks
        uint32_t y[GR];
#pragma unroll
        for (int r = 0; r < GR; ++r) {
          int pb = blk, po = off + r * 32 + lane;
          if (po >= MT_N) { po -= MT_N; ++pb; }
          y[r] = mt_temper(ring[pb & (RING - 1)][po]);
        }
        int used = 32 * GR;
#pragma unroll
        for (int r = 0; r < GR; ++r) {
          const uint32_t m0 = mask_for((uint32_t)i);
          const int sh = 32 - __clz(m0);
          int c = (lane * (i + 1)) >> sh;                  // expected accepts before this lane
          uint32_t bal, v;
          int ip;
          if (i - 31 > (int)(m0 >> 1)) {
            // common case: every index this row can reach shares the mask -> accept iff c <= i - v
            v = y[r] & m0;
            const int t = i - (int)v;
            bal = __ballot_sync(0xffffffffu, c <= t);
            while (true) {
              c = __popc(bal & below);
              const uint32_t bal2 = __ballot_sync(0xffffffffu, c <= t);
              ++n_pass;
              if (bal2 == bal) break;
              bal = bal2;
            }
            ip = i - c;
          } else {
            ip = i - c;
            v = y[r] & mask_for((uint32_t)max(ip, 1));
            bal = __ballot_sync(0xffffffffu, ip >= 1 && v <= (uint32_t)ip);
            while (true) {
              c = __popc(bal & below);
              ip = i - c;
              v = y[r] & mask_for((uint32_t)max(ip, 1));
              const uint32_t bal2 = __ballot_sync(0xffffffffu, ip >= 1 && v <= (uint32_t)ip);
              ++n_pass;
              if (bal2 == bal) break;
              bal = bal2;
            }
          }
          const int acc = __popc(bal);
          if ((bal >> lane) & 1u) jrow[ip] = (int)v;
          ++n_round;
          if (acc >= i) {                                  // the permutation ends at its last accepted draw
            used = r * 32 + (32 - __clz(bal));
            ++job;
            fresh = true;
            break;
          }
          i -= acc;
        }
        off += used;
        if (off > MT_N) { off -= MT_N; ++blk; }            // off == 624 stays: "end of block blk"
      }
      if (lane == 0) { s_blk = blk; s_off = off; s_done = done; }
    }
    __syncthreads();
    blk = s_blk; off = s_off; done = s_done;
    __syncthreads();
  }
  {
    // hand back the block the consumer stopped in (an offset of 624 = "block exhausted", which is
    // also what the host generators hold at a block boundary)
    const uint32_t* fin = ring[blk % RING];
    for (int k = tid; k < MT_N; k += SMP_NT) np_state[k] = fin[k];
    if (tid == 0) {
      np_state[MT_N] = (uint32_t)off;
      status[0] = flag;
      // diagnostics: SM cycles spent in the stream walk, ballot passes, rows
      status[1] = (int)((clock64() - clk0) >> 4);
      status[2] = n_pass;
      status[3] = n_round;
    }
  }
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
