// a1: JSONFileDataLoader.next_one_tf (util/data_loader.py:259-309) on the device, bit-exact.
//
// The reference consumes two independent Mersenne-Twister streams per episode:
//   * CPython `random`:  random.sample(rel2scope.keys(), N)                       (:260)
//   * NumPy legacy global RandomState: np.random.choice(range(lo,hi), K+Q, False)   (:267, one per class)
//     == lo + permutation(hi-lo)[:K+Q], then np.random.permutation(N*Q)            (:296)
// Both are rejection samplers, so the stream position after an episode is data dependent and the
// NumPy draws (~1000 per class) form one serial chain.  The chain is made cheap instead of parallel:
//   G  smp_gen_kernel    the NumPy MT19937 stream for the whole batch (CTA 0) and, concurrently, the few
//                        Python draws - random.sample's pool / set algorithms restated literally - plus
//                        every job's permutation length (CTA 1);
//   T  smp_table_kernel  for every 64-draw block and EVERY Fisher-Yates index the index after the block
//                        (and after the pair of blocks) and the end offsets: job independent, built by
//                        the whole GPU;
//   W  smp_walk_kernel   one warp chains lookups through those tables (staged in shared memory by a
//                        producer warp) and walks the unaligned stretches exactly, 32 draws per row:
//                        lane p's index is i - (#accepted before p), its mask follows from that index,
//                        and "accepted iff masked draw <= index" is a triangular system solved by a
//                        ballot/prefix-count fixed-point iteration (exact across mask changes);
//   F  smp_fill_kernel   replays the walker's logged runs in parallel -> accepted (i, j) pairs (j-lists).
// sampler_trace_kernel turns each j-list into the first K+Q entries of its permutation WITHOUT
// simulating the swaps: the value that ends at position p is found by tracing p backwards through the
// swap list (identity start), one thread per output.  sampler_gather_kernel assembles
// support||permuted-query row ids, labels, and gathers the four [instances, L] int32 tables (:268-271,
// :286-307).
// Stream states use the (key[624], pos) layout of random.getstate() / np.random.get_state(), so the
// host can hand its streams in and take them back.
#include "common.cuh"
#include "indnet_internal.h"
#include "tc_common.cuh"
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
// NumPy random_interval mask: smallest 2^k - 1 >= mx (mx >= 1): all ones below the top set bit.
__device__ __forceinline__ uint32_t mask_for(uint32_t mx) { return 0xFFFFFFFFu >> __clz(mx); }

// status codes written to status[0]
enum { SMP_OK = 0, SMP_SCOPE_TOO_SMALL = 1, SMP_SCOPE_TOO_LARGE = 2, SMP_STREAM_BUDGET = 3 };

constexpr int TB = 64;     // draws per transition-table block

// ---- kernel G: both host streams, two CTAs ------------------------------------------------------
// CTA 0: nb consecutive untempered MT19937 blocks of the NumPy stream.  keys[0..624) = the state handed
// in; beyond that the recurrence runs in rounds of 227 words, two rounds per barrier, out of a 2048-word
// shared ring.
// CTA 1: the Python stream - random.sample(range(R), N) per episode, restated literally (pool / set
// variants of CPython's Random.sample), one thread; then the permutation length of every job.
constexpr int GEN_NT = 256;

// the part of the MT19937 recurrence that mixes two consecutive words: x[n] = x[n-227] ^ mt_g(x[n-624], x[n-623])
__device__ __forceinline__ uint32_t mt_g(uint32_t a, uint32_t b) {
  const uint32_t y = (a & 0x80000000u) | (b & 0x7FFFFFFFu);
  return (y >> 1) ^ ((y & 1u) ? 0x9908B0DFu : 0u);
}

// ---- persistent stream window ------------------------------------------------------------------
// The NumPy stream and its tables live in rings addressed by ABSOLUTE draw / block index (origin = the first
// word of the key block handed in at the last reset), kept in the caller's scratch across calls.  A call
// only needs the window [win_start, win_start + nblk) of blocks to exist before it walks; while it walks,
// the NEXT window is generated and tabulated on a forked stream (the continuation of the stream does not
// depend on where this walk stops), so in steady state generation and tables are off the call's critical
// path.  The header remembers the absolute position; a call whose np_state is not the state the ring
// expects (first call, re-seeded or reloaded stream, different sizes) resets the origin.
struct SmpPlan { long long gen_from, gen_to, tab_from, tab_to; };
struct SmpHdr {
  unsigned long long magic;      // identifies an initialised header for this layout
  long long apos;                // absolute index of the next draw to consume
  long long gen_front;           // words [.., gen_front) are generated
  long long tab_front;           // blocks [.., tab_front) are tabulated
  long long win_start;           // block of apos at the start of the current call
  SmpPlan plan[2];               // [0] what this call must still build, [1] what it builds ahead
  int nsec;
  int use_v2;                    // 1: the parallel path (jump tables + chain + replay) serves this call, 0: the serial walker
  int ntab;                      // v2: number of distinct permutation lengths (class scopes + the query permutation)
  int sizes[4];                  // v2: those lengths
  int uni_zc, uni_zq;            // v2, every relation the same size (FewRel): table index of the class / query
                                 // permutation length - every episode is then the same job sequence, and the chain
                                 // steps one EPISODE per lookup through the composed table (smp_epi_kernel); else -1
  uint32_t py_backup[MT_N + 1];  // Python stream as handed in: a failing call puts it back (streams unchanged on error)
};

// One CTA.  ahead == 0: check that np_state is the state the ring continues from (else reset), open the
// window; then plan what generation / table kernels of this phase have to do.
__global__ void smp_plan_kernel(SmpHdr* __restrict__ h, const uint32_t* __restrict__ np_state,
                                uint32_t* __restrict__ keys, uint32_t kmask, long long nblk, int ahead,
                                unsigned long long magic, const int* __restrict__ scopes, int R, int nq, int allow_v2) {
  if (ahead == 0 && threadIdx.x == 0) {
    // the parallel path tabulates "where does a permutation of length n end when it starts at draw p" for every
    // distinct n: feasible when the relations have at most three distinct sizes (FewRel: one)
    int nt = 0, nt_rel = 0, zq = 0, sz[4] = {0, 0, 0, 0};
    bool ok = allow_v2 != 0;
    for (int r = 0; r <= R && ok; ++r) {
      const int n = r < R ? scopes[2 * r + 1] - scopes[2 * r] : nq;
      bool seen = false;
      for (int k = 0; k < nt; ++k) if (sz[k] == n) { seen = true; zq = k; }
      if (!seen) { if (nt < 4) { zq = nt; sz[nt++] = n; } else ok = false; }
      if (r == R - 1) nt_rel = nt;
    }
    h->use_v2 = ok ? 1 : 0;
    h->ntab = ok ? nt : 0;
    for (int k = 0; k < 4; ++k) h->sizes[k] = sz[k];
    const bool uni = ok && allow_v2 == 1 && nt_rel == 1;
    h->uni_zc = uni ? 0 : -1;
    h->uni_zq = uni ? zq : -1;
  }
  if (ahead == 0) {
    bool ok = h->magic == magic;
    long long blk = 0;
    if (ok) {
      const long long pos = h->apos;
      blk = pos / MT_N;
      int off = (int)(pos % MT_N);
      if (off == 0 && blk > 0) { blk -= 1; off = MT_N; }
      ok = np_state[MT_N] == (uint32_t)off;
      if (ok)
        for (int k = threadIdx.x; k < MT_N; k += blockDim.x)
          if (keys[(uint32_t)(blk * MT_N + k) & kmask] != np_state[k]) ok = false;
    }
    const int all_ok = __syncthreads_and(ok ? 1 : 0);
    if (!all_ok) {                                   // new origin: the handed-in key block is words [0, 624)
      for (int k = threadIdx.x; k < MT_N; k += blockDim.x) keys[k] = np_state[k];
      if (threadIdx.x == 0) {
        h->magic = magic;
        h->apos = (long long)np_state[MT_N];
        h->gen_front = MT_N;
        h->tab_front = 0;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) h->win_start = h->apos >> 6;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const long long target = ((h->win_start + nblk * (1 + ahead)) + 7) & ~7LL;     // whole chunks
    SmpPlan pl;
    pl.tab_from = h->tab_front;
    pl.tab_to = target > pl.tab_from ? target : pl.tab_from;
    h->tab_front = pl.tab_to;
    const long long need = target * TB + MT_N;       // + one key block: the state handed back must be complete
    pl.gen_from = h->gen_front;
    const long long it = need > pl.gen_from ? (need - pl.gen_from + 453) / 454 : 0;
    pl.gen_to = pl.gen_from + it * 454;              // whole double rounds of the generator
    h->gen_front = pl.gen_to;
    h->plan[ahead] = pl;
  }
}

__global__ void __launch_bounds__(GEN_NT)
smp_gen_kernel(SmpHdr* __restrict__ h, int ahead, uint32_t* __restrict__ keys, uint32_t kmask,
               uint32_t* __restrict__ py_state, int R, int B, int N, int* __restrict__ classes,
               const int* __restrict__ scopes, int K, int Q, int n_max, int* __restrict__ jsize) {
  __shared__ uint32_t ring[2048];
  const int tid = threadIdx.x;
  if (blockIdx.x == 0) {
    const long long from = h->plan[ahead].gen_from, total = h->plan[ahead].gen_to;
    if (from >= total) return;
    for (int k = tid; k < MT_N; k += GEN_NT) {
      const long long n = from - MT_N + k;
      ring[n & 2047] = keys[(uint32_t)n & kmask];
    }
    __syncthreads();
    // Rounds of exactly 227 words: x[n-227] is then the word this thread produced one round earlier (kept in
    // a register), and the two mixed words lie 397..624 back, i.e. at least one whole round behind even for
    // the second of two consecutive rounds - so two rounds are produced per barrier.
    const bool act = tid < 227;                              // every thread runs the loop: the barrier stays converged
    uint32_t prev = act ? ring[(from - 227 + tid) & 2047] : 0u;
    for (long long base = from; base < total; base += 2 * 227) {
      if (act) {
        const long long nA = base + tid, nB = nA + 227;
        const uint32_t vA = prev ^ mt_g(ring[(nA - MT_N) & 2047], ring[(nA - MT_N + 1) & 2047]);
        const uint32_t vB = vA ^ mt_g(ring[(nB - MT_N) & 2047], ring[(nB - MT_N + 1) & 2047]);
        ring[nA & 2047] = vA; keys[(uint32_t)nA & kmask] = vA;
        ring[nB & 2047] = vB; keys[(uint32_t)nB & kmask] = vB;
        prev = vB;
      }
      __syncthreads();
    }
    return;
  }
  // ---- CTA 1.  One thread consumes the Python stream, one draw per loop iteration (random.sample's pool / set
  // variants and _randbelow's rejection loop unrolled into a flat state machine), from a block of outputs the
  // whole CTA has tempered; when the block runs out the CTA twists the state cooperatively and the thread
  // resumes where it stopped.
  uint32_t* pkey = ring;                     // 624 words
  uint32_t* pout = ring + 1024;              // their tempered outputs
  __shared__ int pool[1024];
  __shared__ uint32_t seen[32];
  __shared__ int more;
  for (int i = tid; i < MT_N; i += GEN_NT) { const uint32_t v = py_state[i]; pkey[i] = v; pout[i] = mt_temper(v); h->py_backup[i] = v; }
  if (tid == 0) h->py_backup[MT_N] = py_state[MT_N];
  for (int i = tid; i < R; i += GEN_NT) pool[i] = i;
  if (tid < 32) seen[tid] = 0u;
  __syncthreads();
  int pos = (int)py_state[MT_N], eb = 0, ei = 0, nt = 0;     // stream offset, episode, draw within it (thread 0)
  int touched[64];
  bool use_pool = true;
  {
    int setsize = 21;
    if (N > 5) {   // setsize += 4 ** ceil(log(N*3, 4))
      int p4 = 1;
      while (p4 < 3 * N) p4 *= 4;
      setsize += p4;
    }
    use_pool = R <= setsize;
  }
  while (true) {
    if (tid == 0) {
      while (eb < B && pos < MT_N) {
        const int n = use_pool ? R - ei : R;
        const uint32_t r = pout[pos++] >> __clz(n);          // getrandbits(n.bit_length())
        if (r >= (uint32_t)n) continue;                      // _randbelow rejects
        int* out = classes + (long long)eb * N;
        if (use_pool) {
          out[ei] = pool[r];
          pool[r] = pool[R - ei - 1];
          if (nt < 64) touched[nt++] = (int)r;
        } else {
          if (seen[r >> 5] & (1u << (r & 31))) continue;     // already selected: draw again
          seen[r >> 5] |= 1u << (r & 31);
          out[ei] = (int)r;
        }
        if (++ei == N) {                                     // episode complete: back to a fresh pool / empty set
          if (use_pool) {
            if (N <= 64) { for (int q = 0; q < nt; ++q) pool[touched[q]] = touched[q]; }
            else { for (int q = 0; q < R; ++q) pool[q] = q; }
          } else {
            for (int q = 0; q < 32; ++q) seen[q] = 0u;
          }
          ei = 0; nt = 0; ++eb;
        }
      }
      more = eb < B ? 1 : 0;
    }
    __syncthreads();
    if (!more) break;
    mt_twist_block(pkey);                                    // block exhausted with episodes left: next 624 words
    for (int i = tid; i < MT_N; i += GEN_NT) pout[i] = mt_temper(pkey[i]);
    pos = 0;
    __syncthreads();
  }
  if (tid == 0) py_state[MT_N] = (uint32_t)pos;
  for (int i = tid; i < MT_N; i += GEN_NT) py_state[i] = pkey[i];
  __syncthreads();
  // permutation length of every job of the walk (episode-major: N class scopes, then the N*Q query
  // permutation); negative = the error the walker reports when it reaches that job
  for (int j = tid; j < B * (N + 1); j += GEN_NT) {
    const int b = j / (N + 1), ci = j % (N + 1);
    int n = N * Q;
    if (ci < N) {
      const int c = classes[(long long)b * N + ci];
      n = scopes[2 * c + 1] - scopes[2 * c];
      if (n < K + Q) n = -SMP_SCOPE_TOO_SMALL;
      else if (n > n_max) n = -SMP_SCOPE_TOO_LARGE;
    }
    jsize[j] = n;
  }
}

// ---- kernel T: transition tables ----------------------------------------------------------------
// For every 64-draw block b of the stream and every Fisher-Yates index s in [1, n_states):
// F[b][s] = index after the block's draws are offered to a permutation currently at index s
// (accept iff (y & mask(s)) <= s, then s -= 1), saturating at 0 = "the permutation ends inside this
// block"; for s <= 64 the end table E[b][s] holds how many of the block's draws that permutation
// consumes (0xFF: it does not end here).  Even blocks also carry F2[b][s] = F[b+1][F[b][s]], the
// index after the PAIR of blocks (0 if the permutation ends inside either).  All of it is independent of
// which job meets the block, so the whole GPU builds the tables in parallel and the serial walker only
// chains one lookup per 128 (or 64) draws.  The tempered draws are stored too, for the walker's exact
// rows and the fill kernel.
constexpr int E_STRIDE = 80;   // bytes per block of E (65 used; keeps the record 16-byte granular)
// One record per block, contiguous in memory so that a run of blocks is ONE bulk copy for the walker:
//   [ F row: n_states x u16 | F2 row (even blocks): n_states x u16 | 64 tempered draws: u32 | E row: 80 x u8 |
//     E2 row (even blocks): 144 x u8 ]
// E2[b][s], s <= 128: draws of the PAIR (b, b+1) consumed by a permutation entering at index s that ends
// inside the pair (0xFF otherwise).
constexpr int E2_STRIDE = 144;
__host__ __device__ inline long long rec_bytes(long long n_states) { return n_states * 4 + TB * 4 + E_STRIDE + E2_STRIDE; }

// index after the 64 draws y[0..64) starting from s0 >= 1; e = draws consumed if it reaches 0 (s0 <= 64).
// ym = the same draws already masked with mask_for(s0): valid for the whole block when the index cannot
// drop to the next smaller mask within 64 draws (s0 > (mask >> 1) + 64), which is most states - their
// step is then a compare and a predicated decrement.
__device__ __forceinline__ int walk_block(const uint32_t* y, const uint32_t* ym, int s0, int& e) {
  int s = s0;
  uint32_t m = mask_for((uint32_t)s);
  int thr = (int)(m >> 1);
  e = 0xFF;
  if (s0 > thr + TB) {
#pragma unroll 16
    for (int k = 0; k < TB; ++k) s -= (ym[k] <= (uint32_t)s) ? 1 : 0;
  } else if (s0 <= TB) {
    for (int k = 0; k < TB; ++k) {
      if (s <= thr) { m >>= 1; thr >>= 1; }
      const uint32_t v = y[k] & m;
      if (s >= 1 && v <= (uint32_t)s) { if (--s == 0) e = k + 1; }
    }
  } else {
#pragma unroll 8
    for (int k = 0; k < TB; ++k) {
      if (s <= thr) { m >>= 1; thr >>= 1; }       // the index shrinks by at most one per draw
      const uint32_t v = y[k] & m;
      s -= (v <= (uint32_t)s) ? 1 : 0;            // s0 > 64 cannot reach 0 inside the block
    }
  }
  return s;
}

// one CTA per pair of blocks (2p, 2p+1)
__global__ void __launch_bounds__(256)
smp_table_kernel(const SmpHdr* __restrict__ h, int ahead, const uint32_t* __restrict__ keys, uint32_t kmask,
                 unsigned char* __restrict__ tab, int bcap, int n_states) {
  extern __shared__ unsigned short fs[];          // [2][n_states]: F of the even / odd block
  __shared__ uint32_t y[2][TB];
  __shared__ unsigned char es[2][TB + 1];         // E of the even / odd block
  __shared__ uint32_t yms[2][17][TB];             // the draws masked to b bits, b = 0..16
  const long long b0 = h->plan[ahead].tab_from + 2LL * blockIdx.x;      // absolute block (even)
  if (b0 >= h->plan[ahead].tab_to) return;
  const long long rb = rec_bytes(n_states);
  unsigned char* rec0 = tab + (b0 % bcap) * rb;                          // ring slot of b0; b0 + 1 follows it (bcap is even)
  if (threadIdx.x < 2 * TB) {
    const int hh = threadIdx.x / TB, k = threadIdx.x % TB;
    const uint32_t v = mt_temper(keys[(uint32_t)((b0 + hh) * TB + k) & kmask]);
    y[hh][k] = v;
    reinterpret_cast<uint32_t*>(rec0 + hh * rb + (long long)n_states * 4)[k] = v;
  }
  __syncthreads();
  for (int q = threadIdx.x; q < 2 * 17 * TB; q += 256) {      // draws masked to 0..16 bits
    const int h = q / (17 * TB), bits = (q / TB) % 17, k = q % TB;
    yms[h][bits][k] = y[h][k] & ((1u << bits) - 1u);
  }
  __syncthreads();
  for (int s0 = threadIdx.x; s0 < n_states; s0 += 256) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      int e = 0xFF;
      const int s = (s0 >= 1) ? walk_block(y[h], yms[h][32 - __clz(s0)], s0, e) : 0;
      fs[h * n_states + s0] = (unsigned short)s;
      unsigned char* rec = rec0 + h * rb;
      reinterpret_cast<unsigned short*>(rec)[s0] = (unsigned short)s;
      if (s0 < E_STRIDE) rec[(long long)n_states * 4 + TB * 4 + s0] = (unsigned char)((s0 >= 1 && s0 <= TB) ? e : 0xFF);
      if (s0 <= TB) es[h][s0] = (unsigned char)((s0 >= 1) ? e : 0xFF);
    }
  }
  __syncthreads();
  unsigned short* F2 = reinterpret_cast<unsigned short*>(rec0 + (long long)n_states * 2);
  for (int s0 = threadIdx.x; s0 < n_states; s0 += 256) {
    const int s1 = fs[s0];
    F2[s0] = s1 > 0 ? fs[n_states + s1] : (unsigned short)0;
  }
  unsigned char* E2 = rec0 + (long long)n_states * 4 + TB * 4 + E_STRIDE;
  for (int s0 = threadIdx.x; s0 < E2_STRIDE; s0 += 256) {
    int e2 = 0xFF;
    if (s0 >= 1 && s0 <= 2 * TB && s0 < n_states) {
      const int s1 = fs[s0];
      if (s1 == 0) e2 = (s0 <= TB) ? es[0][s0] : 0xFF;           // ends in the even block (then s0 <= 64)
      else if (fs[n_states + s1] == 0 && s1 <= TB) e2 = TB + es[1][s1];   // ends in the odd block (then s1 <= 64)
    }
    E2[s0] = (unsigned char)e2;
  }
}

// ---- kernel W: the serial walk ------------------------------------------------------------------
// Jobs: for each episode its N class permutations, then the query permutation.  At a 64-aligned stream
// position the walker jumps a pair of blocks through F2 or one block through F (and records (index, job)
// per block for the fill kernel); a permutation that ends inside the block it entered aligned takes its
// end offset from E; everything else (the stretch from a permutation's first draw to the next block
// boundary, and permutations that start and end inside one block) is walked exactly in rows of up to 32
// draws (ballot fixed point over the lanes).
// The lookups form one dependent chain, so their latency is the walk time.  The tables of the next
// blocks are therefore staged in shared memory ahead of the walker: the stream is cut into chunks of
// 2^cb_log blocks whose records are contiguous in global memory, and a ring of rc chunk slots is kept
// full with bulk async copies (one mbarrier per slot, complete_tx).  Warp 0 walks; warp 1 is the
// producer: it watches the walker's chunk counter in shared memory and refills the slots the walker
// has left, so a chunk transition costs the walker one counter store and one (already complete)
// barrier test.  A lookup is one LDS.
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// Shared-memory loads by 32-bit shared address.  Written as asm so the walker's loop-carried index
// stays in ordinary registers: left to itself the compiler proves the whole walk warp-uniform and moves
// it to the uniform datapath, whose longer dependent-issue latency is the wrong trade for this chain.
__device__ __forceinline__ uint32_t lds_u16(uint32_t a) {
  unsigned short v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
  return v;
}

#ifdef INDNET_SMP_TIMELINE
#define SMP_T(var) const long long var = clock64()
#define SMP_ACC(acc, a, b) acc += (b) - (a)
#else
#define SMP_T(var) do { } while (0)
#define SMP_ACC(acc, a, b) do { } while (0)
#endif

__device__ __forceinline__ int lds_acquire_s32(uint32_t a) {
  int v;
  asm volatile("ld.acquire.cta.shared.s32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ void sts_release_s32(uint32_t a, int v) {
  asm volatile("st.release.cta.shared.s32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}

// The walker is one warp on an otherwise idle SM: every instruction costs its full dependent latency, so
// the walk time is (instructions on its path) x (a few cycles).  It therefore keeps nothing but the
// chain: for every run of table jumps it logs ONE record {first block, entry index, job}; the fill kernel
// replays those runs in parallel from the same tables to recover the per-block indices and the j-lists.
__global__ void __launch_bounds__(64)
smp_walk_kernel(const uint32_t* __restrict__ keys, uint32_t kmask, const unsigned char* __restrict__ tab, int ccap,
                int n_states, long long nblk, int cb_log, int rc, uint32_t* __restrict__ np_state,
                uint32_t* __restrict__ py_state, const int* __restrict__ jsize, int jobs, int n_max,
                int* __restrict__ jlist, int4* __restrict__ seclog, SmpHdr* __restrict__ hdr, int* __restrict__ status) {
  extern __shared__ __align__(128) unsigned char ring[];
  if (hdr->use_v2) return;                        // the parallel path serves this call
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int CB = 1 << cb_log;
  const uint32_t f_bytes = (uint32_t)n_states * 2u, y_off = 2u * f_bytes, e_off = y_off + TB * 4u, e2_off = e_off + E_STRIDE;
  const uint32_t blk_bytes = (uint32_t)rec_bytes(n_states), slot_bytes = blk_bytes << cb_log;
  const uint32_t ring_s = tc::smem_u32(ring);
  uint64_t* bars = reinterpret_cast<uint64_t*>(ring + (size_t)rc * slot_bytes);
  const uint32_t bars_s = tc::smem_u32(bars);
  volatile uint32_t* zeros = reinterpret_cast<volatile uint32_t*>(bars + rc);   // [32]
  volatile int* ctl = reinterpret_cast<volatile int*>(bars + rc) + 32;          // [0] walker's chunk - c0, [1] done, [2] landed chunks
  const uint32_t ctl_s = tc::smem_u32(const_cast<int*>(ctl));
  // absolute coordinates (see SmpHdr): this call may consume blocks [win_start, win_start + nblk) only
  const long long win_start = hdr->win_start;
  const long long blk_end = (win_start + nblk + 7) & ~7LL;     // tables exist up to here (whole chunks)
  const int nchunks = (int)(blk_end >> cb_log);
  if (threadIdx.x == 0) {
    for (int r = 0; r < rc; ++r) tc::mbar_init(&bars[r], 1);
    tc::fence_barrier_init();
    ctl[0] = 0;
    ctl[1] = 0;
    ctl[2] = 0;
  }
  // A zero the compiler cannot see through, one per lane: adding it to the walk index keeps the chain
  // in ordinary registers (see lds_u16 above).
  if (warp == 0) zeros[lane] = 0u;
  __syncthreads();
  long long pos = hdr->apos;                      // absolute draw index
  const int c0 = (int)(pos >> (6 + cb_log));

  if (warp == 1) {
    // ---- producer (one lane): keep chunks [walker's chunk, +rc) resident.  It also does the waiting on
    // the copies' mbarriers and publishes the number of landed chunks (ctl[2], release), so the walker's
    // chunk transition is one counter store and one counter load.  It sleeps between polls: a spinning
    // neighbour warp would sit in the same shared-memory pipeline as the walker's lookups.
    if (lane == 0) {
      int issued = c0, landed = c0;
      while (true) {
        const int cur = ctl[0], dn = ctl[1];
        const int hi = min(c0 + cur + rc, nchunks);
        while (issued < hi) {
          const uint32_t r = (uint32_t)(issued - c0) % (uint32_t)rc;
          const uint32_t bar = bars_s + 8u * r;
          asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(slot_bytes) : "memory");
          bulk_g2s(ring_s + r * slot_bytes, tab + (long long)(issued % ccap) * slot_bytes, slot_bytes, bar);
          ++issued;
        }
        while (landed < issued) {                  // in issue order; blocks only while a copy is in flight
          const uint32_t rel = (uint32_t)(landed - c0);
          tc::mbar_wait(&bars[rel % (uint32_t)rc], (rel / (uint32_t)rc) & 1u);
          ++landed;
          sts_release_s32(ctl_s + 8u, landed - c0);
        }
        if (dn) break;
        __nanosleep(200);
      }
    }
    return;
  }

  // ---- walker (warp 0)
  const uint32_t below = (1u << lane) - 1u;
  const int lane_zero = (int)zeros[lane];
  const uint32_t fb = f_bytes, pair_bytes = 2u * blk_bytes;
  const long long clk0 = clock64();
  // the shared window base as a value the compiler has to keep (it would otherwise be re-derived from a
  // special register, at that read's latency, at every chunk transition)
  const uint32_t ring_w = ring_s + (uint32_t)lane_zero, ctl_w = ctl_s + (uint32_t)lane_zero;
  int ready = c0 - 1;                             // last chunk whose slot has been waited for
  int slot_i = -1;                                // its slot: (ready - c0) % rc, advanced incrementally
  uint32_t slot_s = ring_w;                       // shared address of that slot
  int flag = SMP_OK, n_sec = 0, n_row = 0;
#ifdef INDNET_SMP_TIMELINE
  long long a_job = 0, a_head = 0, a_wait = 0, a_jump = 0, a_row = 0, n_iter = 0;
#endif
  int n_next = jobs > 0 ? jsize[0] : 0;
  for (int job = 0; job < jobs && flag == SMP_OK; ++job) {
    SMP_T(tj0);
    const int n = n_next;
    if (n < 0) { flag = -n; break; }
    n_next = (job + 1 < jobs) ? jsize[job + 1] : 0;   // in flight while this job is walked
    int i = n - 1 + lane_zero;
    int* jrow = jlist + (long long)job * n_max;
    SMP_T(tj1);
    SMP_ACC(a_job, tj0, tj1);
    while (i > 0) {
      const long long b = pos >> 6;
      const int off = (int)(pos & (TB - 1));
      if (b >= win_start + nblk) { flag = SMP_STREAM_BUDGET; break; }
      const int c = (int)(b >> cb_log);
      SMP_T(tw0);
      while (ready < c) {                          // entering a new chunk: tell the producer, then make sure it landed
        ++ready;
        const int rel = ready - c0;
        if (lane == 0) asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"(ctl_w), "r"(rel) : "memory");
        while (lds_acquire_s32(ctl_w + 8u) <= rel) { }   // chunks [c0, c0 + ctl[2]) have landed
        if (++slot_i == rc) slot_i = 0;
        slot_s = ring_w + (uint32_t)slot_i * slot_bytes;
      }
      SMP_T(tw1);
      SMP_ACC(a_wait, tw0, tw1);
#ifdef INDNET_SMP_TIMELINE
      ++n_iter;
#endif
      const int lb = (int)(b & (CB - 1));
      if (off == 0) {
        // ---- run of table jumps from block b to the end of this chunk (or of the permutation)
        if (lane == 0) seclog[n_sec] = make_int4((int)(b - win_start), i, job, 0);
        ++n_sec;
        const uint32_t cbase = slot_s;                     // block 0 of this chunk
        uint32_t v = (uint32_t)i;
        int lbb = lb;
        uint32_t eaddr = 0;                                // where the end offset sits if the permutation ends in this run
        int eblk = 0;                                      // ... and the block (within the chunk) it is relative to
        if ((lb & 1) || CB == 1) {                         // odd block (or no pairs): one single-block jump first
          const uint32_t ba = cbase + (uint32_t)lb * blk_bytes;
          const uint32_t nx = lds_u16(ba + 2u * v);
          if (nx == 0) { eaddr = ba + e_off + v; eblk = lb; }
          else { v = nx; ++lbb; }
        }
        SMP_T(tc0);
        if (eaddr == 0 && lbb < CB) {
          if (cb_log == 3) {
            // the remaining pairs as one straight-line chain of F2 lookups entered at the current pair;
            // index 0 is absorbing (F2[b][0] = 0), so an ending permutation arrives as 0 and is located after
            uint32_t v0 = 1, v1 = 1, v2 = 1, v3 = 1, v4;
            const uint32_t t0 = cbase + fb;
            switch (lbb >> 1) {
              case 0: v0 = v;
                      v1 = lds_u16(t0 + 2u * v0);
                      if (false) {
              case 1:   v1 = v; }
                      v2 = lds_u16(t0 + pair_bytes + 2u * v1);
                      if (false) {
              case 2:   v2 = v; }
                      v3 = lds_u16(t0 + 2u * pair_bytes + 2u * v2);
                      if (false) {
              default:  v3 = v; }
                      v4 = lds_u16(t0 + 3u * pair_bytes + 2u * v3);
            }
            const int passed = (v1 != 0) + (v2 != 0) + (v3 != 0) + (v4 != 0);   // = index of the pair it ends in, or 4
            if (passed == 4) { v = v4; lbb = CB; }
            else {
              const uint32_t ve = passed == 0 ? v0 : passed == 1 ? v1 : passed == 2 ? v2 : v3;
              eaddr = cbase + (uint32_t)passed * pair_bytes + e2_off + ve;
              eblk = 2 * passed;
            }
          } else {
            while (lbb < CB) {                             // generic chunk size: the same chain as a loop
              const uint32_t ba = cbase + (uint32_t)lbb * blk_bytes;
              const uint32_t nx = lds_u16(ba + fb + 2u * v);
              if (nx == 0) { eaddr = ba + e2_off + v; eblk = lbb; break; }
              v = nx;
              lbb += 2;
            }
          }
        }
        SMP_T(tc1);
        SMP_ACC(a_head, tc0, tc1);
        const long long chunk_pos = (b - lb) << 6;
        if (eaddr != 0) {                                  // the permutation ends inside this run
          pos = chunk_pos + (long long)(eblk * TB + (int)lds_u8(eaddr));
          i = 0;
        } else {
          pos = chunk_pos + (long long)lbb * TB;
          i = (int)v;
        }
        SMP_T(tx1);
        SMP_ACC(a_jump, tw1, tx1);
        continue;
      }
      // ---- exact row: up to 32 draws, not crossing the next block boundary
      const uint32_t blk_s = slot_s + (uint32_t)lb * blk_bytes;
      const int limit = min(32, TB - off);
      const bool valid = lane < limit;
      const uint32_t y = valid ? lds_u32(blk_s + y_off + 4u * (uint32_t)(off + lane)) : 0u;
      const uint32_t m0 = mask_for((uint32_t)i);
      int cn = (lane * (i + 1)) >> (32 - __clz(m0));       // expected accepts before this lane
      int ip = i - cn;
      uint32_t v = y & mask_for((uint32_t)max(ip, 1));
      uint32_t bal = __ballot_sync(0xffffffffu, valid && ip >= 1 && v <= (uint32_t)ip);
      while (true) {
        cn = __popc(bal & below);
        ip = i - cn;
        v = y & mask_for((uint32_t)max(ip, 1));
        const uint32_t bal2 = __ballot_sync(0xffffffffu, valid && ip >= 1 && v <= (uint32_t)ip);
        if (bal2 == bal) break;
        bal = bal2;
      }
      const int acc = __popc(bal);
      if ((bal >> lane) & 1u) jrow[ip] = (int)v;
      int used = limit;
      if (acc >= i) used = 32 - __clz(bal);                // the permutation ends at its last accepted draw
      i -= acc;
      pos += used;
      ++n_row;
      SMP_T(tr1);
      SMP_ACC(a_row, tw1, tr1);
    }
  }
  if (lane == 0) ctl[1] = 1;                               // the producer drains its copies and leaves
  // hand the stream back: the block the walk stopped in (offset 624 = "block exhausted").  A call that fails
  // (status 1-3) leaves BOTH streams exactly as they were handed in, so the host can retry (e.g. in smaller
  // batches) or raise without having lost draws.
  if (flag != SMP_OK) {
    for (int k = lane; k <= MT_N; k += 32) py_state[k] = hdr->py_backup[k];
    if (lane == 0) {
      hdr->nsec = 0;
      status[0] = flag;
      status[1] = (int)((clock64() - clk0) >> 4);
      status[2] = n_sec;
      status[3] = n_row;
    }
  } else {
    long long blk = pos / MT_N;
    int off = (int)(pos % MT_N);
    if (off == 0 && blk > 0) { blk -= 1; off = MT_N; }
    uint32_t tmp[MT_N / 32 + 1];
    int cnt = 0;
    for (int k = lane; k < MT_N; k += 32) tmp[cnt++] = keys[(uint32_t)(blk * MT_N + k) & kmask];
    cnt = 0;
    for (int k = lane; k < MT_N; k += 32) np_state[k] = tmp[cnt++];
    if (lane == 0) {
      np_state[MT_N] = (uint32_t)off;
      hdr->apos = pos;
      hdr->nsec = n_sec;
      status[0] = flag;
      status[1] = (int)((clock64() - clk0) >> 4);     // diagnostics: walk cycles / 16, jump runs, exact rows
      status[2] = n_sec;
      status[3] = n_row;
#ifdef INDNET_SMP_TIMELINE
      status[4] = (int)(a_job >> 4); status[5] = (int)(a_head >> 4); status[6] = (int)(a_wait >> 4);
      status[7] = (int)(a_jump >> 4); status[8] = (int)(a_row >> 4); status[9] = (int)n_iter;
#endif
    }
  }
}

// ---- kernel F: accepted (index, j) pairs of the blocks the walker jumped over --------------------
// One thread per (logged run, block of its chunk): replay the run's chain through F / F2 in global memory
// up to that block (at most CB / 2 + 1 dependent lookups), then offer the block's 64 draws.
__global__ void smp_fill_kernel(const unsigned char* __restrict__ tab, int bcap, int n_states,
                                const int4* __restrict__ seclog, const SmpHdr* __restrict__ hdr, int cb_log, int n_max,
                                int* __restrict__ jlist) {
  const int CB = 1 << cb_log;
  if (hdr->use_v2) return;
  const long long id = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const int sec = (int)(id >> cb_log), k = (int)(id & (CB - 1));
  if (sec >= hdr->nsec) return;
  const int4 lg = seclog[sec];
  const long long b_start = hdr->win_start + lg.x, chunk0 = b_start & ~(long long)(CB - 1), mine = chunk0 + k;
  if (mine < b_start) return;
  const long long rb = rec_bytes(n_states);
  auto F = [&](long long blk, int s) { return (int)reinterpret_cast<const unsigned short*>(tab + (blk % bcap) * rb)[s]; };
  auto F2 = [&](long long blk, int s) { return (int)reinterpret_cast<const unsigned short*>(tab + (blk % bcap) * rb + (long long)n_states * 2)[s]; };
  int s = lg.y;
  long long bb = b_start;
  bool found = false;
  if ((bb & 1) || CB == 1) {
    if (bb == mine) found = true;
    else {
      s = F(bb, s);
      if (s == 0) return;                       // the permutation ended before this thread's block
      ++bb;
    }
  }
  while (!found) {                              // bb even, bb <= mine < chunk end
    if (bb == mine) { found = true; break; }
    if (bb + 1 == mine) {
      s = F(bb, s);
      if (s == 0) return;
      found = true;
      break;
    }
    s = F2(bb, s);
    if (s == 0) return;
    bb += 2;
  }
  int* jrow = jlist + (long long)lg.z * n_max;
  uint32_t m = mask_for((uint32_t)s);
  int thr = (int)(m >> 1);
  const uint32_t* yp = reinterpret_cast<const uint32_t*>(tab + (mine % bcap) * rb + (long long)n_states * 4);
  for (int q = 0; q < TB && s >= 1; ++q) {
    if (s <= thr) { m >>= 1; thr >>= 1; }
    const uint32_t v = yp[q] & m;
    if (v <= (uint32_t)s) { jrow[s] = (int)v; --s; }
  }
}

// ---- the parallel path ("v2") ---------------------------------------------------------------------
// The serial walk exists because a permutation's first draw is the previous permutation's last draw plus one.
// That chain is made trivially short instead of fast: for every draw position p of the window and every distinct
// permutation length n the whole GPU tabulates G_n[p] = the position after a permutation(n) that starts at p (head
// of the first block walked draw by draw, then one transition-table lookup per block pair, then the end table) -
// all independent of where the jobs really start.  The chain is then ONE dependent table lookup per job
// (smp_chain_kernel), and once every job's start is known the jobs are replayed in parallel, one warp each,
// 32 draws per ballot fixed point straight from the generated stream (smp_replay_kernel).
constexpr uint32_t SMP_NOEND = 0xFFFFFFFFu;

__global__ void smp_jump_kernel(const SmpHdr* __restrict__ h, const unsigned char* __restrict__ tab, int bcap, int n_states,
                                long long nblk, uint32_t* __restrict__ G) {
  if (!h->use_v2) return;
  const long long npos = nblk * TB;
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= npos * h->ntab) return;
  const int z = (int)(idx / npos);
  const long long o = idx - (long long)z * npos;           // offset of the start position from the window base
  const int n = h->sizes[z];
  int s = n - 1;                                           // permutation(n): i = n-1 .. 1
  if (s <= 0) { G[idx] = (uint32_t)o; return; }
  const long long rb = rec_bytes(n_states);
  const long long base_blk = h->win_start;
  long long bo = o >> 6;
  const int off = (int)(o & (TB - 1));
  if (off) {                                               // head: the rest of the first block, draw by draw
    const uint32_t* y = reinterpret_cast<const uint32_t*>(tab + ((base_blk + bo) % bcap) * rb + (long long)n_states * 4);
    uint32_t m = mask_for((uint32_t)s);
    int thr = (int)(m >> 1);
    for (int k = off; k < TB; ++k) {
      if (s <= thr) { m >>= 1; thr >>= 1; }
      if ((y[k] & m) <= (uint32_t)s) {
        if (--s == 0) { G[idx] = (uint32_t)((bo << 6) + k + 1); return; }
      }
    }
    ++bo;
  }
  while (bo < nblk) {
    const long long blk = base_blk + bo;
    const unsigned char* rec = tab + (blk % bcap) * rb;
    if (!(blk & 1) && bo + 1 < nblk) {                     // even block: try the pair table first
      const int s2 = (int)reinterpret_cast<const unsigned short*>(rec + (long long)n_states * 2)[s];
      if (s2 != 0) { s = s2; bo += 2; continue; }
    }
    const int s1 = (int)reinterpret_cast<const unsigned short*>(rec)[s];
    if (s1 == 0) {                                         // ends inside this block (then s <= 64): E[s] draws consumed
      G[idx] = (uint32_t)((bo << 6) + rec[(long long)n_states * 4 + TB * 4 + s]);
      return;
    }
    s = s1;
    ++bo;
  }
  G[idx] = SMP_NOEND;                                      // runs past the generated window
}

// Every relation the same size: an episode is always N class permutations of that length followed by the query
// permutation, so E[p] = G_q(G_c(... G_c(p))) - the draw position after one whole EPISODE that starts at p - is
// again job independent: one thread per position, N + 1 dependent lookups each, all positions in parallel.
__global__ void smp_epi_kernel(const SmpHdr* __restrict__ h, const uint32_t* __restrict__ G, long long nblk, int N,
                               uint32_t* __restrict__ E) {
  if (!h->use_v2 || h->uni_zc < 0) return;
  const long long npos = nblk * TB;
  const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (idx >= npos) return;
  const uint32_t* Gc = G + (long long)h->uni_zc * npos;
  const uint32_t* Gq = G + (long long)h->uni_zq * npos;
  uint32_t o = (uint32_t)idx;
  for (int c = 0; c < N && o != SMP_NOEND; ++c) o = ((long long)o < npos) ? Gc[o] : SMP_NOEND;
  if (o != SMP_NOEND) o = ((long long)o < npos) ? Gq[o] : SMP_NOEND;
  E[idx] = o;
}

// One warp.  General case: lane 0 chains the jobs through the jump tables, one dependent lookup per job.  Uniform
// relation sizes: lane 0 chains the EPISODES through E (N + 1 times fewer dependent lookups), then the lanes expand
// the jobs inside the episodes in parallel.  Then the warp hands the stream back.
__global__ void smp_chain_kernel(SmpHdr* __restrict__ hdr, const uint32_t* __restrict__ keys, uint32_t kmask, long long nblk,
                                 const uint32_t* __restrict__ G, const uint32_t* __restrict__ E,
                                 const int* __restrict__ jsize, int jobs, int N,
                                 uint32_t* __restrict__ start, uint32_t* __restrict__ np_state, uint32_t* __restrict__ py_state,
                                 int* __restrict__ status) {
  if (!hdr->use_v2) return;
  const int lane = threadIdx.x;
  const long long clk0 = clock64();
  const long long base = hdr->win_start << 6, npos = nblk * TB;
  long long pos = hdr->apos;
  int flag = SMP_OK;
  bool uni = hdr->uni_zc >= 0;
  if (uni) {                                               // a job the generator flagged (jsize < 0): the serial chain reports it
    bool bad = false;
    for (int j = lane; j < jobs; j += 32) bad = bad || jsize[j] < 0;
    uni = !__any_sync(0xffffffffu, bad);
  }
  if (uni) {
    const int P = N + 1, B = jobs / P;
    if (lane == 0) {
      uint32_t o = (uint32_t)(pos - base);
      for (int b = 0; b < B; ++b) {
        start[(long long)b * P] = o;
        if ((long long)o >= npos) { flag = SMP_STREAM_BUDGET; break; }
        const uint32_t e = E[o];
        if (e == SMP_NOEND) { flag = SMP_STREAM_BUDGET; break; }
        o = e;
      }
      pos = base + o;
    }
    flag = __shfl_sync(0xffffffffu, flag, 0);
    __syncwarp();
    if (flag == SMP_OK) {
      const uint32_t* Gc = G + (long long)hdr->uni_zc * npos;
      for (int b = lane; b < B; b += 32) {                 // every start inside the episode precedes E[.] < npos
        uint32_t o = start[(long long)b * P];
        for (int c = 1; c < P; ++c) { o = Gc[o]; start[(long long)b * P + c] = o; }
      }
    }
  } else if (lane == 0) {
    uint32_t o = (uint32_t)(pos - base);
    for (int j = 0; j < jobs; ++j) {
      const int n = jsize[j];
      if (n < 0) { flag = -n; break; }
      int z = 0;
      while (z < hdr->ntab && hdr->sizes[z] != n) ++z;
      start[j] = o;
      if ((long long)o >= npos) { flag = SMP_STREAM_BUDGET; break; }
      const uint32_t e = G[(long long)z * npos + o];
      if (e == SMP_NOEND) { flag = SMP_STREAM_BUDGET; break; }
      o = e;
    }
    pos = base + o;
  }
  flag = __shfl_sync(0xffffffffu, flag, 0);
  pos = __shfl_sync(0xffffffffu, pos, 0);
  if (flag != SMP_OK) {                                    // leave both streams exactly as they were handed in
    for (int k = lane; k <= MT_N; k += 32) py_state[k] = hdr->py_backup[k];
    if (lane == 0) { hdr->nsec = -1; status[0] = flag; status[1] = (int)((clock64() - clk0) >> 4); status[2] = 0; status[3] = 0; }
    return;
  }
  long long blk = pos / MT_N;
  int off = (int)(pos % MT_N);
  if (off == 0 && blk > 0) { blk -= 1; off = MT_N; }
  uint32_t tmp[MT_N / 32 + 1];
  int cnt = 0;
  for (int k = lane; k < MT_N; k += 32) tmp[cnt++] = keys[(uint32_t)(blk * MT_N + k) & kmask];
  cnt = 0;
  for (int k = lane; k < MT_N; k += 32) np_state[k] = tmp[cnt++];
  if (lane == 0) {
    np_state[MT_N] = (uint32_t)off;
    hdr->apos = pos;
    hdr->nsec = 0;
    status[0] = SMP_OK;
    status[1] = (int)((clock64() - clk0) >> 4);     // chain cycles / 16
    status[2] = jobs;
    status[3] = 0;
  }
}

// One warp per job: the accepted (index, draw) pairs of its permutation, 32 draws per step.  Lane l of a row holds
// draw p + l; its index is i minus the accepts before it, which depends on the lanes below - a ballot fixed point
// settles the row in a few rounds (the same scheme as the serial walker's exact rows).
__global__ void __launch_bounds__(128)
smp_replay_kernel(const SmpHdr* __restrict__ hdr, const uint32_t* __restrict__ keys, uint32_t kmask,
                  const uint32_t* __restrict__ start, const int* __restrict__ jsize, int jobs, int stride,
                  int* __restrict__ jlist) {
  if (!hdr->use_v2 || hdr->nsec < 0) return;
  const int lane = threadIdx.x & 31;
  const int job = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (job >= jobs) return;
  int i = jsize[job] - 1;
  long long pos = (hdr->win_start << 6) + start[job];
  int* jrow = jlist + (long long)job * stride;
  const uint32_t below = (1u << lane) - 1u;
  while (i > 0) {
    const uint32_t y = mt_temper(keys[(uint32_t)(pos + lane) & kmask]);
    const uint32_t m0 = mask_for((uint32_t)i);
    int cn = (lane * (i + 1)) >> (32 - __clz(m0));         // expected accepts before this lane
    int ip = i - cn;
    uint32_t v = y & mask_for((uint32_t)max(ip, 1));
    uint32_t bal = __ballot_sync(0xffffffffu, ip >= 1 && v <= (uint32_t)ip);
    while (true) {
      cn = __popc(bal & below);
      ip = i - cn;
      v = y & mask_for((uint32_t)max(ip, 1));
      const uint32_t bal2 = __ballot_sync(0xffffffffu, ip >= 1 && v <= (uint32_t)ip);
      if (bal2 == bal) break;
      bal = bal2;
    }
    const int acc = __popc(bal);
    if ((bal >> lane) & 1u) jrow[ip] = (int)v;
    int used = 32;
    if (acc >= i) used = 32 - __clz(bal);                  // the permutation ends at its last accepted draw
    i -= acc;
    pos += used;
  }
}

// Value that ends at position p after the Fisher-Yates swaps (i = n-1..1, swap a[i], a[j_i]) applied
// to the identity: walk the swaps backwards (i ascending) moving the tracked position.
__global__ void __launch_bounds__(128)
sampler_trace_kernel(const int* __restrict__ scopes, const int* __restrict__ classes,
                     const int* __restrict__ jlist, int B, int N, int K, int Q, int n_max,
                     int pick_max, int* __restrict__ picks, int smem_ints) {
  extern __shared__ int sj[];                       // the job's swap list, staged once (every thread walks all of it)
  const int job = blockIdx.x;
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
  const int* jl = jlist + (long long)job * n_max;
  const bool staged = n <= smem_ints;
  if (staged) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) sj[i] = jl[i];
    __syncthreads();
    jl = sj;
  }
  for (int p = threadIdx.x; p < want && p < n; p += blockDim.x) {
    int pos = p;
#pragma unroll 4
    for (int i = 1; i < n; ++i) {
      const int j = jl[i];
      pos = (pos == i) ? j : ((pos == j) ? i : pos);
    }
    picks[(long long)job * pick_max + p] = pos;
  }
}

// rows[b][r]: r < N*K -> support (class-major); else query slot p = r - N*K holding pre-permutation
// query perm_q[p] (class = perm_q[p] / Q).  Then gather the four tables.
__global__ void sampler_gather_kernel(const int* __restrict__ scopes, const int* __restrict__ classes,
                                      const int* __restrict__ picks, int pick_max, int B, int N, int K, int Q,
                                      int L, const int* __restrict__ t_word, const int* __restrict__ t_pos1,
                                      const int* __restrict__ t_pos2, const int* __restrict__ t_mask,
                                      int* __restrict__ rows, int* __restrict__ labels, int* __restrict__ o_word,
                                      int* __restrict__ o_pos1, int* __restrict__ o_pos2, int* __restrict__ o_mask,
                                      int b0) {
  const int n_s = N * (K + Q), NK = N * K;
  const int bl = blockIdx.x / n_s, r = blockIdx.x % n_s;     // bl: episode within the output, b: within the walked batch
  const int b = b0 + bl;
  int ci, slot;
  if (r < NK) { ci = r / K; slot = r % K; }
  else {
    const int pq = picks[((long long)b * (N + 1) + N) * pick_max + (r - NK)];
    ci = pq / Q; slot = K + pq % Q;
    if (threadIdx.x == 0) labels[(long long)bl * (N * Q) + (r - NK)] = ci;
  }
  const int c = classes[(long long)b * N + ci];
  const int row = scopes[2 * c] + picks[((long long)b * (N + 1) + ci) * pick_max + slot];
  if (threadIdx.x == 0) rows[(long long)bl * n_s + r] = row;
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

namespace {
// Scratch layout shared by the walk and gather halves (all offsets in bytes, 256-byte aligned).
struct SmpLayout {
  long long jobs, stride, pick_max, n_states, nblk;
  long long kcap, bcap;            // key ring (words, power of two) and table ring (blocks): two windows + slack
  int cb_log, rc;                  // walker staging: 2^cb_log blocks per chunk, rc chunk slots (-1: does not fit)
  long long smem_bytes;
  long long o_hdr, o_classes, o_jsize, o_jlist, o_picks, o_keys, o_tab, o_seclog, o_jump, o_epi, o_start, sec_cap, bytes;
};
static double expected_draws(long long n) {      // E[draws] of NumPy's permutation(n) (rejection sampling)
  double e = 0.0;
  for (long long i = 1; i < n; ++i) {
    unsigned long long m = (unsigned long long)i;
    m |= m >> 1; m |= m >> 2; m |= m >> 4; m |= m >> 8; m |= m >> 16; m |= m >> 32;
    e += (double)(m + 1) / (double)(i + 1);
  }
  return e;
}
// stream budget of one call: E[draws] * (1 + g_budget_pct) + g_budget_slack (indnet_sampler_set_budget: tests
// shrink it to exercise the "budget exceeded -> streams unchanged -> retry in smaller batches" path)
static int g_allow_v2 = 1;          // indnet_sampler_set_path: 0 forces the serial walker, 2 the per-job chain
static double g_budget_pct = 0.03;
static long long g_budget_slack = 8192;
static SmpLayout smp_layout(int B, int N, int K, int Q, int n_max) {
  SmpLayout L;
  L.jobs = (long long)B * (N + 1);
  L.pick_max = (K + Q > N * Q) ? (K + Q) : (N * Q);
  L.stride = (n_max > N * Q) ? n_max : (N * Q);      // j-list stride also holds the N*Q-entry query permutation
  L.n_states = (L.stride + 7) / 8 * 8;
  // stream budget: expectation + 3 % + slack (std of one 700-permutation is ~21 draws; this is > 10 sigma)
  const double e = (double)B * ((double)N * expected_draws(n_max) + expected_draws((long long)N * Q));
  long long budget = (long long)(e * (1.0 + g_budget_pct)) + g_budget_slack + MT_N;
  if (budget < 4 * MT_N) budget = 4 * MT_N;
  L.nblk = (budget + TB - 1) / TB;
  L.nblk = (L.nblk + 15) / 16 * 16;                      // whole chunks for every chunk size the walker may pick
  L.bcap = 2 * L.nblk + 32;
  L.kcap = 1;
  while (L.kcap < L.bcap * TB + 8 * MT_N) L.kcap <<= 1;
  // walker staging ring: the largest chunk / ring that fits in 200 KB of shared memory
  const long long per_block = rec_bytes(L.n_states);
  // chunk = 8 blocks (or fewer for very wide tables), as many slots as fit (<= 8): a deep ring of small chunks
  // keeps the copies well ahead of a walker that may pass a chunk in a few hundred cycles
  L.cb_log = L.rc = -1;
  L.smem_bytes = 0;
  for (int cb = 3; cb >= 0 && L.cb_log < 0; --cb) {
    for (int rc = 8; rc >= 3; --rc) {
      const long long need = (per_block << cb) * rc + 8 * rc + 128 + 64;
      if (need <= 200 * 1024) { L.cb_log = cb; L.rc = rc; L.smem_bytes = need; break; }
    }
  }
  auto al = [](long long x) { return (x + 255) / 256 * 256; };
  long long o = 0;
  L.o_hdr = o;     o = al(o + (long long)sizeof(SmpHdr));
  L.o_classes = o; o = al(o + 4LL * B * N);
  L.o_jsize = o;   o = al(o + 4LL * L.jobs);
  L.o_jlist = o;   o = al(o + 4LL * L.jobs * L.stride);
  L.o_picks = o;   o = al(o + 4LL * L.jobs * L.pick_max);
  L.o_keys = o;    o = al(o + 4LL * L.kcap);
  L.o_tab = o;     o = al(o + rec_bytes(L.n_states) * L.bcap);
  L.sec_cap = (L.nblk >> (L.cb_log < 0 ? 0 : L.cb_log)) + L.jobs + 2;   // one run per chunk entered aligned + one per job
  L.o_seclog = o;  o = al(o + 16LL * L.sec_cap);
  L.o_start = o;   o = al(o + 4LL * L.jobs);
  L.o_jump = o;    o = al(o + 4LL * 4 * L.nblk * TB);        // G_n[p] for up to four lengths n
  L.o_epi = o;     o = al(o + 4LL * L.nblk * TB);            // E[p]: the same for one whole episode
  L.bytes = o;
  return L;
}
}  // namespace

extern "C" long long indnet_sampler_scratch_bytes(int B, int N, int K, int Q, int n_max) {
  if (B < 0 || N <= 0 || K < 0 || Q < 0 || n_max <= 0) return -1;
  return smp_layout(B, N, K, Q, n_max).bytes;
}

static int sample_walk(cudaStream_t st, unsigned int* py_state, unsigned int* np_state, const int* scopes, int R,
                       int B, int N, int K, int Q, int n_max, void* scratch, int* status) {
  IN_REQUIRE(B >= 0 && N > 0 && K > 0 && Q > 0 && R > 0 && n_max > 0, "sample_episodes: bad sizes");
  IN_REQUIRE(N <= R, "sample_episodes: Sample larger than population (N=%d classes of %d relations)", N, R);
  IN_REQUIRE(R <= 1024, "sample_episodes: at most 1024 relations supported, got %d", R);
  IN_REQUIRE(n_max < 65536 && N * Q < 65536, "sample_episodes: scopes above 65535 instances are not supported");
  if (B == 0) return INDNET_OK;
  IN_REQUIRE(py_state && np_state && scopes && scratch && status, "sample_episodes: null pointer");
  const SmpLayout L = smp_layout(B, N, K, Q, n_max);
  char* base = (char*)scratch;
  int* classes = (int*)(base + L.o_classes);
  int* jlist = (int*)(base + L.o_jlist);
  int* picks = (int*)(base + L.o_picks);
  uint32_t* keys = (uint32_t*)(base + L.o_keys);
  unsigned char* tab = (unsigned char*)(base + L.o_tab);
  int* jsize = (int*)(base + L.o_jsize);
  int4* seclog = (int4*)(base + L.o_seclog);
  SmpHdr* hdr = (SmpHdr*)(base + L.o_hdr);
  IN_REQUIRE(L.cb_log >= 0, "sample_episodes: n_max=%d is too large for the walker's shared-memory table ring", n_max);
  IN_REQUIRE(L.bcap <= 2147483647LL && L.kcap <= 2147483648LL, "sample_episodes: stream window too long");
  // per-device: kernel attributes, and the forked stream + events on which the next window is built
  static bool dev_set[64] = {};
  static cudaStream_t side_s[64] = {};
  static cudaEvent_t ev_fork[64] = {}, ev_join[64] = {};
  int dev = 0;
  IN_CUDA(cudaGetDevice(&dev));
  IN_REQUIRE(dev >= 0 && dev < 64, "sample_episodes: device index %d out of range", dev);
  if (!dev_set[dev]) {
    IN_CUDA(cudaFuncSetAttribute(smp_walk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    IN_CUDA(cudaFuncSetAttribute(smp_table_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    IN_CUDA(cudaStreamCreateWithFlags(&side_s[dev], cudaStreamNonBlocking));
    IN_CUDA(cudaEventCreateWithFlags(&ev_fork[dev], cudaEventDisableTiming));
    IN_CUDA(cudaEventCreateWithFlags(&ev_join[dev], cudaEventDisableTiming));
    dev_set[dev] = true;
  }
  cudaStream_t side = side_s[dev];
  const uint32_t kmask = (uint32_t)(L.kcap - 1);
  const int bcap = (int)L.bcap, ccap = (int)(L.bcap >> L.cb_log);
  unsigned long long magic = 0x534d5048ull;           // layout signature: a header left by other sizes is stale
  for (long long v : {(long long)B, (long long)N, (long long)K, (long long)Q, (long long)n_max, L.nblk, L.n_states})
    magic = magic * 1000003ull + (unsigned long long)v;
  const unsigned tgrid = (unsigned)(L.nblk / 2 + 16);
  const size_t tsmem = 4 * (size_t)L.n_states;
  // what this call needs (normally already there: the previous call built it while it walked) ...
  smp_plan_kernel<<<1, 256, 0, st>>>(hdr, np_state, keys, kmask, L.nblk, 0, magic, scopes, R, N * Q, g_allow_v2);
  IN_LAUNCH_CHECK("smp_plan_kernel");
  smp_gen_kernel<<<2, GEN_NT, 0, st>>>(hdr, 0, keys, kmask, py_state, R, B, N, classes, scopes, K, Q, n_max, jsize);
  IN_LAUNCH_CHECK("smp_gen_kernel");
  smp_table_kernel<<<tgrid, 256, tsmem, st>>>(hdr, 0, keys, kmask, tab, bcap, (int)L.n_states);
  IN_LAUNCH_CHECK("smp_table_kernel");
  // ... and, forked, the window after it
  IN_CUDA(cudaEventRecord(ev_fork[dev], st));
  IN_CUDA(cudaStreamWaitEvent(side, ev_fork[dev], 0));
  smp_plan_kernel<<<1, 256, 0, side>>>(hdr, np_state, keys, kmask, L.nblk, 1, magic, scopes, R, N * Q, g_allow_v2);
  IN_LAUNCH_CHECK("smp_plan_kernel");
  smp_gen_kernel<<<1, GEN_NT, 0, side>>>(hdr, 1, keys, kmask, py_state, R, B, N, classes, scopes, K, Q, n_max, jsize);
  IN_LAUNCH_CHECK("smp_gen_kernel");
  smp_table_kernel<<<tgrid, 256, tsmem, side>>>(hdr, 1, keys, kmask, tab, bcap, (int)L.n_states);
  IN_LAUNCH_CHECK("smp_table_kernel");
  IN_CUDA(cudaEventRecord(ev_join[dev], side));
  {
    uint32_t* G = (uint32_t*)(base + L.o_jump);
    uint32_t* jstart = (uint32_t*)(base + L.o_start);
    const long long work = 4 * L.nblk * TB;                  // upper bound (ntab <= 4): surplus threads leave at once
    smp_jump_kernel<<<(unsigned)((work + 255) / 256), 256, 0, st>>>(hdr, tab, bcap, (int)L.n_states, L.nblk, G);
    IN_LAUNCH_CHECK("smp_jump_kernel");
    uint32_t* E = (uint32_t*)(base + L.o_epi);
    smp_epi_kernel<<<(unsigned)((L.nblk * TB + 255) / 256), 256, 0, st>>>(hdr, G, L.nblk, N, E);
    IN_LAUNCH_CHECK("smp_epi_kernel");
    smp_chain_kernel<<<1, 32, 0, st>>>(hdr, keys, kmask, L.nblk, G, E, jsize, (int)L.jobs, N, jstart, np_state, py_state, status);
    IN_LAUNCH_CHECK("smp_chain_kernel");
    smp_replay_kernel<<<ceil_div(L.jobs, 4), 128, 0, st>>>(hdr, keys, kmask, jstart, jsize, (int)L.jobs, (int)L.stride, jlist);
    IN_LAUNCH_CHECK("smp_replay_kernel");
  }
  smp_walk_kernel<<<1, 64, (size_t)L.smem_bytes, st>>>(keys, kmask, tab, ccap, (int)L.n_states, L.nblk, L.cb_log, L.rc,
                                                      np_state, py_state, jsize, (int)L.jobs, (int)L.stride, jlist, seclog,
                                                      hdr, status);
  IN_LAUNCH_CHECK("smp_walk_kernel");
  smp_fill_kernel<<<ceil_div(L.sec_cap << L.cb_log, 128), 128, 0, st>>>(tab, bcap, (int)L.n_states, seclog, hdr, L.cb_log,
                                                                        (int)L.stride, jlist);
  IN_LAUNCH_CHECK("smp_fill_kernel");
  {
    const int tr_ints = L.stride <= 11 * 1024 ? (int)L.stride : 0;     // <= 44 KB: within the default dynamic limit
    sampler_trace_kernel<<<(unsigned)L.jobs, 128, sizeof(int) * (size_t)tr_ints, st>>>(
        scopes, classes, jlist, B, N, K, Q, (int)L.stride, (int)L.pick_max, picks, tr_ints);
  }
  IN_LAUNCH_CHECK("sampler_trace_kernel");
  IN_CUDA(cudaStreamWaitEvent(st, ev_join[dev], 0));    // join: the next call's plan must see the finished build
  return INDNET_OK;
}

static int sample_gather(cudaStream_t st, const int* scopes, const int* t_word, const int* t_pos1, const int* t_pos2,
                         const int* t_mask, int L, int B, int N, int K, int Q, int n_max, const void* scratch,
                         int* rows, int* labels, int* o_word, int* o_pos1, int* o_pos2, int* o_mask, int b0 = 0,
                         int nb = -1) {
  IN_REQUIRE(B >= 0 && N > 0 && K > 0 && Q > 0 && L > 0 && n_max > 0, "sample_gather: bad sizes");
  if (nb < 0) nb = B;
  IN_REQUIRE(b0 >= 0 && nb >= 0 && b0 + nb <= B, "sample_gather: episode range [%d, %d) outside the batch of %d", b0, b0 + nb, B);
  if (B == 0 || nb == 0) return INDNET_OK;
  IN_REQUIRE(scopes && t_word && t_pos1 && t_pos2 && t_mask && scratch && rows && labels && o_word && o_pos1 &&
             o_pos2 && o_mask, "sample_gather: null pointer");
  const SmpLayout lay = smp_layout(B, N, K, Q, n_max);
  const char* base = (const char*)scratch;
  const int* classes = (const int*)(base + lay.o_classes);
  const int* picks = (const int*)(base + lay.o_picks);
  const int pick_max = (int)lay.pick_max;
  sampler_gather_kernel<<<nb * N * (K + Q), 64, 0, st>>>(scopes, classes, picks, pick_max, B, N, K, Q, L, t_word,
                                                        t_pos1, t_pos2, t_mask, rows, labels, o_word, o_pos1, o_pos2,
                                                        o_mask, b0);
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

extern "C" int indnet_sample_gather_range(indnet_stream st, const int* scopes, const int* t_word, const int* t_pos1,
                                          const int* t_pos2, const int* t_mask, int L, int B, int N, int K, int Q,
                                          int n_max, const void* scratch, int b0, int nb, int* rows, int* labels,
                                          int* o_word, int* o_pos1, int* o_pos2, int* o_mask) {
  return sample_gather((cudaStream_t)st, scopes, t_word, t_pos1, t_pos2, t_mask, L, B, N, K, Q, n_max, scratch, rows,
                       labels, o_word, o_pos1, o_pos2, o_mask, b0, nb);
}

extern "C" int indnet_sampler_set_budget(double pct, long long slack) {
  IN_REQUIRE(pct > -1.0 && slack >= 0, "sampler_set_budget: pct must be > -1 and slack >= 0");
  g_budget_pct = pct;
  g_budget_slack = slack;
  return INDNET_OK;
}

extern "C" int indnet_sampler_set_path(int allow_parallel) {
  g_allow_v2 = (allow_parallel == 2) ? 2 : (allow_parallel ? 1 : 0);
  return INDNET_OK;
}
