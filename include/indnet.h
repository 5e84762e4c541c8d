/* libindnet - C-ABI of the B200-native (sm_100a) Induction-Network episode pipeline.
 *
 * The reference (wws0815/Induction-Network-on-FewRel) has no FFI: its boundary is the Python API of
 * model/layer_module.py, model/graph.py, model/base.py and util/data_loader.py, and everything below
 * that API runs inside TensorFlow ops.  Each entry point here replaces the TF op group named in its
 * comment (file:line in /root/reference).  The Python host (induction_b200.model.*, .util.*) binds
 * these with ctypes; INTEGRATION.md shows the stub a maintainer of the reference would add.
 *
 * Conventions
 *  - Plain pointers and sizes only; every pointer is a DEVICE pointer unless the comment says host.
 *    All matrices are row-major, dense unless a leading dimension is given.
 *  - indnet_stream is a cudaStream_t; all work is enqueued on it, nothing synchronises, nothing
 *    allocates.  Scratch, when an op needs it, is passed in by the caller (the one exception is
 *    indnet_lstm_bwd_static, whose scratch is a fixed array inside the library).
 *  - Return 0 on success; non-zero (INDNET_ERR_*) on error with a message in indnet_last_error()
 *    (thread-local).  Nothing throws across this boundary.
 *  - FP32 entry points are the "parity mode" (1e-5 vs the reference graph); *_tc entry points are the
 *    bf16 tensor-core mode (tcgen05/TMEM, FP32 accumulation, 2e-2).
 *  - Symbols: N classes, K shots, Q queries/class, L sequence length, E word dim, P=5 position dim,
 *    D=E+2P, u LSTM units, H=2u, S NTN slices, X sentences, T=X*L tokens.
 */
#ifndef INDNET_H_
#define INDNET_H_

#ifdef __cplusplus
extern "C" {
#endif

#define INDNET_ABI_VERSION 6

typedef void* indnet_stream; /* cudaStream_t */

/* Host string describing the last error on this thread. */
const char* indnet_last_error(void);
int indnet_abi_version(void);
/* Number of CUDA kernels this library has launched in this process (bench.py's gpu_launches). */
long long indnet_launch_count(void);

/* ---- dense building blocks (FP32 parity mode) -------------------------------------------------
 * C = alpha*op(A)*op(B) + beta*C (+ bias[n]) (act: 0 none, 1 tanh, 2 relu).  Two-level batching:
 * batch index z -> (zo = z / inner, zi = z % inner), operand offset zo*s?o + zi*s?i.  splitk <= 0
 * picks a split automatically (atomic accumulation) when the output is small and K long.
 * Replaces tf.matmul / tf.layers.Dense / tf.tensordot at model/layer_module.py:24,30,45,65,76,132 and
 * the LSTM kernel matmul of model/graph.py:44, plus their autodiff counterparts. */
int indnet_sgemm(indnet_stream st, int transA, int transB, int M, int N, int K, float alpha, const float* A, int lda,
                 const float* B, int ldb, float beta, float* C, int ldc, const float* bias, int act, int batch,
                 int inner, long long sAo, long long sAi, long long sBo, long long sBi, long long sCo, long long sCi,
                 int splitk);
/* out[c] (+)= sum_r A[r, c]  - bias gradients (autodiff of the bias adds at layer_module.py:45,65). */
int indnet_colsum(indnet_stream st, const float* A, int rows, int cols, int lda, float* out, int accumulate);
/* out[d0][d2][d1] = in[d0][d1][d2]  - converts TF's M[H,H,S] (layer_module.py:16) to the internal M'[H,S,H]. */
int indnet_permute021(indnet_stream st, const float* in, float* out, int d0, int d1, int d2);

/* ---- dense building block, bf16 tensor-core mode (tcgen05 + TMEM + TMA) -------------------------
 * C[M,N] = A*B (+ bias[n]) (act 0/1 tanh/2 relu) (accumulate: += C), bf16 operands, fp32 accumulation.
 * a_mn = 0: A stored [M][K] row-major (lda); a_mn = 1: A stored [K][M].  b_mn = 0: B stored [N][K]
 * (i.e. B transposed); b_mn = 1: B stored [K][N].  Leading dimensions in elements, multiples of 8.
 * Output fp32 C and/or bf16 Cb (either may be NULL), same ldc.  batch > 1: operand strides sA/sB/sC
 * (elements).  splitk <= 0: automatic split-K (fp32 atomics) when the output grid is small.
 * Same call sites as indnet_sgemm. */
int indnet_tc_gemm(indnet_stream st, int a_mn, int b_mn, int M, int N, int K, const void* A, long long lda,
                   const void* B, long long ldb, float* C, void* Cb, int ldc, const float* bias, int act,
                   int accumulate, int batch, long long sA, long long sB, long long sC, int splitk);
/* Same GEMM (batch 1, no split-K) with a rank-1 term in the epilogue:
 *   C[m][n] = sum_k A*B + rs[m] * v(m / rv_rows)[n]
 * where v(x) is row x of the pair (rv_a, rv_b), N wide, in the support/query split layout of
 * indnet_attn_pool_bwd (group, split).  This is d(loss)/d(h) of the self-attention block in one pass:
 * the Dense path dpre * W_a1^T plus the pooling path alpha_t * denc (autodiff of
 * model/layer_module.py:44-54), written once (fp32 C and/or bf16 Cb) instead of accumulated in HBM. */
int indnet_tc_gemm_rank1(indnet_stream st, int a_mn, int b_mn, int M, int N, int K, const void* A, long long lda,
                         const void* B, long long ldb, float* C, void* Cb, int ldc, const float* rs,
                         const float* rv_a, const float* rv_b, int rv_rows, int group, int split);
/* Both weight gradients of one LSTM direction in one pass over dz (bf16 mode, hidden_size 128, D <= 64):
 *   dkernel[D+u][4u] = [x | h_prev]^T * dz,  x_bf [T][64], hprev_bf [T][u], dz_bf [T][4u]
 * (autodiff of the LSTM kernel matmul, model/graph.py:44).  Overwrites dkernel (fp32, split-K reductions).
 * dbias (optional, [4u], overwritten): row D of the stacked product, i.e. the bias gradient sum_t dz[t] when
 * x carries 1.0 in column D (indnet_embed_fwd_bf16 writes it when D <= 62).  accumulate != 0: the caller has
 * zeroed dkernel / dbias already (one memset for several gradient slots, off the critical path) and the split-K
 * partial sums are added to what is there. */
int indnet_tc_gemm_lstm_dw(indnet_stream st, const void* x_bf, const void* hprev_bf, const void* dz_bf,
                           float* dkernel, float* dbias, long long T, int D, int u, int splitk, int accumulate);
/* FP32-accurate tensor-core mode (the reference graph is fp32, model/graph.py:47): every fp32 operand is split into three
 * bf16 matrices, x = head + middle + tail (24 mantissa bits), by indnet_cast_bf16_split3 - rows [rows][cols] with input
 * stride ld_in are written with the padded stride ld_out (a multiple of 8, zeros in the padding) - and indnet_tc_gemm3
 * issues the six cross products of weight >= 2^-16 per K step into one fp32 accumulator (tanh in the epilogue is
 * tanhf).  Arguments otherwise as indnet_tc_gemm; fp32 output only.  Same call sites as indnet_sgemm. */
int indnet_cast_bf16_split3(indnet_stream st, const float* in, long long ld_in, long long rows, int cols, int ld_out,
                            void* hi, void* mid, void* lo);
int indnet_tc_gemm3(indnet_stream st, int a_mn, int b_mn, int M, int N, int K, const void* Ah, const void* Am,
                    const void* Al, long long lda, const void* Bh, const void* Bm, const void* Bl, long long ldb,
                    float* C, int ldc, const float* bias, int act, int accumulate, int batch, long long sA,
                    long long sB, long long sC, int splitk);
/* out (bf16) = in (fp32), n elements. */
int indnet_cast_bf16(indnet_stream st, const float* in, void* out, long long n);
/* Row-wise cast with a padded output row stride: out[r][0..ldo) = in[r][0..cols), zeros beyond cols
 * (makes rows 16-byte aligned for TMA when cols*2 is not a multiple of 16). */
int indnet_cast_bf16_rows(indnet_stream st, const float* in, void* out, long long rows, int cols, int ldo);

/* cudaMemsetAsync on the stream (gradient regions that accumulate must start at zero). */
int indnet_memset(indnet_stream st, void* dst, int byte_value, long long bytes);

/* ---- a1: JSONFileDataLoader.next_one_tf, util/data_loader.py:259-309 --------------------------
 * Samples B consecutive episodes and gathers their rows, bit-exact with the reference's draws.
 * py_state / np_state: device uint32[625] = MT19937 key[624] + pos, the layouts of random.getstate()
 * and np.random.get_state(); advanced in place.  scopes [R][2] = rel2scope values in dict order;
 * t_* the four [instances][L] int32 tables; n_max >= the largest scope.  scratch: caller-provided,
 * indnet_sampler_scratch_bytes(...) bytes.  Outputs: rows [B][n_s] (row ids, support class-major then
 * permuted queries), labels [B][N*Q], o_* [B][n_s][L].  status (device int[16]): [0] = 0 ok, 1 = a
 * sampled relation has fewer than K+Q instances, 2 = a scope exceeds n_max, 3 = the pre-generated
 * stream budget (expected draws + 3 % + 8192) was exceeded; [1..3] diagnostics (SM cycles / 16 of the
 * stream walk, table-jump runs, exact rows); [4..9] are written only by the timeline build.
 * scratch is PERSISTENT state: it holds the generated stream and its tables in rings across calls (the call
 * builds the next call's window on a forked internal stream while it walks), keyed to np_state - hand the
 * same scratch to consecutive calls of one sampler; a call whose np_state is not the continuation the scratch
 * expects (first call, re-seeded stream, other sizes) rebuilds from np_state.  The first call on a device
 * creates that internal stream: make it outside stream capture.
 * The NumPy stream is consumed through per-block transition tables (csrc/sampler.cu);
 * n_max bounds the table width (at most ~12k instances per relation fit the walker's shared-memory
 * ring, larger values are refused). */
long long indnet_sampler_scratch_bytes(int B, int N, int K, int Q, int n_max);
int indnet_sample_episodes(indnet_stream st, unsigned int* py_state, unsigned int* np_state, const int* scopes, int R,
                           const int* t_word, const int* t_pos1, const int* t_pos2, const int* t_mask, int L, int B,
                           int N, int K, int Q, int n_max, void* scratch, int* rows, int* labels, int* o_word,
                           int* o_pos1, int* o_pos2, int* o_mask, int* status);

/* The same in two halves so the serial stream walk of step t+1 can overlap the train step t on a
 * side stream: walk = draws -> picks in `scratch`; gather = picks -> rows/labels/gathered tables. */
int indnet_sample_walk(indnet_stream st, unsigned int* py_state, unsigned int* np_state, const int* scopes, int R,
                       int B, int N, int K, int Q, int n_max, void* scratch, int* status);
int indnet_sample_gather(indnet_stream st, const int* scopes, const int* t_word, const int* t_pos1, const int* t_pos2,
                         const int* t_mask, int L, int B, int N, int K, int Q, int n_max, const void* scratch,
                         int* rows, int* labels, int* o_word, int* o_pos1, int* o_pos2, int* o_mask);
/* Stream budget of one call = E[draws] * (1 + pct) + slack (defaults 0.03 / 8192, > 40 sigma).  Process-wide;
 * changes the scratch layout, so set it before the first indnet_sampler_scratch_bytes.  Used by the tests to
 * force status 3. */
int indnet_sampler_set_budget(double pct, long long slack);
/* 1 (default): the parallel path - "where does a permutation that starts at draw p end" is tabulated for every draw of
 * the window by the whole GPU, the serial part is one table lookup per permutation, the permutations are then replayed
 * in parallel (needs at most three distinct relation sizes, else the serial walker runs); when every relation has the
 * same size (FewRel) the lookups are composed into one per EPISODE first, so the serial part is B lookups instead of
 * B*(N+1).  2: the parallel path without that composition.  0: always the serial walker.
 * Same draws, same results either way.  Process-wide. */
int indnet_sampler_set_path(int allow_parallel);
/* Episode-sharded data parallelism: every rank walks the SAME streams for the global batch of B
 * episodes and gathers only its own episodes [b0, b0 + nb); outputs are [nb][...].  Episode b of the global
 * batch is therefore bit-identical whatever the number of ranks. */
int indnet_sample_gather_range(indnet_stream st, const int* scopes, const int* t_word, const int* t_pos1,
                               const int* t_pos2, const int* t_mask, int L, int B, int N, int K, int Q, int n_max,
                               const void* scratch, int b0, int nb, int* rows, int* labels, int* o_word, int* o_pos1,
                               int* o_pos2, int* o_mask);

/* ---- a2: Base.get_embedding, model/base.py:76-86 --------------------------------------------- */
int indnet_embed_fwd(indnet_stream st, const int* word, const int* pos1, const int* pos2, const float* word_table,
                     const float* pos1_table, const float* pos2_table, float* x /*[T][D]*/, long long T, int E, int P);
/* bf16 tensor-core mode: same lookup written as bf16 rows of 64 columns (xb [T][64]), the K-major operand
 * of the input-projection MMAs: columns [0, D) the embedding, columns D and D+1 = 1.0 when D <= 62 (they
 * multiply the bias rows of the packed LSTM weights and make row D of the weight-gradient GEMM the bias
 * gradient), zeros after.  Requires E + 2P <= 64. */
int indnet_embed_fwd_bf16(indnet_stream st, const int* word, const int* pos1, const int* pos2,
                          const float* word_table, const float* pos1_table, const float* pos2_table, void* xb,
                          long long T, int E, int P);
/* Scatter-add dx into dpos1/dpos2 [pos_rows][P] (and dword_table [vocab][E] when not NULL, i.e. when
 * the word table is trainable, base.py:62-65).  Accumulates (+=). */
int indnet_embed_bwd(indnet_stream st, const int* word, const int* pos1, const int* pos2, const float* dx,
                     float* dword_table, float* dpos1, float* dpos2, long long T, int E, int P, int pos_rows);
/* Position-table gradients from a split input gradient: g[t][j] = dx_a[t*ld + col0 + j] (+ dx_b[...] when
 * given), j < 2P (pos1 columns then pos2 columns).  Same scatter as indnet_embed_bwd. */
int indnet_embed_pos_bwd2(indnet_stream st, const int* pos1, const int* pos2, const float* dx_a, const float* dx_b,
                          int ld, int col0, float* dpos1, float* dpos2, long long T, int P, int pos_rows);

/* ---- a3: rnn.BasicLSTMCell x2 + bidirectional_dynamic_rnn, model/graph.py:34-49 ---------------
 * xproj [2][X][L][4u] = x*W_x + bias per direction (gate order i,j,f,o); Wh_* [u][4u] (row stride 4u:
 * the last u rows of the TF kernel [D+u,4u]).  Writes hout [X][L][2u] (fw | bw, output dropout applied
 * when keep_prob < 1), and for BPTT: gates [2][X][L][4u] (post-activation), cstate [2][X][L][u],
 * hprev [2][X][L][u] (recurrent input of each step).  Dropout stream = seed + *seed_step * 1000003
 * (seed_step: optional device int64, e.g. global_step, so a replayed CUDA graph draws fresh masks). */
int indnet_lstm_fwd(indnet_stream st, const float* xproj, const float* Wh_fw, const float* Wh_bw, float* hout,
                    float* gates, float* cstate, float* hprev, int X, int L, int u, float keep_prob,
                    unsigned long long seed, const long long* seed_step);
/* BPTT: dz [2][X][L][4u] = dLoss/d(pre-activation gates).  whT_scratch: caller-provided [2][4u][u]
 * floats (transposed recurrent weights, rebuilt on every call). */
int indnet_lstm_bwd(indnet_stream st, const float* dhout, const float* gates, const float* cstate, const float* Wh_fw,
                    const float* Wh_bw, float* whT_scratch, float* dz, int X, int L, int u, float keep_prob,
                    unsigned long long seed, const long long* seed_step);

/* The weight packing step of indnet_lstm_fwd_tc on its own (depends on the weights only): a caller may run it beside
 * the embedding gather and then pass NULL for the four weight pointers of indnet_lstm_fwd_tc. */
int indnet_lstm_pack_tc(indnet_stream st, const float* kernel_fw, const float* bias_fw, const float* kernel_bw,
                        const float* bias_bw, void* wt_scratch, int D, int u);
/* bf16 tensor-core mode of indnet_lstm_fwd with the input projection AND the bias fused into the MMAs
 * (hidden_size 128, E+10 <= 62): x_bf [X][L][64] bf16 embeddings with 1.0 in columns D and D+1
 * (indnet_embed_fwd_bf16), kernel_* the TF LSTM kernels [D+u][4u] fp32, bias_* [4u].  wt_scratch:
 * caller-provided bf16 [2][4u][192] (packed transposed weights incl. the bias rows, rebuilt every call; with the
 * four weight pointers NULL it is taken as already packed by indnet_lstm_pack_tc).
 * hout (fp32 copy) must be NULL.  Writes hout_bf [X][L][2u] bf16 and, when given (training; pass all three
 * or none), for BPTT: gates_bf [2][X][L][u][4] bf16 (i, j, f, o of a unit adjacent), cstate_bf [2][X][L][u]
 * bf16 (the recurrence itself carries c in fp32 registers), hprev_bf [2][X][L][u] bf16.  All buffers
 * 16-byte aligned (they are moved as TMA tiles). */
int indnet_lstm_fwd_tc(indnet_stream st, const void* x_bf, const float* kernel_fw, const float* bias_fw,
                       const float* kernel_bw, const float* bias_bw, void* wt_scratch, float* hout, void* hout_bf,
                       void* gates_bf, void* cstate_bf, void* hprev_bf, int X, int L, int D, int u, float keep_prob,
                       unsigned long long seed, const long long* seed_step);

/* hidden_size 256 (BASELINE configs[4], the encoder + induction inference sweep): forward only, a cluster of four
 * CTAs per 32- or 64-sentence tile, each holding a quarter of the gate rows in tensor memory and exchanging h
 * through distributed shared memory (csrc/tc_lstm256.cu).  Same inputs as indnet_lstm_fwd_tc; wt_scratch bf16
 * [2][4u][320]; writes hout_bf [X][L][2u] bf16 only (eval: no dropout, no BPTT streams).  E+10 <= 62. */
int indnet_lstm256_fwd_tc(indnet_stream st, const void* x_bf, const float* kernel_fw, const float* bias_fw,
                          const float* kernel_bw, const float* bias_bw, void* wt_scratch, void* hout_bf, int X, int L,
                          int D, int u);

/* bf16 tensor-core mode of indnet_lstm_bwd (hidden_size 128): gates_bf [2][X][L][u][4] and cstate_bf as
 * saved by indnet_lstm_fwd_tc; kernel_*_bf = bf16 copies of the TF LSTM kernels [D+u][4u] (W_h = rows
 * D.., staged into tensor memory, no scratch); dhout_bf [X][L][2u] bf16 (see indnet_tc_gemm_rank1).
 * Writes dz_bf [2][X][L][4u] bf16 (the bias gradients are row D of indnet_tc_gemm_lstm_dw's product: x
 * carries 1.0 in column D).  When dx2 is given
 * (fp32 [2][X][L][16]) the kernel also forms the input gradient of features dx_first..D-1 (at most 16:
 * the 2 x 5 position-embedding columns when the word table is frozen) per direction,
 * dx2[dir][tok][f - dx_first] = sum_k dz[dir][tok][k] * kernel[f][k]; the caller adds the two directions
 * (indnet_embed_pos_bwd2). */
int indnet_lstm_bwd_tc(indnet_stream st, const void* dhout_bf, const void* gates_bf, const void* cstate_bf,
                       const void* kernel_fw_bf, const void* kernel_bw_bf, void* dz_bf, float* dx2, int dx_first, int X,
                       int L, int D, int u, float keep_prob, unsigned long long seed, const long long* seed_step);
/* out[c] = sum_r A[r][c] for a dense bf16 matrix [rows][cols] (bias gradients from the bf16 dz stream). */
int indnet_colsum_bf16(indnet_stream st, const void* A, long long rows, int cols, float* out);

/* No counterpart in the reference.  Same computation as indnet_lstm_bwd for callers that must not
 * allocate device memory (custom-op / plugin callbacks that only get input and output buffers,
 * loops that cannot afford cudaMalloc's device synchronisation): the [2][4u][u] scratch is a fixed
 * 512 KiB __device__ array inside the library, so u <= 128 (larger u is rejected with an error).
 * The first call on each device looks up the array's address - make it before any stream capture.
 * All calls on a GPU share the array: calls that can overlap on the device race; serialise them. */
int indnet_lstm_bwd_static(indnet_stream st, const float* dhout, const float* gates, const float* cstate,
                           const float* Wh_fw, const float* Wh_bw, float* dz, int X, int L, int u, float keep_prob,
                           unsigned long long seed, const long long* seed_step);

/* ---- a4/a5: self_attention, model/layer_module.py:49-98 (pooling part; Dense+tanh is a GEMM) and
 * the support/query split of model/graph.py:57-62 ------------------------------------------------
 * proj [X][L][H] = tanh(hout*Wa1+ba1); mask [X][L] raw {0,1,2,3} (binarised as model/graph.py:52).
 * alphas [X][L].  Sentences come in episodes of `group` rows; the first `split` rows of each episode
 * are written densely to enc_a [X/group][split][H], the others to enc_b [X/group][group-split][H]
 * (split == group: everything to enc_a, enc_b may be NULL). */
int indnet_attn_pool_fwd(indnet_stream st, const float* hout, const float* proj, const int* mask, const float* w_a2,
                         float* enc_a, float* enc_b, float* alphas, int X, int L, int H, int group, int split);
/* Reads denc_a/denc_b in the same split layout.  Writes dhout = alpha_t*denc (caller adds
 * dpre*Wa1^T; the bf16 variant accepts dhout = NULL when that term is fused into
 * indnet_tc_gemm_rank1), dpre [X][L][H], dw2_part [X][H]. */
int indnet_attn_pool_bwd(indnet_stream st, const float* denc_a, const float* denc_b, const float* hout,
                         const float* proj, const float* alphas, const float* w_a2, float* dhout, float* dpre,
                         float* dw2_part, int X, int L, int H, int group, int split);

/* bf16 tensor-core mode variants: hout_bf / proj_bf are the bf16 copies [X][L][H]; dpre_bf is written in
 * bf16 (the operand format of the Dense-gradient GEMMs); dsum_part [X][H] (optional) receives the
 * per-sentence fp32 sums of dpre over time (Dense-bias gradient partials); enc_a_bf / enc_b_bf (optional)
 * receive bf16 copies of the encodings in the same split layout; dw2_sum / dsum_sum [H] (optional, H % 8 == 0)
 * receive the sums over sentences of dw2_part / dsum_part directly (the W_a2 and Dense-bias gradients), in
 * which case the partial buffers may be NULL.  Everything else as above. */
/* The whole attention block in one kernel (csrc/attn_fused.cu): tcgen05 GEMM h*Wa1 with bias + tanh + the dot with
 * w_a2 taken while the accumulator is drained, masked softmax and weighted sum inside the same CTA (a CTA owns
 * floor(128 / L) whole sentences).  wa1_bf = bf16 copy of the Dense kernel [H][H] (TF layout [in][out]); proj_bf
 * (optional) receives tanh(h*Wa1 + ba1) in bf16 for the backward pass.  L <= 128; H a multiple of 64 up to 256, or
 * 512 / 768 / 1024.  Outputs as indnet_attn_pool_fwd_bf16. */
int indnet_attn_fused_fwd(indnet_stream st, const void* hout_bf, const void* wa1_bf, const float* ba1, const float* w_a2,
                          const int* mask, void* proj_bf, float* enc_a, float* enc_b, void* enc_a_bf, void* enc_b_bf,
                          float* alphas, int X, int L, int H, int group, int split);
/* Backward of the attention block up to d(loss)/d(h) in one kernel (H a multiple of 64 up to 256, L <= 128): softmax /
 * weighted-sum / score / tanh backward on the CUDA cores, dpre written in place as the A operand of the tcgen05 GEMM
 * dpre * Wa1^T, the pooling-path term alpha_t * denc added in the epilogue.  Writes dpre_bf [X*L][H] (operand of the
 * Dense-kernel gradient GEMM), dhout_bf [X*L][H], and the W_a2 / Dense-bias gradients dw2_sum / dba_sum [H]
 * (overwritten; accumulate != 0: added to, the caller zeroed them).  Autodiff of model/layer_module.py:65-96. */
int indnet_attn_fused_bwd(indnet_stream st, const float* denc_a, const float* denc_b, const void* hout_bf,
                          const void* proj_bf, const float* alphas, const float* w_a2, const void* wa1_bf, void* dpre_bf,
                          void* dhout_bf, float* dw2_sum, float* dba_sum, int X, int L, int H, int group, int split,
                          int accumulate);
int indnet_attn_pool_fwd_bf16(indnet_stream st, const void* hout_bf, const void* proj_bf, const int* mask,
                              const float* w_a2, float* enc_a, float* enc_b, void* enc_a_bf, void* enc_b_bf,
                              float* alphas, int X, int L, int H, int group, int split);
int indnet_attn_pool_bwd_bf16(indnet_stream st, const float* denc_a, const float* denc_b, const void* hout_bf,
                              const void* proj_bf, const float* alphas, const float* w_a2, float* dhout,
                              void* dpre_bf, float* dw2_part, float* dsum_part, float* dw2_sum, float* dsum_sum,
                              int X, int L, int H, int group, int split);

/* ---- a6/a7: dynamic_routing + squash, model/layer_module.py:101-162 ---------------------------
 * ehat [BC][K][H] = e_s*W_s; b_in [BC][K] or NULL (zeros, model/graph.py:71); cvec [BC][H]. K <= 32.
 * cvec_bf / dehat_bf (optional, may be NULL): bf16 copies of the outputs written in the same pass (operands
 * of the bf16-mode GEMMs that follow). */
int indnet_routing_fwd(indnet_stream st, const float* ehat, const float* b_in, float* cvec, void* cvec_bf, int BC,
                       int K, int H, int iters);
int indnet_routing_bwd(indnet_stream st, const float* ehat, const float* b_in, const float* dcvec, float* dehat,
                       void* dehat_bf, int BC, int K, int H, int iters);
int indnet_squash_fwd(indnet_stream st, const float* v, float* out, int rows, int H);
int indnet_squash_bwd(indnet_stream st, const float* v, const float* dout, float* dv, int rows, int H);

/* ---- a8-a10: neural_tensor_layer head + build_loss + build_predict/accuracy ---------------------
 * model/layer_module.py:35-47, model/graph.py:92-94, model/base.py:101-107.
 * V [rows][N*S] pre-ReLU bilinear products, column c*S+k; w [(k*N+c)][N] in the reference's row order;
 * labels [rows]; probs [rows][N]; loss_part [rows] = sum_c (p-y)^2; correct [rows] = argmax hit.
 * out2 / running / ticket (optional, out2 and ticket together): the block that finishes last also does what
 * indnet_loss_acc_reduce does (same row order, hence the same bits); ticket is a zero-initialised device
 * counter owned by the caller, left at zero again on return.
 * dlogit / dV / dV_bf / ld_bf / scale (optional, training): the head's backward for the fused mean-squared-error loss
 * is local to a row, so the same launch also writes what indnet_ntn_head_bwd (dprobs = NULL, dw = db = NULL) would. */
int indnet_ntn_head_fwd(indnet_stream st, const float* V, const float* w, const float* b, const int* labels,
                        float* probs, float* loss_part, int* correct, int rows, int N, int S, float* out2,
                        float* running, unsigned int* ticket, float* dlogit, float* dV, void* dV_bf, int ld_bf,
                        float scale);
/* out2[0] = loss (mean over rows*N), out2[1] = accuracy.  running (optional, device float[3]): += {loss,
 * accuracy, 1} - the running sums of model/base.py:156-158 kept on the device. */
int indnet_loss_acc_reduce(indnet_stream st, const float* loss_part, const int* correct, float* out2, float* running,
                           int rows, int N);
/* Upstream gradient: dprobs [rows][N] when not NULL, else the fused mean-squared-error gradient
 * scale*2(p-y) with scale = 1/(rows*N).  Overwrites dlogit [rows][N], dV [rows][N*S], dw, db; dV_bf (optional)
 * receives a bf16 copy of dV with rows padded to ld_bf elements (zeros past N*S), the TMA operand format.
 * dw = db = NULL leaves the two parameter gradients to a later indnet_ntn_head_dw (reads V and dlogit only,
 * so it can run beside the GEMMs that consume dV). */
int indnet_ntn_head_bwd(indnet_stream st, const float* V, const float* w, const float* probs, const int* labels,
                        const float* dprobs, float* dlogit, float* dV, void* dV_bf, int ld_bf, float* dw, float* db,
                        int rows, int N, int S, float scale);
int indnet_ntn_head_dw(indnet_stream st, const float* V, const float* dlogit, float* dw, float* db, int rows, int N,
                       int S);

/* ---- a11: build_optimize, model/base.py:94-99 ---------------------------------------------------
 * One fused TF-form Adam update over a flat buffer.  global_step: device int64, read for the staircase
 * decay then incremented; hyper: device float[2] receiving {lr_t, lr}.  grad_scale multiplies g.
 * shadow_bf16: optional bf16 copy of params (same indexing) refreshed by the update. */
int indnet_adam_step(indnet_stream st, float* params, const float* grads, float* m, float* v, long long n,
                     long long* global_step, float* hyper, float lr0, int decay_steps, float decay_rate, float beta1,
                     float beta2, float eps, float grad_scale, void* shadow_bf16);
/* The same update in two halves (a slice whose gradient is final early - M - can be updated underneath the
 * rest of the backward pass): indnet_adam_prep computes hyper = {lr_t, lr} from global_step and advances the
 * counter only when advance != 0; indnet_adam_apply updates n elements starting at the given pointers with the
 * lr_t found in hyper; advance_step (optional): the counter, incremented once by this call - pass it with the
 * step's last slice; enable (optional, device): the call is a no-op when *enable == 0 - for a slice whose update
 * is deferred into the next step's graph, where a replay may find nothing pending.
 * Sequence for one step: prep(0), apply(slice), ..., apply(rest, &global_step). */
int indnet_adam_prep(indnet_stream st, long long* global_step, float* hyper, float lr0, int decay_steps,
                     float decay_rate, float beta1, float beta2, int advance);
int indnet_adam_apply(indnet_stream st, float* params, const float* grads, float* m, float* v, long long n,
                      const float* hyper, float beta1, float beta2, float eps, float grad_scale, void* shadow_bf16,
                      long long* advance_step, const int* enable);

#ifdef __cplusplus
}
#endif
#endif /* INDNET_H_ */
