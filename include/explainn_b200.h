/*
 * explainn_b200.h -- C ABI of libexplainn_b200.so (hand-written sm_100a kernels for the ExplaiNN hot path).
 *
 * The reference (wassermanlab/ExplaiNN) is pure Python on torch.nn and has no FFI of its own
 * (SURVEY.md section 8b); every entry point below therefore cites the reference Python lines whose
 * ATen calls it replaces.  Paths are relative to the reference repository root.
 *
 * Conventions
 *   - All pointers are DEVICE pointers unless stated otherwise; the caller owns every buffer.  No entry
 *     point allocates or frees device memory or keeps a pointer after it returns.
 *   - Every call is asynchronous on `stream` and re-entrant (no global mutable state).
 *   - Exceptions to the two rules above, each documented at its declaration: explainn_p2p_alloc / _free / _open /
 *     _close (CUDA-IPC peer buffers are allocated and mapped here), explainn_p2p_error (synchronous 4-byte read) and
 *     explainn_interp_unit_max_noalloc (scratch in a library-owned static device array shared by all calls on a GPU),
 *     and explainn_fc_stack_bwd, which runs its final-Linear weight gradient on a library-owned side stream per device
 *     (forked from and joined back into `stream` inside the call; created on the first call on a device, so make that
 *     first call outside stream capture; do not run the call from two host threads on one device at the same time).
 *   - Return value 0 = success; otherwise an EXPLAINN_ERR_* code and explainn_last_error() (thread-local)
 *     describes the failure.  No C++ exception crosses this boundary.
 *   - Tensors are dense row-major with the shapes given, fp32 unless stated otherwise.
 *   - U = num_cnns, L = input_length, F = filter_size, S = L-F+1, P = floor((S-7)/7)+1, H = 100, C = num_classes.
 */
#ifndef EXPLAINN_B200_H
#define EXPLAINN_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef __CUDA_RUNTIME_H__
typedef struct CUstream_st *cudaStream_t;
#endif

#define EXPLAINN_B200_ABI_VERSION 1

#define EXPLAINN_OK 0
#define EXPLAINN_ERR_ARG 1
#define EXPLAINN_ERR_CUDA 2
#define EXPLAINN_ERR_UNSUPPORTED 3
#define EXPLAINN_ERR_LIMIT 4

#define EXPLAINN_DTYPE_F32 0
#define EXPLAINN_DTYPE_F16 1
#define EXPLAINN_DTYPE_F32_PREACT 2   /* explainn_unit_activations only: z = BN(conv) before the exponential, fp32 */

#define EXPLAINN_LOSS_BCE_WITH_LOGITS 0
#define EXPLAINN_LOSS_MSE 1

/* ---- runtime --------------------------------------------------------------------------------------- */
const char *explainn_last_error(void);
int explainn_abi_version(void);
int explainn_num_sms(void);

/* ---- input encoding ---------------------------------------------------------------------------------
 * Replaces the U-fold channel tiling `x.repeat(1, num_cnns, 1)` (explainn/models/networks.py:153) and fixes
 * the input contract of explainn/utils/tools.py:136-181 (rows A,C,G,T; unknown bases = all-zero column).
 * x: (B,4,L) one-hot, dtype EXPLAINN_DTYPE_F32/F16, batch stride in elements (pass 4*U*L to read an already
 * tiled (B,4U,L) tensor).  codes: (B, ceil(L/16)) 2-bit codes; nmask: (B, ceil(L/32)) bit set = N column.
 * *n_invalid is incremented by the number of columns that are neither one-hot nor all-zero. */
int explainn_pack_onehot(const void *x, int x_dtype, long long B, int L, long long batch_stride,
                         uint32_t *codes, uint32_t *nmask, int *n_invalid, cudaStream_t stream);

/* Reverse complement on the packed form (A<->T, C<->G, order reversed, N stays N): the `s[::-1, ::-1]` of the
 * strand-aware prediction loop (scripts/fasta2predictions.py:160, scripts/tsv2predictions.py). */
int explainn_revcomp_packed(const uint32_t *codes, const uint32_t *nmask, long long B, int L, uint32_t *codes_rc,
                            uint32_t *nmask_rc, cudaStream_t stream);

/* DNA text -> packed codes on the device: the batched counterpart of `dna_one_hot` (explainn/utils/tools.py:136-181).
 * text: concatenated ASCII bytes on the device, sequence b = text[offsets[b] .. offsets[b+1]) (offsets: B+1 int64 on the
 * device).  Case-insensitive A,C,G,T (and, as in the reference, the digits '0'..'3') -> codes 0..3, anything else -> N.
 * Longer than L: centre crop starting at (len-L)/2; shorter: placed at column (L-len)/2, N elsewhere (tools.py:150-173). */
int explainn_encode_dna(const unsigned char *text, const long long *offsets, long long B, int L, uint32_t *codes,
                        uint32_t *nmask, cudaStream_t stream);

/* Packed codes -> one-hot (B,4,L) of dtype EXPLAINN_DTYPE_F32/F16 (what `dna_one_hot` returns, tools.py:163); N columns are zero. */
int explainn_unpack_onehot(const uint32_t *codes, const uint32_t *nmask, long long B, int L, void *x, int x_dtype,
                           cudaStream_t stream);

/* ---- weight preparation -------------------------------------------------------------------------------
 * Folds BatchNorm1d (explainn/models/networks.py:100; eval: running stats, train: batch stats) into the
 * grouped Conv1d filter (networks.py:98-99) and builds the k-mer lookup tables the scan kernels read.
 * w (U,4,F), bias/gamma/beta/mean/var (U).  Outputs: tables (explainn_conv_tables_floats), wq_t
 * (explainn_conv_wq_floats), shift (U), scale (U, may be NULL). */
long long explainn_conv_tables_floats(int U, int F);
long long explainn_conv_wq_floats(int U, int F);
int explainn_conv_prepare(const float *w, const float *bias, const float *gamma, const float *beta,
                          const float *mean, const float *var, float eps, int U, int F, float *tables,
                          float *wq_t, float *shift, float *scale, cudaStream_t stream);

/* ---- fused conv + BN + exp + maxpool(7,7) forward -----------------------------------------------------
 * Replaces linears[0..3] (explainn/models/networks.py:98-102).  pooled: (B,U,P), the reference's layout. */
int explainn_conv_pool_fwd(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                           const float *tables, const float *wq_t, const float *shift, float *pooled,
                           cudaStream_t stream);

/* Training variant: tables built with gamma=1, beta=0 and the batch statistics, so the scan yields
 * x_hat = (y-mu)/sigma; z = gamma*x_hat+beta.  Also returns x_hat at the window maximum and the
 * first-max index inside the window (MaxPool1d backward routing).  pooled/xhat (B,U,P) fp32, idx (B,U,P) u8. */
int explainn_conv_pool_fwd_train(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                 const float *tables, const float *wq_t, const float *shift,
                                 const float *gamma, const float *beta, float *pooled, float *xhat,
                                 unsigned char *idx, cudaStream_t stream);

/* ---- tensor-core variant of the fused conv forward (tcgen05 / TMEM) ----------------------------------------------
 * Same reference lines as explainn_conv_pool_fwd[_train] (explainn/models/networks.py:98-102,153).  The conv is
 * evaluated as an fp16 GEMM (one-hot operand exact; fp32 weights split into two fp16 pieces, fp32 accumulation in
 * tensor memory), so results agree with the table kernel to fp32 rounding, not bit for bit; the interpretation
 * entry points keep the table kernel.  images: explainn_conv_tc_image_halfs(U,F) fp16 values written by
 * explainn_conv_tc_prepare (arguments as explainn_conv_prepare).  Every filter row is stored times a power of two that
 * brings its largest tap into [128, 256) -- no folded weight can overflow fp16 or lose its low piece to the fp16
 * subnormal range -- and times the sign of sign_of[u] (NULL: +1; training passes the BatchNorm gain so that pooling
 * gamma * x_hat + beta is a maximum of the raw accumulators).  shift: (2U) fp32, the shift added in the epilogue
 * followed by the signed inverse row scale the epilogue multiplies by.  Outputs as in the table variants (B,U,P).
 * Training tie rule: the first maximum of sign(gamma) * x_hat inside the window (gamma == 0: position 0). */
long long explainn_conv_tc_image_halfs(int U, int F);
int explainn_conv_tc_prepare(const float *w, const float *bias, const float *gamma, const float *beta,
                             const float *mean, const float *var, float eps, const float *sign_of, int U, int F,
                             void *images, float *shift, cudaStream_t stream);
int explainn_conv_pool_fwd_tc(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                              const void *images, const float *shift, float *pooled, cudaStream_t stream);
/* Stated-tolerance variant (inference only): the weights enter as ONE fp16 piece (the high half of the split), i.e.
 * each tap carries a relative rounding error of at most 2^-11 (row-scaled, so no overflow / underflow); accumulation
 * stays fp32.  Half the tensor-core work of explainn_conv_pool_fwd_tc.  Tolerance (tests/test_gpu_eval.py): pooled
 * activations within 2e-3 relative of the fp32 path, logits within 2e-3 of the tensor scale. */
int explainn_conv_pool_fwd_tc_fp16(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                   const void *images, const float *shift, float *pooled, cudaStream_t stream);
int explainn_conv_pool_fwd_train_tc(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                    const void *images, const float *shift, const float *gamma, const float *beta,
                                    float *pooled, float *xhat, unsigned char *idx, cudaStream_t stream);

/* ---- activation scan ------------------------------------------------------------------------------------
 * Replaces `model.linears[:3](seq.repeat(...))` + `.astype(np.half)`
 * (explainn/interpretation/interpretation.py:143-147).  act: (B,U,S) in act_dtype; EXPLAINN_DTYPE_F32_PREACT writes
 * the pre-activation z instead (input of the general max-pool path below). */
int explainn_unit_activations(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                              const float *tables, const float *wq_t, const float *shift, void *act,
                              int act_dtype, cudaStream_t stream);

/* ---- interpretation without the activation tensor -------------------------------------------------------
 * Pass A: per-unit maximum pre-activation over all sequences/positions, the quantity behind
 * `np.amax(activations, axis=(0, 2))` (interpretation.py:187).  scratch: explainn_interp_unit_max_scratch_floats(U)
 * floats of caller-owned workspace; zmax (U); accumulate != 0 folds into the existing zmax (chunked input). */
long long explainn_interp_unit_max_scratch_floats(int U);
int explainn_interp_unit_max(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                             const float *tables, const float *wq_t, const float *shift, float *scratch,
                             float *zmax, int accumulate, cudaStream_t stream);
/* Same result as explainn_interp_unit_max, for callers that must not allocate device memory (a serving process
 * with a fully pre-partitioned GPU, a plugin inside another framework's stream, a thread that may not take the
 * allocator lock): the partial maxima live in a fixed 640 KB array inside the library image instead of
 * caller scratch.  This entry point has no counterpart in the reference.  Limits: U <= 1024 and a GPU with
 * at most 160 SMs, else EXPLAINN_ERR_LIMIT.  The first call on each device looks the array's device address up
 * (cudaGetSymbolAddress, no allocation); make that first call outside stream capture.  All calls on a GPU share
 * the array: each call resets the part it uses, calls that overlap on the GPU race and must be serialised. */
int explainn_interp_unit_max_noalloc(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                     const float *tables, const float *wq_t, const float *shift, float *zmax,
                                     int accumulate, cudaStream_t stream);
/* float16 threshold 0.5*max (interpretation.py:187) and its pre-activation bracket.  All outputs (U). */
int explainn_interp_thresholds(const float *zmax, int U, float *thr, float *zlo, float *zhi, float *amax,
                               cudaStream_t stream);
/* PWM scoring, `networks.PWM.forward` (explainn/models/networks.py:428-437): out[b][u] = max (use_max=1) or sum
 * (use_max=0) over the S valid positions of the grouped conv output z[b][u][s] (tables/wq_t/shift from
 * explainn_conv_prepare with identity BatchNorm).  accumulate=1 folds the result into `out` (max / +), which is how
 * the reverse strand (explainn_revcomp_packed) joins: torch.cat((o, o_rev), 2) then max / sum.  out: (B,U) fp32.
 * use_max=2 writes exp(max z): the unit output of the num_fc = 0 model, conv -> BN -> exp -> MaxPool1d(S)
 * (networks.py:73-80), with tables built from the model's BatchNorm. */
int explainn_pwm_score(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                       const float *tables, const float *wq_t, const float *shift, int use_max, int accumulate,
                       float *out, cudaStream_t stream);

/* Pass B: counts[u,c,k] += #positions with half(act) > thr[u] whose window has base c at offset k
 * (interpretation.py:203-214); hits[u] += number of such positions.  counts (U,4,F) u64, hits (U) u64,
 * both accumulated into (zero them first).  seq_hit (B,U) u8, optional (NULL to skip): 1 if the sequence has any
 * position above threshold for the unit -- the per-unit sequence sets of get_specific_unit_importance
 * (interpretation.py:311-317).  Three kernels: the scan records one hit bit per position, then the windows are counted
 * with bit-parallel population counts (cost independent of the number of hits).  scratch:
 * explainn_interp_pwm_count_scratch_bytes(B, L, U, F) bytes of caller-owned workspace (hit bits, base masks, flags). */
long long explainn_interp_pwm_count_scratch_bytes(long long B, int L, int U, int F);
int explainn_interp_pwm_count(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                              const float *tables, const float *wq_t, const float *shift, const float *zlo,
                              const float *zhi, const float *thr, unsigned long long *counts,
                              unsigned long long *hits, unsigned char *seq_hit, void *scratch, cudaStream_t stream);

/* ---- the two interpretation passes on the tensor cores, same results --------------------------------------------------
 * Same reference lines and the SAME outputs, bit for bit, as explainn_interp_unit_max / explainn_interp_pwm_count
 * (interpretation.py:187,203-214).  The tcgen05 GEMM of explainn_conv_pool_fwd_tc (images / shift from
 * explainn_conv_tc_prepare with the model's eval BatchNorm, sign_of = NULL) gives every pre-activation to within delta[u]
 * (explainn_interp_tc_bounds: w (U,4,F) raw filter, scale / shift (U) from explainn_conv_prepare); positions further than
 * that from a decision (the running maximum in pass A, the float16 threshold bracket in pass B) are decided by the GEMM,
 * the few others are re-evaluated in the canonical fp32 order (wq_t / shift of explainn_conv_prepare) by a resolve kernel.
 *   zmax (U): in / out, the exact maximum so far (start at -inf); accumulates over chunks.
 *   list_scratch: explainn_interp_tc_list_bytes(U) bytes (or more) of caller workspace for the positions to resolve;
 *   *status (device int, zero it once): bit 0 / bit 1 set if pass A / pass B had more such positions than the list
 *   holds -- the results of that call are then incomplete and the caller must redo it with the table entry points.
 *   scratch (pass B): explainn_interp_pwm_count_scratch_bytes(B, L, U, F) bytes, as for explainn_interp_pwm_count. */
int explainn_interp_tc_bounds(const float *w, const float *scale, const float *shift, int U, int F, float *delta,
                              cudaStream_t stream);
long long explainn_interp_tc_list_bytes(int U);
int explainn_interp_unit_max_tc(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                const void *images, const float *shift_rs, const float *delta, const float *wq_t,
                                const float *shift, float *zmax, void *list_scratch, long long list_bytes, int *status,
                                cudaStream_t stream);
int explainn_interp_pwm_count_tc(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                 const void *images, const float *shift_rs, const float *delta, const float *wq_t,
                                 const float *shift, const float *zlo, const float *zhi, const float *thr,
                                 unsigned long long *counts, unsigned long long *hits, unsigned char *seq_hit, void *scratch,
                                 void *list_scratch, long long list_bytes, int *status, cudaStream_t stream);

/* Per (sequence, unit): the maximum activation exp(z) (float32) and the FIRST position that attains it, the rule of
 * `torch.max(activations, dim=2)` (explainn/models/networks.py:228,281; the reference computes it for its ConvNet
 * baselines, BASELINE.json's north star asks for it from the ExplaiNN scan).  amax (B,U) fp32, apos (B,U) int32,
 * bit-identical to oracle/explainn_oracle.c::orc_seq_argmax. */
int explainn_interp_seq_argmax(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                               const float *tables, const float *wq_t, const float *shift, float *amax, int *apos,
                               cudaStream_t stream);

/* ---- threshold / count stage on activations the caller already holds ------------------------------------------
 * `get_pwms_explainn(activations, sequences, filter_size)` (explainn/interpretation/interpretation.py:176-227) with
 * the (N,U,S) activation array of get_explainn_unit_activations (:126-149) as input; act_dtype EXPLAINN_DTYPE_F16
 * (the reference flow) or _F32.  Nothing is recomputed, so the counts are the reference's integers for that array.
 *   explainn_act_unit_max:   folds one chunk of activations into `state` (2*U uint32, zero-initialised by the caller
 *                            before the first chunk): np.amax(activations, axis=(0, 2)) incl. its NaN propagation (:187)
 *   explainn_act_thresholds: amax (U) and thr (U) = 0.5 * amax, both in the activation dtype (:187)
 *   explainn_act_pwm_count:  counts (U,4,F) u64 / hits (U) u64 += windows at positions with act > thr (:203-214);
 *                            seq_hit (N,U) u8 optional (NULL to skip), as in explainn_interp_pwm_count. */
int explainn_act_unit_max(const void *act, int act_dtype, long long N, int U, int S, uint32_t *state, cudaStream_t stream);
int explainn_act_thresholds(const uint32_t *state, int U, int act_dtype, void *amax, void *thr, cudaStream_t stream);
int explainn_act_pwm_count(const void *act, int act_dtype, const uint32_t *codes, const uint32_t *nmask, long long N, int L,
                           int U, int F, const void *thr, unsigned long long *counts, unsigned long long *hits,
                           unsigned char *seq_hit, cudaStream_t stream);

/* ---- per-unit FC stack, eval mode -----------------------------------------------------------------------
 * Replaces linears[4..13] with BatchNorm in eval mode (explainn/models/networks.py:103-116): grouped 1x1
 * Conv1d(U*P->100U) -> BN -> ReLU -> (Dropout = identity) -> grouped 1x1 Conv1d(100U->U) -> BN -> ReLU.
 * pooled (B,U,P); W1 = linears.6.weight (100U,P); b1 (100U); bn2_* (100U); W2 = linears.10.weight (U,100);
 * b2 (U); bn3_* (U).  unit_out: (B,U), the tensor get_explainn_unit_outputs returns
 * (explainn/interpretation/interpretation.py:82-99). */
int explainn_fc_stack_fwd_eval(const float *pooled, long long B, int U, int P, const float *W1, const float *b1,
                               const float *bn2_w, const float *bn2_b, const float *bn2_rm, const float *bn2_rv,
                               const float *W2, const float *b2, const float *bn3_w, const float *bn3_b,
                               const float *bn3_rm, const float *bn3_rv, float eps, float *unit_out,
                               cudaStream_t stream);

/* ---- num_fc = 1 unit (explainn/models/networks.py:82-95): pooled -> grouped Conv1d(U*P -> U, k=1) -> BatchNorm1d(U) ->
 * ReLU.  W (U,P) = linears.6.weight, b1 (U); bn_* = linears.7.  Eval: running statistics, unit_out (B,U).
 * Train: batch statistics (running buffers updated, may be NULL), h2 (U,B) pre-BatchNorm values, part3 (U, fc_nchunk, 2)
 * scratch, mean3/invstd3 (U), unit (B,U), logits (B,C) = unit @ Wf^T + bf (networks.py:147,161).
 * Backward: glogits (B,C) -> all parameter gradients and gpool (B,U,P) = d loss / d pooled for explainn_conv_pool_bwd;
 * gh2 (U,B) and dW_partials (fc_nchunk, U, P) are scratch; db1 is written as exact zeros (bias in front of a
 * train-mode BatchNorm).  Wf == NULL: glogits is (B,U), the gradient w.r.t. the unit outputs. */
int explainn_fc1_stack_fwd_eval(const float *pooled, long long B, int U, int P, const float *W, const float *b1,
                                const float *bn_w, const float *bn_b, const float *bn_rm, const float *bn_rv, float eps,
                                float *unit_out, cudaStream_t stream);
int explainn_fc1_stack_fwd_train(const float *pooled, long long B, int U, int P, int C, const float *W, const float *b1,
                                 const float *bn_w, const float *bn_b, float *bn_rm, float *bn_rv, float momentum,
                                 float eps, const float *Wf, const float *bf, float *h2, float *part3, float *mean3,
                                 float *invstd3, float *unit, float *logits, cudaStream_t stream);
int explainn_fc1_stack_bwd(const float *glogits, long long B, int U, int P, int C, const float *pooled, const float *h2,
                           const float *unit, const float *W, const float *bn_w, const float *mean3, const float *invstd3,
                           const float *Wf, float *gh2, float *dW_partials, float *gpool, float *dW, float *db1,
                           float *dbn_w, float *dbn_b, float *dWf, float *dbf, cudaStream_t stream);

/* ---- final linear ------------------------------------------------------------------------------------------
 * out (B,C) = x (B,U) @ W(C,U)^T + bias (explainn/models/networks.py:147,161). */
int explainn_linear_fwd(const float *x, long long B, int U, int C, const float *W, const float *bias, float *out,
                        cudaStream_t stream);

/* ==== training ================================================================================================ */

/* ---- input co-occurrence statistics -------------------------------------------------------------------------
 * M[i][j] = sum_{b,s} x[b,c,s+k] x[b,c',s+k'], i = c*F+k, j = c'*F+k' (int32, (4F,4F), zeroed by the call).
 * Because the conv input is one-hot, the train-mode BatchNorm1d statistics of the conv output
 * (explainn/models/networks.py:100) and the dense part of its backward follow from M alone. */
int explainn_input_cooc_stats(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int F, int *M,
                              cudaStream_t stream);
/* Batch mean / biased variance of conv(x)+bias per unit from M; updates running stats like BatchNorm1d in
 * train mode (momentum, unbiased variance; networks.py:100) when run_mean != NULL, and *num_batches_tracked += 1
 * when non-NULL.  n_pos = B*S.  Mw (U,4F) = M w_u is kept for the backward. */
int explainn_bn1_batch_stats(const int *M, const float *w, const float *bias, int U, int F, long long n_pos,
                             float momentum, float *run_mean, float *run_var, long long *num_batches_tracked,
                             float *mean, float *var, float *Mw, cudaStream_t stream);

/* ---- FC stack + final linear, train mode --------------------------------------------------------------------
 * Replaces linears[4..13] + final (networks.py:103-116,147) with BatchNorm batch statistics (running stats
 * updated in place when the rm/rv pointers are non-NULL) and inverted dropout (p = pdrop, mask derived from
 * `seed` (+ the device counter *seed_dev when non-NULL, so a captured CUDA graph draws a new mask on every
 * replay) and the element index; the keep bits are also written to dropmask (U,B,4) u32 (may be NULL, then the
 * backward regenerates them from the seed).  Saved for the backward: h1 (U,B,100),
 * mean2/invstd2 (100U), h2 (U,B), mean3/invstd3 (U), unit (B,U).  part2: (U, nchunk, 2, 100) and part3:
 * (U, nchunk, 2) floats of scratch, nchunk = explainn_fc_nchunk(B).  logits: (B,C). */
int explainn_fc_nchunk(long long B);
int explainn_fc_stack_fwd_train(const float *pooled, long long B, int U, int P, int C, const float *W1, const float *b1,
                                const float *bn2_w, const float *bn2_b, float *bn2_rm, float *bn2_rv, const float *W2,
                                const float *b2, const float *bn3_w, const float *bn3_b, float *bn3_rm, float *bn3_rv,
                                float momentum, float eps, float pdrop, unsigned long long seed,
                                const unsigned long long *seed_dev, uint32_t *dropmask, const float *Wf, const float *bf, float *h1, float *part2, float *mean2, float *invstd2, float *h2,
                                float *part3, float *mean3, float *invstd3, float *unit, float *logits,
                                cudaStream_t stream);

/* ---- backward of FC stack + final linear (autograd of networks.py:103-116,147; explainn/train/train.py:65) ---
 * glogits (B,C).  Scratch: gh2 (U,B), pb2 (U,nchunk,3,100), dW1_partials (nchunk,U,100,P).  Outputs: gpool (B,U,P)
 * = d loss / d pooled, and the parameter gradients (any of db1..dbf may be NULL to skip it).  The gradients of
 * the biases that feed a train-mode BatchNorm (b1, b2) are written as exact zeros.  Wf == NULL: glogits is the
 * gradient w.r.t. the unit outputs (B,U) (caller differentiates `model.linears(x)`), dWf/dbf are not written. */
int explainn_fc_stack_bwd(const float *glogits, long long B, int U, int P, int C, const float *pooled, const float *h1,
                          const float *h2, const float *unit, const float *W1, const float *bn2_w, const float *bn2_b,
                          const float *mean2, const float *invstd2, const float *W2, const float *bn3_w,
                          const float *mean3, const float *invstd3, const float *Wf, float pdrop,
                          unsigned long long seed, const unsigned long long *seed_dev, const uint32_t *dropmask, float *gh2, float *pb2, float *dW1_partials, float *gpool,
                          float *dW1, float *db1, float *dbn2_w, float *dbn2_b, float *dW2, float *db2, float *dbn3_w,
                          float *dbn3_b, float *dWf, float *dbf, cudaStream_t stream);

/* ---- the same training entry points in separately callable phases (synchronised BatchNorm under data parallel) ------
 * The reference is single-device: its BatchNorm statistics cover the whole batch (networks.py:100,108,114).  To reproduce
 * that with the batch split over ranks, the additive statistics are all-reduced BETWEEN the kernels that produce and
 * consume them; these entry points expose those seams (arguments as in the unphased calls, plus a phase bitmask):
 *   forward   1: FC1 -> part2 | all-reduce SUM part2 | 2: BatchNorm-2 .. FC2 -> part3 | all-reduce SUM part3 | 4: BatchNorm-3 ..
 *             logits; batch_total = samples of all ranks (statistics and the unbiased running variance use it)
 *   backward  1: final-Linear gradients, d/d unit, BatchNorm-3 sums (first 2U floats of pb2) | average them over ranks |
 *             2: BatchNorm-3 backward, FC2 backward -> pb2 | average pb2 | 4: FC1 backward
 *   conv bwd  1: sparse scatter -> partial | average partial | 2: dense correction with M, Mw, mean, var of the whole batch
 *             and n_pos_total = its B*S
 * (means of local sums over ranks divided by the local batch equal global sums divided by the global batch, so the
 * backward kernels keep B = the local batch; the input co-occurrence table M is all-reduced with SUM before
 * explainn_bn1_batch_stats, whose n_pos is then the global B*S.) */
int explainn_fc_stack_fwd_train_phase(int phases, long long batch_total, const float *pooled, long long B, int U, int P, int C,
                                      const float *W1, const float *b1, const float *bn2_w, const float *bn2_b, float *bn2_rm,
                                      float *bn2_rv, const float *W2, const float *b2, const float *bn3_w, const float *bn3_b,
                                      float *bn3_rm, float *bn3_rv, float momentum, float eps, float pdrop, unsigned long long seed,
                                      const unsigned long long *seed_dev, uint32_t *dropmask, const float *Wf, const float *bf,
                                      float *h1, float *part2, float *mean2, float *invstd2, float *h2, float *part3, float *mean3,
                                      float *invstd3, float *unit, float *logits, cudaStream_t stream);
int explainn_fc_stack_bwd_phase(int phases, const float *glogits, long long B, int U, int P, int C, const float *pooled,
                                const float *h1, const float *h2, const float *unit, const float *W1, const float *bn2_w,
                                const float *bn2_b, const float *mean2, const float *invstd2, const float *W2, const float *bn3_w,
                                const float *mean3, const float *invstd3, const float *Wf, float pdrop, unsigned long long seed,
                                const unsigned long long *seed_dev, const uint32_t *dropmask, float *gh2, float *pb2,
                                float *dW1_partials, float *gpool, float *dW1, float *db1, float *dbn2_w, float *dbn2_b, float *dW2,
                                float *db2, float *dbn3_w, float *dbn3_b, float *dWf, float *dbf, cudaStream_t stream);
int explainn_conv_pool_bwd_phase(int phases, long long n_pos_total, const uint32_t *codes, const uint32_t *nmask, long long B, int L,
                                 int U, int F, const float *gpool, const float *pooled, const float *xhat, const unsigned char *idx,
                                 const int *M, const float *Mw, const float *bias, const float *mean, const float *var,
                                 const float *gamma, float eps, float *partial, int nsplit, float *dW, float *dbias, float *dgamma,
                                 float *dbeta, cudaStream_t stream);

/* ---- backward of conv + BN + exp + maxpool (autograd of networks.py:98-102) ----------------------------------
 * Only the pool-argmax positions carry gradient (sparse scatter); the dense BatchNorm correction comes from
 * M / Mw.  partial: (nsplit, U, 4F+2) floats of scratch.  nsplit = explainn_conv_bwd_nsplit(U) (one slice per SM)
 * selects the tensor-core scatter (csrc/conv_bwd_tc.cu: the sum over the argmax positions as a per-sequence GEMM, gz
 * in three bfloat16 pieces written straight into tensor memory) for the default pool when (U*P) % 4 == 0 and
 * gpool / pooled / xhat are 16-byte aligned; any smaller nsplit >= 1, or any other shape, runs the shared-memory
 * scatter kernel (csrc/train.cu).  Both are deterministic; they differ in summation order only.
 * Outputs dW (U,4,F), dbias (U, exact zeros), dgamma (U), dbeta (U); each may be NULL. */
int explainn_conv_bwd_nsplit(int U);
int explainn_conv_pool_bwd(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                           const float *gpool, const float *pooled, const float *xhat, const unsigned char *idx,
                           const int *M, const float *Mw, const float *bias, const float *mean, const float *var,
                           const float *gamma, float eps, float *partial, int nsplit, float *dW, float *dbias,
                           float *dgamma, float *dbeta, cudaStream_t stream);

/* ==== model variants beside the default unit (explainn/models/networks.py:48-49,73-80,102,117-145) ====
 * General MaxPool1d(pool_size, pool_stride) over the pre-activation map z (B,U,S) that explainn_unit_activations writes
 * with EXPLAINN_DTYPE_F32_PREACT (networks.py:102; the global pool MaxPool1d(S) of the num_fc = 0 unit, networks.py:77,
 * is pool_size = pool_stride = S).  gamma/beta NULL (eval: z carries the BatchNorm map): pooled (B,U,P) = exp(max z).
 * gamma/beta given (train: the map holds x_hat): first maximum of gamma*x_hat+beta, pooled = exp of it, xhat (B,U,P) the
 * x_hat there, idx (B,U,P) u16 its offset inside the window (xhat / idx may be NULL). */
int explainn_pool_general(const float *z, long long B, int U, int S, int pool_size, int pool_stride, const float *gamma,
                          const float *beta, float *pooled, float *xhat, unsigned short *idx, cudaStream_t stream);
/* explainn_conv_pool_bwd for any (pool_size, pool_stride) with the 16-bit window offsets of explainn_pool_general. */
int explainn_conv_pool_bwd_general(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                   int pool_size, int pool_stride, const float *gpool, const float *pooled, const float *xhat,
                                   const unsigned short *idx16, const int *M, const float *Mw, const float *bias,
                                   const float *mean, const float *var, const float *gamma, float eps, float *partial,
                                   int nsplit, float *dW, float *dbias, float *dgamma, float *dbeta, cudaStream_t stream);
/* Backward of the final Linear alone (networks.py:147,161; the num_fc = 0 unit has nothing between pool and Linear):
 * dWf (C,U), dbf (C), gunit (B,U) = glogits Wf, zeroed where unit <= 0.  dWf / dbf may be NULL. */
int explainn_linear_bwd(const float *glogits, const float *unit, const float *Wf, long long B, int U, int C, float *gunit,
                        float *dWf, float *dbf, cudaStream_t stream);
/* Hidden block of the num_fc > 2 unit (networks.py:117-138): in X (B,U,Pin) -> grouped Conv1d(Pin*U -> 100U, k=1) ->
 * BatchNorm1d(100U) -> ReLU -> Dropout -> out (B,U,100) (the Dropout is the one the reference places at the head of the
 * NEXT layer).  W (100U,Pin), b (100U), bn_* (100U).  Train: batch statistics (running buffers updated when non-NULL),
 * Y (U,B,100) pre-BatchNorm values and mean / invstd (100U) are kept for the backward; dropout mask from `seed`
 * (+ *seed_dev) and the element index.  Eval: running statistics, no dropout; Y and stat (2*100U) are scratch.
 * Backward: gout (B,U,100) -> gin (B,U,Pin) (NULL to skip), dW, db (exact zeros), dbn_w, dbn_b; sums (2*100U) and
 * G (U,B,100) are scratch.  The last hidden layer and last_linear reuse explainn_fc_stack_* with P = 100. */
int explainn_hidden_fwd_train(const float *X, long long B, int U, int Pin, const float *W, const float *b, const float *bn_w,
                              const float *bn_b, float *bn_rm, float *bn_rv, float momentum, float eps, float pdrop,
                              unsigned long long seed, const unsigned long long *seed_dev, float *Y, float *mean,
                              float *invstd, float *out, cudaStream_t stream);
int explainn_hidden_fwd_eval(const float *X, long long B, int U, int Pin, const float *W, const float *b, const float *bn_w,
                             const float *bn_b, const float *bn_rm, const float *bn_rv, float eps, float *Y, float *stat,
                             float *out, cudaStream_t stream);
int explainn_hidden_bwd(const float *gout, long long B, int U, int Pin, const float *X, const float *Y, const float *W,
                        const float *bn_w, const float *bn_b, const float *mean, const float *invstd, float pdrop,
                        unsigned long long seed, const unsigned long long *seed_dev, float *sums, float *G, float *gin,
                        float *dW, float *db, float *dbn_w, float *dbn_b, cudaStream_t stream);

/* ---- fused criterion (mean reduction) and its gradient ---------------------------------------------------------
 * BCEWithLogitsLoss / MSELoss as scripts/train.py:210-219 constructs them (explainn/train/train.py:62).
 * loss: 1 float; dlogits (B,C) = grad_scale * d loss / d logits (may be NULL).  scratch: explainn_loss_scratch_bytes()
 * bytes, zero-initialised ONCE by the caller (the kernel leaves it zeroed again). */
long long explainn_loss_scratch_bytes(void);
int explainn_loss_fwd_bwd(const float *logits, const float *target, long long B, int C, int kind, float grad_scale,
                          float *loss, float *dlogits, void *scratch, cudaStream_t stream);

/* Adds 1 to each non-NULL int64 device counter in one launch: `num_batches_tracked` of the train-mode BatchNorm1d
 * layers (torch.nn.BatchNorm1d increments it per forward; networks.py:107,112) and the per-replay dropout counter of a
 * captured training step.  The four pointers must be distinct. */
int explainn_counters_inc(long long *c0, long long *c1, long long *c2, long long *c3, cudaStream_t stream);

/* ---- fused Adam over one flat buffer ------------------------------------------------------------------------------
 * torch.optim.Adam with its defaults-compatible options (amsgrad=False, weight_decay=0, maximize=False), the
 * optimizer scripts/train.py:246-247 builds (explainn/train/train.py:67).  *step (device float) is incremented
 * first; grad is multiplied by grad_scale (1/world_size after a SUM all-reduce). */
int explainn_adam_step(float *param, const float *grad, float *exp_avg, float *exp_avg_sq, long long n, float *step,
                       float lr, float beta1, float beta2, float eps, float grad_scale, cudaStream_t stream);

/* Data-parallel optimizer step as ONE kernel over NVLink peer memory (no counterpart in the reference, which is
 * single-device; it replaces "NCCL all-reduce of the gradients + Adam", the DDP form of train.py:65-67), in
 * reduce-scatter / all-gather form: rank r sums slice r of all ranks' gradients (rank order), scales by 1/world,
 * applies Adam to slice r of its own exp_avg / exp_avg_sq and stores the updated parameter slice into EVERY rank's
 * parameter buffer.  params[q] / grads[q] / flags[q], q < world: device pointers to rank q's flat parameter buffer,
 * flat gradient buffer and flag buffer (explainn_p2p_flag_bytes() bytes, zero-initialised once), valid in THIS process
 * (the caller opens the peers' buffers through CUDA IPC); index `rank` holds the local ones.  The three arrays
 * themselves are host arrays.  The call updates the floats [offset, offset + n) of the flat buffers (16-byte aligned,
 * offset % 4 == n % 4 == 0), split into `world` slices; *step (device) is incremented first when tick != 0 -- a step may
 * be issued as several calls over disjoint ranges (the FC-stack gradients are final before the conv backward runs, so
 * their update can overlap it), the first with tick = 1, each with its own seq.  Calls on one GPU must not overlap in time.
 * The moments of slices other than `rank` are not touched.  seq: the same strictly increasing positive number on every
 * rank for every step.  All ranks must call it once per step.  A late peer is waited for (never a trap); after
 * timeout_s seconds (0 = wait for ever) the kernel gives up, sets the error word that explainn_p2p_error reads
 * (0 = healthy, 1 + q = gave up waiting for rank q) and finishes with undefined parameters. */
/* Peer buffers: plain cudaMalloc allocations (zero-filled) that CUDA IPC can export -- the only memory this library
 * allocates; export writes a 64-byte handle, open maps a peer's allocation for the current device. */
int explainn_p2p_alloc(long long bytes, void **out);
int explainn_p2p_free(void *p);
int explainn_p2p_export(const void *p, unsigned char *handle);
int explainn_p2p_open(const unsigned char *handle, void **out);
int explainn_p2p_close(void *p);
long long explainn_p2p_flag_bytes(void);
int explainn_p2p_error(const unsigned long long *flags, int *out);
int explainn_p2p_adam_step(float *const *params, float *exp_avg, float *exp_avg_sq, long long offset, long long n,
                           float *step, int tick, const float *const *grads, unsigned long long *const *flags, int world, int rank,
                           unsigned long long seq, float lr, float beta1, float beta2, float eps, double timeout_s,
                           cudaStream_t stream);

/* out[i] = sum_{c<k} partials[c*n + i] */
int explainn_reduce_partials(float *out, const float *partials, long long n, int k, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* EXPLAINN_B200_H */
