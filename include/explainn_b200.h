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
 * Replaces linears[0..3] (explainn/models/networks.py:98-102).  pooled: (U,B,P) -- UNIT-MAJOR, so that the
 * (unit, 128-sequence) tile the FC kernels work on is one contiguous block. */
int explainn_conv_pool_fwd(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                           const float *tables, const float *wq_t, const float *shift, float *pooled,
                           cudaStream_t stream);

/* Training variant: tables built with gamma=1, beta=0 and the batch statistics, so the scan yields
 * x_hat = (y-mu)/sigma; z = gamma*x_hat+beta.  Also returns x_hat at the window maximum and the
 * first-max index inside the window (MaxPool1d backward routing).  pooled/xhat (U,B,P) fp32, idx (U,B,P) u8. */
int explainn_conv_pool_fwd_train(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                 const float *tables, const float *wq_t, const float *shift,
                                 const float *gamma, const float *beta, float *pooled, float *xhat,
                                 unsigned char *idx, cudaStream_t stream);

/* ---- activation scan ------------------------------------------------------------------------------------
 * Replaces `model.linears[:3](seq.repeat(...))` + `.astype(np.half)`
 * (explainn/interpretation/interpretation.py:143-147).  act: (B,U,S) in act_dtype. */
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
/* float16 threshold 0.5*max (interpretation.py:187) and its pre-activation bracket.  All outputs (U). */
int explainn_interp_thresholds(const float *zmax, int U, float *thr, float *zlo, float *zhi, float *amax,
                               cudaStream_t stream);
/* Pass B: counts[u,c,k] += #positions with half(act) > thr[u] whose window has base c at offset k
 * (interpretation.py:203-214); hits[u] += number of such positions.  counts (U,4,F) u64, hits (U) u64,
 * both accumulated into (zero them first). */
int explainn_interp_pwm_count(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                              const float *tables, const float *wq_t, const float *shift, const float *zlo,
                              const float *zhi, const float *thr, unsigned long long *counts,
                              unsigned long long *hits, cudaStream_t stream);

/* ---- per-unit FC stack, eval mode -----------------------------------------------------------------------
 * Replaces linears[4..13] with BatchNorm in eval mode (explainn/models/networks.py:103-116): grouped 1x1
 * Conv1d(U*P->100U) -> BN -> ReLU -> (Dropout = identity) -> grouped 1x1 Conv1d(100U->U) -> BN -> ReLU.
 * pooled (U,B,P); W1 = linears.6.weight (100U,P); b1 (100U); bn2_* (100U); W2 = linears.10.weight (U,100);
 * b2 (U); bn3_* (U).  unit_out: (B,U), the tensor get_explainn_unit_outputs returns
 * (explainn/interpretation/interpretation.py:82-99). */
int explainn_fc_stack_fwd_eval(const float *pooled, long long B, int U, int P, const float *W1, const float *b1,
                               const float *bn2_w, const float *bn2_b, const float *bn2_rm, const float *bn2_rv,
                               const float *W2, const float *b2, const float *bn3_w, const float *bn3_b,
                               const float *bn3_rm, const float *bn3_rv, float eps, float *unit_out,
                               cudaStream_t stream);

/* ---- final linear ------------------------------------------------------------------------------------------
 * out (B,C) = x (B,U) @ W(C,U)^T + bias (explainn/models/networks.py:147,161). */
int explainn_linear_fwd(const float *x, long long B, int U, int C, const float *W, const float *bias, float *out,
                        cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* EXPLAINN_B200_H */
