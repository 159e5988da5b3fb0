/* clsm_b200.h -- C ABI of libclsm_b200.so: the CLSM (convolutional DSSM) hot path of
 * collawolley/dnn_page_vectors `dssm_cnn_v2`, as hand-written sm_100a CUDA kernels.
 *
 * The reference has no FFI of its own: the path is Python that calls Keras-1 layers
 * (dssm_cnn_v2/cnn_dssm_tf.py).  Each entry point below therefore cites the reference
 * LAYER CALL it replaces (file:line under /root/reference).  The Python host package
 * `dnn_page_vectors_b200` binds exactly these symbols through ctypes; INTEGRATION.md
 * shows the stub.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless the name ends in `_h`;
 *  - every call is asynchronous on `stream` (a cudaStream_t passed as void*), takes no
 *    lock and keeps no global state except the last-error string (thread local);
 *  - the library allocates nothing: scratch is a caller-provided workspace whose size
 *    comes from the matching *_workspace_bytes() query;
 *  - return value: 0 = CLSM_OK, negative = error (see clsm_last_error()).
 *  - row-major everywhere; `ld*` are leading dimensions in ELEMENTS.
 */
#ifndef CLSM_B200_H_
#define CLSM_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { CLSM_OK = 0, CLSM_EINVAL = -1, CLSM_ECUDA = -2, CLSM_EWORKSPACE = -3, CLSM_EUNSUPPORTED = -4 };
enum { CLSM_F32 = 0, CLSM_BF16 = 1 };
enum { CLSM_ACT_LINEAR = 0, CLSM_ACT_RELU = 1, CLSM_ACT_TANH = 2 };
/* OR into the `act` argument of clsm_dense_fwd / clsm_dense_bwd: run their GEMMs on the tensor cores with TF32
 * operands (10-bit mantissa, fp32 accumulation) instead of exact fp32 on the CUDA cores.  The host model sets it
 * in its bf16 precision mode and never in the fp32 parity mode. */
#define CLSM_MATH_TF32 0x100
/* conv_pool_fwd implementation selector */
enum { CLSM_IMPL_AUTO = 0, CLSM_IMPL_SIMT = 1, CLSM_IMPL_TCGEN05 = 2 };

int clsm_version(void);
const char* clsm_last_error(void);
/* number of SMs of the current device (grid sizing is done inside the library) */
int clsm_num_sms(void);
/* kernels of this library launched by this process so far (bench.py's gpu_launches) */
unsigned long long clsm_launch_count(void);
/* measurement aid (bench.py's per-kernel roofline): the NEXT clsm_conv_pool_bwd call of this thread records the two
 * given cudaEvent_t handles on its stream, the first right after its dW kernel and the second right before its dX
 * kernel (the weight-gradient reduction and the bias column sum lie between them), then forgets them.  NULL = none. */
void clsm_debug_mark_conv_bwd(void* ev_after_dw, void* ev_before_dx);

/* ---- embedding gather (+ optional inverted dropout) ------------------------------------
 * replaces Embedding(...)            dssm_cnn_v2/cnn_dssm_tf.py:111-116
 *      and Dropout(dropout_prob[0])  dssm_cnn_v2/cnn_dssm_tf.py:117-122
 * X[r, 0:D] = table[tok[r], 0:D] * mask(r, c) / keep ; X[r, D:ldx] = 0.
 * keep >= 1 disables dropout.  mask(r, c) is a counter-based Bernoulli(keep) bit that is a
 * pure function of (seed, r, c) -- clsm_dropout_mask exports it for the oracle.
 * dtype: CLSM_BF16 (table and X bf16) or CLSM_F32 (both fp32).  Rows must be 16-byte aligned
 * (ld_table*esize and ldx*esize multiples of 16).  Token ids are trusted to lie in [0, n_rows), as Keras
 * trusts them (an id outside is an out-of-bounds read); clsm_check_token_ids validates a batch when the feeder is
 * not trusted. */
int clsm_emb_gather(const int32_t* tok, int64_t n_tok, const void* table, int64_t n_rows,
                    int D, int ld_table, int dtype, void* X, int ldx,
                    float keep, uint64_t seed, void* stream);
int clsm_dropout_mask(int64_t n_tok, int D, float keep, uint64_t seed, uint8_t* mask, void* stream);
/* n_bad[0] (device) <- how many of tok[0:n_tok] lie outside [0, n_rows).  Asynchronous; the caller reads the count. */
int clsm_check_token_ids(const int32_t* tok, int64_t n_tok, int64_t n_rows, int32_t* n_bad, void* stream);
/* clsm_emb_gather that also stores the keep decisions it applied, 1 bit per element: mask_bits (n_tok, ldx / 16)
 * uint16, bit c of word (r, b) = column 16 b + c of row r kept.  clsm_emb_scatter_reduce reads them back, so the
 * backward pass does not regenerate the random stream.  mask_bits may be NULL (then identical to clsm_emb_gather).
 * The keep probability is realised in steps of 1/256 (threshold t = round(256 keep), P(keep) = t / 256) and kept
 * values are scaled by 256 / t = clsm_dropout_scale(keep), the inverse of the REALISED probability, so the expectation
 * is exact for every keep; the reference's 0.75 (cnn_dssm_tf.py:37) is t = 192, scale 4/3. */
int clsm_emb_gather_mask(const int32_t* tok, int64_t n_tok, const void* table, int64_t n_rows,
                         int D, int ld_table, int dtype, void* X, int ldx,
                         float keep, uint64_t seed, uint16_t* mask_bits, void* stream);
float clsm_dropout_scale(float keep);
/* The training gather straight from the fp32 MASTER table (n_rows, D), D % 4 == 0, into bf16 X (n_tok, ldx):
 * X = bf16(table[tok] * mask * 256/t), one rounding.  With it the optimizer need not maintain the bf16 compute copy of
 * the table every step (clsm_adam_table with bf16_copy = NULL): 0.64 GB less written per 1M x 300 table and step.
 * Same mask stream and stored bits as clsm_emb_gather_mask.  (The host model uses it on a single GPU; data parallel
 * training keeps the bf16 copy, which is what the ranks broadcast to each other.) */
int clsm_emb_gather_master(const int32_t* tok, int64_t n_tok, const float* table, int64_t n_rows, int D,
                           void* X, int ldx, float keep, uint64_t seed, uint16_t* mask_bits, void* stream);
/* out (n, H) = x * mask / keep with the same counter-based mask (clsm_dropout_mask(n, H, keep, seed) exports it);
 * out may be x.  replaces the SECOND Dropout of the v1 model, Dense -> Dropout(dropout_prob[1]) -> relu
 * (dssm_cnn/cnn_dssm.py:148-150), applied to the semantic vectors (relu commutes with the non-negative scaling) and,
 * in the backward pass, to their gradient. */
int clsm_dropout_rows(const float* x, int64_t n, int H, float keep, uint64_t seed, float* out, void* stream);

/* ---- conv weights: reference layout -> compute layout ------------------------------------
 * W_ref fp32 (k, D, F)  [Keras-1 Convolution1D kernel (k,1,D,F), cnn_dssm_tf.py:85-91]
 * Wt   dtype (Fpad, k*Dpad): Wt[f, j*Dpad + c] = W_ref[j, c, f], zero elsewhere. */
int clsm_pack_conv_w(const float* W_ref, int k, int D, int F, int Dpad, int Fpad,
                     int dtype, void* Wt, void* stream);

/* ---- n-gram window convolution + max over time ---------------------------------------
 * replaces Convolution1D(valid) + MaxPooling1D(L-k+1) + Flatten
 *          dssm_cnn_v2/cnn_dssm_tf.py:84-100
 * X (S*L, ldx) as written by clsm_emb_gather (it must own at least 8 readable rows past the
 * end for the tensor-core path; the gather's callers allocate S*L+8).  Wt as packed above.
 * keys (S, ldk) uint64, column f of this window at keys[s, col0 + f]: on return
 *   keys = (orderable(max_t acc[s,t,f]) << 32) | (0xFFFFFFFF - argmax_t), ties -> lowest t,
 * where acc is the fp32 accumulator before bias/activation (max commutes with both).
 * The call clears its key columns itself.  impl: CLSM_IMPL_*; TCGEN05 needs dtype BF16,
 * Dpad % 64 == 0 and Fpad % 16 == 0, Fpad <= 512. */
int clsm_conv_pool_fwd(const void* X, int dtype, int64_t S, int L, int Dpad,
                       const void* Wt, int k, int F, int Fpad,
                       uint64_t* keys, int ldk, int col0, int impl, void* stream);

/* The same convolution + max over time with the embedding lookup INSIDE the kernel (inference: no dropout):
 * replaces Embedding + Convolution1D + MaxPooling1D, cnn_dssm_tf.py:111-116,84-100, in one launch.
 * tok (S*L) int32 ids, table (n_rows, ld_table = Dpad) bf16 as clsm_table_to_bf16 writes it.  The token tile of the
 * tensor-core kernel is fetched row by row from the table by the copy engine (TMA tile::gather4, the token ids are
 * the row coordinates), so the gathered matrix X is never written to memory or read back.  Results are bit-identical
 * to clsm_emb_gather (keep = 1) + clsm_conv_pool_fwd.  Ids are trusted to lie in [0, n_rows).  CLSM_EUNSUPPORTED for
 * shapes outside the kernel's resident-tile plan (then gather + clsm_conv_pool_fwd). */
int clsm_conv_pool_fwd_tok(const int32_t* tok, const void* table, int64_t n_rows, int ld_table, int64_t S, int L, int D, int Dpad,
                           const void* Wt, int k, int F, int Fpad, uint64_t* keys, int ldk, int col0, void* stream);
/* clsm_conv_pool_fwd told the number of REAL columns D (columns D..Dpad-1 of X and Wt are zero): the tensor-core kernel
 * then skips the K steps that would only multiply padding (D = 300 in Dpad = 320: 19 of 20).  Same results; D = 0
 * means unknown (= clsm_conv_pool_fwd).  clsm_conv_pool_fwd_tok takes the same D. */
int clsm_conv_pool_fwd_d(const void* X, int dtype, int64_t S, int L, int D, int Dpad,
                         const void* Wt, int k, int F, int Fpad,
                         uint64_t* keys, int ldk, int col0, int impl, void* stream);

/* keys -> pooled = act(max + bias) fp32 (S, ldp), argmax int32 (S, ldp) (may be NULL).
 * For relu a column with pooled == 0 reports argmax 0 (all positions tie at 0 after the
 * activation).  replaces the activation of Convolution1D + concat, cnn_dssm_tf.py:88,102-107 */
int clsm_pool_finalize(const uint64_t* keys, int64_t S, int P, int ldk, const float* bias,
                       int act, float* pooled, int32_t* argmax, int ldp, void* stream);

/* ---- dense semantic projection ----------------------------------------------------------
 * replaces Dense(hidden_dims) + Activation('relu')   cnn_dssm_tf.py:124-126
 * sem (S, H) = act(pooled (S, P) . Wd (P, H) + bd). */
int clsm_dense_fwd(const float* pooled, int64_t S, int P, int ldp, const float* Wd, const float* bd,
                   int H, int act, float* sem, void* stream);

/* dense backward.  dpre = dsem * act'(sem);  dWd = pooled^T dpre;  dbd = colsum(dpre);
 * G (S, ldp) = (dpre . Wd^T) * act'(pooled)   (gradient arriving at the pooled maxima).
 * accumulate != 0 adds into dWd/dbd (the doc tower is applied 1+J times with shared weights,
 * cnn_dssm_tf.py:137-152; here that is one batched call, so accumulate is normally 0). */
size_t clsm_dense_bwd_workspace_bytes(int64_t S, int P, int H);
int clsm_dense_bwd(const float* pooled, const float* sem, const float* dsem, int64_t S, int P, int ldp,
                   const float* Wd, int H, int act, float* dWd, float* dbd, float* G,
                   void* ws, size_t ws_bytes, void* stream);

/* ---- cosine / gamma-softmax / loss, forward + gradient in one pass --------------------------
 * replaces RTF (cnn_dssm_tf.py:63-77), x*GAMMA, exp, p0 = e0/sum (:164-176),
 *          binary_crossentropy + accuracy (:182).
 * semQ (B, H); semD (B*(1+J), H), document row b*(1+J)+j, j = 0 the positive.
 * out: cosv, prob (B, 1+J); loss_acc[0] = mean_b -log clip(p0), [1] = mean_b (p0 > 0.5)
 * (means over B; gradients scaled by 1/batch_total so data-parallel ranks pass the global
 * batch);  dSemQ (B, H), dSemD (B*(1+J), H) may be NULL for evaluation. */
int clsm_sim_loss(const float* semQ, const float* semD, int64_t B, int J, int H, float gamma,
                  int64_t batch_total, float* cosv, float* prob, float* loss_acc,
                  float* dSemQ, float* dSemD, void* stream);

/* ---- conv backward (arg-max sparse) -------------------------------------------------------
 * gradient of Convolution1D + global max-pool: only the arg-max window of each (s, f) gets
 * gradient.  G, argmax (S, ldp) at columns [col0, col0+F).  X, Wt as in the forward.
 *   dW_ref (k, D, F) fp32 = sum_s G[s,f] X[s, a+j, c];  db (F) = sum_s G[s,f]
 *   dX (S*L, Dpad) fp32 += / = sum_{f,j: a+j==t} G[s,f] W[j,c,f]
 * first_window != 0 overwrites dX, otherwise accumulates (several window sizes). */
size_t clsm_conv_bwd_workspace_bytes(int64_t S, int L, int k, int D, int F);
int clsm_conv_pool_bwd(const void* X, int dtype, int64_t S, int L, int D, int Dpad,
                       const void* Wt, int k, int F, int Fpad,
                       const float* G, const int32_t* argmax, int ldp, int col0,
                       float* dW_ref, float* db, float* dX, int first_window,
                       void* ws, size_t ws_bytes, void* stream);

/* Same, with a choice of dX format.  dx_dtype CLSM_BF16 stores dX rows as bf16 (S*L, Dpad): the tensor-core dX
 * kernel rounds its fp32 accumulators once on the way out, which halves the largest memory stream of a training step
 * (written here, read once by clsm_emb_scatter_reduce, which still sums in fp32).  Only with dtype CLSM_BF16,
 * first_window = 1 (one window size: later windows ADD into dX and need fp32) and a shape for which
 * clsm_conv_bwd_dx_bf16_ok() returns 1; CLSM_EUNSUPPORTED otherwise.  (No counterpart in the reference: Keras keeps
 * every gradient in the model's float type.) */
int clsm_conv_bwd_dx_bf16_ok(int64_t S, int L, int k, int D, int F);
int clsm_conv_pool_bwd_ex(const void* X, int dtype, int64_t S, int L, int D, int Dpad,
                          const void* Wt, int k, int F, int Fpad,
                          const float* G, const int32_t* argmax, int ldp, int col0,
                          float* dW_ref, float* db, void* dX, int dx_dtype, int first_window,
                          void* ws, size_t ws_bytes, void* stream);
/* The same, one gradient at a time: parts = CLSM_BWD_DW (dW_ref, db), CLSM_BWD_DX (dX) or both.  The two parts read the
 * same inputs, write different outputs and use disjoint regions of ws, so they may be issued on two streams. */
#define CLSM_BWD_DW 1
#define CLSM_BWD_DX 2
int clsm_conv_pool_bwd_parts(const void* X, int dtype, int64_t S, int L, int D, int Dpad,
                             const void* Wt, int k, int F, int Fpad,
                             const float* G, const int32_t* argmax, int ldp, int col0,
                             float* dW_ref, float* db, void* dX, int dx_dtype, int first_window, int parts,
                             void* ws, size_t ws_bytes, void* stream);
/* dX of SEVERAL window sizes in one pass (bf16 tensor-core path): the accumulators stay on chip across the windows and
 * dX (S*L, Dpad), fp32 or bf16, is written once -- one launch per window makes every window after the first add into
 * fp32 rows in HBM.  Wt / ks: HOST arrays of n_windows (2..4) device pointers to the windows' compute weights
 * (Fpad, k_w * Dpad) bf16 and of their widths; window w's columns of G / argmax are col0 + w*F .. col0 + (w+1)*F.
 * clsm_conv_bwd_dx_multi_ok() = 1 when the kernel takes the shape (else use clsm_conv_pool_bwd_parts per window);
 * ws: clsm_conv_bwd_dx_multi_workspace_bytes().  The conv weight gradients still come from
 * clsm_conv_pool_bwd_parts(..., CLSM_BWD_DW, ...) per window. */
int clsm_conv_bwd_dx_multi_ok(int64_t S, int L, int n_windows, const int* ks, int D, int F);
size_t clsm_conv_bwd_dx_multi_workspace_bytes(int n_windows, const int* ks, int D, int F);
int clsm_conv_pool_bwd_dx_multi(int64_t S, int L, int D, int Dpad, const void* const* Wt, const int* ks, int n_windows, int F,
                                const float* G, const int32_t* argmax, int ldp, int col0, void* dX, int dx_dtype,
                                void* ws, size_t ws_bytes, void* stream);

/* ---- embedding gradient: deterministic sort + segmented reduce -------------------------------
 * replaces the gradient of Embedding (AdvancedIncSubtensor1 / UnsortedSegmentSum).
 * dX (n_tok, Dpad) fp32; if keep < 1 the dropout mask of the forward is re-applied.
 * out: n_uniq_out[0] = number of distinct ids; uniq_ids[0:n_uniq]; dE_rows (n_uniq, Dpad) fp32 =
 * sum over the tokens with that id, added in ascending token position, independent of the
 * grid -- the same input gives the same bits on every run and every GPU.
 * row_of_id (n_rows) int32: -1, or the row of dE_rows holding that id's gradient (for Adam). */
size_t clsm_emb_scatter_workspace_bytes(int64_t n_tok, int D);
int clsm_emb_scatter(const int32_t* tok, int64_t n_tok, const float* dX, int Dpad, int D,
                     float keep, uint64_t seed, int64_t n_rows,
                     int32_t* uniq_ids, int32_t* n_uniq_out, float* dE_rows, int32_t* row_of_id,
                     void* ws, size_t ws_bytes, void* stream);

/* The same in two halves, for hosts that overlap them with other work.
 * clsm_emb_scatter_plan: everything that depends on the token ids alone (stable sort, segments, uniq_ids, n_uniq_out,
 *   row_of_id); it needs no gradient, so it can run at the start of the step on another stream, under the forward pass.
 * clsm_emb_scatter_reduce: the segmented sum over dX (dx_dtype CLSM_F32 or CLSM_BF16 rows, fp32 sums) into dE_rows, on
 *   the plan that the SAME workspace holds for the same tokens.  Dropout (keep < 1): mask_bits != NULL reads the
 *   1-bit-per-element mask clsm_emb_gather_mask stored in the forward, NULL regenerates the stream from seed.
 * clsm_emb_scatter == plan + reduce(fp32 dX, regenerated mask): bit-identical results. */
int clsm_emb_scatter_plan(const int32_t* tok, int64_t n_tok, int Dpad, int D, int64_t n_rows,
                          int32_t* uniq_ids, int32_t* n_uniq_out, int32_t* row_of_id,
                          void* ws, size_t ws_bytes, void* stream);
int clsm_emb_scatter_reduce(int64_t n_tok, const void* dX, int dx_dtype, int Dpad, int D,
                            float keep, uint64_t seed, const uint16_t* mask_bits, float* dE_rows,
                            void* ws, size_t ws_bytes, void* stream);

/* ---- optimizer -----------------------------------------------------------------------------
 * replaces optimizer="adam" (Keras-1 Adam, cnn_dssm_tf.py:182):
 *   t = step (1-based); lr_t = lr sqrt(1-b2^t)/(1-b1^t); m = b1 m + (1-b1) g;
 *   v = b2 v + (1-b2) g^2; p -= lr_t m / (sqrt(v) + eps).
 * clsm_adam_dense: flat tensors.  clsm_adam_table: the (n_rows, D) embedding table with its
 * gradient given row-sparsely (row_of_id / dE_rows (n_uniq, ldg) from clsm_emb_scatter); EVERY row is
 * updated (rows without gradient use g = 0), which is what Keras does.  bf16_copy (n_rows,
 * ld_copy) is refreshed in the same pass when not NULL. */
int clsm_adam_dense(float* p, float* m, float* v, const float* g, int64_t n, int step,
                    float lr, float b1, float b2, float eps, void* stream);
int clsm_adam_table(float* p, float* m, float* v, int64_t n_rows, int D,
                    const int32_t* row_of_id, const float* dE_rows, int ldg, int step,
                    float lr, float b1, float b2, float eps,
                    void* bf16_copy, int ld_copy, void* stream);
/* lazy (row-sparse) Adam: only rows in uniq_ids move.  NOT the reference's semantics. */
int clsm_adam_rows(float* p, float* m, float* v, int D, const int32_t* uniq_ids,
                   const int32_t* n_uniq, int64_t max_uniq, const float* dE_rows, int ldg, int step,
                   float lr, float b1, float b2, float eps, void* bf16_copy, int ld_copy, void* stream);
/* Deferred rows: the SAME dense Keras semantics as clsm_adam_table, without streaming the whole table every step.
 * A row without gradient takes its step with g = 0, which reads only the row's own p, m, v and the step number --
 * so it can be taken later, when the row is next needed, by replaying the owed steps in order in registers.
 * last (n_rows) int32: the last step stored p, m, v include; -1 = m and v still exactly zero (such a row owes
 * nothing).  Initialise it to -1 with zeroed m, v, or to the current step when optimizer state is loaded.
 * lr_tab: DEVICE array, lr_tab[s] = lr_t of step s for every s up to the step in use (fill a host array with
 * clsm_adam_lr_table and copy it; entry 0 is unused).  Once every row is settled the table, m and v equal bit for bit
 * what clsm_adam_table would have left.
 *   clsm_adam_defer_settle_tokens: before a batch is gathered at step t, bring the rows its tokens name up to
 *     to_step = t - 1.  ws: clsm_adam_defer_workspace_bytes(n_tok, n_rows).
 *   clsm_adam_defer_rows: the optimizer step `step` of the rows that have a gradient (uniq_ids / n_uniq / dE_rows of
 *     clsm_emb_scatter); rows that still owe steps are settled to step - 1 first.
 *   clsm_adam_defer_settle_all: every row up to to_step -- before the table is read as a whole (export, checkpoint,
 *     inference, a change of lr).
 * bf16_copy (n_rows, ld_copy), when not NULL, receives the bf16 image of every row written. */
int clsm_adam_lr_table(float lr, float b1, float b2, int first_step, int n, float* lr_t_host);
size_t clsm_adam_defer_workspace_bytes(int64_t n_tok, int64_t n_rows);
int clsm_adam_defer_settle_tokens(float* p, float* m, float* v, int64_t n_rows, int D, int32_t* last,
                                  const int32_t* tok, int64_t n_tok, int to_step, const float* lr_tab,
                                  float b1, float b2, float eps, void* bf16_copy, int ld_copy,
                                  void* ws, size_t ws_bytes, void* stream);
int clsm_adam_defer_rows(float* p, float* m, float* v, int D, int32_t* last, const int32_t* uniq_ids,
                         const int32_t* n_uniq, int64_t max_uniq, const float* dE_rows, int ldg, int step,
                         const float* lr_tab, float b1, float b2, float eps, void* bf16_copy, int ld_copy, void* stream);
int clsm_adam_defer_settle_all(float* p, float* m, float* v, int64_t n_rows, int D, int32_t* last, int to_step,
                               const float* lr_tab, float b1, float b2, float eps, void* bf16_copy, int ld_copy, void* stream);
/* fp32 (n_rows, D) -> bf16 (n_rows, ld_copy) with zero pad columns */
int clsm_table_to_bf16(const float* p, int64_t n_rows, int D, void* bf16_copy, int ld_copy, void* stream);

/* ---- data-parallel table optimizer over peer memory ------------------------------------------
 * No counterpart in the reference: its only multi-device code pins tower instances to devices
 * (cnn_dssm_tf.py:139-156).  One process per GPU; rank r owns table rows [r*R, (r+1)*R), R = rows_per_owner.
 * The caller maps every rank's receive buffers and bf16 compute tables into this process (CUDA IPC) and passes
 * HOST arrays of `world` device pointers (world <= 16).
 *
 * clsm_dp_push_rows: cut this rank's row-sparse gradient (output of clsm_emb_scatter: uniq_ids ascending,
 *   n_uniq on the device, compact dE_rows (max_uniq, ldg), row_of_id (n_rows)) at the ownership boundaries and
 *   store owner o's piece into peer_rows[o] (this rank's (cap, ldg) slab of o's receive buffer, rows compact
 *   from 0) and the rebased map slice into peer_map[o] (this rank's (R) slab: compact row inside the piece or
 *   -1).  cap >= min(R, tokens per rank).  seg_scratch: world + 1 ints of device scratch.
 * clsm_adam_table_dp: dense Keras Adam (same arithmetic as clsm_adam_table) over this rank's n_rows owned rows
 *   p/m/v (n_rows, D); the gradient of a row is the sum over sources s = 0..world-1, in that order, of
 *   rows_in[s][maps_in[s][row]] where the map entry is >= 0 (maps_in (world, R), rows_in (world, cap, ldg), both
 *   local); the refreshed bf16 row is stored at the same row offset of every bf16_copies[s] (each rank's view
 *   of its compute table at this rank's first owned row; bf16_copies[self_rank] is this rank's own).  Replicas
 *   must hold identical compute tables on entry: a 4-element group is sent to the peers only when its bf16 image
 *   differs from the local old one.  D, ldg, ld_copy multiples of 4.
 * Ordering is the caller's: a collective (or any cross-rank barrier) on the same stream between the pushes and
 * the Adam call, and another before the next read of the compute tables. */
int clsm_dp_push_rows(const int32_t* uniq_ids, const int32_t* n_uniq, int64_t max_uniq, const float* dE_rows, int ldg, int D,
                      const int32_t* row_of_id, int64_t n_rows, int64_t rows_per_owner, int world,
                      void* const* peer_rows, void* const* peer_map, int64_t cap, int32_t* seg_scratch, void* stream);
int clsm_adam_table_dp(float* p, float* m, float* v, int64_t n_rows, int D, int world,
                       const int32_t* maps_in, int64_t rows_per_owner, const float* rows_in, int64_t cap, int ldg,
                       int step, float lr, float b1, float b2, float eps,
                       void* const* bf16_copies, int self_rank, int ld_copy, void* stream);

/* ---- page vectors + retrieval (no script in the reference; composition of model() and RTF) ---
 * clsm_normalize_rows: vec (N, H) = sem / sqrt(max(|sem|^2, FLT_MIN))   (RTF's norm, :63-72)
 * clsm_cos_topk: scores = Qn . Cn^T over corpus rows [0, N), top-k per query ordered by
 *   (score desc, id asc); ids are id_base + row.  out_score (nQ, k) fp32, out_id (nQ, k) int64.
 * clsm_topk_merge: merge n_parts lists (n_parts, nQ, k) -> (nQ, k), same order.  ids are int64; any negative id
 *   marks an empty slot.  The valid ids of ONE call must span less than 2^32 - 1 (max - min): the merge keys carry
 *   32 bits of id relative to the smallest one.  A wider span cannot be merged exactly; the call then fills
 *   out_id with -2 and out_score with NaN instead of returning truncated ids. */
int clsm_normalize_rows(const float* sem, int64_t N, int H, float* vec, void* stream);
size_t clsm_cos_topk_workspace_bytes(int64_t nQ, int64_t N, int k);
int clsm_cos_topk(const float* Qn, int64_t nQ, const float* Cn, int64_t N, int H, int k,
                  int64_t id_base, float* out_score, int64_t* out_id,
                  void* ws, size_t ws_bytes, void* stream);
size_t clsm_topk_merge_workspace_bytes(int n_parts, int64_t nQ, int k);
int clsm_topk_merge(const float* scores, const int64_t* ids, int n_parts, int64_t nQ, int k,
                    float* out_score, int64_t* out_id, void* ws, size_t ws_bytes, void* stream);

/* bf16 corpus, tensor-core scoring (the production retrieval path): Qn (nQ, H) and Cn (N, H) are
 * L2-normalised bf16 (clsm_normalize_rows_bf16), H = 64 or 128 (other widths: clsm_cos_topk).  The corpus is read once per block of
 * 128 queries by a tcgen05 kernel; selection is the same exact radix select, so the order and the
 * independence from sharding are those of clsm_cos_topk.  The corpus is scanned in segments (4 096 rows, then to 65 536,
 * to 1 Mi, then 1 Mi at a time), each behind the k-th best key of the rows before it; a row that beats it goes to a
 * per-query survivor list that can hold a whole segment, so no corpus order can overflow it.  The workspace
 * therefore grows with N up to about 1 GiB (N >= 1.1 M rows). */
int clsm_normalize_rows_bf16(const float* sem, int64_t N, int H, void* vec, void* stream);
size_t clsm_cos_topk_bf16_workspace_bytes(int64_t nQ, int64_t N, int k);
int clsm_cos_topk_bf16(const void* Qn, int64_t nQ, const void* Cn, int64_t N, int H, int k,
                       int64_t id_base, float* out_score, int64_t* out_id,
                       void* ws, size_t ws_bytes, void* stream);

/* clsm_topk_merge_static: clsm_topk_merge without a workspace argument and without any allocation.
 * (No counterpart in the reference, which has no retrieval code at all.)
 * Who needs it: a caller that is not allowed to obtain device memory at the moment it has to merge --
 * a per-request handler inside an inference server whose rules forbid allocating on the request path
 * (a device allocation can stall every stream of the process), or a host that has already given all
 * free HBM to its own pool and has no raw workspace left to lend.  Such a caller still passes the input
 * lists and the output buffers; only the scratch comes from the library: a fixed array of
 * CLSM_TOPK_MERGE_STATIC_MAX_BYTES that every GPU loading the library carries until the process ends.
 *  - Sizes whose scratch need (clsm_topk_merge_workspace_bytes) exceeds that limit are refused with
 *    CLSM_EWORKSPACE and a message naming both numbers; use clsm_topk_merge for those.
 *  - The first call on each device looks up the array's device address (once, with that device
 *    current).  Make that first call during start-up, before any stream capture is begun.
 *  - All calls on a device share the one array.  Two calls whose kernels can overlap on the GPU (two
 *    streams, two threads) race on it; serialise them on one stream.  Nothing needs clearing between
 *    calls: each call overwrites what it later reads.
 *  - Calls neither cudaMalloc/cudaFree nor any allocating library routine.  Results are identical to
 *    clsm_topk_merge on the same inputs. */
#define CLSM_TOPK_MERGE_STATIC_MAX_BYTES (2 * 1024 * 1024)
int clsm_topk_merge_static(const float* scores, const int64_t* ids, int n_parts, int64_t nQ, int k,
                           float* out_score, int64_t* out_id, void* stream);

/* One strided copy (cudaMemcpy2DAsync, any direction): `rows` rows of width_bytes from src (row pitch src_pitch) to dst
 * (row pitch dst_pitch).  The host model uses it, device to device, to put each document input of a batch -- pos_doc and
 * neg_doc_j arrive as separate (B, Ld) arrays, data_helpers.py:196-197 -- into its slot of the document tower's
 * interleaved (B, 1+J, Ld) token matrix: no transpose kernel.  (From HOST memory make one contiguous copy first: a
 * strided host copy of 400-byte rows was measured at 2 ms for 8 MB.) */
int clsm_copy_rows_strided(const void* src, size_t src_pitch, void* dst, size_t dst_pitch, size_t width_bytes,
                           int64_t rows, void* stream);

/* ---- host-buffer convenience (the e2e / drop-in call): ids on the HOST ------------------------
 * Copies tok_h (pinned or pageable) to tok_d on `stream`, then clsm_emb_gather. */
int clsm_emb_gather_h(const int32_t* tok_h, int32_t* tok_d, int64_t n_tok, const void* table,
                      int64_t n_rows, int D, int ld_table, int dtype, void* X, int ldx,
                      float keep, uint64_t seed, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CLSM_B200_H_ */
