/* b200ps — C ABI of the native sm_100a library (libb200ps.so).
 *
 * Every entry point takes raw device pointers plus the CUDA stream to launch on and
 * returns 0 on success or a negative b200ps error code (b200ps_last_error() gives the
 * text).  Nothing here synchronises the stream.  The Python package reaches these through
 * ctypes (b200ps/ops/native.py); C/C++ embedders include this header and link the .so.
 *
 * Citations of the form (example.py:95) point into the reference tree
 * Amano-Ginji/distributed-tensorflow-example and name the behaviour each kernel replaces.
 */
#ifndef B200PS_H_
#define B200PS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* b200ps_stream_t; /* cudaStream_t */

enum {
  B200PS_OK = 0,
  B200PS_ERR_INVALID = -1, /* bad argument (shape, alignment, dtype, size limit) */
  B200PS_ERR_CUDA = -2,    /* a CUDA runtime/driver call failed */
  B200PS_ERR_UNSUPPORTED = -3
};

/* Text of the last error raised on the calling thread. */
const char* b200ps_last_error(void);
/* Library version string, e.g. "b200ps 0.1 sm_100a". */
const char* b200ps_version(void);
/* Number of kernel launches issued by this library in this process (all threads). */
uint64_t b200ps_launch_count(void);

/* ------------------------------------------------------------------------------------
 * K2/K7: tcgen05 + TMA GEMM with fused epilogues.
 *   C[M,N] = epilogue( alpha * op(A)[M,K] * op(B)[K,N] )
 * A is stored either K-major ([M,K] row-major, a_mn_major = 0) or MN-major ([K,M]
 * row-major, a_mn_major = 1); B either K-major ([N,K] row-major, b_mn_major = 0) or
 * MN-major ([K,N] row-major, b_mn_major = 1).  lda/ldb/ldc/ldaux are row strides in
 * ELEMENTS of the stored 2-D arrays.  Operand rows must be 16-byte aligned.
 * Replaces the Eigen CPU matmuls behind tf.matmul (example.py:95,97) and their
 * autodiff gradients (example.py:118).
 * ---------------------------------------------------------------------------------- */
enum { B200PS_DT_BF16 = 0, B200PS_DT_F32 = 1 }; /* F32 operands run as kind::tf32 */
enum { B200PS_ACT_NONE = 0, B200PS_ACT_SIGMOID = 1 };
enum {
  B200PS_EPI_STORE = 0,      /* C = act(alpha*acc + bias) * dsigmoid(aux)              */
  B200PS_EPI_ATOMIC_ADD = 1, /* C += alpha*acc  (fp32 C, red.global.add; for split-K)  */
  B200PS_EPI_SGD = 2         /* master -= lr*alpha*acc ; lowp = bf16(master) if lowp   */
};

typedef struct b200ps_gemm_args {
  int32_t M, N, K;
  int32_t ab_dtype; /* B200PS_DT_* of A and B */
  int32_t a_mn_major, b_mn_major;
  const void* A;
  int64_t lda;
  const void* B;
  int64_t ldb;
  void* C; /* may be NULL for B200PS_EPI_SGD */
  int64_t ldc;
  int32_t c_dtype; /* B200PS_DT_* */
  int32_t epilogue;
  float alpha;
  const float* bias; /* [N] or NULL */
  int32_t act;
  const void* aux; /* [M,N] activation a (c_dtype); C *= a*(1-a). NULL = off */
  int64_t ldaux;
  int32_t split_k;     /* >= 1; > 1 requires B200PS_EPI_ATOMIC_ADD */
  float* sgd_master;   /* [M,N] fp32, row stride ldc (EPI_SGD) */
  void* sgd_lowp;      /* [M,N] bf16 shadow, row stride ldc, or NULL (EPI_SGD) */
  float sgd_lr;
  int32_t max_ctas;    /* 0 = one CTA per SM */
  /* Resident-dataset addressing: A is rows [*a_row_offset, +M) (K-major) or [*a_row_offset, +K)
   * (MN-major) of a stored array with a_rows_total rows; the offset is read ON THE DEVICE at run
   * time, so one captured launch can walk a dataset.  NULL / 0 = plain A. */
  const int32_t* a_row_offset;
  int64_t a_rows_total;
  /* Deterministic split-K: with B200PS_EPI_STORE and split_k > 1, split s stores its partial
   * product to C + s*c_split_stride (elements); the consumer sums the slabs in a fixed order
   * (b200ps_mlp_tail does).  The actual number of splits is min(split_k, ceil(K / block_k)). */
  int64_t c_split_stride;
  /* K2 — parameter pull fused into the GEMM that consumes the weights.  B ([K,N] row-major bf16,
   * ldb == N) is row-sharded over gather_world ranks: gather_src[r] (DEVICE array of gather_world
   * pointers) is rank r's copy of B, valid in rows [r*K/world, (r+1)*K/world) and peer-mapped;
   * args->B is the local copy.  gather_ctas CTAs (default 8) of the same launch stream the remote
   * row panels into the local copy with bulk-async copies over NVLink and publish them panel by
   * panel; the GEMM CTAs start on the local panels and consume remote ones as they land.  On
   * return of the kernel the local copy is complete (the all-gather happened as a side effect).
   * gather_flags: (K/64 + 1) zero-initialised uint32 in local memory, re-armed by the kernel.
   * The owners' slices must be final before the launch (b200ps_push_sync_sgd's flag round does it). */
  const void* const* gather_src;
  uint32_t* gather_flags;
  int32_t gather_world, gather_rank, gather_ctas;
  /* Fused bias gradient: colsum_out[n] = sum_m C[m,n] of the values this call stores (fp32, overwritten;
   * B200PS_EPI_STORE, M % 128 == 0).  NULL = off. */
  float* colsum_out;
} b200ps_gemm_args;

int b200ps_gemm(const b200ps_gemm_args* args, b200ps_stream_t stream);

/* ------------------------------------------------------------------------------------
 * K6/K8: fused MLP head for narrow outputs (C <= 32, H <= 256): bias1 + sigmoid,
 * head GEMM, stable log-softmax cross-entropy, accuracy count and, when train != 0,
 * dZ2 (for the first-layer wgrad GEMM), dW2, db2, db1.  (example.py:96-103,127-128)
 *   z2 [B,H] is the first-layer pre-activation WITHOUT bias.  loss_sum and correct are
 *   accumulated into (+=) unless B200PS_TAIL_OVERWRITE_STATS.  grad_scale multiplies dZ3
 *   (1/global_batch for a mean loss).  labels are read at labels[*label_offset + row] when
 *   label_offset != NULL (device scalar, resident datasets).
 * ---------------------------------------------------------------------------------- */
enum {
  B200PS_TAIL_OVERWRITE_STATS = 1, /* *loss_sum = loss, *correct = count (no += , no zeroing pass) */
  B200PS_TAIL_CLEAR_Z2 = 2         /* zero z2 after reading it (ready for the next split-K GEMM)  */
};
int b200ps_mlp_tail(const float* z2, const float* b1, const float* w2, const float* b2,
                    const int32_t* labels, float* a2_out, float* dz2, float* dw2, float* db2,
                    float* db1, float* loss_sum, int32_t* correct, float* probs_out, int32_t B,
                    int32_t H, int32_t C, int32_t train, float grad_scale, float loss_scale,
                    int32_t flags, const int32_t* label_offset, int32_t z2_splits,
                    int64_t z2_split_stride, b200ps_stream_t stream);
/* z2_splits > 1: z2 is a stack of split-K partial slabs, z2_split_stride elements apart, summed
 * here in slab order (bitwise reproducible, unlike an atomic split-K). */

/* The same head computation, for callers that may not allocate (no counterpart in the
 * reference, which has no device code at all).  b200ps_mlp_tail draws its cross-CTA scratch
 * from a cudaMalloc'ed, growable library workspace; a host process that has committed all
 * device memory to its own plan, or a latency-bound loop that cannot absorb cudaMalloc's
 * device-wide synchronisation, can call this variant instead: the scratch is a fixed 1 MiB
 * array inside the library's device image (held by every GPU that loads the library).
 * Same arguments and results as b200ps_mlp_tail.  Limits and rules:
 *   - ctas * (H*C + C + H + 2) * 4 bytes must fit in 1 MiB (ctas = ceil(B/8) capped at the SM
 *     count); otherwise B200PS_ERR_INVALID is returned and nothing is launched;
 *   - the array is shared by all calls: launches that could overlap on the GPU must be
 *     serialised by the caller;
 *   - the first call on a device performs a one-time address look-up; make it at start-up,
 *     before beginning any stream capture;
 *   - never calls cudaMalloc/cudaFree or any allocating routine. */
int b200ps_mlp_tail_static(const float* z2, const float* b1, const float* w2, const float* b2,
                           const int32_t* labels, float* a2_out, float* dz2, float* dw2,
                           float* db2, float* db1, float* loss_sum, int32_t* correct,
                           float* probs_out, int32_t B, int32_t H, int32_t C, int32_t train,
                           float grad_scale, float loss_scale, int32_t flags,
                           const int32_t* label_offset, int32_t z2_splits, int64_t z2_split_stride,
                           b200ps_stream_t stream);

/* Wide-head softmax cross-entropy: logits [B,C] (ld = row stride) -> dlogits (same dtype,
 * may alias logits, may be NULL), loss_sum += , correct += .  C <= 16384. */
int b200ps_softmax_ce(const void* logits, const int32_t* labels, void* dlogits, float* loss_sum,
                      int32_t* correct, int32_t B, int32_t C, int64_t ld, int32_t dtype,
                      float grad_scale, float loss_scale, b200ps_stream_t stream);

/* out[n] = scale * sum_m x[m,n]  (bias gradients). out is fp32 and is overwritten. */
int b200ps_colsum(const void* x, float* out, int32_t M, int32_t N, int64_t ld, int32_t dtype,
                  float scale, b200ps_stream_t stream);

/* K1 (single shard): params -= lr*grad_scale*grads ; lowp = bf16(params) if lowp != NULL.
 * (ApplyGradientDescent, example.py:108) */
int b200ps_sgd_apply(float* params, const void* grads, int32_t grad_dtype, void* lowp, uint64_t n,
                     float lr, float grad_scale, b200ps_stream_t stream);

/* K11: Adam with TF's ApplyAdam formula (model_export.py:38). beta_pow = {beta1^t, beta2^t}
 * on the device; call b200ps_adam_advance once per step after all variables are applied. */
int b200ps_adam_apply(float* params, const float* grads, float* m, float* v, uint64_t n, float lr,
                      float beta1, float beta2, float eps, const float* beta_pow,
                      b200ps_stream_t stream);
int b200ps_adam_advance(float* beta_pow, float beta1, float beta2, b200ps_stream_t stream);

/* Linear regression forward + gradients (model_export.py:35-36): loss = sum((y-yp)^2)*inv_n. */
int b200ps_linreg_fwd_bwd(const float* x, const float* y, const float* w, const float* b,
                          float* ypred, float* loss, float* gw, float* gb, int32_t B, int32_t D,
                          float inv_n, b200ps_stream_t stream);

/* ------------------------------------------------------------------------------------
 * K9: on-device synthetic input pipelines (replace example.py:60-62,165; lr2.py:87-155;
 * input_pipeline*.py).  Pure functions of (seed, offset + row).
 * ---------------------------------------------------------------------------------- */
int b200ps_synth_mnist(float* x, float* onehot, int32_t* labels, int32_t B, int32_t D, int32_t C,
                       uint64_t seed, uint64_t offset, b200ps_stream_t stream);
int b200ps_synth_dense_bf16(void* x, int32_t* labels, int32_t B, int32_t D, int32_t C,
                            uint64_t seed, uint64_t offset, b200ps_stream_t stream);
/* CSR batch: row_ptr[B+1], fid/fval sized for B*(nnz_mean+spread). skew = 1 uniform ids. */
int b200ps_synth_sparse(int32_t* row_ptr, int64_t* fid, float* fval, float* labels, int32_t B,
                        int32_t nnz_mean, int32_t spread, int64_t F, float skew, uint64_t seed,
                        uint64_t offset, b200ps_stream_t stream);
/* tf.random_normal initialiser (example.py:84-85, lr2.py:384), Philox + Box-Muller.
 * out[i] is element (elem_offset + i) of stream stream_id, so shards of one variable can be
 * initialised independently (elem_offset % 4 == 0). */
int b200ps_init_normal(float* out, uint64_t n, float mean, float stddev, uint64_t seed,
                       uint64_t stream_id, uint64_t elem_offset, b200ps_stream_t stream);
int b200ps_cast_f32_bf16(const float* in, void* out, uint64_t n, b200ps_stream_t stream);

/* ------------------------------------------------------------------------------------
 * K13: the whole step of the narrow sigmoid MLP (example.py:95-118) in ONE launch: forward, stable
 * log-softmax CE + accuracy, backward, and either the SGD update in place (apply = 1: one replica,
 * also *global_step += 1 and the cursor advance), the gradients for a following push (apply = 0), or the
 * whole synchronous-replica push inside the same launch (apply = 2: gradients leave as {value, epoch}
 * packets for their owners over NVLink, owners reduce + apply + publish, replicas unpack; one rank per
 * GPU, every rank launches the same step; mixes freely with b200ps_push_sync_ll on the same tables), or the
 * asynchronous (Hogwild) worker step with no local parameter copy (apply = 3: parameters are fetched from
 * the owners' masters inside the launch, -lr * gradient is added into them with unlocked red.add over NVLink,
 * the chief's global_step counts the update; needs only masters, bounds, global_steps, world and the offsets).
 * Three phases of one co-resident grid separated by two device-wide barriers; exact fp32 FMA in a
 * fixed order (bit-reproducible).  x/labels may be a resident dataset addressed through the device
 * scalar `cursor` (rows [*cursor, *cursor+B)).  Limits: B <= 128, H <= 128 and H % 4 == 0, C <= 16,
 * D % 16 == 0.  Uses the library workspace (see b200ps_mlp_tail).  loss/correct are overwritten and
 * may point into pinned, device-mapped host memory.
 * ---------------------------------------------------------------------------------- */
/* In-kernel push of b200ps_mlp_fused_step (apply = 2): the tables of b200ps_push_sync_ll for the fp32 group
 * that holds the four variables, this rank's index, and the flat offsets of the variables in the group. */
typedef struct b200ps_fused_push {
  uint32_t* const* pads;          /* [world] signal pads */
  float* const* masters;          /* [world] fp32 master == working copy, flat */
  const int64_t* bounds;          /* [world+1] owner ranges (multiples of 8 elements) */
  void* const* ll_grads;          /* [world] packet landing buffers, world * max_shard_elems packets each */
  void* const* ll_params;         /* [world] packet landing buffers, total elements each */
  uint64_t* const* global_steps;  /* [world] or NULL */
  int64_t max_shard_elems;
  int64_t off_w1, off_b1, off_w2, off_b2;
  int32_t world, rank;
} b200ps_fused_push;
int b200ps_mlp_fused_step_set_timing(void* device_buf_24_int64 /* debug: CTA-0 SM-clock stamps; NULL = off */);
int b200ps_mlp_fused_step(const float* x, const int32_t* labels, const int32_t* cursor, float* w1, float* b1,
                          float* w2, float* b2, float* gw1, float* gb1, float* gw2, float* gb2, float* loss,
                          int32_t* correct, uint64_t* global_step, int32_t* cursor_rw, int32_t cursor_step,
                          int32_t cursor_wrap, int32_t B, int32_t D, int32_t H, int32_t C, int32_t apply,
                          float lr, float grad_scale, float loss_scale, const b200ps_fused_push* push,
                          b200ps_stream_t stream);

/* K14: the same narrow-MLP step (example.py:95-118, single replica: SGD applied in place) as a RESIDENT
 * thread-block cluster: one launch of 16 CTAs runs `steps` consecutive steps; W1 / W2 / b1 / b2 stay in shared
 * memory between steps (written back on the last one), the products run on the tensor cores (mma.sync tf32 with
 * the 3xTF32 split = fp32-grade accuracy; flags bit 0 selects plain tf32) and all cross-CTA exchange is
 * distributed shared memory + barrier.cluster.  Step s reads rows [(cursor + s*cursor_step) % cursor_wrap, +B)
 * of x / labels (cursor = *cursor, or 0 when null) and writes loss[s*stat_stride], correct[s*stat_stride];
 * *global_step += steps; *cursor_rw = advanced cursor.  Uses no scratch memory.  b200ps_mlp_cluster_supported
 * reports whether shape and device (a schedulable 16-CTA cluster with ~200 KB shared memory per CTA) qualify. */
int b200ps_mlp_cluster_supported(int32_t B, int32_t D, int32_t H, int32_t C, int32_t* ok);
int b200ps_mlp_cluster_steps(const float* x, const int32_t* labels, const int32_t* cursor, float* w1, float* b1,
                             float* w2, float* b2, float* loss, int32_t* correct, int32_t stat_stride,
                             uint64_t* global_step, int32_t* cursor_rw, int32_t cursor_step, int32_t cursor_wrap,
                             int32_t B, int32_t D, int32_t H, int32_t C, int32_t steps, float lr, float grad_scale,
                             float loss_scale, int32_t flags, b200ps_stream_t stream);
int b200ps_mlp_cluster_set_timing(void* buf);
/* Debug: random per-CTA delays (< ns nanoseconds) inside every following b200ps_mlp_fused_step launch; 0 = off. */
int b200ps_debug_set_skew_ns(int32_t ns);

/* ------------------------------------------------------------------------------------
 * K4/K5: sparse logistic regression over a row-sharded [F,1] weight (lr2.py:390-396).
 * w_shards: device array of `world` shard base pointers (peer-mapped); feature id f lives
 * at w_shards[f / rows_per_shard][f % rows_per_shard].  do_forward computes z, p, loss, dz;
 * do_update scatters -lr*dz*fval with red.global.add (and -lr*sum(dz) into *bias, and
 * *global_step += 1).  Both set = one fused Hogwild step.  Update-only launches read dz_in.
 * Ids outside [0, features) are refused inside the kernel (they contribute nothing and never
 * touch memory) and counted per device; b200ps_sparse_bad_ids reads (and optionally clears) the
 * count — the host-side analogue of embedding_lookup_sparse's InvalidArgument (it synchronises).
 * K10: streaming-AUC histograms (lr2.py:400): bucket = #thresholds strictly below pred.
 * ---------------------------------------------------------------------------------- */
int b200ps_sparse_lr(const int32_t* row_ptr, const int64_t* fid, const float* fval,
                     const float* labels, float* const* w_shards, float* bias, float* pred,
                     float* dz_out, const float* dz_in, float* loss_sum, uint64_t* global_step,
                     int32_t B, int64_t rows_per_shard, int64_t features, int32_t do_forward,
                     int32_t do_update, float lr, float grad_scale, float loss_scale, b200ps_stream_t stream);
int b200ps_sparse_bad_ids(uint64_t* count, int32_t reset);
/* K4u/K5u: the same step over the batch's UNIQUE ids (lr2.py:390: embedding_lookup_sparse gathers each distinct
 * id once; lr2.py:395-396: one IndexedSlices entry per distinct row goes back).  One cooperative launch, one CTA per
 * SM: (do_unique) hash pass — every id into an epoch-tagged open-addressing table, distinct ids appended to the
 * inserting CTA's segment of `uniq`; gather the distinct weights into wbuf; walk the examples on local buffers
 * accumulating dz*val into gbuf; one red.add per distinct id.  A synchronous step is (do_unique=1, forward, finish=0)
 * ... barrier ... (do_unique=0, update, finish=1) with the same B and cap; `finish` advances the table's epoch, so
 * the table never needs clearing.
 * Scratch, caller-owned, zero-filled once with state[1] = 1: keys [table_size] u64, slot_of [table_size] i32,
 * hpos [cap] i32, uniq / wbuf [list_cap], gbuf [b200ps_sparse_dedup_grad_copies() * list_cap], seg_cnt [2 * #SMs] u32 (distinct ids per CTA segment: their sum is the
 * batch's distinct-id count), state [16] u32.  table_size = b200ps_sparse_dedup_table_size(cap) (a power of two
 * >= 4*cap), list_cap = b200ps_sparse_dedup_list_size(cap).  Ids must fit 40 bits. */
int64_t b200ps_sparse_dedup_list_size(int64_t cap);
int32_t b200ps_sparse_dedup_grad_copies(void);
int64_t b200ps_sparse_dedup_table_size(int64_t cap);
int b200ps_sparse_lr_unique(const int32_t* row_ptr, const int64_t* fid, const float* fval, const float* labels,
                            float* const* w_shards, float* bias, float* pred, float* dz_out, const float* dz_in,
                            float* loss_sum, uint64_t* global_step, int32_t B, int32_t cap, int64_t rows_per_shard,
                            int64_t features, int32_t do_unique, int32_t do_forward, int32_t do_update, int32_t finish,
                            float lr, float grad_scale, float loss_scale, uint64_t* keys, int32_t* slot_of,
                            int64_t table_size, int32_t* hpos, int64_t* uniq, float* wbuf, float* gbuf, int64_t list_cap,
                            uint32_t* seg_cnt, uint32_t* state, b200ps_stream_t stream);
/* K4w/K5w: the wide & deep extension sketched at lr2_debug.py:423-428 —
 *   z = sum_j W[fid_j]*val_j + (sum_j val_j*E[fid_j,:]) . w2 + b
 * over row-sharded tables W [F,1] (optional: wide_shards may be NULL), E [F,dim] (dim 128/256/512) and
 * w2 [dim,1] (rows_per_shard a multiple of 4), addressed like w_shards above.  One fused Hogwild launch
 * (do_forward && do_update) gathers, evaluates the loss and scatters all four updates; forward-only
 * evaluates; update-only re-gathers h and applies dz_in (the synchronous-replica split). */
int b200ps_sparse_wd(const int32_t* row_ptr, const int64_t* fid, const float* fval, const float* labels,
                     float* const* wide_shards, int64_t wide_rows_per_shard, float* const* emb_shards,
                     int64_t emb_rows_per_shard, int32_t dim, float* const* w2_shards,
                     int64_t w2_rows_per_shard, float* bias, float* pred, float* dz_out, const float* dz_in,
                     float* loss_sum, uint64_t* global_step, int32_t B, int32_t do_forward, int32_t do_update,
                     float lr, float grad_scale, float loss_scale, float* const* pend_shards,
                     int64_t pend_rows_per_shard, float* const* pend2_shards, int64_t pend2_rows_per_shard,
                     int64_t features, b200ps_stream_t stream);
/* Deferred mode (pend_shards != NULL above): the embedding update of a step is rank one, dE[fid,:] = c[fid]*w2,
 * so workers add only the scalar c into a pending table P (sharded like W: 4 bytes of remote atomic per
 * non-zero instead of 516) and the w2 gradient into P2; the OWNER of a shard folds them in with
 * b200ps_sparse_wd_apply: phase 1 = its rows (W += c, E[row,:] += c*w2, plain stores), phase 2 = its slice
 * of w2.  Pointers here are local shard bases.  Call it before a step / before reading the tables. */
int b200ps_sparse_wd_apply(float* pend_local, float* wide_local, float* emb_local, int64_t rows_local, int32_t dim,
                           float* const* w2_shards, int64_t w2_rows_per_shard, float* pend2_local, float* w2_local,
                           int64_t w2_rows_local, int32_t phase, float* const* pend2_shards /* phase 1, owner-side
                           step: also add dw2 = sum_f c_f*E_old[f] into the pending w2 vector; NULL otherwise */,
                           b200ps_stream_t stream);
/* Owner-side wide & deep step across GPUs (csrc/kernels/sparse_wd_owner.cu): no embedding row crosses a link.
 * request: bucket this rank's non-zeros by owner, write {local row, non-zero index} into the owners' inboxes
 * (inbox[o] + rank*cap + slot) and publish the counts (inbox_cnt[o][rank]); `count` = world+1 zeroed device words.
 * serve (after a cross-rank barrier): dot every requested LOCAL row with w2, add the wide weight, write the scalar
 * pair into the requester's resp[owner*cap + slot] (inbox order: a warp's answers are one contiguous remote write;
 * resp holds world*cap pairs; `slot_tab`, filled by request, tells the requester where each non-zero's answer lands).
 * loss (after a barrier): z, loss, dz from the response array; with do_update the coefficients -lr*dz*v go as scalar
 * red.adds into the owners' pending table.  fold (after a barrier): the owner folds the coefficients of the rows its
 * inbox names (W += c, dw2 += c*E_old[row], E[row] += c*w2); after one more barrier b200ps_sparse_wd_apply phase 2
 * folds w2 (and the pending bias).  Replaces the row gather / row scatter of embedding_lookup_sparse (lr2.py:390,
 * lr2_debug.py:423-428) by ~24 bytes of posted writes per non-zero. */
int b200ps_wd_owner_request(const int32_t* row_ptr, const int64_t* fid, int32_t B, int32_t world, int32_t rank,
                            int32_t cap, int64_t rows_per_shard, int64_t features, void* const* inbox,
                            uint32_t* const* inbox_cnt, void* const* resp, uint32_t* count,
                            int32_t* slot_tab /* [cap] local: owner*cap + slot of each non-zero's answer */,
                            b200ps_stream_t stream);
int b200ps_wd_owner_serve(int32_t world, int32_t rank, int32_t cap, void* const* inbox, uint32_t* const* inbox_cnt,
                          void* const* resp, const float* emb_local, const float* wide_local, int32_t dim,
                          float* const* w2_shards, int64_t w2_rows_per_shard, b200ps_stream_t stream);
int b200ps_wd_owner_fold(int32_t world, int32_t rank, int32_t cap, void* const* inbox, uint32_t* const* inbox_cnt,
                         float* pend_local, float* emb_local, float* wide_local, int32_t dim, float* const* w2_shards,
                         float* const* pend2_shards, int64_t w2_rows_per_shard, b200ps_stream_t stream);
int b200ps_wd_owner_loss(const int32_t* row_ptr, const int64_t* fid, const float* fval, const float* labels, int32_t B,
                         int32_t world, int32_t rank, int32_t cap, int64_t rows_per_shard, int64_t features,
                         void* const* resp, const int32_t* slot_tab, float* const* pend_shards, float* bias,
                         float* bias_pend /* pending bias update on the bias owner, folded after the round (NULL:
                         update the bias in place) */, float* pred,
                         float* dz_out, float* loss_sum, uint64_t* global_step, int32_t do_update, float lr,
                         float grad_scale, float loss_scale, b200ps_stream_t stream);
int b200ps_auc_hist(const float* pred, const float* labels, int32_t B, const float* thresholds,
                    int32_t T, uint64_t* hist_pos, uint64_t* hist_neg, b200ps_stream_t stream);

/* ------------------------------------------------------------------------------------
 * JPEG decode for the image input pipeline (input_pipeline.py:68-78: read_file + decode_jpeg(channels=3)).
 * Host: b200ps_jpeg_parse entropy-decodes a baseline JPEG into quantised 8x8 coefficient blocks (natural order,
 * component after component) and fills the descriptor.  GPU: b200ps_jpeg_idct_color dequantises, runs the inverse
 * DCT (one CTA of 64 threads per block) into component planes, then one thread per output pixel upsamples the
 * chroma (libjpeg's "fancy" triangle filters for 2x1 / 2x2), converts YCbCr -> RGB and — when out_h x out_w
 * differs from the image size — resizes bilinearly.  out is uint8 [out_h, out_w, 3] (grey images are replicated
 * to three channels, as decode_jpeg(channels=3) does).  coef_dev / planes_dev: device buffers of
 * desc->total_coefs int16 / bytes.
 * ---------------------------------------------------------------------------------- */
typedef struct b200ps_jpeg_comp {
  int32_t h, v;                 /* sampling factors */
  int32_t blocks_w, blocks_h;   /* blocks per row / column (padded to whole MCUs) */
  int64_t coef_offset;          /* first coefficient of this component in the coefficient buffer */
  int64_t plane_offset;         /* first sample of this component's plane (blocks_h*8 rows of blocks_w*8) */
  uint16_t quant[64];           /* quantisation table, natural order */
} b200ps_jpeg_comp;
typedef struct b200ps_jpeg_desc {
  int32_t width, height, ncomp, hmax, vmax, mcus_x, mcus_y, total_blocks;
  int64_t total_coefs;
  b200ps_jpeg_comp comp[3];
} b200ps_jpeg_desc;
int b200ps_jpeg_info(const uint8_t* data, uint64_t len, int32_t* width, int32_t* height, int32_t* ncomp);
int b200ps_jpeg_parse(const uint8_t* data, uint64_t len, int16_t* coef, uint64_t coef_cap, b200ps_jpeg_desc* desc);
int b200ps_jpeg_idct_color(const int16_t* coef_dev, const b200ps_jpeg_desc* desc, uint8_t* planes_dev,
                           uint8_t* out_dev, int32_t out_h, int32_t out_w, b200ps_stream_t stream);

/* ------------------------------------------------------------------------------------
 * K1/K3/K12 + pull: parameter-server traffic as kernels over peer-mapped memory.
 * All pointer tables are DEVICE arrays with `world` entries holding addresses valid in the
 * calling process (local or IPC-mapped).  bounds[world+1] are the owner ranges of the flat
 * parameter space (multiples of 8 elements).  masters[r] must be indexable by the GLOBAL
 * element index inside r's range.  rank_base/local_ranks: one launch drives ranks
 * [rank_base, rank_base+local_ranks) — >1 only for single-GPU emulation of several ranks.
 * pads[r]: b200ps_signal_pad_words() zero-initialised uint32 words per rank.
 * ---------------------------------------------------------------------------------- */
int b200ps_signal_pad_words(void);
/* sync replicas: reduce-scatter + mean + SGD on the owner + broadcast, bracketed by the
 * "gradients ready" / "slices landed" flag exchanges (example.py:110-123,147-152). */
int b200ps_push_sync_sgd(uint32_t* const* pads, float* const* masters, void* const* outs,
                         const void* const* grads, const int64_t* bounds,
                         uint64_t* const* global_steps, int32_t world, int32_t rank_base,
                         int32_t local_ranks, int64_t max_shard_elems, int32_t grad_dtype,
                         int32_t out_dtype, int32_t replicas_to_aggregate, float lr,
                         int32_t max_ctas, int32_t* cursor, int32_t cursor_step, int32_t cursor_wrap,
                         int32_t broadcast, b200ps_stream_t stream);
/* broadcast = 1: the owner also writes the fresh values into every replica's working copy (the
 * next pull is already done).  broadcast = 0: only the owner's own copy is refreshed — use it when
 * the consumers pull inside their GEMM (b200ps_gemm gather_src, K2). */
/* cursor (device int32, may be NULL): after the update has landed, *cursor = (*cursor + cursor_step)
 * % cursor_wrap — the batch pointer of a resident dataset, advanced inside the captured step. */
/* The same synchronous update for SMALL fp32 models, latency-optimised: values cross NVLink as
 * 8-byte {value, epoch} packets that the receiver polls, so there is no flag round trip, no remote
 * read round trip and no fence (2 one-way hops instead of ~5).  ll_grads[r]: world*max_shard_elems
 * uint2 per rank; ll_params[r]: total_elems uint2 per rank (both symmetric, zero-initialised).
 * All replicas are aggregated (replicas_to_aggregate == world). */
/* K1-NVLS: the synchronous push through NVSwitch multicast.  mc_grads / mc_work are MULTICAST addresses of the
 * fp32 gradient and working-copy buffers (memory bound to an NVLS multicast object, e.g. torch symmetric memory:
 * b200ps/parallel/nvls.py); the owner of [lo, hi) (multiples of 4) does multimem.ld_reduce.add (the switch sums
 * the replicas' gradients), master -= step * sum on its fp32 master (indexed by global element), and
 * multimem.st of the fresh values to every replica.  The caller provides the cross-rank barrier before
 * ("gradients written") and after ("parameters landed"). */
int b200ps_push_sync_nvls(const float* mc_grads, float* mc_work, float* master, int64_t lo, int64_t hi, float step,
                          uint64_t* global_step, int32_t max_ctas, b200ps_stream_t stream);
/* The ParameterStore's own NVLS path (both flag exchanges of the synchronous barrier inside the launch, like
 * b200ps_push_sync_sgd): do_reduce = 1 is the synchronous push over ALL replicas — owner issues multimem.ld_reduce on
 * the gradients' multicast address, applies mean-gradient SGD to its fp32 master and multimem.st's the fresh values to
 * every replica (broadcast = 0: into its own copy only); do_reduce = 0 is the collective pull (owners broadcast their
 * master slices).  fp32 groups pass master == work_local; bf16 groups (dtype B200PS_DT_BF16) reduce in the switch with
 * fp32 accumulation.  [lo, hi) = the owned range, multiples of 8.  E1/E2 of SURVEY 2.5.3 (example.py:95-97,108,118). */
int b200ps_store_nvls(uint32_t* const* pads, const void* mc_grads, void* mc_work, void* work_local, float* master,
                      int64_t lo, int64_t hi, int32_t dtype, int32_t world, int32_t rank, float lr, int32_t do_reduce,
                      int32_t broadcast, uint64_t* global_step, int32_t* cursor, int32_t cursor_step,
                      int32_t cursor_wrap, b200ps_stream_t stream);
int b200ps_push_sync_ll(uint32_t* const* pads, float* const* masters, const void* const* grads,
                        const int64_t* bounds, uint64_t* const* global_steps, void* const* ll_grads,
                        void* const* ll_params, int32_t world, int32_t rank_base, int32_t local_ranks,
                        int64_t max_shard_elems, int64_t total_elems, float lr, int32_t* cursor,
                        int32_t cursor_step, int32_t cursor_wrap, b200ps_stream_t stream);
/* async (Hogwild): red.add of -lr*g into every owner's master; no barrier (example.py:118). */
int b200ps_push_async(float* const* masters, const void* const* grads, const int64_t* bounds,
                      uint64_t* const* global_steps, int32_t world, int32_t rank_base,
                      int32_t local_ranks, int64_t total_elems, float lr, int32_t max_ctas,
                      int32_t* cursor, int32_t cursor_step, int32_t cursor_wrap, b200ps_stream_t stream);
/* pull: copy peer-owned slices of the masters into this rank's working copy. */
int b200ps_pull(float* const* masters, void* const* outs, const int64_t* bounds, int32_t world,
                int32_t rank_base, int32_t local_ranks, int64_t max_shard_elems,
                int32_t out_dtype, int32_t include_self, int32_t max_ctas,
                b200ps_stream_t stream);
int b200ps_barrier(uint32_t* const* pads, int32_t world, int32_t rank_base, int32_t local_ranks,
                   b200ps_stream_t stream);

/* ------------------------------------------------------------------------------------
 * N1: symmetric memory over CUDA IPC (replaces tf.train.Server, example.py:38-40).
 * ---------------------------------------------------------------------------------- */
int b200ps_symm_alloc(uint64_t bytes, void** ptr_out, void* handle_out /* 64 bytes or NULL */);
int b200ps_symm_open(const void* handle /* 64 bytes */, void** ptr_out);
int b200ps_symm_close(void* ptr);
int b200ps_symm_free(void* ptr);
int b200ps_device_count(int32_t* n);
int b200ps_can_access_peer(int32_t dev, int32_t peer, int32_t* ok);

/* ------------------------------------------------------------------------------------
 * Runtime: the host side of one end-to-end step in a single call — two optional H2D copies from
 * pinned memory, replay of a captured step graph (cudaGraphExec_t), D2H copy of the result, and
 * an event record.  Replaces the per-step sess.run(feed_dict=...) round trip (example.py:168-170).
 * ---------------------------------------------------------------------------------- */
int b200ps_host_step(void* dst0, const void* src0, uint64_t bytes0, void* dst1, const void* src1,
                     uint64_t bytes1, void* graph_exec, void* result_host, const void* result_dev,
                     uint64_t result_bytes, void* done_event, b200ps_stream_t stream);

/* HostFeeder: a ring of `slots` (staging buffers, step graph, events).  submit() copies two pinned
 * host buffers into the next slot on a private copy stream and replays the slot's graph on `stream`
 * once they have landed; the copy of unit t+1 overlaps the kernels of unit t.  The graphs write their
 * results (loss words) directly into the slot's pinned, device-mapped result block
 * (b200ps_feeder_result_ptr), so no D2H copy is queued.  wait() blocks on one unit only.  Only the
 * last `slots` tickets can be waited for.  Not thread-safe: one submitting thread per feeder. */
/* n 32-bit words device -> device or pinned+mapped host memory, as a kernel (graph-capturable result delivery). */
int b200ps_copy_words(void* dst, const void* src, int64_t n_words, b200ps_stream_t stream);
int b200ps_feeder_create(void** feeder, int32_t slots, uint64_t bytes0, uint64_t bytes1, uint64_t result_bytes);
int b200ps_feeder_result_ptr(void* feeder, int32_t slot, void** ptr);
int b200ps_feeder_bind(void* feeder, int32_t slot, void* graph_exec, void* dst0, void* dst1);
int b200ps_feeder_submit(void* feeder, const void* src0, const void* src1, b200ps_stream_t stream,
                         int64_t* ticket);
int b200ps_feeder_wait(void* feeder, int64_t ticket, void* out /* result_bytes, may be NULL */);
int b200ps_feeder_stats(void* feeder, double* phase_ns /* 5: wait, copies, hand-off, launch, record */,
                        int64_t* submitted);
int b200ps_feeder_destroy(void* feeder);
int b200ps_event_create(void** ev);
int b200ps_event_synchronize(void* ev);
int b200ps_event_destroy(void* ev);

/* ------------------------------------------------------------------------------------
 * Host-side native helpers (no GPU needed).
 * ---------------------------------------------------------------------------------- */
/* CRC-32C used by tfevents / tensor-bundle files (example.py:154, model_export.py:53). */
uint32_t b200ps_crc32c_extend(uint32_t init_crc, const void* data, uint64_t n);
/* Native batch loader (the LoadDataThread / FIFOQueue / tf.train.batch path of lr2.py:87-169, natively):
 * `threads` reader threads take files[t::threads], parse libsvm lines, keep each with probability
 * `sampling_rate`, and publish *packed batches* of exactly `batch_rows` rows into a ring of `ring`
 * buffers (pinned when `pinned` and a CUDA device exists).  A packed batch = 64-byte header (magic, rows,
 * nnz, capacities, section offsets, epoch, thread, sequence number) + row_ptr int32[rows+1] + fid
 * int64[cap_nnz] + fval f32[cap_nnz] + labels f32[rows], sections 16-byte aligned at fixed offsets
 * (b200ps_packed_layout), so ONE H2D copy feeds b200ps_sparse_lr / b200ps_sparse_wd and a captured graph
 * can keep its device pointers.  epochs <= 0 loops over the files forever (queue mode).
 * next(): 0 = batch, 1 = end of data, 2 = timeout.  Rows whose non-zeros would overflow cap_nnz and the
 * incomplete last batch of each thread are dropped and counted (b200ps_loader_stats). */
int b200ps_loader_create(void** loader, const char* const* files, int32_t nfiles, int32_t threads,
                         int32_t batch_rows, int64_t cap_nnz, int32_t ring, double sampling_rate, uint64_t seed,
                         int32_t epochs, int32_t pinned);
int b200ps_loader_layout(void* loader, int64_t* out7, int32_t* pinned);
int b200ps_packed_layout(int64_t batch_rows, int64_t cap_nnz, int64_t* out7);
int b200ps_loader_next(void* loader, void** buf, int32_t timeout_ms);
int b200ps_loader_release(void* loader, void* buf);
int b200ps_loader_stats(void* loader, int64_t* out7);
int b200ps_loader_destroy(void* loader);

/* TensorBoard event files, natively (the EventsWriter behind tf.train.SummaryWriter, example.py:131-135,154,171):
 * add_scalars() encodes Event{wall_time, step, summary{value{tag, simple_value}}}, frames it as a TFRecord and
 * queues it; a writer thread owns the file and flushes at most every flush_secs.  wall_time < 0 = now. */
int b200ps_events_open(void** writer, const char* path, double flush_secs);
int b200ps_events_add_scalars(void* writer, double wall_time, int64_t step, const char* const* tags,
                              const float* values, int32_t n);
int b200ps_events_flush(void* writer);
int b200ps_events_close(void* writer);

/* TensorFlow tensor-bundle (Saver V2) checkpoints, natively (the BundleWriter / BundleReader of tf.train.Saver,
 * model_export.py:53): <prefix>.index (leveldb-style table: "" -> header, name -> {dtype, shape, offset, size,
 * masked crc32c}) + <prefix>.data-00000-of-00001 (raw little-endian bytes).  add() streams the caller's buffer
 * to disk with one CRC pass; finish() writes the sorted index and renames both files into place (the writer
 * is destroyed by finish or abort).  dtype = TensorFlow DataType enum (1 float32, 2 float64, 3 int32, 9 int64,
 * 14 bfloat16, ...).  The reader lists entries sorted by name and verifies block and data checksums. */
int b200ps_bundle_writer_create(void** writer, const char* prefix);
int b200ps_bundle_writer_add(void* writer, const char* name, int32_t dtype, const int64_t* shape, int32_t ndim,
                             const void* data, uint64_t nbytes);
int b200ps_bundle_writer_finish(void* writer);
int b200ps_bundle_writer_abort(void* writer);
int b200ps_bundle_reader_open(void** reader, const char* prefix);
int b200ps_bundle_reader_count(void* reader, int32_t* n);
int b200ps_bundle_reader_entry(void* reader, int32_t i, char* name, int32_t name_cap, int32_t* dtype,
                               int64_t* shape8, int32_t* ndim, uint64_t* nbytes);
int b200ps_bundle_reader_read(void* reader, int32_t i, void* dst, uint64_t nbytes, int32_t verify);
int b200ps_bundle_reader_close(void* reader);

/* libsvm text -> CSR (lr2.py:57-84).  Count first, then parse into caller-sized arrays.
 * parse returns 0 or -(1-based line number) of the first malformed line. */
int b200ps_libsvm_count(const char* buf, uint64_t len, uint64_t* rows, uint64_t* nnz);
int64_t b200ps_libsvm_parse(const char* buf, uint64_t len, float* labels, int32_t* row_ptr,
                            int64_t* fid, float* fval, uint64_t max_rows, uint64_t max_nnz);

#ifdef __cplusplus
}
#endif
#endif /* B200PS_H_ */
