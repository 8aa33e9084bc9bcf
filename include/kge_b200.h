/*
 * kge_b200.h -- C-ABI of the B200-native KGE hot path (libkge_b200.so, sm_100a).
 *
 * Drop-in boundary for the data-parallel hot path of
 * NguyenThaiHoc1/CustomKnowledgeGraphEmbedding.  The reference has no native code and
 * therefore no FFI of its own; each entry point below names the reference *Python*
 * interface whose body it replaces (paths relative to /root/reference; "U<n>" rows are
 * the absent PyTorch submodule KnowledgeGraphEmbedding@2e440e0f).
 * INTEGRATION.md shows the ctypes binding a maintainer adds on the reference side.
 *
 * Conventions
 *   - plain C: pointers + sizes; no torch / C++ types.  `stream` is a cudaStream_t passed
 *     as void* (NULL = legacy default stream).
 *   - all table / score pointers are DEVICE pointers to contiguous row-major fp32;
 *     indices are int64 at the boundary (as torch.LongTensor / tf.int64).  Entry points
 *     whose name ends in `_host` take HOST pointers for the per-step batch and do the
 *     H2D / D2H copies themselves.
 *   - the caller owns every buffer; nothing here allocates except through the caller's
 *     workspace (`ws`, `ws_bytes`; size it with kge_workspace_bytes()).
 *   - return value: 0 = ok, negative = kge_status; kge_last_error() gives the message
 *     (thread-local).  Kernels are enqueued asynchronously; CUDA launch errors are
 *     reported by the call that made them.
 *   - indices are not range-checked on the device (as tf.gather on GPU); the Python host
 *     layer checks them where the reference's index_select would raise.
 */
#ifndef KGE_B200_H
#define KGE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KGE_B200_ABI_VERSION 13

#if defined(__GNUC__)
#define KGE_API __attribute__((visibility("default")))
#else
#define KGE_API
#endif

typedef enum {
    KGE_OK = 0,
    KGE_ERR_INVALID_ARGUMENT = -1, /* bad enum / size / NULL pointer / dims the model forbids */
    KGE_ERR_WORKSPACE = -2,        /* workspace missing or too small */
    KGE_ERR_CUDA = -3,             /* a CUDA runtime call or launch failed */
    KGE_ERR_UNSUPPORTED = -4
} kge_status;

/* score functions: U3-U7 (TransE..pRotatE) and tensorflow_codes/model.py:207-224 (InterHT) */
typedef enum {
    KGE_TRANSE = 0,
    KGE_DISTMULT = 1,
    KGE_COMPLEX = 2,
    KGE_ROTATE = 3,
    KGE_PROTATE = 4,
    KGE_INTERHT = 5
} kge_model_id;

/* KGEModel.forward modes (U2).  The TF wire ints 0 = head-batch, 1 = tail-batch, 3 = single
 * (compress_data/main.py:65,69; supervisor.py:18) are mapped by the host layer. */
typedef enum {
    KGE_MODE_SINGLE = 0,
    KGE_MODE_HEAD_BATCH = 1,
    KGE_MODE_TAIL_BATCH = 2
} kge_mode;

/* What KGEModel.__init__ fixes (U1; tensorflow_codes/model.py:49-112). */
typedef struct {
    int32_t model;         /* kge_model_id */
    int32_t hidden_dim;    /* H */
    int32_t entity_dim;    /* De = H or 2H (-de) */
    int32_t relation_dim;  /* Dr = H, 2H (-dr) or 3H (-tr) */
    int64_t nentity;       /* E: rows of the entity table this pointer addresses */
    int64_t nrelation;     /* R */
    float gamma;           /* margin */
    float embedding_range; /* (gamma + 2) / H, as fp32 */
    float interht_eps;     /* 0 = tf.linalg.normalize (model.py:215-218); 1e-12 = F.normalize */
    int32_t reserved;
} kge_model_desc;

/* Loss options of train_step (U9; tensorflow_codes/model.py:4-44; supervisor.py:19-23). */
typedef struct {
    int32_t negative_adversarial_sampling; /* -adv */
    int32_t uni_weight;                    /* --uni_weight */
    int32_t detach_weights;                /* 1 = PyTorch (.detach()), 0 = TF graph model (model.py:168-171) */
    int32_t reserved;
    float adversarial_temperature;         /* -a */
    float regularization;                  /* -r (L3), 0 = off */
} kge_loss_opts;

/* torch.optim.Adam defaults: lr, betas (.9,.999), eps 1e-8, eps inside the bias-corrected
 * denominator.  keras_epsilon_mode = 1 gives tf.keras.optimizers.Adam (run.py:111):
 * p -= lr*sqrt(1-b2^t)/(1-b1^t) * m / (sqrt(v) + eps). */
typedef struct {
    float lr;
    float beta1;
    float beta2;
    float eps;
    int32_t step; /* t >= 1: the step being applied */
    int32_t keras_epsilon_mode;
} kge_adam_opts;

KGE_API const char *kge_last_error(void);
KGE_API int kge_abi_version(void);

/* Number of SMs / device name of the current device (diagnostics; bench.py). */
KGE_API int kge_device_info(int *sm_count, int *cc_major, int *cc_minor, char *name, size_t name_len);

/*
 * K1  fused gather-and-score.  Replaces KGEModel.forward (U2-U8) and the gather + score part of
 * TFKGEModel.call (tensorflow_codes/model.py:127-199).
 *   mode SINGLE:      positive [B,3]                -> scores [B,1]   (negative ignored, N := 1)
 *   mode HEAD_BATCH:  positive [B,3], negative [B,N] heads  -> scores [B,N]
 *   mode TAIL_BATCH:  positive [B,3], negative [B,N] tails  -> scores [B,N]
 * modulus: device pointer to the pRotatE modulus scalar (U7), NULL otherwise.
 */
KGE_API int kge_score_forward(const kge_model_desc *desc, int mode, const float *entity, const float *relation,
                      const float *modulus, const int64_t *positive, const int64_t *negative, int64_t B,
                      int64_t N, float *scores, void *stream);

/*
 * K3  backward of K1: given dL/dscore [B,N] accumulate (+=) dL/dtable into dense grad tables
 * (what loss.backward() leaves in .grad, U9; tape.gradient, supervisor.py:25).
 *   deterministic = 0: warp-aggregated red.global.add (order of float adds not fixed)
 *   deterministic = 1: occurrences of every row are reduced in a fixed order (needs workspace)
 * grad_modulus: device scalar accumulated for pRotatE, NULL otherwise.
 */
KGE_API int kge_score_backward(const kge_model_desc *desc, int mode, const float *entity, const float *relation,
                       const float *modulus, const int64_t *positive, const int64_t *negative, int64_t B,
                       int64_t N, const float *dscore, float *grad_entity, float *grad_relation,
                       float *grad_modulus, int deterministic, void *ws, size_t ws_bytes, void *stream);

/*
 * K2  fused loss (U9 body; tensorflow_codes/model.py:12-31).  negative_score [B,N], positive_score [B]
 * raw scores, subsampling_weight [B] (may be NULL iff uni_weight).  Writes
 *   loss3 = {positive_sample_loss, negative_sample_loss, (pos+neg)/2}   (device, 3 floats)
 * and, when non-NULL, dL/dscore into dneg [B,N], dpos [B] for loss = loss3[2].
 * ws: B*2+8 floats of scratch.
 */
KGE_API int kge_loss_forward_backward(const float *negative_score, const float *positive_score,
                              const float *subsampling_weight, int64_t B, int64_t N, const kge_loss_opts *opts,
                              float *loss3, float *dneg, float *dpos, void *ws, size_t ws_bytes, void *stream);

/*
 * K2 without any allocation or caller workspace: kge_loss_forward_backward whose 2B+8 floats of scratch live in a
 * file-scope __device__ array inside the library.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same inputs, outputs and results (bit for bit: the same two kernels) as kge_loss_forward_backward.
 *   - limit: B <= KGE_LOSS_STATIC_MAX_B positives (kge_loss_static_max_b() returns it); a larger B is refused with
 *     KGE_ERR_INVALID_ARGUMENT and a message naming the limit.  The array is 2 * 65536 + 8 floats = 512 KB per GPU that
 *     loads the library, held until the process exits.
 *   - the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine); the first call on a device
 *     asks the runtime for the array's device address (cudaGetSymbolAddress, once per device, with that device current)
 *     -- make that first call outside a stream capture.
 *   - every call on a device shares the array, which keeps the last call's contents; K2 writes every term before it
 *     reads it, so nothing needs clearing.  Calls that OVERLAP on the GPU (different streams, or different host threads
 *     on one device) RACE on the array: serialise them, or use kge_loss_forward_backward with separate workspaces.
 */
#define KGE_LOSS_STATIC_MAX_B 65536
KGE_API int64_t kge_loss_static_max_b(void);
KGE_API int kge_loss_forward_backward_static(const float *negative_score, const float *positive_score,
                                             const float *subsampling_weight, int64_t B, int64_t N,
                                             const kge_loss_opts *opts, float *loss3, float *dneg, float *dpos,
                                             void *stream);

/* L3 regulariser value (tensorflow_codes/model.py:33-39): out[0] = sum |x|^3 over n elements (device
 * scalar, overwritten).  ws: 1024 floats. */
KGE_API int kge_l3_sum_cubes(const float *x, int64_t n, float *out, void *ws, size_t ws_bytes, void *stream);

/*
 * K4  fused dense Adam over one table (optimizer.step(), U9; apply_gradients, supervisor.py:26).
 * g may be NULL (all-zero gradient).  l3_lambda != 0 adds d/dp lambda*|p|^3 = 3*lambda*p*|p| to g.
 * zero_grad != 0 writes zeros back to g (optimizer.zero_grad() folded in).
 */
KGE_API int kge_adam_dense(float *p, float *g, float *m, float *v, int64_t n, const kge_adam_opts *opts, float l3_lambda,
                   int zero_grad, void *stream);

/*
 * K5  filtered full-entity ranking.  Replaces the body of KGEModel.test_step (U10): for every query
 * (triple, mode) score the candidates e in [entity_lo, entity_hi) of the table, skip the filtered
 * candidates, and count  n_greater[q] = #{e != true : score_e >  score_true}
 *                        n_equal[q]   = #{e != true : score_e == score_true}
 * so that the reference's rank is 1 + n_greater (+ ties in [0, n_equal]).
 *   queries [Q,3] int64; mode HEAD_BATCH or TAIL_BATCH for all Q
 *   filter_indptr [Q+1], filter_idx: CSR of candidate ids to skip per query (global ids;
 *     the true entity itself need not be listed).  NULL/NULL = raw (unfiltered) ranking.
 *   entity points at row 0 of the FULL id space; rows outside [entity_lo, entity_hi) are not
 *     read except the rows the queries name (h, t) unless `query_rows` is given.
 *   true_score [Q]: written (accumulate = 0); counts are overwritten (accumulate = 0) or
 *     added to (accumulate = 1, used when a shard is visited in several calls).
 * ws: see kge_workspace_bytes(KGE_WS_EVAL, ...).
 */
KGE_API int kge_eval_rank(const kge_model_desc *desc, int mode, const float *entity, const float *relation,
                  const float *modulus, const int64_t *queries, int64_t Q, int64_t entity_lo, int64_t entity_hi,
                  const int64_t *filter_indptr, const int64_t *filter_idx, float *true_score, int64_t *n_greater,
                  int64_t *n_equal, int accumulate, void *ws, size_t ws_bytes, void *stream);

/*
 * K5 over a ROW-SHARDED entity table (one process per GPU): `entity_shard` holds only the rows
 * [shard_row0, shard_row0 + shard_rows) of the table (desc->nentity stays the global E), candidates
 * [entity_lo, entity_hi) must lie inside the shard, and the rows the queries name travel with them:
 * query_rows [Q,2,entity_dim] = (head row, tail row) of every query, assembled by the caller (owners fill their
 * rows, one all-reduce(sum) over zeros elsewhere -- exact).  Every rank computes the same true_score from the same
 * query_rows with the same kernel, so the integer counters of the shards add up to the single-GPU result.
 * Replaces the per-GPU body of test_step under the sharding north_star names; reference mechanism: replicas only
 * (tensorflow_codes/run.py:8-17).
 */
KGE_API int kge_eval_rank_shard(const kge_model_desc *desc, int mode, const float *entity_shard, int64_t shard_row0,
                                int64_t shard_rows, const float *relation, const float *modulus,
                                const int64_t *queries, const float *query_rows, int64_t Q, int64_t entity_lo,
                                int64_t entity_hi, const int64_t *filter_indptr, const int64_t *filter_idx,
                                float *true_score, int64_t *n_greater, int64_t *n_equal, int accumulate, void *ws,
                                size_t ws_bytes, void *stream);

/*
 * DistMult / ComplEx are dot products over entity_dim coordinates (U4, U5), so kge_eval_rank runs them as
 * tcgen05 GEMM tiles (3xTF32 split operands fed by TMA, fp32 accumulation in TMEM, the same counting epilogue;
 * csrc/kge_eval_mma.cu).  kge_eval_set_tensor_path(0) routes them through the fp32 CUDA-core tiles of the
 * distance models instead (the exact-check runs of the tests compare the two); default 1.  Process-wide.
 */
KGE_API int kge_eval_set_tensor_path(int on);

/*
 * Fused train step, tables and batch resident in HBM.  Replaces the body of KGEModel.train_step (U9) /
 * Trainer.train_step_fn (supervisor.py:15-28): both score passes, the loss, the backward and dense Adam
 * on both tables (+ the pRotatE modulus) in one call.  `mode` HEAD_BATCH or TAIL_BATCH.
 *   deterministic: step variant.  0 (default) and 1 run the second-generation step (fused score+loss+dL/dq
 *     sweep, entity-major gradient gather fused with Adam; fixed summation order, bit-reproducible) whenever it
 *     applies, else the first-generation kernels with the atomic (0) / fixed-order (1) backward;
 *     2 / 3 force the first-generation kernels with the atomic / fixed-order backward.
 *   adam_m_*, adam_v_*: optimizer state, same shapes as the parameters (modulus: 1 float each; NULL unless pRotatE).
 *   loss3 (device, 3 floats): positive_sample_loss, negative_sample_loss, loss (+ reg4[0] when -r).
 *   reg1 (device, 1 float, may be NULL when regularization == 0): the regularisation term.
 */
KGE_API int kge_train_step(const kge_model_desc *desc, int mode, float *entity, float *relation, float *modulus,
                   float *adam_m_entity, float *adam_v_entity, float *adam_m_relation, float *adam_v_relation,
                   float *adam_m_modulus, float *adam_v_modulus, const int64_t *positive, const int64_t *negative,
                   const float *subsampling_weight, int64_t B, int64_t N, const kge_loss_opts *loss_opts,
                   const kge_adam_opts *adam_opts, int deterministic, float *loss3, float *reg1, void *ws,
                   size_t ws_bytes, void *stream);

/* Same step with the batch in HOST memory (pinned for async copies): copies positive / negative /
 * weights H2D into the workspace, runs kge_train_step, copies loss3 (+reg) back to host_loss4 and
 * synchronises the stream.  This is the end-to-end call bench.py times as `e2e`. */
KGE_API int kge_train_step_host(const kge_model_desc *desc, int mode, float *entity, float *relation, float *modulus,
                        float *adam_m_entity, float *adam_v_entity, float *adam_m_relation,
                        float *adam_v_relation, float *adam_m_modulus, float *adam_v_modulus, const int64_t *host_positive,
                        const int64_t *host_negative, const float *host_subsampling_weight, int64_t B, int64_t N,
                        const kge_loss_opts *loss_opts, const kge_adam_opts *adam_opts, int deterministic,
                        float *host_loss4, void *ws, size_t ws_bytes, void *stream);

/* The two halves of kge_train_step_host for a caller that keeps the NEXT batch in flight while the current step runs
 * (KGEModel.train_step's input lookahead: upstream's `next(train_iterator)` + `.cuda()`, train_step U9, issued one step
 * early on a copy stream):
 *   kge_stage_batch  : the three H2D copies of a batch (pinned host memory for true asynchrony) on `stream`;
 *                      host_subsampling_weight / dev_subsampling_weight may both be NULL (--uni_weight);
 *   kge_stream_wait_event : `stream` waits for a CUDA event (the one recorded after kge_stage_batch on the copy stream);
 *   kge_fetch_floats : n floats device -> host on `stream`, then waits for the stream (the `.item()` of the loss log). */
KGE_API int kge_stage_batch(const int64_t *host_positive, const int64_t *host_negative,
                            const float *host_subsampling_weight, int64_t *dev_positive, int64_t *dev_negative,
                            float *dev_subsampling_weight, int64_t B, int64_t N, void *stream);
KGE_API int kge_stream_wait_event(void *stream, void *event);
KGE_API int kge_fetch_floats(float *host_dst, const float *dev_src, int64_t n, void *stream);

/*
 * Candidate sweep of the fused step: variant 1 (default) = split sweep, one warp per candidate row with the
 * per-item partial dL/dq merged in fixed order (TransE, DistMult, ComplEx, RotatE, pRotatE with H <= 1024, H % 4 == 0);
 * variant 0 = one CTA per positive over row groups (all models; what the split sweep falls back to).  Process-wide;
 * both are bit-reproducible run to run, they differ from each other in summation order only.
 */
KGE_API int kge_train_set_sweep_variant(int variant);

/*
 * Multi-GPU split of the step (one process per GPU; the caller does the NCCL plumbing between the two calls).
 * The step shards twice: positives are data-parallel in the SWEEP, entity rows are owner-parallel in the APPLY.
 *
 *   kge_train_sweep  : scores, loss terms and dL/dq for the B LOCAL positives; writes the per-positive buffers
 *                      (S [B,N] dL/dscore of the negatives, Q [B,planes*H] staged planes, GH/GT [B,De], GR [B,Dr] compact gradient
 *                      rows, GM [B], stats [B,2], terms [B,2]).  wsum: device scalar = sum of the subsampling
 *                      weights over the GLOBAL batch (all-reduced by the caller; NULL = local sum);
 *                      B_total = positives of the global batch (uniform weighting).
 *   (caller)         : all-gather every buffer, positive and negative over the ranks -> B_all = sum of B.
 *   kge_train_apply  : on the gathered buffers: the per-entity use lists, then for the entity rows
 *                      [entity_lo, entity_hi) this rank owns: gradient gather + Adam (m, v of other rows are not
 *                      touched); relation rows and the pRotatE modulus are replicated and updated identically on
 *                      every rank (fixed summation order).  loss3 / reg1 as in kge_train_step.
 *   (caller)         : all-gather (or broadcast) the updated row ranges of the entity table.
 * Applies when kge_train_step's second-generation path applies; otherwise KGE_ERR_UNSUPPORTED.
 */
typedef struct {
    float *S, *Q, *GH, *GT, *GR, *GM, *stats, *terms;
} kge_step_buffers;

KGE_API int kge_train_sweep(const kge_model_desc *desc, int mode, const float *entity, const float *relation,
                            const float *modulus, const int64_t *positive, const int64_t *negative,
                            const float *subsampling_weight, int64_t B, int64_t N, const kge_loss_opts *loss_opts,
                            const float *wsum, int64_t B_total, const kge_step_buffers *out, void *ws, size_t ws_bytes,
                            void *stream);

KGE_API int kge_train_apply(const kge_model_desc *desc, int mode, float *entity, float *relation, float *modulus,
                            float *adam_m_entity, float *adam_v_entity, float *adam_m_relation,
                            float *adam_v_relation, float *adam_m_modulus, float *adam_v_modulus,
                            const int64_t *positive_all, const int64_t *negative_all, int64_t B_all, int64_t N,
                            const kge_loss_opts *loss_opts, const kge_adam_opts *adam_opts,
                            const kge_step_buffers *gathered, int64_t entity_lo, int64_t entity_hi, float *loss3,
                            float *reg1, void *ws, size_t ws_bytes, void *stream);

/* kge_train_apply that ALSO stores every updated entity row into the replicas of the table on the other GPUs, so that the
 * broadcast of the owned rows that has to follow kge_train_apply under data parallelism rides on the Adam epilogue:
 *   multicast_entity != NULL: ONE NVSwitch multicast mapping of the table (the multicast address of the symmetric-memory
 *     block that holds it on every rank): every 16-byte chunk of an updated row leaves the SM as one `multimem.st`, the
 *     switch replicates it to all ranks, the local replica included (`npeers` / `peer_entity` are ignored);
 *   else `peer_entity`: HOST array of `npeers` (<= 15) device pointers to the peers' tables (peer mappings valid on this
 *     GPU): one NVLink store per peer and chunk.
 * The caller puts a barrier across the ranks on the stream before anything reads the replicas.  With a non-zero
 * `regularization` the L3 sum reads the whole local table while peers would be writing it: use kge_train_apply and a
 * broadcast then (dist.DataParallelStep does). */
KGE_API int kge_train_apply_peers(const kge_model_desc *desc, int mode, float *entity, float *relation, float *modulus,
                                  float *adam_m_entity, float *adam_v_entity, float *adam_m_relation,
                                  float *adam_v_relation, float *adam_m_modulus, float *adam_v_modulus,
                                  const int64_t *positive_all, const int64_t *negative_all, int64_t B_all, int64_t N,
                                  const kge_loss_opts *loss_opts, const kge_adam_opts *adam_opts,
                                  const kge_step_buffers *gathered, int64_t entity_lo, int64_t entity_hi, int npeers,
                                  float *const *peer_entity, float *multicast_entity, float *loss3, float *reg1,
                                  void *ws, size_t ws_bytes, void *stream);

/*
 * Between kge_train_sweep and kge_train_apply: the exchange of the per-positive buffers over NVLink peer memory,
 * in place of an all-gather.  One kernel stores, for each of the B local positives (global slot rank * B + b), the
 * ids, S, Q, GR, GM, stats, terms into EVERY rank's gathered buffers and GH[b] / GT[b] only into the buffers of the
 * rank that owns the head / tail entity (row ranges of ceil(E / world) rows, as dist.shard_range).  `peer_gathered`,
 * `peer_positive`, `peer_negative` are HOST arrays of `world` entries holding device pointers valid on this GPU
 * (symmetric-memory peer mappings; entry `rank` is the local buffer).  The caller puts a barrier across the ranks on
 * the stream before kge_train_apply.  Replaces the dense gradient all-reduce of the replicas (supervisor.py:26-30).
 */
KGE_API int kge_train_exchange(const kge_model_desc *desc, const int64_t *positive, const int64_t *negative, int64_t B,
                               int64_t N, const kge_step_buffers *local, int rank, int world,
                               const kge_step_buffers *peer_gathered, int64_t *const *peer_positive,
                               int64_t *const *peer_negative, void *stream);

/*
 * kge_train_apply over a ROW-SHARDED entity table: `entity_shard`, `adam_m_shard`, `adam_v_shard` hold only the rows
 * [shard_row0, shard_row0 + shard_rows) (desc->nentity stays the global E; ids in positive_all / negative_all are
 * global); [entity_lo, entity_hi) must lie inside the shard.  With kge_train_sweep run on a compact table of the
 * rows this rank's batch touches (fetched from their owners, ids remapped) this is the row-sharded train step of
 * no rank ever holds the full table or its Adam state.
 */
KGE_API int kge_train_apply_shard(const kge_model_desc *desc, int mode, float *entity_shard, int64_t shard_row0,
                                  int64_t shard_rows, float *relation, float *modulus, float *adam_m_shard,
                                  float *adam_v_shard, float *adam_m_relation, float *adam_v_relation,
                                  float *adam_m_modulus, float *adam_v_modulus, const int64_t *positive_all,
                                  const int64_t *negative_all, int64_t B_all, int64_t N,
                                  const kge_loss_opts *loss_opts, const kge_adam_opts *adam_opts,
                                  const kge_step_buffers *gathered, int64_t entity_lo, int64_t entity_hi,
                                  float *loss3, float *reg1, void *ws, size_t ws_bytes, void *stream);

/*
 * The two steps in front of kge_train_sweep on a ROW-SHARDED table ("all-gather of the touched rows"):
 *
 * kge_shard_compact   the entity ids the local batch touches -> `unique_ids` (sorted ascending, *n_unique of them, a
 *   device scalar: nothing returns to the host) and the batch re-indexed by position in that list
 *   (compact_positive [B,3], compact_negative [B,N]; relation ids are copied).  Bitmap mark + popcount scan: no sort.
 *   capacity = rows of `unique_ids` / of the compact table, >= min(nentity, B*N + 2B).
 *   ws: kge_shard_compact_workspace_bytes(nentity).
 * kge_shard_fetch_rows   compact_table[u] = row unique_ids[u] of the global table, read from the shard of the rank that
 *   owns it: shard_ptrs[r] (HOST array of `world` DEVICE pointers) is the base of rank r's rows
 *   [r * rows_per_rank, (r+1) * rows_per_rank) -- the own shard and peer mappings of the others (symmetric memory /
 *   cudaIpc over NVLink).  The caller orders the call after the owners' previous update (a barrier across ranks).
 * Replaces the torch.unique / all-to-all plumbing; the reference keeps full replicas (tensorflow_codes/run.py:8-17).
 */
KGE_API size_t kge_shard_compact_workspace_bytes(int64_t nentity);
KGE_API int kge_shard_compact(const int64_t *positive, const int64_t *negative, int64_t B, int64_t N, int64_t nentity,
                              int64_t *compact_positive, int64_t *compact_negative, int64_t *unique_ids,
                              int64_t capacity, int32_t *n_unique, void *ws, size_t ws_bytes, void *stream);
KGE_API int kge_shard_fetch_rows(const int64_t *unique_ids, const int32_t *n_unique, int64_t capacity,
                                 int64_t entity_dim, const float *const *shard_ptrs, int world, int64_t rows_per_rank,
                                 float *compact_table, void *stream);

/*
 * N1  on-GPU negative sampler.  Replaces TrainDataset.__getitem__ of the PyTorch codes (U11; TF twin
 * compress_data/main.py:64-90): for every positive (h, r, t) draw N corrupted heads (HEAD_BATCH) or tails
 * (TAIL_BATCH) uniformly among the entities that do not form a known training triple.
 *   true_sorted [ntrue]: sorted int64 composites (r * nentity + fixed) * nentity + candidate of the training triples,
 *     fixed = tail and candidate = head for HEAD_BATCH, fixed = head and candidate = tail for TAIL_BATCH.
 *   (seed, step, first_index + b, slot) fully determine a draw (Philox4x32-10): first_index is the position of
 *     positive[0] in the caller's sample stream, so a batch does not depend on how it is split over ranks.
 *   negative [B,N] int64 (device, overwritten).
 */
KGE_API int kge_sample_negatives(const int64_t *positive, int64_t B, int64_t N, int mode, int64_t nentity,
                                 const int64_t *true_sorted, int64_t ntrue, uint64_t seed, uint64_t first_index,
                                 uint32_t step, int64_t *negative, void *stream);

/*
 * N2  the batch wire format without TensorFlow (host code).  The reference stores one tf.train.Example per batch in
 * TFRecord shards: writer compress_data/utils.py:35-42 + compress_data/main.py:97-125, reader
 * tensorflow_codes/run.py:40-66 (tf.data.TFRecordDataset + tf.io.parse_single_example).
 *   kge_tfrecord_next        one frame of a shard held in memory: 0 = ok, 1 = end of file, negative = error
 *                            (truncated, or CRC-32C mismatch when verify_crc != 0)
 *   kge_tfrecord_frame       data -> header | data | footer (out needs data_len + 16 bytes); returns bytes written
 *   kge_record_parse_example serialized Example -> positive_sample, negative_sample (int64), subsampling_weight
 *                            (float), mode (int64); counts[4] = values found per feature; any map order, packed or
 *                            unpacked lists; a feature larger than its buffer is an error
 *   kge_record_build_example the inverse (create_example): map in sorted key order, packed lists; out == NULL
 *                            returns the size; 0 when cap is too small
 */
KGE_API uint32_t kge_crc32c(const void *data, size_t n);
KGE_API int kge_tfrecord_next(const uint8_t *file, size_t len, size_t offset, size_t *data_offset, size_t *data_len,
                              size_t *next_offset, int verify_crc);
KGE_API size_t kge_tfrecord_frame(const uint8_t *data, size_t data_len, uint8_t *out);
KGE_API int kge_record_parse_example(const uint8_t *buf, size_t len, int64_t *positive, size_t cap_positive,
                                     int64_t *negative, size_t cap_negative, float *weight, size_t cap_weight,
                                     int64_t *mode, size_t cap_mode, size_t *counts);
KGE_API size_t kge_record_build_example(const int64_t *positive, size_t n_positive, const int64_t *negative,
                                        size_t n_negative, const float *weight, size_t n_weight, const int64_t *mode,
                                        size_t n_mode, uint8_t *out, size_t cap);

/*
 * Measurement hooks (bench.py): when enabled, kge_train_step* brackets each of its stages with CUDA
 * events on the caller's stream.  kge_profile_read synchronises those events and returns, per stage, the
 * summed device time in ms and the number of kernel launches since the last reset.  Stage names:
 * kge_profile_stage_name(i), i in [0, kge_profile_num_stages()).
 */
KGE_API int kge_profile_enable(int on);
KGE_API int kge_profile_num_stages(void);
KGE_API const char *kge_profile_stage_name(int stage);
KGE_API int kge_profile_read(float *ms_sum, int64_t *launches, int reset);

/*
 * TranSparse.  Replaces TFKGEModel('TranSparse', ...) of /root/reference/tensorflow_codes/model.py:
 * the variables of :96-106 (W [R, D, D] trainable, mask [R, D, D] fixed 0/1, both float), the gathers of
 * :139-142,161-164,187-190 and the score of :226-235:
 *     score = gamma - sum_j | ph_j * rh_j - pt_j |
 *     ph = normalize(head @ (mask_r o W_r)),  pt = normalize(X @ (mask_r o W_r)),  rh = normalize(relation_r)
 * project_tail = 0: X = head, the reference AS WRITTEN (:229 projects `head` a second time; the tail entity is never
 *   read and tail-batch yields ONE score per positive, [B, 1]);  project_tail = 1: X = tail.
 * entity_dim == relation_dim == dim (the matmul of :227 needs it).  The projections run as 3xTF32 tiles on tcgen05.
 *
 * kge_transparse_forward: mode SINGLE -> positive_score [B];  HEAD_BATCH / TAIL_BATCH -> positive_score [B] AND
 *   negative_score [B, kge_transparse_nout(desc, mode, N)] in one pass (the TF train step calls the model twice per
 *   batch, supervisor.py:17-20: both calls are answered here).  negative [B, N] may be NULL when no candidate row is
 *   read (SINGLE; TAIL_BATCH without project_tail).  Ids must lie in range (not checked on the device).
 *   with_backward != 0 also prepares what kge_transparse_backward needs, in `ws`.
 * kge_transparse_backward: gradients of sum(dpositive * positive_score) + sum(dnegative * negative_score);
 *   grad_entity [E, D], grad_relation [R, D], grad_W [R, D, D] are OVERWRITTEN (dense; rows / relations the batch does
 *   not touch get zeros; grad_W is already multiplied by the mask).  `ws` must be the workspace of the preceding
 *   kge_transparse_forward(with_backward = 1) with the same desc / mode / positive / B / N, untouched in between.
 *   The entity gradient is accumulated with float atomics (order of a row's uses not fixed); everything else is
 *   summed in a fixed order.
 */
typedef struct {
    int64_t nentity;
    int64_t nrelation;
    int64_t dim;
    float gamma;
    int32_t project_tail;
} kge_transparse_desc;

KGE_API int64_t kge_transparse_nout(const kge_transparse_desc *desc, int mode, int64_t N);
KGE_API size_t kge_transparse_workspace_bytes(const kge_transparse_desc *desc, int mode, int64_t B, int64_t N,
                                              int with_backward);
KGE_API int kge_transparse_forward(const kge_transparse_desc *desc, int mode, const float *entity, const float *relation,
                                   const float *W, const float *mask, const int64_t *positive, const int64_t *negative,
                                   int64_t B, int64_t N, float *positive_score, float *negative_score, int with_backward,
                                   void *ws, size_t ws_bytes, void *stream);
KGE_API int kge_transparse_backward(const kge_transparse_desc *desc, int mode, const float *relation, const float *mask,
                                    const int64_t *positive, int64_t B, int64_t N, const float *dpositive,
                                    const float *dnegative, float *grad_entity, float *grad_relation, float *grad_W,
                                    void *ws, size_t ws_bytes, void *stream);

/* The reductions of TFKGEModel.call on raw scores (tensorflow_codes/model.py:145 `log_sigmoid(single_score)`; :168-171,
 * :195-198 `reduce_sum(softmax(s * 1, axis=1) * log_sigmoid(-s), axis=1)` -- temperature 1, NO stop-gradient on the
 * softmax):  score [B, N] (N = 1 with single != 0) -> out [B];  backward: dscore [B, N] = gout[b] * d out[b] / d score. */
KGE_API int kge_tf_call_terms(const float *score, int64_t B, int64_t N, int single, float *out, void *stream);
KGE_API int kge_tf_call_terms_backward(const float *score, const float *gout, int64_t B, int64_t N, int single,
                                       float *dscore, void *stream);

/* Optional device-side range check of a batch (no reference counterpart as a call: tf.gather / index_select raise on an
 * out-of-range id, the row gathers here would read -- and the Adam bucket build write -- out of bounds).  positive
 * [B,3], negative [B,N] (may be NULL with N = 0) on the device; scratch32: 32 device bytes.  SYNCHRONOUS: waits for the
 * stream and returns KGE_ERR_INVALID_ARGUMENT with the count and an example id in kge_last_error().  Switched on per
 * model (KGEModel.check_ids / TranSparseModel.check_ids) while a data pipeline is brought up; off by default. */
KGE_API int kge_check_ids(const int64_t *positive, const int64_t *negative, int64_t B, int64_t N, int64_t nentity,
                          int64_t nrelation, void *scratch32, void *stream);

/* Measurement probe (bench.py `roofline.l2`; no reference counterpart): a persistent grid re-reads `bytes` of `buf`
 * (device, 16-byte aligned, should fit L2: 32 MB) `iters` times with L2-only 128-bit loads; *bytes_read (host) = bytes
 * requested from L2 in total.  The caller times the call with events on `stream`.  sink: one device float. */
KGE_API int kge_probe_l2_read(const void *buf, size_t bytes, int iters, float *sink, int64_t *bytes_read, void *stream);

typedef enum {
    KGE_WS_BACKWARD = 0, /* kge_score_backward (deterministic) */
    KGE_WS_LOSS = 1,
    KGE_WS_EVAL = 2,
    KGE_WS_TRAIN_STEP = 3,
    KGE_WS_TRAIN_STEP_HOST = 4,
    KGE_WS_TRAIN_APPLY = 5, /* kge_train_apply: pass B_all as B */
    KGE_WS_TRAIN_SWEEP = 6  /* kge_train_sweep: per-item partials of the split sweep (ws may be NULL: row-group sweep) */
} kge_ws_kind;

/* Bytes of workspace the call of that kind needs for (desc, B, N) -- for KGE_WS_EVAL pass Q as B and the
 * total filter length as N. */
KGE_API size_t kge_workspace_bytes(int kind, const kge_model_desc *desc, int64_t B, int64_t N);

#ifdef __cplusplus
}
#endif
#endif /* KGE_B200_H */
