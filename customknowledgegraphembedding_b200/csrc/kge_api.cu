// C-ABI of libkge_b200.so (see include/kge_b200.h).  Thin: validates, derives the per-launch
// constants, carves the caller's workspace and enqueues the kernels.  No allocation, no torch.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <string>

#include "kge_internal.h"
#include "kge_train.h"

namespace kge {

static thread_local std::string g_last_error;

int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_last_error = buf;
    return code;
}

int fail_cuda(cudaError_t e, const char* where) {
    return fail(KGE_ERR_CUDA, "%s: %s (%s)", where, cudaGetErrorName(e), cudaGetErrorString(e));
}

int sm_count() {
    static thread_local int cached_dev = -1, cached = 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev != cached_dev) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached = n;
        cached_dev = dev;
    }
    return cached;
}

// Validation mirrors KGEModel.__init__ + InterHT's -de -tr requirement.
int make_params(const kge_model_desc* d, Params* P) {
    if (d == nullptr) return fail(KGE_ERR_INVALID_ARGUMENT, "desc is NULL");
    if (d->model < KGE_TRANSE || d->model > KGE_INTERHT)
        return fail(KGE_ERR_INVALID_ARGUMENT, "model %d not supported", d->model);
    if (d->hidden_dim <= 0 || d->nentity <= 0 || d->nrelation <= 0)
        return fail(KGE_ERR_INVALID_ARGUMENT, "hidden_dim / nentity / nrelation must be positive");
    const int H = d->hidden_dim;
    int need_de = H, need_dr = H;
    switch (d->model) {
        case KGE_ROTATE: need_de = 2 * H; need_dr = H; break;
        case KGE_COMPLEX: need_de = 2 * H; need_dr = 2 * H; break;
        case KGE_INTERHT: need_de = 2 * H; need_dr = 3 * H; break;
        default: break;
    }
    if (d->entity_dim != need_de || d->relation_dim != need_dr)
        return fail(KGE_ERR_INVALID_ARGUMENT,
                    "model %d with hidden_dim %d needs entity_dim %d and relation_dim %d (got %d, %d)", d->model, H,
                    need_de, need_dr, d->entity_dim, d->relation_dim);
    P->model = d->model;
    P->H = H;
    P->De = d->entity_dim;
    P->Dr = d->relation_dim;
    P->gamma = d->gamma;
    // U6: pi = 3.14159265358979323846 ; U7: pi = 3.14159262358979323846 (sic)
    const double pi = (d->model == KGE_PROTATE) ? 3.14159262358979323846 : 3.14159265358979323846;
    P->phase_div = (float)((double)d->embedding_range / pi);
    P->inv_phase_div = 1.0f / P->phase_div;
    P->interht_eps = d->interht_eps;
    P->E = d->nentity;
    P->R = d->nrelation;
    return KGE_OK;
}

static int check_mode(int mode) {
    if (mode != KGE_MODE_SINGLE && mode != KGE_MODE_HEAD_BATCH && mode != KGE_MODE_TAIL_BATCH)
        return fail(KGE_ERR_INVALID_ARGUMENT, "mode %d not supported", mode);
    return KGE_OK;
}

}  // namespace kge

using namespace kge;

extern "C" {

const char* kge_last_error(void) { return g_last_error.c_str(); }
int kge_abi_version(void) { return KGE_B200_ABI_VERSION; }

int kge_device_info(int* sms, int* cc_major, int* cc_minor, char* name, size_t name_len) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail_cuda(e, "cudaGetDevice");
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, dev);
    if (e != cudaSuccess) return fail_cuda(e, "cudaGetDeviceProperties");
    if (sms) *sms = prop.multiProcessorCount;
    if (cc_major) *cc_major = prop.major;
    if (cc_minor) *cc_minor = prop.minor;
    if (name && name_len) {
        strncpy(name, prop.name, name_len - 1);
        name[name_len - 1] = 0;
    }
    return KGE_OK;
}

int kge_score_forward(const kge_model_desc* desc, int mode, const float* entity, const float* relation,
                      const float* modulus, const int64_t* positive, const int64_t* negative, int64_t B, int64_t N,
                      float* scores, void* stream) {
    Params P;
    int rc = make_params(desc, &P);
    if (rc) return rc;
    if ((rc = check_mode(mode))) return rc;
    if (B < 0 || N < 0) return fail(KGE_ERR_INVALID_ARGUMENT, "negative batch size");
    if (mode == KGE_MODE_SINGLE) N = 1;
    if (B == 0 || N == 0) return KGE_OK;
    if (!entity || !relation || !positive || !scores) return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    if (mode != KGE_MODE_SINGLE && !negative) return fail(KGE_ERR_INVALID_ARGUMENT, "negative is NULL");
    if (P.model == KGE_PROTATE && !modulus) return fail(KGE_ERR_INVALID_ARGUMENT, "pRotatE needs modulus");
    cudaError_t e = launch_score_forward(P, mode, entity, relation, modulus, positive, negative, B, N, scores,
                                         (cudaStream_t)stream);
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_score_forward");
}

int kge_score_backward(const kge_model_desc* desc, int mode, const float* entity, const float* relation,
                       const float* modulus, const int64_t* positive, const int64_t* negative, int64_t B, int64_t N,
                       const float* dscore, float* grad_entity, float* grad_relation, float* grad_modulus,
                       int deterministic, void* ws, size_t ws_bytes, void* stream) {
    Params P;
    int rc = make_params(desc, &P);
    if (rc) return rc;
    if ((rc = check_mode(mode))) return rc;
    if (B < 0 || N < 0) return fail(KGE_ERR_INVALID_ARGUMENT, "negative batch size");
    if (mode == KGE_MODE_SINGLE) N = 1;
    if (B == 0 || N == 0) return KGE_OK;
    if (!entity || !relation || !positive || !dscore || !grad_entity || !grad_relation)
        return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    if (mode != KGE_MODE_SINGLE && !negative) return fail(KGE_ERR_INVALID_ARGUMENT, "negative is NULL");
    if (P.model == KGE_PROTATE && !modulus) return fail(KGE_ERR_INVALID_ARGUMENT, "pRotatE needs modulus");
    cudaError_t e;
    if (deterministic) {
        const size_t need = backward_det_workspace_bytes(P, B, N);
        if (!ws || ws_bytes < need)
            return fail(KGE_ERR_WORKSPACE, "deterministic backward needs %zu workspace bytes, got %zu", need, ws_bytes);
        e = launch_score_backward_det(P, mode, entity, relation, modulus, positive, negative, B, N, dscore,
                                      grad_entity, grad_relation, grad_modulus, ws, ws_bytes, (cudaStream_t)stream);
    } else {
        e = launch_score_backward_atomic(P, mode, entity, relation, modulus, positive, negative, B, N, dscore,
                                         grad_entity, grad_relation, grad_modulus, (cudaStream_t)stream);
    }
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_score_backward");
}

int kge_loss_forward_backward(const float* negative_score, const float* positive_score,
                              const float* subsampling_weight, int64_t B, int64_t N, const kge_loss_opts* opts,
                              float* loss3, float* dneg, float* dpos, void* ws, size_t ws_bytes, void* stream) {
    if (!opts || !negative_score || !positive_score || !loss3) return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    if (B <= 0 || N <= 0) return fail(KGE_ERR_INVALID_ARGUMENT, "B and N must be positive");
    if (!opts->uni_weight && !subsampling_weight)
        return fail(KGE_ERR_INVALID_ARGUMENT, "subsampling_weight is NULL without uni_weight");
    if (!ws || ws_bytes < loss_workspace_bytes(B))
        return fail(KGE_ERR_WORKSPACE, "loss needs %zu workspace bytes, got %zu", loss_workspace_bytes(B), ws_bytes);
    cudaError_t e = launch_loss(negative_score, positive_score, subsampling_weight, B, N, *opts, loss3, dneg, dpos,
                                (float*)ws, (cudaStream_t)stream);
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_loss_forward_backward");
}

int kge_l3_sum_cubes(const float* x, int64_t n, float* out, void* ws, size_t ws_bytes, void* stream) {
    if (!x || !out || n < 0) return fail(KGE_ERR_INVALID_ARGUMENT, "bad argument");
    if (!ws || ws_bytes < 1024 * sizeof(float)) return fail(KGE_ERR_WORKSPACE, "l3 needs 4096 workspace bytes");
    cudaError_t e = launch_l3_sum_cubes(x, n, out, (float*)ws, (cudaStream_t)stream);
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_l3_sum_cubes");
}

int kge_adam_dense(float* p, float* g, float* m, float* v, int64_t n, const kge_adam_opts* opts, float l3_lambda,
                   int zero_grad, void* stream) {
    if (!p || !m || !v || !opts || n < 0) return fail(KGE_ERR_INVALID_ARGUMENT, "bad argument");
    if (opts->step < 1) return fail(KGE_ERR_INVALID_ARGUMENT, "adam step must be >= 1");
    cudaError_t e = launch_adam(p, g, m, v, n, adam_scalars(*opts), l3_lambda, zero_grad, (cudaStream_t)stream);
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_adam_dense");
}

// shared body of kge_eval_rank / kge_eval_rank_shard.  shard_row0 < 0: `entity` addresses the full id space.
static int eval_rank_impl(const kge_model_desc* desc, int mode, const float* entity, int64_t shard_row0,
                          int64_t shard_rows, const float* relation, const float* modulus, const int64_t* queries,
                          const float* query_rows, int64_t Q, int64_t entity_lo, int64_t entity_hi,
                          const int64_t* filter_indptr, const int64_t* filter_idx, float* true_score,
                          int64_t* n_greater, int64_t* n_equal, int accumulate, void* ws, size_t ws_bytes, void* stream) {
    Params P;
    int rc = make_params(desc, &P);
    if (rc) return rc;
    if (mode != KGE_MODE_HEAD_BATCH && mode != KGE_MODE_TAIL_BATCH)
        return fail(KGE_ERR_INVALID_ARGUMENT, "negative batch mode %d not supported", mode);
    if (Q < 0 || entity_lo < 0 || entity_hi < entity_lo || entity_hi > P.E)
        return fail(KGE_ERR_INVALID_ARGUMENT, "bad query count or entity range");
    if (Q == 0) return KGE_OK;
    if (!entity || !relation || !queries || !true_score || !n_greater || !n_equal)
        return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    if ((filter_indptr == nullptr) != (filter_idx == nullptr) && filter_indptr == nullptr)
        return fail(KGE_ERR_INVALID_ARGUMENT, "filter_idx without filter_indptr");
    if (P.model == KGE_PROTATE && !modulus) return fail(KGE_ERR_INVALID_ARGUMENT, "pRotatE needs modulus");
    if (shard_row0 >= 0) {
        if (!query_rows) return fail(KGE_ERR_INVALID_ARGUMENT, "a row-sharded table needs query_rows");
        if (entity_lo < shard_row0 || entity_hi > shard_row0 + shard_rows)
            return fail(KGE_ERR_INVALID_ARGUMENT, "candidate range [%lld, %lld) outside the shard [%lld, %lld)",
                        (long long)entity_lo, (long long)entity_hi, (long long)shard_row0,
                        (long long)(shard_row0 + shard_rows));
        // the kernels address candidates by global id: rebase so that id e lands on shard row e - shard_row0.
        // Only rows in [entity_lo, entity_hi) are dereferenced (every other row comes from query_rows).
        entity = reinterpret_cast<const float*>(reinterpret_cast<uintptr_t>(entity) -
                                                (uintptr_t)shard_row0 * (uintptr_t)P.De * sizeof(float));
    }
    const size_t need = eval_workspace_bytes(P, Q, 0);
    if (!ws || ws_bytes < need) return fail(KGE_ERR_WORKSPACE, "eval needs %zu workspace bytes, got %zu", need, ws_bytes);
    cudaError_t e = launch_eval_rank(P, mode, entity, relation, modulus, queries, query_rows, Q, entity_lo, entity_hi,
                                     filter_indptr, filter_idx, true_score, n_greater, n_equal, accumulate, ws,
                                     ws_bytes, (cudaStream_t)stream);
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_eval_rank");
}

int kge_eval_rank(const kge_model_desc* desc, int mode, const float* entity, const float* relation,
                  const float* modulus, const int64_t* queries, int64_t Q, int64_t entity_lo, int64_t entity_hi,
                  const int64_t* filter_indptr, const int64_t* filter_idx, float* true_score, int64_t* n_greater,
                  int64_t* n_equal, int accumulate, void* ws, size_t ws_bytes, void* stream) {
    return eval_rank_impl(desc, mode, entity, -1, 0, relation, modulus, queries, nullptr, Q, entity_lo, entity_hi,
                          filter_indptr, filter_idx, true_score, n_greater, n_equal, accumulate, ws, ws_bytes, stream);
}

int kge_eval_rank_shard(const kge_model_desc* desc, int mode, const float* entity_shard, int64_t shard_row0,
                        int64_t shard_rows, const float* relation, const float* modulus, const int64_t* queries,
                        const float* query_rows, int64_t Q, int64_t entity_lo, int64_t entity_hi,
                        const int64_t* filter_indptr, const int64_t* filter_idx, float* true_score, int64_t* n_greater,
                        int64_t* n_equal, int accumulate, void* ws, size_t ws_bytes, void* stream) {
    if (shard_row0 < 0 || shard_rows < 0) return fail(KGE_ERR_INVALID_ARGUMENT, "bad shard");
    return eval_rank_impl(desc, mode, entity_shard, shard_row0, shard_rows, relation, modulus, queries, query_rows, Q,
                          entity_lo, entity_hi, filter_indptr, filter_idx, true_score, n_greater, n_equal, accumulate, ws,
                          ws_bytes, stream);
}

int kge_sample_negatives(const int64_t* positive, int64_t B, int64_t N, int mode, int64_t nentity,
                         const int64_t* true_sorted, int64_t ntrue, uint64_t seed, uint64_t first_index, uint32_t step,
                         int64_t* negative, void* stream) {
    if (mode != KGE_MODE_HEAD_BATCH && mode != KGE_MODE_TAIL_BATCH)
        return fail(KGE_ERR_INVALID_ARGUMENT, "Training batch mode %d not supported", mode);
    if (B < 0 || N < 0 || nentity <= 0 || ntrue < 0) return fail(KGE_ERR_INVALID_ARGUMENT, "bad sizes");
    if (B == 0 || N == 0) return KGE_OK;
    if (!positive || !negative || (ntrue > 0 && !true_sorted)) return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    cudaError_t e = launch_sample_negatives(positive, B, N, mode == KGE_MODE_HEAD_BATCH, nentity, true_sorted, ntrue, seed,
                                            first_index, step, negative, (cudaStream_t)stream);
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_sample_negatives");
}

int kge_eval_set_tensor_path(int on) {
    eval_set_tensor_path(on);
    return KGE_OK;
}

size_t kge_workspace_bytes(int kind, const kge_model_desc* desc, int64_t B, int64_t N) {
    Params P;
    if (kind == KGE_WS_LOSS) return loss_workspace_bytes(B);
    if (make_params(desc, &P)) return 0;
    switch (kind) {
        case KGE_WS_BACKWARD: return backward_det_workspace_bytes(P, B, N);
        case KGE_WS_EVAL: return eval_workspace_bytes(P, B, N);
        case KGE_WS_TRAIN_STEP: return train_step_workspace_bytes(P, B, N, false);
        case KGE_WS_TRAIN_STEP_HOST: return train_step_workspace_bytes(P, B, N, true);
        case KGE_WS_TRAIN_APPLY: return fused_step_workspace_bytes(P, B, N);
        case KGE_WS_TRAIN_SWEEP: return train_sweep_workspace_bytes(P, B, N);
    }
    return 0;
}

}  // extern "C"
