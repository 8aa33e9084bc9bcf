// Internal launcher declarations (host side) shared by the translation units of libkge_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "kge_common.cuh"

namespace kge {

int sm_count();  // SMs of the current device (cached)

// error plumbing + desc validation (kge_api.cu)
int fail(int code, const char* fmt, ...);
int fail_cuda(cudaError_t e, const char* where);
int make_params(const kge_model_desc* d, Params* P);

// true when every row of both tables can be read with 128-bit loads
inline bool vec4_ok(const Params& P, const void* a, const void* b) {
    return (P.H % 4 == 0) && (P.De % 4 == 0) && (P.Dr % 4 == 0) && ((reinterpret_cast<uintptr_t>(a) & 15) == 0) &&
           ((reinterpret_cast<uintptr_t>(b) & 15) == 0);
}

// K1
cudaError_t launch_score_forward(const Params& P, int mode, const float* ent, const float* rel, const float* modulus,
                                 const int64_t* pos, const int64_t* neg, int64_t B, int64_t N, float* scores,
                                 cudaStream_t st);

// K3 (atomic variant): += into dense grad tables
cudaError_t launch_score_backward_atomic(const Params& P, int mode, const float* ent, const float* rel,
                                         const float* modulus, const int64_t* pos, const int64_t* neg, int64_t B,
                                         int64_t N, const float* dscore, float* grad_ent, float* grad_rel,
                                         float* grad_modulus, cudaStream_t st);

// K3 (deterministic variant)
size_t backward_det_workspace_bytes(const Params& P, int64_t B, int64_t N);
cudaError_t launch_score_backward_det(const Params& P, int mode, const float* ent, const float* rel,
                                      const float* modulus, const int64_t* pos, const int64_t* neg, int64_t B,
                                      int64_t N, const float* dscore, float* grad_ent, float* grad_rel,
                                      float* grad_modulus, void* ws, size_t ws_bytes, cudaStream_t st);

cudaError_t launch_histogram(int mode, const int64_t* pos, const int64_t* neg, long long BN, long long B, int* counts,
                             long long e_lo, long long e_hi, cudaStream_t st);
cudaError_t launch_scan_fill(int mode, const int64_t* pos, const int64_t* neg, long long BN, long long B, long long E,
                             int* counts, int* offsets, int* cursor, int* items, int* block_sums, long long e_lo,
                             long long e_hi, cudaStream_t st);

cudaError_t launch_exclusive_scan(const int* in, int* out, int* block_sums, long long n, cudaStream_t st);

// row-sharded train step (kge_shard.cu): touched rows of a batch -> compact ids, rows pulled from the owners' shards
size_t shard_compact_workspace_bytes(int64_t nentity);
cudaError_t launch_shard_compact(const int64_t* pos, const int64_t* neg, int64_t B, int64_t N, int64_t E, int64_t* cpos,
                                 int64_t* cneg, int64_t* uniq, int64_t capacity, int* n_unique, void* ws, cudaStream_t st);
cudaError_t launch_shard_fetch_rows(const int64_t* uniq, const int* n_unique, int64_t capacity, int64_t De,
                                    const float* const* shards, int world, int64_t rows_per_rank, float* compact,
                                    cudaStream_t st);

// K2
cudaError_t launch_loss_final(const float* terms, int64_t B, float* loss3, cudaStream_t st);
size_t loss_workspace_bytes(int64_t B);
cudaError_t launch_loss(const float* neg_score, const float* pos_score, const float* w, int64_t B, int64_t N,
                        const kge_loss_opts& o, float* loss3, float* dneg, float* dpos, float* ws, cudaStream_t st);
cudaError_t launch_l3_sum_cubes(const float* x, int64_t n, float* out, float* ws, cudaStream_t st);

// K4
struct AdamScalars {
    float beta1, beta2, step_size, inv_sqrt_bc2, eps;
};
AdamScalars adam_scalars(const kge_adam_opts& o);
cudaError_t launch_adam(float* p, float* g, float* m, float* v, int64_t n, const AdamScalars& a, float l3_lambda,
                        int zero_grad, cudaStream_t st);

// K5
size_t eval_workspace_bytes(const Params& P, int64_t Q, int64_t nfilter);
cudaError_t launch_eval_rank(const Params& P, int mode, const float* ent, const float* rel, const float* modulus,
                             const int64_t* queries, const float* query_rows, int64_t Q, int64_t e_lo, int64_t e_hi,
                             const int64_t* filter_indptr, const int64_t* filter_idx, float* true_score,
                             int64_t* n_greater, int64_t* n_equal, int accumulate, void* ws, size_t ws_bytes,
                             cudaStream_t st);

// K5-T (kge_eval_mma.cu): tcgen05 tiles for DistMult / ComplEx
cudaError_t launch_eval_stage(const Params& P, bool head, const float* ent, const float* rel, const int64_t* queries,
                              const float* query_rows, int64_t Q, float* Qs, cudaStream_t st);
void eval_set_tensor_path(int on);
bool eval_mma_applies(const Params& P);
size_t eval_mma_workspace_bytes(const Params& P, int64_t Q, int64_t nfilter);
cudaError_t launch_eval_rank_mma(const Params& P, int mode, const float* ent, const float* rel, const int64_t* queries,
                                 const float* query_rows, int64_t Q, int64_t e_lo, int64_t e_hi, const int64_t* indptr, const int64_t* idx,
                                 float* true_score, unsigned long long* n_greater, unsigned long long* n_equal,
                                 int accumulate, void* ws, size_t ws_bytes, cudaStream_t st);

// N1 (kge_sampler.cu)
cudaError_t launch_sample_negatives(const int64_t* positive, int64_t B, int64_t N, int head, int64_t E,
                                    const int64_t* true_sorted, int64_t ntrue, uint64_t seed, uint64_t first_index,
                                    uint32_t step, int64_t* negative, cudaStream_t st);

}  // namespace kge
