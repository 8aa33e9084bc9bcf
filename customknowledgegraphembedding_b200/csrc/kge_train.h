// Fused train step (host-side composition of the kernels) -- see kge_train.cu / kge_train_fused.cu.
#pragma once
#include "kge_internal.h"

namespace kge {
size_t train_step_workspace_bytes(const Params& P, int64_t B, int64_t N, bool host_batch);
void train_set_sweep_variant(int v);  // 1 = warp-per-row split sweep where it applies (default), 0 = row groups per CTA

// second-generation step (kge_train_fused.cu)
size_t fused_step_workspace_bytes(const Params& P, int64_t B, int64_t N);
bool fused_step_applicable(const Params& P, const kge_loss_opts& lo, int64_t B, int64_t N, const void* ent,
                           const void* rel);
cudaError_t launch_train_step_fused(const Params& P, int mode, float* ent, float* rel, float* modulus, float* m_ent,
                                    float* v_ent, float* m_rel, float* v_rel, float* m_mod, float* v_mod,
                                    const int64_t* pos, const int64_t* neg, const float* w, int64_t B, int64_t N,
                                    const kge_loss_opts& lo, const kge_adam_opts& ao, void* ws, size_t ws_bytes,
                                    float** out_small, cudaStream_t st);
cudaError_t launch_train_sweep(const Params& P, int mode, const float* ent, const float* rel, const float* modulus,
                               const int64_t* pos, const int64_t* neg, const float* w, int64_t B, int64_t N,
                               const kge_loss_opts& lo, const float* wsum, int64_t B_total, const kge_step_buffers& o,
                               void* ws, size_t ws_bytes, cudaStream_t st);
size_t train_sweep_workspace_bytes(const Params& P, int64_t B, int64_t N);
cudaError_t launch_train_apply(const Params& P, int mode, float* ent, float* rel, float* modulus, float* m_ent,
                               float* v_ent, float* m_rel, float* v_rel, float* m_mod, float* v_mod,
                               const int64_t* pos_all, const int64_t* neg_all, int64_t B_all, int64_t N,
                               const kge_loss_opts& lo, const kge_adam_opts& ao, const kge_step_buffers& in, int64_t e_lo,
                               int64_t e_hi, int npeers, float* const* peers, float* mc, void* ws,
                               size_t ws_bytes, float** out_small, cudaStream_t st);
cudaError_t launch_exchange(const Params& P, const int64_t* pos, const int64_t* neg, int64_t B, int64_t N,
                            const kge_step_buffers& local, int rank, int world, const kge_step_buffers* peers,
                            int64_t* const* peer_pos, int64_t* const* peer_neg, cudaStream_t st);
}
