// Fused train step: both score passes, loss, backward and dense Adam enqueued by ONE C-ABI call on one
// stream, nothing returning to the host in between.
//
// Replaces the body of KGEModel.train_step and of Trainer.train_step_fn
// (/root/reference/tensorflow_codes/supervisor.py:15-28).
//
// Workspace carve (256-byte aligned regions):
//   [host-batch only] positive i64[B,3] | negative i64[B,N] | weight f32[B]
//   score_neg f32[B,N] | score_pos f32[B] | dneg f32[B,N] | dpos f32[B]
//   loss scratch f32[2B+8] | l3 scratch f32[1024] | small f32[16] (loss3, reg sums, grad_modulus)
//   grad_entity f32[E,De] | grad_relation f32[R,Dr]   (dense, as .grad in the reference)
//   deterministic-backward scratch
#include "kge_prof.h"
#include "kge_train.h"

namespace kge {

bool g_prof_on = false;
std::vector<ProfRec> g_prof_live;
std::vector<cudaEvent_t> g_prof_pool;
float g_prof_ms[ST_COUNT];
int64_t g_prof_launches[ST_COUNT];

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

struct TrainWs {
    int64_t* pos;
    int64_t* neg;
    float* w;
    float *score_neg, *score_pos, *dneg, *dpos, *loss_ws, *l3_ws, *small, *grad_ent, *grad_rel;
    void* det_ws;
    size_t det_bytes;
    void* fused_ws;
    size_t fused_bytes;
    size_t total;
};

static TrainWs carve(const Params& P, int64_t B, int64_t N, bool host_batch, void* base) {
    TrainWs w{};
    size_t off = 0;
    auto take = [&](size_t bytes) {
        void* p = base ? (void*)((char*)base + off) : nullptr;
        off += align256(bytes);
        return p;
    };
    if (host_batch) {
        w.pos = (int64_t*)take((size_t)B * 3 * sizeof(int64_t));
        w.neg = (int64_t*)take((size_t)B * N * sizeof(int64_t));
        w.w = (float*)take((size_t)B * sizeof(float));
    }
    const size_t v1_begin = off;
    w.score_neg = (float*)take((size_t)B * N * sizeof(float));
    w.score_pos = (float*)take((size_t)B * sizeof(float));
    w.dneg = (float*)take((size_t)B * N * sizeof(float));
    w.dpos = (float*)take((size_t)B * sizeof(float));
    w.loss_ws = (float*)take(loss_workspace_bytes(B));
    w.l3_ws = (float*)take(1024 * sizeof(float));
    w.small = (float*)take(16 * sizeof(float));
    w.grad_ent = (float*)take((size_t)P.E * P.De * sizeof(float));
    w.grad_rel = (float*)take((size_t)P.R * P.Dr * sizeof(float));
    w.det_bytes = backward_det_workspace_bytes(P, B, N);
    w.det_ws = take(w.det_bytes);
    // the second-generation step overlays the same region (only one of the two runs in a call)
    w.fused_ws = w.score_neg;
    w.fused_bytes = fused_step_workspace_bytes(P, B, N);
    const size_t v1_bytes = off - v1_begin;
    if (w.fused_bytes > v1_bytes) off = v1_begin + ((w.fused_bytes + 255) & ~(size_t)255);
    w.total = off;
    return w;
}

size_t train_step_workspace_bytes(const Params& P, int64_t B, int64_t N, bool host_batch) {
    return carve(P, B, N, host_batch, nullptr).total;
}

// small[0..2] = loss3 ; small[3] = reg ; small[4] = sum|E|^3 ; small[5] = sum|R|^3 ; small[8] = grad_modulus
__global__ void finalize_reg_kernel(float* small, float lambda) {
    const float reg = lambda * (small[4] + small[5]);
    small[3] = reg;
    small[2] += reg;
}

static int train_step_device(const Params& P, int mode, float* ent, float* rel, float* modulus, float* m_ent,
                             float* v_ent, float* m_rel, float* v_rel, float* m_mod, float* v_mod, const int64_t* pos,
                             const int64_t* neg, const float* w, int64_t B, int64_t N, const kge_loss_opts& lo,
                             const kge_adam_opts& ao, int variant, const TrainWs& ws, cudaStream_t st,
                             float** out_small) {
    cudaError_t e;
    *out_small = ws.small;
    // variant: 0 default / 1 deterministic -> second-generation step when applicable (always fixed-order);
    //          2 / 3 force the first-generation kernels (atomic / deterministic backward)
    if (variant <= 1 && fused_step_applicable(P, lo, B, N, ent, rel)) {
        e = launch_train_step_fused(P, mode, ent, rel, modulus, m_ent, v_ent, m_rel, v_rel, m_mod, v_mod, pos, neg, w, B, N,
                                    lo, ao, ws.fused_ws, ws.fused_bytes, out_small, st);
        return e == cudaSuccess ? KGE_OK : fail_cuda(e, "train_step(fused)");
    }
    const int deterministic = (variant == 1 || variant == 3) ? 1 : 0;
#define KGE_TRY(expr, what)                            \
    do {                                               \
        e = (expr);                                    \
        if (e != cudaSuccess) return fail_cuda(e, what); \
    } while (0)
    {
        ProfScope ps(ST_ZERO, st);
        KGE_TRY(cudaMemsetAsync(ws.grad_ent, 0, (size_t)P.E * P.De * sizeof(float), st), "memset grad_entity");
        KGE_TRY(cudaMemsetAsync(ws.grad_rel, 0, (size_t)P.R * P.Dr * sizeof(float), st), "memset grad_relation");
        KGE_TRY(cudaMemsetAsync(ws.small, 0, 16 * sizeof(float), st), "memset small");
    }
    {
        ProfScope ps(ST_FWD_NEG, st);
        KGE_TRY(launch_score_forward(P, mode, ent, rel, modulus, pos, neg, B, N, ws.score_neg, st), "forward(neg)");
    }
    {
        ProfScope ps(ST_FWD_POS, st);
        KGE_TRY(launch_score_forward(P, KGE_MODE_SINGLE, ent, rel, modulus, pos, nullptr, B, 1, ws.score_pos, st),
                "forward(pos)");
    }
    {
        ProfScope ps(ST_LOSS, st);
        KGE_TRY(launch_loss(ws.score_neg, ws.score_pos, w, B, N, lo, ws.small, ws.dneg, ws.dpos, ws.loss_ws, st), "loss");
    }
    float* gmod = (P.model == KGE_PROTATE) ? ws.small + 8 : nullptr;
    if (deterministic) {
        {
            ProfScope ps(ST_BWD_NEG, st);
            KGE_TRY(launch_score_backward_det(P, mode, ent, rel, modulus, pos, neg, B, N, ws.dneg, ws.grad_ent,
                                              ws.grad_rel, gmod, ws.det_ws, ws.det_bytes, st),
                    "backward_det(neg)");
        }
        ProfScope ps(ST_BWD_POS, st);
        KGE_TRY(launch_score_backward_det(P, KGE_MODE_SINGLE, ent, rel, modulus, pos, nullptr, B, 1, ws.dpos,
                                          ws.grad_ent, ws.grad_rel, gmod, ws.det_ws, ws.det_bytes, st),
                "backward_det(pos)");
    } else {
        {
            ProfScope ps(ST_BWD_NEG, st);
            KGE_TRY(launch_score_backward_atomic(P, mode, ent, rel, modulus, pos, neg, B, N, ws.dneg, ws.grad_ent,
                                                 ws.grad_rel, gmod, st),
                    "backward(neg)");
        }
        ProfScope ps(ST_BWD_POS, st);
        KGE_TRY(launch_score_backward_atomic(P, KGE_MODE_SINGLE, ent, rel, modulus, pos, nullptr, B, 1, ws.dpos,
                                             ws.grad_ent, ws.grad_rel, gmod, st),
                "backward(pos)");
    }
    float l3 = 0.f;
    if (lo.regularization != 0.f) {
        ProfScope ps(ST_REG, st);
        l3 = lo.regularization;
        KGE_TRY(launch_l3_sum_cubes(ent, (int64_t)P.E * P.De, ws.small + 4, ws.l3_ws, st), "l3(entity)");
        KGE_TRY(launch_l3_sum_cubes(rel, (int64_t)P.R * P.Dr, ws.small + 5, ws.l3_ws, st), "l3(relation)");
        finalize_reg_kernel<<<1, 1, 0, st>>>(ws.small, l3);
        KGE_TRY(cudaGetLastError(), "finalize_reg");
    }
    const AdamScalars a = adam_scalars(ao);
    {
        ProfScope ps(ST_ADAM_ENT, st);
        KGE_TRY(launch_adam(ent, ws.grad_ent, m_ent, v_ent, (int64_t)P.E * P.De, a, l3, 0, st), "adam(entity)");
    }
    ProfScope ps(ST_ADAM_REL, st);
    KGE_TRY(launch_adam(rel, ws.grad_rel, m_rel, v_rel, (int64_t)P.R * P.Dr, a, l3, 0, st), "adam(relation)");
    if (P.model == KGE_PROTATE) KGE_TRY(launch_adam(modulus, gmod, m_mod, v_mod, 1, a, 0.f, 0, st), "adam(modulus)");
#undef KGE_TRY
    return KGE_OK;
}

static int check_train_args(const kge_model_desc* desc, Params* P, int mode, const void* ent, const void* rel,
                            const void* modulus, const void* m_ent, const void* v_ent, const void* m_rel,
                            const void* v_rel, const void* m_mod, const void* v_mod, const void* pos, const void* neg, const void* w,
                            int64_t B, int64_t N, const kge_loss_opts* lo, const kge_adam_opts* ao) {
    int rc = make_params(desc, P);
    if (rc) return rc;
    if (mode != KGE_MODE_HEAD_BATCH && mode != KGE_MODE_TAIL_BATCH)
        return fail(KGE_ERR_INVALID_ARGUMENT, "negative batch mode %d not supported", mode);
    if (B <= 0 || N <= 0) return fail(KGE_ERR_INVALID_ARGUMENT, "B and N must be positive");
    if (!ent || !rel || !m_ent || !v_ent || !m_rel || !v_rel || !pos || !neg || !lo || !ao)
        return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    if (!lo->uni_weight && !w) return fail(KGE_ERR_INVALID_ARGUMENT, "subsampling_weight is NULL without uni_weight");
    if (P->model == KGE_PROTATE && (!modulus || !m_mod || !v_mod))
        return fail(KGE_ERR_INVALID_ARGUMENT, "pRotatE needs modulus and its Adam state");
    if (ao->step < 1) return fail(KGE_ERR_INVALID_ARGUMENT, "adam step must be >= 1");
    return KGE_OK;
}

}  // namespace kge

using namespace kge;

extern "C" {

int kge_profile_enable(int on) {
    g_prof_on = on != 0;
    return KGE_OK;
}
int kge_profile_num_stages(void) { return ST_COUNT; }
const char* kge_profile_stage_name(int stage) { return (stage >= 0 && stage < ST_COUNT) ? kStageNames[stage] : ""; }
int kge_profile_read(float* ms_sum, int64_t* launches, int reset) {
    for (const ProfRec& r : g_prof_live) {
        cudaError_t e = cudaEventSynchronize(r.b);
        if (e != cudaSuccess) return fail_cuda(e, "kge_profile_read");
        float ms = 0.f;
        cudaEventElapsedTime(&ms, r.a, r.b);
        g_prof_ms[r.stage] += ms;
        g_prof_launches[r.stage] += r.launches;
        g_prof_pool.push_back(r.a);
        g_prof_pool.push_back(r.b);
    }
    g_prof_live.clear();
    for (int i = 0; i < ST_COUNT; ++i) {
        if (ms_sum) ms_sum[i] = g_prof_ms[i];
        if (launches) launches[i] = g_prof_launches[i];
        if (reset) {
            g_prof_ms[i] = 0.f;
            g_prof_launches[i] = 0;
        }
    }
    return KGE_OK;
}

int kge_train_step(const kge_model_desc* desc, int mode, float* entity, float* relation, float* modulus,
                   float* adam_m_entity, float* adam_v_entity, float* adam_m_relation, float* adam_v_relation,
                   float* adam_m_modulus, float* adam_v_modulus, const int64_t* positive, const int64_t* negative,
                   const float* subsampling_weight, int64_t B, int64_t N, const kge_loss_opts* loss_opts,
                   const kge_adam_opts* adam_opts, int deterministic, float* loss3, float* reg1, void* ws,
                   size_t ws_bytes, void* stream) {
    Params P;
    int rc = check_train_args(desc, &P, mode, entity, relation, modulus, adam_m_entity, adam_v_entity,
                              adam_m_relation, adam_v_relation, adam_m_modulus, adam_v_modulus, positive, negative,
                              subsampling_weight, B, N, loss_opts, adam_opts);
    if (rc) return rc;
    if (!loss3) return fail(KGE_ERR_INVALID_ARGUMENT, "loss3 is NULL");
    const size_t need = train_step_workspace_bytes(P, B, N, false);
    if (!ws || ws_bytes < need) return fail(KGE_ERR_WORKSPACE, "train step needs %zu workspace bytes, got %zu", need, ws_bytes);
    cudaStream_t st = (cudaStream_t)stream;
    const TrainWs w = carve(P, B, N, false, ws);
    float* small = nullptr;
    rc = train_step_device(P, mode, entity, relation, modulus, adam_m_entity, adam_v_entity, adam_m_relation,
                           adam_v_relation, adam_m_modulus, adam_v_modulus, positive, negative, subsampling_weight, B, N, *loss_opts,
                           *adam_opts, deterministic, w, st, &small);
    if (rc) return rc;
    cudaError_t e;
    if (reg1 == loss3 + 3) {  // one contiguous float[4]: a single copy
        e = cudaMemcpyAsync(loss3, small, 4 * sizeof(float), cudaMemcpyDeviceToDevice, st);
    } else {
        e = cudaMemcpyAsync(loss3, small, 3 * sizeof(float), cudaMemcpyDeviceToDevice, st);
        if (e == cudaSuccess && reg1) e = cudaMemcpyAsync(reg1, small + 3, sizeof(float), cudaMemcpyDeviceToDevice, st);
    }
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_train_step(copy loss)");
}

int kge_train_sweep(const kge_model_desc* desc, int mode, const float* entity, const float* relation,
                    const float* modulus, const int64_t* positive, const int64_t* negative,
                    const float* subsampling_weight, int64_t B, int64_t N, const kge_loss_opts* loss_opts,
                    const float* wsum, int64_t B_total, const kge_step_buffers* out, void* ws, size_t ws_bytes,
                    void* stream) {
    Params P;
    int rc = make_params(desc, &P);
    if (rc) return rc;
    if (mode != KGE_MODE_HEAD_BATCH && mode != KGE_MODE_TAIL_BATCH)
        return fail(KGE_ERR_INVALID_ARGUMENT, "negative batch mode %d not supported", mode);
    if (B <= 0 || N <= 0 || B_total < B) return fail(KGE_ERR_INVALID_ARGUMENT, "bad batch sizes");
    if (!entity || !relation || !positive || !negative || !loss_opts || !out)
        return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    if (!loss_opts->uni_weight && !subsampling_weight)
        return fail(KGE_ERR_INVALID_ARGUMENT, "subsampling_weight is NULL without uni_weight");
    if (P.model == KGE_PROTATE && !modulus) return fail(KGE_ERR_INVALID_ARGUMENT, "pRotatE needs modulus");
    if (!fused_step_applicable(P, *loss_opts, B_total, N, entity, relation))
        return fail(KGE_ERR_UNSUPPORTED, "kge_train_sweep: the fused step does not apply to this configuration");
    cudaError_t e = launch_train_sweep(P, mode, entity, relation, modulus, positive, negative, subsampling_weight, B, N,
                                       *loss_opts, wsum, B_total, *out, ws, ws_bytes,
                                       (cudaStream_t)stream);
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_train_sweep");
}

int kge_train_apply(const kge_model_desc* desc, int mode, float* entity, float* relation, float* modulus,
                    float* adam_m_entity, float* adam_v_entity, float* adam_m_relation, float* adam_v_relation,
                    float* adam_m_modulus, float* adam_v_modulus, const int64_t* positive_all,
                    const int64_t* negative_all, int64_t B_all, int64_t N, const kge_loss_opts* loss_opts,
                    const kge_adam_opts* adam_opts, const kge_step_buffers* gathered, int64_t entity_lo,
                    int64_t entity_hi, float* loss3, float* reg1, void* ws, size_t ws_bytes, void* stream) {
    return kge_train_apply_peers(desc, mode, entity, relation, modulus, adam_m_entity, adam_v_entity, adam_m_relation,
                                 adam_v_relation, adam_m_modulus, adam_v_modulus, positive_all, negative_all, B_all, N,
                                 loss_opts, adam_opts, gathered, entity_lo, entity_hi, 0, nullptr, nullptr, loss3, reg1, ws,
                                 ws_bytes, stream);
}

int kge_train_apply_peers(const kge_model_desc* desc, int mode, float* entity, float* relation, float* modulus,
                          float* adam_m_entity, float* adam_v_entity, float* adam_m_relation, float* adam_v_relation,
                          float* adam_m_modulus, float* adam_v_modulus, const int64_t* positive_all,
                          const int64_t* negative_all, int64_t B_all, int64_t N, const kge_loss_opts* loss_opts,
                          const kge_adam_opts* adam_opts, const kge_step_buffers* gathered, int64_t entity_lo,
                          int64_t entity_hi, int npeers, float* const* peer_entity, float* multicast_entity, float* loss3,
                          float* reg1, void* ws, size_t ws_bytes, void* stream) {
    if (npeers < 0 || npeers > 15 || (npeers > 0 && !peer_entity))
        return fail(KGE_ERR_INVALID_ARGUMENT, "npeers must be 0..15 with a pointer table");
    Params P;
    int rc = check_train_args(desc, &P, mode, entity, relation, modulus, adam_m_entity, adam_v_entity, adam_m_relation,
                              adam_v_relation, adam_m_modulus, adam_v_modulus, positive_all, negative_all,
                              (const void*)1, B_all, N, loss_opts, adam_opts);
    if (rc) return rc;
    if (!gathered || !loss3) return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    if (entity_lo < 0 || entity_hi < entity_lo || entity_hi > P.E) return fail(KGE_ERR_INVALID_ARGUMENT, "bad entity range");
    if (!fused_step_applicable(P, *loss_opts, B_all, N, entity, relation))
        return fail(KGE_ERR_UNSUPPORTED, "kge_train_apply: the fused step does not apply to this configuration");
    const size_t need = fused_step_workspace_bytes(P, B_all, N);
    if (!ws || ws_bytes < need) return fail(KGE_ERR_WORKSPACE, "train apply needs %zu workspace bytes, got %zu", need, ws_bytes);
    cudaStream_t st = (cudaStream_t)stream;
    float* small = nullptr;
    cudaError_t e = launch_train_apply(P, mode, entity, relation, modulus, adam_m_entity, adam_v_entity, adam_m_relation,
                                       adam_v_relation, adam_m_modulus, adam_v_modulus, positive_all, negative_all, B_all, N,
                                       *loss_opts, *adam_opts, *gathered, entity_lo, entity_hi, npeers, peer_entity,
                                       multicast_entity, ws, ws_bytes, &small, st);
    if (e != cudaSuccess) return fail_cuda(e, "kge_train_apply");
    if (reg1 == loss3 + 3) {
        e = cudaMemcpyAsync(loss3, small, 4 * sizeof(float), cudaMemcpyDeviceToDevice, st);
    } else {
        e = cudaMemcpyAsync(loss3, small, 3 * sizeof(float), cudaMemcpyDeviceToDevice, st);
        if (e == cudaSuccess && reg1) e = cudaMemcpyAsync(reg1, small + 3, sizeof(float), cudaMemcpyDeviceToDevice, st);
    }
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_train_apply(copy loss)");
}

int kge_train_exchange(const kge_model_desc* desc, const int64_t* positive, const int64_t* negative, int64_t B, int64_t N,
                       const kge_step_buffers* local, int rank, int world, const kge_step_buffers* peer_gathered,
                       int64_t* const* peer_positive, int64_t* const* peer_negative, void* stream) {
    Params P;
    int rc = make_params(desc, &P);
    if (rc) return rc;
    if (!positive || !negative || !local || !peer_gathered || !peer_positive || !peer_negative)
        return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    if (B < 0 || N < 1) return fail(KGE_ERR_INVALID_ARGUMENT, "bad batch shape");
    if (world < 1 || world > 16 || rank < 0 || rank >= world)
        return fail(KGE_ERR_INVALID_ARGUMENT, "rank %d / world %d (at most 16 peers)", rank, world);
    cudaError_t e = launch_exchange(P, positive, negative, B, N, *local, rank, world, peer_gathered, peer_positive,
                                    peer_negative, (cudaStream_t)stream);
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_train_exchange");
}

int kge_stage_batch(const int64_t* host_positive, const int64_t* host_negative, const float* host_subsampling_weight,
                    int64_t* dev_positive, int64_t* dev_negative, float* dev_subsampling_weight, int64_t B, int64_t N,
                    void* stream) {
    if (!host_positive || !host_negative || !dev_positive || !dev_negative) return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    if ((host_subsampling_weight == nullptr) != (dev_subsampling_weight == nullptr))
        return fail(KGE_ERR_INVALID_ARGUMENT, "subsampling weight: host and device pointers must both be given or both be NULL");
    if (B < 0 || N < 1) return fail(KGE_ERR_INVALID_ARGUMENT, "bad batch shape");
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e = cudaMemcpyAsync(dev_positive, host_positive, (size_t)B * 3 * sizeof(int64_t), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(dev_negative, host_negative, (size_t)B * N * sizeof(int64_t), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess && host_subsampling_weight)
        e = cudaMemcpyAsync(dev_subsampling_weight, host_subsampling_weight, (size_t)B * sizeof(float), cudaMemcpyHostToDevice, st);
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_stage_batch");
}

int kge_stream_wait_event(void* stream, void* event) {
    if (!event) return fail(KGE_ERR_INVALID_ARGUMENT, "event is NULL");
    cudaError_t e = cudaStreamWaitEvent((cudaStream_t)stream, (cudaEvent_t)event, 0);
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_stream_wait_event");
}

int kge_fetch_floats(float* host_dst, const float* dev_src, int64_t n, void* stream) {
    if (!host_dst || !dev_src || n < 0) return fail(KGE_ERR_INVALID_ARGUMENT, "bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e = cudaMemcpyAsync(host_dst, dev_src, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_fetch_floats");
}

int kge_train_set_sweep_variant(int variant) {
    train_set_sweep_variant(variant);
    return KGE_OK;
}

int kge_train_apply_shard(const kge_model_desc* desc, int mode, float* entity_shard, int64_t shard_row0,
                          int64_t shard_rows, float* relation, float* modulus, float* adam_m_shard, float* adam_v_shard,
                          float* adam_m_relation, float* adam_v_relation, float* adam_m_modulus, float* adam_v_modulus,
                          const int64_t* positive_all, const int64_t* negative_all, int64_t B_all, int64_t N,
                          const kge_loss_opts* loss_opts, const kge_adam_opts* adam_opts,
                          const kge_step_buffers* gathered, int64_t entity_lo, int64_t entity_hi, float* loss3,
                          float* reg1, void* ws, size_t ws_bytes, void* stream) {
    if (!desc || !entity_shard || !adam_m_shard || !adam_v_shard) return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    if (shard_row0 < 0 || shard_rows < 0 || entity_lo < shard_row0 || entity_hi > shard_row0 + shard_rows)
        return fail(KGE_ERR_INVALID_ARGUMENT, "row range [%lld, %lld) outside the shard [%lld, %lld)", (long long)entity_lo,
                    (long long)entity_hi, (long long)shard_row0, (long long)(shard_row0 + shard_rows));
    // rows are addressed by global id and only [entity_lo, entity_hi) is dereferenced: rebase the three shards
    const uintptr_t off = (uintptr_t)shard_row0 * (uintptr_t)desc->entity_dim * sizeof(float);
    auto rebase = [off](float* p) { return reinterpret_cast<float*>(reinterpret_cast<uintptr_t>(p) - off); };
    return kge_train_apply(desc, mode, rebase(entity_shard), relation, modulus, rebase(adam_m_shard), rebase(adam_v_shard),
                           adam_m_relation, adam_v_relation, adam_m_modulus, adam_v_modulus, positive_all, negative_all,
                           B_all, N, loss_opts, adam_opts, gathered, entity_lo, entity_hi, loss3, reg1, ws, ws_bytes, stream);
}

int kge_train_step_host(const kge_model_desc* desc, int mode, float* entity, float* relation, float* modulus,
                        float* adam_m_entity, float* adam_v_entity, float* adam_m_relation, float* adam_v_relation,
                        float* adam_m_modulus, float* adam_v_modulus, const int64_t* host_positive, const int64_t* host_negative,
                        const float* host_subsampling_weight, int64_t B, int64_t N, const kge_loss_opts* loss_opts,
                        const kge_adam_opts* adam_opts, int deterministic, float* host_loss4, void* ws,
                        size_t ws_bytes, void* stream) {
    Params P;
    int rc = check_train_args(desc, &P, mode, entity, relation, modulus, adam_m_entity, adam_v_entity,
                              adam_m_relation, adam_v_relation, adam_m_modulus, adam_v_modulus, host_positive, host_negative,
                              host_subsampling_weight, B, N, loss_opts, adam_opts);
    if (rc) return rc;
    if (!host_loss4) return fail(KGE_ERR_INVALID_ARGUMENT, "host_loss4 is NULL");
    const size_t need = train_step_workspace_bytes(P, B, N, true);
    if (!ws || ws_bytes < need) return fail(KGE_ERR_WORKSPACE, "train step needs %zu workspace bytes, got %zu", need, ws_bytes);
    cudaStream_t st = (cudaStream_t)stream;
    const TrainWs w = carve(P, B, N, true, ws);
    float* small = nullptr;
    cudaError_t e = cudaMemcpyAsync(w.pos, host_positive, (size_t)B * 3 * sizeof(int64_t), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(w.neg, host_negative, (size_t)B * N * sizeof(int64_t), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess && host_subsampling_weight)
        e = cudaMemcpyAsync(w.w, host_subsampling_weight, (size_t)B * sizeof(float), cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return fail_cuda(e, "kge_train_step_host(H2D)");
    rc = train_step_device(P, mode, entity, relation, modulus, adam_m_entity, adam_v_entity, adam_m_relation,
                           adam_v_relation, adam_m_modulus, adam_v_modulus, w.pos, w.neg, host_subsampling_weight ? w.w : nullptr, B,
                           N, *loss_opts, *adam_opts, deterministic, w, st, &small);
    if (rc) return rc;
    e = cudaMemcpyAsync(host_loss4, small, 4 * sizeof(float), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_train_step_host(D2H)");
}

}  // extern "C"
