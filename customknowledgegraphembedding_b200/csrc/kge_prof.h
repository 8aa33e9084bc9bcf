// Measurement hooks shared by the train-step translation units (see include/kge_b200.h kge_profile_*).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

namespace kge {

enum Stage {
    ST_ZERO = 0, ST_FWD_NEG, ST_FWD_POS, ST_LOSS, ST_BWD_NEG, ST_BWD_POS, ST_REG, ST_ADAM_ENT, ST_ADAM_REL,
    ST_SWEEP, ST_BUCKETS, ST_DSCORE, ST_ENTITY, ST_RELATION, ST_COUNT
};
static const char* const kStageNames[ST_COUNT] = {
    "zero_grad", "score_forward(neg)", "score_forward(pos)", "loss", "score_backward(neg)", "score_backward(pos)",
    "l3_reg", "adam(entity)", "adam(relation+modulus)",
    "fused_sweep(scores+loss+dL/dq)", "buckets(scan+fill+sort)", "loss_final", "entity_gather+adam",
    "relation_gather+adam"};
// kernels launched per stage (memsets are not kernels); a scope whose count depends on the path taken overrides it
// through ProfScope::launches
static const int kStageLaunches[ST_COUNT] = {0, 1, 1, 2, 1, 1, 5, 1, 1, 1, 5, 1, 1, 1};

struct ProfRec {
    int stage;
    int launches;
    cudaEvent_t a, b;
};
extern bool g_prof_on;
extern std::vector<ProfRec> g_prof_live;
extern std::vector<cudaEvent_t> g_prof_pool;
extern float g_prof_ms[ST_COUNT];
extern int64_t g_prof_launches[ST_COUNT];

inline cudaEvent_t prof_event() {
    if (!g_prof_pool.empty()) {
        cudaEvent_t e = g_prof_pool.back();
        g_prof_pool.pop_back();
        return e;
    }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}

struct ProfScope {
    cudaStream_t st;
    ProfRec rec;
    bool on;
    ProfScope(int stage, cudaStream_t s) : st(s), on(g_prof_on) {
        if (!on) return;
        rec.stage = stage;
        rec.launches = kStageLaunches[stage];
        rec.a = prof_event();
        rec.b = prof_event();
        cudaEventRecord(rec.a, st);
    }
    void launches(int n) { rec.launches = n; }
    ~ProfScope() {
        if (!on) return;
        cudaEventRecord(rec.b, st);
        g_prof_live.push_back(rec);
    }
};

}  // namespace kge
