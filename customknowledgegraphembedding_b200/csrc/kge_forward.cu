// K1: fused gather-and-score forward (sm_100a).
//
// grid = (B, S): CTA (b, y) stages the candidate-independent planes of positive b in shared
// memory once, then its warps stream candidate rows n = y*chunk .. with 128-bit loads and
// reduce each row with warp shuffles.  Replaces KGEModel.forward and the
// gather + score part of TFKGEModel.call (/root/reference/tensorflow_codes/model.py:127-199):
// the [B,N,De] gather is never materialised.
#include "kge_internal.h"
#include "kge_score.cuh"

namespace kge {

constexpr int kFwdThreads = 256;

template <int MODEL, int VEC>
__global__ void __launch_bounds__(kFwdThreads)
score_forward_kernel(Params P, int mode, const float* __restrict__ ent, const float* __restrict__ rel,
                     const float* __restrict__ modulus_p, const int64_t* __restrict__ pos,
                     const int64_t* __restrict__ neg, int N, int chunk, float* __restrict__ scores) {
    extern __shared__ __align__(16) float smem[];
    constexpr int planes = model_planes(MODEL);
    float* q = smem;
    float* aux = q + planes * P.H;
    float* red = aux + 4;

    const int b = blockIdx.x;
    const long long h = pos[3 * (long long)b + 0], r = pos[3 * (long long)b + 1], t = pos[3 * (long long)b + 2];
    const bool head = (mode == KGE_MODE_HEAD_BATCH);
    stage_positive<MODEL>(q, aux, red, P, ent, rel, h, r, t, head);

    const float modulus = (MODEL == KGE_PROTATE) ? __ldg(modulus_p) : 0.f;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int n0 = blockIdx.y * chunk;
    const int n1 = min(N, n0 + chunk);
    for (int n = n0 + warp; n < n1; n += nwarps) {
        const long long c = (mode == KGE_MODE_SINGLE) ? t : neg[(long long)b * N + n];
        const float* row = ent + c * (long long)P.De;
        const float s = head ? score_row<MODEL, VEC, true>(q, row, P, lane, modulus)
                             : score_row<MODEL, VEC, false>(q, row, P, lane, modulus);
        if (lane == 0) scores[(long long)b * N + n] = s;
    }
}

template <int MODEL>
static cudaError_t launch_fwd_model(const Params& P, int mode, const float* ent, const float* rel,
                                    const float* modulus, const int64_t* pos, const int64_t* neg, int64_t B,
                                    int64_t N, float* scores, cudaStream_t st) {
    const size_t smem = (size_t)(model_planes(MODEL) * P.H + 4 + 32) * sizeof(float);
    // split the candidates of one positive over several CTAs when B alone cannot fill the chip
    const int nwarps = kFwdThreads / 32;
    const long long target = 4LL * sm_count();
    int S = 1;
    if (B < target) {
        const long long want = (target + B - 1) / B;
        const long long max_split = (N + nwarps - 1) / nwarps;  // at least one row per warp
        S = (int)std::max<long long>(1, std::min<long long>(want, max_split));
    }
    const int chunk = (int)((N + S - 1) / S);
    S = (int)((N + chunk - 1) / chunk);
    dim3 grid((unsigned)B, (unsigned)S);
    const bool v4 = vec4_ok(P, ent, rel);
    if (v4) {
        if (smem > 48 * 1024)
            cudaFuncSetAttribute(score_forward_kernel<MODEL, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        score_forward_kernel<MODEL, 4><<<grid, kFwdThreads, smem, st>>>(P, mode, ent, rel, modulus, pos, neg, (int)N,
                                                                        chunk, scores);
    } else {
        if (smem > 48 * 1024)
            cudaFuncSetAttribute(score_forward_kernel<MODEL, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        score_forward_kernel<MODEL, 1><<<grid, kFwdThreads, smem, st>>>(P, mode, ent, rel, modulus, pos, neg, (int)N,
                                                                        chunk, scores);
    }
    return cudaGetLastError();
}

cudaError_t launch_score_forward(const Params& P, int mode, const float* ent, const float* rel, const float* modulus,
                                 const int64_t* pos, const int64_t* neg, int64_t B, int64_t N, float* scores,
                                 cudaStream_t st) {
    if (B == 0 || N == 0) return cudaSuccess;
    switch (P.model) {
        case KGE_TRANSE: return launch_fwd_model<KGE_TRANSE>(P, mode, ent, rel, modulus, pos, neg, B, N, scores, st);
        case KGE_DISTMULT: return launch_fwd_model<KGE_DISTMULT>(P, mode, ent, rel, modulus, pos, neg, B, N, scores, st);
        case KGE_COMPLEX: return launch_fwd_model<KGE_COMPLEX>(P, mode, ent, rel, modulus, pos, neg, B, N, scores, st);
        case KGE_ROTATE: return launch_fwd_model<KGE_ROTATE>(P, mode, ent, rel, modulus, pos, neg, B, N, scores, st);
        case KGE_PROTATE: return launch_fwd_model<KGE_PROTATE>(P, mode, ent, rel, modulus, pos, neg, B, N, scores, st);
        case KGE_INTERHT: return launch_fwd_model<KGE_INTERHT>(P, mode, ent, rel, modulus, pos, neg, B, N, scores, st);
    }
    return cudaErrorInvalidValue;
}

}  // namespace kge
