// K3: backward of the fused gather-and-score (sm_100a).
//
// Atomic variant: CTA (b, y) re-stages the planes of positive b, its warps re-read the candidate
// rows, rebuild the elementwise residual, and
//   * add dL/d(candidate row) straight into the dense gradient table with vectorised
//     red.global.add.v4.f32 (one 16-byte reduction per lane per 4 coordinates),
//   * keep dL/dq in lane-owned shared-memory accumulators (one set per warp), so the N-fold
//     accumulated gradient of the positive's own rows costs one reduction per coordinate per CTA
//     instead of N.
// Equivalent of loss.backward() through index_select / tape.gradient
// (/root/reference/tensorflow_codes/supervisor.py:25).
//
// Deterministic variant: see kge_backward_det.cu.
#include "kge_internal.h"
#include "kge_score.cuh"

namespace kge {

constexpr int kBwdThreads = 256;

template <int MODEL, int VEC>
__global__ void __launch_bounds__(kBwdThreads)
score_backward_atomic_kernel(Params P, int mode, const float* __restrict__ ent, const float* __restrict__ rel,
                             const float* __restrict__ modulus_p, const int64_t* __restrict__ pos,
                             const int64_t* __restrict__ neg, int N, int chunk, const float* __restrict__ dscore,
                             float* grad_ent, float* grad_rel, float* grad_modulus) {
    extern __shared__ __align__(16) float smem[];
    constexpr int planes = model_planes(MODEL);
    const int qdim = planes * P.H;
    float* q = smem;
    float* aux = q + qdim;
    float* red = aux + 4;
    float* gq_all = red + 32;  // [nwarps][qdim]

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int b = blockIdx.x;
    const long long h = pos[3 * (long long)b + 0], r = pos[3 * (long long)b + 1], t = pos[3 * (long long)b + 2];
    const bool head = (mode == KGE_MODE_HEAD_BATCH);

    for (int k = threadIdx.x; k < nwarps * qdim; k += blockDim.x) gq_all[k] = 0.f;
    stage_positive<MODEL>(q, aux, red, P, ent, rel, h, r, t, head);  // ends with __syncthreads()

    const float modulus = (MODEL == KGE_PROTATE) ? __ldg(modulus_p) : 0.f;
    float* gq = gq_all + warp * qdim;
    float gmod = 0.f;
    const int n0 = blockIdx.y * chunk;
    const int n1 = min(N, n0 + chunk);
    for (int n = n0 + warp; n < n1; n += nwarps) {
        const float g = __ldg(dscore + (long long)b * N + n);
        if (g == 0.f) continue;
        const long long c = (mode == KGE_MODE_SINGLE) ? t : neg[(long long)b * N + n];
        const float* row = ent + c * (long long)P.De;
        float* grow = grad_ent + c * (long long)P.De;
        if (head) backward_row<MODEL, VEC, true, EMIT_ATOMIC, true>(q, row, g, P, lane, modulus, gq, grow, gmod);
        else backward_row<MODEL, VEC, false, EMIT_ATOMIC, true>(q, row, g, P, lane, modulus, gq, grow, gmod);
    }
    __syncthreads();
    // fold the per-warp accumulators into warp 0's in a fixed order
    for (int k = threadIdx.x; k < qdim; k += blockDim.x) {
        float s = gq_all[k];
        for (int w = 1; w < nwarps; ++w) s += gq_all[w * qdim + k];
        gq_all[k] = s;
    }
    __syncthreads();
    chain_positive<MODEL, true>(gq_all, q, aux, red, P, ent, rel, h, r, t, head, grad_ent + h * (long long)P.De,
                                grad_rel + r * (long long)P.Dr, grad_ent + t * (long long)P.De);
    if constexpr (MODEL == KGE_PROTATE) {
        const float tot = block_sum(gmod, red);
        if (threadIdx.x == 0 && grad_modulus != nullptr && tot != 0.f) atomicAdd(grad_modulus, tot);
    }
}

template <int MODEL>
static cudaError_t launch_bwd_model(const Params& P, int mode, const float* ent, const float* rel,
                                    const float* modulus, const int64_t* pos, const int64_t* neg, int64_t B,
                                    int64_t N, const float* dscore, float* grad_ent, float* grad_rel,
                                    float* grad_modulus, cudaStream_t st) {
    // one lane-owned dL/dq accumulator set per warp: drop warps until it fits in shared memory
    int nwarps = kBwdThreads / 32;
    auto smem_for = [&](int w) { return (size_t)(model_planes(MODEL) * P.H * (1 + w) + 4 + 32) * sizeof(float); };
    while (nwarps > 1 && smem_for(nwarps) > 200 * 1024) nwarps >>= 1;
    const size_t smem = smem_for(nwarps);
    if (smem > 227 * 1024) return cudaErrorInvalidValue;
    const int threads = nwarps * 32;
    const long long target = 2LL * sm_count();
    int S = 1;
    if (B < target) {
        const long long want = (target + B - 1) / B;
        const long long max_split = (N + 4 * nwarps - 1) / (4 * nwarps);
        S = (int)std::max<long long>(1, std::min<long long>(want, max_split));
    }
    const int chunk = (int)((N + S - 1) / S);
    S = (int)((N + chunk - 1) / chunk);
    dim3 grid((unsigned)B, (unsigned)S);
    if (vec4_ok(P, ent, rel) && vec4_ok(P, grad_ent, grad_rel)) {
        cudaFuncSetAttribute(score_backward_atomic_kernel<MODEL, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)smem);
        score_backward_atomic_kernel<MODEL, 4><<<grid, threads, smem, st>>>(
            P, mode, ent, rel, modulus, pos, neg, (int)N, chunk, dscore, grad_ent, grad_rel, grad_modulus);
    } else {
        cudaFuncSetAttribute(score_backward_atomic_kernel<MODEL, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)smem);
        score_backward_atomic_kernel<MODEL, 1><<<grid, threads, smem, st>>>(
            P, mode, ent, rel, modulus, pos, neg, (int)N, chunk, dscore, grad_ent, grad_rel, grad_modulus);
    }
    return cudaGetLastError();
}

cudaError_t launch_score_backward_atomic(const Params& P, int mode, const float* ent, const float* rel,
                                         const float* modulus, const int64_t* pos, const int64_t* neg, int64_t B,
                                         int64_t N, const float* dscore, float* grad_ent, float* grad_rel,
                                         float* grad_modulus, cudaStream_t st) {
    if (B == 0 || N == 0) return cudaSuccess;
#define KGE_CASE(M) \
    case M: return launch_bwd_model<M>(P, mode, ent, rel, modulus, pos, neg, B, N, dscore, grad_ent, grad_rel, grad_modulus, st)
    switch (P.model) {
        KGE_CASE(KGE_TRANSE);
        KGE_CASE(KGE_DISTMULT);
        KGE_CASE(KGE_COMPLEX);
        KGE_CASE(KGE_ROTATE);
        KGE_CASE(KGE_PROTATE);
        KGE_CASE(KGE_INTERHT);
    }
#undef KGE_CASE
    return cudaErrorInvalidValue;
}

}  // namespace kge
