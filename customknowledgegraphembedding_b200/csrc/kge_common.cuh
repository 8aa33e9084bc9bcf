// Shared device helpers for the sm_100a KGE kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "kge_b200.h"

namespace kge {

constexpr int kWarp = 32;

// Per-launch constants derived from kge_model_desc on the host (kge_api.cu: make_params).
struct Params {
    int model;
    int H;   // hidden_dim
    int De;  // entity_dim
    int Dr;  // relation_dim
    float gamma;
    float phase_div;      // fp32( double(embedding_range) / pi )  -- U6/U7 divide by this
    float inv_phase_div;  // 1 / phase_div (candidate side of pRotatE)
    float interht_eps;
    long long E;
    long long R;
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Block-wide sum with a fixed combination order (deterministic).  `red` holds >= 32 floats.
// All threads get the result.  Must be called by every thread of the block.
__device__ __forceinline__ float block_sum(float v, float* red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();  // protect `red` from the previous use
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float t = 0.f;
    for (int w = 0; w < nwarps; ++w) t += red[w];
    return t;
}

__device__ __forceinline__ float block_max(float v, float* red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float t = red[0];
    for (int w = 1; w < nwarps; ++w) t = fmaxf(t, red[w]);
    return t;
}

// MUFU.SQRT / MUFU.RSQ: <= 1-2 ulp, far inside the 1e-5 contract.
__device__ __forceinline__ float fast_sqrt(float x) {
    float y;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float fast_rsqrt(float x) {
    float y;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ float sgnf(float x) { return (x > 0.f) ? 1.f : ((x < 0.f) ? -1.f : 0.f); }

// logsigmoid(x) = min(x,0) - log1p(exp(-|x|))   (bitwise F.logsigmoid on CPU)
__device__ __forceinline__ float log_sigmoid(float x) { return fminf(x, 0.f) - log1pf(expf(-fabsf(x))); }
__device__ __forceinline__ float sigmoidf(float x) { return 1.f / (1.f + expf(-x)); }

// ---- VEC-wide (1 or 4 floats) loads / stores / reductions ---------------------------------
template <int VEC>
__device__ __forceinline__ void ld_global(float (&x)[VEC], const float* __restrict__ p) {
    if constexpr (VEC == 4) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(p));
        x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w;
    } else {
        x[0] = __ldg(p);
    }
}
template <int VEC>
__device__ __forceinline__ void ld_shared(float (&x)[VEC], const float* p) {
    if constexpr (VEC == 4) {
        const float4 v = *reinterpret_cast<const float4*>(p);
        x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w;
    } else {
        x[0] = *p;
    }
}
template <int VEC>
__device__ __forceinline__ void st_ptr(float* p, const float (&x)[VEC]) {
    if constexpr (VEC == 4) {
        *reinterpret_cast<float4*>(p) = make_float4(x[0], x[1], x[2], x[3]);
    } else {
        *p = x[0];
    }
}
// p[0..VEC) += x   (plain read-modify-write; used on lane-owned shared-memory accumulators)
template <int VEC>
__device__ __forceinline__ void add_ptr(float* p, const float (&x)[VEC]) {
    float y[VEC];
    ld_shared<VEC>(y, p);
#pragma unroll
    for (int i = 0; i < VEC; ++i) y[i] += x[i];
    st_ptr<VEC>(p, y);
}
// global reduction without return value: RED.E.ADD.F32 (x4 vectorised on sm_90+)
template <int VEC>
__device__ __forceinline__ void red_add(float* p, const float (&x)[VEC]) {
    if constexpr (VEC == 4) {
        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(x[0]), "f"(x[1]), "f"(x[2]),
                     "f"(x[3])
                     : "memory");
    } else {
        asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(x[0]) : "memory");
    }
}

// Adam on one element (torch.optim.Adam / keras via the precomputed scalars, see kge_adam.cu)
__device__ __forceinline__ void adam_update(float& p, float g, float& m, float& v, float beta1, float beta2,
                                            float step_size, float inv_sqrt_bc2, float eps, float l3) {
    if (l3 != 0.f) g = fmaf(3.0f * l3 * p, fabsf(p), g);  // d/dp l3 |p|^3
    m = fmaf(beta1, m, (1.0f - beta1) * g);
    v = fmaf(beta2, v, (1.0f - beta2) * g * g);
    const float denom = fmaf(sqrtf(v), inv_sqrt_bc2, eps);
    p -= step_size * (m / denom);
}

}  // namespace kge
