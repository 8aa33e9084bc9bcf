// K4: fused dense Adam over one table (sm_100a), HBM-streaming: 128-bit loads/stores, grid sized
// as a multiple of the SM count, gradient zeroing and the L3 gradient folded in.
//
// Replaces optimizer.step() of torch.optim.Adam over dense gradients (every row of
// both tables moves every step) and tf.keras Adam apply_gradients
// (/root/reference/tensorflow_codes/supervisor.py:26, run.py:111) via inv_sqrt_bc2 = 1.
//   m = b1 m + (1-b1) g ; v = b2 v + (1-b2) g^2 ; p -= step_size * m / (sqrt(v) * inv_sqrt_bc2 + eps)
#include <math.h>

#include "kge_internal.h"

namespace kge {

AdamScalars adam_scalars(const kge_adam_opts& o) {
    AdamScalars a;
    a.beta1 = o.beta1;
    a.beta2 = o.beta2;
    a.eps = o.eps;
    const double bc1 = 1.0 - pow((double)o.beta1, (double)o.step);
    const double bc2 = 1.0 - pow((double)o.beta2, (double)o.step);
    if (o.keras_epsilon_mode) {
        a.step_size = (float)((double)o.lr * sqrt(bc2) / bc1);
        a.inv_sqrt_bc2 = 1.0f;
    } else {
        a.step_size = (float)((double)o.lr / bc1);
        a.inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
    }
    return a;
}

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamScalars& a, float l3) {
    adam_update(p, g, m, v, a.beta1, a.beta2, a.step_size, a.inv_sqrt_bc2, a.eps, l3);
}

template <bool HAS_G>
__global__ void __launch_bounds__(256)
adam_dense_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                  long long n4, long long n, AdamScalars a, float l3, int zero_grad) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    float4* p4 = reinterpret_cast<float4*>(p);
    float4* g4 = reinterpret_cast<float4*>(g);
    float4* m4 = reinterpret_cast<float4*>(m);
    float4* v4 = reinterpret_cast<float4*>(v);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 pp = p4[i], mm = m4[i], vv = v4[i];
        float4 gg = make_float4(0.f, 0.f, 0.f, 0.f);
        if constexpr (HAS_G) gg = g4[i];
        adam_one(pp.x, gg.x, mm.x, vv.x, a, l3);
        adam_one(pp.y, gg.y, mm.y, vv.y, a, l3);
        adam_one(pp.z, gg.z, mm.z, vv.z, a, l3);
        adam_one(pp.w, gg.w, mm.w, vv.w, a, l3);
        p4[i] = pp;
        m4[i] = mm;
        v4[i] = vv;
        if constexpr (HAS_G)
            if (zero_grad) g4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    // scalar tail (n not a multiple of 4)
    for (long long i = 4 * n4 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float pp = p[i], mm = m[i], vv = v[i];
        float gg = 0.f;
        if constexpr (HAS_G) gg = g[i];
        adam_one(pp, gg, mm, vv, a, l3);
        p[i] = pp;
        m[i] = mm;
        v[i] = vv;
        if constexpr (HAS_G)
            if (zero_grad) g[i] = 0.f;
    }
}

cudaError_t launch_adam(float* p, float* g, float* m, float* v, int64_t n, const AdamScalars& a, float l3_lambda,
                        int zero_grad, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    auto aligned = [](const void* x) { return (reinterpret_cast<uintptr_t>(x) & 15) == 0; };
    const bool v4ok = aligned(p) && aligned(m) && aligned(v) && (g == nullptr || aligned(g));
    const long long n4 = v4ok ? (long long)(n / 4) : 0;
    const long long work = std::max<long long>(n4, 1);
    const long long max_blocks = 8LL * sm_count();
    const int blocks = (int)std::max<long long>(1, std::min<long long>(max_blocks, (work + 255) / 256));
    if (g != nullptr) adam_dense_kernel<true><<<blocks, 256, 0, st>>>(p, g, m, v, n4, (long long)n, a, l3_lambda, zero_grad);
    else adam_dense_kernel<false><<<blocks, 256, 0, st>>>(p, g, m, v, n4, (long long)n, a, l3_lambda, zero_grad);
    return cudaGetLastError();
}

}  // namespace kge
