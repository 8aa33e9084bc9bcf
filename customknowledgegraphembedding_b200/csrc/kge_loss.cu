// K2: fused loss -- logsigmoid, self-adversarial softmax weighting, subsampling / uniform weights,
// and dL/dscore in one pass over the [B,N] score matrix (sm_100a).
//
// Replaces the loss body of KGEModel.train_step, restated by the reference at
// /root/reference/tensorflow_codes/model.py:12-31, and the fused terms of the TF graph model
// (model.py:145, 168-171, 195-198; supervisor.py:19-23) when detach_weights == 0.
//
//   ws layout (floats): [0..B) coef_b * pos_b   [B..2B) coef_b * neg_b
// Every reduction has a fixed order, so the loss is bit-reproducible run to run.
#include "kge_internal.h"

namespace kge {

constexpr int kLossThreads = 128;

__global__ void __launch_bounds__(kLossThreads)
loss_rows_kernel(const float* __restrict__ neg_score, const float* __restrict__ pos_score,
                 const float* __restrict__ w, int B, int N, int adv, int uni, int detach, float alpha,
                 float* __restrict__ terms, float* __restrict__ dneg, float* __restrict__ dpos) {
    __shared__ float red[32];
    const int b = blockIdx.x;
    const float* s = neg_score + (long long)b * N;

    // coef_b = w_b / sum(w)  or 1/B; every CTA reduces the B weights in the same order
    float coef;
    if (uni) {
        coef = 1.0f / (float)B;
    } else {
        float part = 0.f;
        for (int i = threadIdx.x; i < B; i += blockDim.x) part += __ldg(w + i);
        const float wsum = block_sum(part, red);
        coef = __ldg(w + b) / wsum;
    }

    float negterm;
    if (adv) {
        float mx = -INFINITY;
        for (int n = threadIdx.x; n < N; n += blockDim.x) mx = fmaxf(mx, __ldg(s + n) * alpha);
        mx = block_max(mx, red);
        float z = 0.f;
        for (int n = threadIdx.x; n < N; n += blockDim.x) z += expf(__ldg(s + n) * alpha - mx);
        z = block_sum(z, red);
        const float inv_z = 1.0f / z;
        float acc = 0.f;
        for (int n = threadIdx.x; n < N; n += blockDim.x) {
            const float sn = __ldg(s + n);
            acc += expf(sn * alpha - mx) * inv_z * log_sigmoid(-sn);
        }
        negterm = block_sum(acc, red);
        if (dneg != nullptr) {
            // d negterm / d s_n = -p_n sigma(s_n)  [+ alpha p_n (l_n - negterm) when the weights carry gradient]
            for (int n = threadIdx.x; n < N; n += blockDim.x) {
                const float sn = __ldg(s + n);
                const float p = expf(sn * alpha - mx) * inv_z;
                float d = -p * sigmoidf(sn);
                if (!detach) d += alpha * p * (log_sigmoid(-sn) - negterm);
                dneg[(long long)b * N + n] = -0.5f * coef * d;
            }
        }
    } else {
        float acc = 0.f;
        for (int n = threadIdx.x; n < N; n += blockDim.x) acc += log_sigmoid(-__ldg(s + n));
        negterm = block_sum(acc, red) / (float)N;
        if (dneg != nullptr) {
            const float c = 0.5f * coef / (float)N;
            for (int n = threadIdx.x; n < N; n += blockDim.x) dneg[(long long)b * N + n] = c * sigmoidf(__ldg(s + n));
        }
    }
    if (threadIdx.x == 0) {
        const float sp = __ldg(pos_score + b);
        terms[b] = coef * log_sigmoid(sp);
        terms[B + b] = coef * negterm;
        if (dpos != nullptr) dpos[b] = -0.5f * coef * sigmoidf(-sp);
    }
}

__global__ void __launch_bounds__(256) loss_final_kernel(const float* __restrict__ terms, int B, float* loss3) {
    __shared__ float red[32];
    float p = 0.f, n = 0.f;
    for (int i = threadIdx.x; i < B; i += blockDim.x) {
        p += terms[i];
        n += terms[B + i];
    }
    p = block_sum(p, red);
    n = block_sum(n, red);
    if (threadIdx.x == 0) {
        loss3[0] = -p;
        loss3[1] = -n;
        loss3[2] = (-p + -n) / 2.0f;
    }
}

cudaError_t launch_loss_final(const float* terms, int64_t B, float* loss3, cudaStream_t st) {
    loss_final_kernel<<<1, 256, 0, st>>>(terms, (int)B, loss3);
    return cudaGetLastError();
}

size_t loss_workspace_bytes(int64_t B) { return (size_t)(2 * B + 8) * sizeof(float); }

cudaError_t launch_loss(const float* neg_score, const float* pos_score, const float* w, int64_t B, int64_t N,
                        const kge_loss_opts& o, float* loss3, float* dneg, float* dpos, float* ws, cudaStream_t st) {
    loss_rows_kernel<<<(unsigned)B, kLossThreads, 0, st>>>(neg_score, pos_score, w, (int)B, (int)N,
                                                          o.negative_adversarial_sampling, o.uni_weight,
                                                          o.detach_weights, o.adversarial_temperature, ws, dneg, dpos);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    loss_final_kernel<<<1, 256, 0, st>>>(ws, (int)B, loss3);
    return cudaGetLastError();
}

// ---- L3 regulariser value: sum |x|^3 (model.py:33-39) --------------------------------------
constexpr int kL3Blocks = 1024;

__global__ void __launch_bounds__(256) l3_partial_kernel(const float* __restrict__ x, long long n, float* partial) {
    __shared__ float red[32];
    float acc = 0.f;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float a = fabsf(__ldg(x + i));
        acc += a * a * a;
    }
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) partial[blockIdx.x] = acc;
}

__global__ void __launch_bounds__(256) l3_final_kernel(const float* __restrict__ partial, int n, float* out) {
    __shared__ float red[32];
    float acc = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) acc += partial[i];
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) out[0] = acc;
}

cudaError_t launch_l3_sum_cubes(const float* x, int64_t n, float* out, float* ws, cudaStream_t st) {
    l3_partial_kernel<<<kL3Blocks, 256, 0, st>>>(x, (long long)n, ws);
    l3_final_kernel<<<1, 256, 0, st>>>(ws, kL3Blocks, out);
    return cudaGetLastError();
}

}  // namespace kge
