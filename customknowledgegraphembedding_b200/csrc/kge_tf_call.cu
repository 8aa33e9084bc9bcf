// The reductions of TFKGEModel.call (/root/reference/tensorflow_codes/model.py:145,168-171,195-198) on raw scores:
//   mode 3 (single):        out[b] = log_sigmoid(s[b])
//   mode 0 / 1 (negatives): out[b] = sum_n softmax(s[b, :])_n * log_sigmoid(-s[b, n])      (temperature 1, weights attached)
// as one row kernel with its backward -- the gradient flows THROUGH the softmax weights, as in the TF graph model:
//   d out[b] / d s[b, n] = p_n * (sign * sigmoid(-sign * s_n) + l_n - out[b]),   l_n = log_sigmoid(sign * s_n),
// sign = +1 for mode 3 (N = 1, p = 1), -1 for the negatives.  One warp per row, online softmax in fp32.
#include "kge_internal.h"

namespace kge {

__device__ __forceinline__ float tfc_logsigmoid(float x) { return fminf(x, 0.f) - log1pf(__expf(-fabsf(x))); }
__device__ __forceinline__ float tfc_sigmoid(float x) { return 1.f / (1.f + __expf(-x)); }

// BACKWARD = false: out[b];  true: ds[b, n] = gout[b] * d out[b] / d s[b, n] (out[b] recomputed from the row)
template <bool BACKWARD>
__global__ void __launch_bounds__(256)
tf_call_rows_kernel(const float* __restrict__ s, long long B, int N, float sign, const float* __restrict__ gout,
                    float* __restrict__ out, float* __restrict__ ds) {
    const int lane = threadIdx.x & 31;
    const long long b = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (b >= B) return;
    const float* row = s + b * N;
    float m = -INFINITY;
    for (int n = lane; n < N; n += 32) m = fmaxf(m, row[n]);
    m = warp_max(m);
    float z = 0.f, acc = 0.f;
    for (int n = lane; n < N; n += 32) {
        const float e = __expf(row[n] - m);
        z += e;
        acc += e * tfc_logsigmoid(sign * row[n]);
    }
    z = warp_sum(z);
    acc = warp_sum(acc);
    const float o = acc / z;
    if (!BACKWARD) {
        if (lane == 0) out[b] = o;
        return;
    }
    const float g = gout[b] / z;
    for (int n = lane; n < N; n += 32) {
        const float x = row[n];
        ds[b * N + n] = g * __expf(x - m) * (sign * tfc_sigmoid(-sign * x) + tfc_logsigmoid(sign * x) - o);
    }
}

}  // namespace kge

using namespace kge;

extern "C" {

int kge_tf_call_terms(const float* score, int64_t B, int64_t N, int single, float* out, void* stream) {
    if (!score || !out) return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    if (B <= 0 || N <= 0 || N > 0x7fffffff) return fail(KGE_ERR_INVALID_ARGUMENT, "bad sizes");
    const unsigned blocks = (unsigned)((B * 32 + 255) / 256);
    tf_call_rows_kernel<false><<<blocks, 256, 0, (cudaStream_t)stream>>>(score, B, (int)N, single ? 1.f : -1.f, nullptr, out, nullptr);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_tf_call_terms");
}

int kge_tf_call_terms_backward(const float* score, const float* gout, int64_t B, int64_t N, int single, float* dscore,
                               void* stream) {
    if (!score || !gout || !dscore) return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    if (B <= 0 || N <= 0 || N > 0x7fffffff) return fail(KGE_ERR_INVALID_ARGUMENT, "bad sizes");
    const unsigned blocks = (unsigned)((B * 32 + 255) / 256);
    tf_call_rows_kernel<true><<<blocks, 256, 0, (cudaStream_t)stream>>>(score, B, (int)N, single ? 1.f : -1.f, gout, nullptr, dscore);
    cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_tf_call_terms_backward");
}

}  // extern "C"
