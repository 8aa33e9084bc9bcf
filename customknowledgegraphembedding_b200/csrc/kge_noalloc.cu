// kge_loss_forward_backward_static: K2 (fused loss, kge_loss.cu) for callers that can neither allocate nor pass a
// workspace -- code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in
// the reference.
//
// The scratch K2 needs (2B + 8 floats: the per-positive loss terms between its two kernels) lives in ONE file-scope
// __device__ array of KGE_LOSS_STATIC_MAX_B positives (512 KB + 32 B; every GPU that loads the library holds it until
// the process exits).  In HOST code the array's name is a host-side stand-in, not a device address: the address the
// kernels get comes from cudaGetSymbolAddress, asked once per device with that device current.
//
// The call allocates nothing: no cudaMalloc / cudaFree, no allocating library routine.  Every call on a device shares
// the array, which keeps the last call's contents; K2 writes each term it later reads (terms[0..B) and terms[B..2B) by
// the rows kernel, before the final kernel sums them), so nothing has to start at zero and nothing is cleared.  Calls
// that overlap on the GPU (two streams, two host threads on the same device) RACE on the array: serialise them.
#include <mutex>

#include "kge_internal.h"

namespace kge {

constexpr int64_t kLossStaticMaxB = KGE_LOSS_STATIC_MAX_B;
constexpr int kMaxDevices = 64;

static __device__ __align__(256) float g_loss_static_scratch[2 * KGE_LOSS_STATIC_MAX_B + 8];

static float* g_scratch_ptr[kMaxDevices];  // device address of the array on device i, NULL until asked
static std::mutex g_scratch_mu;

// Device address of the scratch array on the CURRENT device (asked once per device, cached).
static cudaError_t loss_static_scratch(float** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_scratch_mu);
    if (g_scratch_ptr[dev] == nullptr) {
        float* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_loss_static_scratch);
        if (e != cudaSuccess) return e;
        if (p == nullptr) return cudaErrorInvalidSymbol;
        g_scratch_ptr[dev] = p;
    }
    *out = g_scratch_ptr[dev];
    return cudaSuccess;
}

}  // namespace kge

using namespace kge;

extern "C" {

int64_t kge_loss_static_max_b(void) { return kLossStaticMaxB; }

int kge_loss_forward_backward_static(const float* negative_score, const float* positive_score,
                                     const float* subsampling_weight, int64_t B, int64_t N, const kge_loss_opts* opts,
                                     float* loss3, float* dneg, float* dpos, void* stream) {
    if (!opts || !negative_score || !positive_score || !loss3) return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    if (B <= 0 || N <= 0) return fail(KGE_ERR_INVALID_ARGUMENT, "B and N must be positive");
    if (B > kLossStaticMaxB)
        return fail(KGE_ERR_INVALID_ARGUMENT,
                    "kge_loss_forward_backward_static: B = %lld exceeds the static scratch (KGE_LOSS_STATIC_MAX_B = %lld "
                    "positives); use kge_loss_forward_backward with a caller-owned workspace",
                    (long long)B, (long long)kLossStaticMaxB);
    if (!opts->uni_weight && !subsampling_weight)
        return fail(KGE_ERR_INVALID_ARGUMENT, "subsampling_weight is NULL without uni_weight");
    float* scratch = nullptr;
    cudaError_t e = loss_static_scratch(&scratch);
    if (e != cudaSuccess) return fail_cuda(e, "kge_loss_forward_backward_static(cudaGetSymbolAddress)");
    e = launch_loss(negative_score, positive_score, subsampling_weight, B, N, *opts, loss3, dneg, dpos, scratch,
                    (cudaStream_t)stream);
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_loss_forward_backward_static");
}

}  // extern "C"
