// kge_check_ids: device-side range check of a batch's ids (ids parsed from a record shard, or a wrong
// --nentity, would otherwise turn into out-of-bounds row gathers where tf.gather / index_select raise).  Optional and
// synchronous by nature: the caller switches it on while bringing a data pipeline up (KGEModel.check_ids).
#include "kge_internal.h"

namespace kge {

// out[0] = entity ids outside [0, E), out[1] = relation ids outside [0, R), out[2] = first offending entity id,
// out[3] = first offending relation id (any one of them)
__global__ void __launch_bounds__(256)
check_ids_kernel(const int64_t* __restrict__ pos, const int64_t* __restrict__ neg, long long B, long long BN, long long E,
                 long long R, unsigned long long* out) {
    unsigned long long bad_e = 0, bad_r = 0;
    long long first_e = 0, first_r = 0;
    const long long total = BN + 3 * B;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const bool is_neg = i < BN;
        const long long v = is_neg ? neg[i] : pos[i - BN];
        const bool is_rel = !is_neg && ((i - BN) % 3 == 1);
        if (is_rel) {
            if (v < 0 || v >= R) { if (!bad_r) first_r = v; ++bad_r; }
        } else {
            if (v < 0 || v >= E) { if (!bad_e) first_e = v; ++bad_e; }
        }
    }
    if (bad_e) {
        atomicAdd(out + 0, bad_e);
        atomicExch(out + 2, (unsigned long long)first_e);
    }
    if (bad_r) {
        atomicAdd(out + 1, bad_r);
        atomicExch(out + 3, (unsigned long long)first_r);
    }
}

}  // namespace kge

using namespace kge;

extern "C" int kge_check_ids(const int64_t* positive, const int64_t* negative, int64_t B, int64_t N, int64_t nentity,
                             int64_t nrelation, void* scratch32, void* stream) {
    if (!positive || !scratch32) return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    if (B <= 0 || N < 0 || (N > 0 && !negative) || nentity <= 0 || nrelation <= 0)
        return fail(KGE_ERR_INVALID_ARGUMENT, "bad sizes");
    cudaStream_t st = (cudaStream_t)stream;
    unsigned long long* out = (unsigned long long*)scratch32;
    cudaError_t e = cudaMemsetAsync(out, 0, 32, st);
    if (e != cudaSuccess) return fail_cuda(e, "kge_check_ids(memset)");
    const long long total = B * N + 3 * B;
    const int blocks = (int)std::min<long long>((total + 255) / 256, 8LL * sm_count());
    check_ids_kernel<<<blocks, 256, 0, st>>>(positive, negative, B, B * N, nentity, nrelation, out);
    if ((e = cudaGetLastError()) != cudaSuccess) return fail_cuda(e, "kge_check_ids");
    unsigned long long host[4] = {0, 0, 0, 0};
    if ((e = cudaMemcpyAsync(host, out, 32, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return fail_cuda(e, "kge_check_ids(copy)");
    if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return fail_cuda(e, "kge_check_ids(sync)");
    if (host[0])
        return fail(KGE_ERR_INVALID_ARGUMENT, "entity id out of range [0, %lld): %llu ids of the batch, e.g. %lld",
                    (long long)nentity, host[0], (long long)host[2]);
    if (host[1])
        return fail(KGE_ERR_INVALID_ARGUMENT, "relation id out of range [0, %lld): %llu ids of the batch, e.g. %lld",
                    (long long)nrelation, host[1], (long long)host[3]);
    return KGE_OK;
}
