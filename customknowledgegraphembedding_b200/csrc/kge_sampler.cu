// N1: on-GPU negative sampler with true-triple rejection (sm_100a).
//
// Replaces TrainDataset.__getitem__ of the PyTorch codes (the TF twin builds the same generators at
// /root/reference/compress_data/main.py:64-90): for a positive (h, r, t) draw N corrupted heads (head-batch) or
// tails (tail-batch) uniformly from the entities that do NOT form a known training triple.  Upstream draws 2N ids
// at a time, drops the known ones with np.in1d and keeps the first N survivors -- i.i.d. uniform over
// E \ true(h,r) or E \ true(r,t); here every slot redraws until it is accepted, which is the same distribution.
//
//   true set   one sorted int64 array per mode: composite key (r * E + fixed) * E + candidate, where `fixed` is the
//              tail (head-batch) or the head (tail-batch).  The CTA of a positive narrows it to the key's range
//              with two binary searches; a candidate is then looked up in that (short) range.
//   randomness Philox4x32-10, counter = (positive index in the sample stream, slot, attempt, step), key = seed:
//              the batch depends on (seed, step, global positive index) only -- not on the launch geometry, the
//              batch split over ranks, or the GPU count.
//   output     negative [B,N] int64 on the device, ready for kge_train_step: the per-step H2D copy of the batch
//              (2.1 MB at the headline config) disappears.
#include "kge_internal.h"

namespace kge {

struct Philox {
    uint32_t c[4];
};

__device__ __forceinline__ Philox philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
    constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        const uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0;
        k1 += W1;
    }
    return Philox{{c0, c1, c2, c3}};
}

// first index in [lo, hi) with a[i] >= x
__device__ __forceinline__ long long lower_bound_ll(const int64_t* __restrict__ a, long long lo, long long hi, int64_t x) {
    while (lo < hi) {
        const long long mid = (lo + hi) >> 1;
        if (__ldg(a + mid) < x) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}

__global__ void __launch_bounds__(128)
sample_negatives_kernel(const int64_t* __restrict__ positive, long long B, int N, int head, long long E,
                        const int64_t* __restrict__ true_sorted, long long ntrue, uint32_t seed_lo, uint32_t seed_hi,
                        unsigned long long first_index, uint32_t step, int64_t* __restrict__ negative) {
    __shared__ long long range[2];
    const long long b = blockIdx.x;
    const int64_t h = positive[3 * b], r = positive[3 * b + 1], t = positive[3 * b + 2];
    const int64_t key0 = (r * E + (head ? t : h)) * E;  // composite of candidate 0
    if (threadIdx.x < 2)  // [key0, key0 + E) is the range of this (relation, fixed entity)
        range[threadIdx.x] = lower_bound_ll(true_sorted, 0, ntrue, key0 + (threadIdx.x ? E : 0));
    __syncthreads();
    const long long lo = range[0], hi = range[1];
    const unsigned long long gidx = first_index + (unsigned long long)b;
    for (int slot = threadIdx.x; slot < N; slot += blockDim.x) {
        long long e = 0;
        for (uint32_t attempt = 0; attempt < 256; ++attempt) {  // two 64-bit draws per Philox block
            const Philox p = philox4x32_10((uint32_t)gidx, (uint32_t)slot, attempt, step, seed_lo,
                                           seed_hi ^ (uint32_t)(gidx >> 32));
            bool done = false;
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                if (done) continue;
                const unsigned long long x = ((unsigned long long)p.c[2 * half + 1] << 32) | p.c[2 * half];
                e = (long long)__umul64hi(x, (unsigned long long)E);  // uniform on [0, E), bias < E / 2^64
                const long long pos = lower_bound_ll(true_sorted, lo, hi, key0 + e);
                done = !(pos < hi && __ldg(true_sorted + pos) == key0 + e);
            }
            if (done) break;
        }
        negative[b * (long long)N + slot] = e;
    }
}

cudaError_t launch_sample_negatives(const int64_t* positive, int64_t B, int64_t N, int head, int64_t E,
                                    const int64_t* true_sorted, int64_t ntrue, uint64_t seed, uint64_t first_index,
                                    uint32_t step, int64_t* negative, cudaStream_t st) {
    if (B == 0 || N == 0) return cudaSuccess;
    sample_negatives_kernel<<<(unsigned)B, 128, 0, st>>>(positive, B, (int)N, head, E, true_sorted, ntrue, (uint32_t)seed,
                                                         (uint32_t)(seed >> 32), first_index, step, negative);
    return cudaGetLastError();
}

}  // namespace kge
