// Exchange of the per-positive buffers of one data-parallel train step by peer stores over NVLink (sm_100a).
//
// Between kge_train_sweep (local positives) and kge_train_apply (owned entity rows) every rank needs, for the GLOBAL
// batch: the ids, S (scores -> dL/ds), the staged planes Q, the relation gradient rows GR and the per-positive
// scalars -- and, ONLY for the positives whose head / tail entity it owns, the compact gradient rows GH / GT.
// The reference gets the same effect from the dense gradient all-reduce of tf.distribute (run.py:8-17,
// supervisor.py:26-30).  An NCCL all-gather ships GH / GT (half of the bytes) to every rank; here one kernel stores
// each row exactly where it is needed, straight into the peers' gathered buffers (symmetric-memory peer pointers):
//
//   block b (a local positive), global slot g = rank * B + b:
//     ids, S[b,:], Q[b], GR[b], GM[b], stats[b], terms[b]  -> every rank p, at slot g
//     GH[b] -> the owner of pos[b].head only;  GT[b] -> the owner of pos[b].tail only
//
// Bytes sent per rank: (G-1) * B * (N*12 + (QP*H + Dr)*4 + 44) + ~2 * B * De * 4 * (G-1)/G, against
// (G-1) * B * (N*12 + (QP*H + 2*De + Dr)*4 + 44) for the all-gather.  The caller orders the apply after a barrier
// across the ranks (stream-ordered symmetric-memory barrier); rows of GH / GT at slots a rank does not own are never
// read by its apply (the entity kernel walks the buckets of owned rows only).
#include "kge_internal.h"
#include "kge_ptx.cuh"
#include "kge_score.cuh"
#include "kge_train.h"

namespace kge {

constexpr int kMaxPeers = 16;

struct PeerTable {
    float* S[kMaxPeers];
    float* Q[kMaxPeers];
    float* GH[kMaxPeers];
    float* GT[kMaxPeers];
    float* GR[kMaxPeers];
    float* GM[kMaxPeers];
    float* stats[kMaxPeers];
    float* terms[kMaxPeers];
    int64_t* pos[kMaxPeers];
    int64_t* neg[kMaxPeers];
};

// n elements of 4 bytes; 16-byte stores when both rows allow it
__device__ __forceinline__ void copy_row4(float* __restrict__ dst, const float* __restrict__ src, int n) {
    const bool vec = ((n & 3) == 0) && (((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 15) == 0);
    if (vec) {
        const float4* s4 = reinterpret_cast<const float4*>(src);
        float4* d4 = reinterpret_cast<float4*>(dst);
        for (int i = threadIdx.x; i < n / 4; i += blockDim.x) d4[i] = s4[i];
    } else {
        for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
    }
}

// one row of n 4-byte elements to the same offset of every rank's buffer: each 16-byte chunk is loaded ONCE and stored
// to all `world` destinations (independent fire-and-forget stores, constant indices into the parameter bank)
template <typename PtrArray>
__device__ __forceinline__ void fan_row4(const PtrArray& dst, long long dst_off, const float* __restrict__ src, int n, int world) {
    bool vec = ((n & 3) == 0) && ((dst_off & 3) == 0) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0);
#pragma unroll
    for (int p = 0; p < kMaxPeers; ++p)
        if (p < world) vec = vec && ((reinterpret_cast<uintptr_t>(dst[p]) & 15) == 0);
    if (vec) {
        const float4* s4 = reinterpret_cast<const float4*>(src);
        for (int i = threadIdx.x; i < n / 4; i += blockDim.x) {
            const float4 v = s4[i];
#pragma unroll
            for (int p = 0; p < kMaxPeers; ++p)
                if (p < world) reinterpret_cast<float4*>(reinterpret_cast<float*>(dst[p]) + dst_off)[i] = v;
        }
    } else {
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            const float v = src[i];
#pragma unroll
            for (int p = 0; p < kMaxPeers; ++p)
                if (p < world) (reinterpret_cast<float*>(dst[p]) + dst_off)[i] = v;
        }
    }
}

__global__ void __launch_bounds__(256)
exchange_kernel(PeerTable T, kge_step_buffers L, const int64_t* __restrict__ pos, const int64_t* __restrict__ neg, int B,
                int N, int q_row, int De, int Dr, int rank, int world, long long rows_per_rank) {
    pdl_enter();
    const long long b = blockIdx.x, g = (long long)rank * B + b;
    const long long h = pos[3 * b], t = pos[3 * b + 2];
    const int own_h = (int)(h / rows_per_rank), own_t = (int)(t / rows_per_rank);
    const int fan = world;
    fan_row4(T.S, g * N, L.S + b * N, N, fan);
    fan_row4(T.GR, g * Dr, L.GR + b * Dr, Dr, fan);
    fan_row4(T.Q, g * q_row, L.Q + b * q_row, q_row, fan);
    fan_row4(T.neg, 2 * g * N, reinterpret_cast<const float*>(neg + b * N), 2 * N, fan);  // int64 ids as float pairs
    fan_row4(T.pos, 6 * g, reinterpret_cast<const float*>(pos + 3 * b), 6, fan);
    fan_row4(T.stats, 2 * g, L.stats + 2 * b, 2, fan);
    fan_row4(T.terms, 2 * g, L.terms + 2 * b, 2, fan);
    fan_row4(T.GM, g, L.GM + b, 1, fan);
    // the compact gradient rows of the head / tail entity go to the rank that owns the row, and only there
#pragma unroll
    for (int p = 0; p < kMaxPeers; ++p) {
        if (p == own_h) copy_row4(T.GH[p] + g * De, L.GH + b * De, De);
        if (p == own_t) copy_row4(T.GT[p] + g * De, L.GT + b * De, De);
    }
}

cudaError_t launch_exchange(const Params& P, const int64_t* pos, const int64_t* neg, int64_t B, int64_t N,
                            const kge_step_buffers& local, int rank, int world, const kge_step_buffers* peers,
                            int64_t* const* peer_pos, int64_t* const* peer_neg, cudaStream_t st) {
    if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world) return cudaErrorInvalidValue;
    if (B == 0) return cudaSuccess;
    PeerTable T{};
    for (int p = 0; p < world; ++p) {
        T.S[p] = peers[p].S; T.Q[p] = peers[p].Q; T.GH[p] = peers[p].GH; T.GT[p] = peers[p].GT;
        T.GR[p] = peers[p].GR; T.GM[p] = peers[p].GM; T.stats[p] = peers[p].stats; T.terms[p] = peers[p].terms;
        T.pos[p] = peer_pos[p]; T.neg[p] = peer_neg[p];
    }
    const long long rows_per_rank = (P.E + world - 1) / world;  // dist.shard_range
    return launch_pdl(exchange_kernel, dim3((unsigned)B), dim3(256), 0, st, T, local, pos, neg, (int)B, (int)N,
                      model_planes(P.model) * P.H, P.De, P.Dr, rank, world, rows_per_rank);
}

}  // namespace kge
