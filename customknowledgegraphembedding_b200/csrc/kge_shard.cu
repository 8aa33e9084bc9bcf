// Row-sharded train step (sm_100a): the pieces around the sweep that make a rank's batch runnable when no rank holds
// the whole entity table (BASELINE configs[4]: RotatE on YAGO3-10 / a 10M-entity graph, table row-sharded
// over 8 GPUs).  The reference's own mechanism is full replicas under tf.distribute
// (/root/reference/tensorflow_codes/run.py:8-17, supervisor.py:26-30); this replaces it.
//
//   kge_shard_compact     the entity ids a local batch touches (heads, tails, candidates) -> sorted unique list `uniq`
//                         and the batch re-indexed by position in that list.  No sort, no host round trip:
//                           mark   bitmap[id >> 5] |= 1 << (id & 31)                 (one atomicOr per first sighting)
//                           scan   wordbase[w] = #set bits in words < w              (exclusive scan of popcounts)
//                           emit   uniq[wordbase[w] + j] = 32 w + (j-th set bit)     (ascending = sorted)
//                           remap  compact(id) = wordbase[id >> 5] + popc(bitmap[id >> 5] & ((1 << (id & 31)) - 1))
//                         n_unique stays on the device; buffers are sized by capacity = min(E, B*N + 2B).
//   kge_shard_fetch_rows  compact[u] = row uniq[u] of the global table, PULLED from the shard of the rank that owns it
//                         through peer mappings (symmetric memory over NVLink; the own shard is an ordinary local
//                         read): one warp per row, 16-byte accesses, all loads of a row in flight together.
//
// The sweep then runs on the compact table (kge_train_sweep with the compact batch), the per-positive buffers go to
// the row owners (kge_train_exchange) and every owner updates its rows in place (kge_train_apply_shard).
#include <algorithm>

#include "kge_internal.h"
#include "kge_ptx.cuh"

namespace kge {

static inline size_t sh_align(size_t x) { return (x + 255) & ~(size_t)255; }

struct ShardWs {
    unsigned* bitmap;  // W words
    int* counts;       // W
    int* wordbase;     // W + 1
    int* block_sums;
    size_t total;
};

static ShardWs shard_carve(int64_t E, void* base) {
    ShardWs w{};
    const size_t W = (size_t)((E + 31) / 32);
    size_t off = 0;
    auto take = [&](size_t bytes) {
        void* p = base ? (void*)((char*)base + off) : nullptr;
        off += sh_align(bytes);
        return p;
    };
    w.bitmap = (unsigned*)take(W * sizeof(unsigned));
    w.counts = (int*)take(W * sizeof(int));
    w.wordbase = (int*)take((W + 1) * sizeof(int));
    w.block_sums = (int*)take(((W + 1023) / 1024 + 2) * sizeof(int));
    w.total = off;
    return w;
}

size_t shard_compact_workspace_bytes(int64_t nentity) { return shard_carve(nentity, nullptr).total + 256; }

__device__ __forceinline__ long long batch_id(long long i, const int64_t* __restrict__ pos, const int64_t* __restrict__ neg,
                                              long long BN, long long B) {
    if (i < BN) return neg[i];
    const long long j = i - BN;
    return (j < B) ? pos[3 * j] : pos[3 * (j - B) + 2];
}

__global__ void __launch_bounds__(256)
shard_mark_kernel(const int64_t* __restrict__ pos, const int64_t* __restrict__ neg, long long BN, long long B,
                  unsigned* bitmap) {
    pdl_enter();
    const long long total = BN + 2 * B;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long id = batch_id(i, pos, neg, BN, B);
        const unsigned bit = 1u << (id & 31);
        unsigned* w = bitmap + (id >> 5);
        if (!(*reinterpret_cast<volatile unsigned*>(w) & bit)) atomicOr(w, bit);  // most ids were seen before
    }
}

__global__ void __launch_bounds__(256) shard_popc_kernel(const unsigned* __restrict__ bitmap, int* counts, long long W) {
    pdl_enter();
    const long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (w < W) counts[w] = __popc(bitmap[w]);
}

__global__ void __launch_bounds__(256)
shard_emit_kernel(const unsigned* __restrict__ bitmap, const int* __restrict__ wordbase, long long W, int64_t* uniq,
                  long long capacity, int* n_unique) {
    pdl_enter();
    const long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (w == 0) *n_unique = (int)min((long long)wordbase[W], capacity);
    if (w >= W) return;
    unsigned m = bitmap[w];
    long long o = wordbase[w];
    while (m) {
        const int b = __ffs(m) - 1;
        if (o < capacity) uniq[o] = 32 * w + b;
        ++o;
        m &= m - 1;
    }
}

__device__ __forceinline__ long long compact_of(long long id, const unsigned* __restrict__ bitmap,
                                                const int* __restrict__ wordbase) {
    const long long w = id >> 5;
    return (long long)wordbase[w] + __popc(bitmap[w] & ((1u << (id & 31)) - 1u));
}

__global__ void __launch_bounds__(256)
shard_remap_kernel(const int64_t* __restrict__ pos, const int64_t* __restrict__ neg, long long BN, long long B,
                   const unsigned* __restrict__ bitmap, const int* __restrict__ wordbase, int64_t* cpos, int64_t* cneg) {
    pdl_enter();
    const long long total = BN + B;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        if (i < BN) {
            cneg[i] = compact_of(neg[i], bitmap, wordbase);
        } else {
            const long long b = i - BN;
            cpos[3 * b] = compact_of(pos[3 * b], bitmap, wordbase);
            cpos[3 * b + 1] = pos[3 * b + 1];
            cpos[3 * b + 2] = compact_of(pos[3 * b + 2], bitmap, wordbase);
        }
    }
}

cudaError_t launch_shard_compact(const int64_t* pos, const int64_t* neg, int64_t B, int64_t N, int64_t E, int64_t* cpos,
                                 int64_t* cneg, int64_t* uniq, int64_t capacity, int* n_unique, void* wsp, cudaStream_t st) {
    char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(wsp) + 255) & ~(uintptr_t)255);
    const ShardWs ws = shard_carve(E, base);
    const long long W = (E + 31) / 32, BN = B * N, total = BN + 2 * B;
    cudaError_t e;
    if ((e = cudaMemsetAsync(ws.bitmap, 0, (size_t)W * sizeof(unsigned), st)) != cudaSuccess) return e;
    const int gb = (int)std::min<long long>((total + 255) / 256, 16LL * sm_count());
    if ((e = launch_pdl(shard_mark_kernel, dim3(gb), dim3(256), 0, st, pos, neg, BN, (long long)B, ws.bitmap)) != cudaSuccess) return e;
    const unsigned wb = (unsigned)((W + 255) / 256);
    if ((e = launch_pdl(shard_popc_kernel, dim3(wb), dim3(256), 0, st, (const unsigned*)ws.bitmap, ws.counts, W)) != cudaSuccess) return e;
    if ((e = launch_exclusive_scan(ws.counts, ws.wordbase, ws.block_sums, W, st)) != cudaSuccess) return e;
    if ((e = launch_pdl(shard_emit_kernel, dim3(wb), dim3(256), 0, st, (const unsigned*)ws.bitmap, (const int*)ws.wordbase, W, uniq,
                        (long long)capacity, n_unique)) != cudaSuccess)
        return e;
    return launch_pdl(shard_remap_kernel, dim3(gb), dim3(256), 0, st, pos, neg, BN, (long long)B, (const unsigned*)ws.bitmap,
                      (const int*)ws.wordbase, cpos, cneg);
}

// ------------------------------------------------------------------------------------------------
// rows from their owners
// ------------------------------------------------------------------------------------------------
constexpr int kMaxShards = 16;
struct ShardPtrs {
    const float* p[kMaxShards];
};

// One warp per row.  VPL float4 per lane are requested before any is stored, so a row of De floats has all its
// 16-byte loads in flight together (remote rows cross NVLink: latency is hidden by the independent loads and by the
// other warps; nothing is cached on the way -- the owner rewrites the row every step).
template <int VPL>
__global__ void __launch_bounds__(256)
shard_fetch_kernel(const int64_t* __restrict__ uniq, const int* __restrict__ n_unique, ShardPtrs S, int world,
                   long long rows_per_rank, int De, float* __restrict__ compact) {
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const int U = *n_unique;
    const int nvec = De >> 2;  // De % 4 == 0 (checked by the caller)
    for (long long u = warp; u < U; u += nwarps) {
        const long long id = uniq[u];
        const int owner = (int)(id / rows_per_rank);
        const float* base = nullptr;
#pragma unroll
        for (int r = 0; r < kMaxShards; ++r)
            if (r == owner) base = S.p[r];  // constant indices: the table stays in the parameter bank
        const float4* src = reinterpret_cast<const float4*>(base + (id - owner * rows_per_rank) * (long long)De);
        float4* dst = reinterpret_cast<float4*>(compact + u * (long long)De);
        for (int v0 = 0; v0 < nvec; v0 += 32 * VPL) {
            float4 x[VPL];
#pragma unroll
            for (int j = 0; j < VPL; ++j) {
                const int v = v0 + 32 * j + lane;
                if (v < nvec)
                    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                                 : "=f"(x[j].x), "=f"(x[j].y), "=f"(x[j].z), "=f"(x[j].w)
                                 : "l"(src + v)
                                 : "memory");
            }
#pragma unroll
            for (int j = 0; j < VPL; ++j) {
                const int v = v0 + 32 * j + lane;
                if (v < nvec) dst[v] = x[j];
            }
        }
    }
}

__global__ void __launch_bounds__(256)
shard_fetch_scalar_kernel(const int64_t* __restrict__ uniq, const int* __restrict__ n_unique, ShardPtrs S, int world,
                          long long rows_per_rank, int De, float* __restrict__ compact) {
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const int U = *n_unique;
    for (long long u = warp; u < U; u += nwarps) {
        const long long id = uniq[u];
        const int owner = (int)(id / rows_per_rank);
        const float* base = nullptr;
#pragma unroll
        for (int r = 0; r < kMaxShards; ++r)
            if (r == owner) base = S.p[r];
        const volatile float* src = base + (id - owner * rows_per_rank) * (long long)De;
        float* dst = compact + u * (long long)De;
        for (int k = lane; k < De; k += 32) dst[k] = src[k];
    }
}

cudaError_t launch_shard_fetch_rows(const int64_t* uniq, const int* n_unique, int64_t capacity, int64_t De,
                                    const float* const* shards, int world, int64_t rows_per_rank, float* compact,
                                    cudaStream_t st) {
    if (world < 1 || world > kMaxShards) return cudaErrorInvalidValue;
    ShardPtrs S{};
    bool vec = (De % 4 == 0) && ((reinterpret_cast<uintptr_t>(compact) & 15) == 0);
    for (int r = 0; r < world; ++r) {
        S.p[r] = shards[r];
        vec = vec && ((reinterpret_cast<uintptr_t>(shards[r]) & 15) == 0);
    }
    // persistent grid: 8 warps per CTA, enough CTAs to keep ~64 warps per SM in flight, never more warps than rows
    const long long want = std::min<long long>((capacity + 7) / 8, 8LL * sm_count());
    const dim3 grid((unsigned)std::max<long long>(want, 1)), block(256);
    if (!vec) return launch_pdl(shard_fetch_scalar_kernel, grid, block, 0, st, uniq, n_unique, S, world, (long long)rows_per_rank, (int)De, compact);
    // a 2000-float row is 500 float4: 16 per lane in one round
    if (De <= 4 * 32 * 4) return launch_pdl(shard_fetch_kernel<4>, grid, block, 0, st, uniq, n_unique, S, world, (long long)rows_per_rank, (int)De, compact);
    if (De <= 4 * 32 * 8) return launch_pdl(shard_fetch_kernel<8>, grid, block, 0, st, uniq, n_unique, S, world, (long long)rows_per_rank, (int)De, compact);
    return launch_pdl(shard_fetch_kernel<16>, grid, block, 0, st, uniq, n_unique, S, world, (long long)rows_per_rank, (int)De, compact);
}

}  // namespace kge

using namespace kge;

extern "C" {

size_t kge_shard_compact_workspace_bytes(int64_t nentity) { return nentity > 0 ? shard_compact_workspace_bytes(nentity) : 0; }

int kge_shard_compact(const int64_t* positive, const int64_t* negative, int64_t B, int64_t N, int64_t nentity,
                      int64_t* compact_positive, int64_t* compact_negative, int64_t* unique_ids, int64_t capacity,
                      int32_t* n_unique, void* ws, size_t ws_bytes, void* stream) {
    if (!positive || !negative || !compact_positive || !compact_negative || !unique_ids || !n_unique)
        return fail(KGE_ERR_INVALID_ARGUMENT, "kge_shard_compact: NULL pointer");
    if (B <= 0 || N <= 0 || nentity <= 0 || capacity <= 0) return fail(KGE_ERR_INVALID_ARGUMENT, "kge_shard_compact: bad sizes");
    if (nentity >= 0x7fffffffLL * 32 || B * N + 2 * B >= 0x7fffffffLL)
        return fail(KGE_ERR_INVALID_ARGUMENT, "kge_shard_compact: batch or table too large for 32-bit compact ids");
    if (capacity < std::min<int64_t>(nentity, B * N + 2 * B))
        return fail(KGE_ERR_INVALID_ARGUMENT, "kge_shard_compact: capacity %lld < min(nentity, B*N + 2B) = %lld",
                    (long long)capacity, (long long)std::min<int64_t>(nentity, B * N + 2 * B));
    const size_t need = shard_compact_workspace_bytes(nentity);
    if (!ws || ws_bytes < need) return fail(KGE_ERR_WORKSPACE, "kge_shard_compact needs %zu workspace bytes, got %zu", need, ws_bytes);
    cudaError_t e = launch_shard_compact(positive, negative, B, N, nentity, compact_positive, compact_negative, unique_ids,
                                         capacity, n_unique, ws, (cudaStream_t)stream);
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_shard_compact");
}

int kge_shard_fetch_rows(const int64_t* unique_ids, const int32_t* n_unique, int64_t capacity, int64_t entity_dim,
                         const float* const* shard_ptrs, int world, int64_t rows_per_rank, float* compact_table,
                         void* stream) {
    if (!unique_ids || !n_unique || !shard_ptrs || !compact_table) return fail(KGE_ERR_INVALID_ARGUMENT, "kge_shard_fetch_rows: NULL pointer");
    if (world < 1 || world > 16) return fail(KGE_ERR_INVALID_ARGUMENT, "kge_shard_fetch_rows: world %d (1..16 shards)", world);
    if (capacity <= 0 || entity_dim <= 0 || rows_per_rank <= 0) return fail(KGE_ERR_INVALID_ARGUMENT, "kge_shard_fetch_rows: bad sizes");
    for (int r = 0; r < world; ++r)
        if (!shard_ptrs[r]) return fail(KGE_ERR_INVALID_ARGUMENT, "kge_shard_fetch_rows: shard pointer %d is NULL", r);
    cudaError_t e = launch_shard_fetch_rows(unique_ids, n_unique, capacity, entity_dim, shard_ptrs, world, rows_per_rank,
                                            compact_table, (cudaStream_t)stream);
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_shard_fetch_rows");
}

}  // extern "C"
