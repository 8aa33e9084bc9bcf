// K3, deterministic-order variant (sm_100a): no floating-point atomics anywhere.
//
//  A  per positive b (one CTA): stage the planes q_b (kept in the workspace for C), sweep the
//     candidates accumulating dL/dq only, fold the warps in a fixed order and chain to compact
//     per-positive rows GH[b], GR[b], GT[b] (+ the pRotatE modulus partial GM[b]).
//  B  bucket every use of an entity row by entity id: candidate uses (flat index i = b*N+n), head
//     uses (BN + b) and tail uses (BN + B + b): integer histogram -> exclusive scan -> fill ->
//     per-bucket sort, so each bucket lists its uses in ascending item order whatever the fill order.
//  C  per entity (one warp): walk the bucket in that order, rebuilding dL/d(row) of each candidate
//     use from (q_b, row, dscore_i) into lane-owned shared-memory accumulators and adding the GH/GT
//     rows; one plain += into the dense grad row at the end.
//  D  per relation (one CTA): sum GR[b] over b ascending with pos[b,1] == r.
//
// Same result as the atomic variant up to float reassociation, and bit-identical run to run.
#include "kge_internal.h"
#include "kge_ptx.cuh"
#include "kge_score.cuh"

namespace kge {

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

struct DetWs {
    float *Q, *GH, *GT, *GR, *GM;
    int *counts, *offsets, *cursor, *items, *block_sums;
    size_t total;
};

constexpr int kScanBlock = 1024;

static DetWs det_carve(const Params& P, int64_t B, int64_t N, void* base) {
    DetWs w{};
    size_t off = 0;
    auto take = [&](size_t bytes) {
        void* p = base ? (void*)((char*)base + off) : nullptr;
        off += align256(bytes);
        return p;
    };
    const size_t qdim = (size_t)model_planes(P.model) * P.H;
    w.Q = (float*)take((size_t)B * qdim * sizeof(float));
    w.GH = (float*)take((size_t)B * P.De * sizeof(float));
    w.GT = (float*)take((size_t)B * P.De * sizeof(float));
    w.GR = (float*)take((size_t)B * P.Dr * sizeof(float));
    w.GM = (float*)take((size_t)B * sizeof(float));
    w.counts = (int*)take((size_t)(P.E + 1) * sizeof(int));
    w.offsets = (int*)take((size_t)(P.E + 1) * sizeof(int));
    w.cursor = (int*)take((size_t)(P.E + 1) * sizeof(int));
    w.items = (int*)take((size_t)(B * N + 2 * B) * sizeof(int));
    w.block_sums = (int*)take((size_t)((P.E + kScanBlock) / kScanBlock + 1) * sizeof(int));
    w.total = off;
    return w;
}

size_t backward_det_workspace_bytes(const Params& P, int64_t B, int64_t N) {
    if (N < 1) N = 1;
    return det_carve(P, B, N, nullptr).total;
}

// ---- A ------------------------------------------------------------------------------------
template <int MODEL, int VEC>
__global__ void __launch_bounds__(256)
det_positive_kernel(Params P, int mode, const float* __restrict__ ent, const float* __restrict__ rel,
                    const float* __restrict__ modulus_p, const int64_t* __restrict__ pos,
                    const int64_t* __restrict__ neg, int N, const float* __restrict__ dscore, float* Q, float* GH,
                    float* GT, float* GR, float* GM) {
    extern __shared__ __align__(16) float smem[];
    constexpr int planes = model_planes(MODEL);
    const int qdim = planes * P.H;
    float* q = smem;
    float* aux = q + qdim;
    float* red = aux + 4;
    float* gq_all = red + 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int b = blockIdx.x;
    const long long h = pos[3 * (long long)b + 0], r = pos[3 * (long long)b + 1], t = pos[3 * (long long)b + 2];
    const bool head = (mode == KGE_MODE_HEAD_BATCH);
    for (int k = threadIdx.x; k < nwarps * qdim; k += blockDim.x) gq_all[k] = 0.f;
    stage_positive<MODEL>(q, aux, red, P, ent, rel, h, r, t, head);
    for (int k = threadIdx.x; k < qdim; k += blockDim.x) Q[(long long)b * qdim + k] = q[k];

    const float modulus = (MODEL == KGE_PROTATE) ? __ldg(modulus_p) : 0.f;
    float* gq = gq_all + warp * qdim;
    float gmod = 0.f;
    for (int n = warp; n < N; n += nwarps) {
        const float g = __ldg(dscore + (long long)b * N + n);
        if (g == 0.f) continue;
        const long long c = (mode == KGE_MODE_SINGLE) ? t : neg[(long long)b * N + n];
        const float* row = ent + c * (long long)P.De;
        if (head) backward_row<MODEL, VEC, true, EMIT_NONE, true>(q, row, g, P, lane, modulus, gq, nullptr, gmod);
        else backward_row<MODEL, VEC, false, EMIT_NONE, true>(q, row, g, P, lane, modulus, gq, nullptr, gmod);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < qdim; k += blockDim.x) {
        float s = gq_all[k];
        for (int w = 1; w < nwarps; ++w) s += gq_all[w * qdim + k];
        gq_all[k] = s;
    }
    // compact per-positive rows start from zero; chain_positive adds into them
    float* gh = GH + (long long)b * P.De;
    float* gt = GT + (long long)b * P.De;
    float* gr = GR + (long long)b * P.Dr;
    for (int k = threadIdx.x; k < P.De; k += blockDim.x) {
        gh[k] = 0.f;
        gt[k] = 0.f;
    }
    for (int k = threadIdx.x; k < P.Dr; k += blockDim.x) gr[k] = 0.f;
    __syncthreads();
    chain_positive<MODEL, false>(gq_all, q, aux, red, P, ent, rel, h, r, t, head, gh, gr, gt);
    if constexpr (MODEL == KGE_PROTATE) {
        const float tot = block_sum(gmod, red);
        if (threadIdx.x == 0) GM[b] = tot;
    }
}

// ---- B ------------------------------------------------------------------------------------
__device__ __forceinline__ long long item_entity(int item, int mode, const int64_t* __restrict__ pos,
                                                 const int64_t* __restrict__ neg, long long BN, long long B) {
    if (item < BN) return (mode == KGE_MODE_SINGLE) ? pos[3 * (long long)item + 2] : neg[item];
    const long long j = item - BN;
    return (j < B) ? pos[3 * j + 0] : pos[3 * (j - B) + 2];
}

// only uses of the entity rows [e_lo, e_hi) are counted / filed: a rank that owns a row range builds the buckets of
// that range alone (the per-bucket sort is quadratic in the bucket length, which grows with the global batch)
__global__ void det_histogram_kernel(int mode, const int64_t* __restrict__ pos, const int64_t* __restrict__ neg,
                                     long long BN, long long B, int* counts, long long e_lo, long long e_hi) {
    pdl_enter();
    const long long total = BN + 2 * B;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int e = item_entity((int)i, mode, pos, neg, BN, B);
        if (e >= e_lo && e < e_hi) atomicAdd(counts + e, 1);
    }
}

// exclusive scan of counts[0..n) -> offsets[0..n], three phases, 1024 elements per block
__global__ void __launch_bounds__(256) scan_block_kernel(const int* __restrict__ in, int* out, int* block_sums, long long n) {
    pdl_enter();
    __shared__ int warp_tot[8];
    const long long base = (long long)blockIdx.x * kScanBlock + threadIdx.x * 4;
    int v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = (base + i < n) ? in[base + i] : 0;
    const int mine = v[0] + v[1] + v[2] + v[3];
    int incl = mine;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    int woff = 0;
    for (int w = 0; w < warp; ++w) woff += warp_tot[w];
    int run = woff + incl - mine;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (base + i < n) out[base + i] = run;
        run += v[i];
    }
    if (threadIdx.x == 255) block_sums[blockIdx.x] = woff + incl;
}

__global__ void scan_sums_kernel(int* block_sums, int nblocks) {  // single thread block, sequential carry
    pdl_enter();
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nblocks; base += blockDim.x) {
        const int i = base + threadIdx.x;
        const int v = (i < nblocks) ? block_sums[i] : 0;
        int incl = v;
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        __shared__ int wt[32];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += y;
        }
        if (lane == 31) wt[warp] = incl;
        __syncthreads();
        int woff = 0;
        for (int w = 0; w < warp; ++w) woff += wt[w];
        const int excl = carry + woff + incl - v;
        if (i < nblocks) block_sums[i] = excl;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) block_sums[nblocks] = carry;
}

__global__ void scan_add_kernel(int* out, const int* __restrict__ block_sums, long long n, int* cursor) {
    pdl_enter();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        const int v = out[i] + block_sums[i / kScanBlock];
        out[i] = v;
        if (cursor != nullptr) cursor[i] = v;
    } else if (i == n) {
        out[n] = block_sums[(n + kScanBlock - 1) / kScanBlock];
    }
}

__global__ void det_fill_kernel(int mode, const int64_t* __restrict__ pos, const int64_t* __restrict__ neg,
                                long long BN, long long B, int* cursor, int* items, long long e_lo, long long e_hi) {
    pdl_enter();
    const long long total = BN + 2 * B;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int e = item_entity((int)i, mode, pos, neg, BN, B);
        if (e >= e_lo && e < e_hi) items[atomicAdd(cursor + e, 1)] = (int)i;
    }
}

// ---- C ------------------------------------------------------------------------------------
__device__ __forceinline__ int warp_min_int(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

template <int MODEL, int VEC>
__global__ void __launch_bounds__(256)
det_entity_kernel(Params P, int mode, const float* __restrict__ ent, const float* __restrict__ modulus_p, int N,
                  const float* __restrict__ dscore, const float* __restrict__ Q, const float* __restrict__ GH,
                  const float* __restrict__ GT, const int* __restrict__ offsets, const int* __restrict__ items,
                  long long BN, long long B, float* grad_ent) {
    extern __shared__ __align__(16) float smem[];
    constexpr int planes = model_planes(MODEL);
    const int qdim = planes * P.H, De = P.De;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    float* acc = smem + warp * De;
    const bool head = (mode == KGE_MODE_HEAD_BATCH);
    const float modulus = (MODEL == KGE_PROTATE) ? __ldg(modulus_p) : 0.f;
    float gmod_unused = 0.f;
    for (long long e = (long long)blockIdx.x * nwarps + warp; e < P.E; e += (long long)gridDim.x * nwarps) {
        const int beg = offsets[e], end = offsets[e + 1];
        if (beg == end) continue;
        for (int k = lane; k < De; k += 32) acc[k] = 0.f;
        __syncwarp();
        const float* row = ent + e * (long long)De;
        int last = -1;
        for (int j = beg; j < end; ++j) {
            // next use in ascending item order (the fill order of the bucket is not fixed)
            int m = 0x7fffffff;
            for (int idx = beg + lane; idx < end; idx += 32) {
                const int it = items[idx];
                if (it > last && it < m) m = it;
            }
            m = warp_min_int(m);
            last = m;
            if (m < BN) {
                const float g = __ldg(dscore + m);
                if (g != 0.f) {
                    const float* qb = Q + (long long)(m / N) * qdim;
                    if (head) backward_row<MODEL, VEC, true, EMIT_PLAIN, false>(qb, row, g, P, lane, modulus, nullptr, acc, gmod_unused);
                    else backward_row<MODEL, VEC, false, EMIT_PLAIN, false>(qb, row, g, P, lane, modulus, nullptr, acc, gmod_unused);
                }
            } else {
                const long long j2 = m - BN;
                const float* src = (j2 < B) ? (GH + j2 * De) : (GT + (j2 - B) * De);
                for (int k = lane; k < De; k += 32) acc[k] += src[k];
            }
            __syncwarp();
        }
        float* out = grad_ent + e * (long long)De;
        for (int k = lane; k < De; k += 32) out[k] += acc[k];
        __syncwarp();
    }
}

// ---- D ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
det_relation_kernel(int Dr, const int64_t* __restrict__ pos, int B, const float* __restrict__ GR, float* grad_rel) {
    const long long r = blockIdx.x;
    for (int k = threadIdx.x; k < Dr; k += blockDim.x) {
        float acc = 0.f;
        bool any = false;
        for (int b = 0; b < B; ++b) {
            if (pos[3 * (long long)b + 1] == r) {
                acc += GR[(long long)b * Dr + k];
                any = true;
            }
        }
        if (any) grad_rel[r * Dr + k] += acc;
    }
}

__global__ void __launch_bounds__(256) det_modulus_kernel(const float* __restrict__ GM, int B, float* grad_modulus) {
    __shared__ float red[32];
    float acc = 0.f;
    for (int b = threadIdx.x; b < B; b += blockDim.x) acc += GM[b];
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) grad_modulus[0] += acc;
}

cudaError_t launch_histogram(int mode, const int64_t* pos, const int64_t* neg, long long BN, long long B, int* counts,
                             long long e_lo, long long e_hi, cudaStream_t st) {
    const long long total = BN + 2 * B;
    const int hb = (int)std::min<long long>((total + 255) / 256, 8LL * sm_count());
    return launch_pdl(det_histogram_kernel, dim3(hb), dim3(256), 0, st, mode, pos, neg, BN, B, counts, e_lo, e_hi);
}

// exclusive scan of in[0..n) -> out[0..n] (out[n] = total); block_sums: (n + 1023) / 1024 + 2 ints of scratch
cudaError_t launch_exclusive_scan(const int* in, int* out, int* block_sums, long long n, cudaStream_t st) {
    const int nblocks = (int)((n + kScanBlock - 1) / kScanBlock);
    cudaError_t e;
    if ((e = launch_pdl(scan_block_kernel, dim3(std::max(nblocks, 1)), dim3(256), 0, st, in, out, block_sums, n)) != cudaSuccess) return e;
    if ((e = launch_pdl(scan_sums_kernel, dim3(1), dim3(1024), 0, st, block_sums, nblocks)) != cudaSuccess) return e;
    return launch_pdl(scan_add_kernel, dim3((unsigned)((n + 1 + 255) / 256)), dim3(256), 0, st, out, (const int*)block_sums, n,
                      (int*)nullptr);
}

// exclusive scan of counts -> offsets (+ a copy in cursor), then scatter every use into its bucket
cudaError_t launch_scan_fill(int mode, const int64_t* pos, const int64_t* neg, long long BN, long long B, long long E,
                             int* counts, int* offsets, int* cursor, int* items, int* block_sums, long long e_lo,
                             long long e_hi, cudaStream_t st) {
    const long long total = BN + 2 * B;
    const int hb = (int)std::min<long long>((total + 255) / 256, 8LL * sm_count());
    const int nblocks = (int)((E + kScanBlock - 1) / kScanBlock);
    cudaError_t e;
    if ((e = launch_pdl(scan_block_kernel, dim3(nblocks), dim3(256), 0, st, counts, offsets, block_sums, E)) != cudaSuccess) return e;
    if ((e = launch_pdl(scan_sums_kernel, dim3(1), dim3(1024), 0, st, block_sums, nblocks)) != cudaSuccess) return e;
    if ((e = launch_pdl(scan_add_kernel, dim3((unsigned)((E + 1 + 255) / 256)), dim3(256), 0, st, offsets, block_sums, E, cursor)) != cudaSuccess)
        return e;
    return launch_pdl(det_fill_kernel, dim3(hb), dim3(256), 0, st, mode, pos, neg, BN, B, cursor, items, e_lo, e_hi);
}

template <int MODEL>
static cudaError_t launch_det_model(const Params& P, int mode, const float* ent, const float* rel,
                                    const float* modulus, const int64_t* pos, const int64_t* neg, int64_t B,
                                    int64_t N, const float* dscore, float* grad_ent, float* grad_rel,
                                    float* grad_modulus, const DetWs& w, cudaStream_t st) {
    const bool v4 = vec4_ok(P, ent, rel) && vec4_ok(P, grad_ent, grad_rel);
    const long long BN = B * N;
    if (BN + 2 * B >= 0x7fffffffLL) return cudaErrorInvalidValue;
    // A
    int nwarps = 8;
    auto smem_for = [&](int nw) { return (size_t)(model_planes(MODEL) * P.H * (1 + nw) + 4 + 32) * sizeof(float); };
    while (nwarps > 1 && smem_for(nwarps) > 200 * 1024) nwarps >>= 1;
    const size_t smemA = smem_for(nwarps);
    if (smemA > 227 * 1024) return cudaErrorInvalidValue;
    if (v4) {
        cudaFuncSetAttribute(det_positive_kernel<MODEL, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemA);
        det_positive_kernel<MODEL, 4><<<(unsigned)B, nwarps * 32, smemA, st>>>(P, mode, ent, rel, modulus, pos, neg, (int)N,
                                                                              dscore, w.Q, w.GH, w.GT, w.GR, w.GM);
    } else {
        cudaFuncSetAttribute(det_positive_kernel<MODEL, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemA);
        det_positive_kernel<MODEL, 1><<<(unsigned)B, nwarps * 32, smemA, st>>>(P, mode, ent, rel, modulus, pos, neg, (int)N,
                                                                              dscore, w.Q, w.GH, w.GT, w.GR, w.GM);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    // B
    e = cudaMemsetAsync(w.counts, 0, (size_t)(P.E + 1) * sizeof(int), st);
    if (e != cudaSuccess) return e;
    const long long total = BN + 2 * B;
    const int hb = (int)std::min<long long>((total + 255) / 256, 8LL * sm_count());
    det_histogram_kernel<<<hb, 256, 0, st>>>(mode, pos, neg, BN, B, w.counts, 0, P.E);
    const int nblocks = (int)((P.E + kScanBlock - 1) / kScanBlock);
    scan_block_kernel<<<nblocks, 256, 0, st>>>(w.counts, w.offsets, w.block_sums, P.E);
    scan_sums_kernel<<<1, 1024, 0, st>>>(w.block_sums, nblocks);
    scan_add_kernel<<<(unsigned)((P.E + 1 + 255) / 256), 256, 0, st>>>(w.offsets, w.block_sums, P.E, w.cursor);
    det_fill_kernel<<<hb, 256, 0, st>>>(mode, pos, neg, BN, B, w.cursor, w.items, 0, P.E);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    // C
    int cw = 8;
    while (cw > 1 && (size_t)cw * P.De * sizeof(float) > 200 * 1024) cw >>= 1;
    const size_t smemC = (size_t)cw * P.De * sizeof(float);
    if (smemC > 227 * 1024) return cudaErrorInvalidValue;
    const int cblocks = (int)std::min<long long>((P.E + cw - 1) / cw, 16LL * sm_count());
    if (v4) {
        cudaFuncSetAttribute(det_entity_kernel<MODEL, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemC);
        det_entity_kernel<MODEL, 4><<<cblocks, cw * 32, smemC, st>>>(P, mode, ent, modulus, (int)N, dscore, w.Q, w.GH, w.GT,
                                                                    w.offsets, w.items, BN, B, grad_ent);
    } else {
        cudaFuncSetAttribute(det_entity_kernel<MODEL, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemC);
        det_entity_kernel<MODEL, 1><<<cblocks, cw * 32, smemC, st>>>(P, mode, ent, modulus, (int)N, dscore, w.Q, w.GH, w.GT,
                                                                    w.offsets, w.items, BN, B, grad_ent);
    }
    // D
    det_relation_kernel<<<(unsigned)P.R, 256, 0, st>>>(P.Dr, pos, (int)B, w.GR, grad_rel);
    if (MODEL == KGE_PROTATE && grad_modulus != nullptr) det_modulus_kernel<<<1, 256, 0, st>>>(w.GM, (int)B, grad_modulus);
    return cudaGetLastError();
}

cudaError_t launch_score_backward_det(const Params& P, int mode, const float* ent, const float* rel,
                                      const float* modulus, const int64_t* pos, const int64_t* neg, int64_t B,
                                      int64_t N, const float* dscore, float* grad_ent, float* grad_rel,
                                      float* grad_modulus, void* ws, size_t ws_bytes, cudaStream_t st) {
    if (B == 0 || N == 0) return cudaSuccess;
    if (ws_bytes < backward_det_workspace_bytes(P, B, N)) return cudaErrorInvalidValue;
    const DetWs w = det_carve(P, B, N, ws);
#define KGE_CASE(M) \
    case M: return launch_det_model<M>(P, mode, ent, rel, modulus, pos, neg, B, N, dscore, grad_ent, grad_rel, grad_modulus, w, st)
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
