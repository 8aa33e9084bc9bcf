// K5: filtered full-entity ranking by counting (sm_100a).
//
// Replaces the body of KGEModel.test_step: instead of materialising the [Q,E]
// score matrix and arg-sorting it, every (query, candidate) score is produced inside a register
// tile and immediately compared with the query's true score; only two integer counters per query
// leave the SM.
//
//   1 stage      per query: the candidate-independent planes q (same maths as K1) -> workspace
//   2 norms      InterHT only: 1/|a_e|, 1/|b_e| per entity
//   3 true       per query (one warp): score of the true entity, accumulated k = 0..H-1 in order
//   4 tile       128 queries x 128 entities per CTA, 8x8 scores per thread, coordinates streamed
//                through shared memory in chunks of 16 (k-major, register-staged double buffer);
//                epilogue counts score > true / score == true, skipping the true entity itself
//   5 filter     per filtered (query, entity) pair (one warp): undo its contribution to the counters
//
// Centred accumulation (distance models): a score is gamma - sum_k term_k with ~H terms of similar size, and a plain
// fp32 running sum loses ~1e-6 of the sum to rounding over H = 1000 adds -- the spacing of neighbouring candidates'
// scores at E >= 1e4, so ranks flip against the exact ranking of the same fp32 tables.  Here the running sum is shifted
// after every chunk of 16 coordinates by a per-query constant cen_q = (true entity's sum) / #chunks: candidates that
// score near the true entity keep |running sum| ~ 0.1 instead of ~ sum, which cuts the rounding error ~30x; candidates
// far from it lose nothing that matters.  Only differences to the true score are compared, and every pair of a query
// is shifted by the same constant the same number of times, so the ranking is that of the unshifted sums.
//
// Bit-consistency: 3, 4 and 5 evaluate a pair with the SAME instruction sequence (explicit
// round-to-nearest intrinsics, no contraction freedom) in the SAME order, so the true entity
// scores exactly its true score inside the tile and a filtered entity is subtracted iff it was
// counted.  This is what makes "rank = 1 + #greater" reproduce the reference's tie handling.
#include "kge_internal.h"
#include "kge_score.cuh"

namespace kge {

__host__ __device__ constexpr int cand_planes(int model) {
    return (model == KGE_COMPLEX || model == KGE_ROTATE || model == KGE_INTERHT) ? 2 : 1;
}

struct EvalConsts {
    float gamma, inv_phase_div, modulus;
};

// raw table values -> what the inner loop consumes
template <int MODEL>
__device__ __forceinline__ void prep_candidate(float& c0, float& c1, float x0, float x1, float inv_div, float ina,
                                               float inb) {
    if constexpr (MODEL == KGE_PROTATE) {
        c0 = __fmul_rn(x0, inv_div);
        c1 = 0.f;
    } else if constexpr (MODEL == KGE_INTERHT) {
        c0 = __fmul_rn(x0, ina);
        c1 = __fadd_rn(__fmul_rn(x1, inb), 1.0f);
    } else {
        c0 = x0;
        c1 = x1;
    }
}

template <int MODEL, bool HEAD>
__device__ __forceinline__ float accum(float acc, float q0, float q1, float q2, float c0, float c1) {
    if constexpr (MODEL == KGE_TRANSE) {
        return __fadd_rn(acc, fabsf(HEAD ? __fadd_rn(c0, q0) : __fsub_rn(q0, c0)));
    } else if constexpr (MODEL == KGE_DISTMULT) {
        return __fmaf_rn(q0, c0, acc);
    } else if constexpr (MODEL == KGE_COMPLEX) {
        return __fmaf_rn(q1, c1, __fmaf_rn(q0, c0, acc));
    } else if constexpr (MODEL == KGE_ROTATE) {
        const float vr = __fsub_rn(q0, c0), vi = __fsub_rn(q1, c1);
        return __fadd_rn(acc, fast_sqrt(__fmaf_rn(vr, vr, __fmul_rn(vi, vi))));
    } else if constexpr (MODEL == KGE_PROTATE) {
        const float v = HEAD ? __fadd_rn(c0, q0) : __fsub_rn(q0, c0);
        return __fadd_rn(acc, fabsf(sinf(v)));
    } else {  // InterHT: q = (A, B+1, mid), c = (a_hat, b_hat + 1)
        const float v = HEAD ? __fmaf_rn(c0, q1, __fmaf_rn(-q0, c1, q2)) : __fmaf_rn(q0, c1, __fmaf_rn(-c0, q1, q2));
        return __fadd_rn(acc, fabsf(v));
    }
}

// Two candidates per instruction on the packed fp32 pipe (fma.rn.f32x2 / FFMA2; a scalar 3-register FFMA issues at
// half rate on sm_100, which is what bounded the tile kernel below the SFU ceiling).  Each component goes through
// exactly the roundings of accum<> above (x - y == fma(y, -1, x); __f*2_rn round per component like __f*_rn), so
// the tile stays bit-consistent with the true-score and filter passes.
template <int MODEL>
__host__ __device__ constexpr bool packed_accum() { return MODEL == KGE_ROTATE || MODEL == KGE_TRANSE; }

template <int MODEL, bool HEAD>
__device__ __forceinline__ float2 accum2(float2 acc, float q0, float q1, float2 c0, float2 c1) {
    const float2 neg1 = make_float2(-1.f, -1.f);
    const float2 qq0 = make_float2(q0, q0);
    if constexpr (MODEL == KGE_TRANSE) {
        float2 v = HEAD ? __fadd2_rn(c0, qq0) : __ffma2_rn(c0, neg1, qq0);
        v.x = fabsf(v.x);
        v.y = fabsf(v.y);
        return __fadd2_rn(acc, v);
    } else {  // RotatE
        const float2 qq1 = make_float2(q1, q1);
        const float2 vr = __ffma2_rn(c0, neg1, qq0), vi = __ffma2_rn(c1, neg1, qq1);
        const float2 n2 = __ffma2_rn(vr, vr, __fmul2_rn(vi, vi));
        return __fadd2_rn(acc, make_float2(fast_sqrt(n2.x), fast_sqrt(n2.y)));
    }
}

// models whose score is gamma - (a sum of H non-negative terms): centred accumulation, chunks of kCenChunk coordinates
template <int MODEL>
__host__ __device__ constexpr bool centered() {
    return MODEL == KGE_TRANSE || MODEL == KGE_ROTATE || MODEL == KGE_PROTATE || MODEL == KGE_INTERHT;
}
constexpr int kCenChunk = 16;  // == kKC of the tile kernel

template <int MODEL>
__device__ __forceinline__ float finalize(float acc, const EvalConsts& c) {
    if constexpr (MODEL == KGE_DISTMULT || MODEL == KGE_COMPLEX) return acc;
    else if constexpr (MODEL == KGE_PROTATE) return __fsub_rn(c.gamma, __fmul_rn(acc, c.modulus));
    else return __fsub_rn(c.gamma, acc);
}

// One WARP, one (query, entity) pair, coordinates accumulated in ascending order: the reference order of K5 (what a
// thread of the tile kernel does for each of its 64 pairs).  A single thread walking the two rows is a chain of H
// dependent-latency global loads (measured 0.25 ms for the true scores and 0.69 ms for the filter entries of 2048
// queries at H = 1000).  Here the lanes read both rows coalesced and evaluate the per-coordinate term -- the value
// accum<> adds, which does not depend on the running sum -- into a warp-private shared-memory strip; lane 0 then adds
// the terms in ascending k with the same __fadd_rn chain.  DistMult / ComplEx accumulate through fma(q, c, acc), whose
// product is not rounded on its own: for them the strip holds the operands and lane 0 runs the fma chain.
constexpr int kPairChunk = 1024;  // coordinates per strip pass (bounds shared memory whatever H is)
constexpr int kPairWarps = 4;
template <int MODEL>
__host__ __device__ constexpr int pair_strips() { return MODEL == KGE_DISTMULT ? 2 : (MODEL == KGE_COMPLEX ? 4 : 1); }
template <int MODEL>
constexpr size_t pair_smem() { return (size_t)kPairWarps * pair_strips<MODEL>() * kPairChunk * sizeof(float); }

template <int MODEL, bool HEAD>
__device__ __forceinline__ float pair_score_warp(const float* __restrict__ qrow, const float* __restrict__ erow, int H,
                                                 const EvalConsts& c, float ina, float inb, float* __restrict__ strip,
                                                 float cen = 0.f, float* raw_acc = nullptr) {
    constexpr int QP = model_planes(MODEL), CP = cand_planes(MODEL);
    const int lane = threadIdx.x & 31;
    float acc = 0.f;
    for (int k0 = 0; k0 < H; k0 += kPairChunk) {
        const int n = min(kPairChunk, H - k0);
        for (int i = lane; i < n; i += 32) {
            const int k = k0 + i;
            const float q0 = qrow[k];
            const float q1 = (QP > 1) ? qrow[H + k] : 0.f;
            const float q2 = (QP > 2) ? qrow[2 * H + k] : 0.f;
            const float x0 = __ldg(erow + k);
            const float x1 = (CP > 1) ? __ldg(erow + H + k) : 0.f;
            float c0, c1;
            prep_candidate<MODEL>(c0, c1, x0, x1, c.inv_phase_div, ina, inb);
            if constexpr (MODEL == KGE_DISTMULT) {
                strip[i] = q0;
                strip[kPairChunk + i] = c0;
            } else if constexpr (MODEL == KGE_COMPLEX) {
                strip[i] = q0;
                strip[kPairChunk + i] = c0;
                strip[2 * kPairChunk + i] = q1;
                strip[3 * kPairChunk + i] = c1;
            } else {
                // accum<>(0, ...) == +0 + term == term for every term these models produce (|x|, sqrt: never -0)
                strip[i] = accum<MODEL, HEAD>(0.f, q0, q1, q2, c0, c1);
            }
        }
        __syncwarp();
        if (lane == 0) {
            if constexpr (MODEL == KGE_DISTMULT) {
                for (int i = 0; i < n; ++i) acc = __fmaf_rn(strip[i], strip[kPairChunk + i], acc);
            } else if constexpr (MODEL == KGE_COMPLEX) {
                for (int i = 0; i < n; ++i)
                    acc = __fmaf_rn(strip[2 * kPairChunk + i], strip[3 * kPairChunk + i],
                                    __fmaf_rn(strip[i], strip[kPairChunk + i], acc));
            } else {
                // chunks of kCenChunk coordinates (kPairChunk is a multiple: strip offsets are absolute chunk boundaries),
                // the running sum shifted by cen after each -- exactly what a thread of the tile kernel does; a partial
                // last chunk (H % 16 != 0) is padded there with +0 terms and shifted too
                for (int i0 = 0; i0 < n; i0 += kCenChunk) {
                    const int m = min(kCenChunk, n - i0);
                    int i = 0;
                    for (; i + 4 <= m; i += 4) {
                        const float4 t = *reinterpret_cast<const float4*>(strip + i0 + i);
                        acc = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(acc, t.x), t.y), t.z), t.w);
                    }
                    for (; i < m; ++i) acc = __fadd_rn(acc, strip[i0 + i]);
                    acc = __fsub_rn(acc, cen);
                }
            }
        }
        __syncwarp();
    }
    if (raw_acc != nullptr) *raw_acc = __shfl_sync(0xffffffffu, acc, 0);
    return __shfl_sync(0xffffffffu, finalize<MODEL>(acc, c), 0);
}

// ---- 1 stage ---------------------------------------------------------------------------------
template <int MODEL>
__global__ void __launch_bounds__(128)
eval_stage_kernel(Params P, int head, const float* __restrict__ ent, const float* __restrict__ rel,
                  const int64_t* __restrict__ queries, const float* __restrict__ qrows, float* __restrict__ Qs) {
    extern __shared__ __align__(16) float smem[];
    constexpr int planes = model_planes(MODEL);
    const int qdim = planes * P.H;
    float* q = smem;
    float* aux = q + qdim;
    float* red = aux + 4;
    const long long b = blockIdx.x;
    // qrows [Q,2,De] (row-sharded tables): the head / tail rows of query b travel with the query
    if (qrows != nullptr) stage_positive<MODEL>(q, aux, red, P, qrows, rel, 2 * b, queries[3 * b + 1], 2 * b + 1, head != 0);
    else stage_positive<MODEL>(q, aux, red, P, ent, rel, queries[3 * b], queries[3 * b + 1], queries[3 * b + 2], head != 0);
    for (int k = threadIdx.x; k < qdim; k += blockDim.x) Qs[b * qdim + k] = q[k];
}

// ---- 2 norms (InterHT) -----------------------------------------------------------------------
__global__ void __launch_bounds__(256)
eval_inv_norm_kernel(Params P, const float* __restrict__ ent, long long e_lo, long long e_hi, float* __restrict__ ina,
                     float* __restrict__ inb) {
    // rows [e_lo, e_hi) of `ent` -> ina / inb [e - e_lo]
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
    ina -= e_lo;
    inb -= e_lo;
    for (long long e = e_lo + warp; e < e_hi; e += nw) {
        const float* row = ent + e * (long long)P.De;
        float sa = 0.f, sb = 0.f;
        for (int k = lane; k < P.H; k += 32) {
            const float a = __ldg(row + k), b = __ldg(row + P.H + k);
            sa = fmaf(a, a, sa);
            sb = fmaf(b, b, sb);
        }
        float na = sqrtf(warp_sum(sa)), nb = sqrtf(warp_sum(sb));
        if (P.interht_eps > 0.f) {
            na = fmaxf(na, P.interht_eps);
            nb = fmaxf(nb, P.interht_eps);
        }
        if (lane == 0) {
            ina[e] = 1.0f / na;
            inb[e] = 1.0f / nb;
        }
    }
}

// ---- 3 true ----------------------------------------------------------------------------------
// row of the true entity of query q: from the table, or from qrows [Q,2,De] when the table is row-sharded
template <bool HEAD>
__device__ __forceinline__ const float* true_row(const Params& P, const float* __restrict__ ent,
                                                 const float* __restrict__ qrows, const int64_t* __restrict__ queries,
                                                 long long q) {
    if (qrows != nullptr) return qrows + (2 * q + (HEAD ? 0 : 1)) * (long long)P.De;
    return ent + (HEAD ? queries[3 * q] : queries[3 * q + 2]) * (long long)P.De;
}

// InterHT: 1/|a|, 1/|b| of every query's true entity (one warp per query; same arithmetic as eval_inv_norm_kernel)
template <bool HEAD>
__global__ void __launch_bounds__(256)
eval_true_norm_kernel(Params P, const float* __restrict__ ent, const float* __restrict__ qrows,
                      const int64_t* __restrict__ queries, long long Q, float* __restrict__ tna, float* __restrict__ tnb) {
    const int lane = threadIdx.x & 31;
    const long long q = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (q >= Q) return;
    const float* row = true_row<HEAD>(P, ent, qrows, queries, q);
    float sa = 0.f, sb = 0.f;
    for (int k = lane; k < P.H; k += 32) {
        const float a = __ldg(row + k), b = __ldg(row + P.H + k);
        sa = fmaf(a, a, sa);
        sb = fmaf(b, b, sb);
    }
    float na = sqrtf(warp_sum(sa)), nb = sqrtf(warp_sum(sb));
    if (P.interht_eps > 0.f) {
        na = fmaxf(na, P.interht_eps);
        nb = fmaxf(nb, P.interht_eps);
    }
    if (lane == 0) {
        tna[q] = 1.0f / na;
        tnb[q] = 1.0f / nb;
    }
}

template <int MODEL, bool HEAD>
__global__ void __launch_bounds__(32 * kPairWarps)
eval_true_kernel(Params P, EvalConsts c, const float* __restrict__ modulus_p, const float* __restrict__ ent,
                 const float* __restrict__ qrows, const float* __restrict__ Qs, const float* __restrict__ tna,
                 const float* __restrict__ tnb, const int64_t* __restrict__ queries, long long Q,
                 float* __restrict__ true_score, float* __restrict__ true_cmp, float* __restrict__ cen) {
    extern __shared__ __align__(16) float smem[];
    float* strip = smem + (size_t)(threadIdx.x >> 5) * pair_strips<MODEL>() * kPairChunk;
    const long long q = (long long)blockIdx.x * kPairWarps + (threadIdx.x >> 5);  // one warp per query
    if (q >= Q) return;
    if (MODEL == KGE_PROTATE) c.modulus = __ldg(modulus_p);
    const int qdim = model_planes(MODEL) * P.H;
    const float a = (MODEL == KGE_INTERHT) ? tna[q] : 0.f, b = (MODEL == KGE_INTERHT) ? tnb[q] : 0.f;
    const float* trow = true_row<HEAD>(P, ent, qrows, queries, q);
    float acc0 = 0.f;
    const float s = pair_score_warp<MODEL, HEAD>(Qs + q * qdim, trow, P.H, c, a, b, strip, 0.f, &acc0);
    float s_cmp = s, cq = 0.f;
    if constexpr (centered<MODEL>()) {
        // the query's shift: the true entity's own mean chunk sum; then the true score again, shifted like every candidate
        cq = __fdiv_rn(acc0, (float)((P.H + kCenChunk - 1) / kCenChunk));
        s_cmp = pair_score_warp<MODEL, HEAD>(Qs + q * qdim, trow, P.H, c, a, b, strip, cq);
    }
    if ((threadIdx.x & 31) == 0) {
        true_score[q] = s;   // the plain score: what the caller sees
        true_cmp[q] = s_cmp; // what candidates are compared with
        cen[q] = cq;
    }
}

// ---- 5 filter --------------------------------------------------------------------------------
// one warp per filter entry (grid-stride; the number of entries lives on the device)
template <int MODEL, bool HEAD>
__global__ void __launch_bounds__(32 * kPairWarps)
eval_filter_kernel(Params P, EvalConsts c, const float* __restrict__ modulus_p, const float* __restrict__ ent,
                   const float* __restrict__ Qs, const float* __restrict__ ina, const float* __restrict__ inb,
                   const int64_t* __restrict__ queries, long long Q, const int64_t* __restrict__ indptr,
                   const int64_t* __restrict__ idx, long long e_lo, long long e_hi,
                   const float* __restrict__ true_score, const float* __restrict__ cen, unsigned long long* n_greater,
                   unsigned long long* n_equal) {
    extern __shared__ __align__(16) float smem[];
    float* strip = smem + (size_t)(threadIdx.x >> 5) * pair_strips<MODEL>() * kPairChunk;
    const int lane = threadIdx.x & 31;
    const long long total = indptr[Q];
    if (MODEL == KGE_PROTATE) c.modulus = __ldg(modulus_p);
    const int qdim = model_planes(MODEL) * P.H;
    for (long long f = (long long)blockIdx.x * kPairWarps + (threadIdx.x >> 5); f < total;
         f += (long long)gridDim.x * kPairWarps) {
        // query owning entry f: last q with indptr[q] <= f
        long long lo = 0, hi = Q;
        while (hi - lo > 1) {
            const long long mid = (lo + hi) >> 1;
            if (indptr[mid] <= f) lo = mid;
            else hi = mid;
        }
        const long long q = lo;
        const long long e = idx[f];
        const long long te = HEAD ? queries[3 * q] : queries[3 * q + 2];
        if (e == te || e < e_lo || e >= e_hi) continue;
        // a candidate listed twice must be undone once: only its first occurrence in the row acts
        bool dup = false;
        for (long long j = indptr[q] + lane; j < f; j += 32) dup |= (idx[j] == e);
        if (__any_sync(0xffffffffu, dup)) continue;
        const float a = (MODEL == KGE_INTERHT) ? ina[e] : 0.f, b = (MODEL == KGE_INTERHT) ? inb[e] : 0.f;
        const float s = pair_score_warp<MODEL, HEAD>(Qs + q * qdim, ent + e * (long long)P.De, P.H, c, a, b, strip,
                                                     centered<MODEL>() ? cen[q] : 0.f);
        if (lane == 0) {
            const float t = true_score[q];
            if (s > t) atomicAdd(n_greater + q, ~0ULL);      // -1
            else if (s == t) atomicAdd(n_equal + q, ~0ULL);  // -1
        }
    }
}

// ---- 4 tile ----------------------------------------------------------------------------------
constexpr int kTQ = 128, kTE = 128, kKC = 16, kEvalThreads = 256;

template <int MODEL>
constexpr size_t eval_tile_smem() {
    return (size_t)2 * (model_planes(MODEL) + cand_planes(MODEL)) * kKC * kTQ * sizeof(float);
}

template <int MODEL, bool HEAD>
__global__ void __launch_bounds__(kEvalThreads)
eval_tile_kernel(Params P, EvalConsts c, const float* __restrict__ modulus_p, const float* __restrict__ ent,
                 const float* __restrict__ Qs, const float* __restrict__ ina, const float* __restrict__ inb,
                 const int64_t* __restrict__ queries, long long Q, long long e_lo, long long e_hi,
                 const float* __restrict__ true_score, const float* __restrict__ cen, unsigned long long* n_greater,
                 unsigned long long* n_equal) {
    constexpr int QP = model_planes(MODEL), CP = cand_planes(MODEL);
    static_assert(kCenChunk == kKC, "the centring chunk is the tile kernel's k chunk");
    extern __shared__ __align__(16) float smem[];
    __shared__ __align__(16) float cen_s[kTQ];  // per-query shift of the running sum (0 for the dot-product models)
    // [buf][plane][k][row]
    float* Qt = smem;
    float* Et = smem + 2 * QP * kKC * kTQ;
    const int H = P.H, qdim = QP * H;
    const int tid = threadIdx.x;
    if (MODEL == KGE_PROTATE) c.modulus = __ldg(modulus_p);

    const long long e0 = e_lo + (long long)blockIdx.x * kTE;
    const long long q0 = (long long)blockIdx.y * kTQ;

    // loader role: one row of each tile, 8 consecutive coordinates per chunk
    const int lrow = tid & 127, lk = (tid >> 7) * 8;
    const long long lq = min(q0 + lrow, Q - 1);
    const long long le = min(e0 + lrow, e_hi - 1);
    const float* qsrc = Qs + lq * qdim;
    const float* esrc = ent + le * (long long)P.De;
    const float l_ina = (MODEL == KGE_INTERHT) ? ina[le] : 0.f, l_inb = (MODEL == KGE_INTERHT) ? inb[le] : 0.f;
    const bool vec = (H % 4 == 0);

    float qreg[QP][8], ereg[CP][8];
    auto fetch = [&](int kbase) {
#pragma unroll
        for (int p = 0; p < QP; ++p) {
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                const int k = kbase + lk + 4 * v;
                if (vec && k + 3 < H) {
                    const float4 x = *reinterpret_cast<const float4*>(qsrc + p * H + k);
                    qreg[p][4 * v] = x.x; qreg[p][4 * v + 1] = x.y; qreg[p][4 * v + 2] = x.z; qreg[p][4 * v + 3] = x.w;
                } else {
#pragma unroll
                    for (int i = 0; i < 4; ++i) qreg[p][4 * v + i] = (k + i < H) ? qsrc[p * H + k + i] : 0.f;
                }
            }
        }
#pragma unroll
        for (int p = 0; p < CP; ++p) {
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                const int k = kbase + lk + 4 * v;
                if (vec && k + 3 < H) {
                    const float4 x = __ldg(reinterpret_cast<const float4*>(esrc + p * H + k));
                    ereg[p][4 * v] = x.x; ereg[p][4 * v + 1] = x.y; ereg[p][4 * v + 2] = x.z; ereg[p][4 * v + 3] = x.w;
                } else {
#pragma unroll
                    for (int i = 0; i < 4; ++i) ereg[p][4 * v + i] = (k + i < H) ? __ldg(esrc + p * H + k + i) : 0.f;
                }
            }
        }
    };
    auto stash = [&](int buf, int kbase) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const bool live = (kbase + lk + i) < H;  // padding coordinates contribute exactly +0
#pragma unroll
            for (int p = 0; p < QP; ++p) Qt[((buf * QP + p) * kKC + lk + i) * kTQ + lrow] = qreg[p][i];
            float c0, c1;
            prep_candidate<MODEL>(c0, c1, ereg[0][i], (CP > 1) ? ereg[CP - 1][i] : 0.f, c.inv_phase_div, l_ina, l_inb);
            if (!live) { c0 = 0.f; c1 = 0.f; }
            Et[((buf * CP + 0) * kKC + lk + i) * kTE + lrow] = c0;
            if (CP > 1) Et[((buf * CP + 1) * kKC + lk + i) * kTE + lrow] = c1;
        }
    };

    // compute role: 8 queries x 8 entities, split 4 + 4 so that LDS.128 is conflict-free
    const int ty = tid >> 4, tx = tid & 15;
    float2 acc[8][4];  // [query][candidate pair]
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = make_float2(0.f, 0.f);

    const int nchunks = (H + kKC - 1) / kKC;
    if (tid < kTQ) cen_s[tid] = (centered<MODEL>() && q0 + tid < Q) ? cen[q0 + tid] : 0.f;
    fetch(0);
    stash(0, 0);
    __syncthreads();
    for (int ch = 0; ch < nchunks; ++ch) {
        const int buf = ch & 1;
        if (ch + 1 < nchunks) fetch((ch + 1) * kKC);
#pragma unroll 4
        for (int k = 0; k < kKC; ++k) {
            float qv[QP][8], ev[CP][8];
#pragma unroll
            for (int p = 0; p < QP; ++p) {
                const float* src = Qt + ((buf * QP + p) * kKC + k) * kTQ;
                const float4 a = *reinterpret_cast<const float4*>(src + ty * 4);
                const float4 b = *reinterpret_cast<const float4*>(src + 64 + ty * 4);
                qv[p][0] = a.x; qv[p][1] = a.y; qv[p][2] = a.z; qv[p][3] = a.w;
                qv[p][4] = b.x; qv[p][5] = b.y; qv[p][6] = b.z; qv[p][7] = b.w;
            }
#pragma unroll
            for (int p = 0; p < CP; ++p) {
                const float* src = Et + ((buf * CP + p) * kKC + k) * kTE;
                const float4 a = *reinterpret_cast<const float4*>(src + tx * 4);
                const float4 b = *reinterpret_cast<const float4*>(src + 64 + tx * 4);
                ev[p][0] = a.x; ev[p][1] = a.y; ev[p][2] = a.z; ev[p][3] = a.w;
                ev[p][4] = b.x; ev[p][5] = b.y; ev[p][6] = b.z; ev[p][7] = b.w;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float q0 = qv[0][i], q1 = (QP > 1) ? qv[QP > 1 ? 1 : 0][i] : 0.f;
                    if constexpr (packed_accum<MODEL>()) {
                        acc[i][j] = accum2<MODEL, HEAD>(acc[i][j], q0, q1, make_float2(ev[0][2 * j], ev[0][2 * j + 1]),
                                                        (CP > 1) ? make_float2(ev[CP - 1][2 * j], ev[CP - 1][2 * j + 1])
                                                                 : make_float2(0.f, 0.f));
                    } else {
                        const float q2 = (QP > 2) ? qv[QP > 2 ? 2 : 0][i] : 0.f;
                        acc[i][j].x = accum<MODEL, HEAD>(acc[i][j].x, q0, q1, q2, ev[0][2 * j], (CP > 1) ? ev[CP - 1][2 * j] : 0.f);
                        acc[i][j].y = accum<MODEL, HEAD>(acc[i][j].y, q0, q1, q2, ev[0][2 * j + 1],
                                                         (CP > 1) ? ev[CP - 1][2 * j + 1] : 0.f);
                    }
                }
        }
        if constexpr (centered<MODEL>()) {  // shift the running sums: acc - cen_q (== acc + (-cen_q), one rounding)
            const float4 ca = *reinterpret_cast<const float4*>(cen_s + ty * 4);
            const float4 cb = *reinterpret_cast<const float4*>(cen_s + 64 + ty * 4);
            const float cq[8] = {ca.x, ca.y, ca.z, ca.w, cb.x, cb.y, cb.z, cb.w};
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float2 m = make_float2(-cq[i], -cq[i]);
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = __fadd2_rn(acc[i][j], m);
            }
        }
        if (ch + 1 < nchunks) {
            stash(buf ^ 1, (ch + 1) * kKC);  // the other buffer was last read in iteration ch-1
            __syncthreads();
        }
    }

    // epilogue: compare with the true score, reduce over the 16 threads that share the query rows
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const long long q = q0 + ((i < 4) ? (ty * 4 + i) : (64 + ty * 4 + i - 4));
        const bool qok = q < Q;
        const float t = qok ? true_score[q] : 0.f;
        const long long te = qok ? (HEAD ? queries[3 * q] : queries[3 * q + 2]) : -1;
        int gt = 0, eq = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const long long e = e0 + ((j < 4) ? (tx * 4 + j) : (64 + tx * 4 + j - 4));
            const float s = finalize<MODEL>((j & 1) ? acc[i][j >> 1].y : acc[i][j >> 1].x, c);
            const bool ok = qok && e < e_hi && e != te;
            gt += (ok && s > t) ? 1 : 0;
            eq += (ok && s == t) ? 1 : 0;
        }
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
            gt += __shfl_xor_sync(0xffffffffu, gt, o);
            eq += __shfl_xor_sync(0xffffffffu, eq, o);
        }
        if (tx == 0 && qok) {
            if (gt) atomicAdd(n_greater + q, (unsigned long long)gt);
            if (eq) atomicAdd(n_equal + q, (unsigned long long)eq);
        }
    }
}

// ---- launcher --------------------------------------------------------------------------------
static inline size_t ev_align(size_t x) { return (x + 255) & ~(size_t)255; }

size_t eval_workspace_bytes(const Params& P, int64_t Q, int64_t nfilter) {
    if (eval_mma_applies(P)) return eval_mma_workspace_bytes(P, Q, nfilter);
    size_t total = ev_align((size_t)Q * model_planes(P.model) * P.H * sizeof(float)) + 2 * ev_align((size_t)Q * sizeof(float));
    if (P.model == KGE_INTERHT) total += 2 * ev_align((size_t)P.E * sizeof(float)) + 2 * ev_align((size_t)Q * sizeof(float));
    return total + 256;
}

template <int MODEL, bool HEAD>
static cudaError_t launch_eval_model(const Params& P, const float* ent, const float* rel, const float* modulus,
                                     const int64_t* queries, const float* qrows, int64_t Q, int64_t e_lo, int64_t e_hi,
                                     const int64_t* indptr, const int64_t* idx, float* true_score,
                                     unsigned long long* n_greater, unsigned long long* n_equal, int accumulate,
                                     void* ws, cudaStream_t st) {
    const int qdim = model_planes(MODEL) * P.H;
    float* Qs = (float*)ws;
    float* true_cmp = (float*)((char*)ws + ev_align((size_t)Q * qdim * sizeof(float)));  // shifted true scores
    float* cen = (float*)((char*)true_cmp + ev_align((size_t)Q * sizeof(float)));        // per-query shifts
    float *ina = nullptr, *inb = nullptr, *tna = nullptr, *tnb = nullptr;
    if (MODEL == KGE_INTERHT) {
        // candidate norms for rows [e_lo, e_hi) only, addressed by global id through an offset base
        float* na = (float*)((char*)cen + ev_align((size_t)Q * sizeof(float)));
        float* nb = (float*)((char*)na + ev_align((size_t)P.E * sizeof(float)));
        tna = (float*)((char*)nb + ev_align((size_t)P.E * sizeof(float)));
        tnb = (float*)((char*)tna + ev_align((size_t)Q * sizeof(float)));
        ina = na - e_lo;
        inb = nb - e_lo;
    }
    EvalConsts c{P.gamma, P.inv_phase_div, 0.f};
    cudaError_t e;
    if (!accumulate) {
        if ((e = cudaMemsetAsync(n_greater, 0, (size_t)Q * sizeof(unsigned long long), st)) != cudaSuccess) return e;
        if ((e = cudaMemsetAsync(n_equal, 0, (size_t)Q * sizeof(unsigned long long), st)) != cudaSuccess) return e;
    }
    const size_t smem_stage = (size_t)(qdim + 4 + 32) * sizeof(float);
    if (smem_stage > 48 * 1024)
        cudaFuncSetAttribute(eval_stage_kernel<MODEL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_stage);
    eval_stage_kernel<MODEL><<<(unsigned)Q, 128, smem_stage, st>>>(P, HEAD ? 1 : 0, ent, rel, queries, qrows, Qs);
    if (MODEL == KGE_INTERHT) {
        if (e_hi > e_lo) {
            const int blocks = (int)std::min<long long>((e_hi - e_lo + 7) / 8, 16LL * sm_count());
            eval_inv_norm_kernel<<<blocks, 256, 0, st>>>(P, ent, e_lo, e_hi, ina + e_lo, inb + e_lo);
        }
        eval_true_norm_kernel<HEAD><<<(unsigned)((Q * 32 + 255) / 256), 256, 0, st>>>(P, ent, qrows, queries, Q, tna, tnb);
    }
    constexpr size_t smem_pair = pair_smem<MODEL>();
    if (smem_pair > 48 * 1024) {
        cudaFuncSetAttribute(eval_true_kernel<MODEL, HEAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_pair);
        cudaFuncSetAttribute(eval_filter_kernel<MODEL, HEAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_pair);
    }
    eval_true_kernel<MODEL, HEAD><<<(unsigned)((Q + kPairWarps - 1) / kPairWarps), 32 * kPairWarps, smem_pair, st>>>(
        P, c, modulus, ent, qrows, Qs, tna, tnb, queries, Q, true_score, true_cmp, cen);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if (e_hi > e_lo) {
        constexpr size_t smem = eval_tile_smem<MODEL>();
        cudaFuncSetAttribute(eval_tile_kernel<MODEL, HEAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        dim3 grid((unsigned)((e_hi - e_lo + kTE - 1) / kTE), (unsigned)((Q + kTQ - 1) / kTQ));
        eval_tile_kernel<MODEL, HEAD><<<grid, kEvalThreads, smem, st>>>(P, c, modulus, ent, Qs, ina, inb, queries, Q, e_lo,
                                                                         e_hi, true_cmp, cen, n_greater, n_equal);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
    }
    if (indptr != nullptr && idx != nullptr) {
        // the number of filter entries lives on the device (indptr[Q]): grid-stride over it, no host sync
        eval_filter_kernel<MODEL, HEAD><<<8 * sm_count(), 32 * kPairWarps, smem_pair, st>>>(
            P, c, modulus, ent, Qs, ina, inb, queries, Q, indptr, idx, e_lo, e_hi, true_cmp, cen, n_greater, n_equal);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
    }
    return cudaSuccess;
}

cudaError_t launch_eval_stage(const Params& P, bool head, const float* ent, const float* rel, const int64_t* queries,
                              const float* qrows, int64_t Q, float* Qs, cudaStream_t st) {
    const size_t smem = (size_t)(model_planes(P.model) * P.H + 4 + 32) * sizeof(float);
#define KGE_CASE(M)                                                                                                  \
    case M:                                                                                                          \
        if (smem > 48 * 1024)                                                                                        \
            cudaFuncSetAttribute(eval_stage_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);      \
        eval_stage_kernel<M><<<(unsigned)Q, 128, smem, st>>>(P, head ? 1 : 0, ent, rel, queries, qrows, Qs);        \
        break
    switch (P.model) {
        KGE_CASE(KGE_TRANSE);
        KGE_CASE(KGE_DISTMULT);
        KGE_CASE(KGE_COMPLEX);
        KGE_CASE(KGE_ROTATE);
        KGE_CASE(KGE_PROTATE);
        KGE_CASE(KGE_INTERHT);
        default: return cudaErrorInvalidValue;
    }
#undef KGE_CASE
    return cudaGetLastError();
}

cudaError_t launch_eval_rank(const Params& P, int mode, const float* ent, const float* rel, const float* modulus,
                             const int64_t* queries, const float* qrows, int64_t Q, int64_t e_lo, int64_t e_hi,
                             const int64_t* filter_indptr, const int64_t* filter_idx, float* true_score,
                             int64_t* n_greater, int64_t* n_equal, int accumulate, void* ws, size_t ws_bytes,
                             cudaStream_t st) {
    auto* ng = reinterpret_cast<unsigned long long*>(n_greater);
    auto* ne = reinterpret_cast<unsigned long long*>(n_equal);
    if (eval_mma_applies(P))
        return launch_eval_rank_mma(P, mode, ent, rel, queries, qrows, Q, e_lo, e_hi, filter_indptr, filter_idx, true_score, ng,
                                    ne, accumulate, ws, ws_bytes, st);
    if (ws_bytes < eval_workspace_bytes(P, Q, 0)) return cudaErrorInvalidValue;
    const bool head = (mode == KGE_MODE_HEAD_BATCH);
#define KGE_CASE(M)                                                                                                  \
    case M:                                                                                                          \
        return head ? launch_eval_model<M, true>(P, ent, rel, modulus, queries, qrows, Q, e_lo, e_hi, filter_indptr, \
                                                 filter_idx, true_score, ng, ne, accumulate, ws, st)                 \
                    : launch_eval_model<M, false>(P, ent, rel, modulus, queries, qrows, Q, e_lo, e_hi, filter_indptr,\
                                                  filter_idx, true_score, ng, ne, accumulate, ws, st)
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
