// K5-T: tcgen05 path of the filtered full-entity ranking for the GEMM-shaped models (sm_100a).
//
// For DistMult and ComplEx the score of (query q, candidate e) is a dot product over K = entity_dim
// coordinates: S = Qs [Q,K] x Ent^T [K,E].  This file runs that product on the 5th-gen
// tensor cores and keeps the ranking-by-counting epilogue of K5 (kge_eval.cu): the [Q,E] matrix never leaves
// the SM, two integer counters per query do.
//
//   precision   3xTF32: every fp32 operand x is split into hi = tf32(x) and lo = tf32(x - hi); the tile
//               accumulates lo*hi + hi*lo + hi*hi in fp32 (TMEM).  Relative error per product <= ~3 * 2^-22,
//               i.e. the same order as fp32 summation-order noise over K terms.
//   operands    split once per call into K-major workspaces [rows, Kpad] (Kpad = K rounded up to 32, zero
//               filled), read by TMA (cp.async.bulk.tensor.2d, 128-byte swizzle) into a 2-stage ring of
//               96 KB stages: A = 128 queries x 32 k (hi, lo), B = 256 entities x 32 k (hi, lo).
//   MMA         one elected thread: 12 x tcgen05.mma.cta_group::1.kind::tf32 (M128 N256 K8) per stage,
//               tcgen05.commit releases the stage / publishes the accumulator.
//   accumulator 2 x 256 TMEM columns (all 512): the epilogue of tile i overlaps the MMAs of tile i+1.
//   epilogue    4 warps, thread = query row (TMEM lane), tcgen05.ld 32x32b.x32; each score is compared with
//               the query's true score; the true entity and the filtered candidates are skipped BY ID
//               (sorted per-query filter list + cursor), so no bit-exact re-evaluation is needed.
//   true score  the same kernel in DIAG mode: B = the gathered true-entity rows of the 128 queries, the
//               diagonal of the tile is the true score -- same arithmetic as every other candidate, so a
//               candidate with an identical row ties exactly (n_equal keeps its meaning).
//   schedule    persistent CTAs (one per SM), tile = (entity tile, query tile) with the query tile fastest so
//               that concurrently running CTAs share the B tile in L2.
//   CTA pairs   (count mode, default) two CTAs of a cluster take two query tiles against one entity tile as ONE
//               256 x 256 tile: tcgen05.mma.cta_group::2 issued by the leader, each CTA stages its queries and half
//               of the entity tile (64 KB per stage and SM, 3 stages) -- see the PAIR comment at the kernel.
#include <climits>
#include <cstdlib>

#include "kge_umma.cuh"

namespace kge {
namespace mma {

constexpr int kThreads = 192;  // warp 0 TMA producer, warp 1 MMA issuer + TMEM owner, warps 2..5 epilogue
constexpr long long kEntityChunk = 32768;  // rows of the split B workspace per pass

// rows of `src` (pitch floats apart; row = gather ? row named by the query : row0 + r * row_stride) -> hi / lo [rows, Kpad]
__global__ void __launch_bounds__(256)
split_rows_kernel(const float* __restrict__ src, long long pitch, long long row0, int row_stride,
                  const int64_t* __restrict__ queries, int gather_col, long long rows, int K, int Kpad,
                  float* __restrict__ hi, float* __restrict__ lo) {
    const int k4n = Kpad >> 2;
    const long long total = rows * k4n;
    const bool vec = (K % 4 == 0) && (pitch % 4 == 0) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / k4n;
        const int k = (int)(i - r * k4n) * 4;
        const long long srow = (gather_col >= 0) ? queries[3 * r + gather_col] : row0 + r * row_stride;
        const float* p = src + srow * pitch + k;
        float x[4];
        if (vec && k + 3 < K) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(p));
            x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w;
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) x[j] = (k + j < K) ? __ldg(p + j) : 0.f;
        }
        float h[4], l[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) split_tf32(x[j], h[j], l[j]);
        *reinterpret_cast<float4*>(hi + r * Kpad + k) = make_float4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<float4*>(lo + r * Kpad + k) = make_float4(l[0], l[1], l[2], l[3]);
    }
}

// ---- per-query sorted filter lists ------------------------------------------------------------------
// One warp per query: rank sort of its CSR segment (ties broken by position) into int32 ids.  Segments are
// short (wn18rr: mean 1.4-2.1, p99 14, max 473 -- SURVEY 8(c)); O(n^2 / 32) per warp.
__global__ void __launch_bounds__(256)
filter_sort_kernel(const int64_t* __restrict__ indptr, const int64_t* __restrict__ idx, long long Q, int* __restrict__ sidx) {
    const int lane = threadIdx.x & 31;
    const long long q = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (q >= Q) return;
    const long long b = indptr[q], n = indptr[q + 1] - b;
    for (long long i = lane; i < n; i += 32) {
        const int64_t x = idx[b + i];
        long long rank = 0;
        for (long long j = 0; j < n; ++j) {
            const int64_t y = idx[b + j];
            rank += (y < x || (y == x && j < i)) ? 1 : 0;
        }
        sidx[b + rank] = (int)x;
    }
}

// ---- the tile kernel ------------------------------------------------------------------------------
struct TileArgs {
    long long Q;
    long long b_row_id0;  // global entity id of row 0 of the B workspace (count mode)
    long long b_rows;     // rows of the B workspace that are candidates
    int nkb;              // Kpad / BK
    int nqt;              // query tiles
    int net;              // entity tiles (count mode)
    int true_col;         // column of `queries` naming the true entity: 0 head-batch, 2 tail-batch
    const int64_t* queries;
    const int64_t* indptr;  // CSR offsets of the filter lists (NULL = unfiltered)
    const int* sidx;        // per-query sorted filter ids
    float* true_score;
    unsigned long long* n_greater;
    unsigned long long* n_equal;
};

// PAIR (count mode only): clusters of two CTAs work on two query tiles against the same entity tile as ONE 256 x 256 tile
// (tcgen05.mma.cta_group::2): each CTA stages its own 128 queries and HALF of the entity tile -- 64 KB instead of 96 KB per
// k-stage and SM, three stages instead of two.  The one-CTA kernel is bound by those L2 -> SM bytes, not by the tensor
// pipe (kge_transparse.cu has the same pipeline with a store epilogue and the measurements).  Protocol: both producers
// signal the LEADER's `full` barrier, the leader issues every MMA, its commits reach both CTAs' `empty` / `tfull` barriers,
// both CTAs' epilogue warps release the accumulator on the leader's `tempty`.
template <bool DIAG, bool PAIR>
__global__ void __launch_bounds__(kThreads, 1)
eval_mma_kernel(const __grid_constant__ CUtensorMap tmAhi, const __grid_constant__ CUtensorMap tmAlo,
                const __grid_constant__ CUtensorMap tmBhi, const __grid_constant__ CUtensorMap tmBlo, const TileArgs a) {
    static_assert(!(DIAG && PAIR), "the true-score pass has no shared B tile");
    constexpr int NST = PAIR ? STAGES2 : STAGES;
    constexpr uint32_t STB = PAIR ? STAGE2_BYTES : STAGE_BYTES;
    constexpr uint32_t BHALF = PAIR ? BH_BYTES : B_BYTES;  // bytes of the B block (hi or lo) this CTA stages
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B atoms are 1024-byte aligned
    const uint32_t bars = base + NST * STB;
    // barriers: full[s] @ +0, empty[s] @ +24, tfull[a] @ +48, tempty[a] @ +64, TMEM base address @ +80
    auto full = [&](int s) { return bars + 8u * s; };
    auto empty = [&](int s) { return bars + 24u + 8u * s; };
    auto tfull = [&](int s) { return bars + 48u + 8u * s; };
    auto tempty = [&](int s) { return bars + 64u + 8u * s; };
    const uint32_t tmem_slot = bars + 80u;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t crank = 0;
    if constexpr (PAIR) crank = cluster_ctarank();
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmAhi);
        tma_prefetch_desc(&tmAlo);
        tma_prefetch_desc(&tmBhi);
        tma_prefetch_desc(&tmBlo);
        for (int s = 0; s < NST; ++s) {
            mbar_init(full(s), 1);   // the (leader's) producer's arrive.expect_tx
            mbar_init(empty(s), 1);  // tcgen05.commit of the MMA thread
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(tfull(s), 1);              // tcgen05.commit after the last k-block
            mbar_init(tempty(s), PAIR ? 8 : 4);  // one lane of each epilogue warp (of both CTAs of a pair)
        }
        fence_barrier_init();
    }
    if (warp == 1) {
        if constexpr (PAIR) tmem_alloc_pair(tmem_slot, TMEM_COLS);
        else tmem_alloc(tmem_slot, TMEM_COLS);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if constexpr (PAIR) cluster_sync_all();  // the peer's barriers and TMEM exist before anything is sent to them
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

    // work items: DIAG one per query tile; count mode one per (entity tile, query tile), the query tile fastest; PAIR one
    // per (entity tile, PAIR of query tiles) and cluster -- this CTA takes query tile 2 * pair + rank (a tile past the last
    // query reads zeros and counts nothing)
    const int hq = (a.nqt + 1) / 2;
    const int nwork = DIAG ? a.nqt : (PAIR ? hq * a.net : a.nqt * a.net);
    const int w0 = PAIR ? (int)cluster_id_x() : (int)blockIdx.x, wstride = PAIR ? (int)cluster_count_x() : (int)gridDim.x;
    auto tile_qt = [&](int w) { return DIAG ? w : (PAIR ? 2 * (w % hq) + (int)crank : w % a.nqt); };
    auto tile_et = [&](int w) { return PAIR ? w / hq : w / a.nqt; };

    if (warp == 0) {
        if (lane == 0) {
            // ===== TMA producer =====
            uint32_t it = 0;
            for (int w = w0; w < nwork; w += wstride) {
                const int qt = tile_qt(w);
                const int brow = DIAG ? qt * BM : tile_et(w) * BN + (PAIR ? (int)crank * (BN / 2) : 0);
                for (int kb = 0; kb < a.nkb; ++kb, ++it) {
                    const int s = it % NST;
                    mbar_wait(empty(s), ((it / NST) & 1u) ^ 1u);
                    const uint32_t st = base + s * STB;
                    if constexpr (PAIR) {
                        if (crank == 0) mbar_arrive_expect_tx(full(s), 2 * STB);  // the bytes of both CTAs
                        const uint32_t lf = map_to_rank(full(s), 0);
                        tma_load_2d_pair(st, &tmAhi, lf, kb * BK, qt * BM);
                        tma_load_2d_pair(st + A_BYTES, &tmAlo, lf, kb * BK, qt * BM);
                        tma_load_2d_pair(st + 2 * A_BYTES, &tmBhi, lf, kb * BK, brow);
                        tma_load_2d_pair(st + 2 * A_BYTES + BHALF, &tmBlo, lf, kb * BK, brow);
                    } else {
                        mbar_arrive_expect_tx(full(s), STB);
                        tma_load_2d(st, &tmAhi, full(s), kb * BK, qt * BM);
                        tma_load_2d(st + A_BYTES, &tmAlo, full(s), kb * BK, qt * BM);
                        tma_load_2d(st + 2 * A_BYTES, &tmBhi, full(s), kb * BK, brow);
                        tma_load_2d(st + 2 * A_BYTES + BHALF, &tmBlo, full(s), kb * BK, brow);
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && crank == 0) {
            // ===== MMA issuer (the leader's, in a pair) =====
            uint32_t it = 0, tcount = 0;
            for (int w = w0; w < nwork; w += wstride, ++tcount) {
                const int acc = tcount & 1u;
                mbar_wait(tempty(acc), ((tcount >> 1) & 1u) ^ 1u);
                tc_fence_after();
                const uint32_t d = tmem_base + acc * BN;
                for (int kb = 0; kb < a.nkb; ++kb, ++it) {
                    const int s = it % NST;
                    mbar_wait(full(s), (it / NST) & 1u);
                    tc_fence_after();
                    const uint32_t st = base + s * STB;
                    const uint64_t ahi = umma_desc_sw128(st), alo = umma_desc_sw128(st + A_BYTES);
                    const uint64_t bhi = umma_desc_sw128(st + 2 * A_BYTES), blo = umma_desc_sw128(st + 2 * A_BYTES + BHALF);
                    // small terms first: lo*hi, hi*lo, then hi*hi
                    if constexpr (PAIR) {
#pragma unroll
                        for (int k = 0; k < BK / UMMA_K; ++k) umma_tf32_pair(d, alo + 2 * k, bhi + 2 * k, (kb | k) != 0);
#pragma unroll
                        for (int k = 0; k < BK / UMMA_K; ++k) umma_tf32_pair(d, ahi + 2 * k, blo + 2 * k, 1u);
#pragma unroll
                        for (int k = 0; k < BK / UMMA_K; ++k) umma_tf32_pair(d, ahi + 2 * k, bhi + 2 * k, 1u);
                        umma_commit_pair(empty(s));
                    } else {
#pragma unroll
                        for (int k = 0; k < BK / UMMA_K; ++k) umma_tf32(d, alo + 2 * k, bhi + 2 * k, (kb | k) != 0);
#pragma unroll
                        for (int k = 0; k < BK / UMMA_K; ++k) umma_tf32(d, ahi + 2 * k, blo + 2 * k, 1u);
#pragma unroll
                        for (int k = 0; k < BK / UMMA_K; ++k) umma_tf32(d, ahi + 2 * k, bhi + 2 * k, 1u);
                        umma_commit(empty(s));  // frees the stage once these MMAs have read it
                    }
                }
                if constexpr (PAIR) umma_commit_pair(tfull(acc));
                else umma_commit(tfull(acc));  // accumulator complete
            }
        }
    } else {
        // ===== epilogue: warps 2..5, TMEM lane quadrant = warp % 4, thread = query row =====
        const int wq = warp & 3;
        const int row = wq * 32 + lane;
        uint32_t tcount = 0;
        for (int w = w0; w < nwork; w += wstride, ++tcount) {
            const int acc = tcount & 1u;
            const int qt = tile_qt(w);
            const long long q = (long long)qt * BM + row;
            const bool qok = q < a.Q;
            const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16) + acc * BN;
            if constexpr (DIAG) {
                mbar_wait(tfull(acc), (tcount >> 1) & 1u);
                tc_fence_after();
                float v[32];
                tmem_ld32(taddr + wq * 32, v);  // columns = the rows this warp owns: the diagonal block
                float s = 0.f;
#pragma unroll
                for (int j = 0; j < 32; ++j) s = (j == lane) ? v[j] : s;
                if (qok) a.true_score[q] = s;
            } else {
                const long long e0 = a.b_row_id0 + (long long)tile_et(w) * BN;  // global id of column 0
                const long long e_end = a.b_row_id0 + a.b_rows;
                float t = 0.f;
                int te = -1, cur = 0, fe = 0, nextf = INT_MAX;
                if (qok) {
                    t = a.true_score[q];
                    te = (int)a.queries[3 * q + a.true_col];
                    if (a.indptr != nullptr) {
                        int lo = (int)a.indptr[q];
                        fe = (int)a.indptr[q + 1];
                        int hi = fe;  // first entry >= e0
                        while (lo < hi) {
                            const int mid = (lo + hi) >> 1;
                            if (a.sidx[mid] < e0) lo = mid + 1;
                            else hi = mid;
                        }
                        cur = lo;
                        if (cur < fe) nextf = a.sidx[cur];
                    }
                }
                mbar_wait(tfull(acc), (tcount >> 1) & 1u);
                tc_fence_after();
                int gt = 0, eq = 0;
#pragma unroll 1
                for (int c = 0; c < BN / 32; ++c) {
                    float v[32];
                    __syncwarp();  // tcgen05.ld is .sync.aligned: reconverge after the per-lane paths below
                    tmem_ld32(taddr + c * 32, v);
                    const long long eb = e0 + c * 32;
                    const bool live = qok && eb < e_end;
                    const bool special = ((unsigned long long)((long long)te - eb) < 32ull) ||
                                         ((unsigned long long)((long long)nextf - eb) < 32ull) || (eb + 32 > e_end);
                    if (!live) {
                        // padding rows / columns past the candidate range: nothing to count
                    } else if (!special) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            gt += (v[j] > t) ? 1 : 0;
                            eq += (v[j] == t) ? 1 : 0;
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const long long e = eb + j;
                            bool ok = (e < e_end) && (e != te);
                            if (e == nextf) {
                                ok = false;
                                do {
                                    ++cur;
                                    nextf = (cur < fe) ? a.sidx[cur] : INT_MAX;
                                } while (nextf == e);
                            }
                            gt += (ok && v[j] > t) ? 1 : 0;
                            eq += (ok && v[j] == t) ? 1 : 0;
                        }
                    }
                }
                if (gt) atomicAdd(a.n_greater + q, (unsigned long long)gt);
                if (eq) atomicAdd(a.n_equal + q, (unsigned long long)eq);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if constexpr (PAIR) mbar_arrive_cluster(map_to_rank(tempty(acc), 0));
                else mbar_arrive(tempty(acc));
            }
        }
    }
    __syncwarp();
    tc_fence_before();
    __syncthreads();
    if constexpr (PAIR) cluster_sync_all();  // no CTA leaves (or frees its TMEM) while its peer may still use it
    if (warp == 1) {
        __syncwarp();
        tc_fence_after();
        if constexpr (PAIR) tmem_dealloc_pair(tmem_base, TMEM_COLS);
        else tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ---- host side ----------------------------------------------------------------------------------
static inline size_t al(size_t x) { return (x + 1023) & ~(size_t)1023; }
static inline int kpad_of(int K) { return (K + BK - 1) / BK * BK; }

struct Layout {
    size_t qs, ahi, alo, thi, tlo, bhi, blo, sidx, total;
};
static Layout layout(const Params& P, int64_t Q, int64_t nfilter) {
    const int Kpad = kpad_of(P.De);
    const size_t qk = al((size_t)Q * Kpad * sizeof(float));
    const size_t bk = al((size_t)std::min<long long>(std::max<long long>(P.E, 1), kEntityChunk) * Kpad * sizeof(float));
    Layout L;
    size_t o = 0;
    L.qs = o;   o += al((size_t)Q * P.De * sizeof(float));
    L.ahi = o;  o += qk;
    L.alo = o;  o += qk;
    L.thi = o;  o += qk;
    L.tlo = o;  o += qk;
    L.bhi = o;  o += bk;
    L.blo = o;  o += bk;
    L.sidx = o; o += al((size_t)std::max<int64_t>(nfilter, 1) * sizeof(int));
    L.total = o + 1024;  // slack to align the caller's pointer
    return L;
}

}  // namespace mma

static int g_tensor_path = 1;
// CTA pairs in the counting pass (default); KGE_EVAL_PAIR=0 selects the one-CTA kernel
static bool eval_pair() {
    static int v = -1;
    if (v < 0) {
        const char* s = getenv("KGE_EVAL_PAIR");
        v = (s != nullptr && s[0] == '0') ? 0 : 1;
    }
    return v != 0;
}
void eval_set_tensor_path(int on) { g_tensor_path = on ? 1 : 0; }
bool eval_mma_applies(const Params& P) {
    return g_tensor_path && (P.model == KGE_DISTMULT || P.model == KGE_COMPLEX) && P.E < (1LL << 31);
}
size_t eval_mma_workspace_bytes(const Params& P, int64_t Q, int64_t nfilter) { return mma::layout(P, Q, nfilter).total; }

cudaError_t launch_eval_rank_mma(const Params& P, int mode, const float* ent, const float* rel, const int64_t* queries,
                                 const float* qrows, int64_t Q, int64_t e_lo, int64_t e_hi, const int64_t* indptr, const int64_t* idx,
                                 float* true_score, unsigned long long* n_greater, unsigned long long* n_equal,
                                 int accumulate, void* ws, size_t ws_bytes, cudaStream_t st) {
    using namespace mma;
    cudaError_t e;
    // the sorted copy of the filter lists lives in the workspace: its length (indptr[Q]) is read back here
    int64_t nfilter = 0;
    if (indptr != nullptr && idx != nullptr) {
        if ((e = cudaMemcpyAsync(&nfilter, indptr + Q, sizeof(int64_t), cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
        if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return e;
    }
    const Layout L = layout(P, Q, nfilter);
    if (ws_bytes < L.total) return cudaErrorInvalidValue;
    char* w = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(ws) + 1023) & ~(uintptr_t)1023);
    float* Qs = (float*)(w + L.qs);
    float *Ahi = (float*)(w + L.ahi), *Alo = (float*)(w + L.alo), *Thi = (float*)(w + L.thi), *Tlo = (float*)(w + L.tlo);
    float *Bhi = (float*)(w + L.bhi), *Blo = (float*)(w + L.blo);
    int* sidx = (int*)(w + L.sidx);
    const int K = P.De, Kpad = kpad_of(K);
    const bool head = (mode == KGE_MODE_HEAD_BATCH);
    if (!accumulate) {
        if ((e = cudaMemsetAsync(n_greater, 0, (size_t)Q * sizeof(unsigned long long), st)) != cudaSuccess) return e;
        if ((e = cudaMemsetAsync(n_equal, 0, (size_t)Q * sizeof(unsigned long long), st)) != cudaSuccess) return e;
    }
    static bool attr_set = false;
    if (!attr_set) {
        if ((e = cudaFuncSetAttribute(eval_mma_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES)) != cudaSuccess) return e;
        if ((e = cudaFuncSetAttribute(eval_mma_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES)) != cudaSuccess) return e;
        if ((e = cudaFuncSetAttribute(eval_mma_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM2_BYTES)) != cudaSuccess) return e;
        attr_set = true;
    }
    const int nsm = sm_count();
    auto blocks_for = [&](long long work4) { return (unsigned)std::min<long long>((work4 + 255) / 256, 8LL * nsm); };

    // 1  q = h o r / r o t (the candidate-independent factor, same kernel as the SIMT path), split into hi / lo
    if ((e = launch_eval_stage(P, head, ent, rel, queries, qrows, Q, Qs, st)) != cudaSuccess) return e;
    split_rows_kernel<<<blocks_for(Q * (Kpad / 4)), 256, 0, st>>>(Qs, K, 0, 1, nullptr, -1, Q, K, Kpad, Ahi, Alo);
    // 2  rows of the true entities, gathered (from the table, or from qrows [Q,2,De] when it is row-sharded) and split
    if (qrows != nullptr)
        split_rows_kernel<<<blocks_for(Q * (Kpad / 4)), 256, 0, st>>>(qrows, P.De, head ? 0 : 1, 2, nullptr, -1, Q, K, Kpad, Thi, Tlo);
    else
        split_rows_kernel<<<blocks_for(Q * (Kpad / 4)), 256, 0, st>>>(ent, P.De, 0, 1, queries, head ? 0 : 2, Q, K, Kpad, Thi, Tlo);
    // 3  sorted filter lists
    const bool filtered = indptr != nullptr && idx != nullptr && nfilter > 0;
    if (filtered) filter_sort_kernel<<<(unsigned)((Q * 32 + 255) / 256), 256, 0, st>>>(indptr, idx, Q, sidx);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;

    CUtensorMap tmAhi, tmAlo, tmThi, tmTlo, tmBhi, tmBlo;
    if (!make_tmap(&tmAhi, Ahi, Q, Kpad, BM) || !make_tmap(&tmAlo, Alo, Q, Kpad, BM) || !make_tmap(&tmThi, Thi, Q, Kpad, BN) ||
        !make_tmap(&tmTlo, Tlo, Q, Kpad, BN))
        return cudaErrorNotSupported;

    TileArgs a{};
    a.Q = Q;
    a.nkb = Kpad / BK;
    a.nqt = (int)((Q + BM - 1) / BM);
    a.true_col = head ? 0 : 2;
    a.queries = queries;
    a.indptr = filtered ? indptr : nullptr;
    a.sidx = sidx;
    a.true_score = true_score;
    a.n_greater = n_greater;
    a.n_equal = n_equal;
    // 4  true scores: the diagonal of Qs x T^T
    a.net = 1;
    eval_mma_kernel<true, false><<<(unsigned)std::min(a.nqt, nsm), kThreads, SMEM_BYTES, st>>>(tmAhi, tmAlo, tmThi, tmTlo, a);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    // 5  candidates, one chunk of the split workspace at a time
    for (long long c0 = e_lo; c0 < e_hi; c0 += kEntityChunk) {
        const long long rows = std::min(kEntityChunk, e_hi - c0);
        split_rows_kernel<<<blocks_for(rows * (Kpad / 4)), 256, 0, st>>>(ent, P.De, c0, 1, nullptr, -1, rows, K, Kpad, Bhi, Blo);
        const bool pair = eval_pair();
        const int bbox = pair ? BN / 2 : BN;  // a CTA of a pair stages half of the entity tile
        if (!make_tmap(&tmBhi, Bhi, rows, Kpad, bbox) || !make_tmap(&tmBlo, Blo, rows, Kpad, bbox)) return cudaErrorNotSupported;
        a.b_row_id0 = c0;
        a.b_rows = rows;
        a.net = (int)((rows + BN - 1) / BN);
        const long long ntiles = (long long)a.nqt * a.net;
        if (pair) {
            // clusters of two CTAs, one work item per (entity tile, pair of query tiles)
            const long long nwork = (long long)((a.nqt + 1) / 2) * a.net;
            cudaLaunchConfig_t cfg{};
            static int pairs = 0;  // CTA pairs the device holds at once
            if (pairs == 0) pairs = max_active_pairs(eval_mma_kernel<false, true>, kThreads, SMEM2_BYTES);
            cfg.gridDim = dim3(2u * (unsigned)std::max<long long>(1, std::min<long long>(nwork, pairs)));
            cfg.blockDim = dim3(kThreads);
            cfg.dynamicSmemBytes = SMEM2_BYTES;
            cfg.stream = st;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = 2;
            attr[0].val.clusterDim.y = 1;
            attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr;
            cfg.numAttrs = 1;
            if ((e = cudaLaunchKernelEx(&cfg, eval_mma_kernel<false, true>, tmAhi, tmAlo, tmBhi, tmBlo, a)) != cudaSuccess) return e;
        } else {
            eval_mma_kernel<false, false><<<(unsigned)std::min<long long>(ntiles, nsm), kThreads, SMEM_BYTES, st>>>(tmAhi, tmAlo, tmBhi, tmBlo, a);
        }
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
    }
    return cudaSuccess;
}

}  // namespace kge
