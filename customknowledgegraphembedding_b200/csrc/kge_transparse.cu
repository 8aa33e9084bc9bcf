// TranSparse (/root/reference/tensorflow_codes/model.py:96-106 variables, :139-142,161-164,187-190
// gathers, :226-235 score): every relation r owns a projection M_r = mask_r o W_r [D, D]; a triple scores
//
//     gamma - sum_j | ph_j * rh_j - pt_j |,     ph = normalize(head @ M_r), pt = normalize(X @ M_r), rh = normalize(r)
//
// with X = head AGAIN in the reference as written (:229 -- the tail entity is never read; project_tail = 0 reproduces
// it, and then tail-batch scores are [B, 1]) or X = tail (project_tail = 1, what :229 evidently meant).
//
// The projections are the work: (rows x D) @ (D x D) per relation, B * N * D^2 multiply-adds per pass.  They run on the
// 5th-gen tensor cores as 3xTF32 tiles (kge_umma.cuh: tcgen05.mma kind::tf32, TMEM accumulators, TMA-fed 128-byte
// swizzled stages), the same pipeline as K5-T with a store epilogue:
//
//   row space     the positives are sorted by relation (stable); sorted positive sb owns rpp consecutive rows:
//                 its N candidate rows (when the mode scores candidates), its head row, its tail row (project_tail).
//                 Rows of one relation are contiguous, so each projection is ONE grouped GEMM over a device-built
//                 tile list (relation, row tile, column tile).
//   forward       plan -> split M_r^T and the gathered rows into hi / lo -> GEMM 1: P = X @ M_r -> score kernel
//                 (one CTA per positive: row norms, normalised relation, the L1 sums).
//   backward      dscore kernel (dL/dP per row, in place over P; per-positive relation gradients) -> split dP ->
//                 GEMM 2: dX = dP @ M_r^T -> rows scattered into the entity gradient;
//                 GEMM 3: dW_r = mask_r o (X_r^T @ dP_r), K = the rows of the relation, read from transposed splits
//                 whose per-relation segments are padded to the 32-row K stage with zeros.
//
// No CPU path; fp32 in and out; the three GEMMs keep fp32-level accuracy through the operand split.
#include <climits>
#include <cstdlib>

#include "kge_umma.cuh"

namespace kge {
namespace ts {
using namespace mma;

constexpr int kGemmThreads = 192;  // warp 0 TMA producer, warp 1 MMA issuer + TMEM owner, warps 2..5 epilogue
constexpr int kPlanThreads = 1024;

struct Tile {
    int a_row, b_row, k0, nkb;       // TMA coordinates: A rows from a_row, B rows from b_row, K from k0 for nkb stages
    int c_row, c_col, rows, cols;    // where the accumulator goes and how much of it is real
};

struct Dims {
    int B, N, R, D, Kpad;
    int rpp;        // rows per positive in the row space: N + 1 + tail
    int tail;       // the tail row is projected (project_tail)
    int cand_side;  // 0: no candidate rows, 1: candidates are heads (head-batch), 2: candidates are tails (tail-batch)
    int neg_copy;   // tail-batch without project_tail: the one negative score per positive equals the positive score
    int nout;       // negative scores per positive
    int nct, nkt;   // ceil(D / 256), ceil(D / 128)
    long long rows, rows_pad;
    int max_tiles1, max_tiles3;
};

static inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

static bool make_dims(const kge_transparse_desc& d, int mode, int64_t B, int64_t N, Dims* o) {
    Dims x{};
    x.B = (int)B;
    x.R = (int)d.nrelation;
    x.D = (int)d.dim;
    x.Kpad = (x.D + BK - 1) / BK * BK;
    x.tail = d.project_tail ? 1 : 0;
    if (mode == KGE_MODE_HEAD_BATCH) {
        x.N = (int)N, x.cand_side = 1, x.nout = (int)N;
    } else if (mode == KGE_MODE_TAIL_BATCH && x.tail) {
        x.N = (int)N, x.cand_side = 2, x.nout = (int)N;
    } else if (mode == KGE_MODE_TAIL_BATCH) {
        x.N = 0, x.cand_side = 0, x.neg_copy = 1, x.nout = 1;
    } else {
        x.N = 0, x.cand_side = 0, x.nout = 0;
    }
    x.rpp = x.N + 1 + x.tail;
    x.rows = (long long)x.B * x.rpp;
    const long long nseg = std::min<long long>(x.R, x.B);
    x.rows_pad = ((x.rows + 31) / 32) * 32 + 32 * nseg + 32;
    x.nct = ceil_div(x.D, BN);
    x.nkt = ceil_div(x.D, BM);
    // tile lists are padded to PAIRS of tiles that share their B tile (the CTA pair of gemm3x_pair_kernel)
    const long long t1 = ((x.rows + BM - 1) / BM + 2 * nseg) * x.nct, t3 = nseg * (x.nkt + 1) * x.nct;
    if (x.rows_pad >= INT_MAX || (long long)x.R * x.D >= INT_MAX || t1 >= INT_MAX || t3 >= INT_MAX) return false;
    x.max_tiles1 = (int)t1;
    x.max_tiles3 = (int)t3;
    *o = x;
    return true;
}

// ---- workspace -------------------------------------------------------------------------------------------------
struct Ws {
    int *cnt, *off, *padoff, *rtoff, *g3off, *cursor, *perm, *row_ent, *ntiles;
    Tile *tiles1, *tiles3;
    float *invn, *rinv;
    float *Mthi, *Mtlo, *Xhi, *Xlo, *P;
    // backward only
    float *Mhi, *Mlo, *XThi, *XTlo, *dPhi, *dPlo, *dPThi, *dPTlo, *dX, *GR;
    size_t total;
};

static Ws carve(const Dims& d, void* base, bool backward) {
    Ws w{};
    size_t o = 0;
    char* p = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(base) + 1023) & ~(uintptr_t)1023);
    auto take = [&](size_t bytes) {
        char* q = p + o;
        o += (bytes + 1023) & ~(size_t)1023;
        return q;
    };
    const size_t R1 = (size_t)d.R + 1;
    w.cnt = (int*)take(R1 * 4), w.off = (int*)take(R1 * 4), w.padoff = (int*)take(R1 * 4), w.rtoff = (int*)take(R1 * 4);
    w.g3off = (int*)take(R1 * 4), w.cursor = (int*)take(R1 * 4);
    w.perm = (int*)take((size_t)d.B * 4);
    w.row_ent = (int*)take((size_t)d.rows * 4);
    w.ntiles = (int*)take(64);
    w.tiles1 = (Tile*)take((size_t)d.max_tiles1 * sizeof(Tile));
    w.tiles3 = (Tile*)take((size_t)d.max_tiles3 * sizeof(Tile));
    w.invn = (float*)take((size_t)d.rows * 4);
    w.rinv = (float*)take((size_t)d.B * 4);
    const size_t mk = (size_t)d.R * d.D * d.Kpad * 4, xk = (size_t)d.rows * d.Kpad * 4;
    const size_t tk = (size_t)d.D * d.rows_pad * 4;
    w.Mthi = (float*)take(mk), w.Mtlo = (float*)take(mk);
    w.Xhi = (float*)take(xk), w.Xlo = (float*)take(xk);
    w.P = (float*)take(xk);
    if (backward) {
        w.Mhi = (float*)take(mk), w.Mlo = (float*)take(mk);
        w.XThi = (float*)take(tk), w.XTlo = (float*)take(tk);
        w.dPhi = (float*)take(xk), w.dPlo = (float*)take(xk);
        w.dPThi = (float*)take(tk), w.dPTlo = (float*)take(tk);
        w.dX = (float*)take(xk);
        w.GR = (float*)take((size_t)d.B * d.D * 4);
    }
    w.total = o + 1024;
    return w;
}

// ---- plan: relation segments, the stable order of the positives, the tile lists ---------------------------------
__device__ __forceinline__ int4 add4(int4 a, int4 b) { return make_int4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }

__device__ __forceinline__ int4 seg_values(int c, int rpp) {
    const long long rows = (long long)c * rpp;
    const int nrt = (int)((rows + BM - 1) / BM);
    return make_int4(c, (nrt + 1) & ~1, (int)(((rows + 31) / 32) * 32), c > 0 ? 1 : 0);
}

// One CTA.  off[r]: first sorted positive of relation r; rtoff[r]: first row tile; padoff[r]: first column of the
// relation's segment in the transposed (K = rows) workspaces; g3off[r]: index among the relations that occur.
__global__ void __launch_bounds__(kPlanThreads)
plan_kernel(const int64_t* __restrict__ pos, Dims d, int* cnt, int* off, int* padoff, int* rtoff, int* g3off,
            int* cursor, int* perm, int* ntiles, Tile* tiles1, Tile* tiles3) {
    __shared__ int4 s_scan[kPlanThreads];
    __shared__ int s_rel[kPlanThreads];
    const int tid = threadIdx.x;
    for (int r = tid; r <= d.R; r += kPlanThreads) cnt[r] = 0;
    __syncthreads();
    for (int b = tid; b < d.B; b += kPlanThreads) atomicAdd(cnt + (int)pos[3 * (long long)b + 1], 1);
    __syncthreads();
    // exclusive scans over the relations: thread t owns a contiguous range
    const int per = (d.R + kPlanThreads - 1) / kPlanThreads;
    const int r0 = min(d.R, tid * per), r1 = min(d.R, r0 + per);
    int4 loc = make_int4(0, 0, 0, 0);
    for (int r = r0; r < r1; ++r) loc = add4(loc, seg_values(cnt[r], d.rpp));
    s_scan[tid] = loc;
    __syncthreads();
    for (int s = 1; s < kPlanThreads; s <<= 1) {
        int4 v = s_scan[tid];
        if (tid >= s) v = add4(v, s_scan[tid - s]);
        __syncthreads();
        s_scan[tid] = v;
        __syncthreads();
    }
    int4 run = s_scan[tid];
    run = make_int4(run.x - loc.x, run.y - loc.y, run.z - loc.z, run.w - loc.w);
    for (int r = r0; r < r1; ++r) {
        off[r] = run.x, rtoff[r] = run.y, padoff[r] = run.z, g3off[r] = run.w, cursor[r] = run.x;
        run = add4(run, seg_values(cnt[r], d.rpp));
    }
    if (tid == kPlanThreads - 1) {
        const int4 tot = s_scan[tid];
        off[d.R] = tot.x, rtoff[d.R] = tot.y, padoff[d.R] = tot.z, g3off[d.R] = tot.w;
        ntiles[0] = tot.y * d.nct;
        ntiles[1] = tot.w * ((d.nkt + 1) & ~1) * d.nct;
    }
    __syncthreads();
    // stable counting sort of the positives by relation, 1024 at a time
    for (int c0 = 0; c0 < d.B; c0 += kPlanThreads) {
        const int b = c0 + tid;
        const int r = (b < d.B) ? (int)pos[3 * (long long)b + 1] : -1;
        s_rel[tid] = r;
        __syncthreads();
        int rank = 0;
        bool last = true;
        if (r >= 0) {
            for (int j = 0; j < tid; ++j) rank += (s_rel[j] == r) ? 1 : 0;
            for (int j = tid + 1; j < kPlanThreads; ++j)
                if (s_rel[j] == r) {
                    last = false;
                    break;
                }
            perm[cursor[r] + rank] = b;
        }
        __syncthreads();
        if (r >= 0 && last) cursor[r] += rank + 1;
        __syncthreads();
    }
    // tiles
    for (int r = tid; r < d.R; r += kPlanThreads) {
        const int c = cnt[r];
        if (c == 0) continue;
        const long long rows = (long long)c * d.rpp, row0 = (long long)off[r] * d.rpp;
        const int nrt = (int)((rows + BM - 1) / BM), nrt2 = (nrt + 1) & ~1;
        Tile* t1 = tiles1 + (long long)rtoff[r] * d.nct;
        for (int ct = 0; ct < d.nct; ++ct)
            for (int rt = 0; rt < nrt2; ++rt) {  // row tile fastest: tiles 2p, 2p + 1 share the B tile
                const int rr = min(rt, nrt - 1);  // (odd count: the pair's second tile repeats the first and stores nothing)
                Tile t;
                t.a_row = (int)(row0 + (long long)rr * BM);
                t.b_row = r * d.D + ct * BN;
                t.k0 = 0;
                t.nkb = d.Kpad / BK;
                t.c_row = t.a_row;
                t.c_col = ct * BN;
                t.rows = (rt < nrt) ? (int)min((long long)BM, rows - (long long)rt * BM) : 0;
                t.cols = min(BN, d.D - ct * BN);
                t1[ct * nrt2 + rt] = t;
            }
        const int nkt2 = (d.nkt + 1) & ~1;
        Tile* t3 = tiles3 + (long long)g3off[r] * nkt2 * d.nct;
        const int kpadrows = (int)(((rows + 31) / 32) * 32);
        for (int jt = 0; jt < d.nct; ++jt)
            for (int kt = 0; kt < nkt2; ++kt) {
                const int kk = min(kt, d.nkt - 1);
                Tile t;
                t.a_row = kk * BM;
                t.b_row = jt * BN;
                t.k0 = padoff[r];
                t.nkb = kpadrows / BK;
                t.c_row = r * d.D + kk * BM;
                t.c_col = jt * BN;
                t.rows = (kt < d.nkt) ? min(BM, d.D - kt * BM) : 0;
                t.cols = min(BN, d.D - jt * BN);
                t3[jt * nkt2 + kt] = t;
            }
    }
}

// ---- operand splits ----------------------------------------------------------------------------------------------
// M_r = mask_r o W_r -> hi / lo of M_r (rows k, K = j: the B operand of dX = dP @ M^T) and of M_r^T (rows j, K = k: the
// B operand of P = X @ M), both [R * D, Kpad] with zeros in the K padding.  Relations absent from the batch are skipped.
__global__ void __launch_bounds__(256)
weight_split_kernel(const float* __restrict__ W, const float* __restrict__ mask, const int* __restrict__ cnt, int D, int Kpad,
                    float* __restrict__ Mhi, float* __restrict__ Mlo, float* __restrict__ Mthi, float* __restrict__ Mtlo) {
    __shared__ float s_hi[32][33], s_lo[32][33];
    const int r = blockIdx.z;
    if (cnt[r] == 0) return;
    const int k0 = blockIdx.y * 32, j0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
    const float* Wr = W + (long long)r * D * D;
    const float* Kr = mask + (long long)r * D * D;
    for (int i = ty; i < 32; i += 8) {
        const int k = k0 + i, j = j0 + tx;
        float m = 0.f;
        if (k < D && j < D) m = __ldg(Wr + (long long)k * D + j) * __ldg(Kr + (long long)k * D + j);
        float h, l;
        split_tf32(m, h, l);
        s_hi[i][tx] = h, s_lo[i][tx] = l;
        if (Mhi != nullptr && k < D) {
            const long long o = ((long long)r * D + k) * Kpad + j;
            Mhi[o] = h, Mlo[o] = l;
        }
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
        const int j = j0 + i, k = k0 + tx;
        if (j < D) {
            const long long o = ((long long)r * D + j) * Kpad + k;
            Mthi[o] = s_hi[tx][i], Mtlo[o] = s_lo[tx][i];
        }
    }
}

// Rows -> hi / lo [rows, Kpad] (zeros in the K padding) and, when T* is given, their transposes [D, rows_pad] with the
// row index moved into its relation's padded segment.  GATHER: row (sb, local) is an entity row (candidate, head or
// tail of positive perm[sb]) and row_ent remembers which; otherwise row i of `src` (pitch floats apart).
template <bool GATHER>
__global__ void __launch_bounds__(256)
rows_split_kernel(const float* __restrict__ src, long long pitch, const int64_t* __restrict__ pos,
                  const int64_t* __restrict__ neg, const int* __restrict__ perm, const int* __restrict__ off,
                  const int* __restrict__ padoff, Dims d, int* __restrict__ row_ent, float* __restrict__ hi,
                  float* __restrict__ lo, float* __restrict__ Thi, float* __restrict__ Tlo) {
    __shared__ float s_hi[32][33], s_lo[32][33];
    __shared__ long long s_src[32];
    __shared__ int s_prow[32];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const long long row0 = (long long)blockIdx.x * 32;
    const int k0 = blockIdx.y * 32;
    if (threadIdx.x < 32) {
        const long long row = row0 + threadIdx.x;
        long long srow = -1;
        int prow = -1;
        if (row < d.rows) {
            const int sb = (int)(row / d.rpp), local = (int)(row - (long long)sb * d.rpp);
            const long long b = perm[sb];
            const int r = (int)pos[3 * b + 1];
            prow = (int)(row + ((long long)padoff[r] - (long long)off[r] * d.rpp));
            if (GATHER) {
                if (local < d.N) srow = neg[b * d.N + local];
                else if (local == d.N) srow = pos[3 * b + 0];
                else srow = pos[3 * b + 2];
                if (blockIdx.y == 0) row_ent[row] = (int)srow;
            } else {
                srow = row;
            }
        }
        s_src[threadIdx.x] = srow;
        s_prow[threadIdx.x] = prow;
    }
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
        const long long row = row0 + i;
        if (row >= d.rows) break;
        const int k = k0 + tx;
        const float x = (k < d.D) ? __ldg(src + s_src[i] * pitch + k) : 0.f;
        float h, l;
        split_tf32(x, h, l);
        hi[row * d.Kpad + k] = h, lo[row * d.Kpad + k] = l;
        s_hi[i][tx] = h, s_lo[i][tx] = l;
    }
    if (Thi == nullptr) return;
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
        const int k = k0 + i;
        if (k < d.D && row0 + tx < d.rows) {
            const long long o = (long long)k * d.rows_pad + s_prow[tx];
            Thi[o] = s_hi[tx][i], Tlo[o] = s_lo[tx][i];
        }
    }
}

// Zeros in the K padding of the transposed workspaces: columns [padoff[r] + rows_r, padoff[r + 1]) of every row k.
// (The GEMM over K = rows reads whole 32-column stages; everything else in a segment is written by rows_split_kernel.)
__global__ void __launch_bounds__(256)
pad_zero_kernel(const int* __restrict__ cnt, const int* __restrict__ padoff, Dims d, float* __restrict__ Thi,
                float* __restrict__ Tlo) {
    const int r = blockIdx.x;
    const int c0 = padoff[r] + cnt[r] * d.rpp, c1 = padoff[r + 1];
    if (c0 >= c1) return;
    const int n = c1 - c0;  // < 32
    for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < d.D * n; i += gridDim.y * blockDim.x) {
        const int k = i / n, c = c0 + (i - k * n);
        Thi[(long long)k * d.rows_pad + c] = 0.f;
        Tlo[(long long)k * d.rows_pad + c] = 0.f;
    }
}

// ---- the grouped GEMM ----------------------------------------------------------------------------------------------
// Persistent CTAs over a device-built tile list.  Per tile: C[c_row .. +rows, c_col .. +cols] = (cmask o) A . B^T
// (+ C when beta), A rows [a_row, +128) and B rows [b_row, +256) of the K-major split workspaces, K columns
// [k0, k0 + 32 * nkb).  Rows / columns of the tile past `rows` / `cols` belong to a neighbouring relation (or are
// out-of-bounds zeros) and are computed but not stored.
struct GemmArgs {
    const Tile* tiles;
    const int* ntiles;
    float* C;
    long long pitch;
    const float* cmask;  // same indexing as C, or NULL
    int beta;
};

__global__ void __launch_bounds__(kGemmThreads, 1)
gemm3x_kernel(const __grid_constant__ CUtensorMap tmAhi, const __grid_constant__ CUtensorMap tmAlo,
              const __grid_constant__ CUtensorMap tmBhi, const __grid_constant__ CUtensorMap tmBlo, const GemmArgs a) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B atoms are 1024-byte aligned
    const uint32_t bars = base + STAGES * STAGE_BYTES;
    auto full = [&](int s) { return bars + 8u * s; };
    auto empty = [&](int s) { return bars + 16u + 8u * s; };
    auto tfull = [&](int s) { return bars + 32u + 8u * s; };
    auto tempty = [&](int s) { return bars + 48u + 8u * s; };
    const uint32_t tmem_slot = bars + 64u;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmAhi);
        tma_prefetch_desc(&tmAlo);
        tma_prefetch_desc(&tmBhi);
        tma_prefetch_desc(&tmBlo);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(full(s), 1);
            mbar_init(empty(s), 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(tfull(s), 1);
            mbar_init(tempty(s), 4);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

    const int ntiles = *a.ntiles;

    if (warp == 0) {
        if (lane == 0) {
            // ===== TMA producer =====
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                const Tile t = a.tiles[tile];
                if (t.rows == 0) continue;  // the filler of an odd pair (gemm3x_pair_kernel needs it)
                for (int kb = 0; kb < t.nkb; ++kb, ++it) {
                    const int s = it % STAGES;
                    mbar_wait(empty(s), ((it / STAGES) & 1u) ^ 1u);
                    mbar_arrive_expect_tx(full(s), STAGE_BYTES);
                    const uint32_t st = base + s * STAGE_BYTES;
                    const int kc = t.k0 + kb * BK;
                    tma_load_2d(st, &tmAhi, full(s), kc, t.a_row);
                    tma_load_2d(st + A_BYTES, &tmAlo, full(s), kc, t.a_row);
                    tma_load_2d(st + 2 * A_BYTES, &tmBhi, full(s), kc, t.b_row);
                    tma_load_2d(st + 2 * A_BYTES + B_BYTES, &tmBlo, full(s), kc, t.b_row);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ===== MMA issuer =====
            uint32_t it = 0, tcount = 0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                if (a.tiles[tile].rows == 0) continue;
                const int nkb = a.tiles[tile].nkb;
                const int acc = tcount & 1u;
                mbar_wait(tempty(acc), ((tcount >> 1) & 1u) ^ 1u);
                tc_fence_after();
                const uint32_t d = tmem_base + acc * BN;
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % STAGES;
                    mbar_wait(full(s), (it / STAGES) & 1u);
                    tc_fence_after();
                    const uint32_t st = base + s * STAGE_BYTES;
                    const uint64_t ahi = umma_desc_sw128(st), alo = umma_desc_sw128(st + A_BYTES);
                    const uint64_t bhi = umma_desc_sw128(st + 2 * A_BYTES), blo = umma_desc_sw128(st + 2 * A_BYTES + B_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) umma_tf32(d, alo + 2 * k, bhi + 2 * k, (kb | k) != 0);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) umma_tf32(d, ahi + 2 * k, blo + 2 * k, 1u);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) umma_tf32(d, ahi + 2 * k, bhi + 2 * k, 1u);
                    umma_commit(empty(s));
                }
                umma_commit(tfull(acc));
                ++tcount;
            }
        }
    } else {
        // ===== epilogue: warps 2..5, TMEM lane quadrant = warp % 4, thread = accumulator row =====
        const int wq = warp & 3;
        const int row = wq * 32 + lane;
        const bool vec = (a.pitch % 4 == 0) && ((reinterpret_cast<uintptr_t>(a.C) & 15) == 0) &&
                         (a.cmask == nullptr || (reinterpret_cast<uintptr_t>(a.cmask) & 15) == 0);
        uint32_t tcount = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            const Tile t = a.tiles[tile];
            if (t.rows == 0) continue;
            const int acc = tcount & 1u;
            const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16) + acc * BN;
            const long long o0 = (long long)(t.c_row + row) * a.pitch + t.c_col;
            const bool rok = row < t.rows;
            mbar_wait(tfull(acc), (tcount >> 1) & 1u);
            tc_fence_after();
#pragma unroll 1
            for (int c = 0; c * 32 < t.cols; ++c) {
                float v[32];
                __syncwarp();
                tmem_ld32(taddr + c * 32, v);
                if (!rok) continue;
                const int ncol = min(32, t.cols - c * 32);
                float* crow = a.C + o0 + c * 32;
                const float* mrow = a.cmask ? a.cmask + o0 + c * 32 : nullptr;
                if (vec && ncol == 32) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        float4 x = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                        if (mrow != nullptr) {
                            const float4 m = __ldg(reinterpret_cast<const float4*>(mrow + j));
                            x.x *= m.x, x.y *= m.y, x.z *= m.z, x.w *= m.w;
                        }
                        if (a.beta) {
                            const float4 y = *reinterpret_cast<const float4*>(crow + j);
                            x.x += y.x, x.y += y.y, x.z += y.z, x.w += y.w;
                        }
                        *reinterpret_cast<float4*>(crow + j) = x;
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        if (j < ncol) {
                            float x = v[j];
                            if (mrow != nullptr) x *= __ldg(mrow + j);
                            if (a.beta) x += crow[j];
                            crow[j] = x;
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty(acc));
            ++tcount;
        }
    }
    __syncwarp();
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// The same GEMM on CTA PAIRS (cta_group::2): the two tiles of a pair (same B tile, two A tiles) are one 256 x 256 tile of the
// pair.  Each CTA stages its own A tile and half of the B tile -- 64 KB per k-stage and SM instead of 96 KB, which is what
// bounds the one-CTA kernel (ncu: 7.2 GB from L2 per GEMM at 10.7 TB/s, tensor pipe 52 % active) -- in a 3-stage ring.
// Both producers signal the LEADER's `full` barrier (the leader issues every MMA); the leader's commits reach both CTAs'
// `empty` / `tfull` barriers; both CTAs' epilogue warps release the accumulator on the leader's `tempty`.

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm3x_pair_kernel(const __grid_constant__ CUtensorMap tmAhi, const __grid_constant__ CUtensorMap tmAlo,
                   const __grid_constant__ CUtensorMap tmBhi, const __grid_constant__ CUtensorMap tmBlo, const GemmArgs a) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bars = base + STAGES2 * STAGE2_BYTES;
    auto full = [&](int s) { return bars + 8u * s; };
    auto empty = [&](int s) { return bars + 24u + 8u * s; };
    auto tfull = [&](int s) { return bars + 48u + 8u * s; };
    auto tempty = [&](int s) { return bars + 64u + 8u * s; };
    const uint32_t tmem_slot = bars + 80u;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t crank = cluster_ctarank();
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmAhi);
        tma_prefetch_desc(&tmAlo);
        tma_prefetch_desc(&tmBhi);
        tma_prefetch_desc(&tmBlo);
        for (int s = 0; s < STAGES2; ++s) {
            mbar_init(full(s), 1);   // the leader's arrive.expect_tx (bytes of both CTAs)
            mbar_init(empty(s), 1);  // the leader's commit, delivered to both CTAs
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(tfull(s), 1);   // the leader's commit, delivered to both CTAs
            mbar_init(tempty(s), 8);  // one lane of each epilogue warp of both CTAs (used on the leader only)
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc_pair(tmem_slot, TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    cluster_sync_all();  // the peer's barriers and TMEM exist before anything is sent to them
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

    const int ntiles = *a.ntiles;  // even: pairs
    const int first = 2 * (int)cluster_id_x(), stride = 2 * (int)cluster_count_x();

    if (warp == 0) {
        if (lane == 0) {
            // ===== TMA producer (both CTAs) =====
            uint32_t it = 0;
            for (int tile0 = first; tile0 < ntiles; tile0 += stride) {
                const Tile t = a.tiles[tile0 + crank];
                for (int kb = 0; kb < t.nkb; ++kb, ++it) {
                    const int s = it % STAGES2;
                    mbar_wait(empty(s), ((it / STAGES2) & 1u) ^ 1u);
                    if (crank == 0) mbar_arrive_expect_tx(full(s), 2 * STAGE2_BYTES);
                    const uint32_t lf = map_to_rank(full(s), 0);
                    const uint32_t st = base + s * STAGE2_BYTES;
                    const int kc = t.k0 + kb * BK;
                    tma_load_2d_pair(st, &tmAhi, lf, kc, t.a_row);
                    tma_load_2d_pair(st + A_BYTES, &tmAlo, lf, kc, t.a_row);
                    tma_load_2d_pair(st + 2 * A_BYTES, &tmBhi, lf, kc, t.b_row + (int)crank * (BN / 2));
                    tma_load_2d_pair(st + 2 * A_BYTES + BH_BYTES, &tmBlo, lf, kc, t.b_row + (int)crank * (BN / 2));
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && crank == 0) {
            // ===== MMA issuer (leader only) =====
            uint32_t it = 0, tcount = 0;
            for (int tile0 = first; tile0 < ntiles; tile0 += stride, ++tcount) {
                const int nkb = a.tiles[tile0].nkb;
                const int acc = tcount & 1u;
                mbar_wait(tempty(acc), ((tcount >> 1) & 1u) ^ 1u);
                tc_fence_after();
                const uint32_t d = tmem_base + acc * BN;
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % STAGES2;
                    mbar_wait(full(s), (it / STAGES2) & 1u);
                    tc_fence_after();
                    const uint32_t st = base + s * STAGE2_BYTES;
                    const uint64_t ahi = umma_desc_sw128(st), alo = umma_desc_sw128(st + A_BYTES);
                    const uint64_t bhi = umma_desc_sw128(st + 2 * A_BYTES), blo = umma_desc_sw128(st + 2 * A_BYTES + BH_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) umma_tf32_pair(d, alo + 2 * k, bhi + 2 * k, (kb | k) != 0);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) umma_tf32_pair(d, ahi + 2 * k, blo + 2 * k, 1u);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) umma_tf32_pair(d, ahi + 2 * k, bhi + 2 * k, 1u);
                    umma_commit_pair(empty(s));
                }
                umma_commit_pair(tfull(acc));
            }
        }
    } else {
        // ===== epilogue (both CTAs): this CTA's 128 accumulator rows = its own tile of the pair =====
        const int wq = warp & 3;
        const int row = wq * 32 + lane;
        const bool vec = (a.pitch % 4 == 0) && ((reinterpret_cast<uintptr_t>(a.C) & 15) == 0) &&
                         (a.cmask == nullptr || (reinterpret_cast<uintptr_t>(a.cmask) & 15) == 0);
        uint32_t tcount = 0;
        for (int tile0 = first; tile0 < ntiles; tile0 += stride, ++tcount) {
            const Tile t = a.tiles[tile0 + crank];
            const int acc = tcount & 1u;
            const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16) + acc * BN;
            const long long o0 = (long long)(t.c_row + row) * a.pitch + t.c_col;
            const bool rok = row < t.rows;
            mbar_wait(tfull(acc), (tcount >> 1) & 1u);
            tc_fence_after();
            if (t.rows > 0) {
#pragma unroll 1
                for (int c = 0; c * 32 < t.cols; ++c) {
                    float v[32];
                    __syncwarp();
                    tmem_ld32(taddr + c * 32, v);
                    if (!rok) continue;
                    const int ncol = min(32, t.cols - c * 32);
                    float* crow = a.C + o0 + c * 32;
                    const float* mrow = a.cmask ? a.cmask + o0 + c * 32 : nullptr;
                    if (vec && ncol == 32) {
#pragma unroll
                        for (int j = 0; j < 32; j += 4) {
                            float4 x = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                            if (mrow != nullptr) {
                                const float4 m = __ldg(reinterpret_cast<const float4*>(mrow + j));
                                x.x *= m.x, x.y *= m.y, x.z *= m.z, x.w *= m.w;
                            }
                            if (a.beta) {
                                const float4 y = *reinterpret_cast<const float4*>(crow + j);
                                x.x += y.x, x.y += y.y, x.z += y.z, x.w += y.w;
                            }
                            *reinterpret_cast<float4*>(crow + j) = x;
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            if (j < ncol) {
                                float x = v[j];
                                if (mrow != nullptr) x *= __ldg(mrow + j);
                                if (a.beta) x += crow[j];
                                crow[j] = x;
                            }
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(map_to_rank(tempty(acc), 0));
        }
    }
    __syncwarp();
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();  // no CTA leaves (or frees its TMEM) while its peer may still use it
    if (warp == 1) {
        __syncwarp();
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, TMEM_COLS);
    }
}

// ---- score / dscore: one CTA per (sorted) positive ---------------------------------------------------------------
// (warp_sum / block_sum: kge_common.cuh -- fixed combination order)
__device__ __forceinline__ float sgn(float x) { return (x > 0.f) ? 1.f : ((x < 0.f) ? -1.f : 0.f); }

struct ScoreArgs {
    const float* rel;
    const int64_t* pos;
    const int* perm;
    float* P;            // [rows, Kpad]: the projections; dscore overwrites each row with dL/dP
    float* invn;         // [rows]  1 / ||P_row||
    float* rinv;         // [B]     1 / ||r|| per sorted positive
    float* pos_score;    // [B]        (original order)
    float* neg_score;    // [B, nout]
    const float* dpos;   // upstream gradients (dscore)
    const float* dneg;
    float* GR;           // [B, D] per-positive relation gradient, sorted order
    float gamma;
};

__global__ void __launch_bounds__(256) score_kernel(Dims d, ScoreArgs a) {
    extern __shared__ float sm[];
    __shared__ float red[33];
    float *rhat = sm, *sh_h = sm + d.D, *sh_t = sm + 2 * d.D;
    const int sb = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nw = blockDim.x >> 5;
    const long long b = a.perm[sb];
    const int r = (int)a.pos[3 * b + 1];
    const long long row0 = (long long)sb * d.rpp;
    float ss = 0.f;
    for (int j = tid; j < d.D; j += blockDim.x) {
        const float x = __ldg(a.rel + (long long)r * d.D + j);
        rhat[j] = x;
        ss += x * x;
    }
    ss = block_sum(ss, red);
    const float ri = 1.f / sqrtf(ss);
    for (int j = tid; j < d.D; j += blockDim.x) rhat[j] *= ri;
    if (tid == 0) a.rinv[sb] = ri;
    for (int lr = warp; lr < d.rpp; lr += nw) {
        const float* p = a.P + (row0 + lr) * d.Kpad;
        float s = 0.f;
        for (int j = lane; j < d.D; j += 32) s += p[j] * p[j];
        s = warp_sum(s);
        if (lane == 0) a.invn[row0 + lr] = 1.f / sqrtf(s);
    }
    __syncthreads();
    const float ih = a.invn[row0 + d.N], it = d.tail ? a.invn[row0 + d.N + 1] : 0.f;
    for (int j = tid; j < d.D; j += blockDim.x) {
        const float h = a.P[(row0 + d.N) * d.Kpad + j] * ih;
        sh_h[j] = h;
        sh_t[j] = d.tail ? a.P[(row0 + d.N + 1) * d.Kpad + j] * it : h;
    }
    __syncthreads();
    float s = 0.f;
    for (int j = tid; j < d.D; j += blockDim.x) s += fabsf(sh_h[j] * rhat[j] - sh_t[j]);
    s = block_sum(s, red);
    if (tid == 0) {
        a.pos_score[b] = a.gamma - s;
        if (d.neg_copy) a.neg_score[b] = a.gamma - s;
    }
    if (d.cand_side == 0) return;
    for (int n = warp; n < d.N; n += nw) {
        const float* p = a.P + (row0 + n) * d.Kpad;
        const float inv = a.invn[row0 + n];
        float t = 0.f;
        for (int j = lane; j < d.D; j += 32) {
            const float y = p[j] * inv;
            const float ph = (d.cand_side == 1) ? y : sh_h[j];
            const float pt = (d.cand_side == 1) ? (d.tail ? sh_t[j] : y) : y;
            t += fabsf(ph * rhat[j] - pt);
        }
        t = warp_sum(t);
        if (lane == 0) a.neg_score[b * d.N + n] = a.gamma - t;
    }
}

// dL/dP for every row of the positive (in place over P) and its relation gradient.  With y = P / ||P|| and A = dL/dy
// summed over the uses of the row:  dL/dP = (A - y (y . A)) / ||P||.
__global__ void __launch_bounds__(256) dscore_kernel(Dims d, ScoreArgs a) {
    extern __shared__ float sm[];
    __shared__ float red[33];
    const int sb = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, nw = blockDim.x >> 5;
    float *rhat = sm, *sh_h = sm + d.D, *sh_t = sm + 2 * d.D;
    float *accH = sm + 3 * d.D, *accT = accH + (size_t)nw * d.D, *accR = accT + (size_t)nw * d.D;
    const long long b = a.perm[sb];
    const int r = (int)a.pos[3 * b + 1];
    const long long row0 = (long long)sb * d.rpp;
    const float ri = a.rinv[sb];
    const float ih = a.invn[row0 + d.N], it = d.tail ? a.invn[row0 + d.N + 1] : 0.f;
    for (int j = tid; j < d.D; j += blockDim.x) {
        rhat[j] = __ldg(a.rel + (long long)r * d.D + j) * ri;
        const float h = a.P[(row0 + d.N) * d.Kpad + j] * ih;
        sh_h[j] = h;
        sh_t[j] = d.tail ? a.P[(row0 + d.N + 1) * d.Kpad + j] * it : h;
    }
    for (int i = tid; i < 3 * nw * d.D; i += blockDim.x) accH[i] = 0.f;
    __syncthreads();
    float* aH = accH + (size_t)warp * d.D;
    float* aT = accT + (size_t)warp * d.D;
    float* aR = accR + (size_t)warp * d.D;
    if (warp == 0) {
        const float g = a.dpos[b] + (d.neg_copy ? a.dneg[b] : 0.f);
        for (int j = lane; j < d.D; j += 32) {
            const float s = sgn(sh_h[j] * rhat[j] - sh_t[j]);
            if (d.tail) {
                aH[j] += -g * s * rhat[j];
                aT[j] += g * s;
            } else {
                aH[j] += -g * s * (rhat[j] - 1.f);
            }
            aR[j] += -g * s * sh_h[j];
        }
    }
    if (d.cand_side != 0) {
        for (int n = warp; n < d.N; n += nw) {
            float* p = a.P + (row0 + n) * d.Kpad;
            const float inv = a.invn[row0 + n];
            const float g = a.dneg[b * d.N + n];
            float dot = 0.f;
            for (int j = lane; j < d.D; j += 32) {
                const float y = p[j] * inv;
                float aj;
                if (d.cand_side == 1) {
                    const float s = sgn(y * rhat[j] - (d.tail ? sh_t[j] : y));
                    aj = d.tail ? -g * s * rhat[j] : -g * s * (rhat[j] - 1.f);
                } else {
                    aj = g * sgn(sh_h[j] * rhat[j] - y);
                }
                dot += y * aj;
            }
            dot = warp_sum(dot);
            for (int j = lane; j < d.D; j += 32) {
                const float y = p[j] * inv;
                float aj;
                if (d.cand_side == 1) {
                    const float s = sgn(y * rhat[j] - (d.tail ? sh_t[j] : y));
                    aj = d.tail ? -g * s * rhat[j] : -g * s * (rhat[j] - 1.f);
                    if (d.tail) aT[j] += g * s;
                    aR[j] += -g * s * y;
                } else {
                    const float s = sgn(sh_h[j] * rhat[j] - y);
                    aj = g * s;
                    aH[j] += -g * s * rhat[j];
                    aR[j] += -g * s * sh_h[j];
                }
                p[j] = (aj - y * dot) * inv;
            }
        }
    }
    __syncthreads();
    // the warps' partial sums, in warp order
    for (int j = tid; j < d.D; j += blockDim.x) {
        float h = 0.f, t = 0.f, rr = 0.f;
        for (int w = 0; w < nw; ++w) {
            h += accH[(size_t)w * d.D + j];
            t += accT[(size_t)w * d.D + j];
            rr += accR[(size_t)w * d.D + j];
        }
        accH[j] = h, accT[j] = t, accR[j] = rr;
    }
    float v = 0.f;
    for (int j = tid; j < d.D; j += blockDim.x) v += sh_h[j] * accH[j];
    const float dotH = block_sum(v, red);
    for (int j = tid; j < d.D; j += blockDim.x) a.P[(row0 + d.N) * d.Kpad + j] = (accH[j] - sh_h[j] * dotH) * ih;
    if (d.tail) {
        v = 0.f;
        for (int j = tid; j < d.D; j += blockDim.x) v += sh_t[j] * accT[j];
        const float dotT = block_sum(v, red);
        for (int j = tid; j < d.D; j += blockDim.x) a.P[(row0 + d.N + 1) * d.Kpad + j] = (accT[j] - sh_t[j] * dotT) * it;
    }
    v = 0.f;
    for (int j = tid; j < d.D; j += blockDim.x) v += rhat[j] * accR[j];
    const float dotR = block_sum(v, red);
    for (int j = tid; j < d.D; j += blockDim.x) a.GR[(long long)sb * d.D + j] = (accR[j] - rhat[j] * dotR) * ri;
}

// relation gradient: the positives of a relation are contiguous in sorted order; summed in that order
__global__ void __launch_bounds__(256)
relation_grad_kernel(const float* __restrict__ GR, const int* __restrict__ off, int D, float* __restrict__ grad_rel) {
    const int r = blockIdx.x;
    const int s0 = off[r], s1 = off[r + 1];
    for (int j = threadIdx.x; j < D; j += blockDim.x) {
        float s = 0.f;
        for (int sb = s0; sb < s1; ++sb) s += GR[(long long)sb * D + j];
        grad_rel[(long long)r * D + j] = s;
    }
}

// dX rows -> entity gradient rows (float atomics: the order of the uses of one entity is not fixed); four coordinates
// per reduction (red.global.add.v4.f32) when the rows allow 16-byte accesses
__global__ void __launch_bounds__(256)
scatter_rows_kernel(const float* __restrict__ dX, int Kpad, const int* __restrict__ row_ent, long long rows, int D,
                    float* grad_ent) {
    if ((D & 3) == 0 && (reinterpret_cast<uintptr_t>(grad_ent) & 15) == 0) {
        const int d4 = D >> 2;
        const long long total = rows * d4;
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
            const long long row = i / d4;
            const int j = (int)(i - row * d4) * 4;
            const float4 v = *reinterpret_cast<const float4*>(dX + row * Kpad + j);
            float* dst = grad_ent + (long long)row_ent[row] * D + j;
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
        }
        return;
    }
    const long long total = rows * D;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long row = i / D;
        const int j = (int)(i - row * D);
        atomicAdd(grad_ent + (long long)row_ent[row] * D + j, dX[row * Kpad + j]);
    }
}

// ---- host side ---------------------------------------------------------------------------------------------------
static int g_pair = -1;  // 1 (default): gemm3x_pair_kernel on CTA pairs; 0 (KGE_TS_PAIR=0): one CTA per tile
static int gemm_pair() {
    if (g_pair < 0) {
        const char* v = getenv("KGE_TS_PAIR");
        g_pair = (v != nullptr && v[0] == '0') ? 0 : 1;
    }
    return g_pair;
}

static cudaError_t run_gemm(const float* Ahi, const float* Alo, long long a_rows, long long a_inner, const float* Bhi,
                            const float* Blo, long long b_rows, long long b_inner, const Tile* tiles, const int* ntiles,
                            int max_tiles, float* C, long long pitch, const float* cmask, int beta, cudaStream_t st) {
    static bool attr_set = false;
    cudaError_t e;
    if (!attr_set) {
        if ((e = cudaFuncSetAttribute(gemm3x_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES)) != cudaSuccess) return e;
        if ((e = cudaFuncSetAttribute(gemm3x_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM2_BYTES)) != cudaSuccess) return e;
        attr_set = true;
    }
    const bool pair = gemm_pair() != 0;
    CUtensorMap tAhi, tAlo, tBhi, tBlo;
    const int bbox = pair ? BN / 2 : BN;  // a CTA of a pair stages half of the B tile
    if (!make_tmap(&tAhi, Ahi, a_rows, (int)a_inner, BM) || !make_tmap(&tAlo, Alo, a_rows, (int)a_inner, BM) ||
        !make_tmap(&tBhi, Bhi, b_rows, (int)b_inner, bbox) || !make_tmap(&tBlo, Blo, b_rows, (int)b_inner, bbox))
        return cudaErrorNotSupported;
    GemmArgs g{tiles, ntiles, C, pitch, cmask, beta};
    if (pair) {
        static int pairs = 0;  // CTA pairs the device holds at once
        if (pairs == 0) pairs = max_active_pairs(gemm3x_pair_kernel, kGemmThreads, SMEM2_BYTES);
        const int grid = 2 * std::max(1, std::min(max_tiles / 2, pairs));
        gemm3x_pair_kernel<<<grid, kGemmThreads, SMEM2_BYTES, st>>>(tAhi, tAlo, tBhi, tBlo, g);
    } else {
        const int grid = std::max(1, std::min(max_tiles, sm_count()));
        gemm3x_kernel<<<grid, kGemmThreads, SMEM_BYTES, st>>>(tAhi, tAlo, tBhi, tBlo, g);
    }
    return cudaGetLastError();
}

static cudaError_t big_smem(const void* fn, size_t bytes) {
    if (bytes <= 48 * 1024) return cudaSuccess;
    return cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
}

static cudaError_t forward(const Dims& d, const Ws& w, const float* ent, const float* rel, const float* W, const float* mask,
                           const int64_t* pos, const int64_t* neg, float gamma, float* pos_score, float* neg_score,
                           bool backward, cudaStream_t st) {
    cudaError_t e;
    plan_kernel<<<1, kPlanThreads, 0, st>>>(pos, d, w.cnt, w.off, w.padoff, w.rtoff, w.g3off, w.cursor, w.perm, w.ntiles,
                                            w.tiles1, w.tiles3);
    const int kt = d.Kpad / 32;
    weight_split_kernel<<<dim3(kt, kt, d.R), 256, 0, st>>>(W, mask, w.cnt, d.D, d.Kpad, backward ? w.Mhi : nullptr, w.Mlo,
                                                           w.Mthi, w.Mtlo);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if (backward) pad_zero_kernel<<<dim3(d.R, 8), 256, 0, st>>>(w.cnt, w.padoff, d, w.XThi, w.XTlo);
    rows_split_kernel<true><<<dim3((unsigned)((d.rows + 31) / 32), kt), 256, 0, st>>>(
        ent, d.D, pos, neg, w.perm, w.off, w.padoff, d, w.row_ent, w.Xhi, w.Xlo, backward ? w.XThi : nullptr, w.XTlo);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    // GEMM 1: P = X @ M_r
    if ((e = run_gemm(w.Xhi, w.Xlo, d.rows, d.Kpad, w.Mthi, w.Mtlo, (long long)d.R * d.D, d.Kpad, w.tiles1, w.ntiles,
                      d.max_tiles1, w.P, d.Kpad, nullptr, 0, st)) != cudaSuccess)
        return e;
    const size_t smem = 3 * (size_t)d.D * sizeof(float);
    if (smem > 200 * 1024) return cudaErrorInvalidValue;
    if ((e = big_smem((const void*)score_kernel, smem)) != cudaSuccess) return e;
    ScoreArgs a{rel, pos, w.perm, w.P, w.invn, w.rinv, pos_score, neg_score, nullptr, nullptr, nullptr, gamma};
    score_kernel<<<d.B, 256, smem, st>>>(d, a);
    return cudaGetLastError();
}

static cudaError_t backward(const Dims& d, const Ws& w, int64_t E, const float* rel, const float* mask, const int64_t* pos,
                            const float* dpos, const float* dneg, float* g_ent, float* g_rel, float* g_W, cudaStream_t st) {
    cudaError_t e;
    int nw = 8;
    auto smem_for = [&](int n) { return (size_t)(3 + 3 * n) * d.D * sizeof(float); };
    while (nw > 1 && smem_for(nw) > 200 * 1024) nw >>= 1;
    if (smem_for(nw) > 200 * 1024) return cudaErrorInvalidValue;
    if ((e = big_smem((const void*)dscore_kernel, smem_for(nw))) != cudaSuccess) return e;
    ScoreArgs a{rel, pos, w.perm, w.P, w.invn, w.rinv, nullptr, nullptr, dpos, dneg, w.GR, 0.f};
    dscore_kernel<<<d.B, 32 * nw, smem_for(nw), st>>>(d, a);
    relation_grad_kernel<<<d.R, 256, 0, st>>>(w.GR, w.off, d.D, g_rel);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    pad_zero_kernel<<<dim3(d.R, 8), 256, 0, st>>>(w.cnt, w.padoff, d, w.dPThi, w.dPTlo);
    const int kt = d.Kpad / 32;
    rows_split_kernel<false><<<dim3((unsigned)((d.rows + 31) / 32), kt), 256, 0, st>>>(
        w.P, d.Kpad, pos, nullptr, w.perm, w.off, w.padoff, d, nullptr, w.dPhi, w.dPlo, w.dPThi, w.dPTlo);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    // GEMM 2: dX = dP @ M_r^T
    if ((e = run_gemm(w.dPhi, w.dPlo, d.rows, d.Kpad, w.Mhi, w.Mlo, (long long)d.R * d.D, d.Kpad, w.tiles1, w.ntiles,
                      d.max_tiles1, w.dX, d.Kpad, nullptr, 0, st)) != cudaSuccess)
        return e;
    if ((e = cudaMemsetAsync(g_ent, 0, (size_t)E * d.D * sizeof(float), st)) != cudaSuccess) return e;
    const long long total = d.rows * d.D;
    scatter_rows_kernel<<<(unsigned)std::min<long long>((total + 255) / 256, 32LL * sm_count()), 256, 0, st>>>(
        w.dX, d.Kpad, w.row_ent, d.rows, d.D, g_ent);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    // GEMM 3: dW_r = mask_r o (X_r^T @ dP_r); relations absent from the batch keep a zero gradient
    if ((e = cudaMemsetAsync(g_W, 0, (size_t)d.R * d.D * d.D * sizeof(float), st)) != cudaSuccess) return e;
    return run_gemm(w.XThi, w.XTlo, d.D, d.rows_pad, w.dPThi, w.dPTlo, d.D, d.rows_pad, w.tiles3, w.ntiles + 1, d.max_tiles3,
                    g_W, d.D, mask, 0, st);
}

static int check_desc(const kge_transparse_desc* desc, int mode, int64_t B, int64_t N, Dims* d) {
    if (!desc) return fail(KGE_ERR_INVALID_ARGUMENT, "NULL desc");
    if (desc->nentity <= 0 || desc->nrelation <= 0 || desc->dim <= 0)
        return fail(KGE_ERR_INVALID_ARGUMENT, "TranSparse: nentity, nrelation and dim must be positive");
    if (desc->nrelation > 65535) return fail(KGE_ERR_UNSUPPORTED, "TranSparse: at most 65535 relations");
    if (mode != KGE_MODE_SINGLE && mode != KGE_MODE_HEAD_BATCH && mode != KGE_MODE_TAIL_BATCH)
        return fail(KGE_ERR_INVALID_ARGUMENT, "bad mode %d", mode);
    if (B <= 0 || B >= INT_MAX || (mode != KGE_MODE_SINGLE && (N <= 0 || N >= INT_MAX)))
        return fail(KGE_ERR_INVALID_ARGUMENT, "bad batch sizes");
    if (!make_dims(*desc, mode, B, N, d)) return fail(KGE_ERR_UNSUPPORTED, "TranSparse: sizes exceed the 32-bit tile coordinates");
    return KGE_OK;
}

}  // namespace ts
}  // namespace kge

using namespace kge;

extern "C" {

int64_t kge_transparse_nout(const kge_transparse_desc* desc, int mode, int64_t N) {
    if (mode == KGE_MODE_HEAD_BATCH) return N;
    if (mode == KGE_MODE_TAIL_BATCH) return (desc && desc->project_tail) ? N : 1;
    return 0;
}

size_t kge_transparse_workspace_bytes(const kge_transparse_desc* desc, int mode, int64_t B, int64_t N, int with_backward) {
    ts::Dims d;
    if (ts::check_desc(desc, mode, B, N, &d) != KGE_OK) return 0;
    return ts::carve(d, nullptr, with_backward != 0).total;
}

int kge_transparse_forward(const kge_transparse_desc* desc, int mode, const float* entity, const float* relation,
                           const float* W, const float* mask, const int64_t* positive, const int64_t* negative, int64_t B,
                           int64_t N, float* positive_score, float* negative_score, int with_backward, void* ws,
                           size_t ws_bytes, void* stream) {
    ts::Dims d;
    int rc = ts::check_desc(desc, mode, B, N, &d);
    if (rc) return rc;
    if (!entity || !relation || !W || !mask || !positive || !positive_score)
        return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    if (mode != KGE_MODE_SINGLE && !negative_score) return fail(KGE_ERR_INVALID_ARGUMENT, "negative_score is NULL");
    if (d.cand_side != 0 && !negative) return fail(KGE_ERR_INVALID_ARGUMENT, "negative is NULL");
    const ts::Ws w = ts::carve(d, ws, with_backward != 0);
    if (!ws || ws_bytes < w.total)
        return fail(KGE_ERR_WORKSPACE, "TranSparse forward needs %zu workspace bytes, got %zu", w.total, ws_bytes);
    cudaError_t e = ts::forward(d, w, entity, relation, W, mask, positive, negative, desc->gamma, positive_score,
                                negative_score, with_backward != 0, (cudaStream_t)stream);
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_transparse_forward");
}

int kge_transparse_backward(const kge_transparse_desc* desc, int mode, const float* relation, const float* mask,
                            const int64_t* positive, int64_t B, int64_t N, const float* dpositive, const float* dnegative,
                            float* grad_entity, float* grad_relation, float* grad_W, void* ws, size_t ws_bytes,
                            void* stream) {
    ts::Dims d;
    int rc = ts::check_desc(desc, mode, B, N, &d);
    if (rc) return rc;
    if (!relation || !mask || !positive || !dpositive || !grad_entity || !grad_relation || !grad_W)
        return fail(KGE_ERR_INVALID_ARGUMENT, "NULL pointer");
    if (mode != KGE_MODE_SINGLE && !dnegative) return fail(KGE_ERR_INVALID_ARGUMENT, "dnegative is NULL");
    const ts::Ws w = ts::carve(d, ws, true);
    if (!ws || ws_bytes < w.total)
        return fail(KGE_ERR_WORKSPACE, "TranSparse backward needs %zu workspace bytes, got %zu", w.total, ws_bytes);
    cudaError_t e = ts::backward(d, w, desc->nentity, relation, mask, positive, dpositive, dnegative, grad_entity,
                                 grad_relation, grad_W, (cudaStream_t)stream);
    return e == cudaSuccess ? KGE_OK : fail_cuda(e, "kge_transparse_backward");
}

}  // extern "C"
