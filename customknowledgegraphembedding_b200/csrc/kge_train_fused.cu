// Fused train step, second generation (sm_100a): no dense gradient table, no float atomics.
//
//  F1 sweep      one CTA per positive b.  Every thread owns fixed coordinates of every row, so the
//                staged planes q_b and the dL/dq accumulators live in registers.  ONE pass over the N
//                candidate rows produces the scores AND dL/dq: the self-adversarial softmax weights are
//                folded in with the online-softmax recurrence (running max m, rescale on change), so the
//                rows are never re-read.  Row sums cross threads through one shared-memory reduction per
//                group of RPI rows.  Writes: scores S[b,:], planes Q[b], row stats (m_b, scale_b), loss
//                terms, compact per-positive gradient rows GH[b], GR[b], GT[b]; counts every row use.
//  F2 buckets    integer histogram (done in F1) -> exclusive scan -> fill -> per-bucket sort: for each
//                entity the list of its uses in ascending order.
//  F2b dscore    S[b,n] <- dL/ds[b,n] from (S, m_b, scale_b); loss terms -> loss3.
//  F3 entity     one CTA per entity row (persistent grid): dL/d(row) = sum over its uses, rebuilt from
//                (Q[b], row, dL/ds) coordinate by coordinate in registers in that fixed order, plus the
//                GH/GT rows of positives it heads/tails; then Adam on the row in the same registers.
//                The (dL/ds, b) pairs of the uses are staged in shared memory, the Q rows stream through a
//                cp.async ring (3 rows in flight per CTA), RotatE math is packed (FFMA2).
//                Rows without uses take the g = 0 Adam update (dense-Adam semantics of the reference).
//  F4 relation   one CTA per relation row: sum GR[b] over b ascending, Adam; pRotatE modulus likewise.
//
// Replaces the body of KGEModel.train_step (loss restated at
// /root/reference/tensorflow_codes/model.py:4-44).  Results are bit-reproducible run to run.
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "kge_coord.cuh"
#include "kge_internal.h"
#include "kge_prof.h"
#include "kge_ptx.cuh"
#include "kge_train.h"
#include <type_traits>

namespace kge {

// all threads get sum over the block of each v[i]; red: NV*32 floats.  Fixed order.
template <int NV>
__device__ __forceinline__ void block_reduce_vals(float (&v)[NV], float* red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        v[i] = warp_sum(v[i]);
        if (lane == 0) red[i * 32 + warp] = v[i];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        float t = 0.f;
        for (int w = 0; w < nw; ++w) t += red[i * 32 + w];
        v[i] = t;
    }
    __syncthreads();
}

template <int CP, int CH>
__device__ __forceinline__ void load_coords(float (&x)[CP][CH][4], const float* __restrict__ row, int H, int tid, int T,
                                            int plane_stride) {
#pragma unroll
    for (int p = 0; p < CP; ++p)
#pragma unroll
        for (int j = 0; j < CH; ++j) {
            const int k0 = 4 * (tid + j * T);
            float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
            if (k0 < H) val = __ldg(reinterpret_cast<const float4*>(row + p * plane_stride + k0));
            x[p][j][0] = val.x; x[p][j][1] = val.y; x[p][j][2] = val.z; x[p][j][3] = val.w;
        }
}

// dL/ds of one negative score: scale_b * exp(alpha*s - m_b) * sigma(s)   (or scale_b * sigma(s) without -adv).
// Attached softmax weights (the TF graph model, /root/reference/tensorflow_codes/model.py:168-171,195-198: no
// stop-gradient on softmax(s)): d/ds_n sum_m p_m l_m = p_n (l'_n + alpha (l_n - Lbar)), l = logsigmoid(-s),
// Lbar = sum_m p_m l_m  ->  scale_b * e_n * (sigma(s_n) - alpha (l_n - Lbar)).
__device__ __forceinline__ float dscore_value(float s, float m_b, float scale_b, int adv, float alpha, bool att = false,
                                              float lbar = 0.f) {
    if (!adv) return scale_b * sigmoidf(s);
    const float e = expf(s * alpha - m_b);
    if (att) return scale_b * e * (sigmoidf(s) - alpha * (log_sigmoid(-s) - lbar));
    return scale_b * e * sigmoidf(s);
}

// 1 / |x| of an InterHT half-row from its sum of squares.  eps = 0: tf.linalg.normalize (model.py:215-218, no clamp);
// eps > 0: F.normalize(x, 2, -1) -- x / max(|x|, eps).  proj = 0 where the clamp is active: there y = x / eps and the
// backward has no projection term.
__device__ __forceinline__ float inv_norm(float ss, float eps, float& proj) {
    const float n = sqrtf(ss);
    proj = (eps > 0.f && n < eps) ? 0.f : 1.f;
    return 1.0f / ((eps > 0.f) ? fmaxf(n, eps) : n);
}

struct SweepArgs {
    const float* ent;
    const float* rel;
    const float* modulus;
    const int64_t* pos;
    const int64_t* neg;
    const float* w;
    int B, N, adv, uni;
    float alpha;
    float *S, *Q, *GH, *GT, *GR, *GM, *stats, *terms;
    int* counts;         // NULL: uses are counted later (multi-rank apply phase)
    const float* wsum;   // NULL: sum the B local weights; else device scalar with the global sum
    int B_total;         // positives of the global batch (uniform weights: coef = 1 / B_total)
    // Split sweep (F1w, train_sweep_rows_kernel): part 0 = everything in this kernel (row groups of RPI rows);
    // 1 = stage Q[b] and return; 2 = the candidate rows were swept by F1w: merge its partial dL/dq, finish the
    // softmax statistics from S, then the positive score and the chain rule as usual.
    int part;
    int parts;           // items per positive of F1w
    const float* PA;     // [B*parts][QP*H] partial dL/dq, relative to the item's own running max
    const float* PM;     // [B*parts][2]    (running max, pRotatE modulus partial)
    int* item_counter;   // part 1: work counter of F1w, reset here (stream order puts it before F1w)
    // InterHT: (1/|a_e|, 1/|b_e|, proj_a, proj_b) of every row of `ent`, computed once per step by
    // interht_inv_norm_kernel -- the sweep reads a candidate's two norms instead of reducing them over the CTA for every
    // row group (that was 3 of the 4 block barriers and 8 of the 12 warp reductions per group of 4 rows).
    const float4* invn;
};

// One warp per entity row: sums of squares of the two half rows -> inverse norms (tf.linalg.normalize / F.normalize eps).
__global__ void __launch_bounds__(256)
interht_inv_norm_kernel(const float* __restrict__ ent, long long E, int H, float eps, float4* __restrict__ invn) {
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long e = warp; e < E; e += nw) {
        const float* row = ent + e * 2LL * H;
        float sa = 0.f, sb = 0.f;
        if ((H & 3) == 0) {
            for (int k = 4 * lane; k < H; k += 128) {
                const float4 a = __ldg(reinterpret_cast<const float4*>(row + k));
                const float4 b = __ldg(reinterpret_cast<const float4*>(row + H + k));
                sa = fmaf(a.x, a.x, fmaf(a.y, a.y, fmaf(a.z, a.z, fmaf(a.w, a.w, sa))));
                sb = fmaf(b.x, b.x, fmaf(b.y, b.y, fmaf(b.z, b.z, fmaf(b.w, b.w, sb))));
            }
        } else {
            for (int k = lane; k < H; k += 32) {
                const float a = __ldg(row + k), b = __ldg(row + H + k);
                sa = fmaf(a, a, sa);
                sb = fmaf(b, b, sb);
            }
        }
        sa = warp_sum(sa);
        sb = warp_sum(sb);
        if (lane == 0) {
            float pa, pb;
            const float ia = inv_norm(sa, eps, pa), ib = inv_norm(sb, eps, pb);
            invn[e] = make_float4(ia, ib, pa, pb);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// F1
// ------------------------------------------------------------------------------------------------
// PART (SweepArgs.part) is a template parameter so that the staging and merge passes of the split sweep do not carry
// the register footprint of the row-group sweep they skip.
// ATT: attached softmax weights (TF graph-model gradient) -- a second dL/dq accumulator (sum_n e_n ds_n/dq, the factor
// of alpha * Lbar, known only after the last row) rides along; row-group sweep (PART 0) only.
template <int MODEL, int CH, int RPI, int PART, bool ATT = false>
__global__ void __launch_bounds__(256)
train_sweep_kernel(Params P, CoordConsts kc, int head_i, SweepArgs A) {
    pdl_enter();
    constexpr int QP = model_planes(MODEL), CP = cand_planes_of(MODEL), RP = rel_planes_of(MODEL);
    __shared__ float red[4 * 32 * 2];      // block reductions: up to 8 values x 32 warps
    __shared__ float red2[2][RPI * 32];    // row-sum partials of the sweep, double-buffered by row group
    __shared__ float sh_fin[8];  // m, z, l, g, g2

    const int tid = threadIdx.x, T = blockDim.x, lane = tid & 31, warp = tid >> 5, nw = (T + 31) >> 5;
    const int H = P.H, N = A.N;
    const bool head = head_i != 0;
    const long long b = blockIdx.x;
    const long long h = A.pos[3 * b], r = A.pos[3 * b + 1], t = A.pos[3 * b + 2];
    const float* hrow = A.ent + h * (long long)P.De;
    const float* trow = A.ent + t * (long long)P.De;
    const float* rrow = A.rel + r * (long long)P.Dr + rel_offset<MODEL>(H, 0);
    if (MODEL == KGE_PROTATE) kc.modulus = __ldg(A.modulus);

    // ---- coef_b = w_b / sum(w) or 1/B -------------------------------------------------------------
    float coef;
    if (A.uni) {
        coef = 1.0f / (float)A.B_total;
    } else if (A.wsum != nullptr) {
        coef = __ldg(A.w + b) / __ldg(A.wsum);
    } else {
        float part[1] = {0.f};
        for (int i = tid; i < A.B; i += T) part[0] += __ldg(A.w + i);
        block_reduce_vals<1>(part, red);
        coef = __ldg(A.w + b) / part[0];
    }
    if (tid == 0 && A.counts != nullptr && PART != 1) {  // every use of an entity row is counted for the bucket build
        atomicAdd(A.counts + h, 1);
        atomicAdd(A.counts + t, 1);
    }

    // ---- stage the candidate-independent planes in registers ------------------------------------------
    float q[QP][CH][4];
    {
        float fe[CP][CH][4], rv[RP][CH][4];
        load_coords<CP, CH>(fe, head ? trow : hrow, H, tid, T, H);
        load_coords<RP, CH>(rv, rrow, H, tid, T, H);
        float nrm[2] = {0.f, 0.f};
        if constexpr (MODEL == KGE_INTERHT) {
            const float4 nv = __ldg(A.invn + (head ? t : h));
            nrm[0] = nv.x;
            nrm[1] = nv.y;
        }
#pragma unroll
        for (int j = 0; j < CH; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float qq[3], e2[2] = {fe[0][j][i], (CP > 1) ? fe[CP - 1][j][i] : 0.f};
                float r2[2] = {rv[0][j][i], (RP > 1) ? rv[RP - 1][j][i] : 0.f};
                float cr, sr;
                stage_coord<MODEL>(qq, e2, r2, head, kc, nrm, cr, sr);
#pragma unroll
                for (int p = 0; p < QP; ++p) q[p][j][i] = qq[p];
            }
        {  // (the merge pass rewrites the Q[b] the staging pass wrote: the store leaves the row in L2 for the entity
           //  kernel that follows -- without it that kernel measured 4 us slower, more than the store costs)
#pragma unroll
            for (int p = 0; p < QP; ++p)
#pragma unroll
                for (int j = 0; j < CH; ++j) {
                    const int k0 = 4 * (tid + j * T);
                    if (k0 < H)
                        *reinterpret_cast<float4*>(A.Q + (b * QP + p) * H + k0) =
                            make_float4(q[p][j][0], q[p][j][1], q[p][j][2], q[p][j][3]);
                }
        }
    }

    if constexpr (PART == 1) {  // staging pass of the split sweep: Q[b] is all F1w needs
        if (blockIdx.x == 0 && threadIdx.x == 0) *A.item_counter = 0;
        return;
    }

    // ---- the sweep -------------------------------------------------------------------------------
    float acc[QP][CH][4];
#pragma unroll
    for (int p = 0; p < QP; ++p)
#pragma unroll
        for (int j = 0; j < CH; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[p][j][i] = 0.f;
    float acc2[ATT ? QP : 1][ATT ? CH : 1][4];  // ATT: sum_n e_n ds_n/dq
#pragma unroll
    for (int p = 0; p < (ATT ? QP : 1); ++p)
#pragma unroll
        for (int j = 0; j < (ATT ? CH : 1); ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i) acc2[p][j][i] = 0.f;
    // online-softmax state, replicated in the lanes of warp 0
    float m_run = -INFINITY, z_run = 0.f, l_run = 0.f, g_run = 0.f, g2_run = 0.f;

    // every candidate use is counted for the bucket build (integer atomics, off the critical path)
    if (A.counts != nullptr)
        for (int n = tid; n < N; n += T) atomicAdd(A.counts + __ldg(A.neg + b * N + n), 1);
    int cid[RPI];
#pragma unroll
    for (int rr = 0; rr < RPI; ++rr) cid[rr] = (rr < N) ? (int)__ldg(A.neg + b * N + rr) : 0;

    const int n_sweep = (PART == 0) ? N : 0;
    for (int n0 = 0; n0 < n_sweep; n0 += RPI) {
        float c[RPI][CP][CH][4];
        float sg[(MODEL == KGE_INTERHT) ? RPI : 1][CH][4];
#pragma unroll
        for (int rr = 0; rr < RPI; ++rr) load_coords<CP, CH>(c[rr], A.ent + (long long)cid[rr] * P.De, H, tid, T, H);
        float cnv[(MODEL == KGE_INTERHT) ? RPI : 1][2];
        if constexpr (MODEL == KGE_INTERHT) {  // the candidates' inverse norms: one broadcast load per row
#pragma unroll
            for (int rr = 0; rr < RPI; ++rr) {
                const float4 nv = __ldg(A.invn + cid[rr]);
                cnv[rr][0] = nv.x;
                cnv[rr][1] = nv.y;
            }
        }
        // candidate ids of the NEXT row group: their latency hides behind this group's arithmetic
#pragma unroll
        for (int rr = 0; rr < RPI; ++rr) {
            const int n = n0 + RPI + rr;
            cid[rr] = (n < N) ? (int)__ldg(A.neg + b * N + n) : 0;
        }
        // phase a: row sums; c is rewritten into the backward direction
        float* redp = red2[(n0 / RPI) & 1];
        float part[RPI];
#pragma unroll
        for (int rr = 0; rr < RPI; ++rr) {
            part[rr] = 0.f;
            float cn[2] = {0.f, 0.f};
            if constexpr (MODEL == KGE_INTERHT) {
                cn[0] = cnv[rr][0];
                cn[1] = cnv[rr][1];
            }
#pragma unroll
            for (int j = 0; j < CH; ++j)
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    float qq[3] = {q[0][j][i], (QP > 1) ? q[QP > 1 ? 1 : 0][j][i] : 0.f, (QP > 2) ? q[QP > 2 ? 2 : 0][j][i] : 0.f};
                    float cc[2] = {c[rr][0][j][i], (CP > 1) ? c[rr][CP - 1][j][i] : 0.f};
                    float s1 = 0.f;
                    const bool live = 4 * (tid + j * T) < H;
                    const float term = head ? fwd_coord<MODEL, true>(qq, cc, kc, cn, s1) : fwd_coord<MODEL, false>(qq, cc, kc, cn, s1);
                    part[rr] += live ? term : 0.f;
                    c[rr][0][j][i] = cc[0];
                    if (CP > 1) c[rr][CP - 1][j][i] = cc[1];
                    if constexpr (MODEL == KGE_INTERHT) sg[rr][j][i] = live ? s1 : 0.f;
                }
            part[rr] = warp_sum(part[rr]);
            if (lane == 0) redp[rr * 32 + warp] = part[rr];
        }
        __syncthreads();
        // Every warp finishes the rows itself (lane rr takes row rr) and advances its own copy of the softmax
        // recurrence: identical arithmetic in every warp, so no second barrier and no serial section.
        float wgt_lane, wgt2_lane = 0.f, sc = 1.0f;
        {
            const int n = n0 + lane;
            const bool valid = lane < RPI && n < N;
            float tot = 0.f;
            if (lane < RPI)
                for (int w = 0; w < nw; ++w) tot += redp[lane * 32 + w];
            const float s = score_from_sum<MODEL>(tot, kc);
            if (valid && warp == 0) A.S[b * N + n] = s;
            const float ls = valid ? log_sigmoid(-s) : 0.f;
            const float sig = valid ? sigmoidf(s) : 0.f;
            if (A.adv) {
                const float a = valid ? s * A.alpha : -INFINITY;
                const float m_new = fmaxf(m_run, warp_max(a));
                sc = (m_run == -INFINITY) ? 0.f : expf(m_run - m_new);
                const float e = valid ? expf(a - m_new) : 0.f;
                z_run = z_run * sc + warp_sum(e);
                l_run = l_run * sc + warp_sum(e * ls);
                wgt_lane = e * sig;
                if constexpr (ATT) {
                    wgt_lane = e * (sig - A.alpha * ls);
                    wgt2_lane = e;
                    if (MODEL == KGE_PROTATE) g2_run = g2_run * sc + warp_sum(e * tot);
                }
                if (MODEL == KGE_PROTATE) g_run = g_run * sc + warp_sum(wgt_lane * tot);
                m_run = m_new;
            } else {
                l_run += warp_sum(ls);
                if (MODEL == KGE_PROTATE) g_run += warp_sum(sig * tot);
                wgt_lane = sig;
            }
        }
        // phase b: acc = acc * sc + sum_r wgt_r * ds_r/dq
        if (sc != 1.0f) {
#pragma unroll
            for (int p = 0; p < QP; ++p)
#pragma unroll
                for (int j = 0; j < CH; ++j)
#pragma unroll
                    for (int i = 0; i < 4; ++i) acc[p][j][i] *= sc;
            if constexpr (ATT) {
#pragma unroll
                for (int p = 0; p < QP; ++p)
#pragma unroll
                    for (int j = 0; j < CH; ++j)
#pragma unroll
                        for (int i = 0; i < 4; ++i) acc2[p][j][i] *= sc;
            }
        }
#pragma unroll
        for (int rr = 0; rr < RPI; ++rr) {
            const float wgt = __shfl_sync(0xffffffffu, wgt_lane, rr);
            const float wgt2 = ATT ? __shfl_sync(0xffffffffu, wgt2_lane, rr) : 0.f;
#pragma unroll
            for (int j = 0; j < CH; ++j)
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    float aa[3] = {acc[0][j][i], (QP > 1) ? acc[QP > 1 ? 1 : 0][j][i] : 0.f, (QP > 2) ? acc[QP > 2 ? 2 : 0][j][i] : 0.f};
                    const float dd[2] = {c[rr][0][j][i], (CP > 1) ? c[rr][CP - 1][j][i] : 0.f};
                    const float s1 = (MODEL == KGE_INTERHT) ? sg[(MODEL == KGE_INTERHT) ? rr : 0][j][i] : 0.f;
                    if (head) dq_accumulate<MODEL, true>(aa, dd, s1, wgt, kc);
                    else dq_accumulate<MODEL, false>(aa, dd, s1, wgt, kc);
#pragma unroll
                    for (int p = 0; p < QP; ++p) acc[p][j][i] = aa[p];
                    if constexpr (ATT) {
                        float a2[3] = {acc2[0][j][i], (QP > 1) ? acc2[QP > 1 ? 1 : 0][j][i] : 0.f, (QP > 2) ? acc2[QP > 2 ? 2 : 0][j][i] : 0.f};
                        if (head) dq_accumulate<MODEL, true>(a2, dd, s1, wgt2, kc);
                        else dq_accumulate<MODEL, false>(a2, dd, s1, wgt2, kc);
#pragma unroll
                        for (int p = 0; p < QP; ++p) acc2[p][j][i] = a2[p];
                    }
                }
        }
    }

    if (PART == 2) {
        // ---- merge of the split sweep: scores S[b,:] and the per-item partials are in global memory --------
        float m = -INFINITY;
        if (A.adv)
            for (int p = 0; p < A.parts; ++p) m = fmaxf(m, __ldg(A.PM + 2 * (b * A.parts + p)));
        float zl[2] = {0.f, 0.f};
        for (int n = tid; n < N; n += T) {
            const float s = A.S[b * N + n];
            const float ls = log_sigmoid(-s);
            if (A.adv) {
                const float e = expf(s * A.alpha - m);
                zl[0] += e;
                zl[1] += e * ls;
            } else {
                zl[1] += ls;
            }
        }
        block_reduce_vals<2>(zl, red);
        m_run = m;
        z_run = zl[0];
        l_run = zl[1];
        g_run = 0.f;
        // fixed order: bit-reproducible.  The partials of four items are requested together (they do not depend on
        // one another: one L2 round trip instead of four), then folded in ascending order.
        constexpr int MU = (CH == 1) ? 4 : 2;
        for (int p0 = 0; p0 < A.parts; p0 += MU) {
            float4 v[MU][QP][CH];
            float mp[MU], gp[MU];
#pragma unroll
            for (int u = 0; u < MU; ++u) {
                const long long item = b * A.parts + min(p0 + u, A.parts - 1);
                mp[u] = __ldg(A.PM + 2 * item);
                gp[u] = __ldg(A.PM + 2 * item + 1);
#pragma unroll
                for (int pl = 0; pl < QP; ++pl)
#pragma unroll
                    for (int j = 0; j < CH; ++j) {
                        const int k0 = 4 * (tid + j * T);
                        v[u][pl][j] = (k0 < H) ? __ldg(reinterpret_cast<const float4*>(A.PA + (item * QP + pl) * H + k0))
                                               : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
            }
#pragma unroll
            for (int u = 0; u < MU; ++u) {
                if (p0 + u >= A.parts) break;
                const float f = A.adv ? ((mp[u] == -INFINITY) ? 0.f : expf(mp[u] - m)) : 1.0f;
                g_run = fmaf(gp[u], f, g_run);
#pragma unroll
                for (int pl = 0; pl < QP; ++pl)
#pragma unroll
                    for (int j = 0; j < CH; ++j) {
                        if (4 * (tid + j * T) < H) {
                            acc[pl][j][0] = fmaf(v[u][pl][j].x, f, acc[pl][j][0]);
                            acc[pl][j][1] = fmaf(v[u][pl][j].y, f, acc[pl][j][1]);
                            acc[pl][j][2] = fmaf(v[u][pl][j].z, f, acc[pl][j][2]);
                            acc[pl][j][3] = fmaf(v[u][pl][j].w, f, acc[pl][j][3]);
                        }
                    }
            }
        }
    }

    // ---- row statistics -> every thread ------------------------------------------------------------
    if (tid == 0) {
        sh_fin[0] = m_run;
        sh_fin[1] = z_run;
        sh_fin[2] = l_run;
        sh_fin[3] = g_run;
        sh_fin[4] = g2_run;
    }
    __syncthreads();
    const float m_b = sh_fin[0];
    const float denom = A.adv ? sh_fin[1] : (float)N;
    const float negterm = sh_fin[2] / denom;
    const float scale_b = 0.5f * coef / denom;  // dL/ds_n = scale_b * e_n * sigma(s_n)
    const bool att = ATT && A.adv;
    const float att_k = att ? A.alpha * negterm : 0.f;  // weight of the second accumulator: alpha * Lbar
    const float gsum = sh_fin[3] + att_k * sh_fin[4];
    if constexpr (ATT) {
#pragma unroll
        for (int p = 0; p < QP; ++p)
#pragma unroll
            for (int j = 0; j < CH; ++j)
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[p][j][i] = fmaf(att_k, acc2[p][j][i], acc[p][j][i]);
    }
    // F2b folded in: S[b,:] <- dL/ds.  The raw scores are not read again by anybody; every later consumer (the entity
    // kernel of this rank, or of the row owners after the exchange) wants the derivative, and here the work stays with
    // the rank that owns the positive instead of being repeated on the global batch by every rank.
    for (int n = tid; n < N; n += T) {
        const long long i = b * N + n;
        A.S[i] = dscore_value(A.S[i], m_b, scale_b, A.adv, A.alpha, att, negterm);
    }
    // dL/dq of the negative part
#pragma unroll
    for (int p = 0; p < QP; ++p)
#pragma unroll
        for (int j = 0; j < CH; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[p][j][i] *= scale_b;

    // ---- positive score in 'single' form, its gradient, and the chain to (h, r, t) ------------------
    float he[CP][CH][4], te[CP][CH][4], rv[RP][CH][4];
    load_coords<CP, CH>(he, hrow, H, tid, T, H);
    load_coords<CP, CH>(te, trow, H, tid, T, H);
    load_coords<RP, CH>(rv, rrow, H, tid, T, H);
    float nh[2] = {0.f, 0.f}, nt[2] = {0.f, 0.f};
    float pjh[2] = {1.f, 1.f}, pjt[2] = {1.f, 1.f};  // InterHT: 0 where the F.normalize clamp is active (no projection)
    if constexpr (MODEL == KGE_INTERHT) {  // the same table values the staging above and the candidates use
        const float4 vh = __ldg(A.invn + h), vt = __ldg(A.invn + t);
        nh[0] = vh.x; nh[1] = vh.y; pjh[0] = vh.z; pjh[1] = vh.w;
        nt[0] = vt.x; nt[1] = vt.y; pjt[0] = vt.z; pjt[1] = vt.w;
    }
    // q1 = planes in single form (fixed entity = head); equal to q in tail-batch mode
    float q1[QP][CH][4], crs[CH][4], srs[CH][4];
    float d1[CP][CH][4], sg1[CH][4];
    float ptot[1] = {0.f};
#pragma unroll
    for (int j = 0; j < CH; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float qq[3], e2[2] = {he[0][j][i], (CP > 1) ? he[CP - 1][j][i] : 0.f};
            float r2[2] = {rv[0][j][i], (RP > 1) ? rv[RP - 1][j][i] : 0.f};
            float cr = 0.f, sr = 0.f;
            stage_coord<MODEL>(qq, e2, r2, false, kc, nh, cr, sr);
            crs[j][i] = cr;
            srs[j][i] = sr;
#pragma unroll
            for (int p = 0; p < QP; ++p) q1[p][j][i] = qq[p];
            float cc[2] = {te[0][j][i], (CP > 1) ? te[CP - 1][j][i] : 0.f};
            float s1 = 0.f;
            const float term = fwd_coord<MODEL, false>(qq, cc, kc, nt, s1);
            ptot[0] += (4 * (tid + j * T) < H) ? term : 0.f;
            d1[0][j][i] = cc[0];
            if (CP > 1) d1[CP - 1][j][i] = cc[1];
            sg1[j][i] = (4 * (tid + j * T) < H) ? s1 : 0.f;
        }
    block_reduce_vals<1>(ptot, red);
    const float s_pos = score_from_sum<MODEL>(ptot[0], kc);
    const float dpos = -0.5f * coef * sigmoidf(-s_pos);

    // dL/dq1 from the positive score; dL/dt (candidate side of the positive score)
    float gq1[QP][CH][4], gt_c[CP][CH][4];
    float dots[2] = {0.f, 0.f};  // InterHT: <a_hat, d a_hat>, <b_hat, d b_hat> of the candidate t
#pragma unroll
    for (int j = 0; j < CH; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float aa[3] = {0.f, 0.f, 0.f};
            const float dd[2] = {d1[0][j][i], (CP > 1) ? d1[CP - 1][j][i] : 0.f};
            dq_accumulate<MODEL, false>(aa, dd, sg1[j][i], dpos, kc);
#pragma unroll
            for (int p = 0; p < QP; ++p) gq1[p][j][i] = aa[p];
            float gc[2] = {0.f, 0.f};
            if constexpr (MODEL == KGE_INTERHT) {
                // v = A*bc - ac*Bv + M ; u = dL/dv ; d ac = -u*Bv ; d bc = u*A
                const float u = -dpos * sg1[j][i];
                const float d_ac = -u * q1[1][j][i], d_bc = u * q1[0][j][i];
                gc[0] = d_ac;
                gc[1] = d_bc;
                dots[0] = fmaf(dd[0], d_ac, dots[0]);
                dots[1] = fmaf(dd[1] - 1.0f, d_bc, dots[1]);
            } else {
                const float qq[3] = {q1[0][j][i], (QP > 1) ? q1[QP > 1 ? 1 : 0][j][i] : 0.f, 0.f};
                const float cc[2] = {te[0][j][i], (CP > 1) ? te[CP - 1][j][i] : 0.f};
                dc_accumulate<MODEL, false>(gc, qq, cc, dpos, kc);
            }
            gt_c[0][j][i] = gc[0];
            if (CP > 1) gt_c[CP - 1][j][i] = gc[1];
        }
    if constexpr (MODEL == KGE_INTERHT) {
        block_reduce_vals<2>(dots, red);
        dots[0] *= pjt[0];
        dots[1] *= pjt[1];
#pragma unroll
        for (int j = 0; j < CH; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                gt_c[0][j][i] = (gt_c[0][j][i] - d1[0][j][i] * dots[0]) * nt[0];
                gt_c[1][j][i] = (gt_c[1][j][i] - (d1[1][j][i] - 1.0f) * dots[1]) * nt[1];
            }
    }

    // chain: mode-form planes (q, acc) and single-form planes (q1, gq1)
    if (!head) {  // same planes: add the two dL/dq
#pragma unroll
        for (int p = 0; p < QP; ++p)
#pragma unroll
            for (int j = 0; j < CH; ++j)
#pragma unroll
                for (int i = 0; i < 4; ++i) gq1[p][j][i] += acc[p][j][i];
    }
    float dA[4] = {0.f, 0.f, 0.f, 0.f};  // InterHT: <A,dA>,<Bh,dB> for the single-form chain and the mode-form chain
    if constexpr (MODEL == KGE_INTERHT) {
#pragma unroll
        for (int j = 0; j < CH; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                dA[0] = fmaf(q1[0][j][i], gq1[0][j][i], dA[0]);
                dA[1] = fmaf(q1[1][j][i] - 1.0f, gq1[1][j][i], dA[1]);
                dA[2] = fmaf(q[0][j][i], acc[0][j][i], dA[2]);
                dA[3] = fmaf(q[1][j][i] - 1.0f, acc[1][j][i], dA[3]);
            }
        block_reduce_vals<4>(dA, red);
        dA[0] *= pjh[0]; dA[1] *= pjh[1];  // single-form chain: fixed entity = h
        dA[2] *= pjt[0]; dA[3] *= pjt[1];  // mode-form chain (head-batch): fixed entity = t
    }
    float* gh = A.GH + b * (long long)P.De;
    float* gt = A.GT + b * (long long)P.De;
    float* gr = A.GR + b * (long long)P.Dr;
#pragma unroll
    for (int j = 0; j < CH; ++j) {
        const int k0 = 4 * (tid + j * T);
        float oh[CP][4], ot[CP][4], orl[RP][4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            // single-form chain: fixed entity = h
            float ge[2], grr[2];
            const float g1[3] = {gq1[0][j][i], (QP > 1) ? gq1[QP > 1 ? 1 : 0][j][i] : 0.f, (QP > 2) ? gq1[QP > 2 ? 2 : 0][j][i] : 0.f};
            const float qa[3] = {q1[0][j][i], (QP > 1) ? q1[QP > 1 ? 1 : 0][j][i] : 0.f, (QP > 2) ? q1[QP > 2 ? 2 : 0][j][i] : 0.f};
            const float e2[2] = {he[0][j][i], (CP > 1) ? he[CP - 1][j][i] : 0.f};
            const float r2[2] = {rv[0][j][i], (RP > 1) ? rv[RP - 1][j][i] : 0.f};
            chain_coord<MODEL>(ge, grr, g1, qa, e2, r2, false, kc, crs[j][i], srs[j][i], dA[0], dA[1], nh);
            oh[0][i] = ge[0];
            if (CP > 1) oh[CP - 1][i] = ge[1];
            orl[0][i] = grr[0];
            if (RP > 1) orl[RP - 1][i] = grr[1];
            ot[0][i] = gt_c[0][j][i];
            if (CP > 1) ot[CP - 1][i] = gt_c[CP - 1][j][i];
            if (head) {  // mode-form chain: fixed entity = t
                const float g2[3] = {acc[0][j][i], (QP > 1) ? acc[QP > 1 ? 1 : 0][j][i] : 0.f, (QP > 2) ? acc[QP > 2 ? 2 : 0][j][i] : 0.f};
                const float qb[3] = {q[0][j][i], (QP > 1) ? q[QP > 1 ? 1 : 0][j][i] : 0.f, (QP > 2) ? q[QP > 2 ? 2 : 0][j][i] : 0.f};
                const float t2[2] = {te[0][j][i], (CP > 1) ? te[CP - 1][j][i] : 0.f};
                chain_coord<MODEL>(ge, grr, g2, qb, t2, r2, true, kc, crs[j][i], srs[j][i], dA[2], dA[3], nt);
                ot[0][i] += ge[0];
                if (CP > 1) ot[CP - 1][i] += ge[1];
                orl[0][i] += grr[0];
                if (RP > 1) orl[RP - 1][i] += grr[1];
            }
        }
        if (k0 < H) {
#pragma unroll
            for (int p = 0; p < CP; ++p) {
                *reinterpret_cast<float4*>(gh + p * H + k0) = make_float4(oh[p][0], oh[p][1], oh[p][2], oh[p][3]);
                *reinterpret_cast<float4*>(gt + p * H + k0) = make_float4(ot[p][0], ot[p][1], ot[p][2], ot[p][3]);
            }
#pragma unroll
            for (int p = 0; p < RP; ++p)
                *reinterpret_cast<float4*>(gr + rel_offset<MODEL>(H, p) + k0) =
                    make_float4(orl[p][0], orl[p][1], orl[p][2], orl[p][3]);
            if constexpr (MODEL == KGE_INTERHT) {  // the outer thirds of the relation row get no gradient
                *reinterpret_cast<float4*>(gr + k0) = make_float4(0.f, 0.f, 0.f, 0.f);
                *reinterpret_cast<float4*>(gr + 2 * H + k0) = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
    }
    if (tid == 0) {
        A.stats[2 * b] = m_b;
        A.stats[2 * b + 1] = scale_b;
        A.terms[2 * b] = coef * log_sigmoid(s_pos);
        A.terms[2 * b + 1] = coef * negterm;
        if (MODEL == KGE_PROTATE) A.GM[b] = -(scale_b * gsum + dpos * ptot[0]);
    }
}

// ------------------------------------------------------------------------------------------------
// F1w: the candidate sweep, one WARP per row (split sweep, parts 1 + F1w + 2)
//
// Work item = (positive b, part p): rows [p*rpp, (p+1)*rpp) of b's candidates.  A warp owns an item: lane l owns
// the coordinates 4*(l + 32*j), j < CH, of every row, so a row sum is ONE warp reduction -- no block barrier, no
// shared-memory round trip, and the online-softmax recurrence advances once per row inside the warp instead of
// once per row group in every warp of a CTA.  q_b comes from Q[b] (staged by part 1) through shared memory, the
// dL/dq accumulator stays in registers relative to the item's own running max and is written once per item
// (PA / PM); part 2 merges the items of a positive in fixed order.  Items are handed out by an atomic counter
// (8 items per positive: 8192 items for 1776 resident warps), results do not depend on who ran which item.
// The next row is prefetched into L2 while the current one is processed.
// ------------------------------------------------------------------------------------------------
struct RowsArgs {
    const float* ent;
    const int64_t* neg;
    const float* Q;
    const float* modulus;
    float* S;
    float* PA;
    float* PM;
    int* counter;
    int B, N, parts, rpp, adv;
    float alpha;
};

// Candidate rows reach the warp through a warp-private ring of R shared-memory slots filled by 1-D bulk copies (TMA,
// one instruction per 8 KB row, completion on an mbarrier): the copy of row n + R starts as soon as row n has been
// lifted into registers, so it overlaps the reduction and the accumulation pass of rows n .. n + R - 1.
//
// Round 2: a work item belongs to a CTA, not to a warp.  Item = (positive b, part p) = up to 128 consecutive candidates
// of b (fewer, larger parts than before: ceil(N / 128), at most 8); the 4 warps of the CTA take a quarter each.  q_b is
// staged ONCE per item for the four warps (8 kB per CTA instead of per warp), which leaves room for R = 2 row slots per
// warp at the same 3 CTAs per SM: two rows in flight per warp.  At the end of the item the four partial dL/dq (each
// relative to its warp's running max) are merged in fixed order through the row slots, so one partial per 128 rows
// leaves the SM instead of one per 32: a quarter of the PA traffic, of the Q re-reads and of the merge pass's work.
__host__ __device__ constexpr uint32_t rows_warp_region(uint32_t slot_bytes, int R, int QP, int CH) {
    return (R * slot_bytes > (uint32_t)(QP * CH * 512)) ? R * slot_bytes : (uint32_t)(QP * CH * 512);
}

template <int MODEL, int CH, bool HEAD, int R>
__global__ void __launch_bounds__(128, 3)
train_sweep_rows_kernel(Params P, CoordConsts kc, RowsArgs A) {
    pdl_enter();
    constexpr int QP = model_planes(MODEL), CP = cand_planes_of(MODEL);
    constexpr int QV = QP * CH * 32;     // float4 per staged q_b (and per partial dL/dq)
    extern __shared__ float4 f1w_q[];    // [QP][CH][32] q_b | [warp] region (R row slots / merge scratch) | [warp][R] mbarriers
    __shared__ int sh_item;
    __shared__ float sh_pm[4][2];        // per warp: running max, modulus partial
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float4* myq = f1w_q;
    const uint32_t row_bytes = (uint32_t)P.De * 4u, slot_bytes = (row_bytes + 127u) & ~127u;
    const uint32_t region = rows_warp_region(slot_bytes, R, QP, CH);
    const uint32_t regions_s = smem_u32(f1w_q + QV);
    const uint32_t ring_s = regions_s + warp * region;
    const uint32_t bars_s = regions_s + 4 * region + warp * R * 8;
    if (lane == 0) {
#pragma unroll
        for (int r = 0; r < R; ++r) mbar_init(bars_s + 8 * r, 1);
        fence_barrier_init();
    }
    __syncthreads();
    uint32_t issued = 0, consumed = 0;  // rows of this warp so far: slot = count % R, parity = (count / R) & 1
    const int H = P.H, N = A.N;
    if (MODEL == KGE_PROTATE) kc.modulus = __ldg(A.modulus);
    const int total = A.B * A.parts;
    const int jfull = H / 128;  // chunks whose 32 lanes all hold live coordinates
    const float cn[2] = {0.f, 0.f};
    const int rpw = (A.rpp + 3) / 4;  // rows of a part per warp

    for (;;) {
        __syncthreads();  // the previous item is over: q_b and the merge scratch may be rewritten
        if (tid == 0) sh_item = atomicAdd(A.counter, 1);
        __syncthreads();
        const int item = sh_item;
        if (item >= total) break;
        const long long b = item / A.parts;
        const int p0 = (item % A.parts) * A.rpp, p1 = min(N, p0 + A.rpp);
        const int n0 = min(p1, p0 + warp * rpw), n1 = min(p1, n0 + rpw);  // this warp's quarter (may be empty)

        for (int idx = tid; idx < QV; idx += 128) {
            const int pl = idx / (CH * 32), k0 = 4 * ((idx & 31) + 32 * ((idx >> 5) % CH));
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);  // padding coordinates: q = 0 and c = 0 give a zero term
            if (k0 < H) v = __ldg(reinterpret_cast<const float4*>(A.Q + (b * QP + pl) * H + k0));
            f1w_q[idx] = v;
        }
        __syncthreads();

        float acc[QP][CH][4];
#pragma unroll
        for (int pl = 0; pl < QP; ++pl)
#pragma unroll
            for (int j = 0; j < CH; ++j)
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[pl][j][i] = 0.f;
        float m_run = -INFINITY, g_run = 0.f;

        auto issue_row = [&](long long rid) {  // lane 0 only
            const uint32_t slot = issued % R, bar = bars_s + 8 * slot;
            mbar_arrive_expect_tx(bar, row_bytes);
            bulk_g2s(ring_s + slot * slot_bytes, A.ent + rid * (long long)P.De, row_bytes, bar);
        };
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (n0 + r < n1) {
                const long long rid = __ldg(A.neg + b * N + n0 + r);
                if (lane == 0) issue_row(rid);
                ++issued;
            }
        for (int n = n0; n < n1; ++n) {
            // id of the row that will take this slot next; its latency hides behind the wait below
            const long long id_ahead = (n + R < n1) ? __ldg(A.neg + b * N + n + R) : 0;
            const uint32_t slot = consumed % R;
            mbar_wait(bars_s + 8 * slot, (consumed / R) & 1u);
            ++consumed;
            const uint32_t row_s = ring_s + slot * slot_bytes + 16u * lane;
            float c[CP][CH][4];
#pragma unroll
            for (int pl = 0; pl < CP; ++pl)
#pragma unroll
                for (int j = 0; j < CH; ++j) {
                    const int k0 = 4 * (lane + 32 * j);
                    c[pl][j][0] = c[pl][j][1] = c[pl][j][2] = c[pl][j][3] = 0.f;
                    if (j < jfull || k0 < H)  // j < jfull is warp-uniform: every lane of the chunk is inside the row
                        lds128(c[pl][j], row_s + (uint32_t)(pl * H + 128 * j) * 4u);
                }
            // The slot is refilled by the async proxy (TMA) while it was read through the generic proxy: the proxy fence
            // orders every lane's reads before the refill that lane 0 issues after the warp barrier.  Without it the
            // refill was seen to overtake a lane's reads about once in 10^5 rows (config 2: three of 8192 items mixed the
            // coordinates of two consecutive rows -- scratch/diag_sweep.py, tests/test_gpu_baseline_sizes.py).
            fence_proxy_async();
            __syncwarp();  // every lane has lifted its part of the slot
            if (n + R < n1) {
                if (lane == 0) issue_row(id_ahead);
                ++issued;
            }
            float part = 0.f;
            if constexpr (MODEL == KGE_ROTATE) {
                // RotatE on packed fp32 (fma.rn.f32x2 / FFMA2: two coordinates per instruction): v = q - c,
                // |v|, direction v / |v| left in c for the accumulation below.  rsqrt(max(n2, tiny)) gives
                // direction 0 and norm 0 at v = 0 (torch: d|v|/dv = 0 there) without a select.
                const float2 neg1 = make_float2(-1.f, -1.f);
                float2 part2 = make_float2(0.f, 0.f);
#pragma unroll
                for (int j = 0; j < CH; ++j) {
                    const float4 q0 = myq[(0 * CH + j) * 32 + lane];
                    const float4 q1 = myq[(1 * CH + j) * 32 + lane];
#pragma unroll
                    for (int h2 = 0; h2 < 2; ++h2) {
                        const float2 qr = h2 ? make_float2(q0.z, q0.w) : make_float2(q0.x, q0.y);
                        const float2 qi = h2 ? make_float2(q1.z, q1.w) : make_float2(q1.x, q1.y);
                        const float2 cr = make_float2(c[0][j][2 * h2], c[0][j][2 * h2 + 1]);
                        const float2 ci = make_float2(c[1][j][2 * h2], c[1][j][2 * h2 + 1]);
                        const float2 vr = __ffma2_rn(cr, neg1, qr), vi = __ffma2_rn(ci, neg1, qi);
                        const float2 n2 = __ffma2_rn(vr, vr, __fmul2_rn(vi, vi));
                        const float2 rs = make_float2(fast_rsqrt(fmaxf(n2.x, 1e-37f)), fast_rsqrt(fmaxf(n2.y, 1e-37f)));
                        const float2 dr = __fmul2_rn(vr, rs), di = __fmul2_rn(vi, rs);
                        part2 = __ffma2_rn(n2, rs, part2);
                        c[0][j][2 * h2] = dr.x; c[0][j][2 * h2 + 1] = dr.y;
                        c[1][j][2 * h2] = di.x; c[1][j][2 * h2 + 1] = di.y;
                    }
                }
                part = part2.x + part2.y;
            } else
#pragma unroll
            for (int j = 0; j < CH; ++j) {
                const float4 q0 = myq[(0 * CH + j) * 32 + lane];
                const float4 q1 = (QP > 1) ? myq[((QP > 1 ? 1 : 0) * CH + j) * 32 + lane] : make_float4(0.f, 0.f, 0.f, 0.f);
                const float q0a[4] = {q0.x, q0.y, q0.z, q0.w}, q1a[4] = {q1.x, q1.y, q1.z, q1.w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float qq[3] = {q0a[i], q1a[i], 0.f};
                    float cc[2] = {c[0][j][i], (CP > 1) ? c[CP - 1][j][i] : 0.f};
                    float s1 = 0.f;
                    part += fwd_coord<MODEL, HEAD>(qq, cc, kc, cn, s1);
                    c[0][j][i] = cc[0];
                    if (CP > 1) c[CP - 1][j][i] = cc[1];
                }
            }
            const float tot = warp_sum(part);
            const float s = score_from_sum<MODEL>(tot, kc);
            if (lane == 0) A.S[b * N + n] = s;
            float wgt = sigmoidf(s);
            if (A.adv) {
                const float a = s * A.alpha;
                if (a > m_run) {  // warp-uniform; rare once the running max has settled
                    const float sc = (m_run == -INFINITY) ? 0.f : expf(m_run - a);
                    m_run = a;
                    g_run *= sc;
#pragma unroll
                    for (int pl = 0; pl < QP; ++pl)
#pragma unroll
                        for (int j = 0; j < CH; ++j)
#pragma unroll
                            for (int i = 0; i < 4; ++i) acc[pl][j][i] *= sc;
                }
                wgt *= expf(a - m_run);
            }
            if (MODEL == KGE_PROTATE) g_run = fmaf(wgt, tot, g_run);
            if constexpr (MODEL == KGE_ROTATE) {
                const float2 nw = make_float2(-wgt, -wgt);  // ds/dq = -direction
#pragma unroll
                for (int pl = 0; pl < 2; ++pl)
#pragma unroll
                    for (int j = 0; j < CH; ++j)
#pragma unroll
                        for (int h2 = 0; h2 < 2; ++h2) {
                            const float2 a2 = __ffma2_rn(make_float2(c[pl][j][2 * h2], c[pl][j][2 * h2 + 1]), nw,
                                                         make_float2(acc[pl][j][2 * h2], acc[pl][j][2 * h2 + 1]));
                            acc[pl][j][2 * h2] = a2.x;
                            acc[pl][j][2 * h2 + 1] = a2.y;
                        }
            } else
#pragma unroll
            for (int j = 0; j < CH; ++j)
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    float aa[3] = {acc[0][j][i], (QP > 1) ? acc[QP > 1 ? 1 : 0][j][i] : 0.f, 0.f};
                    const float dd[2] = {c[0][j][i], (CP > 1) ? c[CP - 1][j][i] : 0.f};
                    dq_accumulate<MODEL, HEAD>(aa, dd, 0.f, wgt, kc);
                    acc[0][j][i] = aa[0];
                    if (QP > 1) acc[QP > 1 ? 1 : 0][j][i] = aa[1];
                }
        }
        // ---- merge of the four warps' partials, fixed order (warp 0, 1, 2, 3) --------------------------------------
        // every copy into this warp's slots has landed and has been read (each row was waited for and lifted): the
        // region is free to hold the warp's partial, lane-owned layout [(pl, j)][lane]
        {
            float4* mine = f1w_q + QV + warp * (region / 16);
#pragma unroll
            for (int pl = 0; pl < QP; ++pl)
#pragma unroll
                for (int j = 0; j < CH; ++j)
                    mine[(pl * CH + j) * 32 + lane] = make_float4(acc[pl][j][0], acc[pl][j][1], acc[pl][j][2], acc[pl][j][3]);
            if (lane == 0) {
                sh_pm[warp][0] = m_run;
                sh_pm[warp][1] = g_run;
            }
        }
        __syncthreads();
        {
            float m = -INFINITY, f[4], g = 0.f;
            if (A.adv) {
#pragma unroll
                for (int w = 0; w < 4; ++w) m = fmaxf(m, sh_pm[w][0]);
            }
#pragma unroll
            for (int w = 0; w < 4; ++w) {
                f[w] = A.adv ? ((sh_pm[w][0] == -INFINITY) ? 0.f : expf(sh_pm[w][0] - m)) : 1.0f;
                g = fmaf(sh_pm[w][1], f[w], g);
            }
            for (int idx = tid; idx < QV; idx += 128) {
                const int pl = idx / (CH * 32), k0 = 4 * ((idx & 31) + 32 * ((idx >> 5) % CH));
                float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int w = 0; w < 4; ++w) {
                    const float4 v = f1w_q[QV + w * (region / 16) + idx];
                    o.x = fmaf(v.x, f[w], o.x);
                    o.y = fmaf(v.y, f[w], o.y);
                    o.z = fmaf(v.z, f[w], o.z);
                    o.w = fmaf(v.w, f[w], o.w);
                }
                if (k0 < H) *reinterpret_cast<float4*>(A.PA + ((long long)item * QP + pl) * H + k0) = o;
            }
            if (tid == 0) {
                A.PM[2 * (long long)item] = m;
                A.PM[2 * (long long)item + 1] = g;
            }
        }
        // the regions were written and read through the generic proxy; the next item's row copies (async proxy) come
        // after this fence and the barrier at the top of the loop
        fence_proxy_async();
    }
}

// ------------------------------------------------------------------------------------------------
// F2: per-bucket sort (one warp per entity): ascending item order whatever order the fill produced
// ------------------------------------------------------------------------------------------------
// Buckets of up to kSortSmem uses are staged in a warp-private shared-memory strip and sorted there (bitonic network on
// the strip padded to a power of two: O(n log^2 n) instead of the n^2 rank counting of round 1, which read the bucket
// from global memory once per item -- at 8 ranks a row of the global batch has 8x the uses and the quadratic pass was the
// largest part of the apply's bucket build).  Item ids are unique, so the order is total.  Longer buckets keep the
// rank-counting path (correct for any length).
constexpr int kSortSmem = 1024;
__global__ void __launch_bounds__(256)
bucket_sort_kernel(const int* __restrict__ offsets, const int* __restrict__ in, int* __restrict__ out,
                   int* __restrict__ out_b, long long E, int N) {
    pdl_enter();
    __shared__ int strip_all[8][kSortSmem];
    const int lane = threadIdx.x & 31;
    int* strip = strip_all[threadIdx.x >> 5];
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long e = warp; e < E; e += nwarps) {
        const int beg = offsets[e], end = offsets[e + 1], len = end - beg;
        if (len <= 0) continue;
        if (len <= kSortSmem) {
            int P = 32;
            while (P < len) P <<= 1;
            for (int i = lane; i < P; i += 32) strip[i] = (i < len) ? in[beg + i] : 0x7fffffff;
            __syncwarp();
            for (int k = 2; k <= P; k <<= 1)
                for (int j = k >> 1; j > 0; j >>= 1) {
                    for (int i = lane; i < P; i += 32) {
                        const int ixj = i ^ j;
                        if (ixj > i) {
                            const int a = strip[i], b = strip[ixj];
                            if ((a > b) == ((i & k) == 0)) {
                                strip[i] = b;
                                strip[ixj] = a;
                            }
                        }
                    }
                    __syncwarp();
                }
            for (int i = lane; i < len; i += 32) {
                const int it = strip[i];
                out[beg + i] = it;
                out_b[beg + i] = it / N;  // positive index of a candidate use (unused for head / tail uses)
            }
            __syncwarp();  // the strip is reused by the warp's next bucket
        } else {
            for (int idx = beg + lane; idx < end; idx += 32) {
                const int it = in[idx];
                int rank = 0;
                for (int j = beg; j < end; ++j) rank += (in[j] < it) ? 1 : 0;
                out[beg + rank] = it;
                out_b[beg + rank] = it / N;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// F3: entity-major gradient gather fused with Adam
// ------------------------------------------------------------------------------------------------
// One relation row: dL/d(row) = sum of GR[b] over the positives b that use it, in ascending b; then Adam.  Called by a
// whole CTA (any size); match / nmatch are shared-memory scratch (relation_scratch_ints(B) ints, one int).
__host__ __device__ inline int relation_scratch_ints(int B) { return B + (B + 31) / 32; }
__device__ __forceinline__ void relation_row_update(long long r, int Dr, const int64_t* __restrict__ pos, int B,
                                                    const float* __restrict__ GR, float* rel, float* m, float* v,
                                                    const AdamScalars& ad, float l3, int* match, int* nmatch) {
    // every warp tests 32 positives at a time and leaves one bitmap word per group -- four groups per round, so that
    // four independent loads are in flight per lane; warp 0 then expands the bitmap into the list of matching positives
    // in ascending b (each lane owns a run of consecutive words; a warp scan of the run totals places them)
    unsigned* bits = reinterpret_cast<unsigned*>(match + B);
    const int words = (B + 31) / 32, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int w0 = warp; w0 < words; w0 += 4 * nw) {
        long long rid[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int b = 32 * (w0 + u * nw) + lane;
            rid[u] = (b < B) ? pos[3 * (long long)b + 1] : -1;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int w = w0 + u * nw;
            if (w < words) {  // warp-uniform
                const unsigned mask = __ballot_sync(0xffffffffu, rid[u] == r);
                if (lane == 0) bits[w] = mask;
            }
        }
    }
    __syncthreads();
    if (threadIdx.x < 32) {
        const int per = (words + 31) / 32, w_lo = min(lane * per, words), w_hi = min(w_lo + per, words);
        int cnt = 0;
        for (int w = w_lo; w < w_hi; ++w) cnt += __popc(bits[w]);
        int off = cnt;  // inclusive scan of the run totals
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, off, o);
            if (lane >= o) off += v;
        }
        if (lane == 31) *nmatch = off;
        off -= cnt;
        for (int w = w_lo; w < w_hi; ++w) {
            unsigned mask = bits[w];
            while (mask) {
                match[off++] = 32 * w + (__ffs(mask) - 1);
                mask &= mask - 1;
            }
        }
    }
    __syncthreads();
    const int nm = *nmatch;
    // dL/d(row) = sum of GR[match[j]] in ascending j; a thread owns up to four coordinates (stride blockDim) and keeps
    // 16 independent loads in flight (4 positives x 4 coordinates); the additions of a coordinate stay in ascending j
    for (int kb = threadIdx.x; kb < Dr; kb += 4 * blockDim.x) {
        float g[4] = {0.f, 0.f, 0.f, 0.f};
        int j = 0;
        for (; j + 4 <= nm; j += 4) {
            float x[4][4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float* src = GR + (long long)match[j + u] * Dr;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int k = kb + c * blockDim.x;
                    x[u][c] = (k < Dr) ? src[k] : 0.f;
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int c = 0; c < 4; ++c) g[c] += x[u][c];
        }
        for (; j < nm; ++j) {
            const float* src = GR + (long long)match[j] * Dr;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int k = kb + c * blockDim.x;
                if (k < Dr) g[c] += src[k];
            }
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int k = kb + c * blockDim.x;
            if (k < Dr) {
                const long long i = r * Dr + k;
                float pp = rel[i], mm = m[i], vv = v[i];
                adam_update(pp, g[c], mm, vv, ad.beta1, ad.beta2, ad.step_size, ad.inv_sqrt_bc2, ad.eps, l3);
                rel[i] = pp;
                m[i] = mm;
                v[i] = vv;
            }
        }
    }
    __syncthreads();  // match / nmatch may be rewritten by the caller's next row
}

constexpr int kMaxRowPeers = 15;

struct EntityArgs {
    float* ent;
    float* m;
    float* v;
    const float* Q;
    const float* G;  // dL/ds [B,N]
    const float* GH;
    const float* GT;
    const int* offsets;
    const int* items;
    const int* items_b;
    int N;
    long long BN, B;
    long long e_lo, e_hi;  // rows this rank owns
};
// replicas of the table on the other ranks (peer mappings): the updated row is stored there as well, so that the
// broadcast of the owned rows rides on the Adam epilogue instead of following it.  A separate, LAST kernel parameter
// that is an empty struct in the single-GPU instantiation: with the pointer table inside EntityArgs the parameter
// block grew by 128 bytes and ptxas' different allocation for the very same loop cost 13 % (0.240 -> 0.269 ms).
struct EntityPeers {
    int npeers;
    float* peer[kMaxRowPeers];
    // NVSwitch multicast mapping of the table (symmetric memory): ONE multimem.st per chunk reaches every replica -- the
    // SM issues a seventh of the remote stores of the per-peer loop at 8 GPUs.  NULL: per-peer stores.
    float* mc;
};
struct NoPeers {};

// One use of the row by a negative slot: gc += dL/ds * ds/dc, with q of the using positive.
template <int MODEL, int CH, bool HEAD>
__device__ __forceinline__ void entity_use(float (&gc)[cand_planes_of(MODEL)][CH][4],
                                           const float (&c)[cand_planes_of(MODEL)][CH][4],
                                           const float (&q)[model_planes(MODEL)][CH][4], float g, const CoordConsts& kc) {
    constexpr int QP = model_planes(MODEL), CP = cand_planes_of(MODEL);
    if constexpr (MODEL == KGE_ROTATE) {
        // two coordinates per instruction (FFMA2 / FMUL2).  rsqrt(|v|^2 + tiny): at v = 0 the product with v is 0,
        // which is the d|v|/dv = 0 torch uses there; tiny = 1e-37 is below half an ulp of any |v|^2 > 2e-30
        const float2 neg1 = make_float2(-1.f, -1.f), g2 = make_float2(g, g), tiny = make_float2(1e-37f, 1e-37f);
#pragma unroll
        for (int j = 0; j < CH; ++j)
#pragma unroll
            for (int h2 = 0; h2 < 2; ++h2) {
                const float2 qr = make_float2(q[0][j][2 * h2], q[0][j][2 * h2 + 1]);
                const float2 qi = make_float2(q[1][j][2 * h2], q[1][j][2 * h2 + 1]);
                const float2 cr = make_float2(c[0][j][2 * h2], c[0][j][2 * h2 + 1]);
                const float2 ci = make_float2(c[1][j][2 * h2], c[1][j][2 * h2 + 1]);
                const float2 vr = __ffma2_rn(cr, neg1, qr), vi = __ffma2_rn(ci, neg1, qi);
                const float2 n2 = __ffma2_rn(vr, vr, __ffma2_rn(vi, vi, tiny));
                const float2 wv = __fmul2_rn(g2, make_float2(fast_rsqrt(n2.x), fast_rsqrt(n2.y)));
                const float2 a = __ffma2_rn(wv, vr, make_float2(gc[0][j][2 * h2], gc[0][j][2 * h2 + 1]));
                const float2 b = __ffma2_rn(wv, vi, make_float2(gc[1][j][2 * h2], gc[1][j][2 * h2 + 1]));
                gc[0][j][2 * h2] = a.x; gc[0][j][2 * h2 + 1] = a.y;
                gc[1][j][2 * h2] = b.x; gc[1][j][2 * h2 + 1] = b.y;
            }
    } else {
#pragma unroll
        for (int j = 0; j < CH; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float qq[3] = {q[0][j][i], (QP > 1) ? q[QP > 1 ? 1 : 0][j][i] : 0.f, 0.f};
                const float cc[2] = {c[0][j][i], (CP > 1) ? c[CP - 1][j][i] : 0.f};
                float g2[2] = {gc[0][j][i], (CP > 1) ? gc[CP - 1][j][i] : 0.f};
                dc_accumulate<MODEL, HEAD>(g2, qq, cc, g, kc);
                gc[0][j][i] = g2[0];
                if (CP > 1) gc[CP - 1][j][i] = g2[1];
            }
    }
}

// InterHT: one use of the row.  v = +-(A (b_hat + 1) - a_hat B) + M with a_hat = a / |a|, b_hat = b / |b| of THIS row;
// u = dL/dv; d a_hat, d b_hat; then the two row dots of the normalisation backward -- one block reduction per use, ONE
// barrier: `red` alternates between two buffers from use to use, so the reads of use i are ordered before the writes of
// use i + 2 by the barrier of use i + 1.
template <int CH, bool HEAD>
__device__ __forceinline__ void interht_entity_use(float (&gc)[2][CH][4], const float (&c)[2][CH][4], const float (&q)[3][CH][4],
                                                   float g, const float (&cn)[2], const float (&cpj)[2], float* red, int tid,
                                                   int T, int H) {
    float dac[CH][4], dbc[CH][4], dots[2] = {0.f, 0.f};
#pragma unroll
    for (int j = 0; j < CH; ++j)
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float ac = c[0][j][k] * cn[0], bh = c[1][j][k] * cn[1], bc = bh + 1.0f;
            const float Aq = q[0][j][k], Bv = q[1][j][k], Mq = q[2][j][k];
            const float v = HEAD ? (ac * Bv - Aq * bc + Mq) : (Aq * bc - ac * Bv + Mq);
            const float u = (4 * (tid + j * T) < H) ? -g * sgnf(v) : 0.f;
            dac[j][k] = HEAD ? (u * Bv) : (-u * Bv);
            dbc[j][k] = HEAD ? (-u * Aq) : (u * Aq);
            dots[0] = fmaf(ac, dac[j][k], dots[0]);
            dots[1] = fmaf(bh, dbc[j][k], dots[1]);
        }
    const int lane = tid & 31, warp = tid >> 5, nw = (T + 31) >> 5;
    dots[0] = warp_sum(dots[0]);
    dots[1] = warp_sum(dots[1]);
    if (lane == 0) {
        red[warp] = dots[0];
        red[32 + warp] = dots[1];
    }
    __syncthreads();
    float d0 = 0.f, d1 = 0.f;
    for (int w = 0; w < nw; ++w) {
        d0 += red[w];
        d1 += red[32 + w];
    }
    d0 *= cpj[0];  // clamped half-row (F.normalize eps): y = x / eps, no projection term
    d1 *= cpj[1];
#pragma unroll
    for (int j = 0; j < CH; ++j)
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float ac = c[0][j][k] * cn[0], bh = c[1][j][k] * cn[1];
            gc[0][j][k] += (dac[j][k] - ac * d0) * cn[0];
            gc[1][j][k] += (dbc[j][k] - bh * d1) * cn[1];
        }
}

// One CTA per entity row, each thread owning CH float4 chunks of every plane.  The uses of the row (bucket of the
// row: negatives in ascending slot order, then the positives) are staged kUseChunk at a time into shared memory as
// (dL/ds, source) pairs, so the loop over uses carries no dependent global loads: the q rows stream through a
// kRing-deep cp.async ring of thread-private shared-memory slots and are applied in bucket order.
// Measured (B200, headline config): 0.281 ms (one use at a time, dependent loads) -> 0.241 ms.
constexpr int kUseChunk = 128;
constexpr int kRing = 4;

__device__ __forceinline__ void cp_async16(uint32_t smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_dst), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// PEERS: the multi-GPU instantiation that also stores the updated row into the other ranks' replicas.  It is a template
// parameter because the mere presence of the (empty) peer loop in the Adam epilogue cost the single-GPU kernel 13 %
// (0.237 -> 0.269 ms: stores through unknown pointers between the m / v loads of consecutive chunks).
template <int MODEL, int CH, bool HEAD, int TMAX, bool PEERS>
__global__ void __launch_bounds__(TMAX, TMAX == 128 ? (MODEL == KGE_INTERHT ? 4 : 6) : 2)  // InterHT: 3 q planes, 48 KB ring
train_entity_adam_kernel(Params P, CoordConsts kc, const float* __restrict__ modulus, EntityArgs A, AdamScalars ad,
                         float l3, std::conditional_t<PEERS, EntityPeers, NoPeers> X) {
    pdl_enter();
    constexpr int QP = model_planes(MODEL), CP = cand_planes_of(MODEL);
    __shared__ float red[2 * 32];
    __shared__ float use_g[kUseChunk];
    __shared__ int use_src[kUseChunk];  // >= 0: positive b whose q row applies; < 0: -1 - (compact gradient row)
    __shared__ float red_ih[2][64];     // InterHT: the per-use reduction, double-buffered (interht_entity_use)
    int ih_parity = 0;
    const int tid = threadIdx.x, T = blockDim.x, H = P.H;
    if (MODEL == KGE_PROTATE) kc.modulus = __ldg(modulus);
    extern __shared__ float4 ring[];  // [kRing][QP][CH][T]: slot (d, p, j) of thread tid -- only tid touches it
    int koff[CH];   // element offset of this thread's chunks inside a plane
    bool kval[CH];
#pragma unroll
    for (int j = 0; j < CH; ++j) {
        koff[j] = 4 * (tid + j * T);
        kval[j] = koff[j] < H;
    }
    // ring slot of chunk k = (plane, j) of this thread: shared byte address in slot 0, and the byte offset of the
    // chunk inside a q row
    uint32_t slot_s[QP * CH];
    const char* qbase[QP * CH];
    const uint32_t slot_bytes = (uint32_t)(QP * CH) * T * 16u;
    const unsigned q_row_bytes = (unsigned)(QP * H) * (unsigned)sizeof(float);
#pragma unroll
    for (int k = 0; k < QP * CH; ++k) {
        slot_s[k] = (uint32_t)__cvta_generic_to_shared(ring + k * T + tid);
        qbase[k] = reinterpret_cast<const char*>(A.Q + ((k / CH) * H + koff[k % CH]));
    }
    {  // chunks past H are never copied: they read as zeros
        for (int d = 0; d < kRing; ++d)
#pragma unroll
            for (int k = 0; k < QP * CH; ++k)
                if (!kval[k % CH]) ring[(d * QP * CH + k) * T + tid] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    for (long long e = A.e_lo + blockIdx.x; e < A.e_hi; e += gridDim.x) {
        const int beg = A.offsets[e], end = A.offsets[e + 1];
        float* row = A.ent + e * (long long)P.De;
        float c[CP][CH][4], gc[CP][CH][4];
#pragma unroll
        for (int p = 0; p < CP; ++p)
#pragma unroll
            for (int j = 0; j < CH; ++j) {
                const int k0 = 4 * (tid + j * T);
                float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
                if (k0 < H) val = *reinterpret_cast<const float4*>(row + p * H + k0);
                c[p][j][0] = val.x; c[p][j][1] = val.y; c[p][j][2] = val.z; c[p][j][3] = val.w;
#pragma unroll
                for (int i = 0; i < 4; ++i) gc[p][j][i] = 0.f;
            }
        if (beg < end) {  // the Adam state of the row is needed after the uses: start it towards L2 now
            const char* mline = reinterpret_cast<const char*>(A.m + e * (long long)P.De);
            const char* vline = reinterpret_cast<const char*>(A.v + e * (long long)P.De);
            const int lines = (P.De * 4 + 127) / 128;
            for (int l = tid; l < 2 * lines; l += T)
                asm volatile("prefetch.global.L2 [%0];" ::"l"((l < lines ? mline : vline - (size_t)lines * 128) + (size_t)l * 128));
        }
        float cn[2] = {0.f, 0.f}, cpj[2] = {1.f, 1.f};
        if constexpr (MODEL == KGE_INTERHT) {
            if (beg < end) {  // uniform over the block
                float ss[2] = {0.f, 0.f};
#pragma unroll
                for (int j = 0; j < CH; ++j)
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        ss[0] = fmaf(c[0][j][i], c[0][j][i], ss[0]);
                        ss[1] = fmaf(c[1][j][i], c[1][j][i], ss[1]);
                    }
                block_reduce_vals<2>(ss, red);
                cn[0] = inv_norm(ss[0], P.interht_eps, cpj[0]);
                cn[1] = inv_norm(ss[1], P.interht_eps, cpj[1]);
            }
        }
        for (int base = beg; base < end; base += kUseChunk) {
            const int n = min(kUseChunk, end - base);
            __syncthreads();  // the previous chunk (or row) is consumed
            for (int i = tid; i < n; i += T) {
                const int item = A.items[base + i];
                const bool slot = item < A.BN;
                use_g[i] = slot ? A.G[item] : 0.f;
                use_src[i] = slot ? A.items_b[base + i] : (-1 - (int)(item - A.BN));
            }
            __syncthreads();
            auto add_compact = [&](int s) {  // this entity is the head / tail of a positive: its compact gradient row
                const long long j2 = -1 - s;
                const float* src = (j2 < A.B) ? (A.GH + j2 * P.De) : (A.GT + (j2 - A.B) * P.De);
                float x[CP][CH][4];
                load_coords<CP, CH>(x, src, H, tid, T, H);
#pragma unroll
                for (int p = 0; p < CP; ++p)
#pragma unroll
                    for (int j = 0; j < CH; ++j)
#pragma unroll
                        for (int k = 0; k < 4; ++k) gc[p][j][k] += x[p][j][k];
            };
            {
                // q rows travel through a kRing-deep ring of thread-private shared-memory slots filled by cp.async:
                // kRing - 1 rows are in flight while one is applied; uses stay in bucket order.  All addresses are
                // thread constants (slot_s, goff) plus one per-use term.
                auto fetch = [&](int slot, int s) {
                    const uint32_t dst = slot * slot_bytes;
#pragma unroll
                    for (int k = 0; k < QP * CH; ++k)  // one 32 x 32 + 64 bit multiply-add per address
                        if (kval[k % CH]) cp_async16(slot_s[k] + dst, qbase[k] + (unsigned long long)(unsigned)s * q_row_bytes);
                };
#pragma unroll
                for (int d = 0; d < kRing - 1; ++d) {
                    if (d < n && use_src[d] >= 0) fetch(d, use_src[d]);
                    cp_async_commit();
                }
                for (int i = 0; i < n; ++i) {
                    const int ahead = i + kRing - 1;  // its slot held use i - 1, applied in the previous iteration
                    if (ahead < n) {
                        const int sa = use_src[ahead];
                        if (sa >= 0) fetch(ahead % kRing, sa);
                    }
                    cp_async_commit();
                    cp_async_wait<kRing - 1>();  // the group of use i has landed
                    const int sc = use_src[i];
                    if (sc >= 0) {
                        float q[QP][CH][4];
                        const uint32_t src = (i % kRing) * slot_bytes;
#pragma unroll
                        for (int k = 0; k < QP * CH; ++k) lds128(q[k / CH][k % CH], slot_s[k] + src);
                        if constexpr (MODEL == KGE_INTERHT) {
                            interht_entity_use<CH, HEAD>(gc, c, q, use_g[i], cn, cpj, red_ih[ih_parity], tid, T, H);
                            ih_parity ^= 1;
                        } else {
                            entity_use<MODEL, CH, HEAD>(gc, c, q, use_g[i], kc);
                        }
                    } else {
                        add_compact(sc);
                    }
                }
                cp_async_wait<0>();
            }
        }
        // Adam on the row, in the registers that hold it (rows without uses: g = 0)
        float* mrow = A.m + e * (long long)P.De;
        float* vrow = A.v + e * (long long)P.De;
#pragma unroll
        for (int p = 0; p < CP; ++p)
#pragma unroll
            for (int j = 0; j < CH; ++j) {
                const int k0 = 4 * (tid + j * T);
                if (k0 < H) {
                    float4 mm = *reinterpret_cast<float4*>(mrow + p * H + k0);
                    float4 vv = *reinterpret_cast<float4*>(vrow + p * H + k0);
                    adam_update(c[p][j][0], gc[p][j][0], mm.x, vv.x, ad.beta1, ad.beta2, ad.step_size, ad.inv_sqrt_bc2, ad.eps, l3);
                    adam_update(c[p][j][1], gc[p][j][1], mm.y, vv.y, ad.beta1, ad.beta2, ad.step_size, ad.inv_sqrt_bc2, ad.eps, l3);
                    adam_update(c[p][j][2], gc[p][j][2], mm.z, vv.z, ad.beta1, ad.beta2, ad.step_size, ad.inv_sqrt_bc2, ad.eps, l3);
                    adam_update(c[p][j][3], gc[p][j][3], mm.w, vv.w, ad.beta1, ad.beta2, ad.step_size, ad.inv_sqrt_bc2, ad.eps, l3);
                    *reinterpret_cast<float4*>(row + p * H + k0) = make_float4(c[p][j][0], c[p][j][1], c[p][j][2], c[p][j][3]);
                    *reinterpret_cast<float4*>(mrow + p * H + k0) = mm;
                    *reinterpret_cast<float4*>(vrow + p * H + k0) = vv;
                }
            }
        if constexpr (PEERS) {
            // NVLink stores into the other ranks' replicas, after the local row / m / v traffic has been issued: the
            // updated row is still in registers
            if (X.mc != nullptr) {
                float* mrow_all = X.mc + e * (long long)P.De;
#pragma unroll
                for (int p = 0; p < CP; ++p)
#pragma unroll
                    for (int j = 0; j < CH; ++j) {
                        const int k0 = 4 * (tid + j * T);
                        if (k0 < H)
                            asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mrow_all + p * H + k0),
                                         "f"(c[p][j][0]), "f"(c[p][j][1]), "f"(c[p][j][2]), "f"(c[p][j][3])
                                         : "memory");
                    }
            } else
#pragma unroll
            for (int q = 0; q < kMaxRowPeers; ++q) {  // constant indices: the pointer table stays in the parameter bank
                if (q >= X.npeers) break;
                float* prow = X.peer[q] + e * (long long)P.De;
#pragma unroll
                for (int p = 0; p < CP; ++p)
#pragma unroll
                    for (int j = 0; j < CH; ++j) {
                        const int k0 = 4 * (tid + j * T);
                        if (k0 < H)
                            *reinterpret_cast<float4*>(prow + p * H + k0) = make_float4(c[p][j][0], c[p][j][1], c[p][j][2], c[p][j][3]);
                    }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// F4: relation rows (sum GR[b] over b ascending) + Adam; pRotatE modulus
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
train_relation_adam_kernel(int Dr, const int64_t* __restrict__ pos, int B, const float* __restrict__ GR, float* rel,
                           float* m, float* v, AdamScalars ad, float l3, int* scratch) {
    pdl_enter();
    extern __shared__ int match[];  // up to B entries
    __shared__ int nmatch;
    relation_row_update(blockIdx.x, Dr, pos, B, GR, rel, m, v, ad, l3, match, &nmatch);
}

__global__ void __launch_bounds__(256)
train_modulus_adam_kernel(const float* __restrict__ GM, int B, float* modulus, float* m, float* v, AdamScalars ad) {
    pdl_enter();
    __shared__ float red[32];
    float acc = 0.f;
    for (int b = threadIdx.x; b < B; b += blockDim.x) acc += GM[b];
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) {
        float pp = modulus[0], mm = m[0], vv = v[0];
        adam_update(pp, acc, mm, vv, ad.beta1, ad.beta2, ad.step_size, ad.inv_sqrt_bc2, ad.eps, 0.f);
        modulus[0] = pp;
        m[0] = mm;
        v[0] = vv;
    }
}

// loss terms [B][2] (pos, neg) -> loss3, fixed order (one CTA of 1024 threads: the kernel sits on the critical path of
// the step and is pure latency, so both sums go through one reduction and the loads of a thread are independent)
__global__ void __launch_bounds__(1024) loss_final2_kernel(const float* __restrict__ terms, int B, float* loss3) {
    pdl_enter();
    __shared__ float red[32][2];
    float p = 0.f, n = 0.f;
    if ((reinterpret_cast<uintptr_t>(terms) & 7) == 0) {
        const float2* t2 = reinterpret_cast<const float2*>(terms);
#pragma unroll 4
        for (int i = threadIdx.x; i < B; i += blockDim.x) {
            const float2 t = t2[i];
            p += t.x;
            n += t.y;
        }
    } else {
        for (int i = threadIdx.x; i < B; i += blockDim.x) {
            p += terms[2 * i];
            n += terms[2 * i + 1];
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        p += __shfl_xor_sync(0xffffffffu, p, o);
        n += __shfl_xor_sync(0xffffffffu, n, o);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
        red[warp][0] = p;
        red[warp][1] = n;
    }
    __syncthreads();
    if (warp == 0) {
        const int nw = blockDim.x >> 5;
        p = (lane < nw) ? red[lane][0] : 0.f;
        n = (lane < nw) ? red[lane][1] : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            p += __shfl_xor_sync(0xffffffffu, p, o);
            n += __shfl_xor_sync(0xffffffffu, n, o);
        }
        if (lane == 0) {
            loss3[0] = -p;
            loss3[1] = -n;
            loss3[2] = (-p + -n) / 2.0f;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static inline size_t fz_align(size_t x) { return (x + 255) & ~(size_t)255; }

struct FusedWs {
    float *S, *Q, *GH, *GT, *GR, *GM, *stats, *terms, *small, *l3_ws;
    int *counts, *offsets, *cursor, *items, *items_sorted, *items_b, *block_sums;
    float *PA, *PM;      // split sweep: per-item partial dL/dq and (running max, modulus partial)
    int* item_counter;
    float4* invn;        // InterHT: inverse norms of every row of the table the sweep reads
    size_t total;
};
static inline size_t invn_bytes(const Params& P) { return P.model == KGE_INTERHT ? (size_t)P.E * sizeof(float4) : 0; }

// split sweep (F1w): items per positive and rows per item
static inline int f1w_parts(int64_t N) { return (int)std::max<int64_t>(1, std::min<int64_t>(8, (N + 127) / 128)); }
static int g_sweep_variant = 1;  // 1 = warp per row (F1w) where it applies, 0 = row groups per CTA
void train_set_sweep_variant(int v) { g_sweep_variant = v ? 1 : 0; }
static inline bool att_weights(const kge_loss_opts& lo) { return !lo.detach_weights && lo.negative_adversarial_sampling; }
static inline bool f1w_applies(const Params& P, const kge_loss_opts& lo) {
    // (attached softmax weights carry a second dL/dq accumulator: row-group sweep, where a thread owns few coordinates)
    return g_sweep_variant == 1 && P.model != KGE_INTERHT && P.H % 4 == 0 && P.H <= 1024 && !att_weights(lo);
}

static FusedWs fused_carve(const Params& P, int64_t B, int64_t N, void* base) {
    FusedWs w{};
    size_t off = 0;
    auto take = [&](size_t bytes) {
        void* p = base ? (void*)((char*)base + off) : nullptr;
        off += fz_align(bytes);
        return p;
    };
    const size_t qdim = (size_t)model_planes(P.model) * P.H;
    w.S = (float*)take((size_t)B * N * sizeof(float));
    w.Q = (float*)take((size_t)B * qdim * sizeof(float));
    w.GH = (float*)take((size_t)B * P.De * sizeof(float));
    w.GT = (float*)take((size_t)B * P.De * sizeof(float));
    w.GR = (float*)take((size_t)B * P.Dr * sizeof(float));
    w.GM = (float*)take((size_t)B * sizeof(float));
    w.stats = (float*)take((size_t)2 * B * sizeof(float));
    w.terms = (float*)take((size_t)(2 * B + 8) * sizeof(float));
    w.small = (float*)take(16 * sizeof(float));
    w.l3_ws = (float*)take(1024 * sizeof(float));
    w.counts = (int*)take((size_t)(P.E + 1) * sizeof(int));
    w.offsets = (int*)take((size_t)(P.E + 1) * sizeof(int));
    w.cursor = (int*)take((size_t)(P.E + 1) * sizeof(int));
    w.items = (int*)take((size_t)(B * N + 2 * B) * sizeof(int));
    w.items_sorted = (int*)take((size_t)(B * N + 2 * B) * sizeof(int));
    w.items_b = (int*)take((size_t)(B * N + 2 * B) * sizeof(int));
    w.block_sums = (int*)take((size_t)((P.E + 1024) / 1024 + 2) * sizeof(int));
    w.PA = (float*)take((size_t)B * f1w_parts(N) * qdim * sizeof(float));
    w.PM = (float*)take((size_t)B * f1w_parts(N) * 2 * sizeof(float));
    w.item_counter = (int*)take(64);
    w.invn = (float4*)take(invn_bytes(P));
    w.total = off;
    return w;
}

size_t fused_step_workspace_bytes(const Params& P, int64_t B, int64_t N) { return fused_carve(P, B, N, nullptr).total; }
size_t train_sweep_workspace_bytes(const Params& P, int64_t B, int64_t N) {
    const size_t qdim = (size_t)model_planes(P.model) * P.H;
    return fz_align((size_t)B * f1w_parts(N) * qdim * sizeof(float)) + fz_align((size_t)B * f1w_parts(N) * 2 * sizeof(float)) + 512 +
           fz_align(invn_bytes(P));
}

bool fused_step_applicable(const Params& P, const kge_loss_opts& lo, int64_t B, int64_t N, const void* ent,
                           const void* rel) {
    if (!vec4_ok(P, ent, rel)) return false;
    if (P.H > 4096) return false;
    if (B * N + 2 * B >= 0x7fffffffLL || P.E >= 0x7fffffffLL) return false;
    return true;
}

__global__ void fused_finalize_reg_kernel(float* small, float lambda) {
    pdl_enter();
    const float reg = lambda * (small[4] + small[5]);
    small[3] = reg;
    small[2] += reg;
}

// One step = sweep (data-parallel over positives) + apply (entity-row-parallel over [e_lo, e_hi)).
// Single GPU: both phases on the local batch.  Multi rank: sweep on the local batch, the per-positive
// buffers are all-gathered by the caller, apply runs on the global batch for the rows this rank owns.
struct FusedCall {
    int phase;  // 0 both, 1 sweep only, 2 apply only
    // sweep
    const int64_t* pos;
    const int64_t* neg;
    const float* w;
    int64_t B;
    const float* wsum;
    int64_t B_total;
    // apply
    const int64_t* pos_all;
    const int64_t* neg_all;
    int64_t B_all;
    long long e_lo, e_hi;
    // per-positive buffers (sweep: written for the local batch; apply: read for the global batch)
    float *S, *Q, *GH, *GT, *GR, *GM, *stats, *terms;
    // kge_train_apply_peers: replicas of the entity table on the other ranks
    int npeers = 0;
    float* const* peers = nullptr;  // host array of npeers device pointers
    float* mc = nullptr;            // NVSwitch multicast mapping of the entity table (NULL: per-peer stores)
};

template <int MODEL, int CHW, bool HEAD, int R>
static cudaError_t launch_sweep_rows_inst(const Params& P, const CoordConsts& kc, const RowsArgs& A, cudaStream_t st) {
    auto kernel = train_sweep_rows_kernel<MODEL, CHW, HEAD, R>;
    // per CTA: the staged q planes; per warp: a region of R row slots (128-byte aligned) and R mbarriers
    const uint32_t slot_bytes = ((uint32_t)P.De * 4u + 127u) & ~127u;
    const size_t smem = (size_t)model_planes(MODEL) * CHW * 32 * sizeof(float4) +
                        4 * (size_t)rows_warp_region(slot_bytes, R, model_planes(MODEL), CHW) + 4 * R * 8;
    static size_t grid_for = 0;
    static int grid = 0;  // persistent: every CTA the device can hold
    if (grid_for != smem) {
        cudaError_t e;
        if (smem > 40 * 1024 &&
            (e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess)
            return e;
        int per_sm = 0;
        if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, 128, smem)) != cudaSuccess) return e;
        grid = std::max(1, per_sm) * sm_count();
        grid_for = smem;
    }
    const int items = A.B * A.parts;
    return launch_pdl(kernel, dim3(std::min(grid, items)), dim3(128), smem, st, P, kc, A);
}

template <int MODEL>
static cudaError_t launch_sweep_rows(const Params& P, const CoordConsts& kc, bool head, const RowsArgs& R, cudaStream_t st) {
    if constexpr (MODEL == KGE_INTERHT) {
        return cudaErrorInvalidValue;
    } else {
        // R = 2 row slots per warp: with q_b staged once per CTA, 8 + 4 * 16 kB per CTA still keeps 3 CTAs (12 warps) on
        // an SM at H = 1000
#define KGE_ROWS(CHW)                                                                                                   \
    return head ? launch_sweep_rows_inst<MODEL, CHW, true, 2>(P, kc, R, st)                                             \
                : launch_sweep_rows_inst<MODEL, CHW, false, 2>(P, kc, R, st)
        if (P.H <= 256) { KGE_ROWS(2); }
        if (P.H <= 512) { KGE_ROWS(4); }
        KGE_ROWS(8);
#undef KGE_ROWS
    }
}

template <int MODEL, int ECH, bool HEAD, int TMAX, bool PEERS>
static cudaError_t launch_entity_inst(const Params& P, const CoordConsts& kc, const float* modulus, const EntityArgs& A,
                                      const EntityPeers& X, const AdamScalars& ad, float l3, int ethreads, cudaStream_t st) {
    cudaError_t e;
    auto kernel = train_entity_adam_kernel<MODEL, ECH, HEAD, TMAX, PEERS>;
    const size_t ring_bytes = (size_t)kRing * model_planes(MODEL) * ECH * ethreads * sizeof(float4);
    static size_t allowed = 0;
    if (ring_bytes > 40 * 1024 && allowed < ring_bytes) {  // static shared memory counts towards the 48 KB default
        if ((e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ring_bytes)) != cudaSuccess) return e;
        allowed = ring_bytes;
    }
    // persistent grid: as many CTAs as are resident at once, rows dealt round-robin
    static int resident_for = -1, resident_cached = 0;  // the occupancy query is a slow host call: once per shape
    if (resident_for != ethreads) {
        if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident_cached, kernel, ethreads, ring_bytes)) != cudaSuccess) return e;
        resident_for = ethreads;
    }
    const int blocks = (int)std::min<long long>(A.e_hi - A.e_lo, (long long)std::max(1, resident_cached) * sm_count());
    if constexpr (PEERS) return launch_pdl(kernel, dim3(blocks), dim3(ethreads), ring_bytes, st, P, kc, modulus, A, ad, l3, X);
    else return launch_pdl(kernel, dim3(blocks), dim3(ethreads), ring_bytes, st, P, kc, modulus, A, ad, l3, NoPeers{});
}

template <int MODEL, int ECH>
static cudaError_t launch_entity(const Params& P, const CoordConsts& kc, int head, const float* modulus, const EntityArgs& A,
                                 const EntityPeers& X, const AdamScalars& ad, float l3, cudaStream_t st) {
    const int ethreads = std::max(32, (((P.H / 4 + ECH - 1) / ECH + 31) / 32) * 32);
#define KGE_ENT(HD, TM)                                                                                       \
    ((X.npeers > 0 || X.mc != nullptr) ? launch_entity_inst<MODEL, ECH, HD, TM, true>(P, kc, modulus, A, X, ad, l3, ethreads, st)   \
                  : launch_entity_inst<MODEL, ECH, HD, TM, false>(P, kc, modulus, A, X, ad, l3, ethreads, st))
    if (ethreads <= 128) return head ? KGE_ENT(true, 128) : KGE_ENT(false, 128);
    return head ? KGE_ENT(true, 256) : KGE_ENT(false, 256);
#undef KGE_ENT
}

template <int MODEL, int CH, int RPI>
static cudaError_t fused_launch(const Params& P, int mode, float* ent, float* rel, float* modulus, float* m_ent,
                                float* v_ent, float* m_rel, float* v_rel, float* m_mod, float* v_mod,
                                const FusedCall& fc, int64_t N, const kge_loss_opts& lo, const AdamScalars& ad,
                                const FusedWs& ws, cudaStream_t st) {
    cudaError_t e;
    const int head = (mode == KGE_MODE_HEAD_BATCH) ? 1 : 0;
    CoordConsts kc{P.gamma, P.phase_div, P.inv_phase_div, 0.f};
    const int chunks = (P.H / 4 + CH - 1) / CH;
    const int threads = std::max(32, ((chunks + 31) / 32) * 32);
    if (fc.phase != 2) {
        ProfScope ps(ST_SWEEP, st);
        if (fc.phase == 0 && (e = cudaMemsetAsync(ws.counts, 0, (size_t)(P.E + 1) * sizeof(int), st)) != cudaSuccess) return e;
        if (fc.phase == 0 && (e = cudaMemsetAsync(ws.small, 0, 16 * sizeof(float), st)) != cudaSuccess) return e;
        SweepArgs A{ent, rel, modulus, fc.pos, fc.neg, fc.w, (int)fc.B, (int)N, lo.negative_adversarial_sampling,
                    lo.uni_weight, lo.adversarial_temperature, fc.S, fc.Q, fc.GH, fc.GT, fc.GR, fc.GM, fc.stats,
                    fc.terms, fc.phase == 0 ? ws.counts : nullptr, fc.wsum, (int)fc.B_total};
        if constexpr (MODEL == KGE_INTERHT) {
            if (ws.invn == nullptr) return cudaErrorInvalidValue;
            const int nb = (int)std::min<long long>((P.E + 7) / 8, 16LL * sm_count());
            if ((e = launch_pdl(interht_inv_norm_kernel, dim3(nb), dim3(256), 0, st, (const float*)ent, (long long)P.E, P.H,
                                P.interht_eps, ws.invn)) != cudaSuccess)
                return e;
            A.invn = ws.invn;
            ps.launches(2);
        }
        if (fc.phase != 2 && ws.PA != nullptr && f1w_applies(P, lo)) {
            if constexpr (MODEL != KGE_INTERHT) {
                // split sweep: stage Q | one warp per candidate row | merge + positive score + chain rule
                ps.launches(3);
                A.part = 1;
                A.item_counter = ws.item_counter;
                if ((e = launch_pdl(train_sweep_kernel<MODEL, CH, RPI, 1>, dim3((unsigned)fc.B), dim3(threads), 0, st, P, kc, head, A)) != cudaSuccess) return e;
                const int parts = f1w_parts(N);
                RowsArgs R{ent, fc.neg, fc.Q, modulus, fc.S, ws.PA, ws.PM, ws.item_counter, (int)fc.B, (int)N, parts,
                           (int)((N + parts - 1) / parts), lo.negative_adversarial_sampling, lo.adversarial_temperature};
                if ((e = launch_sweep_rows<MODEL>(P, kc, head != 0, R, st)) != cudaSuccess) return e;
                A.part = 2;
                A.parts = parts;
                A.PA = ws.PA;
                A.PM = ws.PM;
                if ((e = launch_pdl(train_sweep_kernel<MODEL, CH, RPI, 2>, dim3((unsigned)fc.B), dim3(threads), 0, st, P, kc, head, A)) != cudaSuccess) return e;
            }
        } else {
            if (att_weights(lo))
                e = launch_pdl(train_sweep_kernel<MODEL, CH, RPI, 0, true>, dim3((unsigned)fc.B), dim3(threads), 0, st, P, kc, head, A);
            else
                e = launch_pdl(train_sweep_kernel<MODEL, CH, RPI, 0, false>, dim3((unsigned)fc.B), dim3(threads), 0, st, P, kc, head, A);
            if (e != cudaSuccess) return e;
        }
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if (fc.phase == 1) return cudaSuccess;
    }
    const int64_t B = fc.B_all;
    const long long BN = B * N;
    const float l3 = lo.regularization;
    {
        {
            ProfScope ps(ST_BUCKETS, st);
            // apply-only calls build the buckets of the rows [e_lo, e_hi) alone; a full step counted every use in F1
            const long long b_lo = (fc.phase == 2) ? fc.e_lo : 0, b_hi = (fc.phase == 2) ? fc.e_hi : P.E;
            // (ws.small was cleared before the sweep when this call runs one: the sweep's last kernel is then followed
            //  directly by the scan, kernel to kernel)
            if (fc.phase == 2 && (e = cudaMemsetAsync(ws.small, 0, 16 * sizeof(float), st)) != cudaSuccess) return e;
            if (fc.phase == 2) {
                if ((e = cudaMemsetAsync(ws.counts, 0, (size_t)(P.E + 1) * sizeof(int), st)) != cudaSuccess) return e;
                if ((e = launch_histogram(mode, fc.pos_all, fc.neg_all, BN, B, ws.counts, b_lo, b_hi, st)) != cudaSuccess) return e;
            }
            if ((e = launch_scan_fill(mode, fc.pos_all, fc.neg_all, BN, B, P.E, ws.counts, ws.offsets, ws.cursor, ws.items,
                                      ws.block_sums, b_lo, b_hi, st)) != cudaSuccess)
                return e;
            const int sb = (int)std::min<long long>((P.E + 7) / 8, 16LL * sm_count());
            if ((e = launch_pdl(bucket_sort_kernel, dim3(sb), dim3(256), 0, st, ws.offsets, ws.items, ws.items_sorted, ws.items_b, P.E, (int)N)) != cudaSuccess) return e;
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
        }
        {
            ProfScope ps(ST_DSCORE, st);
            // (S already holds dL/ds: the sweep's last pass writes it per positive)
            if ((e = launch_pdl(loss_final2_kernel, dim3(1), dim3(1024), 0, st, fc.terms, (int)B, ws.small)) != cudaSuccess) return e;
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
        }
        if (lo.regularization != 0.f) {
            ProfScope ps(ST_REG, st);
            if ((e = launch_l3_sum_cubes(ent, (int64_t)P.E * P.De, ws.small + 4, ws.l3_ws, st)) != cudaSuccess) return e;
            if ((e = launch_l3_sum_cubes(rel, (int64_t)P.R * P.Dr, ws.small + 5, ws.l3_ws, st)) != cudaSuccess) return e;
            fused_finalize_reg_kernel<<<1, 1, 0, st>>>(ws.small, l3);
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
        }
    }
    if (fc.e_hi > fc.e_lo) {
        ProfScope ps(ST_ENTITY, st);
        EntityArgs A{ent, m_ent, v_ent, fc.Q, fc.S, fc.GH, fc.GT, ws.offsets, ws.items_sorted, ws.items_b, (int)N, BN, B,
                     fc.e_lo, fc.e_hi};
        EntityPeers X{fc.npeers, {}, fc.mc};
        for (int q = 0; q < fc.npeers; ++q) X.peer[q] = fc.peers[q];
        // wider per-thread tiles than the sweep: the per-use overhead is paid once per ECH float4 chunks
        if ((e = launch_entity<MODEL, (CH == 1) ? 2 : CH>(P, kc, head, modulus, A, X, ad, l3, st)) != cudaSuccess) return e;
    }
    {
        // relation rows and the modulus are replicated: every rank applies the same fixed-order update.  (Folding the
        // relation rows into the tail of the entity kernel was tried: the extra code slowed that kernel by more than
        // the launch it saves.)
        ProfScope ps(ST_RELATION, st);
        ps.launches(MODEL == KGE_PROTATE ? 2 : 1);
        if ((e = launch_pdl(train_relation_adam_kernel, dim3((unsigned)P.R), dim3(256),
                            (size_t)relation_scratch_ints((int)B) * sizeof(int), st, P.Dr, fc.pos_all, (int)B, fc.GR, rel,
                            m_rel, v_rel, ad, l3, nullptr)) != cudaSuccess)
            return e;
        if (MODEL == KGE_PROTATE) train_modulus_adam_kernel<<<1, 256, 0, st>>>(fc.GM, (int)B, modulus, m_mod, v_mod, ad);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
    }
    return cudaSuccess;
}

template <int MODEL>
static cudaError_t fused_dispatch_ch(const Params& P, int mode, float* ent, float* rel, float* modulus, float* m_ent,
                                     float* v_ent, float* m_rel, float* v_rel, float* m_mod, float* v_mod,
                                     const FusedCall& fc, int64_t N, const kge_loss_opts& lo, const AdamScalars& ad,
                                     const FusedWs& ws, cudaStream_t st) {
    constexpr bool IH = (MODEL == KGE_INTERHT);
#define KGE_GO(CH, RPI) \
    return fused_launch<MODEL, CH, RPI>(P, mode, ent, rel, modulus, m_ent, v_ent, m_rel, v_rel, m_mod, v_mod, fc, N, lo, ad, ws, st)
    if (P.H <= 1024) { KGE_GO(1, (IH ? 4 : 8)); }
    if (P.H <= 2048) { KGE_GO(2, (IH ? 2 : 4)); }
    KGE_GO(4, (IH ? 1 : 2));
#undef KGE_GO
}

static cudaError_t fused_dispatch(const Params& P, int mode, float* ent, float* rel, float* modulus, float* m_ent,
                                  float* v_ent, float* m_rel, float* v_rel, float* m_mod, float* v_mod,
                                  const FusedCall& fc, int64_t N, const kge_loss_opts& lo, const AdamScalars& ad,
                                  const FusedWs& ws, cudaStream_t st) {
#define KGE_CASE(M) \
    case M: return fused_dispatch_ch<M>(P, mode, ent, rel, modulus, m_ent, v_ent, m_rel, v_rel, m_mod, v_mod, fc, N, lo, ad, ws, st)
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

// out_small: device float[16]: [0..2] loss3, [3] reg
cudaError_t launch_train_step_fused(const Params& P, int mode, float* ent, float* rel, float* modulus, float* m_ent,
                                    float* v_ent, float* m_rel, float* v_rel, float* m_mod, float* v_mod,
                                    const int64_t* pos, const int64_t* neg, const float* w, int64_t B, int64_t N,
                                    const kge_loss_opts& lo, const kge_adam_opts& ao, void* wsp, size_t ws_bytes,
                                    float** out_small, cudaStream_t st) {
    if (ws_bytes < fused_step_workspace_bytes(P, B, N)) return cudaErrorInvalidValue;
    const FusedWs ws = fused_carve(P, B, N, wsp);
    *out_small = ws.small;
    FusedCall fc{0, pos, neg, w, B, nullptr, B, pos, neg, B, 0, P.E, ws.S, ws.Q, ws.GH, ws.GT, ws.GR, ws.GM, ws.stats, ws.terms};
    return fused_dispatch(P, mode, ent, rel, modulus, m_ent, v_ent, m_rel, v_rel, m_mod, v_mod, fc, N, lo, adam_scalars(ao),
                          ws, st);
}

// multi-rank pieces (kge_train_sweep / kge_train_apply)
cudaError_t launch_train_sweep(const Params& P, int mode, const float* ent, const float* rel, const float* modulus,
                               const int64_t* pos, const int64_t* neg, const float* w, int64_t B, int64_t N,
                               const kge_loss_opts& lo, const float* wsum, int64_t B_total, const kge_step_buffers& o,
                               void* wsp, size_t ws_bytes, cudaStream_t st) {
    // the sweep alone needs a workspace only for the split (warp-per-row) variant: per-item partials + item counter
    FusedWs ws{};
    if (wsp != nullptr && ws_bytes >= train_sweep_workspace_bytes(P, B, N)) {
        char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(wsp) + 255) & ~(uintptr_t)255);
        const size_t qdim = (size_t)model_planes(P.model) * P.H;
        ws.PA = (float*)base;
        ws.PM = (float*)(base + fz_align((size_t)B * f1w_parts(N) * qdim * sizeof(float)));
        ws.item_counter = (int*)((char*)ws.PM + fz_align((size_t)B * f1w_parts(N) * 2 * sizeof(float)));
        ws.invn = (float4*)((char*)ws.item_counter + 256);
    } else if (P.model == KGE_INTERHT) {
        return cudaErrorInvalidValue;  // InterHT keeps its inverse-norm table in the sweep workspace
    }
    FusedCall fc{1, pos, neg, w, B, wsum, B_total, nullptr, nullptr, 0, 0, 0, o.S, o.Q, o.GH, o.GT, o.GR, o.GM, o.stats, o.terms};
    kge_adam_opts ao{0.f, 0.9f, 0.999f, 1e-8f, 1, 0};
    return fused_dispatch(P, mode, const_cast<float*>(ent), const_cast<float*>(rel), const_cast<float*>(modulus), nullptr,
                          nullptr, nullptr, nullptr, nullptr, nullptr, fc, N, lo, adam_scalars(ao), ws, st);
}

cudaError_t launch_train_apply(const Params& P, int mode, float* ent, float* rel, float* modulus, float* m_ent,
                               float* v_ent, float* m_rel, float* v_rel, float* m_mod, float* v_mod,
                               const int64_t* pos_all, const int64_t* neg_all, int64_t B_all, int64_t N,
                               const kge_loss_opts& lo, const kge_adam_opts& ao, const kge_step_buffers& in, int64_t e_lo,
                               int64_t e_hi, int npeers, float* const* peers, float* mc, void* wsp,
                               size_t ws_bytes, float** out_small, cudaStream_t st) {
    if (npeers < 0 || npeers > kMaxRowPeers) return cudaErrorInvalidValue;
    if (ws_bytes < fused_step_workspace_bytes(P, B_all, N)) return cudaErrorInvalidValue;
    const FusedWs ws = fused_carve(P, B_all, N, wsp);
    *out_small = ws.small;
    FusedCall fc{2, nullptr, nullptr, nullptr, 0, nullptr, B_all, pos_all, neg_all, B_all, e_lo, e_hi,
                 in.S, in.Q, in.GH, in.GT, in.GR, in.GM, in.stats, in.terms, npeers, peers};
    fc.mc = mc;
    return fused_dispatch(P, mode, ent, rel, modulus, m_ent, v_ent, m_rel, v_rel, m_mod, v_mod, fc, N, lo, adam_scalars(ao),
                          ws, st);
}

}  // namespace kge
