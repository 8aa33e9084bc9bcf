// Per-coordinate score math for the fused train step (kge_train_fused.cu).
//
// In the fused kernels a thread OWNS a fixed set of coordinates k of every row it touches, so all
// per-positive state (staged planes q, dL/dq accumulators) lives in registers and nothing but the
// row-level reductions crosses threads.  Each functor below is the scalar restatement of the same
// maths as kge_score.cuh (/root/reference/tensorflow_codes/model.py:207-224).
//
//   QP = staged planes per coordinate, CP = candidate planes per coordinate
//   stage_coord   (h, r, t coordinate values)            -> q[QP]
//   fwd_coord     (q, c) -> term of the row sum; rewrites c in place into what bwd needs ("direction")
//   dq_coord      (q, direction)                         -> ds/dq[QP]          (for dL/dq += g * ds/dq)
//   dc_coord      (q, c raw)                             -> ds/dc[CP]          (entity-major pass)
//   chain_coord   (dL/dq, h, r, t coordinate values)     -> dL/dh, dL/dr, dL/dt contributions
#pragma once
#include "kge_common.cuh"
#include "kge_score.cuh"

namespace kge {

__host__ __device__ constexpr int cand_planes_of(int model) {
    return (model == KGE_COMPLEX || model == KGE_ROTATE || model == KGE_INTERHT) ? 2 : 1;
}
__host__ __device__ constexpr int rel_planes_of(int model) {  // relation values read per coordinate
    return (model == KGE_COMPLEX) ? 2 : 1;                      // InterHT: only the middle third
}

struct CoordConsts {
    float gamma, phase_div, inv_phase_div, modulus;
};

// relation plane offset: InterHT reads the middle third (model.py:209)
template <int MODEL>
__device__ __forceinline__ int rel_offset(int H, int plane) {
    if constexpr (MODEL == KGE_INTERHT) return H;
    else return plane * H;
}

// ---- staging ----------------------------------------------------------------------------------
// e[] = fixed entity coordinate planes (h for tail-batch/single, t for head-batch), rv[] = relation planes.
// InterHT: nrm[0..1] = 1/|a_f|, 1/|b_f| of the fixed entity.
template <int MODEL>
__device__ __forceinline__ void stage_coord(float (&q)[3], const float (&e)[2], const float (&rv)[2], bool head,
                                            const CoordConsts& c, const float (&nrm)[2], float& cr_out, float& sr_out) {
    q[0] = q[1] = q[2] = 0.f;
    if constexpr (MODEL == KGE_TRANSE) {
        q[0] = head ? (rv[0] - e[0]) : (e[0] + rv[0]);
    } else if constexpr (MODEL == KGE_DISTMULT) {
        q[0] = head ? (rv[0] * e[0]) : (e[0] * rv[0]);
    } else if constexpr (MODEL == KGE_COMPLEX) {
        if (head) {
            q[0] = rv[0] * e[0] + rv[1] * e[1];
            q[1] = rv[0] * e[1] - rv[1] * e[0];
        } else {
            q[0] = e[0] * rv[0] - e[1] * rv[1];
            q[1] = e[0] * rv[1] + e[1] * rv[0];
        }
    } else if constexpr (MODEL == KGE_ROTATE) {
        float sr, cr;
        sincosf(rv[0] / c.phase_div, &sr, &cr);
        cr_out = cr;
        sr_out = sr;
        if (head) {
            q[0] = cr * e[0] + sr * e[1];
            q[1] = cr * e[1] - sr * e[0];
        } else {
            q[0] = e[0] * cr - e[1] * sr;
            q[1] = e[0] * sr + e[1] * cr;
        }
    } else if constexpr (MODEL == KGE_PROTATE) {
        const float pr = rv[0] / c.phase_div;
        q[0] = head ? (pr - e[0] / c.phase_div) : (e[0] / c.phase_div + pr);
    } else {  // InterHT
        q[0] = e[0] * nrm[0];
        q[1] = e[1] * nrm[1] + 1.0f;
        q[2] = rv[0];
    }
}

// ---- forward term; c[] is rewritten to the "direction" the backward needs ----------------------
// InterHT: cn[0..1] = 1/|a_c|, 1/|b_c| of this candidate; sg receives sign(v).
template <int MODEL, bool HEAD>
__device__ __forceinline__ float fwd_coord(const float (&q)[3], float (&c)[2], const CoordConsts& k, const float (&cn)[2],
                                           float& sg) {
    if constexpr (MODEL == KGE_TRANSE) {
        const float v = HEAD ? (c[0] + q[0]) : (q[0] - c[0]);
        c[0] = sgnf(v);
        return fabsf(v);
    } else if constexpr (MODEL == KGE_DISTMULT) {
        return q[0] * c[0];
    } else if constexpr (MODEL == KGE_COMPLEX) {
        return q[0] * c[0] + q[1] * c[1];
    } else if constexpr (MODEL == KGE_ROTATE) {
        const float vr = q[0] - c[0], vi = q[1] - c[1];
        const float n2 = vr * vr + vi * vi;
        const float rs = (n2 > 0.f) ? fast_rsqrt(n2) : 0.f;  // d|v|/dv = 0 at v = 0 (torch)
        c[0] = vr * rs;
        c[1] = vi * rs;
        return n2 * rs;
    } else if constexpr (MODEL == KGE_PROTATE) {
        const float pc = c[0] * k.inv_phase_div;
        float sv, cv;
        sincosf(HEAD ? (pc + q[0]) : (q[0] - pc), &sv, &cv);
        c[0] = sgnf(sv) * cv;
        return fabsf(sv);
    } else {  // InterHT
        const float ac = c[0] * cn[0], bc = c[1] * cn[1] + 1.0f;
        const float v = HEAD ? (ac * q[1] - q[0] * bc + q[2]) : (q[0] * bc - ac * q[1] + q[2]);
        sg = sgnf(v);
        c[0] = ac;
        c[1] = bc;
        return fabsf(v);
    }
}

// score from the row sum
template <int MODEL>
__device__ __forceinline__ float score_from_sum(float tot, const CoordConsts& k) {
    if constexpr (MODEL == KGE_DISTMULT || MODEL == KGE_COMPLEX) return tot;
    else if constexpr (MODEL == KGE_PROTATE) return k.gamma - tot * k.modulus;
    else return k.gamma - tot;
}

// ---- acc[QP] += wgt * ds/dq, from the direction left by fwd_coord --------------------------------
template <int MODEL, bool HEAD>
__device__ __forceinline__ void dq_accumulate(float (&acc)[3], const float (&d)[2], float sg, float wgt,
                                              const CoordConsts& k) {
    if constexpr (MODEL == KGE_TRANSE) {
        acc[0] -= wgt * d[0];
    } else if constexpr (MODEL == KGE_DISTMULT) {
        acc[0] += wgt * d[0];
    } else if constexpr (MODEL == KGE_COMPLEX) {
        acc[0] += wgt * d[0];
        acc[1] += wgt * d[1];
    } else if constexpr (MODEL == KGE_ROTATE) {
        acc[0] -= wgt * d[0];
        acc[1] -= wgt * d[1];
    } else if constexpr (MODEL == KGE_PROTATE) {
        acc[0] -= wgt * k.modulus * d[0];
    } else {  // InterHT: d = (a_hat_c, b_hat_c + 1); s = gamma - sum |v|
        const float u = -wgt * sg;  // wgt * ds/dv
        acc[0] += HEAD ? (-u * d[1]) : (u * d[1]);
        acc[1] += HEAD ? (u * d[0]) : (-u * d[0]);
        acc[2] += u;
    }
}

// ---- entity-major pass: gc[CP] += g * ds/dc for one use (q of the using positive, raw row c) -------
// InterHT is handled separately (needs row-level dots).
template <int MODEL, bool HEAD>
__device__ __forceinline__ void dc_accumulate(float (&gc)[2], const float (&q)[3], const float (&c)[2], float g,
                                              const CoordConsts& k) {
    if constexpr (MODEL == KGE_TRANSE) {
        const float v = HEAD ? (c[0] + q[0]) : (q[0] - c[0]);
        const float u = -g * sgnf(v);
        gc[0] += HEAD ? u : -u;
    } else if constexpr (MODEL == KGE_DISTMULT) {
        gc[0] += g * q[0];
    } else if constexpr (MODEL == KGE_COMPLEX) {
        gc[0] += g * q[0];
        gc[1] += g * q[1];
    } else if constexpr (MODEL == KGE_ROTATE) {
        const float vr = q[0] - c[0], vi = q[1] - c[1];
        const float n2 = vr * vr + vi * vi;
        const float wv = (n2 > 0.f) ? g * fast_rsqrt(n2) : 0.f;
        gc[0] += wv * vr;
        gc[1] += wv * vi;
    } else if constexpr (MODEL == KGE_PROTATE) {
        const float pc = c[0] * k.inv_phase_div;
        float sv, cv;
        sincosf(HEAD ? (pc + q[0]) : (q[0] - pc), &sv, &cv);
        const float u = -g * k.modulus * sgnf(sv) * cv;
        gc[0] += (HEAD ? u : -u) * k.inv_phase_div;
    }
}

// ---- chain rule from dL/dq to the positive's rows, one coordinate ------------------------------------
// gq: dL/dq planes.  e[]: fixed-entity coordinate (t for head-batch, h otherwise); rv: relation planes;
// cr/sr: cos/sin of the RotatE phase.  Outputs: ge[2] = dL/d(fixed entity), gr[2] = dL/d(relation planes).
// InterHT: dotA = <A, dA>, dotB = <B_hat, dB> over the row; nrm = 1/|a_f|, 1/|b_f|.
template <int MODEL>
__device__ __forceinline__ void chain_coord(float (&ge)[2], float (&gr)[2], const float (&gq)[3], const float (&q)[3],
                                            const float (&e)[2], const float (&rv)[2], bool head, const CoordConsts& c,
                                            float cr, float sr, float dotA, float dotB, const float (&nrm)[2]) {
    ge[0] = ge[1] = gr[0] = gr[1] = 0.f;
    if constexpr (MODEL == KGE_TRANSE) {
        gr[0] = gq[0];
        ge[0] = head ? -gq[0] : gq[0];
    } else if constexpr (MODEL == KGE_DISTMULT) {
        gr[0] = gq[0] * e[0];
        ge[0] = gq[0] * rv[0];
    } else if constexpr (MODEL == KGE_COMPLEX) {
        const float ur = gq[0], ui = gq[1];
        if (head) {  // q_re = rr*tr + ri*ti ; q_im = rr*ti - ri*tr
            gr[0] = ur * e[0] + ui * e[1];
            gr[1] = ur * e[1] - ui * e[0];
            ge[0] = ur * rv[0] - ui * rv[1];
            ge[1] = ur * rv[1] + ui * rv[0];
        } else {  // q_re = hr*rr - hi*ri ; q_im = hr*ri + hi*rr
            ge[0] = ur * rv[0] + ui * rv[1];
            ge[1] = -ur * rv[1] + ui * rv[0];
            gr[0] = ur * e[0] + ui * e[1];
            gr[1] = -ur * e[1] + ui * e[0];
        }
    } else if constexpr (MODEL == KGE_ROTATE) {
        const float ur = gq[0], ui = gq[1];
        float gcr, gsr;
        if (head) {  // q_re = cr*tr + sr*ti ; q_im = cr*ti - sr*tr
            gcr = ur * e[0] + ui * e[1];
            gsr = ur * e[1] - ui * e[0];
            ge[0] = ur * cr - ui * sr;
            ge[1] = ur * sr + ui * cr;
        } else {  // q_re = hr*cr - hi*sr ; q_im = hr*sr + hi*cr
            gcr = ur * e[0] + ui * e[1];
            gsr = -ur * e[1] + ui * e[0];
            ge[0] = ur * cr + ui * sr;
            ge[1] = -ur * sr + ui * cr;
        }
        gr[0] = (gsr * cr - gcr * sr) / c.phase_div;
    } else if constexpr (MODEL == KGE_PROTATE) {
        const float u = gq[0] / c.phase_div;
        gr[0] = u;
        ge[0] = head ? -u : u;
    } else {  // InterHT
        ge[0] = (gq[0] - q[0] * dotA) * nrm[0];
        ge[1] = (gq[1] - (q[1] - 1.0f) * dotB) * nrm[1];
        gr[0] = gq[2];
    }
}

}  // namespace kge
