// Per-model score math shared by the forward, backward and train-step kernels.
//
// Vocabulary: for one positive triple (h, r, t) and a batch mode, everything that does not depend
// on the candidate entity is folded once into "staged" planes q (H floats each, shared memory):
//
//   model      planes   tail-batch / single (candidate c = tail)     head-batch (candidate c = head)
//   TransE     1        q = h + r            v = q - c               q = r - t        v = c + q
//   DistMult   1        q = h * r            s = sum q*c             q = r * t
//   ComplEx    2        q = h o r                                    q = conj(r) o t  s = Re<q, c>
//   RotatE     2        q = h o e^{i phase}  v = q - c               q = e^{-i phase} o t
//   pRotatE    1        q = ph_h + ph_r      v = q - ph_c            q = ph_r - ph_t  v = ph_c + q
//   InterHT    3        (A, B+1, mid) of the fixed entity; candidate normalised on the fly
//
// Reference math: SURVEY.md 8(a) U3-U7 (absent PyTorch submodule) and
// /root/reference/tensorflow_codes/model.py:207-224 (InterHT).
#pragma once
#include "kge_common.cuh"

namespace kge {

__host__ __device__ constexpr int model_planes(int model) {
    return (model == KGE_COMPLEX || model == KGE_ROTATE) ? 2 : (model == KGE_INTERHT ? 3 : 1);
}

// ------------------------------------------------------------------------------------------
// Staging: block-wide.  q: planes*H floats (shared).  aux: 4 floats (shared) -- InterHT norms.
// red: >= 32 floats scratch (shared).  Ends with __syncthreads().
// ------------------------------------------------------------------------------------------
template <int MODEL>
__device__ __forceinline__ void stage_positive(float* q, float* aux, float* red, const Params& P,
                                               const float* __restrict__ ent, const float* __restrict__ rel,
                                               long long h, long long r, long long t, bool head) {
    const int H = P.H;
    const float* hrow = ent + h * (long long)P.De;
    const float* trow = ent + t * (long long)P.De;
    const float* rrow = rel + r * (long long)P.Dr;
    if constexpr (MODEL == KGE_TRANSE) {
        for (int k = threadIdx.x; k < H; k += blockDim.x)
            q[k] = head ? (__ldg(rrow + k) - __ldg(trow + k)) : (__ldg(hrow + k) + __ldg(rrow + k));
    } else if constexpr (MODEL == KGE_DISTMULT) {
        for (int k = threadIdx.x; k < H; k += blockDim.x)
            q[k] = head ? (__ldg(rrow + k) * __ldg(trow + k)) : (__ldg(hrow + k) * __ldg(rrow + k));
    } else if constexpr (MODEL == KGE_COMPLEX) {
        for (int k = threadIdx.x; k < H; k += blockDim.x) {
            const float rr = __ldg(rrow + k), ri = __ldg(rrow + H + k);
            if (head) {
                const float tr = __ldg(trow + k), ti = __ldg(trow + H + k);
                q[k] = rr * tr + ri * ti;
                q[H + k] = rr * ti - ri * tr;
            } else {
                const float hr = __ldg(hrow + k), hi = __ldg(hrow + H + k);
                q[k] = hr * rr - hi * ri;
                q[H + k] = hr * ri + hi * rr;
            }
        }
    } else if constexpr (MODEL == KGE_ROTATE) {
        for (int k = threadIdx.x; k < H; k += blockDim.x) {
            float sr, cr;
            sincosf(__ldg(rrow + k) / P.phase_div, &sr, &cr);
            if (head) {
                const float tr = __ldg(trow + k), ti = __ldg(trow + H + k);
                q[k] = cr * tr + sr * ti;
                q[H + k] = cr * ti - sr * tr;
            } else {
                const float hr = __ldg(hrow + k), hi = __ldg(hrow + H + k);
                q[k] = hr * cr - hi * sr;
                q[H + k] = hr * sr + hi * cr;
            }
        }
    } else if constexpr (MODEL == KGE_PROTATE) {
        for (int k = threadIdx.x; k < H; k += blockDim.x) {
            const float pr = __ldg(rrow + k) / P.phase_div;
            q[k] = head ? (pr - __ldg(trow + k) / P.phase_div) : (__ldg(hrow + k) / P.phase_div + pr);
        }
    } else {  // InterHT: fixed entity f = tail (head-batch) or head (tail-batch / single)
        const float* frow = head ? trow : hrow;
        float sa = 0.f, sb = 0.f;
        for (int k = threadIdx.x; k < H; k += blockDim.x) {
            const float a = __ldg(frow + k), b = __ldg(frow + H + k);
            sa += a * a;
            sb += b * b;
        }
        sa = block_sum(sa, red);
        sb = block_sum(sb, red);
        float na = sqrtf(sa), nb = sqrtf(sb);
        if (P.interht_eps > 0.f) {
            na = fmaxf(na, P.interht_eps);
            nb = fmaxf(nb, P.interht_eps);
        }
        if (threadIdx.x == 0) {
            aux[0] = na;
            aux[1] = nb;
        }
        for (int k = threadIdx.x; k < H; k += blockDim.x) {
            q[k] = __ldg(frow + k) / na;
            q[H + k] = __ldg(frow + H + k) / nb + 1.0f;  // u = 1 (model.py:94-95, 219-220)
            q[2 * H + k] = __ldg(rrow + H + k);          // middle third (model.py:209)
        }
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------------------
// Forward: one warp scores one candidate row against the staged planes.  Result on all lanes.
// ------------------------------------------------------------------------------------------
template <int MODEL, int VEC, bool HEAD>
__device__ __forceinline__ float score_row(const float* q, const float* __restrict__ row, const Params& P, int lane,
                                           float modulus) {
    const int H = P.H;
    float acc = 0.f;
    if constexpr (MODEL == KGE_TRANSE) {
#pragma unroll 4
        for (int k = lane * VEC; k < H; k += 32 * VEC) {
            float qq[VEC], c[VEC];
            ld_global<VEC>(c, row + k);
            ld_shared<VEC>(qq, q + k);
#pragma unroll
            for (int i = 0; i < VEC; ++i) acc += fabsf(HEAD ? (c[i] + qq[i]) : (qq[i] - c[i]));
        }
        return P.gamma - warp_sum(acc);
    } else if constexpr (MODEL == KGE_DISTMULT) {
#pragma unroll 4
        for (int k = lane * VEC; k < H; k += 32 * VEC) {
            float qq[VEC], c[VEC];
            ld_global<VEC>(c, row + k);
            ld_shared<VEC>(qq, q + k);
#pragma unroll
            for (int i = 0; i < VEC; ++i) acc = fmaf(HEAD ? c[i] : qq[i], HEAD ? qq[i] : c[i], acc);
        }
        return warp_sum(acc);
    } else if constexpr (MODEL == KGE_COMPLEX) {
#pragma unroll 4
        for (int k = lane * VEC; k < H; k += 32 * VEC) {
            float qr[VEC], qi[VEC], cr[VEC], ci[VEC];
            ld_global<VEC>(cr, row + k);
            ld_global<VEC>(ci, row + H + k);
            ld_shared<VEC>(qr, q + k);
            ld_shared<VEC>(qi, q + H + k);
#pragma unroll
            for (int i = 0; i < VEC; ++i) acc += qr[i] * cr[i] + qi[i] * ci[i];
        }
        return warp_sum(acc);
    } else if constexpr (MODEL == KGE_ROTATE) {
#pragma unroll 4
        for (int k = lane * VEC; k < H; k += 32 * VEC) {
            float qr[VEC], qi[VEC], cr[VEC], ci[VEC];
            ld_global<VEC>(cr, row + k);
            ld_global<VEC>(ci, row + H + k);
            ld_shared<VEC>(qr, q + k);
            ld_shared<VEC>(qi, q + H + k);
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                const float vr = qr[i] - cr[i], vi = qi[i] - ci[i];
                acc += fast_sqrt(vr * vr + vi * vi);
            }
        }
        return P.gamma - warp_sum(acc);
    } else if constexpr (MODEL == KGE_PROTATE) {
#pragma unroll 2
        for (int k = lane * VEC; k < H; k += 32 * VEC) {
            float qq[VEC], c[VEC];
            ld_global<VEC>(c, row + k);
            ld_shared<VEC>(qq, q + k);
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                const float pc = c[i] * P.inv_phase_div;
                acc += fabsf(sinf(HEAD ? (pc + qq[i]) : (qq[i] - pc)));
            }
        }
        return P.gamma - warp_sum(acc) * modulus;
    } else {  // InterHT
        float sa = 0.f, sb = 0.f;
#pragma unroll 4
        for (int k = lane * VEC; k < H; k += 32 * VEC) {
            float a[VEC], b[VEC];
            ld_global<VEC>(a, row + k);
            ld_global<VEC>(b, row + H + k);
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                sa = fmaf(a[i], a[i], sa);
                sb = fmaf(b[i], b[i], sb);
            }
        }
        float na = sqrtf(warp_sum(sa)), nb = sqrtf(warp_sum(sb));
        if (P.interht_eps > 0.f) {
            na = fmaxf(na, P.interht_eps);
            nb = fmaxf(nb, P.interht_eps);
        }
        const float ina = 1.0f / na, inb = 1.0f / nb;
#pragma unroll 2
        for (int k = lane * VEC; k < H; k += 32 * VEC) {
            float a[VEC], b[VEC], A[VEC], Bv[VEC], M[VEC];
            ld_global<VEC>(a, row + k);
            ld_global<VEC>(b, row + H + k);
            ld_shared<VEC>(A, q + k);
            ld_shared<VEC>(Bv, q + H + k);
            ld_shared<VEC>(M, q + 2 * H + k);
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                const float ac = a[i] * ina, bc = b[i] * inb + 1.0f;
                const float v = HEAD ? (ac * Bv[i] - A[i] * bc + M[i]) : (A[i] * bc - ac * Bv[i] + M[i]);
                acc += fabsf(v);
            }
        }
        return P.gamma - warp_sum(acc);
    }
}

// ------------------------------------------------------------------------------------------
// Backward of one (staged planes, candidate row) pair given g = dL/dscore.
//   grad_row : dL/d(candidate row) is ADDED here (EMIT_ATOMIC: red.global.add on the dense grad
//              table; EMIT_PLAIN: plain += on lane-owned storage; EMIT_NONE: not produced)
//   gq       : lane-owned accumulator planes (same layout as q) receiving dL/dq, or unused
//              when !WANT_GQ
//   gmod     : per-lane partial of dL/dmodulus (pRotatE)
// ------------------------------------------------------------------------------------------
enum { EMIT_NONE = 0, EMIT_ATOMIC = 1, EMIT_PLAIN = 2 };

template <int MODEL, int VEC, bool HEAD, int EMIT, bool WANT_GQ>
__device__ __forceinline__ void backward_row(const float* q, const float* __restrict__ row, float g, const Params& P,
                                             int lane, float modulus, float* gq, float* grad_row, float& gmod) {
    const int H = P.H;
    auto emit = [&](float* dst, const float(&x)[VEC]) {
        if constexpr (EMIT == EMIT_ATOMIC) red_add<VEC>(dst, x);
        else if constexpr (EMIT == EMIT_PLAIN) add_ptr<VEC>(dst, x);
    };
    if constexpr (MODEL == KGE_TRANSE) {
#pragma unroll 2
        for (int k = lane * VEC; k < H; k += 32 * VEC) {
            float qq[VEC], c[VEC], gc[VEC], u[VEC];
            ld_global<VEC>(c, row + k);
            ld_shared<VEC>(qq, q + k);
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                const float v = HEAD ? (c[i] + qq[i]) : (qq[i] - c[i]);
                u[i] = -g * sgnf(v);
                gc[i] = HEAD ? u[i] : -u[i];
            }
            emit(grad_row + k, gc);
            if constexpr (WANT_GQ) add_ptr<VEC>(gq + k, u);
        }
    } else if constexpr (MODEL == KGE_DISTMULT) {
#pragma unroll 2
        for (int k = lane * VEC; k < H; k += 32 * VEC) {
            float qq[VEC], c[VEC], gc[VEC], u[VEC];
            ld_global<VEC>(c, row + k);
            ld_shared<VEC>(qq, q + k);
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                gc[i] = g * qq[i];
                u[i] = g * c[i];
            }
            emit(grad_row + k, gc);
            if constexpr (WANT_GQ) add_ptr<VEC>(gq + k, u);
        }
    } else if constexpr (MODEL == KGE_COMPLEX) {
#pragma unroll 2
        for (int k = lane * VEC; k < H; k += 32 * VEC) {
            float qr[VEC], qi[VEC], cr[VEC], ci[VEC], a[VEC], b[VEC];
            ld_global<VEC>(cr, row + k);
            ld_global<VEC>(ci, row + H + k);
            ld_shared<VEC>(qr, q + k);
            ld_shared<VEC>(qi, q + H + k);
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                a[i] = g * qr[i];
                b[i] = g * qi[i];
            }
            emit(grad_row + k, a);
            emit(grad_row + H + k, b);
            if constexpr (WANT_GQ) {
#pragma unroll
                for (int i = 0; i < VEC; ++i) {
                    a[i] = g * cr[i];
                    b[i] = g * ci[i];
                }
                add_ptr<VEC>(gq + k, a);
                add_ptr<VEC>(gq + H + k, b);
            }
        }
    } else if constexpr (MODEL == KGE_ROTATE) {
#pragma unroll 2
        for (int k = lane * VEC; k < H; k += 32 * VEC) {
            float qr[VEC], qi[VEC], cr[VEC], ci[VEC], ur[VEC], ui[VEC];
            ld_global<VEC>(cr, row + k);
            ld_global<VEC>(ci, row + H + k);
            ld_shared<VEC>(qr, q + k);
            ld_shared<VEC>(qi, q + H + k);
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                const float vr = qr[i] - cr[i], vi = qi[i] - ci[i];
                const float n2 = vr * vr + vi * vi;
                const float w = (n2 > 0.f) ? g * fast_rsqrt(n2) : 0.f;  // d|v|/dv = 0 at v = 0 (torch)
                ur[i] = w * vr;  // = +dL/dc ; dL/dq = -ur
                ui[i] = w * vi;
            }
            emit(grad_row + k, ur);
            emit(grad_row + H + k, ui);
            if constexpr (WANT_GQ) {
#pragma unroll
                for (int i = 0; i < VEC; ++i) {
                    ur[i] = -ur[i];
                    ui[i] = -ui[i];
                }
                add_ptr<VEC>(gq + k, ur);
                add_ptr<VEC>(gq + H + k, ui);
            }
        }
    } else if constexpr (MODEL == KGE_PROTATE) {
        float sabs = 0.f;
        for (int k = lane * VEC; k < H; k += 32 * VEC) {
            float qq[VEC], c[VEC], gc[VEC], u[VEC];
            ld_global<VEC>(c, row + k);
            ld_shared<VEC>(qq, q + k);
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                const float pc = c[i] * P.inv_phase_div;
                const float v = HEAD ? (pc + qq[i]) : (qq[i] - pc);
                float sv, cv;
                sincosf(v, &sv, &cv);
                sabs += fabsf(sv);
                u[i] = -g * modulus * sgnf(sv) * cv;  // dL/dv
                gc[i] = (HEAD ? u[i] : -u[i]) * P.inv_phase_div;
            }
            emit(grad_row + k, gc);
            if constexpr (WANT_GQ) add_ptr<VEC>(gq + k, u);
        }
        gmod += -g * sabs;
    } else {  // InterHT
        float sa = 0.f, sb = 0.f;
#pragma unroll 4
        for (int k = lane * VEC; k < H; k += 32 * VEC) {
            float a[VEC], b[VEC];
            ld_global<VEC>(a, row + k);
            ld_global<VEC>(b, row + H + k);
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                sa = fmaf(a[i], a[i], sa);
                sb = fmaf(b[i], b[i], sb);
            }
        }
        float na = sqrtf(warp_sum(sa)), nb = sqrtf(warp_sum(sb));
        bool clamp_a = false, clamp_b = false;
        if (P.interht_eps > 0.f) {
            clamp_a = na < P.interht_eps;
            clamp_b = nb < P.interht_eps;
            na = fmaxf(na, P.interht_eps);
            nb = fmaxf(nb, P.interht_eps);
        }
        const float ina = 1.0f / na, inb = 1.0f / nb;
        // pass 2: dots <a_hat, d a_hat>, <b_hat, d b_hat>, and the staged-side gradient
        float da = 0.f, db = 0.f;
        for (int k = lane * VEC; k < H; k += 32 * VEC) {
            float a[VEC], b[VEC], A[VEC], Bv[VEC], M[VEC], gA[VEC], gB[VEC], gM[VEC];
            ld_global<VEC>(a, row + k);
            ld_global<VEC>(b, row + H + k);
            ld_shared<VEC>(A, q + k);
            ld_shared<VEC>(Bv, q + H + k);
            ld_shared<VEC>(M, q + 2 * H + k);
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                const float ac = a[i] * ina, bh = b[i] * inb, bc = bh + 1.0f;
                const float v = HEAD ? (ac * Bv[i] - A[i] * bc + M[i]) : (A[i] * bc - ac * Bv[i] + M[i]);
                const float u = -g * sgnf(v);  // dL/dv
                const float d_ac = HEAD ? (u * Bv[i]) : (-u * Bv[i]);
                const float d_bc = HEAD ? (-u * A[i]) : (u * A[i]);
                da = fmaf(ac, d_ac, da);
                db = fmaf(bh, d_bc, db);
                gA[i] = HEAD ? (-u * bc) : (u * bc);
                gB[i] = HEAD ? (u * ac) : (-u * ac);
                gM[i] = u;
            }
            if constexpr (WANT_GQ) {
                add_ptr<VEC>(gq + k, gA);
                add_ptr<VEC>(gq + H + k, gB);
                add_ptr<VEC>(gq + 2 * H + k, gM);
            }
        }
        da = clamp_a ? 0.f : warp_sum(da);
        db = clamp_b ? 0.f : warp_sum(db);
        // pass 3: d a = (d a_hat - a_hat <a_hat, d a_hat>) / |a|
        for (int k = lane * VEC; k < H; k += 32 * VEC) {
            float a[VEC], b[VEC], A[VEC], Bv[VEC], M[VEC], ga[VEC], gb[VEC];
            ld_global<VEC>(a, row + k);
            ld_global<VEC>(b, row + H + k);
            ld_shared<VEC>(A, q + k);
            ld_shared<VEC>(Bv, q + H + k);
            ld_shared<VEC>(M, q + 2 * H + k);
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                const float ac = a[i] * ina, bh = b[i] * inb, bc = bh + 1.0f;
                const float v = HEAD ? (ac * Bv[i] - A[i] * bc + M[i]) : (A[i] * bc - ac * Bv[i] + M[i]);
                const float u = -g * sgnf(v);
                const float d_ac = HEAD ? (u * Bv[i]) : (-u * Bv[i]);
                const float d_bc = HEAD ? (-u * A[i]) : (u * A[i]);
                ga[i] = (d_ac - ac * da) * ina;
                gb[i] = (d_bc - bh * db) * inb;
            }
            emit(grad_row + k, ga);
            emit(grad_row + H + k, gb);
        }
    }
}

// ------------------------------------------------------------------------------------------
// Chain rule from dL/dq (gq, planes*H floats in shared memory, already summed over the block)
// to the three rows of the positive triple.  Block-wide.  `out_h/out_r/out_t` are row pointers
// (De / Dr / De floats) that receive += (ATOMIC: red.global.add; else plain store-add).
// ------------------------------------------------------------------------------------------
template <int MODEL, bool ATOMIC>
__device__ __forceinline__ void chain_positive(const float* gq, const float* q, const float* aux, float* red,
                                               const Params& P, const float* __restrict__ ent,
                                               const float* __restrict__ rel, long long h, long long r, long long t,
                                               bool head, float* out_h, float* out_r, float* out_t) {
    const int H = P.H;
    const float* hrow = ent + h * (long long)P.De;
    const float* trow = ent + t * (long long)P.De;
    const float* rrow = rel + r * (long long)P.Dr;
    auto add1 = [&](float* dst, float x) {
        float xx[1] = {x};
        if constexpr (ATOMIC) red_add<1>(dst, xx);
        else *dst += x;
    };
    if constexpr (MODEL == KGE_TRANSE) {
        for (int k = threadIdx.x; k < H; k += blockDim.x) {
            const float u = gq[k];
            add1(out_r + k, u);
            if (head) add1(out_t + k, -u);
            else add1(out_h + k, u);
        }
    } else if constexpr (MODEL == KGE_DISTMULT) {
        for (int k = threadIdx.x; k < H; k += blockDim.x) {
            const float u = gq[k], rr = __ldg(rrow + k);
            if (head) {
                add1(out_r + k, u * __ldg(trow + k));
                add1(out_t + k, u * rr);
            } else {
                add1(out_r + k, u * __ldg(hrow + k));
                add1(out_h + k, u * rr);
            }
        }
    } else if constexpr (MODEL == KGE_COMPLEX) {
        for (int k = threadIdx.x; k < H; k += blockDim.x) {
            const float ur = gq[k], ui = gq[H + k];
            const float rr = __ldg(rrow + k), ri = __ldg(rrow + H + k);
            if (head) {  // q_re = rr*tr + ri*ti ; q_im = rr*ti - ri*tr
                const float tr = __ldg(trow + k), ti = __ldg(trow + H + k);
                add1(out_r + k, ur * tr + ui * ti);
                add1(out_r + H + k, ur * ti - ui * tr);
                add1(out_t + k, ur * rr - ui * ri);
                add1(out_t + H + k, ur * ri + ui * rr);
            } else {  // q_re = hr*rr - hi*ri ; q_im = hr*ri + hi*rr
                const float hr = __ldg(hrow + k), hi = __ldg(hrow + H + k);
                add1(out_h + k, ur * rr + ui * ri);
                add1(out_h + H + k, -ur * ri + ui * rr);
                add1(out_r + k, ur * hr + ui * hi);
                add1(out_r + H + k, -ur * hi + ui * hr);
            }
        }
    } else if constexpr (MODEL == KGE_ROTATE) {
        for (int k = threadIdx.x; k < H; k += blockDim.x) {
            const float ur = gq[k], ui = gq[H + k];
            float sr, cr;
            sincosf(__ldg(rrow + k) / P.phase_div, &sr, &cr);
            float gcr, gsr;
            if (head) {  // q_re = cr*tr + sr*ti ; q_im = cr*ti - sr*tr
                const float tr = __ldg(trow + k), ti = __ldg(trow + H + k);
                gcr = ur * tr + ui * ti;
                gsr = ur * ti - ui * tr;
                add1(out_t + k, ur * cr - ui * sr);
                add1(out_t + H + k, ur * sr + ui * cr);
            } else {  // q_re = hr*cr - hi*sr ; q_im = hr*sr + hi*cr
                const float hr = __ldg(hrow + k), hi = __ldg(hrow + H + k);
                gcr = ur * hr + ui * hi;
                gsr = -ur * hi + ui * hr;
                add1(out_h + k, ur * cr + ui * sr);
                add1(out_h + H + k, -ur * sr + ui * cr);
            }
            add1(out_r + k, (gsr * cr - gcr * sr) / P.phase_div);
        }
    } else if constexpr (MODEL == KGE_PROTATE) {
        for (int k = threadIdx.x; k < H; k += blockDim.x) {
            const float u = gq[k] / P.phase_div;
            add1(out_r + k, u);
            if (head) add1(out_t + k, -u);
            else add1(out_h + k, u);
        }
    } else {  // InterHT: planes (dA, dB, dM); fixed entity f = tail (head-batch) or head
        float da = 0.f, db = 0.f;
        for (int k = threadIdx.x; k < H; k += blockDim.x) {
            da = fmaf(q[k], gq[k], da);
            db = fmaf(q[H + k] - 1.0f, gq[H + k], db);
        }
        da = block_sum(da, red);
        db = block_sum(db, red);
        const float na = aux[0], nb = aux[1];
        if (P.interht_eps > 0.f) {  // clamped norm: y = x / eps, no projection term
            const float* frow = head ? trow : hrow;
            float sa = 0.f, sb = 0.f;
            for (int k = threadIdx.x; k < H; k += blockDim.x) {
                const float a = __ldg(frow + k), b = __ldg(frow + H + k);
                sa += a * a;
                sb += b * b;
            }
            sa = block_sum(sa, red);
            sb = block_sum(sb, red);
            if (sqrtf(sa) < P.interht_eps) da = 0.f;
            if (sqrtf(sb) < P.interht_eps) db = 0.f;
        }
        float* out_f = head ? out_t : out_h;
        for (int k = threadIdx.x; k < H; k += blockDim.x) {
            add1(out_f + k, (gq[k] - q[k] * da) / na);
            add1(out_f + H + k, (gq[H + k] - (q[H + k] - 1.0f) * db) / nb);
            add1(out_r + H + k, gq[2 * H + k]);
        }
    }
}

}  // namespace kge
