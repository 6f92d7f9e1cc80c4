// eval_common.cuh — record layouts, work-list decoding and pair arithmetic shared by the forward and
// backward evaluation kernels.
//
// Math (reference: cuda_sampler/forward.cu:168-275, backward.cu:108-416), per (Gaussian g, sample s):
//   X = mu_g - x_s  (torus-wrapped, forward.cu:147-158)
//   a = Conic_g X ; q = X.a ; alpha = exp(-q/2)
//   value  v*alpha ; gradient v*alpha*a ; Hessian v*alpha*(a a^T - Conic) ;
//   third  v*alpha*(c_ij a_k + c_ik a_j + c_jk a_i - a_i a_j a_k)
// Backward with upstream weights G (per channel, symmetric components summed):
//   S = sum_w G_w f_w ; p = dS/da ; E = explicit dS/dConic
//   dv += alpha*S ; dmu += w*(Conic p - S a) ; dConic_ij += w*(-S X_i X_j/2 + sym(p X^T)_ij + E_ij)
// which expands to the term-by-term closed forms of backward.cu (verified symbolically against
// autograd for all four functors; the single exception — the 1-D third-derivative conic gradient,
// backward.cu:326-329, which is not the derivative of its forward — is reproduced as written).
#pragma once
#include "common.cuh"

namespace pigs {

constexpr float KH = -0.72134752044448170368f;  // -0.5 * log2(e)
constexpr int EV_THREADS = 256;

// ---- compile-time sizes ------------------------------------------------------------------------
// unique (symmetric) components per channel of each function
// PIGS_FN_LAPL as a KERNEL mask bit (2-D, one channel; masks 16 and 19 = value + gradient + Laplacian): the trace of
// the Hessian as ONE accumulator — a0^2 + a1^2 - tr(Conic) — instead of three (16 instead of 19 forward lane-ops per
// pair, 31.5 instead of 36.5 backward).  It takes the Hessian's place in every component order; never both.
__host__ __device__ constexpr int nsym(int D, unsigned bit)
{
    return bit == PIGS_FN_VALUE  ? 1
           : bit == PIGS_FN_DERIV ? D
           : bit == PIGS_FN_HESS  ? D * (D + 1) / 2
           : bit == PIGS_FN_LAPL  ? 1
                                  : (D == 2 ? 4 : 1);
}
__host__ __device__ constexpr int nacc(int D, unsigned mask)
{
    return ((mask & PIGS_FN_VALUE) ? nsym(D, PIGS_FN_VALUE) : 0) + ((mask & PIGS_FN_DERIV) ? nsym(D, PIGS_FN_DERIV) : 0) +
           ((mask & PIGS_FN_HESS) ? nsym(D, PIGS_FN_HESS) : 0) + ((mask & PIGS_FN_LAPL) ? nsym(D, PIGS_FN_LAPL) : 0) +
           ((mask & PIGS_FN_THIRD) ? nsym(D, PIGS_FN_THIRD) : 0);
}
__host__ __device__ constexpr int nconic(int D) { return D * (D + 1) / 2; }
// Gaussian record: D means, D(D+1)/2 conics, C values, 1 wrap-shift code (SHIFT instances), padded to float4s
__host__ __device__ constexpr int grec_floats(int D, int C) { return D + nconic(D) + C + 1; }
__host__ __device__ constexpr int grec_v4(int D, int C) { return (grec_floats(D, C) + 3) / 4; }
// forward record: D means, K = KH*Conic (so that log2(alpha) = X.(K X) needs no scaling multiply), M = KH^2*Conic
// (the constant of the scaled Hessian factor a'a'^T - M with a' = K X = KH*a), C values, 1 wrap-shift code.
// Accumulating in the scaled variables makes the gradient come out times KH, the Hessian times KH^2 and the
// third derivative times KH^3 — constants applied once per output element in the epilogue.
__host__ __device__ constexpr int frec_floats(int D, int C) { return D + 2 * nconic(D) + C + 1; }
__host__ __device__ constexpr int frec_v4(int D, int C) { return (frec_floats(D, C) + 3) / 4; }
// sample record (backward): D coords + C * nacc upstream weights, padded to float4s
__host__ __device__ constexpr int srec_floats(int D, int C, unsigned mask) { return D + C * nacc(D, mask); }
__host__ __device__ constexpr int srec_v4(int D, int C, unsigned mask) { return (srec_floats(D, C, mask) + 3) / 4; }
// samples per thread (forward) / Gaussians per thread (backward)
#ifndef PIGS_FWD_KS16
#define PIGS_FWD_KS16 0  // experiment: 16 samples per thread (one CTA per SM, 255 registers) for the fused 2-D kernel
#endif
__host__ __device__ constexpr int fwd_ks(int D, int C, unsigned mask)
{
    return (PIGS_FWD_KS16 && D == 2 && C == 1 && mask == 7u) ? 16 : ((D + C * nacc(D, mask)) <= 8 ? 8 : 4);
}
__host__ __device__ constexpr int bwd_kg(int D, int C, unsigned mask) { return 4; }

// ---- work lists --------------------------------------------------------------------------------
// Built on the device by worklist_kernel (pack.cu); decoded by every CTA of the persistent
// evaluation kernels.  Every (tile, Gaussian) instance is in one of three classes (pack.cu flag_kernel):
//   seg 0 FAST    : no torus wrap possible for any sample of the tile, safely positive-definite conic
//   seg 1 SHIFT   : per dimension the reference's wrap is the SAME constant (0, -2 or +2) for every sample
//                   of the tile, so X = fl(mu - x) + shift reproduces forward.cu:149-157 exactly
//   seg 2 GENERAL : anything else — per-pair wrap (fmod) and the power > 0 skip
constexpr int EV_FAST = 0, EV_SHIFT = 1, EV_GENERAL = 2, EV_NSEG = 3;
struct TileSeg {
    uint32_t g_begin, g_count;  // packed-record range
    uint32_t s_begin, s_count;  // sorted-sample range
    uint32_t split;             // number of splits of the streamed dimension
    uint32_t item_begin;        // exclusive prefix of items over tiles
};

struct WorkHeader {
    uint32_t counter[4];  // dynamic work counters per class
    uint32_t items[4];    // total items per class
    unsigned long long pairs;
    uint32_t neg_zero;  // bit pattern of -0.0f: a run-time addend ptxas cannot fold (pk2_once)
    uint32_t pad[5];
};
static_assert(sizeof(WorkHeader) == 64, "WorkHeader is one 64-byte line");

struct WorkItem {
    uint32_t g_begin, g_count, s_begin, s_count;
};

// one thread decodes item i of segment list `segs[0..blocks)`.  chunk = size of the stationary
// dimension per item; stream_is_gauss: forward streams Gaussians (stationary samples), backward
// streams samples (stationary Gaussians).
__device__ __forceinline__ WorkItem decode_item(const TileSeg* __restrict__ segs, int blocks, uint32_t i, uint32_t chunk,
                                                bool stream_is_gauss)
{
    int lo = 0, hi = blocks;  // upper_bound on item_begin
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (segs[mid].item_begin <= i)
            lo = mid + 1;
        else
            hi = mid;
    }
    const TileSeg s = segs[lo - 1];
    const uint32_t local = i - s.item_begin;
    const uint32_t c = local / s.split, sp = local % s.split;
    WorkItem w;
    const uint32_t stat_n = stream_is_gauss ? s.s_count : s.g_count;
    const uint32_t strm_n = stream_is_gauss ? s.g_count : s.s_count;
    const uint32_t st0 = c * chunk;
    const uint32_t stc = min(chunk, stat_n - st0);
    const uint32_t b0 = (uint32_t)(((unsigned long long)strm_n * sp) / s.split);
    const uint32_t b1 = (uint32_t)(((unsigned long long)strm_n * (sp + 1)) / s.split);
    if (stream_is_gauss) {
        w.s_begin = s.s_begin + st0;
        w.s_count = stc;
        w.g_begin = s.g_begin + b0;
        w.g_count = b1 - b0;
    } else {
        w.g_begin = s.g_begin + st0;
        w.g_count = stc;
        w.s_begin = s.s_begin + b0;
        w.s_count = b1 - b0;
    }
    return w;
}

// ---- pair arithmetic ---------------------------------------------------------------------------
template <int D, int C>
struct GaussReg {
    float m[D];
    float c[nconic(D)];
    float v[C];
    float sh[D];  // SHIFT class: constant wrap per dimension (0, -2, +2); unused otherwise
};

// wrap-shift code stored (as int bits) in the record slot behind the values: 2 bits per dimension,
// 0 -> 0, 1 -> -2 (1 < X < 2), 2 -> +2 (-2 < X < -1)
__host__ __device__ __forceinline__ float shift_of_code(int code, int d)
{
    const int k = (code >> (2 * d)) & 3;
    return k == 0 ? 0.0f : (k == 1 ? -2.0f : 2.0f);
}

template <int NV>
__device__ __forceinline__ void load_v4(const float4* __restrict__ src, float* __restrict__ f)
{
#pragma unroll
    for (int i = 0; i < NV; i++) {
        const float4 t = src[i];
        f[4 * i] = t.x;
        f[4 * i + 1] = t.y;
        f[4 * i + 2] = t.z;
        f[4 * i + 3] = t.w;
    }
}

template <int D, int C>
__device__ __forceinline__ GaussReg<D, C> load_gauss(const float4* __restrict__ rec)
{
    constexpr int NV = grec_v4(D, C);
    float f[NV * 4];
    load_v4<NV>(rec, f);
    GaussReg<D, C> g;
#pragma unroll
    for (int d = 0; d < D; d++) g.m[d] = f[d];
#pragma unroll
    for (int k = 0; k < nconic(D); k++) g.c[k] = f[D + k];
#pragma unroll
    for (int ch = 0; ch < C; ch++) g.v[ch] = f[D + nconic(D) + ch];
    const int code = __float_as_int(f[D + nconic(D) + C]);
#pragma unroll
    for (int d = 0; d < D; d++) g.sh[d] = shift_of_code(code, d);
    return g;
}

// forward-kernel view of a record (layout: frec_floats)
template <int D, int C>
struct GaussFwd {
    float m[D];
    float k[nconic(D)];   // KH * conic
    float mm[nconic(D)];  // KH^2 * conic
    float v[C];
    float sh[D];
};

template <int D, int C>
__device__ __forceinline__ GaussFwd<D, C> load_gauss_fwd(const float4* __restrict__ rec)
{
    constexpr int NV = frec_v4(D, C);
    constexpr int NC = nconic(D);
    float f[NV * 4];
    load_v4<NV>(rec, f);
    GaussFwd<D, C> g;
#pragma unroll
    for (int d = 0; d < D; d++) g.m[d] = f[d];
#pragma unroll
    for (int k = 0; k < NC; k++) g.k[k] = f[D + k];
#pragma unroll
    for (int k = 0; k < NC; k++) g.mm[k] = f[D + NC + k];
#pragma unroll
    for (int ch = 0; ch < C; ch++) g.v[ch] = f[D + 2 * NC + ch];
    const int code = __float_as_int(f[D + 2 * NC + C]);
#pragma unroll
    for (int d = 0; d < D; d++) g.sh[d] = shift_of_code(code, d);
    return g;
}

// reference torus wrap, forward.cu:149-157 (period 2 hard-wired)
__device__ __forceinline__ float torus_wrap(float X)
{
    const float ax = fabsf(X);
    if (ax > 1.0f) {
        // 1 < |X| < 2: fmod(X, 2) == X exactly, so the wrap is X -/+ 2 (exact); the fmod path only beyond
        if (ax < 2.0f)
            X = X - copysignf(2.0f, X);
        else
            X = (X >= 0.0f) ? fmodf(X, 2.0f) - 2.0f : fmodf(X, 2.0f) + 2.0f;
    }
    return X;
}

// geometry of one pair: offsets X, a = Conic X, alpha.  On the GENERAL path alpha = 0 when the
// reference would skip the pair (power > 0, forward.cu:171).
template <int D, int MODE>
struct PairGeom {
    static constexpr bool GENERAL = MODE == EV_GENERAL;
    float X[D], a[D], alpha;
    template <int C>
    __device__ __forceinline__ void compute(const GaussReg<D, C>& g, const float* __restrict__ x)
    {
#pragma unroll
        for (int d = 0; d < D; d++) {
            X[d] = g.m[d] - x[d];
            if constexpr (MODE == EV_SHIFT) X[d] = __fadd_rn(X[d], g.sh[d]);  // never contracted into the subtract
            if constexpr (GENERAL) X[d] = torus_wrap(X[d]);
        }
        float q;
        if constexpr (D == 1) {
            a[0] = g.c[0] * X[0];
            q = a[0] * X[0];
        } else {
            a[0] = fmaf(g.c[1], X[1], g.c[0] * X[0]);
            a[1] = fmaf(g.c[2], X[1], g.c[1] * X[0]);
            q = fmaf(a[1], X[1], a[0] * X[0]);
        }
#ifdef PIGS_DBG_NOMUFU  // issue-rate experiment only (tools/quick_perf.py --variants): same FP32 count, no MUFU
        alpha = __fadd_rn(__fmul_rn(q, KH), 1.0f);
#else
        alpha = ex2_approx(q * KH);
#endif
        if constexpr (GENERAL) {
            // power = -q/2 > 0  <=>  q < 0 ; NaN falls through like the reference's comparison
            if (q < 0.0f) alpha = 0.0f;
        }
    }
};

// polynomial factors f_w(a, Conic) of every requested function, unique components only, order:
// [value][deriv d0,d1][hess 00,01,11 | laplacian][third 000,001,011,111]
template <int D, unsigned MASK>
__device__ __forceinline__ void pair_factors(const float* __restrict__ a, const float* __restrict__ c, float* __restrict__ f)
{
    int k = 0;
    if constexpr ((MASK & PIGS_FN_VALUE) != 0) f[k++] = 1.0f;
    if constexpr ((MASK & PIGS_FN_DERIV) != 0) {
#pragma unroll
        for (int d = 0; d < D; d++) f[k++] = a[d];
    }
    if constexpr ((MASK & PIGS_FN_HESS) != 0) {
        if constexpr (D == 1) {
            f[k++] = fmaf(a[0], a[0], -c[0]);
        } else {
            f[k++] = fmaf(a[0], a[0], -c[0]);
            f[k++] = fmaf(a[0], a[1], -c[1]);
            f[k++] = fmaf(a[1], a[1], -c[2]);
        }
    }
    if constexpr ((MASK & PIGS_FN_LAPL) != 0) {
        static_assert(!((MASK & PIGS_FN_LAPL) && (MASK & PIGS_FN_HESS)), "Laplacian takes the Hessian's place");
        if constexpr (D == 1)
            f[k++] = fmaf(a[0], a[0], -c[0]);
        else
            f[k++] = fmaf(a[0], a[0], fmaf(a[1], a[1], -(c[0] + c[2])));  // trace of a a^T - C
    }
    if constexpr ((MASK & PIGS_FN_THIRD) != 0) {
        if constexpr (D == 1) {
            f[k++] = a[0] * fmaf(-a[0], a[0], 3.0f * c[0]);  // 3 c a - a^3 (forward.cu:246)
        } else {
            const float a1 = a[0], a2 = a[1], c0 = c[0], c1 = c[1], c2 = c[2];
            const float a11 = a1 * a1, a22 = a2 * a2, tc1 = 2.0f * c1;
            f[k++] = a1 * fmaf(3.0f, c0, -a11);      // xxx: 3 c0 a1 - a1^3
            f[k++] = fmaf(a2, c0 - a11, tc1 * a1);   // xxy: 2 c1 a1 - a1^2 a2 + c0 a2
            f[k++] = fmaf(a1, c2 - a22, tc1 * a2);   // xyy: 2 c1 a2 - a1 a2^2 + c2 a1
            f[k++] = a2 * fmaf(3.0f, c2, -a22);      // yyy: 3 c2 a2 - a2^3
        }
    }
}

// (S, p, E) of one channel from its upstream weights G (same component order as pair_factors)
template <int D, unsigned MASK>
__device__ __forceinline__ void upstream_spe(const float* __restrict__ G, const float* __restrict__ f,
                                             const float* __restrict__ a, const float* __restrict__ c, float& S,
                                             float* __restrict__ p, float* __restrict__ E)
{
    int k = 0;
    S = 0.0f;
#pragma unroll
    for (int d = 0; d < D; d++) p[d] = 0.0f;
#pragma unroll
    for (int i = 0; i < nconic(D); i++) E[i] = 0.0f;
    if constexpr ((MASK & PIGS_FN_VALUE) != 0) S = G[k++];
    if constexpr ((MASK & PIGS_FN_DERIV) != 0) {
#pragma unroll
        for (int d = 0; d < D; d++) {
            S = fmaf(G[k], f[k], S);
            p[d] = G[k];
            k++;
        }
    }
    if constexpr ((MASK & PIGS_FN_HESS) != 0) {
        if constexpr (D == 1) {
            S = fmaf(G[k], f[k], S);
            p[0] = fmaf(2.0f * G[k], a[0], p[0]);
            E[0] = -G[k];
            k++;
        } else {
            const float g00 = G[k], g01 = G[k + 1], g11 = G[k + 2];
            S = fmaf(g00, f[k], S);
            S = fmaf(g01, f[k + 1], S);
            S = fmaf(g11, f[k + 2], S);
            p[0] = fmaf(2.0f * g00, a[0], fmaf(g01, a[1], p[0]));
            p[1] = fmaf(2.0f * g11, a[1], fmaf(g01, a[0], p[1]));
            E[0] = -g00;
            E[1] = -g01;
            E[2] = -g11;
            k += 3;
        }
    }
    if constexpr ((MASK & PIGS_FN_LAPL) != 0) {
        const float gl = G[k];  // = G00 = G11, G01 = 0
        S = fmaf(gl, f[k], S);
#pragma unroll
        for (int d = 0; d < D; d++) p[d] = fmaf(2.0f * gl, a[d], p[d]);
        E[0] = -gl;
        if constexpr (D == 2) E[2] = -gl;
        k++;
    }
    if constexpr ((MASK & PIGS_FN_THIRD) != 0) {
        if constexpr (D == 1) {
            const float T = G[k];
            S = fmaf(T, f[k], S);
            p[0] = fmaf(T, 3.0f * fmaf(-a[0], a[0], c[0]), p[0]);
            E[0] = fmaf(3.0f * a[0], T, E[0]);
            k++;
        } else {
            const float T0 = G[k], T1 = G[k + 1], T2 = G[k + 2], T3 = G[k + 3];
            const float a1 = a[0], a2 = a[1], c0 = c[0], c1 = c[1], c2 = c[2];
            S = fmaf(T0, f[k], S);
            S = fmaf(T1, f[k + 1], S);
            S = fmaf(T2, f[k + 2], S);
            S = fmaf(T3, f[k + 3], S);
            const float h00 = fmaf(-a1, a1, c0), h01 = fmaf(-a1, a2, c1), h11 = fmaf(-a2, a2, c2);
            p[0] += 3.0f * T0 * h00 + 2.0f * T1 * h01 + T2 * h11;
            p[1] += T1 * h00 + 2.0f * T2 * h01 + 3.0f * T3 * h11;
            E[0] += 3.0f * a1 * T0 + a2 * T1;
            E[1] += 2.0f * (a1 * T1 + a2 * T2);
            E[2] += a1 * T2 + 3.0f * a2 * T3;
            k += 4;
        }
    }
}

}  // namespace pigs
