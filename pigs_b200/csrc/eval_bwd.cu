// eval_bwd.cu — fused backward of value / gradient / Hessian (/ third derivative): gradients with
// respect to means, conics and values (replaces the four backward renderCUDA<F> instantiations of
// cuda_sampler/backward.cu:27-499).
//
// Design (B200, sm_100a): the mirror image of the forward kernel.  GAUSSIAN-stationary: every thread
// owns KG Gaussians of a tile (parameters + 6..10 gradient accumulators in registers) and streams the
// tile's samples — coordinates and upstream gradients packed as float4 records by pack.cu — through
// double-buffered shared memory with broadcast LDS.128.  There is NO atomic and NO shuffle in the
// inner loop; each work item ends with one RED.ADD per Gaussian parameter.  (The reference issues
// 6+ same-address atomics per PAIR with a block-uniform target, backward.cu:131-147 — fully
// serialised.)  Work items = (tile, chunk of 256*KG Gaussians, slice of the tile's samples), pulled
// from the device-side work list by a persistent grid.
#include "layouts.cuh"

namespace pigs {

constexpr int BWD_SB = 128;  // sample records per shared-memory batch

#ifndef PIGS_BWD_OPMAJOR
#define PIGS_BWD_OPMAJOR 1
#endif

// One sample record against the KG Gaussians a thread owns — 2-D, one channel, any subset of value / gradient /
// Hessian.  Same arithmetic as the generic path below, but written OP-MAJOR over the KG Gaussians wherever an
// operand of the sample record (warp-uniform) takes part, so that it sits in the same operand slot of consecutive
// instructions (reuse-cache hits: an FP32 instruction needing three register-file reads costs two dispatch cycles
// on B200, tools/mb_issue.cu), and GAUSSIAN-major where the shared operand is the pair's alpha.
template <unsigned MASK, int MODE, int KG>
__device__ __forceinline__ void bwd_sample_2d(const GaussReg<2, 1> (&g)[KG], const float* __restrict__ r,
                                              float (&dm)[KG][2], float (&dc)[KG][3], float (&dv)[KG][1])
{
    constexpr bool HV = (MASK & PIGS_FN_VALUE) != 0, HD = (MASK & PIGS_FN_DERIV) != 0;
    constexpr bool HL = (MASK & PIGS_FN_LAPL) != 0;  // Laplacian = Hessian with G00 = G11 = gl, G01 = 0 (scalar path: as such)
    constexpr bool HH = (MASK & PIGS_FN_HESS) != 0 || HL;
    constexpr int OD = 2 + (HV ? 1 : 0), OH = OD + (HD ? 2 : 0);
    const float x0 = r[0], x1 = r[1];
    const float gu = HV ? r[2] : 0.0f;
    const float g0 = HD ? r[OD] : 0.0f, g1 = HD ? r[OD + 1] : 0.0f;
    const float G00 = HH ? r[OH] : 0.0f, G01 = (HH && !HL) ? r[OH + 1] : 0.0f, G11 = HL ? r[OH] : (HH ? r[OH + 2] : 0.0f);
    float X0[KG], X1[KG], a0[KG], a1[KG], al[KG], S[KG], p0[KG], p1[KG];
#pragma unroll
    for (int j = 0; j < KG; j++) X0[j] = g[j].m[0] - x0;
#pragma unroll
    for (int j = 0; j < KG; j++) X1[j] = g[j].m[1] - x1;
    if constexpr (MODE == EV_SHIFT) {
#pragma unroll
        for (int j = 0; j < KG; j++) X0[j] = __fadd_rn(X0[j], g[j].sh[0]);
#pragma unroll
        for (int j = 0; j < KG; j++) X1[j] = __fadd_rn(X1[j], g[j].sh[1]);
    }
    if constexpr (MODE == EV_GENERAL) {
#pragma unroll
        for (int j = 0; j < KG; j++) {
            X0[j] = torus_wrap(X0[j]);
            X1[j] = torus_wrap(X1[j]);
        }
    }
#pragma unroll
    for (int j = 0; j < KG; j++) a0[j] = g[j].c[0] * X0[j];
#pragma unroll
    for (int j = 0; j < KG; j++) a1[j] = g[j].c[1] * X0[j];
#pragma unroll
    for (int j = 0; j < KG; j++) a0[j] = fmaf(g[j].c[1], X1[j], a0[j]);
#pragma unroll
    for (int j = 0; j < KG; j++) a1[j] = fmaf(g[j].c[2], X1[j], a1[j]);
#pragma unroll
    for (int j = 0; j < KG; j++) al[j] = a0[j] * X0[j];
#pragma unroll
    for (int j = 0; j < KG; j++) al[j] = fmaf(a1[j], X1[j], al[j]);
#pragma unroll
    for (int j = 0; j < KG; j++) {
        const float q = al[j];
        al[j] = ex2_approx(q * KH);
        if constexpr (MODE == EV_GENERAL) {
            if (q < 0.0f) al[j] = 0.0f;  // power > 0: the reference skips the pair (backward.cu:81)
        }
    }
    // S = sum_w G_w f_w and p = dS/da, the sample's upstream weights staying in one operand slot across the KG Gaussians
#pragma unroll
    for (int j = 0; j < KG; j++) S[j] = HD ? fmaf(g0, a0[j], gu) : gu;
    if constexpr (HD) {
#pragma unroll
        for (int j = 0; j < KG; j++) S[j] = fmaf(g1, a1[j], S[j]);
    }
    if constexpr (HH) {
        float h[KG];
#pragma unroll
        for (int j = 0; j < KG; j++) h[j] = fmaf(a0[j], a0[j], -g[j].c[0]);
#pragma unroll
        for (int j = 0; j < KG; j++) S[j] = fmaf(G00, h[j], S[j]);
#pragma unroll
        for (int j = 0; j < KG; j++) h[j] = fmaf(a0[j], a1[j], -g[j].c[1]);
#pragma unroll
        for (int j = 0; j < KG; j++) S[j] = fmaf(G01, h[j], S[j]);
#pragma unroll
        for (int j = 0; j < KG; j++) h[j] = fmaf(a1[j], a1[j], -g[j].c[2]);
#pragma unroll
        for (int j = 0; j < KG; j++) S[j] = fmaf(G11, h[j], S[j]);
        const float G00x2 = G00 + G00, G11x2 = G11 + G11;
#pragma unroll
        for (int j = 0; j < KG; j++) p0[j] = fmaf(G00x2, a0[j], g0);
#pragma unroll
        for (int j = 0; j < KG; j++) p1[j] = fmaf(G11x2, a1[j], g1);
#pragma unroll
        for (int j = 0; j < KG; j++) p0[j] = fmaf(G01, a1[j], p0[j]);
#pragma unroll
        for (int j = 0; j < KG; j++) p1[j] = fmaf(G01, a0[j], p1[j]);
    } else {
#pragma unroll
        for (int j = 0; j < KG; j++) {
            p0[j] = g0;
            p1[j] = g1;
        }
    }
    // per Gaussian: r = p - S X (-> dL/dmu = w C r, conic applied in the epilogue), e = p - S X / 2,
    // dConic_ij += alpha (e_i X_j sym + E_ij), E = -G; the six accumulations share alpha
#pragma unroll
    for (int j = 0; j < KG; j++) {
        const float Sj = S[j], hS = -0.5f * Sj;
        float r0, r1, e1, e2;
        if constexpr (HD || HH) {
            r0 = fmaf(-Sj, X0[j], p0[j]);
            e1 = fmaf(hS, X0[j], p0[j]);
            r1 = fmaf(-Sj, X1[j], p1[j]);
            e2 = fmaf(hS, X1[j], p1[j]);
        } else {
            r0 = -Sj * X0[j];
            e1 = hS * X0[j];
            r1 = -Sj * X1[j];
            e2 = hS * X1[j];
        }
        float T0, T1, T2;
        if constexpr (HH) {
            T0 = fmaf(e1, X0[j], -G00);
            T1 = fmaf(e1, X1[j], fmaf(e2, X0[j], -G01));
            T2 = fmaf(e2, X1[j], -G11);
        } else {
            T0 = e1 * X0[j];
            T1 = fmaf(e1, X1[j], e2 * X0[j]);
            T2 = e2 * X1[j];
        }
        const float w = al[j];
        dv[j][0] = fmaf(w, Sj, dv[j][0]);
        dm[j][0] = fmaf(w, r0, dm[j][0]);
        dm[j][1] = fmaf(w, r1, dm[j][1]);
        dc[j][0] = fmaf(w, T0, dc[j][0]);
        dc[j][1] = fmaf(w, T1, dc[j][1]);
        dc[j][2] = fmaf(w, T2, dc[j][2]);
    }
}

#ifndef PIGS_BWD_PACKED
#define PIGS_BWD_PACKED 1
#endif

// The KG Gaussians a thread owns, as KG/2 register PAIRS (2-D, one channel): parameters and gradient accumulators.
template <int KP>
struct GaussPairs {
    f32x2 m0[KP], m1[KP], c0[KP], c1[KP], c2[KP], sh0[KP], sh1[KP], v[KP];
    f32x2 tc[KP];  // c0 + c2 (trace-only Laplacian kernels)
    f32x2 dv[KP], dm0[KP], dm1[KP], dc0[KP], dc1[KP], dc2[KP];
};

// One sample record against the Gaussian pairs of a thread, in packed FP32x2 arithmetic (same IEEE operations per
// lane as the scalar path).  Why packed: on B200 an FP32 instruction with three register operands that are not in
// the operand-reuse cache takes two dispatch cycles, and with the Gaussians stationary most of this kernel's FMAs
// are of that kind (two per-pair temporaries plus an accumulator or a Gaussian parameter).  One FFMA2 on three
// register pairs does TWO such FMAs in three cycles and needs no reuse-cache luck (tools/mb_issue.cu).  The sample
// record's values are warp-uniform scalars and enter as broadcast operands.
template <unsigned MASK, int MODE, int KP>
__device__ __forceinline__ void bwd_sample_2d_packed(GaussPairs<KP>& G, const float* __restrict__ r)
{
    constexpr bool HV = (MASK & PIGS_FN_VALUE) != 0, HD = (MASK & PIGS_FN_DERIV) != 0, HH = (MASK & PIGS_FN_HESS) != 0;
    constexpr bool HL = (MASK & PIGS_FN_LAPL) != 0;
    constexpr int OD = 2 + (HV ? 1 : 0), OH = OD + (HD ? 2 : 0);
    const float x0 = r[0], x1 = r[1];
    const float gu = HV ? r[2] : 0.0f;
    const float g0 = HD ? r[OD] : 0.0f, g1 = HD ? r[OD + 1] : 0.0f;
    const float G00 = HH ? r[OH] : 0.0f, G01 = HH ? r[OH + 1] : 0.0f, G11 = HH ? r[OH + 2] : 0.0f;
    const float gl = HL ? r[OH] : 0.0f;
    f32x2 X0[KP], X1[KP], a0[KP], a1[KP], al[KP], S[KP], p0[KP], p1[KP];
#pragma unroll
    for (int k = 0; k < KP; k++) X0[k] = add2(G.m0[k], bc2(-x0));
#pragma unroll
    for (int k = 0; k < KP; k++) X1[k] = add2(G.m1[k], bc2(-x1));
    if constexpr (MODE == EV_SHIFT) {
#pragma unroll
        for (int k = 0; k < KP; k++) X0[k] = add2(X0[k], G.sh0[k]);
#pragma unroll
        for (int k = 0; k < KP; k++) X1[k] = add2(X1[k], G.sh1[k]);
    }
    if constexpr (MODE == EV_GENERAL) {
#pragma unroll
        for (int k = 0; k < KP; k++) {
            float lo, hi;
            upk2(X0[k], lo, hi);
            X0[k] = pk2(torus_wrap(lo), torus_wrap(hi));
            upk2(X1[k], lo, hi);
            X1[k] = pk2(torus_wrap(lo), torus_wrap(hi));
        }
    }
#pragma unroll
    for (int k = 0; k < KP; k++) a0[k] = mul2(G.c0[k], X0[k]);
#pragma unroll
    for (int k = 0; k < KP; k++) a1[k] = mul2(G.c1[k], X0[k]);
#pragma unroll
    for (int k = 0; k < KP; k++) a0[k] = fma2(G.c1[k], X1[k], a0[k]);
#pragma unroll
    for (int k = 0; k < KP; k++) a1[k] = fma2(G.c2[k], X1[k], a1[k]);
#pragma unroll
    for (int k = 0; k < KP; k++) al[k] = mul2(a0[k], X0[k]);
#pragma unroll
    for (int k = 0; k < KP; k++) al[k] = fma2(a1[k], X1[k], al[k]);
#pragma unroll
    for (int k = 0; k < KP; k++) {
        float qlo, qhi;
        upk2(al[k], qlo, qhi);
        float alo = ex2_approx(qlo * KH), ahi = ex2_approx(qhi * KH);
        if constexpr (MODE == EV_GENERAL) {
            if (qlo < 0.0f) alo = 0.0f;  // power > 0: the reference skips the pair (backward.cu:81)
            if (qhi < 0.0f) ahi = 0.0f;
        }
        al[k] = pk2(alo, ahi);
    }
#pragma unroll
    for (int k = 0; k < KP; k++) S[k] = HD ? fma2(bc2(g0), a0[k], bc2(gu)) : bc2(gu);
    if constexpr (HD) {
#pragma unroll
        for (int k = 0; k < KP; k++) S[k] = fma2(bc2(g1), a1[k], S[k]);
    }
    if constexpr (HH) {
        f32x2 h[KP];
#pragma unroll
        for (int k = 0; k < KP; k++) h[k] = fma2(a0[k], a0[k], neg2(G.c0[k]));
#pragma unroll
        for (int k = 0; k < KP; k++) S[k] = fma2(bc2(G00), h[k], S[k]);
#pragma unroll
        for (int k = 0; k < KP; k++) h[k] = fma2(a0[k], a1[k], neg2(G.c1[k]));
#pragma unroll
        for (int k = 0; k < KP; k++) S[k] = fma2(bc2(G01), h[k], S[k]);
#pragma unroll
        for (int k = 0; k < KP; k++) h[k] = fma2(a1[k], a1[k], neg2(G.c2[k]));
#pragma unroll
        for (int k = 0; k < KP; k++) S[k] = fma2(bc2(G11), h[k], S[k]);
        const float G00x2 = G00 + G00, G11x2 = G11 + G11;
#pragma unroll
        for (int k = 0; k < KP; k++) p0[k] = fma2(bc2(G00x2), a0[k], bc2(g0));
#pragma unroll
        for (int k = 0; k < KP; k++) p1[k] = fma2(bc2(G11x2), a1[k], bc2(g1));
#pragma unroll
        for (int k = 0; k < KP; k++) p0[k] = fma2(bc2(G01), a1[k], p0[k]);
#pragma unroll
        for (int k = 0; k < KP; k++) p1[k] = fma2(bc2(G01), a0[k], p1[k]);
    } else if constexpr (HL) {
        // trace only: S += gl (a0^2 + a1^2 - tr C), p = g + 2 gl a
        f32x2 h[KP];
#pragma unroll
        for (int k = 0; k < KP; k++) h[k] = fma2(a0[k], a0[k], neg2(G.tc[k]));
#pragma unroll
        for (int k = 0; k < KP; k++) h[k] = fma2(a1[k], a1[k], h[k]);
#pragma unroll
        for (int k = 0; k < KP; k++) S[k] = fma2(bc2(gl), h[k], S[k]);
        const float glx2 = gl + gl;
#pragma unroll
        for (int k = 0; k < KP; k++) p0[k] = fma2(bc2(glx2), a0[k], bc2(g0));
#pragma unroll
        for (int k = 0; k < KP; k++) p1[k] = fma2(bc2(glx2), a1[k], bc2(g1));
    } else {
#pragma unroll
        for (int k = 0; k < KP; k++) {
            p0[k] = bc2(g0);
            p1[k] = bc2(g1);
        }
    }
    // r = p - S X (-> dL/dmu = w C r, conic applied in the epilogue), e = p - S X / 2,
    // dConic_ij += alpha (e_i X_j sym + E_ij), E = -G
#pragma unroll
    for (int k = 0; k < KP; k++) {
        const f32x2 Sk = S[k], hS = mul2(Sk, bc2(-0.5f));
        f32x2 r0, r1, e1, e2, T0, T1, T2;
        if constexpr (HD || HH || HL) {
            r0 = fma2(neg2(Sk), X0[k], p0[k]);
            e1 = fma2(hS, X0[k], p0[k]);
            r1 = fma2(neg2(Sk), X1[k], p1[k]);
            e2 = fma2(hS, X1[k], p1[k]);
        } else {
            r0 = mul2(neg2(Sk), X0[k]);
            e1 = mul2(hS, X0[k]);
            r1 = mul2(neg2(Sk), X1[k]);
            e2 = mul2(hS, X1[k]);
        }
        if constexpr (HH) {
            T0 = fma2(e1, X0[k], bc2(-G00));
            T1 = fma2(e1, X1[k], fma2(e2, X0[k], bc2(-G01)));
            T2 = fma2(e2, X1[k], bc2(-G11));
        } else if constexpr (HL) {
            T0 = fma2(e1, X0[k], bc2(-gl));
            T1 = fma2(e1, X1[k], mul2(e2, X0[k]));
            T2 = fma2(e2, X1[k], bc2(-gl));
        } else {
            T0 = mul2(e1, X0[k]);
            T1 = fma2(e1, X1[k], mul2(e2, X0[k]));
            T2 = mul2(e2, X1[k]);
        }
        const f32x2 w = al[k];
        G.dv[k] = fma2(w, Sk, G.dv[k]);
        G.dm0[k] = fma2(w, r0, G.dm0[k]);
        G.dm1[k] = fma2(w, r1, G.dm1[k]);
        G.dc0[k] = fma2(w, T0, G.dc0[k]);
        G.dc1[k] = fma2(w, T1, G.dc1[k]);
        G.dc2[k] = fma2(w, T2, G.dc2[k]);
    }
}

// ---- generalised packed path: 2-D, one or TWO channels, any subset of value / gradient / Hessian / third derivative ----
// (NAVIER_STOKES' shape: model_pn.py:654,778 samples C = 2 with the third derivative; backward.cu:300-416.)
// Same formulation as bwd_sample_2d_packed — S, p = dS/da, E = explicit dS/dConic, then dL/dmu ~ alpha (p - S X),
// dL/dConic ~ alpha (sym(e X^T) + E) — with
//   * the polynomial factors shared by the channels (h = a a^T - C and the four third-derivative factors, written in
//     terms of h: f_xxx = a0 (2 c0 - h00), f_xxy = 2 c1 a0 - a1 h00, f_xyy = 2 c1 a1 - a0 h11, f_yyy = a1 (2 c2 - h11)),
//   * per channel only S_c (for dL/dv_c = alpha S_c),
//   * p and E formed from the VALUE-WEIGHTED upstream weights W = sum_c v_c G_c (one FFMA2 + FMUL2 per weight for two
//     channels; for one channel W is the sample's scalar itself and v is applied once per Gaussian in the epilogue).
template <int KP, int C>
struct GaussPairsG {
    f32x2 m0[KP], m1[KP], c0[KP], c1[KP], c2[KP], sh0[KP], sh1[KP], v[C][KP];
    f32x2 t0[KP], t1[KP], t2[KP];  // 2 c (third-derivative factors)
    f32x2 dv[C][KP], dm0[KP], dm1[KP], dc0[KP], dc1[KP], dc2[KP];
};

template <unsigned MASK, int MODE, int KP, int C>
__device__ __forceinline__ void bwd_sample_2d_packed_gen(GaussPairsG<KP, C>& G, const float* __restrict__ r)
{
    constexpr bool HV = (MASK & PIGS_FN_VALUE) != 0, HD = (MASK & PIGS_FN_DERIV) != 0, HH = (MASK & PIGS_FN_HESS) != 0;
    constexpr bool HT = (MASK & PIGS_FN_THIRD) != 0;
    static_assert((MASK & PIGS_FN_LAPL) == 0 && C >= 1 && C <= 2, "value / gradient / Hessian / third, one or two channels");
    constexpr int NA = nacc(2, MASK);
    constexpr int OD = HV ? 1 : 0, OH = OD + (HD ? 2 : 0), OT = OH + (HH ? 3 : 0);
    const float x0 = r[0], x1 = r[1];
    f32x2 X0[KP], X1[KP], a0[KP], a1[KP], al[KP];
#pragma unroll
    for (int k = 0; k < KP; k++) X0[k] = add2(G.m0[k], bc2(-x0));
#pragma unroll
    for (int k = 0; k < KP; k++) X1[k] = add2(G.m1[k], bc2(-x1));
    if constexpr (MODE == EV_SHIFT) {
#pragma unroll
        for (int k = 0; k < KP; k++) X0[k] = add2(X0[k], G.sh0[k]);
#pragma unroll
        for (int k = 0; k < KP; k++) X1[k] = add2(X1[k], G.sh1[k]);
    }
    if constexpr (MODE == EV_GENERAL) {
#pragma unroll
        for (int k = 0; k < KP; k++) {
            float lo, hi;
            upk2(X0[k], lo, hi);
            X0[k] = pk2(torus_wrap(lo), torus_wrap(hi));
            upk2(X1[k], lo, hi);
            X1[k] = pk2(torus_wrap(lo), torus_wrap(hi));
        }
    }
#pragma unroll
    for (int k = 0; k < KP; k++) a0[k] = mul2(G.c0[k], X0[k]);
#pragma unroll
    for (int k = 0; k < KP; k++) a1[k] = mul2(G.c1[k], X0[k]);
#pragma unroll
    for (int k = 0; k < KP; k++) a0[k] = fma2(G.c1[k], X1[k], a0[k]);
#pragma unroll
    for (int k = 0; k < KP; k++) a1[k] = fma2(G.c2[k], X1[k], a1[k]);
#pragma unroll
    for (int k = 0; k < KP; k++) al[k] = mul2(a0[k], X0[k]);
#pragma unroll
    for (int k = 0; k < KP; k++) al[k] = fma2(a1[k], X1[k], al[k]);
#pragma unroll
    for (int k = 0; k < KP; k++) {
        float qlo, qhi;
        upk2(al[k], qlo, qhi);
        float alo = ex2_approx(qlo * KH), ahi = ex2_approx(qhi * KH);
        if constexpr (MODE == EV_GENERAL) {
            if (qlo < 0.0f) alo = 0.0f;  // power > 0: the reference skips the pair (backward.cu:81)
            if (qhi < 0.0f) ahi = 0.0f;
        }
        al[k] = pk2(alo, ahi);
    }
    // factors shared by the channels
    f32x2 h00[KP], h01[KP], h11[KP], f3[4][KP];
    if constexpr (HH || HT) {
#pragma unroll
        for (int k = 0; k < KP; k++) h00[k] = fma2(a0[k], a0[k], neg2(G.c0[k]));
#pragma unroll
        for (int k = 0; k < KP; k++) h01[k] = fma2(a0[k], a1[k], neg2(G.c1[k]));
#pragma unroll
        for (int k = 0; k < KP; k++) h11[k] = fma2(a1[k], a1[k], neg2(G.c2[k]));
    }
    if constexpr (HT) {
#pragma unroll
        for (int k = 0; k < KP; k++) f3[0][k] = mul2(a0[k], add2(G.t0[k], neg2(h00[k])));
#pragma unroll
        for (int k = 0; k < KP; k++) f3[1][k] = fma2(neg2(a1[k]), h00[k], mul2(G.t1[k], a0[k]));
#pragma unroll
        for (int k = 0; k < KP; k++) f3[2][k] = fma2(neg2(a0[k]), h11[k], mul2(G.t1[k], a1[k]));
#pragma unroll
        for (int k = 0; k < KP; k++) f3[3][k] = mul2(a1[k], add2(G.t2[k], neg2(h11[k])));
    }
    // per channel: S_c = sum_w G_w f_w, dL/dv_c += alpha S_c
    f32x2 Sc[C][KP];
#pragma unroll
    for (int ch = 0; ch < C; ch++) {
        const float* __restrict__ g = r + 2 + ch * NA;
#pragma unroll
        for (int k = 0; k < KP; k++) {
            f32x2 S = HV ? bc2(g[0]) : bc2(0.0f);
            if constexpr (HD) {
                S = fma2(bc2(g[OD]), a0[k], S);
                S = fma2(bc2(g[OD + 1]), a1[k], S);
            }
            if constexpr (HH) {
                S = fma2(bc2(g[OH]), h00[k], S);
                S = fma2(bc2(g[OH + 1]), h01[k], S);
                S = fma2(bc2(g[OH + 2]), h11[k], S);
            }
            if constexpr (HT) {
#pragma unroll
                for (int i = 0; i < 4; i++) S = fma2(bc2(g[OT + i]), f3[i][k], S);
            }
            Sc[ch][k] = S;
            G.dv[ch][k] = fma2(al[k], S, G.dv[ch][k]);
        }
    }
    // value-weighted upstream weight q of the pair block k (C == 1: the sample's scalar, `scale` folded in on the scalar side)
    auto W = [&](int q, int k, float scale) -> f32x2 {
        if constexpr (C == 1) {
            return bc2(scale == 1.0f ? r[2 + q] : scale * r[2 + q]);
        } else {
            const f32x2 w = fma2(G.v[1][k], bc2(r[2 + NA + q]), mul2(G.v[0][k], bc2(r[2 + q])));
            return scale == 1.0f ? w : (scale == 2.0f ? add2(w, w) : mul2(w, bc2(scale)));
        }
    };
#pragma unroll
    for (int k = 0; k < KP; k++) {
        f32x2 S;
        if constexpr (C == 1)
            S = Sc[0][k];
        else
            S = fma2(G.v[1][k], Sc[1][k], mul2(G.v[0][k], Sc[0][k]));
        f32x2 p0 = HD ? W(OD, k, 1.0f) : bc2(0.0f), p1 = HD ? W(OD + 1, k, 1.0f) : bc2(0.0f);
        f32x2 E00 = bc2(0.0f), E01 = bc2(0.0f), E11 = bc2(0.0f);
        if constexpr (HH) {
            const f32x2 W00 = W(OH, k, 1.0f), W01 = W(OH + 1, k, 1.0f), W11 = W(OH + 2, k, 1.0f);
            const f32x2 a0x2 = add2(a0[k], a0[k]), a1x2 = add2(a1[k], a1[k]);
            p0 = fma2(W00, a0x2, p0);
            p1 = fma2(W11, a1x2, p1);
            p0 = fma2(W01, a1[k], p0);
            p1 = fma2(W01, a0[k], p1);
            E00 = neg2(W00);
            E01 = neg2(W01);
            E11 = neg2(W11);
        }
        if constexpr (HT) {
            // dS/da and explicit dS/dC of the third-derivative factors (eval_common.cuh upstream_spe), h = a a^T - C
            const f32x2 T0x3 = W(OT, k, 3.0f), T1 = W(OT + 1, k, 1.0f), T2 = W(OT + 2, k, 1.0f), T3x3 = W(OT + 3, k, 3.0f);
            const f32x2 T1x2 = add2(T1, T1), T2x2 = add2(T2, T2);
            p0 = fma2(neg2(T0x3), h00[k], p0);
            p1 = fma2(neg2(T1), h00[k], p1);
            p0 = fma2(neg2(T1x2), h01[k], p0);
            p1 = fma2(neg2(T2x2), h01[k], p1);
            p0 = fma2(neg2(T2), h11[k], p0);
            p1 = fma2(neg2(T3x3), h11[k], p1);
            E00 = fma2(T0x3, a0[k], E00);
            E01 = fma2(T1x2, a0[k], E01);
            E11 = fma2(T2, a0[k], E11);
            E00 = fma2(T1, a1[k], E00);
            E01 = fma2(T2x2, a1[k], E01);
            E11 = fma2(T3x3, a1[k], E11);
        }
        const f32x2 hS = mul2(S, bc2(-0.5f));
        f32x2 r0, r1, e1, e2;
        if constexpr (HD || HH || HT) {
            r0 = fma2(neg2(S), X0[k], p0);
            e1 = fma2(hS, X0[k], p0);
            r1 = fma2(neg2(S), X1[k], p1);
            e2 = fma2(hS, X1[k], p1);
        } else {
            r0 = mul2(neg2(S), X0[k]);
            e1 = mul2(hS, X0[k]);
            r1 = mul2(neg2(S), X1[k]);
            e2 = mul2(hS, X1[k]);
        }
        const f32x2 T0 = fma2(e1, X0[k], E00);
        const f32x2 T1 = fma2(e1, X1[k], fma2(e2, X0[k], E01));
        const f32x2 T2 = fma2(e2, X1[k], E11);
        const f32x2 w = al[k];
        G.dm0[k] = fma2(w, r0, G.dm0[k]);
        G.dm1[k] = fma2(w, r1, G.dm1[k]);
        G.dc0[k] = fma2(w, T0, G.dc0[k]);
        G.dc1[k] = fma2(w, T1, G.dc1[k]);
        G.dc2[k] = fma2(w, T2, G.dc2[k]);
    }
}

template <int D, int C, unsigned MASK, int MODE>
__global__ void __launch_bounds__(EV_THREADS, 2)
    eval_bwd_kernel(int blocks, const TileSeg* __restrict__ segs, WorkHeader* __restrict__ wh,
                    const float4* __restrict__ recs, const uint32_t* __restrict__ rec_id,
                    const float4* __restrict__ srecs, int Ctot, int ch0, float* __restrict__ dmeans,
                    float* __restrict__ dvalues, float* __restrict__ dconics)
{
    static_assert(!(D == 1 && (MASK & PIGS_FN_THIRD) && MASK != PIGS_FN_THIRD), "1-D third runs alone");
    constexpr int NV = grec_v4(D, C);
    constexpr int SV = srec_v4(D, C, MASK);
    constexpr int NA = nacc(D, MASK);
    constexpr int KG = bwd_kg(D, C, MASK);
    constexpr int NC = nconic(D);
    constexpr int SEG = MODE;
    constexpr int LD_PER_THREAD = (BWD_SB * SV + EV_THREADS - 1) / EV_THREADS;
    __shared__ float4 sm[2][BWD_SB * SV];
    __shared__ WorkItem s_item;
    __shared__ int s_valid;
    const int tid = threadIdx.x;

    for (;;) {
        if (tid == 0) {
            const uint32_t i = atomicAdd(&wh->counter[SEG], 1u);
            s_valid = i < wh->items[SEG];
            if (s_valid) s_item = decode_item(segs + (size_t)SEG * blocks, blocks, i, EV_THREADS * KG, false);
        }
        __syncthreads();
        if (!s_valid) break;
        const WorkItem it = s_item;

        // ---- stationary operand: KG Gaussians per thread ----
        GaussReg<D, C> g[KG];
        bool act[KG];
        float dm[KG][D], dc[KG][NC], dv[KG][C];
#pragma unroll
        for (int j = 0; j < KG; j++) {
            const uint32_t idx = tid + j * EV_THREADS;
            act[j] = idx < it.g_count;
            if (act[j]) {
                g[j] = load_gauss<D, C>(recs + (size_t)(it.g_begin + idx) * NV);
            } else {
#pragma unroll
                for (int d = 0; d < D; d++) g[j].m[d] = 0.0f;
#pragma unroll
                for (int k = 0; k < NC; k++) g[j].c[k] = 0.0f;
#pragma unroll
                for (int ch = 0; ch < C; ch++) g[j].v[ch] = 0.0f;
            }
#pragma unroll
            for (int d = 0; d < D; d++) dm[j][d] = 0.0f;
#pragma unroll
            for (int k = 0; k < NC; k++) dc[j][k] = 0.0f;
#pragma unroll
            for (int ch = 0; ch < C; ch++) dv[j][ch] = 0.0f;
        }

        // packed view of the stationary Gaussians (2-D, one channel)
        constexpr bool PACKED = PIGS_BWD_PACKED && D == 2 && C == 1 && (MASK & PIGS_FN_THIRD) == 0 && (KG % 2) == 0;
        GaussPairs<PACKED ? KG / 2 : 1> GP;
        if constexpr (PACKED) {
            const float nz = __uint_as_float(wh->neg_zero);
#pragma unroll
            for (int k = 0; k < KG / 2; k++) {
                const GaussReg<D, C>&ga = g[2 * k], &gb = g[2 * k + 1];
                GP.m0[k] = pk2_once(ga.m[0], gb.m[0], nz);
                GP.m1[k] = pk2_once(ga.m[D - 1], gb.m[D - 1], nz);
                GP.c0[k] = pk2_once(ga.c[0], gb.c[0], nz);
                GP.c1[k] = pk2_once(ga.c[NC > 1 ? 1 : 0], gb.c[NC > 1 ? 1 : 0], nz);
                GP.c2[k] = pk2_once(ga.c[NC - 1], gb.c[NC - 1], nz);
                GP.sh0[k] = pk2_once(ga.sh[0], gb.sh[0], nz);
                GP.sh1[k] = pk2_once(ga.sh[D - 1], gb.sh[D - 1], nz);
                GP.v[k] = pk2_once(ga.v[0], gb.v[0], nz);
                if constexpr ((MASK & PIGS_FN_LAPL) != 0)
                    GP.tc[k] = pk2_once(ga.c[0] + ga.c[NC - 1], gb.c[0] + gb.c[NC - 1], nz);
                GP.dv[k] = GP.dm0[k] = GP.dm1[k] = GP.dc0[k] = GP.dc1[k] = GP.dc2[k] = bc2(0.0f);
            }
        }

        // generalised packed view (2-D; two channels and / or the third derivative)
        constexpr bool PACKED_GEN = PIGS_BWD_PACKED && !PACKED && D == 2 && C <= 2 && (KG % 2) == 0 &&
                                    (MASK & PIGS_FN_LAPL) == 0 && (C == 2 || (MASK & PIGS_FN_THIRD) != 0);
        GaussPairsG<PACKED_GEN ? KG / 2 : 1, PACKED_GEN ? C : 1> GG;
        if constexpr (PACKED_GEN) {
            const float nz = __uint_as_float(wh->neg_zero);
#pragma unroll
            for (int k = 0; k < KG / 2; k++) {
                const GaussReg<D, C>&ga = g[2 * k], &gb = g[2 * k + 1];
                GG.m0[k] = pk2_once(ga.m[0], gb.m[0], nz);
                GG.m1[k] = pk2_once(ga.m[D - 1], gb.m[D - 1], nz);
                GG.c0[k] = pk2_once(ga.c[0], gb.c[0], nz);
                GG.c1[k] = pk2_once(ga.c[NC > 1 ? 1 : 0], gb.c[NC > 1 ? 1 : 0], nz);
                GG.c2[k] = pk2_once(ga.c[NC - 1], gb.c[NC - 1], nz);
                GG.sh0[k] = pk2_once(ga.sh[0], gb.sh[0], nz);
                GG.sh1[k] = pk2_once(ga.sh[D - 1], gb.sh[D - 1], nz);
#pragma unroll
                for (int ch = 0; ch < C; ch++) {
                    GG.v[ch][k] = pk2_once(ga.v[ch], gb.v[ch], nz);
                    GG.dv[ch][k] = bc2(0.0f);
                }
                if constexpr ((MASK & PIGS_FN_THIRD) != 0) {
                    GG.t0[k] = pk2_once(2.0f * ga.c[0], 2.0f * gb.c[0], nz);
                    GG.t1[k] = pk2_once(2.0f * ga.c[NC > 1 ? 1 : 0], 2.0f * gb.c[NC > 1 ? 1 : 0], nz);
                    GG.t2[k] = pk2_once(2.0f * ga.c[NC - 1], 2.0f * gb.c[NC - 1], nz);
                }
                GG.dm0[k] = GG.dm1[k] = GG.dc0[k] = GG.dc1[k] = GG.dc2[k] = bc2(0.0f);
            }
        }

        // ---- streamed operand: sample records through shared memory ----
        const float4* __restrict__ src = srecs + (size_t)it.s_begin * SV;
        const uint32_t total_v4 = it.s_count * SV;
        const int nbatch = (it.s_count + BWD_SB - 1) / BWD_SB;
        float4 pre[LD_PER_THREAD];
#pragma unroll
        for (int k = 0; k < LD_PER_THREAD; k++) {
            const uint32_t o = tid + k * EV_THREADS;
            pre[k] = o < total_v4 ? src[o] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int k = 0; k < LD_PER_THREAD; k++) {
            const uint32_t o = tid + k * EV_THREADS;
            if (o < BWD_SB * SV) sm[0][o] = pre[k];
        }
        __syncthreads();

        for (int b = 0; b < nbatch; b++) {
            const int buf = b & 1;
            if (b + 1 < nbatch) {
#pragma unroll
                for (int k = 0; k < LD_PER_THREAD; k++) {
                    const uint32_t l = tid + k * EV_THREADS;
                    const uint32_t o = (uint32_t)(b + 1) * BWD_SB * SV + l;
                    pre[k] = (l < BWD_SB * SV && o < total_v4) ? src[o] : make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
            const int cnt = min((int)BWD_SB, (int)it.s_count - b * BWD_SB);
            const float4* __restrict__ batch = sm[buf];
#pragma unroll 1
            for (int s = 0; s < cnt; s++) {
                float r[SV * 4];
                load_v4<SV>(batch + s * SV, r);
                if constexpr (PACKED) {
                    bwd_sample_2d_packed<MASK, MODE, KG / 2>(GP, r);
                    continue;
                }
                if constexpr (PACKED_GEN) {
                    bwd_sample_2d_packed_gen<MASK, MODE, KG / 2, C>(GG, r);
                    continue;
                }
                if constexpr (PIGS_BWD_OPMAJOR && D == 2 && C == 1 && (MASK & PIGS_FN_THIRD) == 0) {
                    bwd_sample_2d<MASK, MODE, KG>(g, r, dm, dc, dv);
                    continue;
                }
#pragma unroll
                for (int j = 0; j < KG; j++) {
                    PairGeom<D, MODE> pg;
                    pg.compute(g[j], r);
                    float f[NA];
                    pair_factors<D, MASK>(pg.a, g[j].c, f);
                    float S, p[D], E[NC];
                    if constexpr (C == 1) {
                        upstream_spe<D, MASK>(r + D, f, pg.a, g[j].c, S, p, E);
                        dv[j][0] = fmaf(pg.alpha, S, dv[j][0]);
                    } else {
                        S = 0.0f;
#pragma unroll
                        for (int d = 0; d < D; d++) p[d] = 0.0f;
#pragma unroll
                        for (int k = 0; k < NC; k++) E[k] = 0.0f;
#pragma unroll
                        for (int ch = 0; ch < C; ch++) {
                            float Sc, pc[D], Ec[NC];
                            upstream_spe<D, MASK>(r + D + ch * NA, f, pg.a, g[j].c, Sc, pc, Ec);
                            dv[j][ch] = fmaf(pg.alpha, Sc, dv[j][ch]);
                            S = fmaf(g[j].v[ch], Sc, S);
#pragma unroll
                            for (int d = 0; d < D; d++) p[d] = fmaf(g[j].v[ch], pc[d], p[d]);
#pragma unroll
                            for (int k = 0; k < NC; k++) E[k] = fmaf(g[j].v[ch], Ec[k], E[k]);
                        }
                    }
                    // Gradient accumulation.  With r = p - S X:  dL/dmu = w C r  (because a = C X and C is
                    // symmetric: C p - S a = C (p - S X)), so only alpha*r is accumulated per pair and the
                    // conic (and, for one channel, the value v) is applied ONCE per Gaussian in the epilogue.
                    // dL/dConic_ij += w (e_i X_j sym + E_ij),  e = p - S X / 2.
                    const float w = pg.alpha;
                    if constexpr (D == 1) {
                        const float X = pg.X[0], a = pg.a[0], c0 = g[j].c[0];
                        dm[j][0] = fmaf(w, fmaf(-S, X, p[0]), dm[j][0]);
                        if constexpr (MASK == PIGS_FN_THIRD) {
                            // backward.cu:326-329 as written: with T the (value-weighted) upstream weight,
                            // dconic += (2X^2 - 2a^2 X - (2Xa - X)X^2/2 + (a^2 - c)aX^2/2) * T * alpha
                            float Tv;
                            if constexpr (C == 1)
                                Tv = r[D];
                            else {
                                Tv = 0.0f;
#pragma unroll
                                for (int ch = 0; ch < C; ch++) Tv = fmaf(g[j].v[ch], r[D + ch * NA], Tv);
                            }
                            const float X2 = X * X;
                            const float qq = 2.0f * X2 - 2.0f * a * a * X - 0.5f * (2.0f * X * a - X) * X2 +
                                             0.5f * (a * a - c0) * a * X2;
                            dc[j][0] = fmaf(w, qq * Tv, dc[j][0]);
                        } else {
                            const float e = fmaf(-0.5f * S, X, p[0]);
                            dc[j][0] = fmaf(w, fmaf(e, X, E[0]), dc[j][0]);
                        }
                    } else {
                        const float X0 = pg.X[0], X1 = pg.X[1];
                        dm[j][0] = fmaf(w, fmaf(-S, X0, p[0]), dm[j][0]);
                        dm[j][1] = fmaf(w, fmaf(-S, X1, p[1]), dm[j][1]);
                        const float hS = -0.5f * S;
                        const float e1 = fmaf(hS, X0, p[0]);
                        const float e2 = fmaf(hS, X1, p[1]);
                        dc[j][0] = fmaf(w, fmaf(e1, X0, E[0]), dc[j][0]);
                        dc[j][1] = fmaf(w, fmaf(e1, X1, fmaf(e2, X0, E[1])), dc[j][1]);
                        dc[j][2] = fmaf(w, fmaf(e2, X1, E[2]), dc[j][2]);
                    }
                }
            }
            if (b + 1 < nbatch) {
#pragma unroll
                for (int k = 0; k < LD_PER_THREAD; k++) {
                    const uint32_t o = tid + k * EV_THREADS;
                    if (o < BWD_SB * SV) sm[buf ^ 1][o] = pre[k];
                }
            }
            __syncthreads();
        }

        // conic / value of each owned Gaussian for the epilogue (the packed path keeps no scalar copy alive in the loop)
        float ec[KG][NC], ev[KG];
        if constexpr (PACKED) {
#pragma unroll
            for (int k = 0; k < KG / 2; k++) {
                upk2(GP.dv[k], dv[2 * k][0], dv[2 * k + 1][0]);
                upk2(GP.dm0[k], dm[2 * k][0], dm[2 * k + 1][0]);
                upk2(GP.dm1[k], dm[2 * k][D - 1], dm[2 * k + 1][D - 1]);
                upk2(GP.dc0[k], dc[2 * k][0], dc[2 * k + 1][0]);
                upk2(GP.dc1[k], dc[2 * k][NC > 1 ? 1 : 0], dc[2 * k + 1][NC > 1 ? 1 : 0]);
                upk2(GP.dc2[k], dc[2 * k][NC - 1], dc[2 * k + 1][NC - 1]);
                upk2(GP.c0[k], ec[2 * k][0], ec[2 * k + 1][0]);
                upk2(GP.c1[k], ec[2 * k][NC > 1 ? 1 : 0], ec[2 * k + 1][NC > 1 ? 1 : 0]);
                upk2(GP.c2[k], ec[2 * k][NC - 1], ec[2 * k + 1][NC - 1]);
                upk2(GP.v[k], ev[2 * k], ev[2 * k + 1]);
            }
        } else if constexpr (PACKED_GEN) {
#pragma unroll
            for (int k = 0; k < KG / 2; k++) {
#pragma unroll
                for (int ch = 0; ch < C; ch++) upk2(GG.dv[ch][k], dv[2 * k][ch], dv[2 * k + 1][ch]);
                upk2(GG.dm0[k], dm[2 * k][0], dm[2 * k + 1][0]);
                upk2(GG.dm1[k], dm[2 * k][D - 1], dm[2 * k + 1][D - 1]);
                upk2(GG.dc0[k], dc[2 * k][0], dc[2 * k + 1][0]);
                upk2(GG.dc1[k], dc[2 * k][NC > 1 ? 1 : 0], dc[2 * k + 1][NC > 1 ? 1 : 0]);
                upk2(GG.dc2[k], dc[2 * k][NC - 1], dc[2 * k + 1][NC - 1]);
                upk2(GG.c0[k], ec[2 * k][0], ec[2 * k + 1][0]);
                upk2(GG.c1[k], ec[2 * k][NC > 1 ? 1 : 0], ec[2 * k + 1][NC > 1 ? 1 : 0]);
                upk2(GG.c2[k], ec[2 * k][NC - 1], ec[2 * k + 1][NC - 1]);
                upk2(GG.v[0][k], ev[2 * k], ev[2 * k + 1]);
            }
        } else {
#pragma unroll
            for (int j = 0; j < KG; j++) {
#pragma unroll
                for (int k = 0; k < NC; k++) ec[j][k] = g[j].c[k];
                ev[j] = g[j].v[0];
            }
        }
        // ---- epilogue: one RED.ADD per Gaussian parameter per item ----
#pragma unroll
        for (int j = 0; j < KG; j++) {
            if (!act[j]) continue;
            const size_t id = rec_id[it.g_begin + tid + j * EV_THREADS];
            const float vs = (C == 1) ? ev[j] : 1.0f;  // C > 1: v is already inside S, p, E
            if constexpr (D == 1) {
                atomicAdd(dmeans + id, vs * (ec[j][0] * dm[j][0]));
            } else {
                atomicAdd(dmeans + id * 2, vs * fmaf(ec[j][1], dm[j][1], ec[j][0] * dm[j][0]));
                atomicAdd(dmeans + id * 2 + 1, vs * fmaf(ec[j][2], dm[j][1], ec[j][1] * dm[j][0]));
            }
#pragma unroll
            for (int k = 0; k < NC; k++) atomicAdd(dconics + id * NC + k, vs * dc[j][k]);
#pragma unroll
            for (int ch = 0; ch < C; ch++) atomicAdd(dvalues + id * Ctot + ch0 + ch, dv[j][ch]);
        }
        __syncthreads();
    }
}

template <int D, int C, unsigned MASK, int MODE>
static void launch_bwd_m(const EvalCall& c, const TileSeg* segs, int ch0, const EvalWs& ws, BwdOut o, cudaStream_t st)
{
    eval_bwd_kernel<D, C, MASK, MODE><<<num_sms() * 2, EV_THREADS, 0, st>>>(c.blocks, segs, ws.wh, ws.recs, ws.rec_id, ws.srecs,
                                                                           c.C, ch0, o.dmeans, o.dvalues, o.dconics);
}

// the three class kernels run concurrently on forked streams (see eval_fwd.cu launch_fwd_t)
template <int D, int C, unsigned MASK>
static int launch_bwd_t(const EvalCall& c, int ch0, const EvalWs& ws, BwdOut o, cudaStream_t st)
{
    const TileSeg* segs = c.segs + (size_t)WL_BWD * EV_NSEG * c.blocks;
    if (c.single_class) {  // small problem: every instance is GENERAL (pack.cu stage_classes) -> one launch, no fork / join
        launch_bwd_m<D, C, MASK, EV_GENERAL>(c, segs, ch0, ws, o, st);
        PIGS_LAUNCH_CHECK();
        return 0;
    }
    ClassFork fk;
    int rc = fk.fork(st);
    if (rc) return rc;
    launch_bwd_m<D, C, MASK, EV_FAST>(c, segs, ch0, ws, o, fk.stream(0));
    PIGS_LAUNCH_CHECK();
    launch_bwd_m<D, C, MASK, EV_SHIFT>(c, segs, ch0, ws, o, fk.stream(1));
    PIGS_LAUNCH_CHECK();
    launch_bwd_m<D, C, MASK, EV_GENERAL>(c, segs, ch0, ws, o, fk.stream(2));
    PIGS_LAUNCH_CHECK();
    return fk.join();
}

int launch_backward(const EvalCall& c, int C, unsigned mask, int ch0, const EvalWs& ws, BwdOut out, cudaStream_t st)
{
#define BW(D_, C_, M_) \
    if (c.D == D_ && C == C_ && mask == M_) return launch_bwd_t<D_, C_, M_>(c, ch0, ws, out, st);
#define BW_ALL(D_, C_) BW(D_, C_, 1u) BW(D_, C_, 2u) BW(D_, C_, 4u) BW(D_, C_, 7u) BW(D_, C_, 8u)
    BW_ALL(2, 1) BW_ALL(2, 2) BW_ALL(1, 1) BW_ALL(1, 2)
    BW(2, 1, 16u) BW(2, 1, 19u)  // trace-only Laplacian kernels
#undef BW_ALL
#undef BW
    return set_error(-1, "unsupported (D, C, mask)", __FILE__, __LINE__);
}

}  // namespace pigs
