// eval_fwd.cu — fused value / gradient / Hessian (/ third-derivative) evaluation of the mixture at
// every sample (replaces the four renderCUDA<F> instantiations of cuda_sampler/forward.cu:88-343).
//
// Design (B200, sm_100a):
//   * persistent grid (2 CTAs x 256 threads per SM), work items pulled from a device-side list with
//     one atomic per item: an item is (tile, chunk of 256*KS samples, slice of the tile's Gaussians).
//     The reference runs ONE CTA per tile — 16 CTAs on a [-1,1]^2 domain.
//   * each thread keeps KS samples and their accumulators in registers;
//   * the tile's Gaussians stream through shared memory as packed float4 records (pack.cu),
//     double-buffered; every record is read with broadcast LDS.128 and reused for KS pairs;
//   * exp() is one MUFU.EX2 (ex2.approx.ftz) with -log2(e)/2 folded into the argument;
//   * each output element is written once per item with a fire-and-forget RED.ADD (the reference does
//     a global read-modify-write per pair per component, forward.cu:175,182,...).
// The path is FP32-FMA-issue bound (no dense contraction => no tensor cores).
#include "layouts.cuh"

#ifndef PIGS_FWD_PACKED
#define PIGS_FWD_PACKED 1
#endif

namespace pigs {

constexpr int FWD_GB = 256;  // Gaussian records per shared-memory batch

// One Gaussian against the KS samples a thread holds, written OP-MAJOR: every step is applied to all KS samples
// before the next step starts.  On B200 an FP32 instruction that needs three register-file reads occupies the
// dispatch port for two cycles (tools/mb_issue.cu; operands served by the reuse cache, immediates and constants
// are free), so the order is chosen to keep the Gaussian's parameter in the same operand slot of consecutive
// instructions (reuse-cache hits) in the geometry phase, and the pair's weight w in the accumulation phase.
// Scaled variables: a' = K X (= KH a), e = X.a' = log2(alpha); factors in (a', M) come out times KH^order.
template <int D, int C, unsigned MASK, int MODE, int KS, int S0, int HS>
__device__ __forceinline__ void fwd_gaussian_block(const GaussFwd<D, C>& g, const float (&x)[KS][D],
                                                   const f32x2 (&xp)[D][(KS + 1) / 2], float (&acc)[KS][C][nacc(D, MASK)])
{
    constexpr int NA = nacc(D, MASK);
    constexpr bool PACKED = PIGS_FWD_PACKED && D == 2 && (HS % 2) == 0 && (S0 % 2) == 0;
    float X[D][HS], a[D][HS], al[HS];
    if constexpr (PACKED) {
        // geometry in packed FP32x2 over sample pairs: the Gaussian's parameters enter as scalar-broadcast operands,
        // so no instruction reads three register pairs except the last FMA of the quadratic form (3 cycles per 2)
        constexpr int HP = HS / 2, P0 = S0 / 2;
        f32x2 X0[HP], X1[HP], A0[HP], A1[HP], E[HP];
#pragma unroll
        for (int p = 0; p < HP; p++) X0[p] = add2(bc2(g.m[0]), neg2(xp[0][P0 + p]));
#pragma unroll
        for (int p = 0; p < HP; p++) X1[p] = add2(bc2(g.m[1]), neg2(xp[1][P0 + p]));
        if constexpr (MODE == EV_SHIFT) {
#pragma unroll
            for (int p = 0; p < HP; p++) X0[p] = add2(X0[p], bc2(g.sh[0]));
#pragma unroll
            for (int p = 0; p < HP; p++) X1[p] = add2(X1[p], bc2(g.sh[1]));
        }
        if constexpr (MODE == EV_GENERAL) {
#pragma unroll
            for (int p = 0; p < HP; p++) {
                float lo, hi;
                upk2(X0[p], lo, hi);
                X0[p] = pk2(torus_wrap(lo), torus_wrap(hi));
                upk2(X1[p], lo, hi);
                X1[p] = pk2(torus_wrap(lo), torus_wrap(hi));
            }
        }
#pragma unroll
        for (int p = 0; p < HP; p++) A0[p] = mul2(bc2(g.k[0]), X0[p]);
#pragma unroll
        for (int p = 0; p < HP; p++) A1[p] = mul2(bc2(g.k[1]), X0[p]);
#pragma unroll
        for (int p = 0; p < HP; p++) A0[p] = fma2(bc2(g.k[1]), X1[p], A0[p]);
#pragma unroll
        for (int p = 0; p < HP; p++) A1[p] = fma2(bc2(g.k[2]), X1[p], A1[p]);
#pragma unroll
        for (int p = 0; p < HP; p++) E[p] = mul2(A0[p], X0[p]);
#pragma unroll
        for (int p = 0; p < HP; p++) E[p] = fma2(A1[p], X1[p], E[p]);
#pragma unroll
        for (int p = 0; p < HP; p++) {
            upk2(A0[p], a[0][2 * p], a[0][2 * p + 1]);
            upk2(A1[p], a[1][2 * p], a[1][2 * p + 1]);
            upk2(E[p], al[2 * p], al[2 * p + 1]);
        }
    } else {
#pragma unroll
    for (int d = 0; d < D; d++) {
#pragma unroll
        for (int s = 0; s < HS; s++) X[d][s] = g.m[d] - x[S0 + s][d];
        if constexpr (MODE == EV_SHIFT) {
#pragma unroll
            for (int s = 0; s < HS; s++) X[d][s] = __fadd_rn(X[d][s], g.sh[d]);  // never contracted into the subtract
        }
        if constexpr (MODE == EV_GENERAL) {
#pragma unroll
            for (int s = 0; s < HS; s++) X[d][s] = torus_wrap(X[d][s]);
        }
    }
    if constexpr (D == 1) {
#pragma unroll
        for (int s = 0; s < HS; s++) a[0][s] = g.k[0] * X[0][s];
#pragma unroll
        for (int s = 0; s < HS; s++) al[s] = a[0][s] * X[0][s];
    } else {
#pragma unroll
        for (int s = 0; s < HS; s++) a[0][s] = g.k[0] * X[0][s];
#pragma unroll
        for (int s = 0; s < HS; s++) a[1][s] = g.k[1] * X[0][s];
#pragma unroll
        for (int s = 0; s < HS; s++) a[0][s] = fmaf(g.k[1], X[1][s], a[0][s]);
#pragma unroll
        for (int s = 0; s < HS; s++) a[1][s] = fmaf(g.k[2], X[1][s], a[1][s]);
#pragma unroll
        for (int s = 0; s < HS; s++) al[s] = a[0][s] * X[0][s];
#pragma unroll
        for (int s = 0; s < HS; s++) al[s] = fmaf(a[1][s], X[1][s], al[s]);
    }
    }
#pragma unroll
    for (int s = 0; s < HS; s++) {
        const float e = al[s];  // = KH * q = log2(alpha)
#ifdef PIGS_DBG_NOMUFU  // issue-rate experiment only
        al[s] = e + 1.0f;
#else
        al[s] = ex2_approx(e);
#endif
        if constexpr (MODE == EV_GENERAL) {
            // power = -q/2 > 0  <=>  e > 0 (KH < 0); NaN falls through like the reference's comparison (forward.cu:171)
            if (e > 0.0f) al[s] = 0.0f;
        }
    }
    // accumulation, sample-major: factors then the NA accumulators of one (sample, channel) back to back
#pragma unroll
    for (int s = 0; s < HS; s++) {
        float av[D], f[NA];
#pragma unroll
        for (int d = 0; d < D; d++) av[d] = a[d][s];
        pair_factors<D, MASK>(av, g.mm, f);
#pragma unroll
        for (int ch = 0; ch < C; ch++) {
            const float w = g.v[ch] * al[s];
#pragma unroll
            for (int k = 0; k < NA; k++) {
                if (k == 0 && (MASK & PIGS_FN_VALUE))
                    acc[S0 + s][ch][0] += w;
                else
                    acc[S0 + s][ch][k] = fmaf(w, f[k], acc[S0 + s][ch][k]);
            }
        }
    }
}

#ifndef PIGS_FWD_BLOCKED
#define PIGS_FWD_BLOCKED 1
#endif
#ifndef PIGS_FWD_PACKED_ACC
#define PIGS_FWD_PACKED_ACC 0
#endif
// Fully packed variant (2-D, one channel, no third derivative): geometry AND accumulation on sample PAIRS, accumulators
// kept as register pairs.  Every FFMA2 of the accumulation phase reads three distinct pairs (w, factor, accumulator) =
// 3 issue cycles per 2 FMAs unless w sits in the operand-reuse cache (tools/mb_issue.cu) — tools/sass_issue_model.py
// scores the loop ptxas produces; see DESIGN.md §4 for the comparison with the scalar accumulation.
template <unsigned MASK, int MODE, int KS>
__device__ __forceinline__ void fwd_gaussian_pairs(const GaussFwd<2, 1>& g, const f32x2 (&xp)[2][(KS + 1) / 2],
                                                   f32x2 (&accp)[KS / 2][nacc(2, MASK)])
{
    constexpr int HP = KS / 2;
    f32x2 X0[HP], X1[HP], A0[HP], A1[HP], E[HP], W[HP];
#pragma unroll
    for (int p = 0; p < HP; p++) X0[p] = add2(bc2(g.m[0]), neg2(xp[0][p]));
#pragma unroll
    for (int p = 0; p < HP; p++) X1[p] = add2(bc2(g.m[1]), neg2(xp[1][p]));
    if constexpr (MODE == EV_SHIFT) {
#pragma unroll
        for (int p = 0; p < HP; p++) X0[p] = add2(X0[p], bc2(g.sh[0]));
#pragma unroll
        for (int p = 0; p < HP; p++) X1[p] = add2(X1[p], bc2(g.sh[1]));
    }
    if constexpr (MODE == EV_GENERAL) {
#pragma unroll
        for (int p = 0; p < HP; p++) {
            float lo, hi;
            upk2(X0[p], lo, hi);
            X0[p] = pk2(torus_wrap(lo), torus_wrap(hi));
            upk2(X1[p], lo, hi);
            X1[p] = pk2(torus_wrap(lo), torus_wrap(hi));
        }
    }
#pragma unroll
    for (int p = 0; p < HP; p++) A0[p] = mul2(bc2(g.k[0]), X0[p]);
#pragma unroll
    for (int p = 0; p < HP; p++) A1[p] = mul2(bc2(g.k[1]), X0[p]);
#pragma unroll
    for (int p = 0; p < HP; p++) A0[p] = fma2(bc2(g.k[1]), X1[p], A0[p]);
#pragma unroll
    for (int p = 0; p < HP; p++) A1[p] = fma2(bc2(g.k[2]), X1[p], A1[p]);
#pragma unroll
    for (int p = 0; p < HP; p++) E[p] = mul2(A0[p], X0[p]);
#pragma unroll
    for (int p = 0; p < HP; p++) E[p] = fma2(A1[p], X1[p], E[p]);
#pragma unroll
    for (int p = 0; p < HP; p++) {
        float elo, ehi;
        upk2(E[p], elo, ehi);
        float alo = ex2_approx(elo), ahi = ex2_approx(ehi);
        if constexpr (MODE == EV_GENERAL) {
            if (elo > 0.0f) alo = 0.0f;  // power > 0: skipped by the reference (forward.cu:171)
            if (ehi > 0.0f) ahi = 0.0f;
        }
        W[p] = mul2(bc2(g.v[0]), pk2(alo, ahi));
    }
#pragma unroll
    for (int p = 0; p < HP; p++) {
        int k = 0;
        if constexpr ((MASK & PIGS_FN_VALUE) != 0) {
            accp[p][k] = add2(accp[p][k], W[p]);
            k++;
        }
        if constexpr ((MASK & PIGS_FN_DERIV) != 0) {
            accp[p][k] = fma2(W[p], A0[p], accp[p][k]);
            accp[p][k + 1] = fma2(W[p], A1[p], accp[p][k + 1]);
            k += 2;
        }
        if constexpr ((MASK & PIGS_FN_HESS) != 0) {
            accp[p][k] = fma2(W[p], fma2(A0[p], A0[p], bc2(-g.mm[0])), accp[p][k]);
            accp[p][k + 1] = fma2(W[p], fma2(A0[p], A1[p], bc2(-g.mm[1])), accp[p][k + 1]);
            accp[p][k + 2] = fma2(W[p], fma2(A1[p], A1[p], bc2(-g.mm[2])), accp[p][k + 2]);
            k += 3;
        }
        if constexpr ((MASK & PIGS_FN_LAPL) != 0) {
            accp[p][k] = fma2(W[p], fma2(A0[p], A0[p], fma2(A1[p], A1[p], bc2(-(g.mm[0] + g.mm[2])))), accp[p][k]);
            k++;
        }
    }
}

template <int D, int C, unsigned MASK, int MODE, int KS>
__device__ __forceinline__ void fwd_one_gaussian(const GaussFwd<D, C>& g, const float (&x)[KS][D],
                                                 const f32x2 (&xp)[D][(KS + 1) / 2], float (&acc)[KS][C][nacc(D, MASK)])
{
#ifndef PIGS_FWD_HS
#define PIGS_FWD_HS 8
#endif
    constexpr int HS = KS < PIGS_FWD_HS ? KS : PIGS_FWD_HS;
    static_assert(KS <= 2 * HS, "two blocks cover KS");
    fwd_gaussian_block<D, C, MASK, MODE, KS, 0, HS>(g, x, xp, acc);
    if constexpr (KS > HS) fwd_gaussian_block<D, C, MASK, MODE, KS, HS, HS>(g, x, xp, acc);
}

template <int D, int C, unsigned MASK, int MODE>
__global__ void __launch_bounds__(EV_THREADS, (fwd_ks(D, C, MASK) == 16 ? 1 : 2))
    eval_fwd_kernel(int blocks, const TileSeg* __restrict__ segs, WorkHeader* __restrict__ wh,
                    const float4* __restrict__ recs, const uint32_t* __restrict__ slist,
                    const float* __restrict__ samples, int Ctot, int ch0, float* __restrict__ out_u,
                    float* __restrict__ out_du, float* __restrict__ out_ddu, float* __restrict__ out_dddu, int lap)
{
    constexpr int NV = frec_v4(D, C);
    constexpr int NA = nacc(D, MASK);
    constexpr int KS = fwd_ks(D, C, MASK);
    constexpr int SEG = MODE;
    __shared__ float4 sm[2][FWD_GB * NV];
    __shared__ WorkItem s_item;
    __shared__ int s_valid;
    const int tid = threadIdx.x;

    for (;;) {
        if (tid == 0) {
            const uint32_t i = atomicAdd(&wh->counter[SEG], 1u);
            s_valid = i < wh->items[SEG];
            if (s_valid) s_item = decode_item(segs + (size_t)SEG * blocks, blocks, i, EV_THREADS * KS, true);
        }
        __syncthreads();
        if (!s_valid) break;
        const WorkItem it = s_item;

        // ---- stationary operand: KS samples per thread.  BLOCKED assignment (thread t owns samples t*KS .. t*KS+KS-1 of
        // the chunk): in a part-filled chunk (the last one of every tile; a 125 K-sample shard fills its chunks to 95 %) the
        // unused slots fall into WHOLE warps, whose batch loops run with a zero count and leave their issue slots to the SM's
        // other warps, instead of every thread computing 6-7 live and 1-2 dummy samples.  Measured (tools/fwd_shard_probe.py,
        // profiles/r03r_fwd_shard_probe.json): 125 K samples 11.27 -> 10.88 ms, 500 K 40.85 -> 40.69, 4 M unchanged (319.3) ----
        float x[KS][D];
        uint32_t sid[KS];
        float acc[KS][C][NA];
#pragma unroll
        for (int s = 0; s < KS; s++) {
#if PIGS_FWD_BLOCKED
            const uint32_t idx = tid * KS + s;
#else
            const uint32_t idx = tid + s * EV_THREADS;
#endif
            sid[s] = idx < it.s_count ? slist[it.s_begin + idx] : 0xffffffffu;
#pragma unroll
            for (int d = 0; d < D; d++) x[s][d] = sid[s] != 0xffffffffu ? samples[(size_t)sid[s] * D + d] : 0.0f;
#pragma unroll
            for (int ch = 0; ch < C; ch++)
#pragma unroll
                for (int k = 0; k < NA; k++) acc[s][ch][k] = 0.0f;
        }

        f32x2 xp[D][(KS + 1) / 2];
        if constexpr (PIGS_FWD_PACKED && D == 2 && (KS % 2) == 0) {
            const float nz = __uint_as_float(wh->neg_zero);
#pragma unroll
            for (int d = 0; d < D; d++)
#pragma unroll
                for (int p = 0; p < KS / 2; p++) xp[d][p] = pk2_once(x[2 * p][d], x[2 * p + 1][d], nz);
        }

        constexpr bool PACC = PIGS_FWD_PACKED_ACC && PIGS_FWD_PACKED && D == 2 && C == 1 && (KS % 2) == 0 &&
                              (MASK & PIGS_FN_THIRD) == 0;
        f32x2 accp[PACC ? KS / 2 : 1][NA];
        if constexpr (PACC) {
#pragma unroll
            for (int p = 0; p < KS / 2; p++)
#pragma unroll
                for (int k = 0; k < NA; k++) accp[p][k] = bc2(0.0f);
        }

#if PIGS_FWD_BLOCKED
        // whole warps without samples (part-filled chunk, blocked assignment) run the batch loop with a zero count
        const int live_mask = (uint32_t)((tid & ~31) * KS) < it.s_count ? -1 : 0;
#endif
        // ---- streamed operand: Gaussian records through shared memory, double-buffered ----
        const float4* __restrict__ src = recs + (size_t)it.g_begin * NV;
        const uint32_t total_v4 = it.g_count * NV;
        const int nbatch = (it.g_count + FWD_GB - 1) / FWD_GB;
        float4 pre[NV];
#pragma unroll
        for (int k = 0; k < NV; k++) {
            const uint32_t o = tid + k * EV_THREADS;
            pre[k] = o < total_v4 ? src[o] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int k = 0; k < NV; k++) sm[0][tid + k * EV_THREADS] = pre[k];
        __syncthreads();

        for (int b = 0; b < nbatch; b++) {
            const int buf = b & 1;
            if (b + 1 < nbatch) {
#pragma unroll
                for (int k = 0; k < NV; k++) {
                    const uint32_t o = (uint32_t)(b + 1) * FWD_GB * NV + tid + k * EV_THREADS;
                    pre[k] = o < total_v4 ? src[o] : make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
#if PIGS_FWD_BLOCKED
            const int cnt = min((int)FWD_GB, (int)it.g_count - b * FWD_GB) & live_mask;
#else
            const int cnt = min((int)FWD_GB, (int)it.g_count - b * FWD_GB);
#endif
            const float4* __restrict__ batch = sm[buf];
#pragma unroll 1
            for (int j = 0; j < cnt; j++) {
                const GaussFwd<D, C> g = load_gauss_fwd<D, C>(batch + j * NV);
                if constexpr (PACC)
                    fwd_gaussian_pairs<MASK, MODE, KS>(g, xp, accp);
                else
                    fwd_one_gaussian<D, C, MASK, MODE, KS>(g, x, xp, acc);
            }
            if (b + 1 < nbatch) {
#pragma unroll
                for (int k = 0; k < NV; k++) sm[buf ^ 1][tid + k * EV_THREADS] = pre[k];
            }
            __syncthreads();
        }

        if constexpr (PACC) {
#pragma unroll
            for (int p = 0; p < KS / 2; p++)
#pragma unroll
                for (int k = 0; k < NA; k++) upk2(accp[p][k], acc[2 * p][0][k], acc[2 * p + 1][0][k]);
        }
        // ---- epilogue: undo the KH^order scaling, one RED.ADD per output element per item ----
        constexpr float IK1 = 1.0f / KH, IK2 = IK1 * IK1, IK3 = IK2 * IK1;
#pragma unroll
        for (int s = 0; s < KS; s++) {
            if (sid[s] == 0xffffffffu) continue;
            const size_t id = sid[s];
#pragma unroll
            for (int ch = 0; ch < C; ch++) {
                const int c = ch0 + ch;
                int k = 0;
                if constexpr ((MASK & PIGS_FN_VALUE) != 0) {
                    if (out_u) atomicAdd(out_u + id * Ctot + c, acc[s][ch][k]);
                    k++;
                }
                if constexpr ((MASK & PIGS_FN_DERIV) != 0) {
#pragma unroll
                    for (int d = 0; d < D; d++) {
                        if (out_du) atomicAdd(out_du + (id * D + d) * Ctot + c, IK1 * acc[s][ch][k]);
                        k++;
                    }
                }
                if constexpr ((MASK & PIGS_FN_HESS) != 0) {
                    if constexpr (D == 1) {
                        if (out_ddu) atomicAdd(out_ddu + id * Ctot + c, IK2 * acc[s][ch][k]);
                        k++;
                    } else {
                        if (out_ddu && lap) {
                            // PIGS_FN_LAPL: trace of the Hessian, one element per (sample, channel)
                            atomicAdd(out_ddu + id * Ctot + c, IK2 * (acc[s][ch][k] + acc[s][ch][k + 2]));
                        } else if (out_ddu) {
                            float* o = out_ddu + id * 4 * Ctot + c;
                            const float hxy = IK2 * acc[s][ch][k + 1];
                            atomicAdd(o, IK2 * acc[s][ch][k]);
                            atomicAdd(o + Ctot, hxy);
                            atomicAdd(o + 2 * Ctot, hxy);
                            atomicAdd(o + 3 * Ctot, IK2 * acc[s][ch][k + 2]);
                        }
                        k += 3;
                    }
                }
                if constexpr ((MASK & PIGS_FN_LAPL) != 0) {
                    if (out_ddu) atomicAdd(out_ddu + id * Ctot + c, IK2 * acc[s][ch][k]);
                    k++;
                }
                if constexpr ((MASK & PIGS_FN_THIRD) != 0) {
                    if constexpr (D == 1) {
                        if (out_dddu) atomicAdd(out_dddu + id * Ctot + c, IK3 * acc[s][ch][k]);
                        k++;
                    } else {
                        if (out_dddu) {
                            // flat [x|y]^3 index -> unique component (forward.cu:264-271)
                            float* o = out_dddu + id * 8 * Ctot + c;
                            const float t0 = IK3 * acc[s][ch][k], t1 = IK3 * acc[s][ch][k + 1];
                            const float t2 = IK3 * acc[s][ch][k + 2], t3 = IK3 * acc[s][ch][k + 3];
                            atomicAdd(o, t0);
                            atomicAdd(o + 1 * Ctot, t1);
                            atomicAdd(o + 2 * Ctot, t1);
                            atomicAdd(o + 3 * Ctot, t2);
                            atomicAdd(o + 4 * Ctot, t1);
                            atomicAdd(o + 5 * Ctot, t2);
                            atomicAdd(o + 6 * Ctot, t2);
                            atomicAdd(o + 7 * Ctot, t3);
                        }
                        k += 4;
                    }
                }
            }
        }
        __syncthreads();  // s_item / sm reuse
    }
}

template <int D, int C, unsigned MASK, int MODE>
static void launch_fwd_m(const EvalCall& c, const TileSeg* segs, int ch0, const EvalWs& ws, FwdOut o, cudaStream_t st)
{
    eval_fwd_kernel<D, C, MASK, MODE><<<num_sms() * (fwd_ks(D, C, MASK) == 16 ? 1 : 2), EV_THREADS, 0, st>>>(c.blocks, segs, ws.wh, ws.recs, c.slist, c.samples,
                                                                           c.C, ch0, o.u, o.du, o.ddu, o.dddu, o.lap);
}

// The three class kernels of one call are independent (own work counters, outputs combined with RED.ADD), so they are
// launched on three streams forked from / joined back into the caller's: the persistent CTAs of the SHIFT and GENERAL
// kernels move in as the FAST kernel's CTAs run out of items, and the call has ONE tail instead of three back-to-back
// ones — what a small shard (N = 8: a few ms of forward per rank) loses most on.  Deterministic mode keeps them in order.
template <int D, int C, unsigned MASK>
static int launch_fwd_t(const EvalCall& c, int wl, int ch0, const EvalWs& ws, FwdOut o, cudaStream_t st)
{
    const TileSeg* segs = c.segs + (size_t)wl * EV_NSEG * c.blocks;
    if (c.single_class) {  // small problem: every instance is GENERAL (pack.cu stage_classes) -> one launch, no fork / join
        launch_fwd_m<D, C, MASK, EV_GENERAL>(c, segs, ch0, ws, o, st);
        PIGS_LAUNCH_CHECK();
        return 0;
    }
    ClassFork fk;
    int rc = fk.fork(st);
    if (rc) return rc;
    launch_fwd_m<D, C, MASK, EV_FAST>(c, segs, ch0, ws, o, fk.stream(0));
    PIGS_LAUNCH_CHECK();
    launch_fwd_m<D, C, MASK, EV_SHIFT>(c, segs, ch0, ws, o, fk.stream(1));
    PIGS_LAUNCH_CHECK();
    launch_fwd_m<D, C, MASK, EV_GENERAL>(c, segs, ch0, ws, o, fk.stream(2));
    PIGS_LAUNCH_CHECK();
    return fk.join();
}

int launch_forward(const EvalCall& c, int C, unsigned mask, int wl, int ch0, const EvalWs& ws, FwdOut out, cudaStream_t st)
{
#define FW(D_, C_, M_) \
    if (c.D == D_ && C == C_ && mask == M_) return launch_fwd_t<D_, C_, M_>(c, wl, ch0, ws, out, st);
#define FW_ALL(D_, C_) FW(D_, C_, 1u) FW(D_, C_, 2u) FW(D_, C_, 4u) FW(D_, C_, 7u) FW(D_, C_, 8u)
    FW_ALL(2, 1) FW_ALL(2, 2) FW_ALL(1, 1) FW_ALL(1, 2)
    FW(2, 1, 16u) FW(2, 1, 19u)  // trace-only Laplacian kernels (PIGS_FN_LAPL as a kernel mask bit)
#undef FW_ALL
#undef FW
    return set_error(-1, "unsupported (D, C, mask)", __FILE__, __LINE__);
}

}  // namespace pigs
