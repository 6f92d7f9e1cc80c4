// aggregate16.cu — neighbour aggregation forward / backward specialised for the sizes model_pn.py uses
// (L = K = 16 feature / key channels, E = 25 -> D * NF = 12 Fourier frequencies per slot; model_pn.py:44-49,199-230).
// Same arithmetic as the generic kernels in aggregate.cu (reference: aggregate_neighbors.cu:129-321), but:
//   * HALF a warp (16 lanes) owns one Gaussian, so with 16-channel rows every lane is busy both in the per-slot
//     phase (lane = neighbour slot) and in the per-channel phase (lane = channel); slot scalars move between the
//     two phases with width-16 shuffles instead of shared memory;
//   * every loop bound is a compile-time constant (fully unrolled), key / feature rows are read as float4;
//   * sin / cos of f * pi * X use an exact product split + exact reduction modulo 2 and two degree-3 polynomials in
//     u^2 on [-1/4, 1/4] (max abs error 7e-8) instead of sincospif's general path — 12 of them per slot were 36 %
//     of the generic kernel's instructions;
//   * the backward keeps the 2E + NF shared-parameter gradients and dL/dT in REGISTERS across all rows of a thread
//     and reduces them once per CTA (the generic kernel does a shared-memory read-modify-write per term).
// The kernels are instruction-issue bound, not HBM bound: the gathers of neighbour rows hit in L2 (ncu: 0.9 GB of
// DRAM traffic against 2.7 GB algorithmic at P = 1M, 16 neighbours).
#include "common.cuh"

#include <stdlib.h>

#ifndef A16_VEC_RED
#define A16_VEC_RED 1
#endif

namespace pigs {

constexpr int A16_THREADS = 256;
// resident CTAs per SM the register allocation aims for: both kernels wait on dependent gathers (ncu: long_scoreboard
// is the dominant stall), so occupancy is what hides the latency.  Measured on B200 at P = 1M x 16 neighbours
// (profiles/r01t_agg16_occupancy.log): forward 1.10 ms at the compiler's 128 registers (2 CTAs / SM) -> 0.82 (4) ->
// 0.72 ms at 6 CTAs / SM (40 registers + 112 B of local memory); backward 3.30 ms at 167 registers (1 CTA / SM) ->
// 1.82 ms at 2 (128 registers + 112 B), slower again at 3
#ifndef A16_FWD_MINB
#define A16_FWD_MINB 6
#endif
#ifndef A16_HOIST
#define A16_HOIST 0
#endif
#ifndef A16_BWD_MINB
#define A16_BWD_MINB 2
#endif

// sin(pi t), cos(pi t) for t = f * X formed exactly as p + e, reduced modulo 2 exactly (reference: double sin/cos of
// frequencies[e] * M_PI * X[d], aggregate_neighbors.cu:181-182)
#ifndef A16_MUFU_SINCOS
#define A16_MUFU_SINCOS 1
#endif
__device__ __forceinline__ void sincospi_fx(float f, float X, float& s, float& c)
{
    const float p = __fmul_rn(f, X);
    const float e = __fmaf_rn(f, X, -p);
    const float r = __fmaf_rn(-2.0f, rintf(0.5f * p), p);  // exact: p - 2 rint(p / 2) in [-1, 1]
    const float t = __fadd_rn(r, e);
#if A16_MUFU_SINCOS
    // the argument is reduced EXACTLY to [-pi, pi] above, where MUFU.SIN / MUFU.COS are good to 2^-21 absolute
    // (3.6e-7; the 24 terms of an embedding average it down to ~5e-7 of its magnitude, tolerance 1e-5): 9 instructions
    // per sin / cos pair instead of 22 for the polynomial form — 12 pairs per neighbour slot were a third of the kernels'
    // instructions (ncu: both kernels are issue-bound, not HBM-bound)
    const float ang = t * 3.14159265358979323846f;
    s = __sinf(ang);
    c = __cosf(ang);
#else
    const float q = rintf(t + t);                          // nearest half-integer quadrant, -2 .. 2
    const float u = __fmaf_rn(q, -0.5f, t);                // exact, |u| <= 1/4
    const float u2 = u * u;
    float sp = fmaf(u2, -0.5890768766403198f, 2.549767017364502f);
    sp = fmaf(u2, sp, -5.167707920074463f);
    sp = fmaf(u2, sp, 3.1415927410125732f);
    sp *= u;
    float cp = fmaf(u2, -1.3061283826828003f, 4.057577610015869f);
    cp = fmaf(u2, cp, -4.934788227081299f);
    cp = fmaf(u2, cp, 1.0f);
    const int qi = (int)q;
    const bool swap = qi & 1;
    const float s0 = swap ? cp : sp, c0 = swap ? sp : cp;
    // quadrant 1: (c, -s); 2: (-s, -c); 3: (-c, s)
    s = __uint_as_float(__float_as_uint(s0) ^ (((unsigned)qi & 2u) << 30));
    c = __uint_as_float(__float_as_uint(c0) ^ (((unsigned)(qi + 1) & 2u) << 30));
#endif
}

__device__ __forceinline__ void load16(const float* __restrict__ row, float* __restrict__ v)
{
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const float4 t = reinterpret_cast<const float4*>(row)[i];
        v[4 * i] = t.x;
        v[4 * i + 1] = t.y;
        v[4 * i + 2] = t.z;
        v[4 * i + 3] = t.w;
    }
}

template <int D, int NF>
__global__ void __launch_bounds__(A16_THREADS, A16_FWD_MINB)
    agg_forward16_kernel(int64_t row_lo, int64_t P, const float* __restrict__ features, const float* __restrict__ transform,
                         const float* __restrict__ queries, const float* __restrict__ keys, const float* __restrict__ freqs,
                         const float* __restrict__ dt, const int64_t* __restrict__ indices, const int64_t* __restrict__ ranges,
                         const float* __restrict__ dists, const float* __restrict__ densities,
                         const float* __restrict__ inv_total, float* __restrict__ weights, float* __restrict__ embeddings,
                         float* __restrict__ factors, float* __restrict__ out)
{
    constexpr int STRIDE = 2 * NF, E = D * STRIDE + 1;
    __shared__ float s_dt[2 * E], s_fr[NF], s_T[256];
    for (int i = threadIdx.x; i < 2 * E; i += A16_THREADS) s_dt[i] = dt[i];
    for (int i = threadIdx.x; i < NF; i += A16_THREADS) s_fr[i] = freqs[i];
    for (int i = threadIdx.x; i < 256; i += A16_THREADS) s_T[i] = transform[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, g = lane >> 4, gl = lane & 15;
    // warp index through a broadcast so the compiler knows it (and every loop bound derived from it) is warp-uniform:
    // otherwise each __shfl_sync below is wrapped in WARPSYNC / ENDCOLLECTIVE re-convergence code
    const int64_t warp = ((int64_t)blockIdx.x * A16_THREADS >> 5) + __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
    const int64_t nwarps = ((int64_t)gridDim.x * A16_THREADS) >> 5;

    for (int64_t base = row_lo + warp * 2; base < P; base += nwarps * 2) {  // warp-uniform; rows [row_lo, P)
        const int64_t row = base + g;
        const bool valid = row < P;
        const int64_t start = (valid && row > 0) ? ranges[row - 1] : 0;
        const int n = valid ? (int)(ranges[row] - start) : 0;
        float q[16];
        if (valid) {
            load16(queries + row * 16, q);
        } else {
#pragma unroll
            for (int k = 0; k < 16; k++) q[k] = 0.0f;
        }
        const float inv = valid ? inv_total[row] : 0.0f;
        float z = 0.0f;
        const int nmax = __reduce_max_sync(0xffffffffu, n);  // over both rows of the warp; lands in a uniform register
        for (int c0 = 0; c0 < nmax; c0 += 16) {
            const int64_t s = start + c0 + gl;
            const int64_t j = (c0 + gl < n) ? indices[s] : -1;
            float dwe = 0.0f, dwf = 0.0f;
#if A16_HOIST
            // the 16 feature-row gathers of the second phase depend on the indices only: issue them BEFORE the ~250
            // instructions of the per-slot phase, whose arithmetic then covers their latency
            float fv[16];
            {
                const int jl0 = (int)j;
#pragma unroll
                for (int t = 0; t < 16; t++) {
                    const int jj = __shfl_sync(0xffffffffu, jl0, t, 16);
                    fv[t] = jj >= 0 ? features[(size_t)jj * 16 + gl] : 0.0f;
                }
            }
#endif
            if (j >= 0) {
                float key[16];
                load16(keys + j * 16, key);
                float wgt = 0.0f;
#pragma unroll
                for (int k = 0; k < 16; k++) wgt = fmaf(q[k], key[k], wgt);
                float emb = 0.0f, fac = 0.0f;
#pragma unroll
                for (int d = 0; d < D; d++) {
                    const float X = dists[s * D + d];
#pragma unroll
                    for (int e = 0; e < NF; e++) {
                        float sn, cs;
                        sincospi_fx(s_fr[e], X, sn, cs);
                        const int i0 = d * STRIDE + 2 * e;
                        emb = fmaf(s_dt[i0], sn, emb);
                        emb = fmaf(s_dt[i0 + 1], cs, emb);
                        fac = fmaf(s_dt[E + i0], sn, fac);
                        fac = fmaf(s_dt[E + i0 + 1], cs, fac);
                    }
                }
                emb += s_dt[E - 1];
                fac += s_dt[2 * E - 1];
                weights[s] = wgt;
                embeddings[s] = emb;
                factors[s] = fac;
                const float dw = inv * densities[s] * wgt;
                dwf = dw * fac;
                dwe = dw * emb;
            } else if (c0 + gl < n) {
                // skipped neighbour: every slot of the row is written, so the caller's outputs need no zero fill
                weights[s] = 0.0f;
                embeddings[s] = 0.0f;
                factors[s] = 0.0f;
            }
            const int jl = (int)j;
            // z[lane] += sum_t (dwe_t + dwf_t feat_t[lane]): the first sum does not depend on the channel — one
            // half-warp butterfly instead of a broadcast per slot
            float se = dwe;
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o, 16);
            z += se;
#if A16_HOIST
            (void)jl;
#pragma unroll
            for (int t = 0; t < 16; t++) z = fmaf(__shfl_sync(0xffffffffu, dwf, t, 16), fv[t], z);  // skipped slots: dwf = fv = 0
#else
#pragma unroll
            for (int t = 0; t < 16; t++) {
                const int jj = __shfl_sync(0xffffffffu, jl, t, 16);
                const float a1 = __shfl_sync(0xffffffffu, dwf, t, 16);
                if (jj >= 0) z = fmaf(a1, features[(size_t)jj * 16 + gl], z);
            }
#endif
        }
        // out_i = z . T  (the reference applies T inside the neighbour loop, aggregate_neighbors.cu:201-206)
        float acc = 0.0f;
#pragma unroll
        for (int a = 0; a < 16; a++) acc = fmaf(__shfl_sync(0xffffffffu, z, a, 16), s_T[a * 16 + gl], acc);
        if (valid) out[row * 16 + gl] = acc;
    }
}


// Backward (generic counterpart and formulas: aggregate.cu agg_backward_kernel; reference :210-321).  Per row i with
// upstream g_i:  st = T g_i,  A = sum(st),  and per slot (j):  B = st . feat_j,  dc = density * inv_total_i,
//   dL/dfeat_j += dc w fac st      dL/dkey_j += q_i TE      dL/dq_i += key_j TE      TE = dc (emb A + fac B)
//   dL/dT[a][b] += g_i[b] sum_s w_s dc_s (emb_s + fac_s feat_j[a])
//   dL/ddt, dL/dfreq from dA = dc w A and dB = dc w B (reference :266-281)
template <int D, int NF>
__global__ void __launch_bounds__(A16_THREADS, A16_BWD_MINB)
    agg_backward16_kernel(int64_t row_lo, int64_t P, const float* __restrict__ features, const float* __restrict__ transform,
                          const float* __restrict__ queries, const float* __restrict__ keys, const float* __restrict__ freqs,
                          const float* __restrict__ dt, const int64_t* __restrict__ indices, const int64_t* __restrict__ ranges,
                          const float* __restrict__ dists, const float* __restrict__ densities,
                          const float* __restrict__ weights, const float* __restrict__ embeddings,
                          const float* __restrict__ factors, const float* __restrict__ inv_total,
                          const float* __restrict__ gout, float* __restrict__ g_features, float* __restrict__ g_transform,
                          float* __restrict__ g_queries, float* __restrict__ g_keys, float* __restrict__ g_freqs,
                          float* __restrict__ g_dt)
{
    constexpr int STRIDE = 2 * NF, E = D * STRIDE + 1, NCOL = 2 * E + NF;
    __shared__ float s_dt[2 * E], s_fr[NF], s_T[256];
    __shared__ float s_red[NCOL + 256];  // CTA totals of dL/ddt (2E), dL/dfreq (NF), dL/dT (256)
    for (int i = threadIdx.x; i < 2 * E; i += A16_THREADS) s_dt[i] = dt[i];
    for (int i = threadIdx.x; i < NF; i += A16_THREADS) s_fr[i] = freqs[i];
    for (int i = threadIdx.x; i < 256; i += A16_THREADS) s_T[(i & 15) * 16 + (i >> 4)] = transform[i];  // s_T[b][a] = T[a][b]
    for (int i = threadIdx.x; i < NCOL + 256; i += A16_THREADS) s_red[i] = 0.0f;
    __syncthreads();
    const int lane = threadIdx.x & 31, g = lane >> 4, gl = lane & 15;
    // warp index through a broadcast so the compiler knows it (and every loop bound derived from it) is warp-uniform:
    // otherwise each __shfl_sync below is wrapped in WARPSYNC / ENDCOLLECTIVE re-convergence code
    const int64_t warp = ((int64_t)blockIdx.x * A16_THREADS >> 5) + __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
    const int64_t nwarps = ((int64_t)gridDim.x * A16_THREADS) >> 5;
    const float PI = 3.14159265358979323846f;
    // per-thread partial sums, kept in registers over every row this thread touches
    float c_sn[D * NF][2], c_cs[D * NF][2], c_fr[NF], c_bias[2];  // [..][0]: dt row (emb), [..][1]: dt row E.. (fac)
    float c_T[16];                                                 // dL/dT[a][gl], a = 0..15
#pragma unroll
    for (int i = 0; i < D * NF; i++) c_sn[i][0] = c_sn[i][1] = c_cs[i][0] = c_cs[i][1] = 0.0f;
#pragma unroll
    for (int e = 0; e < NF; e++) c_fr[e] = 0.0f;
    c_bias[0] = c_bias[1] = 0.0f;
#pragma unroll
    for (int a = 0; a < 16; a++) c_T[a] = 0.0f;

    for (int64_t base = row_lo + warp * 2; base < P; base += nwarps * 2) {  // warp-uniform; rows [row_lo, P)
        const int64_t row = base + g;
        const bool valid = row < P;
        const int64_t start = (valid && row > 0) ? ranges[row - 1] : 0;
        const int n = valid ? (int)(ranges[row] - start) : 0;
        const int nmax = __reduce_max_sync(0xffffffffu, n);  // over both rows of the warp; lands in a uniform register
        if (nmax == 0) {
            if (valid) g_queries[row * 16 + gl] = 0.0f;  // every row of the window is written: no zero fill needed
            continue;
        }
        const float q_l = valid ? queries[row * 16 + gl] : 0.0f;  // lane = channel views
        const float g_l = valid ? gout[row * 16 + gl] : 0.0f;
        // st[a] = sum_b T[a][b] g[b]  (lane a), A = sum_a st[a]
        float st_l = 0.0f;
#pragma unroll
        for (int b = 0; b < 16; b++) st_l = fmaf(s_T[b * 16 + gl], __shfl_sync(0xffffffffu, g_l, b, 16), st_l);
        float A = st_l;
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) A += __shfl_xor_sync(0xffffffffu, A, o, 16);
        float st[16];  // every lane needs the whole vector for B = st . feat_j
#pragma unroll
        for (int a = 0; a < 16; a++) st[a] = __shfl_sync(0xffffffffu, st_l, a, 16);
        const float inv = valid ? inv_total[row] : 0.0f;
        float wz = 0.0f, gq = 0.0f;
#if A16_VEC_RED
        // scatter layout of the half-warp: lane = (neighbour t mod 4 -> tq, channel quad cq); one 16-byte vector RED
        // per (neighbour, quad) instead of four scalar ones (sm_90+: red.global.add.v4.f32)
        const int cq = gl & 3, tq = gl >> 2;
        float4 sv, qv;
        sv.x = __shfl_sync(0xffffffffu, st_l, 4 * cq + 0, 16), sv.y = __shfl_sync(0xffffffffu, st_l, 4 * cq + 1, 16);
        sv.z = __shfl_sync(0xffffffffu, st_l, 4 * cq + 2, 16), sv.w = __shfl_sync(0xffffffffu, st_l, 4 * cq + 3, 16);
        qv.x = __shfl_sync(0xffffffffu, q_l, 4 * cq + 0, 16), qv.y = __shfl_sync(0xffffffffu, q_l, 4 * cq + 1, 16);
        qv.z = __shfl_sync(0xffffffffu, q_l, 4 * cq + 2, 16), qv.w = __shfl_sync(0xffffffffu, q_l, 4 * cq + 3, 16);
#endif
        for (int c0 = 0; c0 < nmax; c0 += 16) {
            const int64_t s = start + c0 + gl;
            const int64_t j = (c0 + gl < n) ? indices[s] : -1;
            float te = 0.f, c1 = 0.f, wgt = 0.f, dce = 0.f, dcf = 0.f;
            if (j >= 0) {
                float f[16];
                load16(features + j * 16, f);
                float B = 0.0f;
#pragma unroll
                for (int a = 0; a < 16; a++) B = fmaf(st[a], f[a], B);
                wgt = weights[s];
                const float emb = embeddings[s], fac = factors[s];
                const float dc = densities[s] * inv;
                const float dcw = dc * wgt;
                dce = dc * emb;
                dcf = dc * fac;
                te = fmaf(dce, A, dcf * B);
                c1 = dcw * fac;
                const float dA = dcw * A, dB = dcw * B;
#pragma unroll
                for (int d = 0; d < D; d++) {
                    const float X = dists[s * D + d];
#pragma unroll
                    for (int e = 0; e < NF; e++) {
                        float sn, cs;
                        sincospi_fx(s_fr[e], X, sn, cs);
                        const int i0 = d * STRIDE + 2 * e, ci = d * NF + e;
                        c_sn[ci][0] = fmaf(dA, sn, c_sn[ci][0]);
                        c_sn[ci][1] = fmaf(dB, sn, c_sn[ci][1]);
                        c_cs[ci][0] = fmaf(dA, cs, c_cs[ci][0]);
                        c_cs[ci][1] = fmaf(dB, cs, c_cs[ci][1]);
                        const float ts = fmaf(s_dt[E + i0], dB, s_dt[i0] * dA);
                        const float tc = fmaf(s_dt[E + i0 + 1], dB, s_dt[i0 + 1] * dA);
                        c_fr[e] += PI * X * fmaf(cs, ts, -sn * tc);
                    }
                }
                c_bias[0] += dA;
                c_bias[1] += dB;
            }
            const int jl = (int)j;
#if A16_VEC_RED
#pragma unroll
            for (int it = 0; it < 4; it++) {
                const int t = 4 * it + tq;
                const int jj = __shfl_sync(0xffffffffu, jl, t, 16);
                const float t_te = __shfl_sync(0xffffffffu, te, t, 16);
                const float t_c1 = __shfl_sync(0xffffffffu, c1, t, 16);
                if (jj >= 0) {
                    const size_t o = (size_t)jj * 16 + 4 * cq;
                    atomicAdd(reinterpret_cast<float4*>(g_features + o),
                              make_float4(t_c1 * sv.x, t_c1 * sv.y, t_c1 * sv.z, t_c1 * sv.w));
                    atomicAdd(reinterpret_cast<float4*>(g_keys + o),
                              make_float4(qv.x * t_te, qv.y * t_te, qv.z * t_te, qv.w * t_te));
                }
            }
#endif
#pragma unroll
            for (int t = 0; t < 16; t++) {
                const int jj = __shfl_sync(0xffffffffu, jl, t, 16);
                const float t_te = __shfl_sync(0xffffffffu, te, t, 16);
#if !A16_VEC_RED
                const float t_c1 = __shfl_sync(0xffffffffu, c1, t, 16);
#endif
                const float t_w = __shfl_sync(0xffffffffu, wgt, t, 16);
                const float t_dce = __shfl_sync(0xffffffffu, dce, t, 16);
                const float t_dcf = __shfl_sync(0xffffffffu, dcf, t, 16);
                if (jj >= 0) {
                    const size_t o = (size_t)jj * 16 + gl;
#if !A16_VEC_RED
                    atomicAdd(g_features + o, t_c1 * st_l);
#endif
                    wz = fmaf(t_w, fmaf(t_dcf, features[o], t_dce), wz);
                    gq = fmaf(keys[o], t_te, gq);
#if !A16_VEC_RED
                    atomicAdd(g_keys + o, q_l * t_te);
#endif
                }
            }
        }
        if (valid) g_queries[row * 16 + gl] = gq;
        // dL/dT[a][b] += wz[a] g[b]: lane = b
#pragma unroll
        for (int a = 0; a < 16; a++) c_T[a] = fmaf(__shfl_sync(0xffffffffu, wz, a, 16), g_l, c_T[a]);
    }

    // ---- CTA reduction of the register partial sums, then one atomic per address per CTA ----
    // dt / freq columns: every lane of the CTA holds a partial -> warp butterfly, lane 0 adds to shared memory
    auto warp_sum_to = [&](float v, int col) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0 && v != 0.0f) atomicAdd(&s_red[col], v);
    };
#pragma unroll
    for (int d = 0; d < D; d++)
#pragma unroll
        for (int e = 0; e < NF; e++) {
            const int i0 = d * STRIDE + 2 * e, ci = d * NF + e;
            warp_sum_to(c_sn[ci][0], i0);
            warp_sum_to(c_cs[ci][0], i0 + 1);
            warp_sum_to(c_sn[ci][1], E + i0);
            warp_sum_to(c_cs[ci][1], E + i0 + 1);
        }
    warp_sum_to(c_bias[0], E - 1);
    warp_sum_to(c_bias[1], 2 * E - 1);
#pragma unroll
    for (int e = 0; e < NF; e++) warp_sum_to(c_fr[e], 2 * E + e);
    // dL/dT[a][b]: lanes gl and gl + 16 hold the same b
#pragma unroll
    for (int a = 0; a < 16; a++) {
        const float v = c_T[a] + __shfl_xor_sync(0xffffffffu, c_T[a], 16);
        if (g == 0 && v != 0.0f) atomicAdd(&s_red[NCOL + a * 16 + gl], v);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < NCOL + 256; i += A16_THREADS) {
        const float v = s_red[i];
        if (v == 0.0f) continue;
        if (i < 2 * E)
            atomicAdd(g_dt + i, v);
        else if (i < NCOL)
            atomicAdd(g_freqs + (i - 2 * E), v);
        else
            atomicAdd(g_transform + (i - NCOL), v);
    }
}

// ---- host dispatch (called from aggregate.cu); returns 1 if the sizes were handled, 0 to fall back, <0 / >0 on error
int launch_aggregate16_forward(int64_t row_lo, int64_t P, int D, int L, int K, int E, const float* features, const float* transform,
                               const float* queries, const float* keys, const float* freqs, const float* dt,
                               const int64_t* indices, const int64_t* ranges, const float* dists, const float* densities,
                               const float* inv_total, float* weights, float* embeddings, float* factors, float* out,
                               cudaStream_t st)
{
    if (L != 16 || K != 16 || E != 25) return 0;
    if ((reinterpret_cast<uintptr_t>(features) & 15) != 0) return 0;  // 16-byte row loads
    int64_t grid = (P - row_lo + 15) / 16;  // 16 rows per CTA (8 warps x 2)
    if (grid < 1) grid = 1;
    // CTAs per SM the grid-stride loop is cut into.  Measured on B200 at P = 1M x 16 neighbours (profiles/r03h_*): MORE,
    // smaller CTAs are faster (one resident wave of 6 per SM 0.597 ms, 12: 0.575, 16: 0.558, 32: 0.555, 64: 0.551, 128: 0.562, one per 16 rows: 0.618) — the hardware's
    // CTA scheduler balances the SMs better than equal shares of rows do
    static int per_sm = 0;
    if (!per_sm) {
        const char* e = getenv("PIGS_A16_FWD_CTAS_PER_SM");  // tuning hook
        per_sm = e ? atoi(e) : 64;
        if (per_sm < 1) per_sm = 64;
    }
    const int64_t cap = (int64_t)num_sms() * per_sm;
    if (grid > cap) grid = cap;
    if (D == 2)
        agg_forward16_kernel<2, 6><<<(unsigned)grid, A16_THREADS, 0, st>>>(row_lo, P, features, transform, queries, keys, freqs, dt,
                                                                            indices, ranges, dists, densities, inv_total,
                                                                            weights, embeddings, factors, out);
    else
        agg_forward16_kernel<1, 12><<<(unsigned)grid, A16_THREADS, 0, st>>>(row_lo, P, features, transform, queries, keys, freqs, dt,
                                                                             indices, ranges, dists, densities, inv_total,
                                                                             weights, embeddings, factors, out);
    PIGS_LAUNCH_CHECK();
    return 1;
}

int launch_aggregate16_backward(int64_t row_lo, int64_t P, int D, int L, int K, int E, const float* features, const float* transform,
                                const float* queries, const float* keys, const float* freqs, const float* dt,
                                const int64_t* indices, const int64_t* ranges, const float* dists, const float* densities,
                                const float* weights, const float* embeddings, const float* factors, const float* inv_total,
                                const float* gout, float* g_features, float* g_transform, float* g_queries, float* g_keys,
                                float* g_freqs, float* g_dt, cudaStream_t st)
{
    if (L != 16 || K != 16 || E != 25) return 0;
    // 16-byte row loads and vector REDs: rows of an unaligned view go through the generic kernels
    if (((reinterpret_cast<uintptr_t>(features) | reinterpret_cast<uintptr_t>(g_features) | reinterpret_cast<uintptr_t>(g_keys)) & 15) != 0) return 0;
    int64_t grid = (P - row_lo + 15) / 16;
    if (grid < 1) grid = 1;
    const int64_t cap = (int64_t)num_sms() * 8;  // few CTAs: every CTA ends with ~360 atomics on the shared parameters
    if (grid > cap) grid = cap;
    if (D == 2)
        agg_backward16_kernel<2, 6><<<(unsigned)grid, A16_THREADS, 0, st>>>(
            row_lo, P, features, transform, queries, keys, freqs, dt, indices, ranges, dists, densities, weights, embeddings, factors,
            inv_total, gout, g_features, g_transform, g_queries, g_keys, g_freqs, g_dt);
    else
        agg_backward16_kernel<1, 12><<<(unsigned)grid, A16_THREADS, 0, st>>>(
            row_lo, P, features, transform, queries, keys, freqs, dt, indices, ranges, dists, densities, weights, embeddings, factors,
            inv_total, gout, g_features, g_transform, g_queries, g_keys, g_freqs, g_dt);
    PIGS_LAUNCH_CHECK();
    return 1;
}

}  // namespace pigs
