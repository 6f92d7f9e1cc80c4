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

        // ---- epilogue: one RED.ADD per Gaussian parameter per item ----
#pragma unroll
        for (int j = 0; j < KG; j++) {
            if (!act[j]) continue;
            const size_t id = rec_id[it.g_begin + tid + j * EV_THREADS];
            const float vs = (C == 1) ? g[j].v[0] : 1.0f;  // C > 1: v is already inside S, p, E
            if constexpr (D == 1) {
                atomicAdd(dmeans + id, vs * (g[j].c[0] * dm[j][0]));
            } else {
                atomicAdd(dmeans + id * 2, vs * fmaf(g[j].c[1], dm[j][1], g[j].c[0] * dm[j][0]));
                atomicAdd(dmeans + id * 2 + 1, vs * fmaf(g[j].c[2], dm[j][1], g[j].c[1] * dm[j][0]));
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
static void launch_bwd_m(const EvalCall& c, int ch0, const EvalWs& ws, BwdOut o, cudaStream_t st)
{
    eval_bwd_kernel<D, C, MASK, MODE><<<num_sms() * 2, EV_THREADS, 0, st>>>(c.blocks, ws.segs, ws.wh, ws.recs, ws.rec_id, ws.srecs,
                                                                           c.C, ch0, o.dmeans, o.dvalues, o.dconics);
}

template <int D, int C, unsigned MASK>
static int launch_bwd_t(const EvalCall& c, int mode, int ch0, const EvalWs& ws, BwdOut o, cudaStream_t st)
{
    if (mode == EV_FAST)
        launch_bwd_m<D, C, MASK, EV_FAST>(c, ch0, ws, o, st);
    else if (mode == EV_SHIFT)
        launch_bwd_m<D, C, MASK, EV_SHIFT>(c, ch0, ws, o, st);
    else
        launch_bwd_m<D, C, MASK, EV_GENERAL>(c, ch0, ws, o, st);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int launch_backward(const EvalCall& c, int C, unsigned mask, int mode, int ch0, const EvalWs& ws, BwdOut out,
                    cudaStream_t st)
{
#define BW(D_, C_, M_) \
    if (c.D == D_ && C == C_ && mask == M_) return launch_bwd_t<D_, C_, M_>(c, mode, ch0, ws, out, st);
#define BW_ALL(D_, C_) BW(D_, C_, 1u) BW(D_, C_, 2u) BW(D_, C_, 4u) BW(D_, C_, 7u) BW(D_, C_, 8u)
    BW_ALL(2, 1) BW_ALL(2, 2) BW_ALL(1, 1) BW_ALL(1, 2)
#undef BW_ALL
#undef BW
    return set_error(-1, "unsupported (D, C, mask)", __FILE__, __LINE__);
}

}  // namespace pigs
