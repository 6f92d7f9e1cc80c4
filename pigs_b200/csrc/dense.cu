// dense.cu — the reference's dense all-pairs sampling path as CUDA kernels:
//   gaussians.py:48-58   sample_gaussians      u[N,C]      = sum_g v_g exp(-1/2 x^T K_g x),        x = sample - mean_g
//   gaussians.py:89-101  gaussian_derivative   ux[N,D,C]   = sum_g -(K_g x) alpha v_g
//   gaussians.py:103-116 gaussian_derivative2  uxx[N,D,D,C] = sum_g ((K_g x)(K_g x)^T - K_g) alpha v_g
// with FULL [n, D, D] conics (not assumed symmetric), every Gaussian against every sample: no tile cutoff, no torus wrap,
// no power > 0 skip — the semantics of the pure-PyTorch path the reference's CPU drivers (test_no_mlp_1d.py,
// test_derivatives.py:110-143) use, without its O(N n d^2) broadcast tensors.  The backward is what torch.autograd
// gives those three functions: gradients to means, full conics and values.
// Forward: 1-4 samples per thread, Gaussian records broadcast from shared memory; backward: 1-2 Gaussians per thread,
// sample + upstream-gradient records broadcast from shared memory, sums carried in registers.  The other operand is cut
// into chunks over blockIdx.y so that small problems still fill the 148 SMs; a chunked launch adds into zeroed outputs
// with one RED per output element and chunk, an unchunked one stores.
// FP32-issue bound like the tiled kernels (per pair, D = 2, C = 1, all three outputs: 22 FP32 ops + 1 ex2 forward).
#include "common.cuh"

#include <stdlib.h>

namespace pigs {

constexpr int DN_FWD_THREADS = 256, DN_BWD_THREADS = 128, DN_BATCH = 256;
constexpr float DN_KH = -0.72134752044448170368f;  // -0.5 * log2(e)

__device__ __forceinline__ float dn_ex2(float x)
{
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// KS samples per thread (strided by the block size, so loads and stores stay coalesced): one set of shared-memory reads
// of a Gaussian's record serves KS pairs (ncu on the one-sample version: 0.83 issue slots busy per cycle with the LSU pipe
// at 40 % — the record reads, not the FMAs, filled the issue slots).
template <int D, int C, int KS>
__global__ void __launch_bounds__(DN_FWD_THREADS)
    dense_fwd_kernel(unsigned mask, int64_t n, int64_t N, int64_t per_chunk, const float* __restrict__ means,
                     const float* __restrict__ conics, const float* __restrict__ values, const float* __restrict__ samples,
                     float* __restrict__ out_u, float* __restrict__ out_ux, float* __restrict__ out_uxx)
{
    constexpr int REC = D + D * D + C;
    __shared__ float rec[DN_BATCH * REC];
    int64_t sidx[KS];
    float x[KS][D];
    float au[KS][C], ax[KS][D][C], ah[KS][D][D][C];
#pragma unroll
    for (int k = 0; k < KS; k++) {
        sidx[k] = ((int64_t)blockIdx.x * KS + k) * DN_FWD_THREADS + threadIdx.x;
#pragma unroll
        for (int d = 0; d < D; d++) x[k][d] = sidx[k] < N ? samples[sidx[k] * D + d] : 0.0f;
#pragma unroll
        for (int c = 0; c < C; c++) {
            au[k][c] = 0.0f;
#pragma unroll
            for (int i = 0; i < D; i++) {
                ax[k][i][c] = 0.0f;
#pragma unroll
                for (int j = 0; j < D; j++) ah[k][i][j][c] = 0.0f;
            }
        }
    }
    const int64_t g0 = (int64_t)blockIdx.y * per_chunk, g1 = min(n, g0 + per_chunk);
    for (int64_t b = g0; b < g1; b += DN_BATCH) {
        const int nb = (int)min((int64_t)DN_BATCH, g1 - b);
        __syncthreads();
        for (int i = threadIdx.x; i < nb * D; i += DN_FWD_THREADS) rec[(i / D) * REC + i % D] = means[b * D + i];
        for (int i = threadIdx.x; i < nb * D * D; i += DN_FWD_THREADS)
            rec[(i / (D * D)) * REC + D + i % (D * D)] = conics[b * D * D + i];
        for (int i = threadIdx.x; i < nb * C; i += DN_FWD_THREADS) rec[(i / C) * REC + D + D * D + i % C] = values[b * C + i];
        __syncthreads();
        for (int g = 0; g < nb; g++) {
            float r[REC];
#pragma unroll
            for (int i = 0; i < REC; i++) r[i] = rec[g * REC + i];
#pragma unroll
            for (int k = 0; k < KS; k++) {
                float X[D], a[D];
#pragma unroll
                for (int d = 0; d < D; d++) X[d] = x[k][d] - r[d];
                float q = 0.0f;
#pragma unroll
                for (int i = 0; i < D; i++) {
                    a[i] = 0.0f;
#pragma unroll
                    for (int j = 0; j < D; j++) a[i] = fmaf(r[D + i * D + j], X[j], a[i]);
                    q = fmaf(X[i], a[i], q);
                }
                const float alpha = dn_ex2(DN_KH * q);
#pragma unroll
                for (int c = 0; c < C; c++) {
                    const float w = alpha * r[D + D * D + c];
                    au[k][c] += w;
#pragma unroll
                    for (int i = 0; i < D; i++) {
                        ax[k][i][c] = fmaf(-a[i], w, ax[k][i][c]);
#pragma unroll
                        for (int j = 0; j < D; j++)
                            ah[k][i][j][c] = fmaf(fmaf(a[i], a[j], -r[D + i * D + j]), w, ah[k][i][j][c]);
                    }
                }
            }
        }
    }
    const bool add = gridDim.y > 1;
    auto put = [add](float* p, float v) {
        if (add)
            atomicAdd(p, v);
        else
            *p = v;
    };
#pragma unroll
    for (int k = 0; k < KS; k++) {
        const int64_t s = sidx[k];
        if (s >= N) continue;
#pragma unroll
        for (int c = 0; c < C; c++) {
            if (mask & PIGS_FN_VALUE) put(out_u + s * C + c, au[k][c]);
#pragma unroll
            for (int i = 0; i < D; i++) {
                if (mask & PIGS_FN_DERIV) put(out_ux + (s * D + i) * C + c, ax[k][i][c]);
#pragma unroll
                for (int j = 0; j < D; j++)
                    if (mask & PIGS_FN_HESS) put(out_uxx + ((s * D + i) * D + j) * C + c, ah[k][i][j][c]);
            }
        }
    }
}

// ---- 2-D, one channel, large N: the tiled forward kernel's recipe (eval_fwd.cu) on the dense pair set ----------------
// EIGHT samples per thread as four register pairs; every Gaussian is staged ONCE per CTA as three float4
// {mu, Ks = kappa K (4 entries, not assumed symmetric), Ms = kappa^2 K, v} with kappa = -1/2 log2 e folded in, so that
// log2(alpha) = X . (Ks X) needs no scaling multiply and the record costs 3 broadcast LDS.128 per 8 pairs (the scalar
// kernel above: 7 LDS.32 per 4 pairs, LSU pipe at 40 % next to the FMAs); geometry written op-major in packed FP32x2
// over sample pairs with the Gaussian's parameters as scalar-broadcast operands, accumulation sample-major with the
// pair's weight in one operand slot.  19 FP32 lane-ops + 1 ex2 per pair for u + grad u + the four Hessian entries
// (H10 is carried as H01 plus the sum of w (Ms01 - Ms10), which vanishes for symmetric conics).
constexpr int DN2_KS = 8, DN2_REC4 = 3;
static_assert(DN_BATCH == DN_FWD_THREADS, "dense_fwd2_kernel stages one Gaussian per thread");
static int g_dense_ks = 0, g_dense_kg = 0;  // pigs_set_dense_slots: 0 = chosen from the problem size

__global__ void __launch_bounds__(DN_FWD_THREADS, 2)
    dense_fwd2_kernel(unsigned mask, int64_t n, int64_t N, int64_t per_chunk, const float* __restrict__ means,
                      const float* __restrict__ conics, const float* __restrict__ values, const float* __restrict__ samples,
                      float* __restrict__ out_u, float* __restrict__ out_ux, float* __restrict__ out_uxx)
{
    constexpr int KS = DN2_KS, HP = KS / 2;
    __shared__ float4 rec[DN_BATCH * DN2_REC4];
    int64_t sidx[KS];
    f32x2 xp0[HP], xp1[HP];
    float acc[KS][7];
    {
        float x0[KS], x1[KS];
#pragma unroll
        for (int k = 0; k < KS; k++) {
            sidx[k] = ((int64_t)blockIdx.x * KS + k) * DN_FWD_THREADS + threadIdx.x;
            x0[k] = sidx[k] < N ? samples[sidx[k] * 2] : 0.0f;
            x1[k] = sidx[k] < N ? samples[sidx[k] * 2 + 1] : 0.0f;
#pragma unroll
            for (int a = 0; a < 7; a++) acc[k][a] = 0.0f;
        }
#pragma unroll
        for (int p = 0; p < HP; p++) {
            xp0[p] = pk2(x0[2 * p], x0[2 * p + 1]);
            xp1[p] = pk2(x1[2 * p], x1[2 * p + 1]);
        }
    }
    const int64_t g0 = (int64_t)blockIdx.y * per_chunk, g1 = min(n, g0 + per_chunk);
    for (int64_t b = g0; b < g1; b += DN_BATCH) {
        const int nb = (int)min((int64_t)DN_BATCH, g1 - b);
        __syncthreads();
        if ((int)threadIdx.x < nb) {  // DN_BATCH == DN_FWD_THREADS: one Gaussian per thread
            const int64_t g = b + threadIdx.x;
            const float k00 = conics[g * 4], k01 = conics[g * 4 + 1], k10 = conics[g * 4 + 2], k11 = conics[g * 4 + 3];
            const float s00 = __fmul_rn(DN_KH, k00), s01 = __fmul_rn(DN_KH, k01), s10 = __fmul_rn(DN_KH, k10),
                        s11 = __fmul_rn(DN_KH, k11);
            rec[threadIdx.x * DN2_REC4] = make_float4(means[g * 2], means[g * 2 + 1], s00, s01);
            rec[threadIdx.x * DN2_REC4 + 1] = make_float4(s10, s11, __fmul_rn(DN_KH, s00), __fmul_rn(DN_KH, s01));
            // third slot: Ms01 - Ms10 (zero for symmetric conics): H10 = H01 + sum w (Ms01 - Ms10) saves a factor per pair
            const float m01 = __fmul_rn(DN_KH, s01), m10 = __fmul_rn(DN_KH, s10);
            rec[threadIdx.x * DN2_REC4 + 2] = make_float4(__fsub_rn(m01, m10), __fmul_rn(DN_KH, s11), values[g], 0.0f);
        }
        __syncthreads();
#pragma unroll 1
        for (int g = 0; g < nb; g++) {
            const float4 r0 = rec[g * DN2_REC4], r1 = rec[g * DN2_REC4 + 1], r2 = rec[g * DN2_REC4 + 2];
            // x = sample - mean (gaussians.py:50), a' = Ks x, e = x . a' = log2(alpha)
            f32x2 X0[HP], X1[HP], A0[HP], A1[HP], E[HP];
#pragma unroll
            for (int p = 0; p < HP; p++) X0[p] = add2(xp0[p], bc2(-r0.x));
#pragma unroll
            for (int p = 0; p < HP; p++) X1[p] = add2(xp1[p], bc2(-r0.y));
#pragma unroll
            for (int p = 0; p < HP; p++) A0[p] = mul2(bc2(r0.z), X0[p]);
#pragma unroll
            for (int p = 0; p < HP; p++) A1[p] = mul2(bc2(r1.x), X0[p]);
#pragma unroll
            for (int p = 0; p < HP; p++) A0[p] = fma2(bc2(r0.w), X1[p], A0[p]);
#pragma unroll
            for (int p = 0; p < HP; p++) A1[p] = fma2(bc2(r1.y), X1[p], A1[p]);
#pragma unroll
            for (int p = 0; p < HP; p++) E[p] = mul2(A0[p], X0[p]);
#pragma unroll
            for (int p = 0; p < HP; p++) E[p] = fma2(A1[p], X1[p], E[p]);
            float a0[KS], a1[KS], al[KS];
#pragma unroll
            for (int p = 0; p < HP; p++) {
                upk2(A0[p], a0[2 * p], a0[2 * p + 1]);
                upk2(A1[p], a1[2 * p], a1[2 * p + 1]);
                upk2(E[p], al[2 * p], al[2 * p + 1]);
            }
#pragma unroll
            for (int k = 0; k < KS; k++) al[k] = dn_ex2(al[k]);
#pragma unroll
            for (int k = 0; k < KS; k++) {
                const float w = r2.z * al[k];
                const float f00 = fmaf(a0[k], a0[k], -r1.z), f01 = fmaf(a0[k], a1[k], -r1.w);
                const float f11 = fmaf(a1[k], a1[k], -r2.y);
                acc[k][0] += w;
                acc[k][1] = fmaf(w, a0[k], acc[k][1]);
                acc[k][2] = fmaf(w, a1[k], acc[k][2]);
                acc[k][3] = fmaf(w, f00, acc[k][3]);
                acc[k][4] = fmaf(w, f01, acc[k][4]);
                acc[k][5] = fmaf(w, r2.x, acc[k][5]);  // sum w (Ms01 - Ms10)
                acc[k][6] = fmaf(w, f11, acc[k][6]);
            }
        }
    }
    const bool add = gridDim.y > 1;
    auto put = [add](float* p, float v) {
        if (add)
            atomicAdd(p, v);
        else
            *p = v;
    };
    // undo the kappa scaling: grad u = -(1/kappa) sum w a', Hessian = (1/kappa^2) sum w (a' a'^T - Ms)
    constexpr float IK1 = -1.0f / DN_KH, IK2 = 1.0f / (DN_KH * DN_KH);
#pragma unroll
    for (int k = 0; k < KS; k++) {
        const int64_t s = sidx[k];
        if (s >= N) continue;
        if (mask & PIGS_FN_VALUE) put(out_u + s, acc[k][0]);
        if (mask & PIGS_FN_DERIV) {
            put(out_ux + s * 2, IK1 * acc[k][1]);
            put(out_ux + s * 2 + 1, IK1 * acc[k][2]);
        }
        if (mask & PIGS_FN_HESS) {
            put(out_uxx + s * 4, IK2 * acc[k][3]);
            put(out_uxx + s * 4 + 1, IK2 * acc[k][4]);
            put(out_uxx + s * 4 + 2, IK2 * (acc[k][4] + acc[k][5]));
            put(out_uxx + s * 4 + 3, IK2 * acc[k][6]);
        }
    }
}

// per pair, with S_c = dL/dw_c (w_c = alpha v_c):
//   S_c   = g_c - sum_i gx_ic a_i + sum_ij gH_ijc (a_i a_j - K_ij)
//   dv_c += alpha S_c ;  dq = -1/2 alpha sum_c v_c S_c
//   dA_i  = sum_c w_c (-gx_ic + sum_j (gH_ijc + gH_jic) a_j) + dq X_i          (a = K X, q = X . a)
//   dK_ij += dA_i X_j - sum_c w_c gH_ijc ;  dX_j = dq a_j + sum_i dA_i K_ij ;  dmean -= dX
template <int D, int C, int KG>
__global__ void __launch_bounds__(DN_BWD_THREADS)
    dense_bwd_kernel(unsigned mask, int64_t n, int64_t N, int64_t per_chunk, const float* __restrict__ means,
                     const float* __restrict__ conics, const float* __restrict__ values, const float* __restrict__ samples,
                     const float* __restrict__ g_u, const float* __restrict__ g_ux, const float* __restrict__ g_uxx,
                     float* __restrict__ d_means, float* __restrict__ d_conics, float* __restrict__ d_values)
{
    constexpr int REC = D + C + D * C + D * D * C;  // sample, g, gx, gH
    constexpr int SB = 128;
    __shared__ float rec[SB * REC];
    // KG Gaussians per thread (strided by the block size): one read of a sample record serves KG pairs
    int64_t gidx[KG];
    float mu[KG][D], K[KG][D][D], v[KG][C];
    float dmu[KG][D], dK[KG][D][D], dv[KG][C];
#pragma unroll
    for (int t = 0; t < KG; t++) {
        gidx[t] = ((int64_t)blockIdx.x * KG + t) * DN_BWD_THREADS + threadIdx.x;
        const bool live = gidx[t] < n;
#pragma unroll
        for (int i = 0; i < D; i++) {
            mu[t][i] = live ? means[gidx[t] * D + i] : 0.0f;
            dmu[t][i] = 0.0f;
#pragma unroll
            for (int j = 0; j < D; j++) {
                K[t][i][j] = live ? conics[(gidx[t] * D + i) * D + j] : 0.0f;
                dK[t][i][j] = 0.0f;
            }
        }
#pragma unroll
        for (int c = 0; c < C; c++) {
            v[t][c] = live ? values[gidx[t] * C + c] : 0.0f;
            dv[t][c] = 0.0f;
        }
    }
    const int64_t s0 = (int64_t)blockIdx.y * per_chunk, s1 = min(N, s0 + per_chunk);
    for (int64_t b = s0; b < s1; b += SB) {
        const int nb = (int)min((int64_t)SB, s1 - b);
        __syncthreads();
        for (int i = threadIdx.x; i < nb * D; i += DN_BWD_THREADS) rec[(i / D) * REC + i % D] = samples[b * D + i];
        for (int i = threadIdx.x; i < nb * C; i += DN_BWD_THREADS)
            rec[(i / C) * REC + D + i % C] = (mask & PIGS_FN_VALUE) ? g_u[b * C + i] : 0.0f;
        for (int i = threadIdx.x; i < nb * D * C; i += DN_BWD_THREADS)
            rec[(i / (D * C)) * REC + D + C + i % (D * C)] = (mask & PIGS_FN_DERIV) ? g_ux[b * D * C + i] : 0.0f;
        for (int i = threadIdx.x; i < nb * D * D * C; i += DN_BWD_THREADS)
            rec[(i / (D * D * C)) * REC + D + C + D * C + i % (D * D * C)] =
                (mask & PIGS_FN_HESS) ? g_uxx[b * D * D * C + i] : 0.0f;
        __syncthreads();
        for (int k = 0; k < nb; k++) {
            float r[REC];
#pragma unroll
            for (int i = 0; i < REC; i++) r[i] = rec[k * REC + i];
            const float* gu = r + D;
            const float* gx = r + D + C;          // [i][c]
            const float* gh = r + D + C + D * C;  // [i][j][c]
#pragma unroll
            for (int t = 0; t < KG; t++) {
                float X[D], a[D];
#pragma unroll
                for (int d = 0; d < D; d++) X[d] = r[d] - mu[t][d];
                float q = 0.0f;
#pragma unroll
                for (int i = 0; i < D; i++) {
                    a[i] = 0.0f;
#pragma unroll
                    for (int j = 0; j < D; j++) a[i] = fmaf(K[t][i][j], X[j], a[i]);
                    q = fmaf(X[i], a[i], q);
                }
                const float alpha = dn_ex2(DN_KH * q);
                float dalpha = 0.0f, dA[D], wH[D][D];
#pragma unroll
                for (int i = 0; i < D; i++) {
                    dA[i] = 0.0f;
#pragma unroll
                    for (int j = 0; j < D; j++) wH[i][j] = 0.0f;
                }
#pragma unroll
                for (int c = 0; c < C; c++) {
                    float S = gu[c];
                    const float w = alpha * v[t][c];
#pragma unroll
                    for (int i = 0; i < D; i++) {
                        S = fmaf(-gx[i * C + c], a[i], S);
                        dA[i] = fmaf(-gx[i * C + c], w, dA[i]);
#pragma unroll
                        for (int j = 0; j < D; j++) {
                            const float h = gh[(i * D + j) * C + c];
                            S = fmaf(h, fmaf(a[i], a[j], -K[t][i][j]), S);
                            wH[i][j] = fmaf(w, h, wH[i][j]);
                        }
                    }
                    dv[t][c] = fmaf(alpha, S, dv[t][c]);
                    dalpha = fmaf(v[t][c], S, dalpha);
                }
                const float dq = -0.5f * alpha * dalpha;
#pragma unroll
                for (int i = 0; i < D; i++) {
#pragma unroll
                    for (int j = 0; j < D; j++) dA[i] = fmaf(wH[i][j] + wH[j][i], a[j], dA[i]);
                    dA[i] = fmaf(dq, X[i], dA[i]);
                }
#pragma unroll
                for (int j = 0; j < D; j++) {
                    float dX = dq * a[j];
#pragma unroll
                    for (int i = 0; i < D; i++) {
                        dX = fmaf(dA[i], K[t][i][j], dX);
                        dK[t][i][j] += fmaf(dA[i], X[j], -wH[i][j]);
                    }
                    dmu[t][j] -= dX;
                }
            }
        }
    }
    const bool add = gridDim.y > 1;
    auto put = [add](float* p, float val) {
        if (add)
            atomicAdd(p, val);
        else
            *p = val;
    };
#pragma unroll
    for (int t = 0; t < KG; t++) {
        const int64_t g = gidx[t];
        if (g >= n) continue;
#pragma unroll
        for (int i = 0; i < D; i++) {
            put(d_means + g * D + i, dmu[t][i]);
#pragma unroll
            for (int j = 0; j < D; j++) put(d_conics + (g * D + i) * D + j, dK[t][i][j]);
        }
#pragma unroll
        for (int c = 0; c < C; c++) put(d_values + g * C + c, dv[t][c]);
    }
}


// Chunks of the streamed operand (grid.y).  Two aims: enough CTAs to fill the machine even for small stationary counts
// (~4 per SM), and WHOLE WAVES: every CTA of these kernels does the same work, so nx * ny CTAs on `slots` resident ones
// take ceil(nx * ny / slots) rounds — 489 x 2 CTAs on 296 slots are 3.3 waves paid as 4 (measured: 76 ms at 100K x 1M
// where the tiled kernel's pair rate predicts 63).  Among the chunk counts from "fills the machine" up to a few more, take
// the one that wastes the least of its last wave (ties: fewer chunks = fewer output atomics).
static int64_t dn_chunks(int64_t stationary_blocks, int64_t streamed, int batch, int64_t slots)
{
    // at least ~4 CTAs per SM, and about eight waves where the streamed operand is long enough: the CTAs do not run in
    // lockstep, and the finer the last wave's CTAs the shorter the tail (100K x 1M forward: 2 / 3 / 6 / 12 chunks ->
    // 76.0 / 72.3 / 72.1 / 71.9 ms; forward + backward 325.6 / 290.8 / 284.4 / 281.0 ms)
    int64_t want = 4 * (int64_t)num_sms();
    if (8 * slots > want) want = 8 * slots;
    int64_t ny0 = (want + stationary_blocks - 1) / stationary_blocks;
    int64_t most = (streamed + 4 * batch - 1) / (4 * batch);  // at least four batches per chunk
    if (most > 65535) most = 65535;
    if (most < 1) most = 1;
    if (ny0 > most) ny0 = most;
    if (ny0 < 1) ny0 = 1;
    if (const char* e = getenv("PIGS_DENSE_NY")) {  // tuning hook
        const int64_t v = atoll(e);
        if (v >= 1) return v < most ? v : most;
    }
    if (slots <= 0) return ny0;
    int64_t best = ny0;
    double best_eff = -1.0;
    for (int64_t ny = ny0; ny <= most && ny <= ny0 + 4; ny++) {
        const double w = (double)(stationary_blocks * ny) / (double)slots;
        const double eff = w / (double)((int64_t)(w + 0.999999));
        if (eff > best_eff + 0.01) {  // a later candidate must be clearly better
            best_eff = eff;
            best = ny;
        }
    }
    return best;
}

// resident CTAs of a kernel on this device (occupancy x SMs)
template <typename K>
static int64_t dn_slots(K kernel, int threads)
{
    int nb = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, threads, 0) != cudaSuccess || nb <= 0) return 0;
    return (int64_t)nb * num_sms();
}

using DnFwdKernel = void (*)(unsigned, int64_t, int64_t, int64_t, const float*, const float*, const float*, const float*, float*,
                             float*, float*);
using DnBwdKernel = void (*)(unsigned, int64_t, int64_t, int64_t, const float*, const float*, const float*, const float*,
                             const float*, const float*, const float*, float*, float*, float*);

template <int D>
static int dn_fwd(int C, unsigned mask, int64_t n, int64_t N, const float* means, const float* conics, const float* values,
                  const float* samples, float* u, float* ux, float* uxx, cudaStream_t st)
{
    // samples per thread: 8 (packed kernel) for 2-D one-channel problems with enough samples, else 4 for one channel, 2 for
    // two, 1 beyond (register budget); one where N is small anyway
    int ks = (N < (int64_t)DN_FWD_THREADS * 4 * num_sms()) ? 1 : (C == 1 ? 4 : (C == 2 ? 2 : 1));
    if (D == 2 && C == 1 && N >= (int64_t)DN_FWD_THREADS * DN2_KS * num_sms()) ks = DN2_KS;
    if (g_dense_ks > 0) ks = g_dense_ks == 1 ? 1 : (g_dense_ks >= DN2_KS && D == 2 && C == 1 ? DN2_KS : (C == 1 ? 4 : (C == 2 ? 2 : 1)));
    DnFwdKernel kern = nullptr;
    if (D == 2 && C == 1 && ks == DN2_KS)
        kern = dense_fwd2_kernel;
    else if (C == 1 && ks == 4)
        kern = dense_fwd_kernel<D, 1, 4>;
    else if (C == 1)
        kern = dense_fwd_kernel<D, 1, 1>;
    else if (C == 2 && ks == 2)
        kern = dense_fwd_kernel<D, 2, 2>;
    else if (C == 2)
        kern = dense_fwd_kernel<D, 2, 1>;
    else if (C == 3)
        kern = dense_fwd_kernel<D, 3, 1>;
    else if (C == 4)
        kern = dense_fwd_kernel<D, 4, 1>;
    else
        return -1;
    const int64_t nx = (N + (int64_t)DN_FWD_THREADS * ks - 1) / ((int64_t)DN_FWD_THREADS * ks);
    const int64_t ny = dn_chunks(nx, n, DN_BATCH, dn_slots(kern, DN_FWD_THREADS));
    const int64_t per = ((n + ny - 1) / ny + DN_BATCH - 1) / DN_BATCH * DN_BATCH;
    const dim3 grid((unsigned)nx, (unsigned)((n + per - 1) / per));
    if (grid.y > 1) {  // chunked: the kernel adds into zeroed outputs
        if ((mask & PIGS_FN_VALUE) && cudaMemsetAsync(u, 0, sizeof(float) * N * C, st) != cudaSuccess) return -2;
        if ((mask & PIGS_FN_DERIV) && cudaMemsetAsync(ux, 0, sizeof(float) * N * D * C, st) != cudaSuccess) return -2;
        if ((mask & PIGS_FN_HESS) && cudaMemsetAsync(uxx, 0, sizeof(float) * N * D * D * C, st) != cudaSuccess) return -2;
    }
    kern<<<grid, DN_FWD_THREADS, 0, st>>>(mask, n, N, per, means, conics, values, samples, u, ux, uxx);
    return 0;
}

template <int D>
static int dn_bwd(int C, unsigned mask, int64_t n, int64_t N, const float* means, const float* conics, const float* values,
                  const float* samples, const float* gu, const float* gux, const float* guxx, float* dm, float* dc, float* dv,
                  cudaStream_t st)
{
    int kg = (C == 1 && n >= (int64_t)DN_BWD_THREADS * 2 * num_sms()) ? 2 : 1;  // Gaussians per thread
    if (g_dense_kg > 0) kg = (g_dense_kg > 1 && C == 1) ? 2 : 1;
    DnBwdKernel kern = nullptr;
    if (C == 1 && kg == 2)
        kern = dense_bwd_kernel<D, 1, 2>;
    else if (C == 1)
        kern = dense_bwd_kernel<D, 1, 1>;
    else if (C == 2)
        kern = dense_bwd_kernel<D, 2, 1>;
    else if (C == 3)
        kern = dense_bwd_kernel<D, 3, 1>;
    else if (C == 4)
        kern = dense_bwd_kernel<D, 4, 1>;
    else
        return -1;
    const int64_t nx = (n + (int64_t)DN_BWD_THREADS * kg - 1) / ((int64_t)DN_BWD_THREADS * kg);
    const int64_t ny = dn_chunks(nx, N, 128, dn_slots(kern, DN_BWD_THREADS));
    const int64_t per = ((N + ny - 1) / ny + 127) / 128 * 128;
    const dim3 grid((unsigned)nx, (unsigned)((N + per - 1) / per));
    if (grid.y > 1) {
        if (cudaMemsetAsync(dm, 0, sizeof(float) * n * D, st) != cudaSuccess) return -2;
        if (cudaMemsetAsync(dc, 0, sizeof(float) * n * D * D, st) != cudaSuccess) return -2;
        if (cudaMemsetAsync(dv, 0, sizeof(float) * n * C, st) != cudaSuccess) return -2;
    }
    kern<<<grid, DN_BWD_THREADS, 0, st>>>(mask, n, N, per, means, conics, values, samples, gu, gux, guxx, dm, dc, dv);
    return 0;
}

}  // namespace pigs

extern "C" {

void pigs_set_dense_slots(int samples_per_thread, int gaussians_per_thread)
{
    pigs::g_dense_ks = samples_per_thread;
    pigs::g_dense_kg = gaussians_per_thread;
}

int pigs_dense_forward(unsigned fn_mask, int64_t n, int D, int C, int64_t N, const float* means, const float* conics,
                       const float* values, const float* samples, float* out_u, float* out_ux, float* out_uxx, void* stream)
{
    using namespace pigs;
    PIGS_ARG(n > 0 && N > 0 && (D == 1 || D == 2) && C >= 1 && C <= 4);
    PIGS_ARG(fn_mask != 0 && (fn_mask & ~(PIGS_FN_VALUE | PIGS_FN_DERIV | PIGS_FN_HESS)) == 0);
    PIGS_ARG(means && conics && values && samples);
    PIGS_ARG(!(fn_mask & PIGS_FN_VALUE) || out_u);
    PIGS_ARG(!(fn_mask & PIGS_FN_DERIV) || out_ux);
    PIGS_ARG(!(fn_mask & PIGS_FN_HESS) || out_uxx);
    cudaStream_t st = (cudaStream_t)stream;
    const int rc = D == 1 ? dn_fwd<1>(C, fn_mask, n, N, means, conics, values, samples, out_u, out_ux, out_uxx, st)
                          : dn_fwd<2>(C, fn_mask, n, N, means, conics, values, samples, out_u, out_ux, out_uxx, st);
    PIGS_ARG(rc == 0);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int pigs_dense_backward(unsigned fn_mask, int64_t n, int D, int C, int64_t N, const float* means, const float* conics,
                        const float* values, const float* samples, const float* dL_du, const float* dL_dux,
                        const float* dL_duxx, float* dL_dmeans, float* dL_dconics, float* dL_dvalues, void* stream)
{
    using namespace pigs;
    PIGS_ARG(n > 0 && N > 0 && (D == 1 || D == 2) && C >= 1 && C <= 4);
    PIGS_ARG(fn_mask != 0 && (fn_mask & ~(PIGS_FN_VALUE | PIGS_FN_DERIV | PIGS_FN_HESS)) == 0);
    PIGS_ARG(means && conics && values && samples && dL_dmeans && dL_dconics && dL_dvalues);
    PIGS_ARG(!(fn_mask & PIGS_FN_VALUE) || dL_du);
    PIGS_ARG(!(fn_mask & PIGS_FN_DERIV) || dL_dux);
    PIGS_ARG(!(fn_mask & PIGS_FN_HESS) || dL_duxx);
    cudaStream_t st = (cudaStream_t)stream;
    const int rc = D == 1 ? dn_bwd<1>(C, fn_mask, n, N, means, conics, values, samples, dL_du, dL_dux, dL_duxx, dL_dmeans,
                                      dL_dconics, dL_dvalues, st)
                          : dn_bwd<2>(C, fn_mask, n, N, means, conics, values, samples, dL_du, dL_dux, dL_duxx, dL_dmeans,
                                      dL_dconics, dL_dvalues, st);
    PIGS_ARG(rc == 0);
    PIGS_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
