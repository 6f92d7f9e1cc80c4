// residual.cu — PDE right-hand side + residual loss assembly fused into one HBM-bound pass each way
// (SURVEY.md §8f rank 2).  Replaces the torch chain of model_pn.py:612-625 (`pde_rhs`) and :822-834
// (`rhs = dt * pde_rhs(...)`, `pde_loss += mean((ut - rhs)**2)`) that sits directly behind
// GaussianSampler.sample_gaussians / _derivative / _laplacian in every PDE driver.  It consumes the Laplacian
// [N, C] that PIGS_FN_LAPL produces (the trace the reference takes of its [N, D, D, C] Hessian).
//
//   DIFFUSION (model_pn.py:613-614)  rhs = lap
//   BURGERS   (:616-617)             rhs = nu * lap - u * ux[:, 0]
//   WAVE      (:623-627, C == 2)     rhs = (u[:, 1],  10 * lap[:, 0] - 0.1 * u[:, 1])
//   residual  (:822-834)             r = ut - dt * rhs ;  loss = mean(r^2)   (WAVE: 0.01 * mean(r0^2) + mean(r1^2))
//
// Algorithmic bytes per element (C = 1): forward reads u, ux0, lap, ut (16 B; DIFFUSION 8 B), writes nothing but
// per-CTA partial sums; backward re-reads the same and writes the four gradients (16 B; DIFFUSION 8 B).
// The loss is reduced deterministically: per-CTA partials in a fixed order, then one CTA sums them in double.
#include "common.cuh"

namespace pigs {

constexpr int RES_THREADS = 256;

// Time integration (model_pn.py:794-811, 820): with a previous time level (u0 / ux0 / lap0 non-NULL) the kernels form
//   TRAPEZOID  x = ts * x[-1] + (1 - ts) * x[-2]   (ts [N], one value per sample; fp32 op by op like the torch line)
//   FORWARD    x = x[-2]  (ts == NULL, ts_const = 0)      BACKWARD  x = x[-1]  (ts == NULL, ts_const = 1)
//   ut = u[-1] - u[-2]
// themselves instead of reading pre-blended tensors: +12 B read per sample-channel, no [N,C] / [N,D,C] / [N,D,D,C]
// temporaries.  u / ux / lap are the current level [-1]; without a previous level `ut` is read as given.
struct ResArgs {
    int problem, D, C;
    int64_t N;
    float nu, dt;
    const float *u, *ux, *lap, *ut;
    const float *ts = nullptr;
    float ts_const = 1.0f;
    const float *u0 = nullptr, *ux0 = nullptr, *lap0 = nullptr;
};

__device__ __forceinline__ float res_ts(const ResArgs& a, int64_t n) { return a.ts ? a.ts[n] : a.ts_const; }

__device__ __forceinline__ float res_blend(float t, float x1, float x0)
{
    return __fadd_rn(__fmul_rn(t, x1), __fmul_rn(__fsub_rn(1.0f, t), x0));
}

// value of one blended input; `prev` == NULL means no previous level (the tensor is used as it is)
__device__ __forceinline__ float res_in(const float* cur, const float* prev, int64_t i, float t)
{
    return prev ? res_blend(t, cur[i], prev[i]) : cur[i];
}

__device__ __forceinline__ float res_ut(const ResArgs& a, int64_t e)
{
    return a.u0 ? __fsub_rn(a.u[e], a.u0[e]) : a.ut[e];
}

// residual(s) of one sample-channel element e = n * C + c (DIFFUSION / BURGERS) or of one sample n (WAVE: two values)
__device__ __forceinline__ float res_scalar(const ResArgs& a, int64_t e)
{
    const int64_t n = e / a.C;
    const float t = res_ts(a, n);
    float rhs = res_in(a.lap, a.lap0, e, t);
    if (a.problem == PIGS_PDE_BURGERS) {
        const int c = (int)(e - n * a.C);
        const float ux0 = res_in(a.ux, a.ux0, (n * a.D) * a.C + c, t);
        rhs = __fsub_rn(__fmul_rn(a.nu, rhs), __fmul_rn(res_in(a.u, a.u0, e, t), ux0));
    }
    return __fsub_rn(res_ut(a, e), __fmul_rn(a.dt, rhs));
}

__device__ __forceinline__ void res_wave(const ResArgs& a, int64_t n, float& r0, float& r1)
{
    const float t = res_ts(a, n);
    const float u1 = res_in(a.u, a.u0, 2 * n + 1, t);
    const float rhs0 = u1;
    const float rhs1 = __fsub_rn(__fmul_rn(10.0f, res_in(a.lap, a.lap0, 2 * n, t)), __fmul_rn(0.1f, u1));
    r0 = __fsub_rn(res_ut(a, 2 * n), __fmul_rn(a.dt, rhs0));
    r1 = __fsub_rn(res_ut(a, 2 * n + 1), __fmul_rn(a.dt, rhs1));
}

__device__ __forceinline__ double block_sum(double v)
{
    __shared__ double sh[RES_THREADS / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int w = 0; w < RES_THREADS / 32; w++) t += sh[w];
    }
    return t;  // valid on thread 0
}

// pass 1: grid-stride over the elements; optional residual write-out; one partial sum per CTA
__global__ void __launch_bounds__(RES_THREADS) residual_forward_kernel(ResArgs a, float* __restrict__ residual,
                                                                       double* __restrict__ partial)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    double acc = 0.0;
    if (a.problem == PIGS_PDE_WAVE) {
        for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < a.N; n += stride) {
            float r0, r1;
            res_wave(a, n, r0, r1);
            if (residual) {
                residual[2 * n] = r0;
                residual[2 * n + 1] = r1;
            }
            acc += 0.01 * (double)(r0 * r0) + (double)(r1 * r1);
        }
    } else {
        const int64_t E = a.N * a.C;
        for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += stride) {
            const float r = res_scalar(a, e);
            if (residual) residual[e] = r;
            acc += (double)(r * r);
        }
    }
    const double t = block_sum(acc);
    if (threadIdx.x == 0) partial[blockIdx.x] = t;
}

// pass 2: one CTA, fixed summation order -> loss (fp32 device scalar)
__global__ void __launch_bounds__(RES_THREADS) residual_finish_kernel(int nparts, const double* __restrict__ partial,
                                                                      double inv_count, float* __restrict__ loss)
{
    double acc = 0.0;
    for (int i = threadIdx.x; i < nparts; i += RES_THREADS) acc += partial[i];
    const double t = block_sum(acc);
    if (threadIdx.x == 0) *loss = (float)(t * inv_count);
}

// backward: g = dL/dloss (device scalar).  dr = 2 r g / count; chain through r = ut - dt * rhs.  With a previous time
// level the gradient of a blended input x = t x1 + (1 - t) x0 goes t : (1 - t) to the two levels and ut = u1 - u0 adds
// +dr / -dr to d_u / d_u0 (d_ut is then not an output).
struct ResGrads {
    float *d_u, *d_ux, *d_lap, *d_ut, *d_u0, *d_ux0, *d_lap0;
};

__device__ __forceinline__ void res_put(float* cur, float* prev, int64_t i, float t, float v, bool blended)
{
    if (cur) cur[i] = blended ? t * v : v;
    if (prev) prev[i] = (1.0f - t) * v;
}

__global__ void __launch_bounds__(RES_THREADS) residual_backward_kernel(ResArgs a, const float* __restrict__ g_loss,
                                                                        float inv_count, ResGrads o)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const float k = 2.0f * g_loss[0] * inv_count;
    const bool bl = a.u0 != nullptr || a.lap0 != nullptr;  // a previous level exists
    if (a.problem == PIGS_PDE_WAVE) {
        for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < a.N; n += stride) {
            float r0, r1;
            res_wave(a, n, r0, r1);
            const float t = res_ts(a, n);
            const float dr0 = 0.01f * k * r0, dr1 = k * r1;
            if (o.d_ut) {
                o.d_ut[2 * n] = dr0;
                o.d_ut[2 * n + 1] = dr1;
            }
            // r0 = ut0 - dt u1 ; r1 = ut1 - dt (10 lap0 - 0.1 u1)
            const float du1 = -a.dt * dr0 + 0.1f * a.dt * dr1;
            if (bl) {
                if (o.d_u) {
                    o.d_u[2 * n] = dr0;
                    o.d_u[2 * n + 1] = t * du1 + dr1;
                }
                if (o.d_u0) {
                    o.d_u0[2 * n] = -dr0;
                    o.d_u0[2 * n + 1] = (1.0f - t) * du1 - dr1;
                }
            } else if (o.d_u) {
                o.d_u[2 * n] = 0.0f;
                o.d_u[2 * n + 1] = du1;
            }
            res_put(o.d_lap, o.d_lap0, 2 * n, t, -10.0f * a.dt * dr1, bl);
            res_put(o.d_lap, o.d_lap0, 2 * n + 1, t, 0.0f, bl);
        }
        return;
    }
    const int64_t E = a.N * a.C;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += stride) {
        const int64_t n = e / a.C;
        const float t = res_ts(a, n);
        const float dr = k * res_scalar(a, e);
        const float drhs = -a.dt * dr;
        if (o.d_ut) o.d_ut[e] = dr;
        float du = 0.0f;  // gradient w.r.t. the (blended) u entering pde_rhs
        if (a.problem == PIGS_PDE_BURGERS) {
            const int c = (int)(e - n * a.C);
            const int64_t ix = (n * a.D) * a.C + c;
            res_put(o.d_lap, o.d_lap0, e, t, a.nu * drhs, bl);
            du = -res_in(a.ux, a.ux0, ix, t) * drhs;
            res_put(o.d_ux, o.d_ux0, ix, t, -res_in(a.u, a.u0, e, t) * drhs, bl);
            for (int d = 1; d < a.D; d++) res_put(o.d_ux, o.d_ux0, ix + (int64_t)d * a.C, t, 0.0f, bl);
        } else {
            res_put(o.d_lap, o.d_lap0, e, t, drhs, bl);
        }
        if (bl) {
            if (o.d_u) o.d_u[e] = t * du + dr;
            if (o.d_u0) o.d_u0[e] = (1.0f - t) * du - dr;
        } else if (o.d_u) {
            o.d_u[e] = du;
        }
    }
}

// ---- NAVIER_STOKES (vorticity form), model_pn.py:629-631 pde_rhs and :821,843-845 loss ----------------------
//   rhs  = nu * (wxx[:,0,0] + wxx[:,1,1]) - (u[:,0] * wx[:,0] + u[:,1] * wx[:,1])
//   loss = mean((ux[:,0,0] + ux[:,1,1])^2) + mean((wt - dt * rhs)^2)
// u [N,2] (velocity, C = 2), ux [N,2,2] = [n, d, c], wx [N,2], wxx [N,2,2] (vorticity derivatives, model_pn.py:779-781),
// wt [N].  52 B read per sample forward; backward re-reads them and writes 52 B.  Whole-row vector loads.
// With a previous time level (u0 non-NULL; model_pn.py:794-805,818) u, ux, wx, wxx are blended per sample as in ResArgs
// above and wt = w[-1] - w[-2] is formed from the two vorticity tensors passed as wt / w0.
struct NsArgs {
    int64_t N;
    float nu, dt;
    const float2* u;
    const float4* ux;
    const float2* wx;
    const float4* wxx;
    const float* wt;
    const float* ts = nullptr;
    float ts_const = 1.0f;
    const float2* u0 = nullptr;
    const float4* ux0 = nullptr;
    const float2* wx0 = nullptr;
    const float4* wxx0 = nullptr;
    const float* w0 = nullptr;
};

__device__ __forceinline__ float ns_ts(const NsArgs& a, int64_t n) { return a.ts ? a.ts[n] : a.ts_const; }

__device__ __forceinline__ void ns_sample(const NsArgs& a, int64_t n, float2& u, float2& wx, float& div, float& r)
{
    const float4 ux = a.ux[n], wxx = a.wxx[n];
    u = a.u[n];
    wx = a.wx[n];
    float d0 = ux.x, d1 = ux.w, l0 = wxx.x, l1 = wxx.w, wt = a.wt[n];
    if (a.u0) {
        const float t = ns_ts(a, n);
        const float2 up = a.u0[n], wxp = a.wx0[n];
        const float4 uxp = a.ux0[n], wxxp = a.wxx0[n];
        u = make_float2(res_blend(t, u.x, up.x), res_blend(t, u.y, up.y));
        wx = make_float2(res_blend(t, wx.x, wxp.x), res_blend(t, wx.y, wxp.y));
        d0 = res_blend(t, d0, uxp.x), d1 = res_blend(t, d1, uxp.w);
        l0 = res_blend(t, l0, wxxp.x), l1 = res_blend(t, l1, wxxp.w);
        wt = __fsub_rn(wt, a.w0[n]);
    }
    div = __fadd_rn(d0, d1);  // ux[n,0,0] + ux[n,1,1]
    const float lap = __fadd_rn(l0, l1);
    const float adv = __fadd_rn(__fmul_rn(u.x, wx.x), __fmul_rn(u.y, wx.y));
    const float rhs = __fsub_rn(__fmul_rn(a.nu, lap), adv);
    r = __fsub_rn(wt, __fmul_rn(a.dt, rhs));
}

__global__ void __launch_bounds__(RES_THREADS) ns_residual_forward_kernel(NsArgs a, float* __restrict__ residual,
                                                                          float* __restrict__ divergence,
                                                                          double* __restrict__ partial)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    double acc = 0.0;
    for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < a.N; n += stride) {
        float2 u, wx;
        float div, r;
        ns_sample(a, n, u, wx, div, r);
        if (residual) residual[n] = r;
        if (divergence) divergence[n] = div;
        acc += (double)(div * div) + (double)(r * r);
    }
    const double t = block_sum(acc);
    if (threadIdx.x == 0) partial[blockIdx.x] = t;
}

struct NsGrads {
    float2 *d_u, *d_u0;
    float4 *d_ux, *d_ux0;
    float2 *d_wx, *d_wx0;
    float4 *d_wxx, *d_wxx0;
    float *d_wt, *d_w0;  // with a previous level d_wt is dL/dw[-1] and d_w0 = -d_wt
};

__global__ void __launch_bounds__(RES_THREADS) ns_residual_backward_kernel(NsArgs a, const float* __restrict__ g_loss,
                                                                           float inv_count, NsGrads o)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const float k = 2.0f * g_loss[0] * inv_count;
    const bool bl = a.u0 != nullptr;
    for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < a.N; n += stride) {
        float2 u, wx;
        float div, r;
        ns_sample(a, n, u, wx, div, r);
        const float dr = k * r, drhs = -a.dt * dr, dd = k * div, dl = a.nu * drhs;
        const float t = bl ? ns_ts(a, n) : 1.0f, s = 1.0f - t;
        if (o.d_wt) o.d_wt[n] = dr;
        if (o.d_ux) o.d_ux[n] = make_float4(t * dd, 0.0f, 0.0f, t * dd);
        if (o.d_wxx) o.d_wxx[n] = make_float4(t * dl, 0.0f, 0.0f, t * dl);
        if (o.d_u) o.d_u[n] = make_float2(t * (-wx.x * drhs), t * (-wx.y * drhs));
        if (o.d_wx) o.d_wx[n] = make_float2(t * (-u.x * drhs), t * (-u.y * drhs));
        if (bl) {
            if (o.d_w0) o.d_w0[n] = -dr;
            if (o.d_ux0) o.d_ux0[n] = make_float4(s * dd, 0.0f, 0.0f, s * dd);
            if (o.d_wxx0) o.d_wxx0[n] = make_float4(s * dl, 0.0f, 0.0f, s * dl);
            if (o.d_u0) o.d_u0[n] = make_float2(s * (-wx.x * drhs), s * (-wx.y * drhs));
            if (o.d_wx0) o.d_wx0[n] = make_float2(s * (-u.x * drhs), s * (-u.y * drhs));
        }
    }
}

// ---- vorticity and its derivatives from the sampler's outputs at C = 2, model_pn.py:779-781 ----------------
//   w = ux[:,0,1] - ux[:,1,0]   wx = uxx[...,0,1] - uxx[...,1,0]   wxx = uxxx[...,0,1] - uxxx[...,1,0]
// i.e. with the trailing two axes (last spatial axis e, channel c): out[..] = in[.., e=0, c=1] - in[.., e=1, c=0].
// One float4 {(0,0),(0,1),(1,0),(1,1)} in -> one float out; ux, uxx, uxxx are 1, 2, 4 such groups per sample
// (112 B read + 28 B written per sample); the backward writes {0, g, -g, 0}.
__global__ void __launch_bounds__(RES_THREADS) vorticity_forward_kernel(int64_t G, const float4* __restrict__ in,
                                                                        float* __restrict__ out)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < G; i += stride) {
        const float4 v = in[i];
        out[i] = __fsub_rn(v.y, v.z);
    }
}

__global__ void __launch_bounds__(RES_THREADS) vorticity_backward_kernel(int64_t G, const float* __restrict__ g,
                                                                         float4* __restrict__ d_in)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < G; i += stride) {
        const float v = g[i];
        d_in[i] = make_float4(0.0f, v, -v, 0.0f);
    }
}

static inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
static inline bool al8(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 7) == 0; }

static int res_grid(int64_t work)
{
    int64_t nb = (work + RES_THREADS - 1) / RES_THREADS;
    const int64_t cap = (int64_t)num_sms() * 8;  // 8 resident CTAs of 256 threads per SM, grid-stride beyond
    if (nb > cap) nb = cap;
    if (nb < 1) nb = 1;
    return (int)nb;
}

static int res_check(const ResArgs& a)
{
    PIGS_ARG(a.problem == PIGS_PDE_DIFFUSION || a.problem == PIGS_PDE_BURGERS || a.problem == PIGS_PDE_WAVE);
    PIGS_ARG(a.N > 0 && a.C >= 1 && (a.D == 1 || a.D == 2));
    PIGS_ARG(a.lap && (a.ut || a.u0));
    PIGS_ARG(!a.u0 || (a.u && a.lap0));                                   // a previous level needs u, u0, lap0 ...
    PIGS_ARG(!a.u0 || a.problem != PIGS_PDE_BURGERS || a.ux0);            // ... and ux0 where ux enters
    PIGS_ARG(a.u0 || (!a.lap0 && !a.ux0 && !a.ts));
    PIGS_ARG(a.problem != PIGS_PDE_BURGERS || (a.u && a.ux));
    PIGS_ARG(a.problem != PIGS_PDE_WAVE || (a.u && a.C == 2));
    return 0;
}

}  // namespace pigs

extern "C" {

size_t pigs_pde_residual_workspace_bytes(int64_t N, int C)
{
    (void)N;
    (void)C;
    return (size_t)pigs::num_sms() * 8 * sizeof(double);
}

int pigs_pde_residual_forward(int problem, int64_t N, int D, int C, float nu, float dt, const float* u, const float* ux,
                              const float* lap, const float* ut, float* residual, float* loss, void* workspace,
                              void* stream)
{
    using namespace pigs;
    ResArgs a{problem, D, C, N, nu, dt, u, ux, lap, ut};
    int rc = res_check(a);
    if (rc) return rc;
    PIGS_ARG(loss && workspace);
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t work = problem == PIGS_PDE_WAVE ? N : N * C;
    const int nb = res_grid(work);
    residual_forward_kernel<<<nb, RES_THREADS, 0, st>>>(a, residual, (double*)workspace);
    PIGS_LAUNCH_CHECK();
    residual_finish_kernel<<<1, RES_THREADS, 0, st>>>(nb, (const double*)workspace, 1.0 / (double)work, loss);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int pigs_pde_residual_backward(int problem, int64_t N, int D, int C, float nu, float dt, const float* u, const float* ux,
                               const float* lap, const float* ut, const float* dL_dloss, float* dL_du, float* dL_dux,
                               float* dL_dlap, float* dL_dut, void* stream)
{
    using namespace pigs;
    ResArgs a{problem, D, C, N, nu, dt, u, ux, lap, ut};
    int rc = res_check(a);
    if (rc) return rc;
    PIGS_ARG(dL_dloss);
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t work = problem == PIGS_PDE_WAVE ? N : N * C;
    ResGrads o{dL_du, dL_dux, dL_dlap, dL_dut, nullptr, nullptr, nullptr};
    residual_backward_kernel<<<res_grid(work), RES_THREADS, 0, st>>>(a, dL_dloss, (float)(1.0 / (double)work), o);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int pigs_pde_residual_timelevels_forward(int problem, int64_t N, int D, int C, float nu, float dt, const float* ts,
                                         float ts_const, const float* u1, const float* u0, const float* ux1,
                                         const float* ux0, const float* lap1, const float* lap0, float* residual,
                                         float* loss, void* workspace, void* stream)
{
    using namespace pigs;
    ResArgs a{problem, D, C, N, nu, dt, u1, ux1, lap1, nullptr};
    a.ts = ts, a.ts_const = ts_const, a.u0 = u0, a.ux0 = ux0, a.lap0 = lap0;
    PIGS_ARG(u0 != nullptr);
    int rc = res_check(a);
    if (rc) return rc;
    PIGS_ARG(loss && workspace);
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t work = problem == PIGS_PDE_WAVE ? N : N * C;
    const int nb = res_grid(work);
    residual_forward_kernel<<<nb, RES_THREADS, 0, st>>>(a, residual, (double*)workspace);
    PIGS_LAUNCH_CHECK();
    residual_finish_kernel<<<1, RES_THREADS, 0, st>>>(nb, (const double*)workspace, 1.0 / (double)work, loss);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int pigs_pde_residual_timelevels_backward(int problem, int64_t N, int D, int C, float nu, float dt, const float* ts,
                                          float ts_const, const float* u1, const float* u0, const float* ux1,
                                          const float* ux0, const float* lap1, const float* lap0, const float* dL_dloss,
                                          float* dL_du1, float* dL_du0, float* dL_dux1, float* dL_dux0, float* dL_dlap1,
                                          float* dL_dlap0, void* stream)
{
    using namespace pigs;
    ResArgs a{problem, D, C, N, nu, dt, u1, ux1, lap1, nullptr};
    a.ts = ts, a.ts_const = ts_const, a.u0 = u0, a.ux0 = ux0, a.lap0 = lap0;
    PIGS_ARG(u0 != nullptr);
    int rc = res_check(a);
    if (rc) return rc;
    PIGS_ARG(dL_dloss);
    const int64_t work = problem == PIGS_PDE_WAVE ? N : N * C;
    ResGrads o{dL_du1, dL_dux1, dL_dlap1, nullptr, dL_du0, dL_dux0, dL_dlap0};
    residual_backward_kernel<<<res_grid(work), RES_THREADS, 0, (cudaStream_t)stream>>>(a, dL_dloss,
                                                                                       (float)(1.0 / (double)work), o);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int pigs_pde_ns_residual_forward(int64_t N, float nu, float dt, const float* u, const float* ux, const float* wx,
                                 const float* wxx, const float* wt, float* residual, float* divergence, float* loss,
                                 void* workspace, void* stream)
{
    using namespace pigs;
    PIGS_ARG(N > 0 && u && ux && wx && wxx && wt && loss && workspace);
    PIGS_ARG(al8(u) && al16(ux) && al8(wx) && al16(wxx));  // whole-row vector loads
    NsArgs a{N, nu, dt, (const float2*)u, (const float4*)ux, (const float2*)wx, (const float4*)wxx, wt};
    cudaStream_t st = (cudaStream_t)stream;
    const int nb = res_grid(N);
    ns_residual_forward_kernel<<<nb, RES_THREADS, 0, st>>>(a, residual, divergence, (double*)workspace);
    PIGS_LAUNCH_CHECK();
    residual_finish_kernel<<<1, RES_THREADS, 0, st>>>(nb, (const double*)workspace, 1.0 / (double)N, loss);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int pigs_pde_ns_residual_backward(int64_t N, float nu, float dt, const float* u, const float* ux, const float* wx,
                                  const float* wxx, const float* wt, const float* dL_dloss, float* dL_du, float* dL_dux,
                                  float* dL_dwx, float* dL_dwxx, float* dL_dwt, void* stream)
{
    using namespace pigs;
    PIGS_ARG(N > 0 && u && ux && wx && wxx && wt && dL_dloss);
    PIGS_ARG(al8(u) && al16(ux) && al8(wx) && al16(wxx) && al8(dL_du) && al16(dL_dux) && al8(dL_dwx) && al16(dL_dwxx));
    NsArgs a{N, nu, dt, (const float2*)u, (const float4*)ux, (const float2*)wx, (const float4*)wxx, wt};
    NsGrads o{(float2*)dL_du, nullptr, (float4*)dL_dux, nullptr, (float2*)dL_dwx, nullptr, (float4*)dL_dwxx, nullptr, dL_dwt,
              nullptr};
    ns_residual_backward_kernel<<<res_grid(N), RES_THREADS, 0, (cudaStream_t)stream>>>(a, dL_dloss, (float)(1.0 / (double)N), o);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int pigs_pde_ns_residual_timelevels_forward(int64_t N, float nu, float dt, const float* ts, float ts_const, const float* u1,
                                            const float* u0, const float* ux1, const float* ux0, const float* wx1,
                                            const float* wx0, const float* wxx1, const float* wxx0, const float* w1,
                                            const float* w0, float* residual, float* divergence, float* loss,
                                            void* workspace, void* stream)
{
    using namespace pigs;
    PIGS_ARG(N > 0 && u1 && u0 && ux1 && ux0 && wx1 && wx0 && wxx1 && wxx0 && w1 && w0 && loss && workspace);
    PIGS_ARG(al8(u1) && al8(u0) && al16(ux1) && al16(ux0) && al8(wx1) && al8(wx0) && al16(wxx1) && al16(wxx0));
    NsArgs a{N, nu, dt, (const float2*)u1, (const float4*)ux1, (const float2*)wx1, (const float4*)wxx1, w1};
    a.ts = ts, a.ts_const = ts_const, a.u0 = (const float2*)u0, a.ux0 = (const float4*)ux0, a.wx0 = (const float2*)wx0;
    a.wxx0 = (const float4*)wxx0, a.w0 = w0;
    cudaStream_t st = (cudaStream_t)stream;
    const int nb = res_grid(N);
    ns_residual_forward_kernel<<<nb, RES_THREADS, 0, st>>>(a, residual, divergence, (double*)workspace);
    PIGS_LAUNCH_CHECK();
    residual_finish_kernel<<<1, RES_THREADS, 0, st>>>(nb, (const double*)workspace, 1.0 / (double)N, loss);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int pigs_pde_ns_residual_timelevels_backward(int64_t N, float nu, float dt, const float* ts, float ts_const, const float* u1,
                                             const float* u0, const float* ux1, const float* ux0, const float* wx1,
                                             const float* wx0, const float* wxx1, const float* wxx0, const float* w1,
                                             const float* w0, const float* dL_dloss, float* dL_du1, float* dL_du0,
                                             float* dL_dux1, float* dL_dux0, float* dL_dwx1, float* dL_dwx0, float* dL_dwxx1,
                                             float* dL_dwxx0, float* dL_dw1, float* dL_dw0, void* stream)
{
    using namespace pigs;
    PIGS_ARG(N > 0 && u1 && u0 && ux1 && ux0 && wx1 && wx0 && wxx1 && wxx0 && w1 && w0 && dL_dloss);
    PIGS_ARG(al8(u1) && al8(u0) && al16(ux1) && al16(ux0) && al8(wx1) && al8(wx0) && al16(wxx1) && al16(wxx0));
    PIGS_ARG(al8(dL_du1) && al8(dL_du0) && al16(dL_dux1) && al16(dL_dux0) && al8(dL_dwx1) && al8(dL_dwx0) && al16(dL_dwxx1) && al16(dL_dwxx0));
    NsArgs a{N, nu, dt, (const float2*)u1, (const float4*)ux1, (const float2*)wx1, (const float4*)wxx1, w1};
    a.ts = ts, a.ts_const = ts_const, a.u0 = (const float2*)u0, a.ux0 = (const float4*)ux0, a.wx0 = (const float2*)wx0;
    a.wxx0 = (const float4*)wxx0, a.w0 = w0;
    NsGrads o{(float2*)dL_du1, (float2*)dL_du0, (float4*)dL_dux1, (float4*)dL_dux0, (float2*)dL_dwx1, (float2*)dL_dwx0,
              (float4*)dL_dwxx1, (float4*)dL_dwxx0, dL_dw1, dL_dw0};
    ns_residual_backward_kernel<<<res_grid(N), RES_THREADS, 0, (cudaStream_t)stream>>>(a, dL_dloss, (float)(1.0 / (double)N), o);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int pigs_vorticity_forward(int64_t groups, const float* in, float* out, void* stream)
{
    using namespace pigs;
    PIGS_ARG(groups > 0 && in && out);
    PIGS_ARG(al16(in));
    vorticity_forward_kernel<<<res_grid(groups), RES_THREADS, 0, (cudaStream_t)stream>>>(groups, (const float4*)in, out);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int pigs_vorticity_backward(int64_t groups, const float* dL_dout, float* dL_din, void* stream)
{
    using namespace pigs;
    PIGS_ARG(groups > 0 && dL_dout && dL_din);
    PIGS_ARG(al16(dL_din));
    vorticity_backward_kernel<<<res_grid(groups), RES_THREADS, 0, (cudaStream_t)stream>>>(groups, dL_dout, (float4*)dL_din);
    PIGS_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
