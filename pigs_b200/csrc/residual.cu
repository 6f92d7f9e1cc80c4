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

struct ResArgs {
    int problem, D, C;
    int64_t N;
    float nu, dt;
    const float *u, *ux, *lap, *ut;
};

// residual(s) of one sample-channel element e = n * C + c (DIFFUSION / BURGERS) or of one sample n (WAVE: two values)
__device__ __forceinline__ float res_scalar(const ResArgs& a, int64_t e)
{
    float rhs = a.lap[e];
    if (a.problem == PIGS_PDE_BURGERS) {
        const int64_t n = e / a.C;
        const int c = (int)(e - n * a.C);
        const float ux0 = a.ux[(n * a.D) * a.C + c];
        rhs = __fsub_rn(__fmul_rn(a.nu, rhs), __fmul_rn(a.u[e], ux0));
    }
    return __fsub_rn(a.ut[e], __fmul_rn(a.dt, rhs));
}

__device__ __forceinline__ void res_wave(const ResArgs& a, int64_t n, float& r0, float& r1)
{
    const float u1 = a.u[2 * n + 1];
    const float rhs0 = u1;
    const float rhs1 = __fsub_rn(__fmul_rn(10.0f, a.lap[2 * n]), __fmul_rn(0.1f, u1));
    r0 = __fsub_rn(a.ut[2 * n], __fmul_rn(a.dt, rhs0));
    r1 = __fsub_rn(a.ut[2 * n + 1], __fmul_rn(a.dt, rhs1));
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

// backward: g = dL/dloss (device scalar).  dr = 2 r g / count; chain through r = ut - dt * rhs.
__global__ void __launch_bounds__(RES_THREADS) residual_backward_kernel(ResArgs a, const float* __restrict__ g_loss,
                                                                        float inv_count, float* __restrict__ d_u,
                                                                        float* __restrict__ d_ux, float* __restrict__ d_lap,
                                                                        float* __restrict__ d_ut)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const float k = 2.0f * g_loss[0] * inv_count;
    if (a.problem == PIGS_PDE_WAVE) {
        for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < a.N; n += stride) {
            float r0, r1;
            res_wave(a, n, r0, r1);
            const float dr0 = 0.01f * k * r0, dr1 = k * r1;
            if (d_ut) {
                d_ut[2 * n] = dr0;
                d_ut[2 * n + 1] = dr1;
            }
            // r0 = ut0 - dt u1 ; r1 = ut1 - dt (10 lap0 - 0.1 u1)
            if (d_u) {
                d_u[2 * n] = 0.0f;
                d_u[2 * n + 1] = -a.dt * dr0 + 0.1f * a.dt * dr1;
            }
            if (d_lap) {
                d_lap[2 * n] = -10.0f * a.dt * dr1;
                d_lap[2 * n + 1] = 0.0f;
            }
        }
        return;
    }
    const int64_t E = a.N * a.C;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += stride) {
        const float dr = k * res_scalar(a, e);
        const float drhs = -a.dt * dr;
        if (d_ut) d_ut[e] = dr;
        if (a.problem == PIGS_PDE_BURGERS) {
            const int64_t n = e / a.C;
            const int c = (int)(e - n * a.C);
            const int64_t ix = (n * a.D) * a.C + c;
            if (d_lap) d_lap[e] = a.nu * drhs;
            if (d_u) d_u[e] = -a.ux[ix] * drhs;
            if (d_ux) {
                d_ux[ix] = -a.u[e] * drhs;
                for (int d = 1; d < a.D; d++) d_ux[ix + (int64_t)d * a.C] = 0.0f;
            }
        } else {
            if (d_lap) d_lap[e] = drhs;
            if (d_u) d_u[e] = 0.0f;
        }
    }
}

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
    PIGS_ARG(a.lap && a.ut);
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
    residual_backward_kernel<<<res_grid(work), RES_THREADS, 0, st>>>(a, dL_dloss, (float)(1.0 / (double)work), dL_du,
                                                                     dL_dux, dL_dlap, dL_dut);
    PIGS_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
