// covariance.cu — covariance / conic construction from (scaling, transform) and its backward, fused into one
// HBM-bound kernel each.  Replaces the torch chain of gaussians.py:163-193
// (`build_covariances`: tanh, prod, sqrt, diag_embed, two indexed assignments, batched torch.inverse, reshape +
// index) that sits directly in front of `GaussianSampler.preprocess` in every driver
// (test_no_mlp.py:99-102, model_pn.py:695-698, test_pde.py:29-35).
//
//   tau   = tanh(t) * sqrt(s0 * s1)                         gaussians.py:164
//   Sigma = [[s0, tau], [tau, s1]]                          :165,173-177
//   conic = Sigma^-1 = [[s1, -tau], [-tau, s0]] / det       :181 (closed form of the 2x2 inverse)
//   flattened [xx, xy, yy] = entries [0, 1, 3]              :186-189
// 1-D: Sigma = [s], conic = [1/s] (the flatten of :186 is 2-D only; 1-D callers use the full matrices).
//
// Algorithmic bytes: forward 12 B read + 24 B written per Gaussian, backward 36 B read + 12 B written.
#include "common.cuh"

namespace pigs {

template <int D>
__global__ void __launch_bounds__(256) cov_forward_kernel(int64_t P, const float* __restrict__ s, const float* __restrict__ t,
                                                          float* __restrict__ cov, float* __restrict__ conic)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    if constexpr (D == 1) {
        const float s0 = s[i];
        cov[i] = s0;
        conic[i] = 1.0f / s0;
    } else {
        const float s0 = s[2 * i], s1 = s[2 * i + 1];
        const float tau = tanhf(t[i]) * sqrtf(s0 * s1);
        const float det = fmaf(s0, s1, -tau * tau);
        const float inv = 1.0f / det;
        cov[3 * i] = s0;
        cov[3 * i + 1] = tau;
        cov[3 * i + 2] = s1;
        conic[3 * i] = s1 * inv;
        conic[3 * i + 1] = -tau * inv;
        conic[3 * i + 2] = s0 * inv;
    }
}

// Backward of the reference graph.  With K = Sigma^-1 and the upstream gradient of the FLATTENED conic placed at
// the matrix entries the flatten reads (Gk = [[g0, g1], [0, g2]]), autograd of torch.inverse gives
// dL/dSigma = -K^T Gk K^T; the flattened covariance reads entries [0,0], [0,1], [1,1] of Sigma, and both
// off-diagonal entries of Sigma are the same tau (gaussians.py:176-177), so
//   dL/dtau = (dL/dSigma)[0,1] + (dL/dSigma)[1,0] + gcov_xy,   dL/ds_d = (dL/dSigma)[d,d] + gcov_dd + dL/dtau * dtau/ds_d.
template <int D>
__global__ void __launch_bounds__(256) cov_backward_kernel(int64_t P, const float* __restrict__ s, const float* __restrict__ t,
                                                           const float* __restrict__ gcov, const float* __restrict__ gcon,
                                                           float* __restrict__ ds, float* __restrict__ dt)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    if constexpr (D == 1) {
        const float s0 = s[i];
        const float k = 1.0f / s0;
        ds[i] = (gcov ? gcov[i] : 0.0f) - (gcon ? gcon[i] : 0.0f) * k * k;
        if (dt) dt[i] = 0.0f;
    } else {
        const float s0 = s[2 * i], s1 = s[2 * i + 1];
        const float u = tanhf(t[i]);
        const float r = sqrtf(s0 * s1);
        const float tau = u * r;
        const float inv = 1.0f / fmaf(s0, s1, -tau * tau);
        const float k0 = s1 * inv, k1 = -tau * inv, k2 = s0 * inv;
        float d00 = 0.0f, d01 = 0.0f, d11 = 0.0f;  // dL/dSigma[0,0], [0,1] + [1,0], [1,1]
        if (gcon) {
            const float g0 = gcon[3 * i], g1 = gcon[3 * i + 1], g2 = gcon[3 * i + 2];
            // A = Gk K^T (K symmetric): rows of Gk times K
            const float a00 = g0 * k0 + g1 * k1, a01 = g0 * k1 + g1 * k2;
            const float a10 = g2 * k1, a11 = g2 * k2;
            // dSigma = -K A
            d00 = -(k0 * a00 + k1 * a10);
            d11 = -(k1 * a01 + k2 * a11);
            d01 = -(k0 * a01 + k1 * a11) - (k1 * a00 + k2 * a10);
        }
        if (gcov) {
            d00 += gcov[3 * i];
            d01 += gcov[3 * i + 1];
            d11 += gcov[3 * i + 2];
        }
        // tau = tanh(t) sqrt(s0 s1): dtau/dt = (1 - u^2) r, dtau/ds0 = u s1 / (2 r), dtau/ds1 = u s0 / (2 r)
        const float hr = 0.5f * u / r;
        ds[2 * i] = fmaf(d01, hr * s1, d00);
        ds[2 * i + 1] = fmaf(d01, hr * s0, d11);
        dt[i] = d01 * (1.0f - u * u) * r;
    }
}

}  // namespace pigs

extern "C" {

int pigs_build_covariances_forward(int64_t P, int D, const float* scaling, const float* transform, float* covariances,
                                   float* conics, void* stream)
{
    using namespace pigs;
    PIGS_ARG(P >= 0 && (D == 1 || D == 2));
    if (P == 0) return 0;
    PIGS_ARG(scaling && covariances && conics && (D == 1 || transform));
    const unsigned nb = (unsigned)((P + 255) / 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (D == 2)
        cov_forward_kernel<2><<<nb, 256, 0, st>>>(P, scaling, transform, covariances, conics);
    else
        cov_forward_kernel<1><<<nb, 256, 0, st>>>(P, scaling, transform, covariances, conics);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int pigs_build_covariances_backward(int64_t P, int D, const float* scaling, const float* transform,
                                    const float* dL_dcovariances, const float* dL_dconics, float* dL_dscaling,
                                    float* dL_dtransform, void* stream)
{
    using namespace pigs;
    PIGS_ARG(P >= 0 && (D == 1 || D == 2));
    if (P == 0) return 0;
    PIGS_ARG(scaling && dL_dscaling && (D == 1 || (transform && dL_dtransform)));
    const unsigned nb = (unsigned)((P + 255) / 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (D == 2)
        cov_backward_kernel<2><<<nb, 256, 0, st>>>(P, scaling, transform, dL_dcovariances, dL_dconics, dL_dscaling, dL_dtransform);
    else
        cov_backward_kernel<1><<<nb, 256, 0, st>>>(P, scaling, transform, dL_dcovariances, dL_dconics, dL_dscaling, dL_dtransform);
    PIGS_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
