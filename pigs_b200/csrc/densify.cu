// densify.cu — split / prune of the Gaussian set as a device-side compaction.  Replaces the
// torch chains that change P between training steps and force the re-binning:
//   model_pn.py:578-610  Model.split   rows[~indices] ++ two children per split row at mu -/+ lambda_max * v_max
//                                      (torch.linalg.eig + max + gather + cat + repeat_interleave, u halved)
//   model_pn.py:703-716  prune         every per-Gaussian array indexed by keep_indices
//   test_no_mlp_1d.py:203-247          rows[keep_mask] ++ one displaced clone per split row, Adam moments of the clones 0
// as ONE plan (two masks -> the parent index of every output row) + row gathers:
//   output rows = [ rows with keep != 0, in order | m appended rows per row with extend != 0, in order, child 0..m-1 ]
// All HBM-bound: plan 2 B read + 16 B scanned per row, gather 2 * 4 W B per output row, split 20 B + 8 B per row.
#include "common.cuh"

namespace pigs {

struct CompactLayout {
    int64_t* totals;      // [2] number of kept rows, number of extended rows (device; the host reads them once)
    uint32_t* keep_off;   // [P+1] exclusive prefix of the keep flags
    uint32_t* ext_off;    // [P+1] exclusive prefix of the extend flags
    char* scan_ws;
    size_t bytes;
    CompactLayout(void* base, int64_t P)
    {
        Carver c(base);
        totals = c.take<int64_t>(2);
        keep_off = c.take<uint32_t>((size_t)P + 1);
        ext_off = c.take<uint32_t>((size_t)P + 1);
        scan_ws = c.take<char>(scan_ws_bytes(P));
        bytes = c.total();
    }
};

__global__ void __launch_bounds__(256) compact_flags_kernel(int64_t P, const uint8_t* __restrict__ keep,
                                                            const uint8_t* __restrict__ extend, uint32_t* __restrict__ kf,
                                                            uint32_t* __restrict__ ef)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    kf[i] = keep ? (keep[i] != 0) : 1u;
    ef[i] = extend ? (extend[i] != 0) : 0u;
}

__global__ void compact_totals_kernel(int64_t P, const uint32_t* __restrict__ kf, const uint32_t* __restrict__ ef,
                                      int64_t* __restrict__ totals)
{
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        totals[0] = (int64_t)kf[P];
        totals[1] = (int64_t)ef[P];
    }
}

// parent row of every output row
__global__ void __launch_bounds__(256) compact_fill_kernel(int64_t P, const uint8_t* __restrict__ keep,
                                                           const uint8_t* __restrict__ extend, int m,
                                                           const uint32_t* __restrict__ kf, const uint32_t* __restrict__ ef,
                                                           int64_t n_keep, int64_t* __restrict__ src)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    if (keep ? (keep[i] != 0) : true) src[kf[i]] = i;
    if (extend && extend[i] != 0) {
        const int64_t base = n_keep + (int64_t)ef[i] * m;
        for (int j = 0; j < m; j++) src[base + j] = i;
    }
}

// out[i, :] = in[src[i], :] * (i < n_keep ? 1 : appended_scale); one thread per element, rows of W floats
__global__ void __launch_bounds__(256) gather_rows_kernel(int64_t total, int W, int64_t n_keep,
                                                          const int64_t* __restrict__ src, const float* __restrict__ in,
                                                          float* __restrict__ out, float appended_scale)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= total) return;
    const int64_t row = e / W;
    const int col = (int)(e - row * W);
    const float v = in[src[row] * W + col];
    out[e] = row < n_keep ? v : (appended_scale == 0.0f ? 0.0f : v * appended_scale);
}

// means of the output rows of Model.split: kept rows copy; appended rows (m = 2) sit at mu -/+ lambda_max * v_max of the
// parent's covariance (model_pn.py:585-592: the eigenVALUE, not its root, times the unit eigenvector).  The sign of
// v_max is LAPACK's choice in the reference; here v = (lambda - c2, c1) or (c1, lambda - c0), whichever is longer,
// normalised — the two children of a parent are the same SET of two points either way.
template <int D>
__global__ void __launch_bounds__(256) split_means_kernel(int64_t Pout, int64_t n_keep, const int64_t* __restrict__ src,
                                                          const float* __restrict__ means, const float* __restrict__ cov,
                                                          float* __restrict__ out)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= Pout) return;
    const int64_t p = src[i];
    if (i < n_keep) {
#pragma unroll
        for (int d = 0; d < D; d++) out[i * D + d] = means[p * D + d];
        return;
    }
    const float sign = ((i - n_keep) & 1) ? 1.0f : -1.0f;  // child 0: minus, child 1: plus
    if constexpr (D == 1) {
        out[i] = fmaf(sign, fabsf(cov[p]), means[p]);
    } else {
        const float c0 = cov[3 * p], c1 = cov[3 * p + 1], c2 = cov[3 * p + 2];
        const float mid = 0.5f * (c0 + c2), hd = 0.5f * (c0 - c2);
        const float root = sqrtf(fmaf(hd, hd, c1 * c1));
        // eigenvalues mid +/- root; the reference takes the one of largest magnitude
        const float lam = (mid >= 0.0f) ? mid + root : mid - root;
        float vx, vy;
        if (fabsf(lam - c2) >= fabsf(lam - c0)) {
            vx = lam - c2;
            vy = c1;
        } else {
            vx = c1;
            vy = lam - c0;
        }
        float n = sqrtf(fmaf(vx, vx, vy * vy));
        if (n == 0.0f) {  // isotropic: every direction is an eigenvector; LAPACK returns e_x for a diagonal matrix
            vx = 1.0f;
            vy = 0.0f;
            n = 1.0f;
        }
        const float s = sign * fabsf(lam) / n;
        out[2 * i] = fmaf(s, vx, means[2 * p]);
        out[2 * i + 1] = fmaf(s, vy, means[2 * p + 1]);
    }
}

}  // namespace pigs

using namespace pigs;

extern "C" {

size_t pigs_compact_workspace_bytes(int64_t P) { return CompactLayout(nullptr, P).bytes; }

int pigs_compact_count(int64_t P, const uint8_t* keep, const uint8_t* extend, void* workspace, void* stream)
{
    PIGS_ARG(P > 0 && workspace);
    cudaStream_t st = (cudaStream_t)stream;
    CompactLayout L(workspace, P);
    const unsigned nb = (unsigned)((P + 255) / 256);
    compact_flags_kernel<<<nb, 256, 0, st>>>(P, keep, extend, L.keep_off, L.ext_off);
    PIGS_LAUNCH_CHECK();
    int rc = exclusive_scan_u32(L.keep_off, L.keep_off, P, true, L.scan_ws, st);
    if (rc) return rc;
    rc = exclusive_scan_u32(L.ext_off, L.ext_off, P, true, L.scan_ws, st);
    if (rc) return rc;
    compact_totals_kernel<<<1, 32, 0, st>>>(P, L.keep_off, L.ext_off, L.totals);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int pigs_compact_fill(int64_t P, const uint8_t* keep, const uint8_t* extend, int m, const void* workspace, int64_t n_keep,
                      int64_t* src, void* stream)
{
    PIGS_ARG(P > 0 && workspace && src && m >= 0 && n_keep >= 0);
    cudaStream_t st = (cudaStream_t)stream;
    CompactLayout L(const_cast<void*>(workspace), P);
    compact_fill_kernel<<<(unsigned)((P + 255) / 256), 256, 0, st>>>(P, keep, extend, m, L.keep_off, L.ext_off, n_keep, src);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int pigs_gather_rows(int64_t Pout, int W, int64_t n_keep, const int64_t* src, const float* in, float* out,
                     float appended_scale, void* stream)
{
    PIGS_ARG(Pout >= 0 && W >= 1 && n_keep >= 0);
    if (Pout == 0) return 0;
    PIGS_ARG(src && in && out);
    const int64_t total = Pout * W;
    gather_rows_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(total, W, n_keep, src, in, out,
                                                                                          appended_scale);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int pigs_split_means(int64_t Pout, int D, int64_t n_keep, const int64_t* src, const float* means, const float* covariances,
                     float* out_means, void* stream)
{
    PIGS_ARG(Pout >= 0 && (D == 1 || D == 2) && n_keep >= 0);
    if (Pout == 0) return 0;
    PIGS_ARG(src && means && covariances && out_means);
    const unsigned nb = (unsigned)((Pout + 255) / 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (D == 2)
        split_means_kernel<2><<<nb, 256, 0, st>>>(Pout, n_keep, src, means, covariances, out_means);
    else
        split_means_kernel<1><<<nb, 256, 0, st>>>(Pout, n_keep, src, means, covariances, out_means);
    PIGS_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
