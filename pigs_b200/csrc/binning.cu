// binning.cu — cutoff-radius tile binning, bit-exact with the reference's
// (cuda_sampler/forward.cu:24-83 preprocessCUDA, auxiliary.h:21-31 getRect,
//  sampler_impl.cu:54-189 duplicateWithKeys / identifyTileRanges / sampleWithKeys,
//  sample_points.cu:70-77 tile grid) but with none of its host round trips inside the pipeline,
// no device-side new/delete, 32-bit tile keys and a hand-written stable sort.
//
// Floating-point expressions that decide a bin are spelled with explicit round-to-nearest
// intrinsics in the exact operation order of the reference's SASS, so
// nvcc cannot contract or reorder them.
#include "common.cuh"

#include <stdlib.h>
#include "layouts.cuh"
#include "worklist.cuh"

namespace pigs {

// ---------------------------------------------------------------------------------------------
// S1: bounds
// ---------------------------------------------------------------------------------------------
__global__ void bounds_init(float* mm, int D)
{
    if (threadIdx.x < D) {
        mm[threadIdx.x] = __int_as_float(0x7f800000);      // +inf
        mm[D + threadIdx.x] = __int_as_float(0xff800000);  // -inf
    }
}

template <int D>
__global__ void __launch_bounds__(256) bounds_kernel(const float* __restrict__ samples, int64_t N, float* __restrict__ mm,
                                                     bool pair_loads, bool direct)
{
    float lo[D], hi[D];
#pragma unroll
    for (int d = 0; d < D; d++) {
        lo[d] = __int_as_float(0x7f800000);
        hi[d] = __int_as_float(0xff800000);
    }
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += stride) {
        if (D == 2) {
            // 8-byte loads when the caller's tensor is 8-byte aligned (a view at an odd float offset is not)
            const float2 p = pair_loads ? reinterpret_cast<const float2*>(samples)[i]
                                        : make_float2(samples[2 * i], samples[2 * i + 1]);
            lo[0] = fminf(lo[0], p.x);
            hi[0] = fmaxf(hi[0], p.x);
            lo[D - 1] = fminf(lo[D - 1], p.y);
            hi[D - 1] = fmaxf(hi[D - 1], p.y);
        } else {
            const float p = samples[i];
            lo[0] = fminf(lo[0], p);
            hi[0] = fmaxf(hi[0], p);
        }
    }
#pragma unroll
    for (int d = 0; d < D; d++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[d] = fminf(lo[d], __shfl_xor_sync(0xffffffffu, lo[d], o));
            hi[d] = fmaxf(hi[d], __shfl_xor_sync(0xffffffffu, hi[d], o));
        }
    }
    __shared__ float slo[8][D], shi[8][D];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0)
#pragma unroll
        for (int d = 0; d < D; d++) {
            slo[w][d] = lo[d];
            shi[w][d] = hi[d];
        }
    __syncthreads();
    if (threadIdx.x < D) {
        const int d = threadIdx.x;
        float l = slo[0][d], h = shi[0][d];
        for (int i = 1; i < 8; i++) {
            l = fminf(l, slo[i][d]);
            h = fmaxf(h, shi[i][d]);
        }
        if (direct) {  // single-CTA launch (small N): no initialisation launch needed
            mm[d] = l + 0.0f;
            mm[D + d] = h + 0.0f;
        } else {
            atomic_min_float(mm + d, l + 0.0f);
            atomic_max_float(mm + D + d, h + 0.0f);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// header / tile grid (sample_points.cu:73-74)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void make_header(const float* __restrict__ mm, int D, int64_t P, int64_t N, int recip, int subdiv,
                                            float cull, pigs_header& h)
{
    memset(&h, 0, sizeof(h));
    h.magic = PIGS_MAGIC;
    h.D = D;
    h.grid[0] = h.grid[1] = 1;
    h.P = P;
    h.N = N;
    int blocks = 1;
    const float inv = __fdiv_rn(1.0f, PIGS_TILE_WIDTH);
    for (int d = 0; d < D; d++) {
        const float lo = mm[d], hi = mm[D + d];
        const float span = __fadd_rn(__fsub_rn(hi, lo), 1e-6f);
        const float q = recip ? __fmul_rn(span, inv) : __fdiv_rn(span, PIGS_TILE_WIDTH);
        int g = (int)ceilf(q);
        h.ref_grid[d] = g;
        if (subdiv > 1) g *= subdiv;  // culled mode: every reference tile cut subdiv x subdiv
        h.grid[d] = g;
        h.min_b[d] = lo;
        h.max_b[d] = hi;
        blocks *= g;
    }
    if (D == 1) h.ref_grid[1] = 1;
    h.subdiv = subdiv > 1 ? subdiv : 1;
    h.cull_sigmas = subdiv > 1 ? cull : 0.0f;
    h.tile_width = subdiv > 1 ? __fdiv_rn(PIGS_TILE_WIDTH, (float)subdiv) : PIGS_TILE_WIDTH;
    h.blocks = blocks;
    h.nslots = blocks + h.grid[0] + 2;
}

// ---------------------------------------------------------------------------------------------
// S2/S3: radius, rect
// ---------------------------------------------------------------------------------------------
template <int D>
__device__ __forceinline__ float gaussian_radius(const float* __restrict__ cov, bool& degenerate)
{
    degenerate = false;
    if (D == 1) return __fmul_rn(3.0f, __fsqrt_rn(cov[0]));
    const float c0 = cov[0], c1 = cov[1], c2 = cov[2];
    const float t = __fmul_rn(c1, c1);
    const float det = __fmaf_rn(c0, c2, -t);
    if (det == 0.0f) {
        degenerate = true;
        return 0.0f;
    }
    const float mid = __fmul_rn(0.5f, __fadd_rn(c0, c2));
    const float disc = __fmaf_rn(mid, mid, -det);
    const double root = __dsqrt_rn(fmax(1e-6, (double)disc));
    const float lambda = __double2float_rn(__dadd_rn((double)mid, root));
    return __fmul_rn(3.0f, __fsqrt_rn(lambda));
}

__device__ __forceinline__ void tile_rect(float p, float off, float r, int& lo, int& hi)
{
    const float c = __fsub_rn(p, off);
    lo = (int)floorf(__fdiv_rn(__fsub_rn(c, r), PIGS_TILE_WIDTH));
    hi = (int)ceilf(__fdiv_rn(__fadd_rn(c, r), PIGS_TILE_WIDTH));
}

// ---- culled mode (include/pigs_b200.h pigs_preprocess_count_culled) -----------------------------------
__device__ __forceinline__ int wrap_tile(int x, int g) { return x < 0 ? g + (x % g) : x % g; }

template <int D>
__device__ __forceinline__ float cull_radius(const float* __restrict__ cov, float sigmas)
{
    if (D == 1) return __fmul_rn(sigmas, __fsqrt_rn(cov[0]));
    const float c0 = cov[0], c1 = cov[1], c2 = cov[2];
    const float det = __fmaf_rn(c0, c2, -__fmul_rn(c1, c1));
    const float mid = __fmul_rn(0.5f, __fadd_rn(c0, c2));
    const float disc = __fmaf_rn(mid, mid, -det);
    const double root = __dsqrt_rn(fmax(0.0, (double)disc));
    const float lambda = __double2float_rn(__dadd_rn((double)mid, root));
    return __fmul_rn(sigmas, __fsqrt_rn(lambda));
}

// distance from m (any of its period-2 images) to the interval [a, b]
__device__ __forceinline__ float torus_interval_dist(float m, float a, float b)
{
    float best = __int_as_float(0x7f800000);
#pragma unroll
    for (int s = -1; s <= 1; s++) {
        const float ms = __fadd_rn(m, 2.0f * (float)s);
        const float dd = fmaxf(0.0f, fmaxf(__fsub_rn(a, ms), __fsub_rn(ms, b)));
        best = fminf(best, dd);
    }
    return best;
}

// one dimension of one Gaussian: the reference rect [lo, hi) in reference tiles and the test for fine column j of
// reference column x
struct CullDim {
    int lo, hi, g, k;
    float mu, min_b, rc, h;
    __device__ __forceinline__ void init(float mu_, float min_b_, float r_ref, float rc_, int g_, int k_, float h_)
    {
        mu = mu_;
        min_b = min_b_;
        rc = rc_;
        g = g_;
        k = k_;
        h = h_;
        const float c = __fsub_rn(mu, min_b);
        lo = (int)floorf(__fdiv_rn(__fsub_rn(c, r_ref), PIGS_TILE_WIDTH));
        hi = (int)ceilf(__fdiv_rn(__fadd_rn(c, r_ref), PIGS_TILE_WIDTH));
        if (hi - lo >= g) {
            lo = 0;
            hi = g;
        }
    }
    __device__ __forceinline__ bool kept(int x, int j, int& fx) const
    {
        const int tx = wrap_tile(x, g);
        if (tx >= g) return false;  // the reference's alias slot: never evaluated there either
        fx = k * tx + j;
        const float a = __fadd_rn(min_b, __fmul_rn((float)fx, h));
        const float b = __fadd_rn(min_b, __fmul_rn((float)(fx + 1), h));
        return torus_interval_dist(mu, a, b) <= rc;
    }
    __device__ __forceinline__ int count() const
    {
        int n = 0, fx;
        for (int x = lo; x < hi; x++)
            for (int j = 0; j < k; j++) n += kept(x, j, fx) ? 1 : 0;
        return n;
    }
};

// S2 / S3 of ONE Gaussian (forward.cu:44-82): radius, tile rectangle, number of tiles touched — shared by the grid-wide
// count_kernel and the single-CTA small_count_kernel
template <int D>
__device__ __forceinline__ void count_one(int64_t g, const pigs_header* hdr, const float* __restrict__ means,
                                          const float* __restrict__ cov, float* __restrict__ radii,
                                          uint32_t* __restrict__ touched, float* __restrict__ cull_r)
{
    constexpr int K = D * (D + 1) / 2;
    bool degenerate;
    const float r = gaussian_radius<D>(cov + g * K, degenerate);
    uint32_t n = 0;
    if (!degenerate) {
        int lo, hi;
        tile_rect(means[g * D], hdr->min_b[0], r, lo, hi);
        n = (uint32_t)min(hi - lo, hdr->ref_grid[0]);
        if (D == 2) {
            tile_rect(means[g * D + 1], hdr->min_b[1], r, lo, hi);
            n = (uint32_t)min(hi - lo, hdr->ref_grid[1]) * n;
        }
    }
    radii[g] = n ? r : 0.0f;  // the reference radii in both modes
    const int k = hdr->subdiv;
    if (k > 1 && n) {
        // culled mode: instances = kept fine columns x kept fine rows
        const float rc = cull_radius<D>(cov + g * K, hdr->cull_sigmas);
        cull_r[g] = rc;
        CullDim cd;
        cd.init(means[g * D], hdr->min_b[0], r, rc, hdr->ref_grid[0], k, hdr->tile_width);
        n = (uint32_t)cd.count();
        if (D == 2 && n) {
            cd.init(means[g * D + 1], hdr->min_b[1], r, rc, hdr->ref_grid[1], k, hdr->tile_width);
            n *= (uint32_t)cd.count();
        }
    }
    touched[g] = n;
}

// S1 (tile grid, sample_points.cu:73-74) + S2/S3 in ONE launch: every CTA derives the header from the sample bounds
// in shared memory (a dozen flops); block 0 publishes it for the host readback and the later kernels.
template <int D>
__global__ void __launch_bounds__(256) count_kernel(int64_t P, int64_t N, const float* __restrict__ means,
                                                    const float* __restrict__ cov, const float* __restrict__ mm, int recip,
                                                    int subdiv, float cull, pigs_header* __restrict__ hdr_out,
                                                    float* __restrict__ radii, uint32_t* __restrict__ touched,
                                                    float* __restrict__ cull_r)
{
    __shared__ pigs_header s_hdr;
    if (threadIdx.x == 0) {
        make_header(mm, D, P, N, recip, subdiv, cull, s_hdr);
        if (blockIdx.x == 0) *hdr_out = s_hdr;
    }
    __syncthreads();
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= P) return;
    count_one<D>(g, &s_hdr, means, cov, radii, touched, cull_r);
}

// culled mode: one (fine tile, id) pair per kept fine tile, id-major
template <int D, typename KT, typename VT>
__device__ __forceinline__ void emit_culled(int64_t g, float r, const float* __restrict__ means,
                                            const float* __restrict__ cull_r, const uint32_t* __restrict__ offsets,
                                            const pigs_header* __restrict__ hdr, KT* __restrict__ keys, VT* __restrict__ vals)
{
    uint32_t off = offsets[g];
    const uint32_t end = offsets[g + 1];
    if (off == end) return;
    const int k = hdr->subdiv;
    const float rc = cull_r[g], h = hdr->tile_width;
    const int fg0 = hdr->grid[0];
    CullDim cx, cy;
    cx.init(means[g * D], hdr->min_b[0], r, rc, hdr->ref_grid[0], k, h);
    if (D == 2) {
        cy.init(means[g * D + 1], hdr->min_b[1], r, rc, hdr->ref_grid[1], k, h);
        for (int y = cy.lo; y < cy.hi; y++)
            for (int jy = 0; jy < k; jy++) {
                int fy;
                if (!cy.kept(y, jy, fy)) continue;
                for (int x = cx.lo; x < cx.hi; x++)
                    for (int jx = 0; jx < k; jx++) {
                        int fx;
                        if (!cx.kept(x, jx, fx)) continue;
                        keys[off] = (KT)(fy * fg0 + fx);
                        vals[off] = (VT)g;
                        off++;
                    }
            }
    } else {
        for (int x = cx.lo; x < cx.hi; x++)
            for (int jx = 0; jx < k; jx++) {
                int fx;
                if (!cx.kept(x, jx, fx)) continue;
                keys[off] = (KT)fx;
                vals[off] = (VT)g;
                off++;
            }
    }
}

// S4 of ONE Gaussian: one (tile, id) pair per covered tile, id-major (sampler_impl.cu:54-129).  KT / VT: uint32 in the
// grid-wide kernel, narrow shared-memory types in small_bin_kernel.
template <int D, typename KT, typename VT>
__device__ __forceinline__ void emit_one(int64_t g, const float* __restrict__ means, const float* __restrict__ radii,
                                         const float* __restrict__ cull_r, const uint32_t* __restrict__ offsets,
                                         const pigs_header* __restrict__ hdr, KT* __restrict__ keys, VT* __restrict__ vals)
{
    const float r = radii[g];
    if (!(r > 0.0f)) return;
    if (hdr->subdiv > 1) {  // culled mode (grid-uniform branch)
        emit_culled<D, KT, VT>(g, r, means, cull_r, offsets, hdr, keys, vals);
        return;
    }
    uint32_t off = offsets[g];
    const int g0 = hdr->grid[0], g1 = hdr->grid[1];
    int lo0, hi0, lo1 = 0, hi1 = 1;
    tile_rect(means[g * D], hdr->min_b[0], r, lo0, hi0);
    if (hi0 - lo0 >= g0) {
        lo0 = 0;
        hi0 = g0;
    }
    if (D == 2) {
        tile_rect(means[g * D + 1], hdr->min_b[1], r, lo1, hi1);
        if (hi1 - lo1 >= g1) {
            lo1 = 0;
            hi1 = g1;
        }
    }
    for (int y = lo1; y < hi1; y++) {
        const int ty = D == 2 ? wrap_tile(y, g1) : 0;
        for (int x = lo0; x < hi0; x++) {
            const int tx = wrap_tile(x, g0);
            keys[off] = (KT)(ty * g0 + tx);
            vals[off] = (VT)g;
            off++;
        }
    }
}

template <int D>
__global__ void __launch_bounds__(256) emit_kernel(int64_t P, const float* __restrict__ means,
                                                   const float* __restrict__ radii, const float* __restrict__ cull_r,
                                                   const uint32_t* __restrict__ offsets,
                                                   const pigs_header* __restrict__ hdr, uint32_t* __restrict__ keys,
                                                   uint32_t* __restrict__ vals, float* __restrict__ bbox, int nbox,
                                                   pigs_header* __restrict__ hdr_b, pigs_header* __restrict__ hdr_s)
{
    if (blockIdx.x == 0) {
        // side jobs of the first CTA (they used to be a launch and two copies): empty sample boxes for
        // sample_key_kernel, and the header at the front of the two opaque buffers
        for (int k = threadIdx.x; k < nbox * 2 * D; k += blockDim.x)
            bbox[k] = (k % (2 * D)) < D ? __int_as_float(0x7f800000) : __int_as_float(0xff800000);
        if (threadIdx.x < (int)(sizeof(pigs_header) / 4)) {
            const uint32_t w = reinterpret_cast<const uint32_t*>(hdr)[threadIdx.x];
            reinterpret_cast<uint32_t*>(hdr_b)[threadIdx.x] = w;
            reinterpret_cast<uint32_t*>(hdr_s)[threadIdx.x] = w;
        }
    }
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= P) return;
    emit_one<D, uint32_t, uint32_t>(g, means, radii, cull_r, offsets, hdr, keys, vals);
}

// S6 of ONE sample: its tile (sampler_impl.cu:155-189; reference tile with the alias quirk, or the fine tile of the culled
// mode) — shared by sample_key_kernel and small_bin_kernel
template <int D>
__device__ __forceinline__ uint32_t sample_tile_of(const float (&x)[D], const pigs_header* __restrict__ hdr,
                                                   const float (&mn)[D], const int (&gr)[D], int g0, int k, float h, int rblocks)
{
    int t[D];
#pragma unroll
    for (int d = 0; d < D; d++) {
        const int q = (int)__fdiv_rn(__fsub_rn(x[d], mn[d]), PIGS_TILE_WIDTH);
        t[d] = min(gr[d], max(0, q));
    }
    uint32_t tile = D == 2 ? (uint32_t)t[D - 1] * (uint32_t)g0 + (uint32_t)t[0] : (uint32_t)t[0];
    if (k > 1) {
        // culled mode: the reference tile (alias quirk included) cut k x k; not a real tile -> dump slot
        const uint32_t rt = tile;
        tile = (uint32_t)hdr->blocks;
        if ((int)rt < rblocks) {
            const int par[2] = {(int)(rt % (uint32_t)g0), (int)(rt / (uint32_t)g0)};
            int f[D];
#pragma unroll
            for (int d = 0; d < D; d++) {
                const int q = (int)__fdiv_rn(__fsub_rn(x[d], mn[d]), h) - k * par[d];
                f[d] = k * par[d] + min(k - 1, max(0, q));
            }
            tile = D == 2 ? (uint32_t)f[D - 1] * (uint32_t)hdr->grid[0] + (uint32_t)f[0] : (uint32_t)f[0];
        }
    }
    return tile;
}

// S6: tile of each sample (sampler_impl.cu:155-189) + per-tile bounding box of the samples.
// SMEM = true (few tile slots, the usual 4x4 grid): the boxes are reduced in shared memory by a grid-stride CTA and
// merged with one global atomic per slot component per CTA — with uniformly scattered samples every lane of a warp
// hits a different tile, so per-sample global atomics on a few dozen addresses serialise (6.8 ms at 4M samples).
// Each CTA walks whole sort tiles (SORT_TILE_ELEMS consecutive samples) and leaves the per-tile digit histogram of the
// sort's first pass behind (ghist[digit * ntiles + tile]), so the sort starts at its scan.
template <int D, bool SMEM>
__global__ void __launch_bounds__(256) sample_key_kernel(int64_t N, const float* __restrict__ samples,
                                                         const pigs_header* __restrict__ hdr, uint32_t* __restrict__ keys,
                                                         uint32_t* __restrict__ vals, float* __restrict__ bbox /* [nslots][2*D] */,
                                                         int nslots, uint32_t* __restrict__ ghist, int64_t ntiles,
                                                         uint32_t digit_mask)
{
    extern __shared__ float s_box[];  // [nslots][2*D]
    __shared__ uint32_t s_hist[256];
    if constexpr (SMEM) {
        for (int k = threadIdx.x; k < nslots * 2 * D; k += blockDim.x)
            s_box[k] = (k % (2 * D)) < D ? __int_as_float(0x7f800000) : __int_as_float(0xff800000);
        __syncthreads();
    }
    const int g0 = hdr->ref_grid[0];
    const int k = hdr->subdiv;
    const float h = hdr->tile_width;
    const int rblocks = D == 2 ? hdr->ref_grid[0] * hdr->ref_grid[1] : hdr->ref_grid[0];
    float mn[D];
    int gr[D];
#pragma unroll
    for (int d = 0; d < D; d++) {
        mn[d] = hdr->min_b[d];
        gr[d] = hdr->ref_grid[d];
    }
    for (int64_t st = blockIdx.x; st < ntiles; st += gridDim.x) {
        s_hist[threadIdx.x] = 0;
        __syncthreads();
        for (int r = 0; r < SORT_TILE_ELEMS / 256; r++) {
            const int64_t i = st * SORT_TILE_ELEMS + r * 256 + threadIdx.x;
            const bool ok = i < N;
            uint32_t tile = 0xffffffffu;
            float x[D];
            if (ok) {
#pragma unroll
                for (int d = 0; d < D; d++) x[d] = samples[i * D + d];
                tile = sample_tile_of<D>(x, hdr, mn, gr, g0, k, h, rblocks);
                keys[i] = tile;
                vals[i] = (uint32_t)i;
            }
            // one histogram atomic per distinct tile per warp (a 4x4 grid has 16); every lane takes part in the match
            const uint32_t dg = ok ? (tile & digit_mask) : 0u;
            uint32_t peers = __ballot_sync(0xffffffffu, ok);
            for (uint32_t bm = digit_mask, b = 0; bm; bm >>= 1, b++) {  // ballots over the digit's bits (primitives.cu match_digit)
                const bool bit = (dg >> b) & 1u;
                const uint32_t m = __ballot_sync(0xffffffffu, bit);
                peers &= bit ? m : ~m;
            }
            if (ok && (threadIdx.x & 31) == (__ffs(peers) - 1)) atomicAdd(&s_hist[dg], (uint32_t)__popc(peers));
            if (ok) {
                float* box = (SMEM ? s_box : bbox) + (size_t)tile * 2 * D;
#pragma unroll
                for (int d = 0; d < D; d++) {
                    const float v = x[d] + 0.0f;  // -0 -> +0 so the integer-ordered atomics below are consistent
                    // the boxes only ever grow, so a stale read can at worst let a redundant atomic through
                    if (v < *(volatile float*)(box + d)) atomic_min_float(box + d, v);
                    if (v > *(volatile float*)(box + D + d)) atomic_max_float(box + D + d, v);
                }
            }
        }
        __syncthreads();
        ghist[(int64_t)threadIdx.x * ntiles + st] = s_hist[threadIdx.x];
        __syncthreads();
    }
    if constexpr (SMEM) {
        __syncthreads();
        for (int k = threadIdx.x; k < nslots * 2 * D; k += blockDim.x) {
            const float v = s_box[k];
            if ((k % (2 * D)) < D) {
                if (v != __int_as_float(0x7f800000)) atomic_min_float(bbox + k, v);
            } else {
                if (v != __int_as_float(0xff800000)) atomic_max_float(bbox + k, v);
            }
        }
    }
}

__global__ void bbox_init(float* bbox, int64_t nslots, int D)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nslots * 2 * D) return;
    const int k = (int)(i % (2 * D));
    bbox[i] = k < D ? __int_as_float(0x7f800000) : __int_as_float(0xff800000);
}

// S7: identifyTileRanges (sampler_impl.cu:134-151) on the sorted 32-bit tile keys
__global__ void __launch_bounds__(256) ranges_kernel(int64_t L, const uint32_t* __restrict__ keys, uint2* __restrict__ ranges,
                                                     int max_slot)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= L) return;
    const uint32_t cur = keys[i];
    if (i == 0) {
        if ((int)cur <= max_slot) ranges[cur].x = 0;
    } else {
        const uint32_t prev = keys[i - 1];
        if (cur != prev) {
            if ((int)prev <= max_slot) ranges[prev].y = (uint32_t)i;
            if ((int)cur <= max_slot) ranges[cur].x = (uint32_t)i;
        }
    }
    if (i == L - 1 && (int)cur <= max_slot) ranges[cur].y = (uint32_t)L;
}

// ---------------------------------------------------------------------------------------------
// Small problems: the whole count stage / the whole binning stage in ONE single-CTA launch each
// ---------------------------------------------------------------------------------------------
// The reference's own drivers run 100 .. 10 K Gaussians against 4 K .. 65 K points: there the twelve dependent launches of
// the preprocess chain (2 - 15 us each, launch + drain + host issue) are the step, not the arithmetic.  Below the limits
// here the chain is two launches either side of the one header readback:
//   small_count_kernel : sample bounds -> tile grid -> radii / tiles touched -> exclusive scan (+ num_rendered)
//   small_bin_kernel   : Gaussian keys -> stable sort by tile -> ranges;  sample tiles + boxes -> stable sort -> ranges;
//                        class prefix (single-class mode) -> the three work lists
// Both call the same per-element device functions as the grid-wide kernels (count_one, emit_one, sample_tile_of,
// worklist_body); the sort is a stable counting sort whose result any stable sort shares, so every output is the
// grid-wide chain's bit for bit (tests run both).  pigs_set_small_problem_pairs(0) switches both off.
constexpr int SMALL_THREADS = 1024;
constexpr int SMALL_WARPS = SMALL_THREADS / 32;
static_assert(SMALL_WARPS == 32, "small_block_scan scans the warp totals with one warp");
constexpr int64_t SMALL_P_MAX = 16384;   // Gaussian ids travel as uint16 through shared memory
constexpr int64_t SMALL_N_MAX = 16384;   // sample tiles as uint8 in shared memory; beyond ~16 K samples one CTA is slower
                                         // than the grid-wide chain (tools/small_probe.py: 100 x 40 000 0.277 vs 0.252 ms)
constexpr int64_t SMALL_R_MAX = 8192;    // instances: uint8 tile + uint16 id in shared memory (break-even near 10 K)
constexpr int SMALL_SLOTS_MAX = 256;     // tile ids (incl. alias slots) fit one byte / one radix digit

// development switch: PIGS_SMALL_FUSED=0 keeps the grid-wide chains for small problems too (tools/small_probe.py)
static bool small_fused_enabled()
{
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("PIGS_SMALL_FUSED");
        v = (e && e[0] == '0') ? 0 : 1;
    }
    return v != 0 && small_pairs_threshold() > 0;
}

// exclusive scan of one value per thread over the SMALL_THREADS threads of the CTA; *total = block sum.
// sh: SMALL_WARPS + 1 words.
__device__ __forceinline__ uint32_t small_block_scan(uint32_t v, uint32_t* total, uint32_t* sh)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    __syncthreads();  // sh may still be read from the previous call
    if (lane == 31) sh[w] = inc;
    __syncthreads();
    if (w == 0) {  // SMALL_WARPS == 32: one warp scans the warp totals
        const uint32_t c = sh[lane];
        uint32_t wi = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += t;
        }
        sh[lane] = wi - c;
        if (lane == 31) sh[SMALL_WARPS] = wi;
    }
    __syncthreads();
    *total = sh[SMALL_WARPS];
    return inc - v + sh[w];
}

template <int D>
__global__ void __launch_bounds__(SMALL_THREADS) small_count_kernel(int64_t P, int64_t N, const float* __restrict__ samples,
                                                                    const float* __restrict__ means,
                                                                    const float* __restrict__ cov, int have_bounds,
                                                                    float* __restrict__ mm /* in (have_bounds) or out */,
                                                                    int recip, int subdiv, float cull,
                                                                    pigs_header* __restrict__ hdr_out, float* __restrict__ radii,
                                                                    uint32_t* __restrict__ touched, float* __restrict__ cull_r)
{
    __shared__ pigs_header s_hdr;
    __shared__ float s_mm[4];
    __shared__ float slo[SMALL_WARPS][D], shi[SMALL_WARPS][D];
    __shared__ uint32_t s_scan[SMALL_WARPS + 1];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    // ---- S1: sample bounds (bounds_kernel's reduction: min / max are exact in any order; -0 -> +0 at the end) ----
    if (!have_bounds) {
        float lo[D], hi[D];
#pragma unroll
        for (int d = 0; d < D; d++) {
            lo[d] = __int_as_float(0x7f800000);
            hi[d] = __int_as_float(0xff800000);
        }
        for (int64_t i = tid; i < N; i += SMALL_THREADS) {
#pragma unroll
            for (int d = 0; d < D; d++) {
                const float p = samples[i * D + d];
                lo[d] = fminf(lo[d], p);
                hi[d] = fmaxf(hi[d], p);
            }
        }
#pragma unroll
        for (int d = 0; d < D; d++) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                lo[d] = fminf(lo[d], __shfl_xor_sync(0xffffffffu, lo[d], o));
                hi[d] = fmaxf(hi[d], __shfl_xor_sync(0xffffffffu, hi[d], o));
            }
        }
        if (lane == 0)
#pragma unroll
            for (int d = 0; d < D; d++) {
                slo[w][d] = lo[d];
                shi[w][d] = hi[d];
            }
        __syncthreads();
        if (tid < D) {
            float l = slo[0][tid], h = shi[0][tid];
            for (int i = 1; i < SMALL_WARPS; i++) {
                l = fminf(l, slo[i][tid]);
                h = fmaxf(h, shi[i][tid]);
            }
            s_mm[tid] = l + 0.0f;
            s_mm[D + tid] = h + 0.0f;
            mm[tid] = l + 0.0f;
            mm[D + tid] = h + 0.0f;
        }
    } else if (tid < 2 * D) {
        s_mm[tid] = mm[tid];
    }
    __syncthreads();
    // ---- tile grid ----
    if (tid == 0) {
        make_header(s_mm, D, P, N, recip, subdiv, cull, s_hdr);
        *hdr_out = s_hdr;  // num_rendered follows below (same thread)
    }
    __syncthreads();
    // ---- S2 / S3 ----
    for (int64_t g = tid; g < P; g += SMALL_THREADS) count_one<D>(g, &s_hdr, means, cov, radii, touched, cull_r);
    __syncthreads();
    // ---- exclusive scan of the tiles touched, in place; total -> touched[P] and the header word the host reads back ----
    constexpr int ITEMS = 8;
    uint32_t carry = 0;
    for (int64_t tile = 0; tile < P; tile += SMALL_THREADS * ITEMS) {
        const int64_t base = tile + (int64_t)tid * ITEMS;
        uint32_t v[ITEMS], sum = 0;
#pragma unroll
        for (int i = 0; i < ITEMS; i++) {
            v[i] = base + i < P ? touched[base + i] : 0u;
            sum += v[i];
        }
        uint32_t tot;
        uint32_t ex = small_block_scan(sum, &tot, s_scan) + carry;
#pragma unroll
        for (int i = 0; i < ITEMS; i++) {
            if (base + i < P) touched[base + i] = ex;
            ex += v[i];
        }
        carry += tot;
    }
    if (tid == 0) {
        touched[P] = carry;
        hdr_out->num_rendered = (int64_t)carry;
    }
}

// lanes of the warp whose (valid) element has the same tile as mine (primitives.cu match_digit)
__device__ __forceinline__ uint32_t small_match(uint32_t d, int nbits, bool ok)
{
    uint32_t peers = __ballot_sync(0xffffffffu, ok);
    for (int b = 0; b < nbits; b++) {
        const bool bit = (d >> b) & 1u;
        const uint32_t m = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? m : ~m;
    }
    return peers;
}

// Stable counting sort of n (tile u8, value) pairs by tile inside one CTA; value = v16[i], or i itself when v16 == NULL.
// Every warp owns a consecutive range and walks it 32 elements at a time (ranks among equal tiles from ballots), so the
// order of equal tiles is the input order.  Also writes the tile ranges [first, last + 1) of the tiles 0..max_slot
// ((0, 0) for absent ones: identifyTileRanges, sampler_impl.cu:134-151).
__device__ __forceinline__ void small_sort(const uint8_t* k8, const uint16_t* v16, int n, int nbits,
                                           uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                                           uint2* __restrict__ ranges, int max_slot, uint32_t (*wcount)[SMALL_SLOTS_MAX],
                                           uint32_t* s_scan)
{
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    for (int i = tid; i < SMALL_WARPS * SMALL_SLOTS_MAX; i += SMALL_THREADS) (&wcount[0][0])[i] = 0;
    __syncthreads();
    const int per = ((n + SMALL_THREADS - 1) / SMALL_THREADS) * 32;  // elements per warp, whole rounds of 32
    const int wb = w * per;
    for (int i = wb; i < wb + per && i < n; i += 32) {
        const int idx = i + lane;
        const bool ok = idx < n;
        const uint32_t d = ok ? k8[idx] : 0u;
        const uint32_t peers = small_match(d, nbits, ok);
        if (ok && lane == (__ffs(peers) - 1)) wcount[w][d] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    {
        const int d = tid;  // threads 0..255: one per tile id; the others take part in the scan with zeros
        uint32_t tot = 0;
        if (d < SMALL_SLOTS_MAX) {
#pragma unroll 8
            for (int ww = 0; ww < SMALL_WARPS; ww++) tot += wcount[ww][d];
        }
        uint32_t all;
        const uint32_t start = small_block_scan(tot, &all, s_scan);
        if (d < SMALL_SLOTS_MAX) {
            if (d <= max_slot) ranges[d] = tot > 0 ? make_uint2(start, start + tot) : make_uint2(0u, 0u);
            uint32_t run = start;
#pragma unroll 8
            for (int ww = 0; ww < SMALL_WARPS; ww++) {
                const uint32_t c = wcount[ww][d];
                wcount[ww][d] = run;
                run += c;
            }
        }
    }
    __syncthreads();
    for (int i = wb; i < wb + per && i < n; i += 32) {
        const int idx = i + lane;
        const bool ok = idx < n;
        const uint32_t d = ok ? k8[idx] : 0u;
        const uint32_t peers = small_match(d, nbits, ok);
        if (ok) {
            const uint32_t pos = wcount[w][d] + __popc(peers & ((1u << lane) - 1u));
            keys_out[pos] = d;
            vals_out[pos] = v16 ? (uint32_t)v16[idx] : (uint32_t)idx;
        }
        __syncwarp();
        if (ok && lane == (__ffs(peers) - 1)) wcount[w][d] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
}

template <int D>
__global__ void __launch_bounds__(SMALL_THREADS) small_bin_kernel(int64_t P, int64_t N, int blocks, int64_t R, int nbits,
                                                                  const float* __restrict__ means,
                                                                  const float* __restrict__ samples,
                                                                  const float* __restrict__ radii,
                                                                  const float* __restrict__ cull_r,
                                                                  const uint32_t* __restrict__ offsets,
                                                                  const pigs_header* __restrict__ hdr,
                                                                  pigs_header* __restrict__ hdr_b, pigs_header* __restrict__ hdr_s,
                                                                  uint32_t* __restrict__ g_tiles, uint32_t* __restrict__ g_list,
                                                                  uint2* __restrict__ ranges, uint32_t* __restrict__ s_tiles,
                                                                  uint32_t* __restrict__ s_list, uint2* __restrict__ sample_ranges,
                                                                  float* __restrict__ bbox, int nslots,
                                                                  unsigned long long* __restrict__ cls, uint32_t* __restrict__ scode,
                                                                  WorkListCfg cfg, TileSeg* __restrict__ segs,
                                                                  WorkTemplate* __restrict__ wt)
{
    extern __shared__ __align__(16) unsigned char small_raw[];  // max(3 * R, N) bytes
    __shared__ uint32_t wcount[SMALL_WARPS][SMALL_SLOTS_MAX];
    __shared__ uint32_t s_scan[SMALL_WARPS + 1];
    __shared__ float s_box[SMALL_SLOTS_MAX * 2 * D];
    const int tid = threadIdx.x;
    if (tid < (int)(sizeof(pigs_header) / 4)) {  // the header at the front of the two opaque buffers
        const uint32_t wv = reinterpret_cast<const uint32_t*>(hdr)[tid];
        reinterpret_cast<uint32_t*>(hdr_b)[tid] = wv;
        reinterpret_cast<uint32_t*>(hdr_s)[tid] = wv;
    }
    // ---- S4 / S5 / S7: Gaussian keys, stable sort by tile, ranges ----
    {
        uint16_t* v16 = reinterpret_cast<uint16_t*>(small_raw);
        uint8_t* k8 = small_raw + 2 * (size_t)R;
        for (int64_t g = tid; g < P; g += SMALL_THREADS) emit_one<D, uint8_t, uint16_t>(g, means, radii, cull_r, offsets, hdr, k8, v16);
        __syncthreads();
        small_sort(k8, v16, (int)R, nbits, g_tiles, g_list, ranges, blocks, wcount, s_scan);
    }
    // ---- S6 / S5 / S7: sample tiles + per-tile boxes, stable sort, ranges ----
    {
        uint8_t* sk8 = small_raw;
        for (int k = tid; k < nslots * 2 * D; k += SMALL_THREADS)
            s_box[k] = (k % (2 * D)) < D ? __int_as_float(0x7f800000) : __int_as_float(0xff800000);
        __syncthreads();
        const int g0 = hdr->ref_grid[0];
        const int kdiv = hdr->subdiv;
        const float h = hdr->tile_width;
        const int rblocks = D == 2 ? hdr->ref_grid[0] * hdr->ref_grid[1] : hdr->ref_grid[0];
        float mn[D];
        int gr[D];
#pragma unroll
        for (int d = 0; d < D; d++) {
            mn[d] = hdr->min_b[d];
            gr[d] = hdr->ref_grid[d];
        }
        // four samples per thread and step: their loads are issued together (one CTA has little else to hide an L2 round
        // trip per sample behind)
        constexpr int UNR = 4;
        for (int64_t i0 = tid; i0 < N; i0 += (int64_t)SMALL_THREADS * UNR) {
            float xs[UNR][D];
#pragma unroll
            for (int u = 0; u < UNR; u++) {
                const int64_t i = i0 + (int64_t)u * SMALL_THREADS;
#pragma unroll
                for (int d = 0; d < D; d++) xs[u][d] = i < N ? samples[i * D + d] : 0.0f;
            }
#pragma unroll
            for (int u = 0; u < UNR; u++) {
                const int64_t i = i0 + (int64_t)u * SMALL_THREADS;
                if (i >= N) break;
                float x[D];
#pragma unroll
                for (int d = 0; d < D; d++) x[d] = xs[u][d];
                const uint32_t tile = sample_tile_of<D>(x, hdr, mn, gr, g0, kdiv, h, rblocks);
                sk8[i] = (uint8_t)tile;
                float* box = s_box + (size_t)tile * 2 * D;
#pragma unroll
                for (int d = 0; d < D; d++) {
                    const float v = x[d] + 0.0f;  // -0 -> +0 (sample_key_kernel)
                    // the boxes only ever grow: a stale read can at worst let a redundant atomic through; without the
                    // check every sample of a tile serialises on the same two shared-memory words
                    if (v < *(volatile float*)(box + d)) atomic_min_float(box + d, v);
                    if (v > *(volatile float*)(box + D + d)) atomic_max_float(box + D + d, v);
                }
            }
        }
        __syncthreads();
        for (int k = tid; k < nslots * 2 * D; k += SMALL_THREADS) bbox[k] = s_box[k];
        small_sort(sk8, nullptr, (int)N, nbits, s_tiles, s_list, sample_ranges, blocks, wcount, s_scan);
    }
    // ---- single-class staging: the scanned class prefix in closed form (pack.cu flag_all_general_kernel) ----
    for (int64_t e = tid; e <= R; e += SMALL_THREADS) {
        cls[e] = (unsigned long long)e * CLS_GENERAL;
        if (e < R) scode[e] = 0u;
    }
    __syncthreads();
    // ---- the three work lists (pack.cu worklist_kernel runs one CTA per list) ----
    if (tid < 256) {  // worklist_body synchronises its 256 threads with a named barrier
        for (int wl = 0; wl < WL_COUNT; wl++) worklist_body(wl, blocks, ranges, sample_ranges, cls, cfg, segs, wt);
    }
}

static int bit_length(uint32_t n)
{
    int b = 0;
    while (n) {
        b++;
        n >>= 1;
    }
    return b;
}

}  // namespace pigs

using namespace pigs;

extern "C" {

int pigs_sample_bounds(const float* samples, int64_t N, int D, float* d_minmax, void* stream)
{
    PIGS_ARG(D == 1 || D == 2);
    PIGS_ARG(N > 0 && samples && d_minmax);
    cudaStream_t st = (cudaStream_t)stream;
    int64_t nb = (N + 255) / 256;
    const int cap = num_sms() * 8;
    if (nb > cap) nb = cap;
    const bool direct = N <= 64 * 1024;  // one CTA writes the result itself: one launch instead of two
    PIGS_STAGE(st, "origin");
    if (direct) {
        nb = 1;
    } else {
        bounds_init<<<1, 32, 0, st>>>(d_minmax, D);
        PIGS_LAUNCH_CHECK();
    }
    if (D == 2)
        bounds_kernel<2><<<(unsigned)nb, 256, 0, st>>>(samples, N, d_minmax, (reinterpret_cast<uintptr_t>(samples) & 7) == 0, direct);
    else
        bounds_kernel<1><<<(unsigned)nb, 256, 0, st>>>(samples, N, d_minmax, false, direct);
    PIGS_LAUNCH_CHECK();
    PIGS_STAGE(st, "bounds");
    return 0;
}

size_t pigs_geom_workspace_bytes(int64_t P) { return GeomLayout(nullptr, P).bytes; }

int pigs_preprocess_count(int64_t P, int D, int64_t N, const float* means, const float* covariances,
                          const float* d_minmax, float* radii, void* geom_ws, void* stream)
{
    return pigs_preprocess_count_culled(P, D, N, means, covariances, d_minmax, radii, geom_ws, 1, 0.0f, stream);
}

int pigs_preprocess_count_culled(int64_t P, int D, int64_t N, const float* means, const float* covariances,
                                 const float* d_minmax, float* radii, void* geom_ws, int subdiv, float cull_sigmas,
                                 void* stream)
{
    PIGS_ARG(D == 1 || D == 2);
    PIGS_ARG(P > 0 && N > 0 && means && covariances && d_minmax && radii && geom_ws);
    PIGS_ARG(subdiv >= 1 && subdiv <= 64 && (subdiv == 1 || cull_sigmas > 0.0f));
    cudaStream_t st = (cudaStream_t)stream;
    GeomLayout g(geom_ws, P);
    const unsigned nb = (unsigned)((P + 255) / 256);
    if (D == 2)
        count_kernel<2><<<nb, 256, 0, st>>>(P, N, means, covariances, d_minmax, g_recip_mode, subdiv, cull_sigmas, g.hdr, radii,
                                            g.touched, g.cull_r);
    else
        count_kernel<1><<<nb, 256, 0, st>>>(P, N, means, covariances, d_minmax, g_recip_mode, subdiv, cull_sigmas, g.hdr, radii,
                                            g.touched, g.cull_r);
    PIGS_LAUNCH_CHECK();
    PIGS_STAGE(st, "count (grid + radii + tiles touched)");
    // offsets + the total straight into header.num_rendered (the word the host reads back)
    const int rc = exclusive_scan_u32(g.touched, g.touched, P, true, g.scan_ws, st, (long long*)&g.hdr->num_rendered);
    PIGS_STAGE(st, "scan of tiles touched");
    return rc;
}

int pigs_preprocess_count_fused(int64_t P, int D, int64_t N, const float* means, const float* covariances,
                                const float* samples, float* d_minmax, int have_bounds, float* radii, void* geom_ws,
                                int subdiv, float cull_sigmas, void* stream)
{
    PIGS_ARG(D == 1 || D == 2);
    PIGS_ARG(P > 0 && N > 0 && means && covariances && d_minmax && radii && geom_ws);
    PIGS_ARG(have_bounds || samples);
    PIGS_ARG(subdiv >= 1 && subdiv <= 64 && (subdiv == 1 || cull_sigmas > 0.0f));
    if (small_fused_enabled() && P <= SMALL_P_MAX && N <= SMALL_N_MAX) {
        // small problem: bounds + grid + radii / tiles touched + scan in ONE single-CTA launch
        cudaStream_t st = (cudaStream_t)stream;
        GeomLayout g(geom_ws, P);
        PIGS_STAGE(st, "origin");
        if (D == 2)
            small_count_kernel<2><<<1, SMALL_THREADS, 0, st>>>(P, N, samples, means, covariances, have_bounds ? 1 : 0, d_minmax,
                                                                g_recip_mode, subdiv, cull_sigmas, g.hdr, radii, g.touched, g.cull_r);
        else
            small_count_kernel<1><<<1, SMALL_THREADS, 0, st>>>(P, N, samples, means, covariances, have_bounds ? 1 : 0, d_minmax,
                                                                g_recip_mode, subdiv, cull_sigmas, g.hdr, radii, g.touched, g.cull_r);
        PIGS_LAUNCH_CHECK();
        PIGS_STAGE(st, "small count (bounds + grid + radii + tiles touched + scan)");
        return 0;
    }
    if (!have_bounds) {
        const int rc = pigs_sample_bounds(samples, N, D, d_minmax, stream);
        if (rc) return rc;
    }
    return pigs_preprocess_count_culled(P, D, N, means, covariances, d_minmax, radii, geom_ws, subdiv, cull_sigmas, stream);
}

size_t pigs_binning_bytes(int64_t num_rendered, int blocks) { return BinLayout(nullptr, num_rendered, blocks).bytes; }
size_t pigs_sample_binning_bytes(int64_t N, int blocks, int D) { return SampleBinLayout(nullptr, N, blocks, D).bytes; }
size_t pigs_sort_workspace_bytes(int64_t num_rendered, int64_t N)
{
    const int64_t n = num_rendered > N ? num_rendered : N;
    return sort_ws_bytes(n) + 2 * align_up((size_t)(n > 0 ? n : 1) * sizeof(uint32_t));
}

int pigs_preprocess_bin(int64_t P, int D, int64_t N, int blocks, int64_t num_rendered, const float* means,
                        const float* conics, const float* samples, const float* radii, const void* geom_ws,
                        void* binning_buffer, void* sample_binning_buffer, void* ranges, void* sample_ranges, void* sort_ws,
                        void* stream)
{
    PIGS_ARG(D == 1 || D == 2);
    PIGS_ARG(P > 0 && N > 0 && blocks > 0 && num_rendered >= 0);
    PIGS_ARG(means && conics && samples && radii && geom_ws && binning_buffer && sample_binning_buffer && ranges &&
             sample_ranges && sort_ws);
    cudaStream_t st = (cudaStream_t)stream;
    GeomLayout g(const_cast<void*>(geom_ws), P);
    BinLayout b(binning_buffer, num_rendered, blocks);
    SampleBinLayout s(sample_binning_buffer, N, blocks, D);
    const int64_t nmax = num_rendered > N ? num_rendered : N;
    Carver c(sort_ws);
    uint32_t* keys_u = c.take<uint32_t>((size_t)(nmax > 0 ? nmax : 1));
    uint32_t* vals_u = c.take<uint32_t>((size_t)(nmax > 0 ? nmax : 1));
    char* sws = c.take<char>(sort_ws_bytes(nmax));
    // tile ids are <= blocks + grid[0] (alias slots) <= 2*blocks
    const int bits = bit_length((uint32_t)(2 * blocks + 1));
    const int64_t nslots = 2 * (int64_t)blocks + 4;

    PIGS_STAGE(st, "host readback + allocations (gap)");
    // evaluation staging mode of this binning (pack.cu): small problems put every instance into the GENERAL class
    const bool single_class = num_rendered > 0 && small_problem(num_rendered, N, blocks);
    small_binning_note(binning_buffer, num_rendered, N, blocks, single_class);
    if (single_class && small_fused_enabled() && num_rendered <= SMALL_R_MAX && N <= SMALL_N_MAX && P <= 65535 &&
        nslots <= SMALL_SLOTS_MAX) {
        // small problem: keys, both sorts, ranges, sample boxes, class prefix and work lists in ONE single-CTA launch
        const size_t need = (size_t)(3 * num_rendered > N ? 3 * num_rendered : N);
        const size_t dyn = (need + 15) & ~(size_t)15;
        const WorkListCfg cfg = make_worklist_cfg(num_rendered, N, blocks, true);
        static bool attr_set[2][64] = {};
        int dev = 0;
        PIGS_CHECK(cudaGetDevice(&dev));
        if (dev >= 0 && dev < 64 && !attr_set[D - 1][dev]) {
            if (D == 2)
                PIGS_CHECK(cudaFuncSetAttribute(small_bin_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 49152));
            else
                PIGS_CHECK(cudaFuncSetAttribute(small_bin_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 49152));
            attr_set[D - 1][dev] = true;
        }
        if (D == 2)
            small_bin_kernel<2><<<1, SMALL_THREADS, dyn, st>>>(P, N, blocks, num_rendered, bits, means, samples, radii, g.cull_r,
                                                               g.touched, g.hdr, b.hdr, s.hdr, b.tiles, b.list, (uint2*)ranges,
                                                               s.tiles, s.list, (uint2*)sample_ranges, s.bbox, (int)nslots, b.cls,
                                                               b.scode, cfg, b.segs, b.wt);
        else
            small_bin_kernel<1><<<1, SMALL_THREADS, dyn, st>>>(P, N, blocks, num_rendered, bits, means, samples, radii, g.cull_r,
                                                               g.touched, g.hdr, b.hdr, s.hdr, b.tiles, b.list, (uint2*)ranges,
                                                               s.tiles, s.list, (uint2*)sample_ranges, s.bbox, (int)nslots, b.cls,
                                                               b.scode, cfg, b.segs, b.wt);
        PIGS_LAUNCH_CHECK();
        PIGS_STAGE(st, "small bin (keys + sorts + ranges + boxes + classes + work lists)");
        return 0;
    }
    if (num_rendered > 0) {
        const unsigned nb = (unsigned)((P + 255) / 256);
        if (D == 2)
            emit_kernel<2><<<nb, 256, 0, st>>>(P, means, radii, g.cull_r, g.touched, g.hdr, keys_u, vals_u, s.bbox, (int)nslots,
                                               b.hdr, s.hdr);
        else
            emit_kernel<1><<<nb, 256, 0, st>>>(P, means, radii, g.cull_r, g.touched, g.hdr, keys_u, vals_u, s.bbox, (int)nslots,
                                               b.hdr, s.hdr);
        PIGS_LAUNCH_CHECK();
        PIGS_STAGE(st, "emit (Gaussian keys)");
        bool done = false;
        SortExtras ex;
        ex.ranges = (uint2*)ranges;
        ex.max_slot = blocks;
        ex.ranges_written = &done;
        int rc = stable_sort_pairs_u32(keys_u, vals_u, b.tiles, b.list, num_rendered, bits, sws, st, &ex);
        if (rc) return rc;
        if (!done) {
            PIGS_CHECK(cudaMemsetAsync(ranges, 0, (size_t)(blocks + 1) * sizeof(uint2), st));
            ranges_kernel<<<(unsigned)((num_rendered + 255) / 256), 256, 0, st>>>(num_rendered, b.tiles, (uint2*)ranges, blocks);
            PIGS_LAUNCH_CHECK();
        }
    } else {
        PIGS_CHECK(cudaMemcpyAsync(b.hdr, g.hdr, sizeof(pigs_header), cudaMemcpyDeviceToDevice, st));
        PIGS_CHECK(cudaMemcpyAsync(s.hdr, g.hdr, sizeof(pigs_header), cudaMemcpyDeviceToDevice, st));
        PIGS_CHECK(cudaMemsetAsync(ranges, 0, (size_t)(blocks + 1) * sizeof(uint2), st));
        bbox_init<<<(unsigned)((nslots * 2 * D + 255) / 256), 256, 0, st>>>(s.bbox, nslots, D);
        PIGS_LAUNCH_CHECK();
    }
    {
        const int64_t ntiles = sort_tiles(N);
        const size_t box_bytes = (size_t)nslots * 2 * D * sizeof(float);
        const bool smem = box_bytes <= 32 * 1024;
        const unsigned nb = (unsigned)(ntiles < (int64_t)num_sms() * 8 ? ntiles : (int64_t)num_sms() * 8);
        const uint32_t dmask = (1u << (bits >= 8 ? 8 : bits)) - 1u;
        uint32_t* gh = sort_ghist(sws);
        if (D == 2 && smem)
            sample_key_kernel<2, true><<<nb, 256, box_bytes, st>>>(N, samples, g.hdr, keys_u, vals_u, s.bbox, (int)nslots, gh, ntiles, dmask);
        else if (D == 2)
            sample_key_kernel<2, false><<<nb, 256, 0, st>>>(N, samples, g.hdr, keys_u, vals_u, s.bbox, (int)nslots, gh, ntiles, dmask);
        else if (smem)
            sample_key_kernel<1, true><<<nb, 256, box_bytes, st>>>(N, samples, g.hdr, keys_u, vals_u, s.bbox, (int)nslots, gh, ntiles, dmask);
        else
            sample_key_kernel<1, false><<<nb, 256, 0, st>>>(N, samples, g.hdr, keys_u, vals_u, s.bbox, (int)nslots, gh, ntiles, dmask);
        PIGS_LAUNCH_CHECK();
        PIGS_STAGE(st, "sample keys + boxes + histogram");
        bool done = false;
        SortExtras ex;
        ex.hist_done = true;
        ex.ranges = (uint2*)sample_ranges;
        ex.max_slot = blocks;
        ex.ranges_written = &done;
        int rc = stable_sort_pairs_u32(keys_u, vals_u, s.tiles, s.list, N, bits, sws, st, &ex);
        if (rc) return rc;
        if (!done) {
            PIGS_CHECK(cudaMemsetAsync(sample_ranges, 0, (size_t)(blocks + 1) * sizeof(uint2), st));
            ranges_kernel<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(N, s.tiles, (uint2*)sample_ranges, blocks);
            PIGS_LAUNCH_CHECK();
        }
    }
    // evaluation staging that depends on the binning only: instance classes, their scan, the three work lists
    if (num_rendered > 0) {
        int rc = stage_classes(D, blocks, num_rendered, N, means, conics, s.bbox, (const uint2*)ranges,
                               (const uint2*)sample_ranges, b, st, single_class);
        if (rc) return rc;
    }
    return 0;
}

const uint32_t* pigs_point_list(const void* binning_buffer) { return BinLayout(const_cast<void*>(binning_buffer), 1).list; }
const uint32_t* pigs_point_tiles(const void* binning_buffer, int64_t num_rendered)
{
    return BinLayout(const_cast<void*>(binning_buffer), num_rendered).tiles;
}
const uint32_t* pigs_sample_list(const void* sample_binning_buffer)
{
    return SampleBinLayout(const_cast<void*>(sample_binning_buffer), 1, 1, 1).list;
}

}  // extern "C"
