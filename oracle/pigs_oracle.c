/*
 * pigs_oracle.c — TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C, CPU restatement of the differentiable Gaussian-mixture sampling
 * path of kr4b/pigs (`diff-gaussian-sampling`), written from the algorithm's
 * description, used as the parity checker for the CUDA product in
 * pigs_b200/csrc.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this file.  The product path
 * never links, imports or calls it.
 *
 * Parity status: PINNED.  (i) Golden dumps of the unmodified reference CUDA
 * extension (compiled by oracle/build_ref.sh, run on a B200 box by
 * tests/golden/make_refcuda_golden.py and make_refcuda_agg_golden.py ->
 * tests/golden/refcuda_*.npz, 18 recipes) are held in tests/test_oracle_cpu.py:
 * bins and neighbour lists bit for bit, values / gradients to 1e-5.  (ii) The
 * caller-side pieces are pinned to vectors produced by importing the real
 * gaussians.py / model_pn.py on the CPU (tests/golden/make_dense_golden.py,
 * make_covariance_golden.py, make_pde_golden.py, make_densify_golden.py).
 * (iii) On the GPU box the same extension is compared LIVE with the product
 * in tests/test_gpu_vs_reference.py.
 *
 * Every function cites the reference file:line it follows
 * (paths relative to /root/reference/diff-gaussian-sampling unless noted).
 *
 * Arithmetic notes.  The reference mixes float and double (unsuffixed
 * literals such as -0.5, 2.0, 3.0 promote sub-expressions to double); this
 * file mirrors the float/double type of every sub-expression but cannot
 * mirror nvcc's FMA contraction of the fp32 evaluation functors, so values
 * agree with the CUDA reference to fp32 rounding, not bit-for-bit.  The
 * BINNING arithmetic (radius, rect, tile ids) IS written with the explicit
 * fmaf()/op order seen in the reference's SASS and must agree bit-for-bit.
 * Build with -ffp-contract=off.
 */
#define _GNU_SOURCE
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif
#define TILE_W 0.51f /* config.h:18 BLOCK_SIZE */

/* ------------------------------------------------------------------------ */
/* S1 — tile grid.  sample_points.cu:70-74.                                   */
/* torch CUDA: min/max over dim 0; (max - min + 1e-6f) / 0.51f is evaluated   */
/* by ATen's CUDA div-by-CPU-scalar fast path as a multiply by fl(1/0.51f);   */
/* `recip_mode` = 1 selects that, 0 selects IEEE division (CPU torch).        */
/* ------------------------------------------------------------------------ */
void pigs_oracle_bounds(const float *samples, int64_t N, int D, float *min_b, float *max_b)
{
    for (int d = 0; d < D; d++) {
        min_b[d] = INFINITY;
        max_b[d] = -INFINITY;
    }
    for (int64_t i = 0; i < N; i++)
        for (int d = 0; d < D; d++) {
            float x = samples[i * D + d];
            if (x < min_b[d]) min_b[d] = x;
            if (x > max_b[d]) max_b[d] = x;
        }
}

int pigs_oracle_tile_grid(const float *min_b, const float *max_b, int D, int recip_mode, int *grid)
{
    int blocks = 1;
    const float inv = 1.0f / TILE_W;
    for (int d = 0; d < D; d++) {
        float span = (max_b[d] - min_b[d]) + 1e-6f;
        float q = recip_mode ? span * inv : span / TILE_W;
        grid[d] = (int)ceilf(q);
        blocks *= grid[d];
    }
    return blocks;
}

/* ------------------------------------------------------------------------ */
/* S2 — radius.  cuda_sampler/forward.cu:52-61.                               */
/* ------------------------------------------------------------------------ */
static float radius_of(int D, const float *cov, int *degenerate)
{
    *degenerate = 0;
    if (D == 1) {
        /* 3.0 * sqrt(cov[0]): double product of a float sqrt, rounded to float
         * == correctly rounded fp32 product */
        return (float)(3.0 * (double)sqrtf(cov[0]));
    }
    float t = cov[1] * cov[1];
    float det = fmaf(cov[0], cov[2], -t);
    if (det == 0.0f) {
        *degenerate = 1;
        return 0.0f;
    }
    float mid = 0.5f * (cov[0] + cov[2]);
    float disc = fmaf(mid, mid, -det);
    double root = sqrt(fmax(1e-6, (double)disc));
    float lambda = (float)((double)mid + root);
    return (float)(3.0 * (double)sqrtf(lambda));
}

/* S3 — tile rectangle, TORUS build.  cuda_sampler/auxiliary.h:21-31. */
static void rect_of(int D, const float *p, float r, const float *min_b, int *lo, int *hi)
{
    for (int d = 0; d < D; d++) {
        lo[d] = (int)floorf(((p[d] - min_b[d]) - r) / TILE_W);
        hi[d] = (int)ceilf(((p[d] - min_b[d]) + r) / TILE_W);
    }
}

static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

/* preprocessCUDA, cuda_sampler/forward.cu:24-83: radii + tiles_touched. */
void pigs_oracle_radii(int64_t P, int D, const float *means, const float *cov, const float *min_b,
                       const int *grid, float *radii, uint32_t *touched)
{
    const int K = D * (D + 1) / 2;
    for (int64_t g = 0; g < P; g++) {
        radii[g] = 0.0f;
        touched[g] = 0;
        int degenerate;
        float r = radius_of(D, cov + g * K, &degenerate);
        if (degenerate) continue;
        int lo[2], hi[2];
        rect_of(D, means + g * D, r, min_b, lo, hi);
        uint32_t n;
        if (D == 1)
            n = (uint32_t)imin(hi[0] - lo[0], grid[0]);
        else
            n = (uint32_t)imin(hi[1] - lo[1], grid[1]) * (uint32_t)imin(hi[0] - lo[0], grid[0]);
        if (n == 0) continue;
        radii[g] = r;
        touched[g] = n;
    }
}

/* wrapped tile coordinate, cuda_sampler/sampler_impl.cu:89,108-109 (C remainder,
 * including the x == -grid -> grid alias). */
static int wrap_tile(int x, int g) { return x < 0 ? g + (x % g) : x % g; }

static int cmp_u64(const void *a, const void *b)
{
    uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
    return x < y ? -1 : (x > y ? 1 : 0);
}

/* bit_length, restating getHigherMsb, cuda_sampler/sampler_impl.cu:39-50 */
static int bit_length(uint32_t n)
{
    int b = 0;
    while (n) {
        b++;
        n >>= 1;
    }
    return b;
}

/* identifyTileRanges, cuda_sampler/sampler_impl.cu:134-151 */
static void ranges_of(int64_t L, const uint64_t *keys, uint32_t *ranges /* uint2[] */)
{
    for (int64_t i = 0; i < L; i++) {
        uint32_t cur = (uint32_t)(keys[i] >> 32);
        if (i == 0)
            ranges[2 * cur] = 0;
        else {
            uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
            if (cur != prev) {
                ranges[2 * prev + 1] = (uint32_t)i;
                ranges[2 * cur] = (uint32_t)i;
            }
        }
        if (i == L - 1) ranges[2 * cur + 1] = (uint32_t)L;
    }
}

/* S4,S5,S7 — duplicateWithKeys + SortPairs + identifyTileRanges,
 * cuda_sampler/sampler_impl.cu:54-129,279-290.
 * point_list must hold sum(touched) entries; ranges holds (blocks+1) uint2, pre-zeroed here.
 * Returns num_rendered, or -1 if a key's tile id would fall outside ranges (the reference
 * writes out of bounds there; such inputs are excluded from fixtures). */
int64_t pigs_oracle_bin_gaussians(int64_t P, int D, const float *means, const float *radii, const float *min_b,
                                  const int *grid, int blocks, uint32_t *point_list, uint32_t *ranges)
{
    int64_t R = 0, cap = 1024;
    uint64_t *keys = (uint64_t *)malloc(cap * sizeof(uint64_t));
    int bad = 0;
    for (int64_t g = 0; g < P; g++) {
        if (!(radii[g] > 0.0f)) continue;
        int lo[2] = {0, 0}, hi[2] = {1, 1};
        rect_of(D, means + g * D, radii[g], min_b, lo, hi);
        for (int d = 0; d < D; d++)
            if (hi[d] - lo[d] >= grid[d]) {
                lo[d] = 0;
                hi[d] = grid[d];
            }
        int y0 = D == 2 ? lo[1] : 0, y1 = D == 2 ? hi[1] : 1;
        for (int y = y0; y < y1; y++)
            for (int x = lo[0]; x < hi[0]; x++) {
                uint64_t tile;
                if (D == 1)
                    tile = (uint64_t)wrap_tile(x, grid[0]);
                else
                    tile = (uint64_t)(wrap_tile(y, grid[1]) * grid[0] + wrap_tile(x, grid[0]));
                if (tile > (uint64_t)blocks) bad = 1;
                if (R == cap) {
                    cap *= 2;
                    keys = (uint64_t *)realloc(keys, cap * sizeof(uint64_t));
                }
                keys[R++] = (tile << 32) | (uint32_t)g;
            }
    }
    if (bad) {
        free(keys);
        return -1;
    }
    /* radix sort on bits [0, 32+bit_length(blocks)) — keys are unique there */
    (void)bit_length;
    qsort(keys, (size_t)R, sizeof(uint64_t), cmp_u64);
    memset(ranges, 0, (size_t)(blocks + 1) * 2 * sizeof(uint32_t));
    ranges_of(R, keys, ranges);
    for (int64_t i = 0; i < R; i++) point_list[i] = (uint32_t)(keys[i] & 0xffffffffu);
    free(keys);
    return R;
}

/* S6 — sampleWithKeys + SortPairs + identifyTileRanges,
 * cuda_sampler/sampler_impl.cu:155-189,316-327. */
void pigs_oracle_bin_samples(int64_t N, int D, const float *samples, const float *min_b, const int *grid, int blocks,
                             uint32_t *sample_list, uint32_t *sample_tiles, uint32_t *sample_ranges)
{
    uint64_t *keys = (uint64_t *)malloc((size_t)(N > 0 ? N : 1) * sizeof(uint64_t));
    for (int64_t i = 0; i < N; i++) {
        uint32_t t[2] = {0, 0};
        for (int d = 0; d < D; d++)
            t[d] = (uint32_t)imin(grid[d], imax(0, (int)((samples[i * D + d] - min_b[d]) / TILE_W)));
        uint64_t tile = D == 1 ? t[0] : (uint64_t)t[1] * (uint64_t)grid[0] + t[0];
        keys[i] = (tile << 32) | (uint32_t)i;
        if (sample_tiles) sample_tiles[i] = (uint32_t)tile;
    }
    qsort(keys, (size_t)N, sizeof(uint64_t), cmp_u64);
    memset(sample_ranges, 0, (size_t)(blocks + 1) * 2 * sizeof(uint32_t));
    /* a tile id can exceed `blocks` only through the `== grid` alias on the last row; the reference
     * would write past its buffer — we skip such keys when building ranges */
    int64_t L = N;
    while (L > 0 && (keys[L - 1] >> 32) > (uint64_t)blocks) L--;
    ranges_of(L, keys, sample_ranges);
    for (int64_t i = 0; i < N; i++) sample_list[i] = (uint32_t)(keys[i] & 0xffffffffu);
    free(keys);
}

/* ------------------------------------------------------------------------ */
/* Opt-in CULLED binning of the product — NOT reference */
/* semantics; restates pigs_b200/csrc/binning.cu's culled branch so that its   */
/* bins can be checked bit for bit.  Pair set = reference pair set (S8)        */
/* INTERSECTED with { fine tile of the sample within r_c of the mean per       */
/* dimension, period-2 wrap }, fine tiles = every reference tile cut k x k,    */
/* r_c = sigmas * sqrt(lambda_max) (no 1e-6 floor).                            */
/* ------------------------------------------------------------------------ */
static float cull_radius(int D, const float *cov, float sigmas)
{
    if (D == 1) return (float)((double)sigmas * (double)sqrtf(cov[0]));
    float t = cov[1] * cov[1];
    float det = fmaf(cov[0], cov[2], -t);
    float mid = 0.5f * (cov[0] + cov[2]);
    float disc = fmaf(mid, mid, -det);
    double root = sqrt(fmax(0.0, (double)disc));
    float lambda = (float)((double)mid + root);
    return (float)((double)sigmas * (double)sqrtf(lambda));
}

/* distance from m (any of its period-2 images) to the interval [a, b] */
static float torus_interval_dist(float m, float a, float b)
{
    float best = INFINITY;
    for (int s = -1; s <= 1; s++) {
        float ms = m + 2.0f * (float)s;
        float da = a - ms, db = ms - b;
        float dd = fmaxf(0.0f, fmaxf(da, db));
        if (dd < best) best = dd;
    }
    return best;
}

/* fine columns of one dimension kept for one Gaussian; returns the count, writes the fine indices if out != NULL */
static int cull_kept(float mu, float min_b, float r_ref, float r_c, int g, int k, float h, int *out)
{
    int lo = (int)floorf(((mu - min_b) - r_ref) / TILE_W);
    int hi = (int)ceilf(((mu - min_b) + r_ref) / TILE_W);
    if (hi - lo >= g) {
        lo = 0;
        hi = g;
    }
    int n = 0;
    for (int x = lo; x < hi; x++) {
        int tx = wrap_tile(x, g);
        if (tx >= g) continue; /* the reference's alias slot: never evaluated there either */
        for (int j = 0; j < k; j++) {
            int fx = k * tx + j;
            float a = min_b + (float)fx * h;
            float b = min_b + (float)(fx + 1) * h;
            if (torus_interval_dist(mu, a, b) <= r_c) {
                if (out) out[n] = fx;
                n++;
            }
        }
    }
    return n;
}

/* point_list == NULL: returns the number of instances.  Otherwise fills point_list [R] (fine tile asc, id asc) and
 * ranges uint2[fine_blocks+1].  radii = the reference radii (pigs_oracle_radii), ref_grid = the reference grid. */
int64_t pigs_oracle_bin_gaussians_culled(int64_t P, int D, const float *means, const float *cov, const float *radii,
                                         const float *min_b, const int *ref_grid, int k, float sigmas,
                                         uint32_t *point_list, uint32_t *ranges)
{
    const int K = D * (D + 1) / 2;
    const float h = TILE_W / (float)k;
    const int fg0 = k * ref_grid[0];
    const int fblocks = D == 2 ? fg0 * k * ref_grid[1] : fg0;
    int64_t R = 0, cap = 1024;
    uint64_t *keys = point_list ? (uint64_t *)malloc(cap * sizeof(uint64_t)) : NULL;
    int *kx = (int *)malloc(sizeof(int) * (size_t)(k * ref_grid[0] + 1));
    int *ky = (int *)malloc(sizeof(int) * (size_t)(k * (D == 2 ? ref_grid[1] : 1) + 1));
    for (int64_t g = 0; g < P; g++) {
        if (!(radii[g] > 0.0f)) continue;
        float rc = cull_radius(D, cov + g * K, sigmas);
        int nx = cull_kept(means[g * D], min_b[0], radii[g], rc, ref_grid[0], k, h, kx);
        int ny = 1;
        ky[0] = 0;
        if (D == 2) ny = cull_kept(means[g * D + 1], min_b[1], radii[g], rc, ref_grid[1], k, h, ky);
        if (!point_list) {
            R += (int64_t)nx * ny;
            continue;
        }
        for (int y = 0; y < ny; y++)
            for (int x = 0; x < nx; x++) {
                uint64_t tile = (uint64_t)ky[y] * (uint64_t)fg0 + (uint64_t)kx[x];
                if (R == cap) {
                    cap *= 2;
                    keys = (uint64_t *)realloc(keys, cap * sizeof(uint64_t));
                }
                keys[R++] = (tile << 32) | (uint32_t)g;
            }
    }
    free(kx);
    free(ky);
    if (!point_list) return R;
    qsort(keys, (size_t)R, sizeof(uint64_t), cmp_u64);
    memset(ranges, 0, (size_t)(fblocks + 1) * 2 * sizeof(uint32_t));
    ranges_of(R, keys, ranges);
    for (int64_t i = 0; i < R; i++) point_list[i] = (uint32_t)(keys[i] & 0xffffffffu);
    free(keys);
    return R;
}

/* fine tile of every sample: its reference tile (S6, alias quirk included) cut k x k; a sample whose reference tile
 * id is not a real tile goes to the dump slot `fine_blocks` (never evaluated, like the reference's). */
void pigs_oracle_bin_samples_culled(int64_t N, int D, const float *samples, const float *min_b, const int *ref_grid,
                                    int k, uint32_t *sample_list, uint32_t *sample_tiles, uint32_t *sample_ranges)
{
    const float h = TILE_W / (float)k;
    const int rblocks = D == 2 ? ref_grid[0] * ref_grid[1] : ref_grid[0];
    const int fg0 = k * ref_grid[0];
    const int fblocks = rblocks * (D == 2 ? k * k : k);
    uint64_t *keys = (uint64_t *)malloc((size_t)(N > 0 ? N : 1) * sizeof(uint64_t));
    for (int64_t i = 0; i < N; i++) {
        uint32_t t[2] = {0, 0};
        for (int d = 0; d < D; d++)
            t[d] = (uint32_t)imin(ref_grid[d], imax(0, (int)((samples[i * D + d] - min_b[d]) / TILE_W)));
        uint32_t rt = D == 1 ? t[0] : t[1] * (uint32_t)ref_grid[0] + t[0];
        uint64_t tile = (uint64_t)fblocks;
        if ((int)rt < rblocks) {
            int par[2] = {(int)(rt % (uint32_t)ref_grid[0]), (int)(rt / (uint32_t)ref_grid[0])};
            int f[2] = {0, 0};
            for (int d = 0; d < D; d++) {
                int q = (int)((samples[i * D + d] - min_b[d]) / h) - k * par[d];
                f[d] = k * par[d] + imin(k - 1, imax(0, q));
            }
            tile = D == 1 ? (uint64_t)f[0] : (uint64_t)f[1] * (uint64_t)fg0 + (uint64_t)f[0];
        }
        keys[i] = (tile << 32) | (uint32_t)i;
        if (sample_tiles) sample_tiles[i] = (uint32_t)tile;
    }
    qsort(keys, (size_t)N, sizeof(uint64_t), cmp_u64);
    memset(sample_ranges, 0, (size_t)(fblocks + 1) * 2 * sizeof(uint32_t));
    ranges_of(N, keys, sample_ranges);
    for (int64_t i = 0; i < N; i++) sample_list[i] = (uint32_t)(keys[i] & 0xffffffffu);
    free(keys);
}

/* ------------------------------------------------------------------------ */
/* S9 — pair arithmetic.  cuda_sampler/forward.cu:147-158 (offset + torus),   */
/* :168-275 (the four functors).                                              */
/* ------------------------------------------------------------------------ */
static inline void offset_of(int D, const float *mean, const float *x, float *X)
{
    for (int k = 0; k < D; k++) {
        X[k] = mean[k] - x[k];
        if (fabsf(X[k]) > 1.0f) { /* abs(X) > 1.0 */
            if (X[k] >= 0)
                X[k] = (float)(fmod((double)X[k], 2.0) - 2.0);
            else
                X[k] = (float)(fmod((double)X[k], 2.0) + 2.0);
        }
    }
}

enum { FN_VALUE = 0, FN_DERIV = 1, FN_HESS = 2, FN_THIRD = 3 };

static int out_width(int fn, int D)
{
    int w = 1;
    for (int i = 0; i < fn; i++) w *= D;
    return w;
}

/* Evaluates exp(power) and the per-component polynomial factors of one (Gaussian, sample)
 * pair; returns 0 if the reference would `return` early (power > 0). f[] has out_width entries. */
static inline int pair_factors(int fn, int D, const float *X, const float *con, float *alpha, float *f)
{
    if (D == 1) {
        float x1, power;
        if (fn == FN_VALUE) {
            /* -0.5 * con[0] * X[0] * X[0]: all double after the first product */
            power = (float)(-0.5 * (double)con[0] * (double)X[0] * (double)X[0]);
            x1 = 0.0f;
        } else {
            x1 = con[0] * X[0];
            power = (float)(-0.5 * (double)x1 * (double)X[0]);
        }
        if (power > 0.0f) return 0;
        *alpha = expf(power);
        switch (fn) {
        case FN_VALUE: f[0] = 1.0f; break;
        case FN_DERIV: f[0] = x1; break;
        case FN_HESS: f[0] = x1 * x1 - con[0]; break;
        default: /* 2.0*con*x1 - x1^3 + con*x1 */
            f[0] = (float)(2.0 * (double)con[0] * (double)x1 - (double)(x1 * x1 * x1) + (double)(con[0] * x1));
        }
        return 1;
    }
    float power, x1 = 0, x2 = 0;
    if (fn == FN_VALUE) {
        float qd = con[0] * X[0] * X[0] + con[2] * X[1] * X[1];
        power = (float)(-0.5 * (double)qd - (double)(con[1] * X[0] * X[1]));
    } else {
        x1 = con[0] * X[0];
        x2 = con[2] * X[1];
        float qd = x1 * X[0] + x2 * X[1];
        power = (float)(-0.5 * (double)qd - (double)(con[1] * X[0] * X[1]));
    }
    if (power > 0.0f) return 0;
    *alpha = expf(power);
    if (fn == FN_VALUE) {
        f[0] = 1.0f;
        return 1;
    }
    const float a1 = x1 + con[1] * X[1];
    const float a2 = x2 + con[1] * X[0];
    if (fn == FN_DERIV) {
        f[0] = a1;
        f[1] = a2;
    } else if (fn == FN_HESS) {
        f[0] = a1 * a1 - con[0];
        f[1] = a1 * a2 - con[1];
        f[2] = f[1];
        f[3] = a2 * a2 - con[2];
    } else {
        float dxxx = (float)(3.0 * (double)con[0] * (double)a1 - (double)(a1 * a1 * a1));
        float dxxy = (float)(2.0 * (double)con[1] * (double)a1 - (double)(a1 * a1 * a2) + (double)(con[0] * a2));
        float dxyy = (float)(2.0 * (double)con[1] * (double)a2 - (double)(a1 * a2 * a2) + (double)(con[2] * a1));
        float dyyy = (float)(3.0 * (double)con[2] * (double)a2 - (double)(a2 * a2 * a2));
        f[0] = dxxx;
        f[1] = dxxy;
        f[2] = dxxy;
        f[3] = dxyy;
        f[4] = dxxy;
        f[5] = dxyy;
        f[6] = dxyy;
        f[7] = dyyy;
    }
    return 1;
}

/* renderCUDA<F>, cuda_sampler/forward.cu:88-166: out[sample] += values[ch] * alpha * f, Gaussians in
 * list order, sequential fp32 accumulation per sample.  out is [N, width, C], pre-zeroed by the
 * caller (sample_points.cu:215,240,265,290). acc64 != 0 accumulates in double instead (a "truth"
 * mode for tolerance studies; not the reference's arithmetic). */
void pigs_oracle_forward(int fn, int D, int C, int blocks, const uint32_t *ranges, const uint32_t *sample_ranges,
                         const uint32_t *point_list, const uint32_t *sample_list, const float *means,
                         const float *values, const float *conics, const float *samples, float *out, int acc64)
{
    const int W = out_width(fn, D), K = D * (D + 1) / 2;
    for (int b = 0; b < blocks; b++) {
        const uint32_t g0 = ranges[2 * b], g1 = ranges[2 * b + 1];
        const uint32_t s0 = sample_ranges[2 * b], s1 = sample_ranges[2 * b + 1];
        if (g1 <= g0) continue;
#pragma omp parallel for schedule(dynamic, 64)
        for (int64_t si = s0; si < (int64_t)s1; si++) {
            const uint32_t sid = sample_list[si];
            const float *x = samples + (size_t)sid * D;
            float *o = out + (size_t)sid * W * C;
            double acc[16 * 8];
            if (acc64)
                for (int i = 0; i < W * C; i++) acc[i] = 0.0;
            for (uint32_t gi = g0; gi < g1; gi++) {
                const uint32_t id = point_list[gi];
                float X[2], alpha, f[8];
                offset_of(D, means + (size_t)id * D, x, X);
                if (!pair_factors(fn, D, X, conics + (size_t)id * K, &alpha, f)) continue;
                const float *v = values + (size_t)id * C;
                for (int w = 0; w < W; w++)
                    for (int ch = 0; ch < C; ch++) {
                        if (acc64 && W * C <= 128)
                            acc[w * C + ch] += (double)v[ch] * (double)alpha * (double)f[w];
                        else
                            o[w * C + ch] += v[ch] * alpha * f[w];
                    }
            }
            if (acc64 && W * C <= 128)
                for (int i = 0; i < W * C; i++) o[i] = (float)acc[i];
        }
    }
}

/* ------------------------------------------------------------------------ */
/* S10 — backward.  cuda_sampler/backward.cu:27-106 (loop), :108-416 (functors). */
/* Per pair: S = sum_w g_w f_w ; p = dS/da ; E = explicit dS/dc ; then           */
/*   dvalue  += alpha * S_ch                                                    */
/*   dmean   += alpha * (C p - S a)           (with g weighted by values)       */
/*   dconic0 += alpha * (-S X0^2/2 + p1 X0 + E0), dconic1 += alpha*(-S X0 X1 + p1 X1 + p2 X0 + E1), ... */
/* which is the closed form the reference spells out term by term.  The one       */
/* place the reference's formula is NOT the derivative of its forward — the 1-D   */
/* third-derivative conic gradient, backward.cu:326-329 — is reproduced as written.*/
/* Sums are accumulated in double (the reference's atomic order is unspecified).  */
/* ------------------------------------------------------------------------ */
void pigs_oracle_backward(int fn, int P, int D, int C, int blocks, const uint32_t *ranges,
                          const uint32_t *sample_ranges, const uint32_t *point_list, const uint32_t *sample_list,
                          const float *means, const float *values, const float *conics, const float *samples,
                          const float *dL_dout, float *dL_dmeans, float *dL_dvalues, float *dL_dconics)
{
    const int W = out_width(fn, D), K = D * (D + 1) / 2;
    double *am = (double *)calloc((size_t)P * D + 1, sizeof(double));
    double *av = (double *)calloc((size_t)P * C + 1, sizeof(double));
    double *ac = (double *)calloc((size_t)P * K + 1, sizeof(double));
    for (int b = 0; b < blocks; b++) {
        const uint32_t g0 = ranges[2 * b], g1 = ranges[2 * b + 1];
        const uint32_t s0 = sample_ranges[2 * b], s1 = sample_ranges[2 * b + 1];
#pragma omp parallel for schedule(dynamic, 16)
        for (int64_t gi = g0; gi < (int64_t)g1; gi++) {
            const uint32_t id = point_list[gi];
            const float *con = conics + (size_t)id * K;
            const float *v = values + (size_t)id * C;
            double dm[2] = {0, 0}, dc[3] = {0, 0, 0}, dv[16];
            for (int ch = 0; ch < C && ch < 16; ch++) dv[ch] = 0;
            for (uint32_t si = s0; si < s1; si++) {
                const uint32_t sid = sample_list[si];
                float Xf[2] = {0, 0}, alphaf, f[8];
                offset_of(D, means + (size_t)id * D, samples + (size_t)sid * D, Xf);
                if (!pair_factors(fn, D, Xf, con, &alphaf, f)) continue;
                const double alpha = alphaf;
                const float *g = dL_dout + (size_t)sid * W * C;
                /* channel-summed upstream weights gv[w] = sum_ch values[ch] * g[w,ch] */
                double gv[8];
                for (int w = 0; w < W; w++) gv[w] = 0;
                for (int ch = 0; ch < C; ch++) {
                    double Sch = 0;
                    for (int w = 0; w < W; w++) {
                        Sch += (double)f[w] * (double)g[w * C + ch];
                        gv[w] += (double)v[ch] * (double)g[w * C + ch];
                    }
                    if (ch < 16) dv[ch] += alpha * Sch;
                }
                double S = 0;
                for (int w = 0; w < W; w++) S += (double)f[w] * gv[w];
                if (D == 1) {
                    const double X = Xf[0], c = con[0], a = (double)(con[0] * Xf[0]);
                    double p, E;
                    switch (fn) {
                    case FN_VALUE: p = 0; E = 0; break;
                    case FN_DERIV: p = gv[0]; E = 0; break;
                    case FN_HESS: p = 2 * a * gv[0]; E = -gv[0]; break;
                    default: p = (3 * c - 3 * a * a) * gv[0]; E = 3 * a * gv[0];
                    }
                    dm[0] += alpha * (c * p - S * a);
                    if (fn == FN_THIRD) {
                        /* backward.cu:326-329, reproduced verbatim in meaning (not the true derivative) */
                        double q = 2.0 * X * X - 2.0 * a * a * X - 0.5 * (2.0 * X * a - X) * X * X +
                                   0.5 * (a * a - c) * a * X * X;
                        dc[0] += q * gv[0] * alpha;
                    } else
                        dc[0] += alpha * (-0.5 * S * X * X + p * X + E);
                } else {
                    const double X0 = Xf[0], X1 = Xf[1], c0 = con[0], c1 = con[1], c2 = con[2];
                    const double a1 = (double)(con[0] * Xf[0]) + c1 * X1, a2 = (double)(con[2] * Xf[1]) + c1 * X0;
                    double p1 = 0, p2 = 0, E0 = 0, E1 = 0, E2 = 0;
                    if (fn == FN_DERIV) {
                        p1 = gv[0];
                        p2 = gv[1];
                    } else if (fn == FN_HESS) {
                        const double gs = gv[1] + gv[2];
                        p1 = 2 * a1 * gv[0] + a2 * gs;
                        p2 = a1 * gs + 2 * a2 * gv[3];
                        E0 = -gv[0];
                        E1 = -gs;
                        E2 = -gv[3];
                    } else if (fn == FN_THIRD) {
                        const double T0 = gv[0], T1 = gv[1] + gv[2] + gv[4], T2 = gv[3] + gv[5] + gv[6], T3 = gv[7];
                        p1 = T0 * (3 * c0 - 3 * a1 * a1) + T1 * (2 * c1 - 2 * a1 * a2) + T2 * (c2 - a2 * a2);
                        p2 = T1 * (c0 - a1 * a1) + T2 * (2 * c1 - 2 * a1 * a2) + T3 * (3 * c2 - 3 * a2 * a2);
                        E0 = 3 * a1 * T0 + a2 * T1;
                        E1 = 2 * a1 * T1 + 2 * a2 * T2;
                        E2 = a1 * T2 + 3 * a2 * T3;
                    }
                    dm[0] += alpha * (c0 * p1 + c1 * p2 - S * a1);
                    dm[1] += alpha * (c1 * p1 + c2 * p2 - S * a2);
                    dc[0] += alpha * (-0.5 * S * X0 * X0 + p1 * X0 + E0);
                    dc[1] += alpha * (-S * X0 * X1 + p1 * X1 + p2 * X0 + E1);
                    dc[2] += alpha * (-0.5 * S * X1 * X1 + p2 * X1 + E2);
                }
            }
            /* distinct list entries of one tile are distinct Gaussians: no race */
            for (int d = 0; d < D; d++) am[(size_t)id * D + d] += dm[d];
            for (int k = 0; k < K; k++) ac[(size_t)id * K + k] += dc[k];
            for (int ch = 0; ch < C && ch < 16; ch++) av[(size_t)id * C + ch] += dv[ch];
        }
    }
    for (int64_t i = 0; i < (int64_t)P * D; i++) dL_dmeans[i] = (float)am[i];
    for (int64_t i = 0; i < (int64_t)P * C; i++) dL_dvalues[i] = (float)av[i];
    for (int64_t i = 0; i < (int64_t)P * K; i++) dL_dconics[i] = (float)ac[i];
    free(am);
    free(av);
    free(ac);
}

/* ------------------------------------------------------------------------ */
/* Dense pure-PyTorch path, /root/reference/gaussians.py:48-58 (value), :89-101  */
/* (derivative), :103-116 (second derivative): all pairs, no cutoff, no torus,    */
/* FULL [n,d,d] conics, x = sample - mean.  fp32 like torch; per-sample sums in   */
/* float in Gaussian order (torch's sum(1) uses a tree — agreement is to rounding).*/
/* which: bit0 value, bit1 derivative, bit2 second derivative.                     */
/* ------------------------------------------------------------------------ */
void pigs_oracle_dense(int which, int64_t n, int d, int C, int64_t N, const float *means, const float *conics_full,
                       const float *values, const float *samples, float *out_u, float *out_du, float *out_ddu)
{
#pragma omp parallel for schedule(static)
    for (int64_t s = 0; s < N; s++) {
        double u[16] = {0}, du[2 * 16] = {0}, ddu[4 * 16] = {0};
        for (int64_t g = 0; g < n; g++) {
            float x[2], ip[2] = {0, 0};
            const float *cf = conics_full + g * d * d;
            for (int i = 0; i < d; i++) x[i] = samples[s * d + i] - means[g * d + i];
            float q = 0;
            for (int i = 0; i < d; i++) {
                for (int j = 0; j < d; j++) ip[i] += cf[i * d + j] * x[j];
                q += x[i] * ip[i];
            }
            const float dens = expf(-0.5f * q);
            for (int ch = 0; ch < C && ch < 16; ch++) {
                const float v = values[g * C + ch];
                if (which & 1) u[ch] += dens * v;
                if (which & 2)
                    for (int i = 0; i < d; i++) du[i * 16 + ch] += (-ip[i] * dens) * v;
                if (which & 4)
                    for (int i = 0; i < d; i++)
                        for (int j = 0; j < d; j++) ddu[(i * d + j) * 16 + ch] += ((ip[i] * ip[j] - cf[i * d + j]) * dens) * v;
            }
        }
        for (int ch = 0; ch < C && ch < 16; ch++) {
            if (which & 1) out_u[s * C + ch] = (float)u[ch];
            if (which & 2)
                for (int i = 0; i < d; i++) out_du[(s * d + i) * C + ch] = (float)du[i * 16 + ch];
            if (which & 4)
                for (int i = 0; i < d * d; i++) out_ddu[(s * d * d + i) * C + ch] = (float)ddu[i * 16 + ch];
        }
    }
}

/* ------------------------------------------------------------------------ */
/* S11 — neighbour lists.  aggregate_neighbors.cu:18-55 (findCollisions),       */
/* :57-127 (preprocess), host :323-367.                                         */
/* ------------------------------------------------------------------------ */
static inline int collides(int D, const float *mi, float ri, const float *mj, float rj)
{
    float dist = 0.0f;
    for (int d = 0; d < D; d++) {
        float dx = mj[d] - mi[d];
        /* dx = min(dx, abs(2.0 - fmod(abs(dx), 2.0))) in double, back to float */
        double w = fabs(2.0 - fmod(fabs((double)dx), 2.0));
        dx = (float)fmin((double)dx, w);
        dist = fmaf(dx, dx, dist); /* the reference's SASS contracts `dist += dx * dx` into FFMA */
    }
    const float rad = ri + rj;
    return !(dist > rad * rad);
}

/* counts[i] = number of neighbours of i (self included); returns the total. */
int64_t pigs_oracle_neighbor_counts(int64_t P, int D, const float *means, const float *radii, int64_t *counts)
{
    int64_t total = 0;
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : total)
    for (int64_t i = 0; i < P; i++) {
        counts[i] = 0;
        const float ri = (float)((double)radii[i] * 0.2);
        if ((double)ri < 1e-6) continue;
        for (int64_t j = 0; j < P; j++) {
            const float rj = (float)((double)radii[j] * 0.2);
            if ((double)rj < 1e-6) continue;
            if (collides(D, means + i * D, ri, means + j * D, rj)) counts[i]++;
        }
        total += counts[i];
    }
    return total;
}

/* ranges = inclusive cumsum of counts (int64).  indices pre-filled with -1, dists/densities with 0. */
void pigs_oracle_neighbor_fill(int64_t P, int D, const float *means, const float *conics, const float *radii,
                               const int64_t *ranges, int64_t *indices, float *dists, float *densities,
                               float *inv_total)
{
    const int K = D * (D + 1) / 2;
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < P; i++) {
        const int64_t start = i == 0 ? 0 : ranges[i - 1];
        const float ri = (float)((double)radii[i] * 0.2);
        const float my_radius = (float)((double)radii[i] * 0.333);
        const float inv_r = (float)(1.0 / ((double)my_radius + 1e-6));
        float total = 0.0f;
        int64_t cur = -1;
        if (!((double)ri < 1e-6))
            for (int64_t j = 0; j < P; j++) {
                const float rj = (float)((double)radii[j] * 0.2);
                if ((double)rj < 1e-6) continue;
                if (!collides(D, means + i * D, ri, means + j * D, rj)) continue;
                cur++;
                float X[2] = {0, 0};
                offset_of(D, means + j * D, means + i * D, X);
                const float *con = conics + j * K;
                float power;
                if (D == 1)
                    power = (float)(-0.5 * (double)con[0] * (double)X[0] * (double)X[0]);
                else
                    power = (float)(-0.5 * (double)(con[0] * X[0] * X[0] + con[2] * X[1] * X[1]) -
                                    (double)(con[1] * X[0] * X[1]));
                for (int d = 0; d < D; d++) dists[(start + cur) * D + d] = X[d] * inv_r;
                if (power > 0) continue;
                densities[start + cur] = expf(power);
                indices[start + cur] = j;
                total += densities[start + cur];
            }
        inv_total[i] = (float)(1.0 / ((double)total + 1e-6));
    }
}

/* The same lists for a SAMPLE of rows (full-size checks: every row still scans all P candidates, P = 1M in
 * tests/test_gpu_fullsize.py).  Row k of the sample owns slots [k*cap, k*cap + counts[k]) of the outputs; a row
 * with more than cap neighbours keeps its count and the first cap entries. */
void pigs_oracle_neighbor_rows(int64_t P, int D, const float *means, const float *conics, const float *radii,
                               int64_t nrows, const int64_t *rows, int64_t cap, int64_t *counts, int64_t *indices,
                               float *dists, float *densities, float *inv_total)
{
    const int K = D * (D + 1) / 2;
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t k = 0; k < nrows; k++) {
        const int64_t i = rows[k];
        const float ri = (float)((double)radii[i] * 0.2);
        const float my_radius = (float)((double)radii[i] * 0.333);
        const float inv_r = (float)(1.0 / ((double)my_radius + 1e-6));
        float total = 0.0f;
        int64_t cur = -1;
        if (!((double)ri < 1e-6))
            for (int64_t j = 0; j < P; j++) {
                const float rj = (float)((double)radii[j] * 0.2);
                if ((double)rj < 1e-6) continue;
                if (!collides(D, means + i * D, ri, means + j * D, rj)) continue;
                cur++;
                float X[2] = {0, 0};
                offset_of(D, means + j * D, means + i * D, X);
                const float *con = conics + j * K;
                float power;
                if (D == 1)
                    power = (float)(-0.5 * (double)con[0] * (double)X[0] * (double)X[0]);
                else
                    power = (float)(-0.5 * (double)(con[0] * X[0] * X[0] + con[2] * X[1] * X[1]) -
                                    (double)(con[1] * X[0] * X[1]));
                const float dens = power > 0 ? 0.0f : expf(power);
                total += dens;
                if (cur < cap) {
                    for (int d = 0; d < D; d++) dists[(k * cap + cur) * D + d] = X[d] * inv_r;
                    densities[k * cap + cur] = dens;
                    indices[k * cap + cur] = power > 0 ? -1 : j;
                }
            }
        counts[k] = cur + 1;
        inv_total[k] = (float)(1.0 / ((double)total + 1e-6));
    }
}

/* ------------------------------------------------------------------------ */
/* S12 — aggregation forward.  aggregate_neighbors.cu:129-207.                  */
/* ------------------------------------------------------------------------ */
void pigs_oracle_aggregate_forward(int64_t P, int D, int L, int K, int E, const float *features,
                                   const float *transform, const float *queries, const float *keys,
                                   const float *freqs, const float *dt, const int64_t *indices,
                                   const int64_t *ranges, const float *dists, const float *densities,
                                   const float *inv_total, float *weights, float *embeddings, float *factors,
                                   float *out)
{
    const int stride = (E - 1) / D, nF = (E - 1) / D / 2;
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < P; i++) {
        const int64_t start = i == 0 ? 0 : ranges[i - 1], end = ranges[i];
        float *o = out + i * L;
        for (int64_t s = start; s < end; s++) {
            const int64_t j = indices[s];
            if (j == -1) continue;
            float w = 0.0f;
            for (int k = 0; k < K; k++) w += queries[i * K + k] * keys[j * K + k];
            weights[s] = w;
            float emb = 0.0f, fac = 0.0f;
            for (int d = 0; d < D; d++)
                for (int e = 0; e < nF; e++) {
                    /* frequencies[e] * M_PI * X[d]: float*double*float -> double argument */
                    const double arg = (double)freqs[e] * M_PI * (double)dists[s * D + d];
                    const float sn = (float)sin(arg), cs = (float)cos(arg);
                    emb += dt[d * stride + e * 2 + 0] * sn;
                    emb += dt[d * stride + e * 2 + 1] * cs;
                    fac += dt[E + d * stride + e * 2 + 0] * sn;
                    fac += dt[E + d * stride + e * 2 + 1] * cs;
                }
            emb += dt[E - 1];
            fac += dt[2 * E - 1];
            embeddings[s] = emb;
            factors[s] = fac;
            const float dw = inv_total[i] * densities[s] * w;
            const float dwf = dw * fac, dwe = dw * emb;
            for (int a = 0; a < L; a++) {
                const float embedded = dwe + dwf * features[j * L + a];
                for (int b = 0; b < L; b++) o[b] += transform[a * L + b] * embedded;
            }
        }
    }
}

/* aggregation backward.  aggregate_neighbors.cu:210-321.  Double accumulators
 * (the reference's atomic order is unspecified). */
void pigs_oracle_aggregate_backward(int64_t P, int D, int L, int K, int E, const float *features,
                                    const float *transform, const float *queries, const float *keys,
                                    const float *freqs, const float *dt, const int64_t *indices,
                                    const int64_t *ranges, const float *dists, const float *densities,
                                    const float *weights, const float *embeddings, const float *factors,
                                    const float *inv_total, const float *dL_dout, float *dL_dfeatures,
                                    float *dL_dtransform, float *dL_dqueries, float *dL_dkeys, float *dL_dfreqs,
                                    float *dL_ddt)
{
    const int stride = (E - 1) / D, nF = (E - 1) / D / 2;
    double *gf = (double *)calloc((size_t)P * L + 1, sizeof(double));
    double *gk = (double *)calloc((size_t)P * K + 1, sizeof(double));
    double *gq = (double *)calloc((size_t)P * K + 1, sizeof(double));
    double *gt = (double *)calloc((size_t)L * L + 1, sizeof(double));
    double *gfr = (double *)calloc((size_t)nF + 1, sizeof(double));
    double *gdt = (double *)calloc((size_t)2 * E + 1, sizeof(double));
    double *st = (double *)malloc((size_t)L * sizeof(double));
    for (int64_t i = 0; i < P; i++) {
        const int64_t start = i == 0 ? 0 : ranges[i - 1], end = ranges[i];
        const float *g = dL_dout + i * L;
        for (int a = 0; a < L; a++) {
            float acc = 0.0f;
            for (int b = 0; b < L; b++) acc += transform[a * L + b] * g[b];
            st[a] = acc;
        }
        for (int64_t s = start; s < end; s++) {
            const int64_t j = indices[s];
            if (j == -1) continue;
            const double dc = (double)densities[s] * (double)inv_total[i];
            const double w = weights[s], emb = embeddings[s], fac = factors[s];
            const double dcw = dc * w;
            for (int d = 0; d < D; d++)
                for (int e = 0; e < nF; e++) {
                    const double X = dists[s * D + d];
                    const double arg = (double)freqs[e] * M_PI * X;
                    const double sn = (float)sin(arg), cs = (float)cos(arg);
                    const int i0 = d * stride + e * 2;
                    for (int a = 0; a < L; a++) {
                        const double dct = dcw * st[a], fa = features[j * L + a];
                        gdt[i0 + 0] += dct * sn;
                        gdt[E + i0 + 0] += dct * sn * fa;
                        gfr[e] += cs * M_PI * X * dct * ((double)dt[i0] + (double)dt[E + i0] * fa);
                        gdt[i0 + 1] += dct * cs;
                        gdt[E + i0 + 1] += dct * cs * fa;
                        gfr[e] += -sn * M_PI * X * dct * ((double)dt[i0 + 1] + (double)dt[E + i0 + 1] * fa);
                    }
                }
            const double dce = dc * emb, dcf = dc * fac;
            for (int a = 0; a < L; a++) {
                const double dct = dcw * st[a], fa = features[j * L + a];
                gdt[E - 1] += dct;
                gdt[2 * E - 1] += dct * fa;
                gf[j * L + a] += dct * fac;
                const double embedded = dce + dcf * fa;
                const double we = w * embedded;
                for (int b = 0; b < L; b++) gt[a * L + b] += we * (double)g[b];
                const double te = st[a] * embedded;
                for (int k = 0; k < K; k++) {
                    gq[i * K + k] += (double)keys[j * K + k] * te;
                    gk[j * K + k] += (double)queries[i * K + k] * te;
                }
            }
        }
    }
    for (int64_t i = 0; i < P * L; i++) dL_dfeatures[i] = (float)gf[i];
    for (int64_t i = 0; i < P * K; i++) dL_dkeys[i] = (float)gk[i];
    for (int64_t i = 0; i < P * K; i++) dL_dqueries[i] = (float)gq[i];
    for (int i = 0; i < L * L; i++) dL_dtransform[i] = (float)gt[i];
    for (int i = 0; i < nF; i++) dL_dfreqs[i] = (float)gfr[i];
    for (int i = 0; i < 2 * E; i++) dL_ddt[i] = (float)gdt[i];
    free(gf);
    free(gk);
    free(gq);
    free(gt);
    free(gfr);
    free(gdt);
    free(st);
}
