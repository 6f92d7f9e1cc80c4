// aggregate.cu — neighbour lists and neighbour feature aggregation for model_pn.py
// (replaces aggregate_neighbors.cu:18-475 of the reference: findCollisions, preprocess,
//  aggregateNeighbors, aggregateNeighborsBackward and their torch host wrappers).
//
// Neighbour search.  The reference tests all P^2 pairs into a dense P x P bool matrix (1 TB at
// P = 1M).  Here the Gaussians are bucketed into a uniform grid (stable sort by cell, so every cell
// list is ascending in id); a warp owns one Gaussian, enumerates only the cells that can hold a
// neighbour under the reference's (asymmetric, period-2) torus distance and applies the reference's
// exact fp32/fp64 test to every candidate, so the accepted set is identical.  The row is then put in
// ascending id order (shared-memory bitonic sort; rows beyond its capacity go through a per-CTA bitmap
// over ids) which reproduces the reference's CSR bit for bit.
// Fast path (nb_cell_*): all Gaussians of one grid cell share one candidate superset (the cells reachable
// from the cell's bounding box at the cell's largest radius).  A CTA owns the cell, sorts that superset by
// id ONCE in shared memory, and every warp then walks it in id order for one row at a time, so accepted
// candidates leave in ascending id order with a ballot prefix and no per-row sort.  Cells whose superset
// does not fit in shared memory fall back to the per-row kernels.
//
// Aggregation.  One warp per Gaussian; lanes over neighbour slots for the per-slot scalars (query.key,
// Fourier embedding), lanes over feature channels for the row accumulation.  The transform is applied
// ONCE per Gaussian (z . T) instead of once per neighbour (L^2 FMAs each in the reference); the backward
// reduces the shared-parameter gradients (transform, frequencies, distance_transform) in shared memory
// and issues one atomic per address per CTA (the reference: one per neighbour per channel).
#include "common.cuh"
#include "eval_common.cuh"

#include <stdlib.h>

#include <mutex>

namespace pigs {

constexpr int NB_THREADS = 128;            // 4 warps, one row each
constexpr int NB_WARPS = NB_THREADS / 32;
constexpr int NB_WARP_CAP = 2048;          // neighbours per row sorted in shared memory by one warp
constexpr int NB_MAX_SPANS = 6;            // merged cell intervals per dimension (direct + torus images)
constexpr int NB_BIG_CTAS = 64;            // persistent CTAs (one id-bitmap each) for rows beyond NB_WARP_CAP
constexpr int NB_BIG_THREADS = 256;
#ifndef PIGS_NBC_THREADS
#define PIGS_NBC_THREADS 384
#endif
constexpr int NBC_THREADS = PIGS_NBC_THREADS;           // cell-cooperative path: 12 warps share one id-sorted candidate list
constexpr int NBC_WARPS = NBC_THREADS / 32;
constexpr int NBC_CAP = 4096;              // candidates per cell in shared memory (id, x, y, r: 16 B each = 64 KB)

struct NbGrid {
    float lo[2], hi[2];    // bounding box of the valid means
    float inv_h[2];
    int n[2];
    int ncells;
    float rmax;            // max effective radius (radii * 0.2)
    uint32_t big_count;    // rows routed to the bitmap path
    uint32_t big_next;
    int pad[4];
};

// rows (Gaussian ids) whose lists this call builds: [lo, hi).  Multi-GPU aggregation shards the CSR by query
// Gaussian; every other row gets count 0 and the no-neighbour total.
struct NbWindow {
    uint32_t lo, hi;
    __host__ __device__ bool has(uint32_t id) const { return id >= lo && id < hi; }
};

struct NbLayout {
    NbGrid* grid;
    float* reff;          // [P] float(radii * 0.2)
    uint32_t *keys_a, *vals_a, *keys_b, *vals_b;  // [P]
    char* sort_ws;
    uint32_t* cell_start;  // [ncells_max + 2]
    float4* srec;          // [P] sorted by (cell, id): {mx, my, reff, id bits}
    uint32_t* big_list;    // [P]
    uint32_t* bitmaps;     // [NB_BIG_CTAS][words]
    uint8_t* cell_over;    // [ncells_max + 2] 1 = candidate superset of the cell exceeds NBC_CAP: per-row kernels
    float4* cpack;         // [P] conics padded to 16 B (one aligned load per emitted neighbour), built by the fill pass
    int64_t ncells_max, words;
    size_t bytes;
    NbLayout(void* base, int64_t P, int D)
    {
        const int64_t Pn = P > 0 ? P : 1;
        ncells_max = cells_cap(P, D);
        words = (Pn + 31) / 32;
        Carver c(base);
        grid = c.take<NbGrid>(1);
        reff = c.take<float>((size_t)Pn);
        keys_a = c.take<uint32_t>((size_t)Pn);
        vals_a = c.take<uint32_t>((size_t)Pn);
        keys_b = c.take<uint32_t>((size_t)Pn);
        vals_b = c.take<uint32_t>((size_t)Pn);
        sort_ws = c.take<char>(sort_ws_bytes(Pn));
        cell_start = c.take<uint32_t>((size_t)ncells_max + 2);
        srec = c.take<float4>((size_t)Pn);
        big_list = c.take<uint32_t>((size_t)Pn);
        bitmaps = c.take<uint32_t>((size_t)NB_BIG_CTAS * words);
        cell_over = c.take<uint8_t>((size_t)ncells_max + 2);
        cpack = c.take<float4>((size_t)Pn);
        bytes = c.total();
    }
    static int per_dim_cap(int64_t P, int D)
    {
        const int64_t budget = 2 * (P > 0 ? P : 1) + 1024;
        if (D == 1) return (int)(budget < (1 << 22) ? budget : (1 << 22));
        int n = 1;
        while ((int64_t)(n + 1) * (n + 1) <= budget && n < 2048) n++;
        return n;
    }
    static int64_t cells_cap(int64_t P, int D)
    {
        const int64_t n = per_dim_cap(P, D);
        return D == 1 ? n : n * n;
    }
};

// ---- stage 0: effective radii, bounding box of the valid means, max radius -----------------------
__global__ void nb_init_kernel(NbGrid* g)
{
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        for (int d = 0; d < 2; d++) {
            g->lo[d] = __int_as_float(0x7f800000);
            g->hi[d] = __int_as_float(0xff800000);
            g->inv_h[d] = 0.f;
            g->n[d] = 1;
        }
        g->ncells = 1;
        g->rmax = 0.f;
        g->big_count = 0;
        g->big_next = 0;
    }
}

template <int D>
__global__ void __launch_bounds__(256) nb_prepare_kernel(int64_t P, const float* __restrict__ means,
                                                         const float* __restrict__ radii, float* __restrict__ reff,
                                                         NbGrid* __restrict__ g)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    float lo[D], hi[D], rm = 0.f;
#pragma unroll
    for (int d = 0; d < D; d++) {
        lo[d] = __int_as_float(0x7f800000);
        hi[d] = __int_as_float(0xff800000);
    }
    if (i < P) {
        // aggregate_neighbors.cu:28: `radii[idx] * 0.2` is a double product rounded to FLOAT
        const float r = __double2float_rn((double)radii[i] * 0.2);
        reff[i] = r;
        if (!((double)r < 1e-6)) {
            rm = r;
#pragma unroll
            for (int d = 0; d < D; d++) lo[d] = hi[d] = means[i * D + d] + 0.0f;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        rm = fmaxf(rm, __shfl_xor_sync(0xffffffffu, rm, o));
#pragma unroll
        for (int d = 0; d < D; d++) {
            lo[d] = fminf(lo[d], __shfl_xor_sync(0xffffffffu, lo[d], o));
            hi[d] = fmaxf(hi[d], __shfl_xor_sync(0xffffffffu, hi[d], o));
        }
    }
    if ((threadIdx.x & 31) == 0 && rm > 0.f) {
        atomic_max_float(&g->rmax, rm);
#pragma unroll
        for (int d = 0; d < D; d++) {
            atomic_min_float(&g->lo[d], lo[d]);
            atomic_max_float(&g->hi[d], hi[d]);
        }
    }
}

// cell edge = rmax / div (div = 2: a row's reach T <= 2 rmax spans 9-10 cells per dimension, and the
// superset shared by a cell is only ~1.4x the accepted set)
__global__ void nb_grid_kernel(NbGrid* g, int D, int per_dim_cap, float div)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    int ncells = 1;
    for (int d = 0; d < D; d++) {
        const float ext = g->hi[d] - g->lo[d];
        int n = 1;
        float inv_h = 0.f;
        if (g->rmax > 0.f && ext > 0.f && ext < 3.0e38f) {
            const float h = fmaxf(g->rmax / div, ext / (float)per_dim_cap);
            const float q = ext / h;
            n = q >= (float)per_dim_cap ? per_dim_cap : max(1, (int)q);
            inv_h = (float)n / ext;
        }
        g->n[d] = n;
        g->inv_h[d] = inv_h;
        ncells *= n;
    }
    g->ncells = ncells;
}

__device__ __forceinline__ int nb_cell(float x, float lo, float inv_h, int n)
{
    const float c = floorf((x - lo) * inv_h);
    return c >= (float)(n - 1) ? n - 1 : (c > 0.f ? (int)c : 0);  // NaN -> 0
}

template <int D>
__global__ void __launch_bounds__(256) nb_key_kernel(int64_t P, const float* __restrict__ means,
                                                     const float* __restrict__ reff, const NbGrid* __restrict__ g,
                                                     uint32_t* __restrict__ keys, uint32_t* __restrict__ vals)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    uint32_t key = (uint32_t)g->ncells;  // invalid Gaussians sort behind every cell
    if (!((double)reff[i] < 1e-6)) {
        int c = nb_cell(means[i * D], g->lo[0], g->inv_h[0], g->n[0]);
        if (D == 2) c += g->n[0] * nb_cell(means[i * D + 1], g->lo[1], g->inv_h[1], g->n[1]);
        key = (uint32_t)c;
    }
    keys[i] = key;
    vals[i] = (uint32_t)i;
}

// cell_start[c] = first sorted position with key >= c, for c in [0, ncells + 1]
__global__ void __launch_bounds__(256) nb_cell_start_kernel(int64_t P, const uint32_t* __restrict__ keys,
                                                            const NbGrid* __restrict__ g, uint32_t* __restrict__ cell_start)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const uint32_t cur = keys[i];
    const uint32_t first = i == 0 ? 0u : keys[i - 1] + 1u;
    for (uint32_t c = first; c <= cur; c++) cell_start[c] = (uint32_t)i;
    if (i == P - 1)
        for (uint32_t c = cur + 1u; c <= (uint32_t)g->ncells + 1u; c++) cell_start[c] = (uint32_t)P;
}

template <int D>
__global__ void __launch_bounds__(256) nb_gather_kernel(int64_t P, const uint32_t* __restrict__ ids,
                                                        const float* __restrict__ means, const float* __restrict__ reff,
                                                        float4* __restrict__ srec)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const uint32_t id = ids[i];
    srec[i] = make_float4(means[(size_t)id * D], D == 2 ? means[(size_t)id * D + 1] : 0.f, reff[id], __uint_as_float(id));
}

// ---- candidate enumeration -------------------------------------------------------------------------
// Reference distance per dimension (aggregate_neighbors.cu:43-49), dx = other - mine:
//   dx_eff = (FLOAT) min((double)dx, fabs(2.0 - fmod(fabs((double)dx), 2.0)))
// For dx <= 1 this is dx itself (negative offsets never wrap); for dx > 1 it is the distance up to the
// next multiple of 2.  A neighbour needs |dx_eff| <= r_i + r_j <= T in every dimension, i.e. per
// dimension other in [mine - T, mine + T]  or  other in [mine + 2k - T, mine + 2k), k = 1, 2, ...
struct Spans {
    int n;
    int lo[NB_MAX_SPANS], hi[NB_MAX_SPANS];
};

__device__ __forceinline__ void span_add(Spans& s, int lo, int hi, int ncell)
{
    if (s.n > 0 && lo <= s.hi[s.n - 1] + 1) {  // starts are generated in ascending order: merge with the last
        if (hi > s.hi[s.n - 1]) s.hi[s.n - 1] = hi;
        return;
    }
    if (s.n == NB_MAX_SPANS) {  // too many torus images (domain much wider than the period): scan everything
        s.n = 1;
        s.lo[0] = 0;
        s.hi[0] = ncell - 1;
        return;
    }
    s.lo[s.n] = lo;
    s.hi[s.n] = hi;
    s.n++;
}

// spans for every `mine` in [mlo, mhi] (a point, or the bounding box of one cell's Gaussians)
__device__ __forceinline__ void dim_spans(float mlo, float mhi, float T, const NbGrid& g, int d, Spans& s)
{
    const float pad = 1e-4f * T + 1e-6f * (1.0f + fmaxf(fabsf(mlo), fabsf(mhi)) + fabsf(g.lo[d]) + fabsf(g.hi[d]));
    const float Tp = T + pad;
    const int n = g.n[d];
    s.n = 0;
    span_add(s, nb_cell(mlo - Tp, g.lo[d], g.inv_h[d], n), nb_cell(mhi + Tp, g.lo[d], g.inv_h[d], n), n);
    for (int k = 1; k < 64; k++) {
        const float a = mlo + 2.0f * (float)k - Tp;
        if (a > g.hi[d]) break;
        span_add(s, nb_cell(a, g.lo[d], g.inv_h[d], n), nb_cell(mhi + 2.0f * (float)k + pad, g.lo[d], g.inv_h[d], n), n);
        if (s.n == 1 && s.lo[0] == 0 && s.hi[0] == n - 1) break;
    }
}

// the reference's collision test, operation for operation (findCollisions, aggregate_neighbors.cu:40-53;
// `dist += dx * dx` is an FFMA in its SASS)
template <int D>
__device__ __forceinline__ bool nb_collides(const float* mi, float ri, const float* mj, float rj)
{
    float dist = 0.0f;
#pragma unroll
    for (int d = 0; d < D; d++) {
        float dx = __fsub_rn(mj[d], mi[d]);
        if (dx > 1.0f) dx = __double2float_rn(fmin((double)dx, fabs(2.0 - fmod(fabs((double)dx), 2.0))));
        dist = __fmaf_rn(dx, dx, dist);
    }
    const float rad = __fadd_rn(ri, rj);
    return !(dist > __fmul_rn(rad, rad));
}

// Calls f(sorted position p, record) for every candidate of row record `me`; the calling group has
// `stride` threads with index `t`.  f is invoked group-uniformly (p may be >= end: valid = false).
template <int D, typename F>
__device__ __forceinline__ void nb_for_candidates(const float4 me, const NbGrid& g, const uint32_t* __restrict__ cell_start,
                                                  const float4* __restrict__ srec, int t, int stride, F f)
{
    const float T = me.z + g.rmax;
    Spans sx, sy;
    dim_spans(me.x, me.x, T, g, 0, sx);
    if (D == 2)
        dim_spans(me.y, me.y, T, g, 1, sy);
    else {
        sy.n = 1;
        sy.lo[0] = sy.hi[0] = 0;
    }
    for (int iy = 0; iy < sy.n; iy++)
        for (int cy = sy.lo[iy]; cy <= sy.hi[iy]; cy++)
            for (int ix = 0; ix < sx.n; ix++) {
                const uint32_t b = cell_start[cy * g.n[0] + sx.lo[ix]];
                const uint32_t e = cell_start[cy * g.n[0] + sx.hi[ix] + 1];
                for (uint32_t base = b; base < e; base += stride) {
                    const uint32_t p = base + t;
                    const bool valid = p < e;
                    float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (valid) o = srec[p];
                    f(valid, o);
                }
            }
}

// ---- stage A: counts -------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(NB_THREADS) nb_count_kernel(int64_t P, const NbGrid* __restrict__ gp,
                                                              const uint32_t* __restrict__ cell_start,
                                                              const float4* __restrict__ srec, int64_t* __restrict__ counts,
                                                              const uint32_t* __restrict__ row_cell,
                                                              const uint8_t* __restrict__ cell_over, NbWindow win)
{
    const int lane = threadIdx.x & 31;
    const int64_t r = (int64_t)blockIdx.x * NB_WARPS + (threadIdx.x >> 5);
    if (r >= (int64_t)cell_start[gp->ncells]) return;  // invalid rows keep count 0
    if (!cell_over[row_cell[r]]) return;                // counted by nb_cell_count_kernel
    const NbGrid g = *gp;
    const float4 me = srec[r];
    if (!win.has(__float_as_uint(me.w))) return;        // row of another shard: count 0
    const float mi[2] = {me.x, me.y};
    uint32_t cnt = 0;
    nb_for_candidates<D>(me, g, cell_start, srec, lane, 32, [&](bool valid, const float4& o) {
        const float mj[2] = {o.x, o.y};
        if (valid && nb_collides<D>(mi, me.z, mj, o.z)) cnt++;
    });
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if (lane == 0) counts[__float_as_uint(me.w)] = (int64_t)cnt;
}

// ---- stage B: fill ---------------------------------------------------------------------------------
// outputs of one CSR slot (preprocess, aggregate_neighbors.cu:82-126), returns the density (0 if skipped)
// NOWRAP: the caller knows |mj - mi| <= 1 in every dimension, where torus_wrap is the identity
// PACKED: `conics` is the float4-padded copy (NbLayout::cpack)
template <int D, bool NOWRAP = false, bool PACKED = false>
__device__ __forceinline__ float nb_emit_core(int64_t slot, uint32_t j, const float* mj, const float* mi, float inv_r,
                                              const float* __restrict__ conics, int64_t* __restrict__ indices,
                                              float* __restrict__ dists, float* __restrict__ densities)
{
    float X[D];
#pragma unroll
    for (int d = 0; d < D; d++) {
        X[d] = __fsub_rn(mj[d], mi[d]);
        if (!NOWRAP) X[d] = torus_wrap(X[d]);
    }
    float power;
    if (D == 1) {
        const float c0 = PACKED ? conics[(size_t)j * 4] : conics[j];
        power = __double2float_rn(-0.5 * (double)c0 * (double)X[0] * (double)X[0]);
    } else {
        float c0, c1, c2;
        if (PACKED) {
            const float4 c = reinterpret_cast<const float4*>(conics)[j];
            c0 = c.x, c1 = c.y, c2 = c.z;
        } else {
            c0 = conics[(size_t)j * 3], c1 = conics[(size_t)j * 3 + 1], c2 = conics[(size_t)j * 3 + 2];
        }
        const float a = __fmul_rn(__fmul_rn(c0, X[0]), X[0]);
        const float q = __fmaf_rn(__fmul_rn(c2, X[D - 1]), X[D - 1], a);
        const float m = __fmul_rn(__fmul_rn(c1, X[0]), X[D - 1]);
        // The reference forms -0.5 * (double)q - (double)m and rounds it to float.  One float FMA forms the same exact
        // value and rounds it once: the two can differ in the last bit of a tie, never in sign, so the `power > 0` skip
        // (which decides an INDEX, bit-exact) is the same; only a result that underflows takes the double detour.
        // (ncu, profiles/r03h_ncu_nb.json: the fill pass is issue bound; the F2F / DFMA / F2F chain and the full-range
        // expf below were a quarter of its instructions.)
        power = __fmaf_rn(-0.5f, q, -m);
        if (fabsf(power) < 1e-30f) power = __double2float_rn(-0.5 * (double)q - (double)m);
    }
    if (D == 2) {
        reinterpret_cast<float2*>(dists)[slot] = make_float2(__fmul_rn(X[0], inv_r), __fmul_rn(X[D - 1], inv_r));
    } else {
#pragma unroll
        for (int d = 0; d < D; d++) dists[slot * D + d] = __fmul_rn(X[d], inv_r);
    }
    if (power > 0.0f) {  // skipped by the reference: index stays -1, density 0
        indices[slot] = -1;
        densities[slot] = 0.0f;
        return 0.0f;
    }
    // exp through MUFU.EX2 (2 ulp; the evaluation kernels' own exp): densities are compared to 1e-5, not bit for bit
    const float dens = D == 2 ? ex2_approx(__fmul_rn(power, 1.4426950408889634f)) : expf(power);
    densities[slot] = dens;
    indices[slot] = (int64_t)j;
    return dens;
}

template <int D>
__device__ __forceinline__ float nb_emit_slot(int64_t slot, uint32_t j, const float* mi, float inv_r,
                                              const float* __restrict__ means, const float* __restrict__ conics,
                                              int64_t* __restrict__ indices, float* __restrict__ dists,
                                              float* __restrict__ densities)
{
    float mj[D];
#pragma unroll
    for (int d = 0; d < D; d++) mj[d] = means[(size_t)j * D + d];
    return nb_emit_core<D>(slot, j, mj, mi, inv_r, conics, indices, dists, densities);
}

__device__ __forceinline__ float nb_inv_radius(float radius)
{
    // aggregate_neighbors.cu:74-75: my_radius = radii * 0.333 (double -> FLOAT); 1.0 / (my_radius + 1e-6) -> FLOAT
    const float my_radius = __double2float_rn((double)radius * 0.333);
    return __double2float_rn(1.0 / ((double)my_radius + 1e-6));
}

template <int D>
__global__ void __launch_bounds__(NB_THREADS) nb_fill_kernel(int64_t P, NbGrid* __restrict__ gp,
                                                             const uint32_t* __restrict__ cell_start,
                                                             const float4* __restrict__ srec, const float* __restrict__ means,
                                                             const float* __restrict__ conics, const float* __restrict__ radii,
                                                             const int64_t* __restrict__ ranges, int64_t* __restrict__ indices,
                                                             float* __restrict__ dists, float* __restrict__ densities,
                                                             float* __restrict__ inv_total, uint32_t* __restrict__ big_list,
                                                             const uint32_t* __restrict__ row_cell,
                                                             const uint8_t* __restrict__ cell_over, NbWindow win)
{
    __shared__ uint32_t s_buf[NB_WARPS][NB_WARP_CAP];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t r = (int64_t)blockIdx.x * NB_WARPS + w;
    if (r >= P) return;
    const bool valid_row = r < (int64_t)cell_start[gp->ncells];
    if (valid_row && !cell_over[row_cell[r]]) return;  // handled by nb_cell_fill_kernel
    const NbGrid g = *gp;
    const float4 me = srec[r];
    const uint32_t id = __float_as_uint(me.w);
    if (!valid_row || !win.has(id)) {  // no neighbours (or another shard's row): total density 0 (aggregate_neighbors.cu:124-125)
        if (lane == 0) inv_total[id] = __double2float_rn(1.0 / (0.0 + 1e-6));
        return;
    }
    const int64_t start = id == 0 ? 0 : ranges[id - 1];
    const int64_t n = ranges[id] - start;
    if (n > NB_WARP_CAP) {
        if (lane == 0) big_list[atomicAdd(&gp->big_count, 1u)] = (uint32_t)r;
        return;
    }
    const float mi[2] = {me.x, me.y};
    uint32_t* buf = s_buf[w];
    uint32_t cnt = 0;
    nb_for_candidates<D>(me, g, cell_start, srec, lane, 32, [&](bool valid, const float4& o) {
        const float mj[2] = {o.x, o.y};
        const bool ok = valid && nb_collides<D>(mi, me.z, mj, o.z);
        const uint32_t m = __ballot_sync(0xffffffffu, ok);
        if (ok) {
            const uint32_t pos = cnt + __popc(m & ((1u << lane) - 1u));
            if (pos < NB_WARP_CAP) buf[pos] = __float_as_uint(o.w);
        }
        cnt += __popc(m);
    });
    // ascending neighbour id (the reference scans i = 0..P-1): bitonic sort of the padded row
    uint32_t n2 = 1;
    while (n2 < cnt) n2 <<= 1;
    for (uint32_t i = cnt + lane; i < n2; i += 32) buf[i] = 0xffffffffu;
    __syncwarp();
    for (uint32_t k = 2; k <= n2; k <<= 1)
        for (uint32_t j = k >> 1; j > 0; j >>= 1) {
            for (uint32_t t = lane; t < (n2 >> 1); t += 32) {
                const uint32_t i0 = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                const uint32_t i1 = i0 | j;
                const uint32_t a = buf[i0], b = buf[i1];
                const bool up = (i0 & k) == 0;
                if ((a > b) == up) {
                    buf[i0] = b;
                    buf[i1] = a;
                }
            }
            __syncwarp();
        }
    const float inv_r = nb_inv_radius(radii[id]);
    float total = 0.0f;
    for (uint32_t s = lane; s < cnt; s += 32)
        total += nb_emit_slot<D>(start + s, buf[s], mi, inv_r, means, conics, indices, dists, densities);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
    if (lane == 0) inv_total[id] = __double2float_rn(1.0 / ((double)total + 1e-6));
}

// rows with more than NB_WARP_CAP neighbours: a persistent CTA marks the accepted ids in a bitmap over
// [0, P) and emits them in ascending order with a block-wide prefix sum of the word popcounts
template <int D>
__global__ void __launch_bounds__(NB_BIG_THREADS) nb_fill_big_kernel(int64_t P, NbGrid* __restrict__ gp,
                                                                     const uint32_t* __restrict__ cell_start,
                                                                     const float4* __restrict__ srec,
                                                                     const float* __restrict__ means,
                                                                     const float* __restrict__ conics,
                                                                     const float* __restrict__ radii,
                                                                     const int64_t* __restrict__ ranges,
                                                                     int64_t* __restrict__ indices, float* __restrict__ dists,
                                                                     float* __restrict__ densities, float* __restrict__ inv_total,
                                                                     const uint32_t* __restrict__ big_list,
                                                                     uint32_t* __restrict__ bitmaps, int64_t words)
{
    __shared__ uint32_t s_scan[NB_BIG_THREADS];
    __shared__ uint32_t s_row, s_base;
    __shared__ float s_total;
    const NbGrid g = *gp;
    uint32_t* bm = bitmaps + (size_t)blockIdx.x * words;
    const int tid = threadIdx.x;
    for (;;) {
        __syncthreads();
        if (tid == 0) {
            s_row = atomicAdd(&gp->big_next, 1u);
            s_base = 0;
            s_total = 0.0f;
        }
        __syncthreads();
        if (s_row >= g.big_count) break;
        const float4 me = srec[big_list[s_row]];
        const uint32_t id = __float_as_uint(me.w);
        const float mi[2] = {me.x, me.y};
        for (int64_t wd = tid; wd < words; wd += NB_BIG_THREADS) bm[wd] = 0u;
        __syncthreads();
        nb_for_candidates<D>(me, g, cell_start, srec, tid, NB_BIG_THREADS, [&](bool valid, const float4& o) {
            const float mj[2] = {o.x, o.y};
            if (valid && nb_collides<D>(mi, me.z, mj, o.z)) {
                const uint32_t j = __float_as_uint(o.w);
                atomicOr(&bm[j >> 5], 1u << (j & 31));
            }
        });
        __syncthreads();
        const int64_t start = id == 0 ? 0 : ranges[id - 1];
        const float inv_r = nb_inv_radius(radii[id]);
        float total = 0.0f;
        for (int64_t w0 = 0; w0 < words; w0 += NB_BIG_THREADS) {
            const int64_t wd = w0 + tid;
            uint32_t bits = wd < words ? bm[wd] : 0u;
            const uint32_t c = __popc(bits);
            s_scan[tid] = c;
            __syncthreads();
            for (int o = 1; o < NB_BIG_THREADS; o <<= 1) {
                const uint32_t v = tid >= o ? s_scan[tid - o] : 0u;
                __syncthreads();
                s_scan[tid] += v;
                __syncthreads();
            }
            int64_t slot = start + s_base + s_scan[tid] - c;
            while (bits) {
                const uint32_t b = __ffs(bits) - 1;
                bits &= bits - 1;
                total += nb_emit_slot<D>(slot++, (uint32_t)(wd * 32 + b), mi, inv_r, means, conics, indices, dists, densities);
            }
            __syncthreads();
            if (tid == NB_BIG_THREADS - 1) s_base += s_scan[tid];
            __syncthreads();
        }
        atomicAdd(&s_total, total);
        __syncthreads();
        if (tid == 0) inv_total[id] = __double2float_rn(1.0 / ((double)s_total + 1e-6));
    }
}

// Conservative version of nb_collides for a whole cell: false only if NO row with mean inside the box [blo, bhi] and
// effective radius <= rc can collide with Gaussian (mj, rj).  Per dimension dx = mj - mi ranges over [a, b]; the
// reference distance is |dx| for dx <= 1 and at most the gap up to the next even number for dx > 1
// (aggregate_neighbors.cu:43-49), so a lower bound is min(distance of [a, b] from 0, gap of b to the next even
// number [0 if an even number >= 2 lies inside the part of the interval beyond 1]).
template <int D>
__device__ __forceinline__ bool nb_box_reach(const float* blo, const float* bhi, float rc, const float* mj, float rj)
{
    const float rad = rc + rj;
    float sum = 0.f;
#pragma unroll
    for (int d = 0; d < D; d++) {
        const float a = mj[d] - bhi[d], b = mj[d] - blo[d];
        float lb = a > 0.f ? a : (b < 0.f ? -b : 0.f);
        if (b > 1.0f) {
            const float e = 2.0f * ceilf(0.5f * b);  // smallest even number >= b
            const float w = (e - 2.0f >= fmaxf(a, 1.0f) && e >= 4.0f) ? 0.f : e - b;
            lb = fminf(lb, w);
        }
        lb = fmaxf(lb - (1e-5f * rad + 2e-6f * (1.0f + fabsf(mj[d]) + fmaxf(fabsf(blo[d]), fabsf(bhi[d])))), 0.f);
        sum = fmaf(lb, lb, sum);
    }
    return !(sum > rad * rad * 1.0001f);
}

template <int D>
__global__ void __launch_bounds__(256) nb_pack_conics_kernel(int64_t P, const float* __restrict__ conics, float4* __restrict__ cpack)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    cpack[i] = D == 2 ? make_float4(conics[i * 3], conics[i * 3 + 1], conics[i * 3 + 2], 0.f) : make_float4(conics[i], 0.f, 0.f, 0.f);
}

// ---- cell-cooperative path: one CTA per grid cell ---------------------------------------------------
struct NbCellSmem {
    uint32_t id[NBC_CAP];  // candidate ids (ascending after the sort of the fill pass)
    float x[NBC_CAP], y[NBC_CAP], r[NBC_CAP];
};

struct NbCellShared {  // small static part
    float red[NBC_WARPS * 5];
    uint32_t n;
    uint32_t wrap;
    uint16_t ring[NBC_WARPS][128];
};

// block-uniform: does the cell hold a row of this call's window?
__device__ __forceinline__ bool nbc_any_row_in(NbWindow win, const float4* __restrict__ srec, uint32_t rs, uint32_t re)
{
    bool any = false;
    for (uint32_t p = rs + threadIdx.x; p < re; p += NBC_THREADS) any |= win.has(__float_as_uint(srec[p].w));
    return __syncthreads_or(any) != 0;
}

// Builds the candidate superset of cell c in shared memory: every Gaussian of the reachable cells that passes
// nb_box_reach against the cell's bounding box and largest radius.  SORT = true orders it by ascending id (bitonic
// sort of the ids, records re-read by id); SORT = false keeps arrival order (enough for counting).
// Returns the number of candidates, or -1 when there are more than `cap` (then the per-row kernels take the cell).
// sh.wrap == 0 afterwards iff every candidate lies within 1 of every point of the box in every dimension: then
// neither the torus branch of the reference distance nor the wrap of the emitted offsets can trigger for this cell.
// The list is padded to a multiple of 64 with records no row collides with when sh.wrap == 0 (x = 3e38: dx^2 = inf).
template <int D, bool SORT>
__device__ __forceinline__ int nbc_load_cell(const NbGrid& g, const uint32_t* __restrict__ cell_start,
                                             const float4* __restrict__ srec, const float* __restrict__ means,
                                             const float* __restrict__ reff, int cap, NbCellSmem& sm, NbCellShared& sh,
                                             uint32_t rs, uint32_t re)
{
    const int tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
    // bounding box and largest radius of the cell's own Gaussians
    float v[5] = {__int_as_float(0x7f800000), __int_as_float(0x7f800000), __int_as_float(0xff800000),
                  __int_as_float(0xff800000), 0.f};
    for (uint32_t p = rs + tid; p < re; p += NBC_THREADS) {
        const float4 o = srec[p];
        v[0] = fminf(v[0], o.x);
        v[1] = fminf(v[1], o.y);
        v[2] = fmaxf(v[2], o.x);
        v[3] = fmaxf(v[3], o.y);
        v[4] = fmaxf(v[4], o.z);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        v[0] = fminf(v[0], __shfl_xor_sync(0xffffffffu, v[0], o));
        v[1] = fminf(v[1], __shfl_xor_sync(0xffffffffu, v[1], o));
        v[2] = fmaxf(v[2], __shfl_xor_sync(0xffffffffu, v[2], o));
        v[3] = fmaxf(v[3], __shfl_xor_sync(0xffffffffu, v[3], o));
        v[4] = fmaxf(v[4], __shfl_xor_sync(0xffffffffu, v[4], o));
    }
    if (lane == 0)
        for (int k = 0; k < 5; k++) sh.red[wp * 5 + k] = v[k];
    if (tid == 0) {
        sh.n = 0;
        sh.wrap = 0;
    }
    __syncthreads();
    for (int w2 = 0; w2 < NBC_WARPS; w2++) {
        v[0] = fminf(v[0], sh.red[w2 * 5 + 0]);
        v[1] = fminf(v[1], sh.red[w2 * 5 + 1]);
        v[2] = fmaxf(v[2], sh.red[w2 * 5 + 2]);
        v[3] = fmaxf(v[3], sh.red[w2 * 5 + 3]);
        v[4] = fmaxf(v[4], sh.red[w2 * 5 + 4]);
    }
    const float T = v[4] + g.rmax;
    Spans sx, sy;
    dim_spans(v[0], v[2], T, g, 0, sx);
    if (D == 2)
        dim_spans(v[1], v[3], T, g, 1, sy);
    else {
        sy.n = 1;
        sy.lo[0] = sy.hi[0] = 0;
    }
    const float blo[2] = {v[0], v[1]}, bhi[2] = {v[2], v[3]};
    bool wrap = false;
    for (int iy = 0; iy < sy.n; iy++)
        for (int cy = sy.lo[iy]; cy <= sy.hi[iy]; cy++)
            for (int ix = 0; ix < sx.n; ix++) {
                const uint32_t b = cell_start[cy * g.n[0] + sx.lo[ix]];
                const uint32_t e = cell_start[cy * g.n[0] + sx.hi[ix] + 1];
                for (uint32_t base = b; base < e; base += NBC_THREADS) {
                    const uint32_t p = base + tid;
                    bool keep = false;
                    float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (p < e) {
                        o = srec[p];
                        const float mj[2] = {o.x, o.y};
                        keep = nb_box_reach<D>(blo, bhi, v[4], mj, o.z);
                    }
                    const uint32_t m = __ballot_sync(0xffffffffu, keep);
                    uint32_t at = 0;
                    if (lane == 0 && m) at = atomicAdd(&sh.n, (uint32_t)__popc(m));
                    at = __shfl_sync(0xffffffffu, at, 0) + __popc(m & ((1u << lane) - 1u));
                    if (keep && at < (uint32_t)cap) {
                        sm.id[at] = __float_as_uint(o.w);
                        if (!SORT) {
                            sm.x[at] = o.x;
                            sm.y[at] = o.y;
                            sm.r[at] = o.z;
                        }
                        wrap |= __fsub_rn(o.x, blo[0]) > 1.0f || __fsub_rn(o.x, bhi[0]) < -1.0f;
                        if (D == 2) wrap |= __fsub_rn(o.y, blo[1]) > 1.0f || __fsub_rn(o.y, bhi[1]) < -1.0f;
                    }
                }
            }
    if (wrap) sh.wrap = 1;
    __syncthreads();
    const uint32_t n = sh.n;
    if (n > (uint32_t)cap) return -1;
    if (SORT) {
        uint32_t n2 = 2;
        while (n2 < n) n2 <<= 1;
        for (uint32_t i = n + tid; i < n2; i += NBC_THREADS) sm.id[i] = 0xffffffffu;
        __syncthreads();
        for (uint32_t k = 2; k <= n2; k <<= 1)
            for (uint32_t j = k >> 1; j > 0; j >>= 1) {
                for (uint32_t t = tid; t < (n2 >> 1); t += NBC_THREADS) {
                    const uint32_t i0 = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                    const uint32_t i1 = i0 | j;
                    const uint32_t a = sm.id[i0], b2 = sm.id[i1];
                    if ((a > b2) == ((i0 & k) == 0)) {
                        sm.id[i0] = b2;
                        sm.id[i1] = a;
                    }
                }
                __syncthreads();
            }
        for (uint32_t i = tid; i < n; i += NBC_THREADS) {
            const uint32_t j = sm.id[i];
            sm.x[i] = means[(size_t)j * D];
            sm.y[i] = D == 2 ? means[(size_t)j * D + 1] : 0.f;
            sm.r[i] = reff[j];
        }
    }
    for (uint32_t i = n + tid; i < ((n + 63u) & ~63u); i += NBC_THREADS) {
        sm.x[i] = 3.0e38f;
        sm.y[i] = 0.f;
        sm.r[i] = 0.f;
    }
    __syncthreads();
    return (int)n;
}

// nb_collides for rows of a cell whose candidates never take the torus branch (sh.wrap == 0): same operations, no branch
template <int D, bool WRAP>
__device__ __forceinline__ bool nbc_collides(const float* mi, float ri, const float* mj, float rj)
{
    if (WRAP) return nb_collides<D>(mi, ri, mj, rj);
    float dist = 0.0f;
#pragma unroll
    for (int d = 0; d < D; d++) {
        const float dx = __fsub_rn(mj[d], mi[d]);
        dist = __fmaf_rn(dx, dx, dist);
    }
    const float rad = __fadd_rn(ri, rj);
    return !(dist > __fmul_rn(rad, rad));
}

template <int D, bool WRAP>
__device__ __forceinline__ void nbc_count_rows(const NbCellSmem& sm, int n, uint32_t rs, uint32_t re,
                                               const float4* __restrict__ srec, int64_t* __restrict__ counts, NbWindow win)
{
    const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
    for (uint32_t r = rs + wp; r < re; r += NBC_WARPS) {
        const float4 me = srec[r];
        if (!win.has(__float_as_uint(me.w))) continue;
        const float mi[2] = {me.x, me.y};
        uint32_t cnt = 0;
        const int nn = WRAP ? n : ((n + 63) & ~63);  // the padding never collides on the branch-free path
#pragma unroll 4
        for (int i = lane; i < nn; i += 32) {
            const float mj[2] = {sm.x[i], sm.y[i]};
            cnt += nbc_collides<D, WRAP>(mi, me.z, mj, sm.r[i]);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        if (lane == 0) counts[__float_as_uint(me.w)] = (int64_t)cnt;
    }
}

template <int D>
__global__ void __launch_bounds__(NBC_THREADS) nb_cell_count_kernel(const NbGrid* __restrict__ gp,
                                                                    const uint32_t* __restrict__ cell_start,
                                                                    const float4* __restrict__ srec,
                                                                    const float* __restrict__ means,
                                                                    const float* __restrict__ reff, int cap,
                                                                    uint8_t* __restrict__ cell_over, int64_t* __restrict__ counts,
                                                                    NbWindow win, int windowed)
{
    extern __shared__ __align__(16) unsigned char nbc_raw[];
    NbCellSmem& sm = *reinterpret_cast<NbCellSmem*>(nbc_raw);
    __shared__ NbCellShared sh;
    const NbGrid g = *gp;
    for (uint32_t c = blockIdx.x; c < (uint32_t)g.ncells; c += gridDim.x) {
        const uint32_t rs = cell_start[c], re = cell_start[c + 1];
        if (rs == re) {
            if (threadIdx.x == 0) cell_over[c] = 0;
            continue;
        }
        if (windowed && !nbc_any_row_in(win, srec, rs, re)) {  // no row of this shard in the cell
            if (threadIdx.x == 0) cell_over[c] = 0;
            continue;
        }
        __syncthreads();  // the previous cell's readers are done with shared memory
        const int n = nbc_load_cell<D, false>(g, cell_start, srec, means, reff, cap, sm, sh, rs, re);
        if (threadIdx.x == 0) cell_over[c] = n < 0;
        if (n < 0) continue;
        if (sh.wrap)
            nbc_count_rows<D, true>(sm, n, rs, re, srec, counts, win);
        else
            nbc_count_rows<D, false>(sm, n, rs, re, srec, counts, win);
    }
}

// Rows of one cell: candidates are walked in ascending id; accepted ones are queued (their position in the list) in
// a 128-entry per-warp ring (two batches of 32 per step) and leave in full groups of 32, so the CSR stores are whole-warp and consecutive.
template <int D, bool WRAP>
__device__ __forceinline__ void nbc_fill_rows(const NbCellSmem& sm, NbCellShared& sh, int n, uint32_t rs, uint32_t re,
                                              const float4* __restrict__ srec, const float* __restrict__ conics,
                                              const float* __restrict__ radii, const int64_t* __restrict__ ranges,
                                              int64_t* __restrict__ indices, float* __restrict__ dists,
                                              float* __restrict__ densities, float* __restrict__ inv_total, NbWindow win)
{
    const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    uint16_t* ring = sh.ring[wp];
    for (uint32_t r = rs + wp; r < re; r += NBC_WARPS) {
        const float4 me = srec[r];
        const uint32_t id = __float_as_uint(me.w);
        if (!win.has(id)) {  // another shard's row: empty list
            if (lane == 0) inv_total[id] = __double2float_rn(1.0 / (0.0 + 1e-6));
            continue;
        }
        const float mi[2] = {me.x, me.y};
        const int64_t start = id == 0 ? 0 : ranges[id - 1];
        const float inv_r = nb_inv_radius(radii[id]);
        float total = 0.0f;
        uint32_t cnt = 0, done = 0;  // accepted so far; emitted so far (multiple of 32)
        const int nn = (n + 63) & ~63;
        for (int base = 0; base < nn; base += 64) {
            const int i0 = base + lane, i1 = i0 + 32;
            const float a0[2] = {sm.x[i0], sm.y[i0]}, a1[2] = {sm.x[i1], sm.y[i1]};
            bool ok0 = nbc_collides<D, WRAP>(mi, me.z, a0, sm.r[i0]);
            bool ok1 = nbc_collides<D, WRAP>(mi, me.z, a1, sm.r[i1]);
            if (WRAP) {  // the padding is only inert on the branch-free path
                ok0 = ok0 && i0 < n;
                ok1 = ok1 && i1 < n;
            }
            const uint32_t m0 = __ballot_sync(0xffffffffu, ok0), m1 = __ballot_sync(0xffffffffu, ok1);
            const uint32_t c0 = cnt + __popc(m0);
            if (ok0) ring[(cnt + __popc(m0 & lt)) & 127u] = (uint16_t)i0;
            if (ok1) ring[(c0 + __popc(m1 & lt)) & 127u] = (uint16_t)i1;
            cnt = c0 + __popc(m1);
            while (cnt - done >= 32u) {  // a group of 32 accepted candidates is complete (warp-uniform)
                __syncwarp();
                const int ii = ring[(done + lane) & 127u];
                const float mj[2] = {sm.x[ii], sm.y[ii]};
                total += nb_emit_core<D, !WRAP, true>(start + done + lane, sm.id[ii], mj, mi, inv_r, conics, indices, dists,
                                                densities);
                done += 32u;
            }
            __syncwarp();
        }
        if (done + lane < cnt) {
            const int ii = ring[(done + lane) & 127u];
            const float mj[2] = {sm.x[ii], sm.y[ii]};
            total += nb_emit_core<D, !WRAP, true>(start + done + lane, sm.id[ii], mj, mi, inv_r, conics, indices, dists, densities);
        }
        __syncwarp();
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
        if (lane == 0) inv_total[id] = __double2float_rn(1.0 / ((double)total + 1e-6));
    }
}

template <int D>
__global__ void __launch_bounds__(NBC_THREADS) nb_cell_fill_kernel(const NbGrid* __restrict__ gp,
                                                                   const uint32_t* __restrict__ cell_start,
                                                                   const float4* __restrict__ srec,
                                                                   const float* __restrict__ means,
                                                                   const float* __restrict__ reff, int cap,
                                                                   const uint8_t* __restrict__ cell_over,
                                                                   const float* __restrict__ conics,
                                                                   const float* __restrict__ radii,
                                                                   const int64_t* __restrict__ ranges,
                                                                   int64_t* __restrict__ indices, float* __restrict__ dists,
                                                                   float* __restrict__ densities, float* __restrict__ inv_total,
                                                                   NbWindow win, int windowed)
{
    extern __shared__ __align__(16) unsigned char nbc_raw[];
    NbCellSmem& sm = *reinterpret_cast<NbCellSmem*>(nbc_raw);
    __shared__ NbCellShared sh;
    const NbGrid g = *gp;
    for (uint32_t c = blockIdx.x; c < (uint32_t)g.ncells; c += gridDim.x) {
        const uint32_t rs = cell_start[c], re = cell_start[c + 1];
        if (rs == re || cell_over[c]) continue;
        if (windowed && !nbc_any_row_in(win, srec, rs, re)) {
            for (uint32_t p = rs + threadIdx.x; p < re; p += NBC_THREADS)
                inv_total[__float_as_uint(srec[p].w)] = __double2float_rn(1.0 / (0.0 + 1e-6));
            continue;
        }
        __syncthreads();
        const int n = nbc_load_cell<D, true>(g, cell_start, srec, means, reff, cap, sm, sh, rs, re);
        if (n < 0) continue;  // cannot happen: same inputs as the count pass
        if (sh.wrap)
            nbc_fill_rows<D, true>(sm, sh, n, rs, re, srec, conics, radii, ranges, indices, dists, densities, inv_total, win);
        else
            nbc_fill_rows<D, false>(sm, sh, n, rs, re, srec, conics, radii, ranges, indices, dists, densities, inv_total, win);
    }
}

// ---- small P: brute force, one launch per stage -------------------------------------------------------
// The reference's own models hold 400 .. a few thousand Gaussians (main_pn.py: nx = ny = 20).  There the dozen dependent
// launches of the grid path (grid build, key sort, cell tables, two cell passes, per-row passes) cost more than testing
// every pair: one warp per row walks ALL candidates in ascending id, 32 per step — the reference's own loop order
// (findCollisions scans i = 0..P-1), so the accepted ones leave in ascending id without any sort.  Same tests and the
// same per-slot arithmetic as the grid path (nb_collides, nb_emit_core): indices / ranges bit for bit.
constexpr int NBS_THREADS = 256;
constexpr int NBS_WARPS = NBS_THREADS / 32;
constexpr int64_t NBS_P_MAX = 4096;

__device__ __forceinline__ float nbs_reff(float radius) { return __double2float_rn((double)radius * 0.2); }  // nb_prepare_kernel

template <int D, bool FILL>
__global__ void __launch_bounds__(NBS_THREADS) nb_small_kernel(int64_t P, const float* __restrict__ means,
                                                               const float* __restrict__ conics,
                                                               const float* __restrict__ radii,
                                                               const int64_t* __restrict__ ranges,
                                                               int64_t* __restrict__ counts, int64_t* __restrict__ indices,
                                                               float* __restrict__ dists, float* __restrict__ densities,
                                                               float* __restrict__ inv_total, NbWindow win)
{
    const int lane = threadIdx.x & 31;
    const int64_t id = (int64_t)blockIdx.x * NBS_WARPS + (threadIdx.x >> 5);
    if (id >= P) return;
    const float ri = nbs_reff(radii[id]);
    const bool row = !((double)ri < 1e-6) && win.has((uint32_t)id);  // degenerate rows / another shard's rows: empty list
    if (!row) {
        if (lane == 0) {
            if (FILL)
                inv_total[id] = __double2float_rn(1.0 / (0.0 + 1e-6));
            else
                counts[id] = 0;
        }
        return;
    }
    float mi[2] = {0.f, 0.f};
#pragma unroll
    for (int d = 0; d < D; d++) mi[d] = means[id * D + d];
    const int64_t start = FILL ? (id == 0 ? 0 : ranges[id - 1]) : 0;
    const float inv_r = FILL ? nb_inv_radius(radii[id]) : 0.f;
    float total = 0.f;
    uint32_t cnt = 0;
    for (int64_t base = 0; base < P; base += 32) {
        const int64_t j = base + lane;
        bool ok = false;
        float mj[2] = {0.f, 0.f};
        if (j < P) {
            const float rj = nbs_reff(radii[j]);
#pragma unroll
            for (int d = 0; d < D; d++) mj[d] = means[j * D + d];
            ok = !((double)rj < 1e-6) && nb_collides<D>(mi, ri, mj, rj);
        }
        const uint32_t m = __ballot_sync(0xffffffffu, ok);
        if (FILL && ok)
            total += nb_emit_core<D>(start + cnt + __popc(m & ((1u << lane) - 1u)), (uint32_t)j, mj, mi, inv_r, conics, indices,
                                     dists, densities);
        cnt += __popc(m);
    }
    if (FILL) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
        if (lane == 0) inv_total[id] = __double2float_rn(1.0 / ((double)total + 1e-6));
    } else if (lane == 0) {
        counts[id] = (int64_t)cnt;
    }
}

// Which path a search took is remembered on the host per workspace (the count and the fill stage must agree even if the
// knob changes between them)
static int64_t g_nb_small_p = -1;  // -1: not read yet (env PIGS_NB_SMALL_P, default NBS_P_MAX); 0 = never
static int64_t nb_small_p()
{
    if (g_nb_small_p < 0) {
        const char* e = getenv("PIGS_NB_SMALL_P");
        g_nb_small_p = e ? atoll(e) : NBS_P_MAX;
        if (g_nb_small_p < 0) g_nb_small_p = 0;
    }
    return g_nb_small_p;
}
namespace {
struct NbNote {
    const void* ws;
    bool small;
};
NbNote g_nb_notes[64];
int g_nb_note_next = 0;
std::mutex g_nb_note_mu;
void nb_note(const void* ws, bool small)
{
    std::lock_guard<std::mutex> lk(g_nb_note_mu);
    for (auto& n : g_nb_notes)
        if (n.ws == ws) {
            n.small = small;
            return;
        }
    g_nb_notes[g_nb_note_next] = NbNote{ws, small};
    g_nb_note_next = (g_nb_note_next + 1) % 64;
}
bool nb_noted_small(const void* ws)
{
    std::lock_guard<std::mutex> lk(g_nb_note_mu);
    for (auto& n : g_nb_notes)
        if (n.ws == ws) return n.small;
    return false;
}
}  // namespace

// run-time knobs of the neighbour search (pigs_set_neighbor_search; development and tests): cell_cap = 0 routes every
// row through the per-row kernels, cell_div = grid cells per largest radius.  Environment defaults:
// PIGS_NB_CELL_CAP, PIGS_NB_CELL_DIV.
static int g_nb_cell_cap = -1;
static float g_nb_cell_div = 0.f;
static int nb_cell_cap()
{
    if (g_nb_cell_cap < 0) {
        const char* e = getenv("PIGS_NB_CELL_CAP");
        g_nb_cell_cap = e ? atoi(e) : NBC_CAP;
        if (g_nb_cell_cap < 0 || g_nb_cell_cap > NBC_CAP) g_nb_cell_cap = NBC_CAP;
    }
    return g_nb_cell_cap;
}
static float nb_cell_div()
{
    if (g_nb_cell_div == 0.f) {
        const char* e = getenv("PIGS_NB_CELL_DIV");
        g_nb_cell_div = e ? (float)atof(e) : 2.0f;
        if (!(g_nb_cell_div >= 0.25f && g_nb_cell_div <= 16.f)) g_nb_cell_div = 2.0f;
    }
    return g_nb_cell_div;
}

// ====================================================================================================
// aggregation (aggregate_neighbors.cu:129-321)
// ====================================================================================================
constexpr int AG_THREADS = 256;
constexpr int AG_WARPS = AG_THREADS / 32;
constexpr int AG_MAX_L = 128, AG_MAX_K = 128;

// sin / cos of (frequency * pi * X).  The reference evaluates the argument and sin/cos in double
// (`sin(frequencies[e] * M_PI * X[d])`, aggregate_neighbors.cu:181-182); here the product f*X is formed
// exactly as p + e (two floats), reduced modulo 2 exactly, and fed to sincospif: abs error ~2e-7.
__device__ __forceinline__ void sincos_fpx(float f, float X, float& s, float& c)
{
    const float p = __fmul_rn(f, X);
    const float e = __fmaf_rn(f, X, -p);
    const float r = __fsub_rn(p, 2.0f * rintf(0.5f * p));
    sincospif(__fadd_rn(r, e), &s, &c);
}

// scalar-type helpers of the generic aggregation kernels: T = float (the product) or double (the reference's
// `typedef double FLOAT` build, cuda_sampler/config.h:20 — what its one live test, torch.autograd.gradcheck of
// aggregate_neighbors, test_neighbor_aggregation.py:94-101, needs)
__device__ __forceinline__ float tfma(float a, float b, float c) { return fmaf(a, b, c); }
__device__ __forceinline__ double tfma(double a, double b, double c) { return fma(a, b, c); }
__device__ __forceinline__ void sincos_fpx_t(float f, float X, float& s, float& c) { sincos_fpx(f, X, s, c); }
__device__ __forceinline__ void sincos_fpx_t(double f, double X, double& s, double& c)
{
    sincos(f * 3.14159265358979323846 * X, &s, &c);  // the reference's own expression (aggregate_neighbors.cu:181-182)
}

struct AggDims {
    int D, L, K, E, nF, stride;  // stride = (E-1)/D, nF = stride/2
};

template <typename T>
__global__ void __launch_bounds__(AG_THREADS)
    agg_forward_kernel(int64_t row_lo, int64_t P, AggDims dm, const T* __restrict__ features, const T* __restrict__ transform,
                       const T* __restrict__ queries, const T* __restrict__ keys, const T* __restrict__ freqs,
                       const T* __restrict__ dt, const int64_t* __restrict__ indices, const int64_t* __restrict__ ranges,
                       const T* __restrict__ dists, const T* __restrict__ densities,
                       const T* __restrict__ inv_total, T* __restrict__ weights, T* __restrict__ embeddings,
                       T* __restrict__ factors, T* __restrict__ out)
{
    extern __shared__ __align__(16) unsigned char agg_raw[];
    T* sm = reinterpret_cast<T*>(agg_raw);
    const int D = dm.D, L = dm.L, K = dm.K, E = dm.E, nF = dm.nF, stride = dm.stride;
    T* s_dt = sm;                 // [2E]
    T* s_fr = s_dt + 2 * E;       // [nF]
    T* wbase = s_fr + nF;
    const int per_warp = K + L + 3 * 32;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    T* s_q = wbase + (size_t)w * per_warp;  // [K]
    T* s_z = s_q + K;                       // [L]
    T* s_dwe = s_z + L;                     // [32]
    T* s_dwf = s_dwe + 32;                  // [32]
    int* s_j = (int*)(s_dwf + 32);              // [32]
    for (int i = threadIdx.x; i < 2 * E; i += AG_THREADS) s_dt[i] = dt[i];
    for (int i = threadIdx.x; i < nF; i += AG_THREADS) s_fr[i] = freqs[i];
    __syncthreads();

    for (int64_t row = row_lo + (int64_t)blockIdx.x * AG_WARPS + w; row < P; row += (int64_t)gridDim.x * AG_WARPS) {  // rows [row_lo, P)
        const int64_t start = row == 0 ? 0 : ranges[row - 1], end = ranges[row];
        for (int k = lane; k < K; k += 32) s_q[k] = queries[row * K + k];
        const T inv = inv_total[row];
        T z[AG_MAX_L / 32];
#pragma unroll
        for (int a = 0; a < AG_MAX_L / 32; a++) z[a] = T(0.0);
        __syncwarp();
        for (int64_t chunk = start; chunk < end; chunk += 32) {
            const int64_t s = chunk + lane;
            const int64_t j = s < end ? indices[s] : -1;
            T dwe = T(0.0), dwf = T(0.0);
            if (j >= 0) {
                const T* key = keys + j * K;
                T wgt = T(0.0);
                for (int k = 0; k < K; k++) wgt = tfma(s_q[k], key[k], wgt);
                T emb = T(0.0), fac = T(0.0);
                for (int d = 0; d < D; d++) {
                    const T X = dists[s * D + d];
                    for (int e = 0; e < nF; e++) {
                        T sn, cs;
                        sincos_fpx_t(s_fr[e], X, sn, cs);
                        const int i0 = d * stride + 2 * e;
                        emb = tfma(s_dt[i0], sn, emb);
                        emb = tfma(s_dt[i0 + 1], cs, emb);
                        fac = tfma(s_dt[E + i0], sn, fac);
                        fac = tfma(s_dt[E + i0 + 1], cs, fac);
                    }
                }
                emb += s_dt[E - 1];
                fac += s_dt[2 * E - 1];
                weights[s] = wgt;
                embeddings[s] = emb;
                factors[s] = fac;
                const T dw = inv * densities[s] * wgt;
                dwf = dw * fac;
                dwe = dw * emb;
            } else if (s < end) {
                // skipped neighbour: every slot of the row is written, so the caller's outputs need no zero fill
                weights[s] = T(0.0);
                embeddings[s] = T(0.0);
                factors[s] = T(0.0);
            }
            s_dwe[lane] = dwe;
            s_dwf[lane] = dwf;
            s_j[lane] = (int)j;
            __syncwarp();
            const int cnt = (int)min((int64_t)32, end - chunk);
            for (int t = 0; t < cnt; t++) {
                const int jj = s_j[t];
                if (jj < 0) continue;
                const T a0 = s_dwe[t], a1 = s_dwf[t];
                const T* f = features + (size_t)jj * L;
#pragma unroll
                for (int a = 0; a < AG_MAX_L / 32; a++)
                    if (a * 32 + lane < L) z[a] += tfma(a1, f[a * 32 + lane], a0);
            }
            __syncwarp();
        }
#pragma unroll
        for (int a = 0; a < AG_MAX_L / 32; a++)
            if (a * 32 + lane < L) s_z[a * 32 + lane] = z[a];
        __syncwarp();
        // out_i = z . T  (the reference applies T inside the neighbour loop, :201-206)
        for (int b = lane; b < L; b += 32) {
            T acc = T(0.0);
            for (int a = 0; a < L; a++) acc = tfma(s_z[a], transform[a * L + b], acc);
            out[row * L + b] = acc;
        }
        __syncwarp();
    }
}

// Backward.  Per row i with upstream g_i:  st = T g_i,  A = sum(st),  and per neighbour slot (j):
//   B = st . feat_j,  dc = density * inv_total_i,  dcw = dc * weight,  TE = dc * (emb * A + fac * B)
//   dL/dfeat_j   += dcw * fac * st                 dL/dkey_j += q_i * TE      dL/dq_i += key_j * TE
//   dL/dT[a][b]  += g_i[b] * sum_s w_s * dc_s * (emb_s + fac_s * feat_j[a])
//   dL/ddt, dL/dfreq: the per-slot sums over channels collapse to A and B (reference :266-281)
// Shared-parameter gradients are accumulated per lane in shared-memory columns (no conflicts), reduced
// once per CTA, then one atomic per address per CTA.
template <typename T>
__global__ void __launch_bounds__(AG_THREADS)
    agg_backward_kernel(int64_t row_lo, int64_t P, AggDims dm, const T* __restrict__ features, const T* __restrict__ transform,
                        const T* __restrict__ queries, const T* __restrict__ keys, const T* __restrict__ freqs,
                        const T* __restrict__ dt, const int64_t* __restrict__ indices, const int64_t* __restrict__ ranges,
                        const T* __restrict__ dists, const T* __restrict__ densities,
                        const T* __restrict__ weights, const T* __restrict__ embeddings,
                        const T* __restrict__ factors, const T* __restrict__ inv_total,
                        const T* __restrict__ gout, T* __restrict__ g_features, T* __restrict__ g_transform,
                        T* __restrict__ g_queries, T* __restrict__ g_keys, T* __restrict__ g_freqs,
                        T* __restrict__ g_dt)
{
    extern __shared__ __align__(16) unsigned char agg_raw[];
    T* sm = reinterpret_cast<T*>(agg_raw);
    const int D = dm.D, L = dm.L, K = dm.K, E = dm.E, nF = dm.nF, stride = dm.stride;
    const int ncol = 2 * E + nF;
    T* s_dt = sm;                     // [2E]
    T* s_fr = s_dt + 2 * E;           // [nF]
    T* s_gt = s_fr + nF;              // [L*L]  CTA accumulator of dL/dT
    T* wbase = s_gt + L * L;
    const int per_warp = K + 3 * L + 6 * 32 + ncol * 32;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    T* s_q = wbase + (size_t)w * per_warp;  // [K]
    T* s_st = s_q + K;                      // [L]
    T* s_g = s_st + L;                      // [L]
    T* s_wz = s_g + L;                      // [L]
    T* s_te = s_wz + L;                     // [32] x5 + j
    T* s_c1 = s_te + 32;
    T* s_w = s_c1 + 32;
    T* s_dce = s_w + 32;
    T* s_dcf = s_dce + 32;
    int* s_j = (int*)(s_dcf + 32);              // [32] ints inside a slot of 32 T
    T* s_col = s_dcf + 64;                      // [ncol][32] per-lane partial sums of dL/ddt, dL/dfreq
    for (int i = threadIdx.x; i < 2 * E; i += AG_THREADS) s_dt[i] = dt[i];
    for (int i = threadIdx.x; i < nF; i += AG_THREADS) s_fr[i] = freqs[i];
    for (int i = threadIdx.x; i < L * L; i += AG_THREADS) s_gt[i] = T(0.0);
    for (int i = lane; i < ncol * 32; i += 32) s_col[i] = T(0.0);
    __syncthreads();
    const T PI = T(3.14159265358979323846);

    for (int64_t row = row_lo + (int64_t)blockIdx.x * AG_WARPS + w; row < P; row += (int64_t)gridDim.x * AG_WARPS) {  // rows [row_lo, P)
        const int64_t start = row == 0 ? 0 : ranges[row - 1], end = ranges[row];
        if (end == start) continue;
        for (int k = lane; k < K; k += 32) s_q[k] = queries[row * K + k];
        for (int b = lane; b < L; b += 32) s_g[b] = gout[row * L + b];
        __syncwarp();
        T Apart = T(0.0);
        for (int a = lane; a < L; a += 32) {
            T acc = T(0.0);
            for (int b = 0; b < L; b++) acc = tfma(transform[a * L + b], s_g[b], acc);
            s_st[a] = acc;
            Apart += acc;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) Apart += __shfl_xor_sync(0xffffffffu, Apart, o);
        const T A = Apart;
        const T inv = inv_total[row];
        T wz[AG_MAX_L / 32], gq[AG_MAX_K / 32];
#pragma unroll
        for (int a = 0; a < AG_MAX_L / 32; a++) wz[a] = T(0.0);
#pragma unroll
        for (int k = 0; k < AG_MAX_K / 32; k++) gq[k] = T(0.0);
        __syncwarp();
        for (int64_t chunk = start; chunk < end; chunk += 32) {
            const int64_t s = chunk + lane;
            const int64_t j = s < end ? indices[s] : -1;
            T te = T(0), c1 = T(0), wgt = T(0), dce = T(0), dcf = T(0);
            if (j >= 0) {
                const T* f = features + j * L;
                T B = T(0.0);
                for (int a = 0; a < L; a++) B = tfma(s_st[a], f[a], B);
                wgt = weights[s];
                const T emb = embeddings[s], fac = factors[s];
                const T dc = densities[s] * inv;
                const T dcw = dc * wgt;
                dce = dc * emb;
                dcf = dc * fac;
                te = tfma(dce, A, dcf * B);
                c1 = dcw * fac;
                const T dA = dcw * A, dB = dcw * B;
                for (int d = 0; d < D; d++) {
                    const T X = dists[s * D + d];
                    for (int e = 0; e < nF; e++) {
                        T sn, cs;
                        sincos_fpx_t(s_fr[e], X, sn, cs);
                        const int i0 = d * stride + 2 * e;
                        s_col[(i0)*32 + lane] += dA * sn;
                        s_col[(E + i0) * 32 + lane] += dB * sn;
                        s_col[(i0 + 1) * 32 + lane] += dA * cs;
                        s_col[(E + i0 + 1) * 32 + lane] += dB * cs;
                        const T ts = tfma(s_dt[E + i0], dB, s_dt[i0] * dA);
                        const T tc = tfma(s_dt[E + i0 + 1], dB, s_dt[i0 + 1] * dA);
                        s_col[(2 * E + e) * 32 + lane] += PI * X * tfma(cs, ts, -sn * tc);
                    }
                }
                s_col[(E - 1) * 32 + lane] += dA;
                s_col[(2 * E - 1) * 32 + lane] += dB;
            }
            s_te[lane] = te;
            s_c1[lane] = c1;
            s_w[lane] = wgt;
            s_dce[lane] = dce;
            s_dcf[lane] = dcf;
            s_j[lane] = (int)j;
            __syncwarp();
            const int cnt = (int)min((int64_t)32, end - chunk);
            for (int t = 0; t < cnt; t++) {
                const int jj = s_j[t];
                if (jj < 0) continue;
                const T t_te = s_te[t], t_c1 = s_c1[t], t_w = s_w[t], t_dce = s_dce[t], t_dcf = s_dcf[t];
                const T* f = features + (size_t)jj * L;
#pragma unroll
                for (int a = 0; a < AG_MAX_L / 32; a++) {
                    const int ai = a * 32 + lane;
                    if (ai < L) {
                        atomicAdd(g_features + (size_t)jj * L + ai, t_c1 * s_st[ai]);
                        wz[a] = tfma(t_w, tfma(t_dcf, f[ai], t_dce), wz[a]);
                    }
                }
                const T* key = keys + (size_t)jj * K;
#pragma unroll
                for (int k = 0; k < AG_MAX_K / 32; k++) {
                    const int ki = k * 32 + lane;
                    if (ki < K) {
                        gq[k] = tfma(key[ki], t_te, gq[k]);
                        atomicAdd(g_keys + (size_t)jj * K + ki, s_q[ki] * t_te);
                    }
                }
            }
            __syncwarp();
        }
#pragma unroll
        for (int k = 0; k < AG_MAX_K / 32; k++)
            if (k * 32 + lane < K) g_queries[row * K + k * 32 + lane] = gq[k];
#pragma unroll
        for (int a = 0; a < AG_MAX_L / 32; a++)
            if (a * 32 + lane < L) s_wz[a * 32 + lane] = wz[a];
        __syncwarp();
        for (int b = lane; b < L; b += 32) {
            const T gb = s_g[b];
            for (int a = 0; a < L; a++) atomicAdd(&s_gt[a * L + b], s_wz[a] * gb);
        }
        __syncwarp();
    }
    __syncthreads();
    for (int i = threadIdx.x; i < L * L; i += AG_THREADS)
        if (s_gt[i] != T(0.0)) atomicAdd(g_transform + i, s_gt[i]);
    for (int i = threadIdx.x; i < ncol; i += AG_THREADS) {
        T acc = T(0.0);
        for (int ww = 0; ww < AG_WARPS; ww++) {
            const T* col = wbase + (size_t)ww * per_warp + (K + 3 * L + 6 * 32) + i * 32;
            for (int l = 0; l < 32; l++) acc += col[l];
        }
        if (acc != T(0.0)) atomicAdd(i < 2 * E ? g_dt + i : g_freqs + (i - 2 * E), acc);
    }
}

}  // namespace pigs

namespace pigs {
// aggregate16.cu: kernels specialised for L = K = 16, E = 25 (model_pn.py); return 1 if handled, 0 to fall back
int launch_aggregate16_forward(int64_t row_lo, int64_t P, int D, int L, int K, int E, const float* features, const float* transform,
                               const float* queries, const float* keys, const float* freqs, const float* dt,
                               const int64_t* indices, const int64_t* ranges, const float* dists, const float* densities,
                               const float* inv_total, float* weights, float* embeddings, float* factors, float* out,
                               cudaStream_t st);
int launch_aggregate16_backward(int64_t row_lo, int64_t P, int D, int L, int K, int E, const float* features, const float* transform,
                                const float* queries, const float* keys, const float* freqs, const float* dt,
                                const int64_t* indices, const int64_t* ranges, const float* dists, const float* densities,
                                const float* weights, const float* embeddings, const float* factors, const float* inv_total,
                                const float* gout, float* g_features, float* g_transform, float* g_queries, float* g_keys,
                                float* g_freqs, float* g_dt, cudaStream_t st);
}  // namespace pigs

using namespace pigs;

namespace {

bool g_agg_fast = true;  // pigs_set_aggregate_specialised(0): always run the generic kernels (tests compare both)

int bit_len(uint32_t n)
{
    int b = 0;
    while (n) {
        b++;
        n >>= 1;
    }
    return b;
}

size_t agg_fwd_smem(int L, int K, int E, int nF, size_t elem = sizeof(float)) { return elem * (2 * E + nF + (size_t)AG_WARPS * (K + L + 3 * 32)); }
size_t agg_bwd_smem(int L, int K, int E, int nF, size_t elem = sizeof(float))
{
    return elem * (2 * E + nF + (size_t)L * L + (size_t)AG_WARPS * (K + 3 * L + 6 * 32 + (2 * E + nF) * 32));
}

int agg_dims(AggDims& dm, int D, int L, int K, int E)
{
    PIGS_ARG(D == 1 || D == 2);
    PIGS_ARG(L >= 1 && L <= AG_MAX_L);
    PIGS_ARG(K >= 1 && K <= AG_MAX_K);
    PIGS_ARG(E >= 1);
    dm.D = D;
    dm.L = L;
    dm.K = K;
    dm.E = E;
    dm.stride = (E - 1) / D;
    dm.nF = (E - 1) / D / 2;
    return 0;
}

}  // namespace

extern "C" {

size_t pigs_neighbor_workspace_bytes(int64_t P, int D) { return NbLayout(nullptr, P, D == 1 ? 1 : 2).bytes; }

int pigs_neighbors_count(int64_t P, int D, const float* means, const float* radii, int64_t* counts, void* workspace,
                         void* stream)
{
    return pigs_neighbors_count_rows(P, D, means, radii, 0, P, counts, workspace, stream);
}

int pigs_neighbors_count_rows(int64_t P, int D, const float* means, const float* radii, int64_t row_lo, int64_t row_hi,
                              int64_t* counts, void* workspace, void* stream)
{
    PIGS_ARG(D == 1 || D == 2);
    PIGS_ARG(P > 0 && P < (1ll << 31) && means && radii && counts && workspace);
    PIGS_ARG(row_lo >= 0 && row_lo <= row_hi && row_hi <= P);
    const NbWindow win{(uint32_t)row_lo, (uint32_t)row_hi};
    const int windowed = row_lo > 0 || row_hi < P;
    cudaStream_t st = (cudaStream_t)stream;
    NbLayout w(workspace, P, D);
    const bool small = P <= nb_small_p();
    nb_note(workspace, small);
    if (small) {  // brute force: one launch, no grid
        const unsigned nbs = (unsigned)((P + NBS_WARPS - 1) / NBS_WARPS);
        if (D == 2)
            nb_small_kernel<2, false><<<nbs, NBS_THREADS, 0, st>>>(P, means, nullptr, radii, nullptr, counts, nullptr, nullptr,
                                                                   nullptr, nullptr, win);
        else
            nb_small_kernel<1, false><<<nbs, NBS_THREADS, 0, st>>>(P, means, nullptr, radii, nullptr, counts, nullptr, nullptr,
                                                                   nullptr, nullptr, win);
        PIGS_LAUNCH_CHECK();
        return 0;
    }
    const unsigned nb = (unsigned)((P + 255) / 256);
    PIGS_CHECK(cudaMemsetAsync(counts, 0, (size_t)P * sizeof(int64_t), st));
    nb_init_kernel<<<1, 32, 0, st>>>(w.grid);
    PIGS_LAUNCH_CHECK();
    if (D == 2)
        nb_prepare_kernel<2><<<nb, 256, 0, st>>>(P, means, radii, w.reff, w.grid);
    else
        nb_prepare_kernel<1><<<nb, 256, 0, st>>>(P, means, radii, w.reff, w.grid);
    PIGS_LAUNCH_CHECK();
    nb_grid_kernel<<<1, 32, 0, st>>>(w.grid, D, NbLayout::per_dim_cap(P, D), nb_cell_div());
    PIGS_LAUNCH_CHECK();
    if (D == 2)
        nb_key_kernel<2><<<nb, 256, 0, st>>>(P, means, w.reff, w.grid, w.keys_a, w.vals_a);
    else
        nb_key_kernel<1><<<nb, 256, 0, st>>>(P, means, w.reff, w.grid, w.keys_a, w.vals_a);
    PIGS_LAUNCH_CHECK();
    // stable sort by cell id: every cell's list stays ascending in Gaussian id
    const int bits = bit_len((uint32_t)w.ncells_max + 1u);
    int rc = stable_sort_pairs_u32(w.keys_a, w.vals_a, w.keys_b, w.vals_b, P, bits, w.sort_ws, st);
    if (rc) return rc;
    nb_cell_start_kernel<<<nb, 256, 0, st>>>(P, w.keys_b, w.grid, w.cell_start);
    PIGS_LAUNCH_CHECK();
    if (D == 2)
        nb_gather_kernel<2><<<nb, 256, 0, st>>>(P, w.vals_b, means, w.reff, w.srec);
    else
        nb_gather_kernel<1><<<nb, 256, 0, st>>>(P, w.vals_b, means, w.reff, w.srec);
    PIGS_LAUNCH_CHECK();
    // cells first (shared id-sorted candidate list), then the rows of the cells that did not fit
    const int cap = nb_cell_cap();
    const unsigned ncta = (unsigned)std::min<int64_t>(w.ncells_max, (int64_t)num_sms() * 64);
    if (D == 2) {
        PIGS_CHECK(cudaFuncSetAttribute(nb_cell_count_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(NbCellSmem)));
        nb_cell_count_kernel<2><<<ncta, NBC_THREADS, sizeof(NbCellSmem), st>>>(w.grid, w.cell_start, w.srec, means, w.reff,
                                                                               cap, w.cell_over, counts, win, windowed);
    } else {
        PIGS_CHECK(cudaFuncSetAttribute(nb_cell_count_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(NbCellSmem)));
        nb_cell_count_kernel<1><<<ncta, NBC_THREADS, sizeof(NbCellSmem), st>>>(w.grid, w.cell_start, w.srec, means, w.reff,
                                                                               cap, w.cell_over, counts, win, windowed);
    }
    PIGS_LAUNCH_CHECK();
    const unsigned nbw = (unsigned)((P + NB_WARPS - 1) / NB_WARPS);
    if (D == 2)
        nb_count_kernel<2><<<nbw, NB_THREADS, 0, st>>>(P, w.grid, w.cell_start, w.srec, counts, w.keys_b, w.cell_over, win);
    else
        nb_count_kernel<1><<<nbw, NB_THREADS, 0, st>>>(P, w.grid, w.cell_start, w.srec, counts, w.keys_b, w.cell_over, win);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int pigs_neighbors_fill(int64_t P, int D, const float* means, const float* conics, const float* radii,
                        const int64_t* ranges, int64_t* indices, float* dists, float* densities,
                        float* inv_total_densities, void* workspace, void* stream)
{
    return pigs_neighbors_fill_rows(P, D, means, conics, radii, 0, P, ranges, indices, dists, densities, inv_total_densities,
                                    workspace, stream);
}

int pigs_neighbors_fill_rows(int64_t P, int D, const float* means, const float* conics, const float* radii, int64_t row_lo,
                             int64_t row_hi, const int64_t* ranges, int64_t* indices, float* dists, float* densities,
                             float* inv_total_densities, void* workspace, void* stream)
{
    PIGS_ARG(D == 1 || D == 2);
    PIGS_ARG(P > 0 && means && conics && radii && ranges && inv_total_densities && workspace);
    PIGS_ARG(row_lo >= 0 && row_lo <= row_hi && row_hi <= P);
    const NbWindow win{(uint32_t)row_lo, (uint32_t)row_hi};
    const int windowed = row_lo > 0 || row_hi < P;
    cudaStream_t st = (cudaStream_t)stream;
    NbLayout w(workspace, P, D);
    if (nb_noted_small(workspace)) {  // the count stage of this workspace ran the brute-force kernel
        const unsigned nbs = (unsigned)((P + NBS_WARPS - 1) / NBS_WARPS);
        if (D == 2)
            nb_small_kernel<2, true><<<nbs, NBS_THREADS, 0, st>>>(P, means, conics, radii, ranges, nullptr, indices, dists,
                                                                  densities, inv_total_densities, win);
        else
            nb_small_kernel<1, true><<<nbs, NBS_THREADS, 0, st>>>(P, means, conics, radii, ranges, nullptr, indices, dists,
                                                                  densities, inv_total_densities, win);
        PIGS_LAUNCH_CHECK();
        return 0;
    }
    const unsigned nbw = (unsigned)((P + NB_WARPS - 1) / NB_WARPS);
    const int cap = nb_cell_cap();
    const unsigned ncta = (unsigned)std::min<int64_t>(w.ncells_max, (int64_t)num_sms() * 64);
    if (D == 2)
        nb_pack_conics_kernel<2><<<(unsigned)((P + 255) / 256), 256, 0, st>>>(P, conics, w.cpack);
    else
        nb_pack_conics_kernel<1><<<(unsigned)((P + 255) / 256), 256, 0, st>>>(P, conics, w.cpack);
    PIGS_LAUNCH_CHECK();
    if (D == 2) {
        PIGS_CHECK(cudaFuncSetAttribute(nb_cell_fill_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(NbCellSmem)));
        nb_cell_fill_kernel<2><<<ncta, NBC_THREADS, sizeof(NbCellSmem), st>>>(w.grid, w.cell_start, w.srec, means, w.reff,
                                                                              cap, w.cell_over, (const float*)w.cpack, radii, ranges,
                                                                              indices, dists, densities, inv_total_densities, win,
                                                                              windowed);
    } else {
        PIGS_CHECK(cudaFuncSetAttribute(nb_cell_fill_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(NbCellSmem)));
        nb_cell_fill_kernel<1><<<ncta, NBC_THREADS, sizeof(NbCellSmem), st>>>(w.grid, w.cell_start, w.srec, means, w.reff,
                                                                              cap, w.cell_over, (const float*)w.cpack, radii, ranges,
                                                                              indices, dists, densities, inv_total_densities, win,
                                                                              windowed);
    }
    PIGS_LAUNCH_CHECK();
    if (D == 2) {
        nb_fill_kernel<2><<<nbw, NB_THREADS, 0, st>>>(P, w.grid, w.cell_start, w.srec, means, conics, radii, ranges, indices,
                                                      dists, densities, inv_total_densities, w.big_list, w.keys_b, w.cell_over, win);
        PIGS_LAUNCH_CHECK();
        nb_fill_big_kernel<2><<<NB_BIG_CTAS, NB_BIG_THREADS, 0, st>>>(P, w.grid, w.cell_start, w.srec, means, conics, radii,
                                                                      ranges, indices, dists, densities, inv_total_densities,
                                                                      w.big_list, w.bitmaps, w.words);
    } else {
        nb_fill_kernel<1><<<nbw, NB_THREADS, 0, st>>>(P, w.grid, w.cell_start, w.srec, means, conics, radii, ranges, indices,
                                                      dists, densities, inv_total_densities, w.big_list, w.keys_b, w.cell_over, win);
        PIGS_LAUNCH_CHECK();
        nb_fill_big_kernel<1><<<NB_BIG_CTAS, NB_BIG_THREADS, 0, st>>>(P, w.grid, w.cell_start, w.srec, means, conics, radii,
                                                                      ranges, indices, dists, densities, inv_total_densities,
                                                                      w.big_list, w.bitmaps, w.words);
    }
    PIGS_LAUNCH_CHECK();
    return 0;
}

size_t pigs_neighbors_scan_workspace_bytes(int64_t P) { return scan_ws_bytes(P); }

int pigs_neighbors_ranges(int64_t P, const int64_t* counts, int64_t* ranges, int64_t* total, void* workspace, void* stream)
{
    PIGS_ARG(P > 0 && counts && ranges && total && workspace);
    return inclusive_scan_i64(counts, ranges, P, workspace, (cudaStream_t)stream, (long long*)total);
}

void pigs_set_aggregate_specialised(int on) { g_agg_fast = on != 0; }

void pigs_set_neighbor_small(int64_t max_P) { g_nb_small_p = max_P < 0 ? 0 : max_P; }

void pigs_set_neighbor_search(int cell_cap, float cell_div)
{
    g_nb_cell_cap = cell_cap < 0 || cell_cap > NBC_CAP ? NBC_CAP : cell_cap;
    g_nb_cell_div = (cell_div >= 0.25f && cell_div <= 16.f) ? cell_div : 2.0f;
}

int pigs_aggregate_forward(int64_t P, int D, int L, int K, int E, const float* features, const float* transform,
                           const float* queries, const float* keys, const float* frequencies,
                           const float* distance_transform, const int64_t* indices, const int64_t* ranges,
                           const float* dists, const float* densities, const float* inv_total_densities, float* weights,
                           float* embeddings, float* factors, float* neighbor_features, void* stream)
{
    return pigs_aggregate_forward_rows(P, D, L, K, E, 0, P, features, transform, queries, keys, frequencies, distance_transform,
                                       indices, ranges, dists, densities, inv_total_densities, weights, embeddings, factors,
                                       neighbor_features, stream);
}

int pigs_aggregate_forward_rows(int64_t P, int D, int L, int K, int E, int64_t row_lo, int64_t row_hi, const float* features, const float* transform,
                           const float* queries, const float* keys, const float* frequencies,
                           const float* distance_transform, const int64_t* indices, const int64_t* ranges,
                           const float* dists, const float* densities, const float* inv_total_densities, float* weights,
                           float* embeddings, float* factors, float* neighbor_features, void* stream)
{
    AggDims dm;
    int rc = agg_dims(dm, D, L, K, E);
    if (rc) return rc;
    PIGS_ARG(P > 0 && features && transform && queries && keys && distance_transform && ranges && inv_total_densities &&
             neighbor_features);
    PIGS_ARG(dm.nF == 0 || frequencies);
    PIGS_ARG(row_lo >= 0 && row_lo <= row_hi && row_hi <= P);
    if (row_lo == row_hi) return 0;
    if (g_agg_fast && frequencies && weights && embeddings && factors && indices && dists && densities) {
        rc = launch_aggregate16_forward(row_lo, row_hi, D, L, K, E, features, transform, queries, keys, frequencies, distance_transform,
                                        indices, ranges, dists, densities, inv_total_densities, weights, embeddings, factors,
                                        neighbor_features, (cudaStream_t)stream);
        if (rc != 0) return rc == 1 ? 0 : rc;
    }
    const size_t smem = agg_fwd_smem(L, K, E, dm.nF);
    PIGS_ARG(smem <= 200 * 1024);
    PIGS_CHECK(cudaFuncSetAttribute(agg_forward_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t grid = (row_hi - row_lo + AG_WARPS - 1) / AG_WARPS;
    const int64_t cap = (int64_t)num_sms() * 16;
    if (grid > cap) grid = cap;
    agg_forward_kernel<float><<<(unsigned)grid, AG_THREADS, smem, (cudaStream_t)stream>>>(
        row_lo, row_hi, dm, features, transform, queries, keys, frequencies, distance_transform, indices, ranges, dists, densities,
        inv_total_densities, weights, embeddings, factors, neighbor_features);
    PIGS_LAUNCH_CHECK();
    return 0;
}

size_t pigs_aggregate_backward_workspace_bytes(int64_t P, int L, int K, int E)
{
    (void)P, (void)L, (void)K, (void)E;
    return 0;  // all partial sums live in shared memory
}

int pigs_aggregate_backward(int64_t P, int D, int L, int K, int E, int64_t M, const float* features,
                            const float* transform, const float* queries, const float* keys, const float* frequencies,
                            const float* distance_transform, const int64_t* indices, const int64_t* ranges,
                            const float* dists, const float* densities, const float* weights, const float* embeddings,
                            const float* factors, const float* inv_total_densities, const float* dL_dneighbor_features,
                            float* dL_dfeatures, float* dL_dtransform, float* dL_dqueries, float* dL_dkeys,
                            float* dL_dfrequencies, float* dL_ddistance_transform, void* workspace, void* stream)
{
    return pigs_aggregate_backward_rows(P, D, L, K, E, M, 0, P, features, transform, queries, keys, frequencies,
                                        distance_transform, indices, ranges, dists, densities, weights, embeddings, factors,
                                        inv_total_densities, dL_dneighbor_features, dL_dfeatures, dL_dtransform, dL_dqueries,
                                        dL_dkeys, dL_dfrequencies, dL_ddistance_transform, workspace, stream);
}

int pigs_aggregate_backward_rows(int64_t P, int D, int L, int K, int E, int64_t M, int64_t row_lo, int64_t row_hi, const float* features,
                            const float* transform, const float* queries, const float* keys, const float* frequencies,
                            const float* distance_transform, const int64_t* indices, const int64_t* ranges,
                            const float* dists, const float* densities, const float* weights, const float* embeddings,
                            const float* factors, const float* inv_total_densities, const float* dL_dneighbor_features,
                            float* dL_dfeatures, float* dL_dtransform, float* dL_dqueries, float* dL_dkeys,
                            float* dL_dfrequencies, float* dL_ddistance_transform, void* workspace, void* stream)
{
    (void)M, (void)workspace;
    AggDims dm;
    int rc = agg_dims(dm, D, L, K, E);
    if (rc) return rc;
    PIGS_ARG(P > 0 && features && transform && queries && keys && distance_transform && ranges && inv_total_densities &&
             dL_dneighbor_features && dL_dfeatures && dL_dtransform && dL_dqueries && dL_dkeys && dL_ddistance_transform);
    PIGS_ARG(dm.nF == 0 || (frequencies && dL_dfrequencies));
    PIGS_ARG(row_lo >= 0 && row_lo <= row_hi && row_hi <= P);
    if (row_lo == row_hi) return 0;
    if (g_agg_fast && frequencies && dL_dfrequencies) {
        rc = launch_aggregate16_backward(row_lo, row_hi, D, L, K, E, features, transform, queries, keys, frequencies, distance_transform,
                                         indices, ranges, dists, densities, weights, embeddings, factors, inv_total_densities,
                                         dL_dneighbor_features, dL_dfeatures, dL_dtransform, dL_dqueries, dL_dkeys,
                                         dL_dfrequencies, dL_ddistance_transform, (cudaStream_t)stream);
        if (rc != 0) return rc == 1 ? 0 : rc;
    }
    const size_t smem = agg_bwd_smem(L, K, E, dm.nF);
    PIGS_ARG(smem <= 200 * 1024);
    PIGS_CHECK(cudaFuncSetAttribute(agg_backward_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t grid = (row_hi - row_lo + AG_WARPS - 1) / AG_WARPS;
    const int64_t cap = (int64_t)num_sms() * 2;
    if (grid > cap) grid = cap;
    agg_backward_kernel<float><<<(unsigned)grid, AG_THREADS, smem, (cudaStream_t)stream>>>(
        row_lo, row_hi, dm, features, transform, queries, keys, frequencies, distance_transform, indices, ranges, dists, densities, weights,
        embeddings, factors, inv_total_densities, dL_dneighbor_features, dL_dfeatures, dL_dtransform, dL_dqueries, dL_dkeys,
        dL_dfrequencies, dL_ddistance_transform);
    PIGS_LAUNCH_CHECK();
    return 0;
}

// ---- float64 instantiation of the generic aggregation kernels (the reference's `typedef double FLOAT` build) -----------
int pigs_aggregate_forward_f64(int64_t P, int D, int L, int K, int E, const double* features, const double* transform,
                               const double* queries, const double* keys, const double* frequencies,
                               const double* distance_transform, const int64_t* indices, const int64_t* ranges,
                               const double* dists, const double* densities, const double* inv_total_densities,
                               double* weights, double* embeddings, double* factors, double* neighbor_features, void* stream)
{
    AggDims dm;
    int rc = agg_dims(dm, D, L, K, E);
    if (rc) return rc;
    PIGS_ARG(P > 0 && features && transform && queries && keys && distance_transform && ranges && inv_total_densities &&
             neighbor_features && weights && embeddings && factors);
    PIGS_ARG(dm.nF == 0 || frequencies);
    const size_t smem = agg_fwd_smem(L, K, E, dm.nF, sizeof(double));
    PIGS_ARG(smem <= 200 * 1024);
    PIGS_CHECK(cudaFuncSetAttribute(agg_forward_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t grid = (P + AG_WARPS - 1) / AG_WARPS;
    const int64_t cap = (int64_t)num_sms() * 8;
    if (grid > cap) grid = cap;
    agg_forward_kernel<double><<<(unsigned)grid, AG_THREADS, smem, (cudaStream_t)stream>>>(
        0, P, dm, features, transform, queries, keys, frequencies, distance_transform, indices, ranges, dists, densities,
        inv_total_densities, weights, embeddings, factors, neighbor_features);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int pigs_aggregate_backward_f64(int64_t P, int D, int L, int K, int E, const double* features, const double* transform,
                                const double* queries, const double* keys, const double* frequencies,
                                const double* distance_transform, const int64_t* indices, const int64_t* ranges,
                                const double* dists, const double* densities, const double* weights, const double* embeddings,
                                const double* factors, const double* inv_total_densities, const double* dL_dneighbor_features,
                                double* dL_dfeatures, double* dL_dtransform, double* dL_dqueries, double* dL_dkeys,
                                double* dL_dfrequencies, double* dL_ddistance_transform, void* stream)
{
    AggDims dm;
    int rc = agg_dims(dm, D, L, K, E);
    if (rc) return rc;
    PIGS_ARG(P > 0 && features && transform && queries && keys && distance_transform && ranges && inv_total_densities &&
             dL_dneighbor_features && dL_dfeatures && dL_dtransform && dL_dqueries && dL_dkeys && dL_ddistance_transform);
    PIGS_ARG(dm.nF == 0 || (frequencies && dL_dfrequencies));
    const size_t smem = agg_bwd_smem(L, K, E, dm.nF, sizeof(double));
    PIGS_ARG(smem <= 200 * 1024);
    PIGS_CHECK(cudaFuncSetAttribute(agg_backward_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t grid = (P + AG_WARPS - 1) / AG_WARPS;
    const int64_t cap = (int64_t)num_sms() * 2;
    if (grid > cap) grid = cap;
    agg_backward_kernel<double><<<(unsigned)grid, AG_THREADS, smem, (cudaStream_t)stream>>>(
        0, P, dm, features, transform, queries, keys, frequencies, distance_transform, indices, ranges, dists, densities, weights,
        embeddings, factors, inv_total_densities, dL_dneighbor_features, dL_dfeatures, dL_dtransform, dL_dqueries, dL_dkeys,
        dL_dfrequencies, dL_ddistance_transform);
    PIGS_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
