// primitives.cu — hand-written exclusive scan and stable LSD radix sort (u32 key/value pairs).
// Replaces cub::DeviceScan::InclusiveSum and cub::DeviceRadixSort::SortPairs used by the reference
// (cuda_sampler/sampler_impl.cu:253,279,316).  HBM-bound; sized for 148 SMs.
#include "common.cuh"

namespace pigs {

thread_local char g_err[512] = {0};
int64_t g_launches = 0;
int g_recip_mode = 1;
int g_debug_sync = 0;

int set_error(int code, const char* what, const char* file, int line)
{
    const char* cs = code > 0 ? cudaGetErrorString((cudaError_t)code) : "invalid argument";
    snprintf(g_err, sizeof(g_err), "%s: %s (%s:%d)", cs, what, file, line);
    return code;
}

int num_sms()
{
    static int n = 0;
    if (!n) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

// ------------------------------------------------------------------------------------------------
// scan
// ------------------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;  // 2048

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) >= o) v += t;
    }
    return v;
}

// block-wide exclusive scan of one value per thread (blockDim = SCAN_THREADS); returns exclusive
// prefix, *total = block sum
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* total)
{
    __shared__ uint32_t wsum[SCAN_THREADS / 32];
    __shared__ uint32_t tot;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t inc = warp_incl_scan(v);
    if (lane == 31) wsum[w] = inc;
    __syncthreads();
    if (w == 0) {
        uint32_t s = lane < SCAN_THREADS / 32 ? wsum[lane] : 0;
        uint32_t si = warp_incl_scan(s);
        if (lane < SCAN_THREADS / 32) wsum[lane] = si - s;
        if (lane == SCAN_THREADS / 32 - 1) tot = si;
    }
    __syncthreads();
    uint32_t r = inc - v + wsum[w];
    *total = tot;
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_block_sums(const uint32_t* __restrict__ in, uint32_t* __restrict__ bsum,
                                                                int64_t n)
{
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        int64_t idx = base + i * SCAN_THREADS + threadIdx.x;
        if (idx < n) s += in[idx];
    }
    uint32_t tot;
    block_excl_scan(s, &tot);
    if (threadIdx.x == 0) bsum[blockIdx.x] = tot;
}

// single block: exclusive scan of bsum[nb] in place; bsum[nb] = grand total
__global__ void __launch_bounds__(SCAN_THREADS) scan_of_sums(uint32_t* bsum, int64_t nb)
{
    uint32_t carry = 0;
    for (int64_t base = 0; base < nb; base += SCAN_THREADS) {
        int64_t idx = base + threadIdx.x;
        uint32_t v = idx < nb ? bsum[idx] : 0;
        uint32_t tot;
        uint32_t ex = block_excl_scan(v, &tot);
        if (idx < nb) bsum[idx] = ex + carry;
        carry += tot;
    }
    if (threadIdx.x == 0) bsum[nb] = carry;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_apply(const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                                                           const uint32_t* __restrict__ bsum, int64_t n, int64_t nb,
                                                           int write_total)
{
    // each thread owns SCAN_ITEMS consecutive elements so the order is preserved
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS];
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        v[i] = (base + i < n) ? in[base + i] : 0;
        s += v[i];
    }
    uint32_t tot;
    uint32_t ex = block_excl_scan(s, &tot) + bsum[blockIdx.x];
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        if (base + i < n) out[base + i] = ex;
        ex += v[i];
    }
    if (write_total && blockIdx.x == 0 && threadIdx.x == 0) out[n] = bsum[nb];
}

// small inputs (the reference's own drivers run 1K-40K Gaussians): ONE CTA walks the tiles with a running carry —
// one launch instead of three; out may alias in (a tile is read into registers before it is written)
constexpr int64_t SCAN_SMALL_MAX = 16 * SCAN_TILE;
__global__ void __launch_bounds__(SCAN_THREADS) scan_small(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int64_t n,
                                                           int write_total)
{
    uint32_t carry = 0;
    for (int64_t tile = 0; tile < n; tile += SCAN_TILE) {
        const int64_t base = tile + (int64_t)threadIdx.x * SCAN_ITEMS;
        uint32_t v[SCAN_ITEMS];
        uint32_t s = 0;
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; i++) {
            v[i] = (base + i < n) ? in[base + i] : 0;
            s += v[i];
        }
        uint32_t tot;
        uint32_t ex = block_excl_scan(s, &tot) + carry;
#pragma unroll
        for (int i = 0; i < SCAN_ITEMS; i++) {
            if (base + i < n) out[base + i] = ex;
            ex += v[i];
        }
        carry += tot;
    }
    if (write_total && threadIdx.x == 0) out[n] = carry;
}

size_t scan_ws_bytes(int64_t n)
{
    int64_t nb = (n + SCAN_TILE - 1) / SCAN_TILE;
    return align_up((size_t)(nb + 2) * sizeof(uint32_t));
}

int exclusive_scan_u32(const uint32_t* in, uint32_t* out, int64_t n, bool write_total, void* ws, cudaStream_t st)
{
    uint32_t* bsum = (uint32_t*)ws;
    if (n <= 0) {
        if (write_total) PIGS_CHECK(cudaMemsetAsync(out, 0, sizeof(uint32_t), st));
        return 0;
    }
    if (n <= SCAN_SMALL_MAX) {
        scan_small<<<1, SCAN_THREADS, 0, st>>>(in, out, n, write_total ? 1 : 0);
        PIGS_LAUNCH_CHECK();
        return 0;
    }
    const int64_t nb = (n + SCAN_TILE - 1) / SCAN_TILE;
    // NB: scan_block_sums reads the tile strided, scan_apply reads it blocked — both sum the same
    // SCAN_TILE elements, so the per-block totals agree.
    scan_block_sums<<<(unsigned)nb, SCAN_THREADS, 0, st>>>(in, bsum, n);
    PIGS_LAUNCH_CHECK();
    scan_of_sums<<<1, SCAN_THREADS, 0, st>>>(bsum, nb);
    PIGS_LAUNCH_CHECK();
    scan_apply<<<(unsigned)nb, SCAN_THREADS, 0, st>>>(in, out, bsum, n, nb, write_total ? 1 : 0);
    PIGS_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------------------------------------
// stable LSD radix sort, 8-bit digits
//   pass = histogram (per CTA, per digit) -> exclusive scan over [digit][cta] -> stable scatter.
// A CTA owns SORT_TILE consecutive elements; each warp owns a consecutive sub-range and walks it
// in order 32 elements at a time, ranking equal digits with __match_any_sync, so the scatter is
// stable without any per-element block barrier.
// ------------------------------------------------------------------------------------------------
constexpr int SORT_THREADS = 256;
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int SORT_ROUNDS = 16;                                // 32-element groups per warp
constexpr int SORT_WARP_ITEMS = 32 * SORT_ROUNDS;              // 512
constexpr int SORT_TILE = SORT_WARP_ITEMS * SORT_WARPS;        // 4096
constexpr int RADIX = 256;

__global__ void __launch_bounds__(SORT_THREADS) radix_hist(const uint32_t* __restrict__ keys, int64_t n, int shift,
                                                           uint32_t mask, uint32_t* __restrict__ ghist, int64_t nb)
{
    __shared__ uint32_t hist[RADIX];
    for (int i = threadIdx.x; i < RADIX; i += SORT_THREADS) hist[i] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * SORT_TILE;
#pragma unroll 4
    for (int i = 0; i < SORT_TILE / SORT_THREADS; i++) {
        int64_t idx = base + i * SORT_THREADS + threadIdx.x;
        if (idx < n) atomicAdd(&hist[(keys[idx] >> shift) & mask], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < RADIX; i += SORT_THREADS) ghist[(int64_t)i * nb + blockIdx.x] = hist[i];
}

__global__ void __launch_bounds__(SORT_THREADS) radix_scatter(const uint32_t* __restrict__ keys,
                                                              const uint32_t* __restrict__ vals, int64_t n, int shift,
                                                              uint32_t mask, const uint32_t* __restrict__ gbase,
                                                              int64_t nb, uint32_t* __restrict__ keys_out,
                                                              uint32_t* __restrict__ vals_out)
{
    __shared__ uint32_t wcount[SORT_WARPS][RADIX];  // per-warp digit counts, then per-warp running offsets
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < SORT_WARPS * RADIX; i += SORT_THREADS) (&wcount[0][0])[i] = 0;
    __syncthreads();

    const int64_t wbase = (int64_t)blockIdx.x * SORT_TILE + (int64_t)w * SORT_WARP_ITEMS;
    uint32_t k[SORT_ROUNDS], v[SORT_ROUNDS];
    // phase 1: load + per-warp digit counts
#pragma unroll
    for (int r = 0; r < SORT_ROUNDS; r++) {
        const int64_t idx = wbase + r * 32 + lane;
        const bool ok = idx < n;
        k[r] = ok ? keys[idx] : 0xffffffffu;
        v[r] = ok ? vals[idx] : 0u;
        const uint32_t d = ok ? ((k[r] >> shift) & mask) : RADIX;  // RADIX = "invalid" bucket
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        if (ok && lane == (__ffs(peers) - 1)) wcount[w][d] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    // phase 2: per digit, exclusive scan over warps, plus the global base of (digit, this CTA)
    for (int d = threadIdx.x; d < RADIX; d += SORT_THREADS) {
        uint32_t run = gbase[(int64_t)d * nb + blockIdx.x];
#pragma unroll
        for (int ww = 0; ww < SORT_WARPS; ww++) {
            uint32_t c = wcount[ww][d];
            wcount[ww][d] = run;
            run += c;
        }
    }
    __syncthreads();
    // phase 3: in-order walk, stable scatter
#pragma unroll
    for (int r = 0; r < SORT_ROUNDS; r++) {
        const int64_t idx = wbase + r * 32 + lane;
        const bool ok = idx < n;
        const uint32_t d = ok ? ((k[r] >> shift) & mask) : RADIX;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        if (ok) {
            const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
            const uint32_t pos = wcount[w][d] + rank;
            keys_out[pos] = k[r];
            vals_out[pos] = v[r];
        }
        __syncwarp();
        if (ok && lane == (__ffs(peers) - 1)) wcount[w][d] += __popc(peers);
        __syncwarp();
    }
}

size_t sort_ws_bytes(int64_t n)
{
    int64_t nb = (n + SORT_TILE - 1) / SORT_TILE;
    if (nb < 1) nb = 1;
    size_t hist = align_up((size_t)(RADIX * nb + 1) * sizeof(uint32_t));
    return hist + scan_ws_bytes(RADIX * nb) + 2 * align_up((size_t)(n > 0 ? n : 1) * sizeof(uint32_t));
}

int stable_sort_pairs_u32(uint32_t* keys_in, uint32_t* vals_in, uint32_t* keys_out, uint32_t* vals_out, int64_t n,
                          int bits, void* ws, cudaStream_t st)
{
    if (n <= 0) return 0;
    const int64_t nb = (n + SORT_TILE - 1) / SORT_TILE;
    Carver c(ws);
    uint32_t* ghist = c.take<uint32_t>(RADIX * nb + 1);
    char* scan_ws = c.take<char>(scan_ws_bytes(RADIX * nb));
    uint32_t* tmp_k = c.take<uint32_t>(n);
    uint32_t* tmp_v = c.take<uint32_t>(n);
    int passes = (bits + 7) / 8;
    if (passes < 1) passes = 1;
    // choose ping-pong so the final pass writes keys_out/vals_out
    uint32_t *src_k = keys_in, *src_v = vals_in;
    for (int p = 0; p < passes; p++) {
        const bool last = (p == passes - 1);
        uint32_t* dst_k = last ? keys_out : (src_k == tmp_k ? keys_in : tmp_k);
        uint32_t* dst_v = last ? vals_out : (src_v == tmp_v ? vals_in : tmp_v);
        if (last && (src_k == keys_out)) return set_error(-1, "sort aliasing", __FILE__, __LINE__);
        const int shift = p * 8;
        const int pb = bits - shift >= 8 ? 8 : (bits - shift > 0 ? bits - shift : 8);
        const uint32_t mask = (1u << pb) - 1u;
        radix_hist<<<(unsigned)nb, SORT_THREADS, 0, st>>>(src_k, n, shift, mask, ghist, nb);
        PIGS_LAUNCH_CHECK();
        int rc = exclusive_scan_u32(ghist, ghist, RADIX * nb, false, scan_ws, st);
        if (rc) return rc;
        radix_scatter<<<(unsigned)nb, SORT_THREADS, 0, st>>>(src_k, src_v, n, shift, mask, ghist, nb, dst_k, dst_v);
        PIGS_LAUNCH_CHECK();
        src_k = dst_k;
        src_v = dst_v;
    }
    return 0;
}

}  // namespace pigs
