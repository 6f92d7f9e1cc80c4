// primitives.cu — hand-written single-pass scan (decoupled look-back) and stable LSD radix sort (u32 key/value pairs).
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

bool g_stage_timing = false;
namespace {
struct StageMark {
    cudaEvent_t ev;
    const char* name;
};
StageMark g_marks[256];
int g_nmarks = 0;
}  // namespace

void stage_mark(cudaStream_t st, const char* name)
{
    if (g_nmarks >= 256) return;
    StageMark& m = g_marks[g_nmarks];
    if (!m.ev) cudaEventCreate(&m.ev);
    m.name = name;
    cudaEventRecord(m.ev, st);
    g_nmarks++;
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

// ---- single-pass scan with decoupled look-back -------------------------------------------------------------------
// One launch for any n (replaces cub::DeviceScan::InclusiveSum, sampler_impl.cu:253, and this file's former
// three-kernel reduce / scan-of-sums / apply chain): every CTA takes the next tile from an atomic ticket (so a tile only
// ever waits for tiles whose CTAs are already running), scans it, publishes its aggregate, and warp 0 walks the
// predecessors' descriptors backwards until it meets an inclusive prefix.  A descriptor is ONE 64-bit word
// (status << 62 | value), written and read atomically, so no fence is needed between status and value.
// ws = [ticket u64][desc u64 x tiles], zeroed by one cudaMemsetAsync before the launch.
constexpr unsigned long long SCAN_AGG = 1ull << 62, SCAN_INC = 2ull << 62, SCAN_VAL = (1ull << 62) - 1ull;

__device__ __forceinline__ unsigned long long ld_desc(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_desc(unsigned long long* p, unsigned long long v)
{
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// T = uint32_t (tile-touch counts, flags, histograms) or int64_t (neighbour counts).  INCLUSIVE: out[i] = sum in[0..i];
// else out[i] = sum in[0..i) and, with write_total, out[n] = grand total.  total_out (optional): the grand total as
// int64 (device), e.g. the word of a header the host reads back.  out may alias in.
// look-back tiles are twice the size of the small-scan tiles: the chain of predecessors a tile may have to walk (32 per
// step, one L2 round trip each) is what a mid-sized scan waits on (15 us for 1M elements with 2048-element tiles)
constexpr int LB_ITEMS = 16;
constexpr int LB_TILE = SCAN_THREADS * LB_ITEMS;  // 4096

// Round 2b: each CTA owns a SPAN of consecutive tiles (at most one CTA per SM), so the look-back chain has <= 148 links
// instead of one per 4096 elements: pass 1 sums the span and publishes the aggregate (all CTAs reach this point at about the
// same time, so the walk back finds every predecessor ready), pass 2 scans the span tile by tile with a running carry
// (second read from L2).  Measured with one tile per CTA: 18 us for 1M u32, 33 us for 2M u64 — the L2 round trips of
// hundreds of tiles walking back through each other's descriptors.
template <typename T, bool INCLUSIVE>
__global__ void __launch_bounds__(SCAN_THREADS) scan_lookback_kernel(const T* in, T* out, int64_t n,
                                                                     int write_total, unsigned long long* __restrict__ ws,
                                                                     long long* __restrict__ total_out, int64_t span_tiles)
{
    __shared__ unsigned long long s_wsum[SCAN_THREADS / 32];
    __shared__ unsigned long long s_prefix;
    __shared__ unsigned int s_tile;
    unsigned long long* desc = ws + 1;
    if (threadIdx.x == 0) s_tile = (unsigned int)atomicAdd(ws, 1ull);  // span index by start order
    __syncthreads();
    const int64_t cta = s_tile;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t s0 = cta * span_tiles * LB_TILE;
    const int64_t s1 = (s0 + span_tiles * LB_TILE) < n ? (s0 + span_tiles * LB_TILE) : n;
    // ---- pass 1: the span's sum ----
    unsigned long long s = 0;
    for (int64_t i = s0 + threadIdx.x; i < s1; i += SCAN_THREADS) s += (unsigned long long)in[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) s_wsum[w] = s;
    __syncthreads();
    unsigned long long agg = 0;
#pragma unroll
    for (int k = 0; k < SCAN_THREADS / 32; k++) agg += s_wsum[k];
    if (w == 0) {
        unsigned long long excl = 0;
        if (cta == 0) {
            if (lane == 0) st_desc(desc, SCAN_INC | agg);
        } else {
            if (lane == 0) st_desc(desc + cta, SCAN_AGG | agg);
            int64_t jb = cta - 1;
            for (;;) {
                const int64_t j = jb - lane;
                unsigned long long d = j >= 0 ? ld_desc(desc + j) : SCAN_INC;  // before span 0: inclusive prefix 0
                while (__any_sync(0xffffffffu, (d >> 62) == 0ull)) {
                    if ((d >> 62) == 0ull) d = ld_desc(desc + j);
                }
                const unsigned m = __ballot_sync(0xffffffffu, (d >> 62) == 2ull);
                const int first = m ? (__ffs(m) - 1) : 31;  // nearest predecessor holding an inclusive prefix
                unsigned long long c = lane <= first ? (d & SCAN_VAL) : 0ull;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
                excl += c;
                if (m) break;
                jb -= 32;
            }
            if (lane == 0) st_desc(desc + cta, SCAN_INC | (excl + agg));
        }
        if (lane == 0) s_prefix = excl;
    }
    __syncthreads();
    // ---- pass 2: scan the span tile by tile ----
    unsigned long long carry = s_prefix;
    for (int64_t t0 = s0; t0 < s1; t0 += LB_TILE) {
        const int64_t base = t0 + (int64_t)threadIdx.x * LB_ITEMS;
        T v[LB_ITEMS];
        unsigned long long ts = 0;
#pragma unroll
        for (int i = 0; i < LB_ITEMS; i++) {
            v[i] = (base + i < s1) ? in[base + i] : T(0);
            ts += (unsigned long long)v[i];
        }
        unsigned long long inc = ts;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        __syncthreads();  // s_wsum of the previous tile (or of pass 1) has been read by everyone
        if (lane == 31) s_wsum[w] = inc;
        __syncthreads();
        unsigned long long woff = 0, tot = 0;
#pragma unroll
        for (int k = 0; k < SCAN_THREADS / 32; k++) {
            const unsigned long long t = s_wsum[k];
            if (k < w) woff += t;
            tot += t;
        }
        unsigned long long run = carry + woff + (inc - ts);
#pragma unroll
        for (int i = 0; i < LB_ITEMS; i++) {
            if (INCLUSIVE) run += (unsigned long long)v[i];
            if (base + i < s1) out[base + i] = (T)run;
            if (!INCLUSIVE) run += (unsigned long long)v[i];
        }
        carry += tot;
    }
    if (s1 == n && s0 < n && threadIdx.x == 0) {  // the CTA that owns the last element: carry = grand total
        if (write_total) out[n] = (T)carry;
        if (total_out) *total_out = (long long)carry;
    }
}

// grid / span of the single-pass scan: at most one CTA per SM, whole tiles per CTA
static void scan_shape(int64_t n, int64_t& ctas, int64_t& span_tiles)
{
    const int64_t ntiles = (n + LB_TILE - 1) / LB_TILE;
    const int64_t cap = num_sms();
    span_tiles = (ntiles + cap - 1) / cap;
    if (span_tiles < 1) span_tiles = 1;
    ctas = (ntiles + span_tiles - 1) / span_tiles;
}

// small inputs (the reference's own drivers run 1K-40K Gaussians): ONE CTA walks the tiles with a running carry —
// one launch instead of three; out may alias in (a tile is read into registers before it is written)
constexpr int64_t SCAN_SMALL_MAX = 16 * SCAN_TILE;
__global__ void __launch_bounds__(SCAN_THREADS) scan_small(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int64_t n,
                                                           int write_total, long long* __restrict__ total_out)
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
    if (total_out && threadIdx.x == 0) *total_out = (long long)carry;
}

size_t scan_ws_bytes(int64_t n)
{
    int64_t nb = (n + LB_TILE - 1) / LB_TILE;
    return align_up((size_t)(nb + 2) * sizeof(unsigned long long));
}

int exclusive_scan_u32(const uint32_t* in, uint32_t* out, int64_t n, bool write_total, void* ws, cudaStream_t st,
                       long long* total_out)
{
    if (n <= 0) {
        if (write_total) PIGS_CHECK(cudaMemsetAsync(out, 0, sizeof(uint32_t), st));
        if (total_out) PIGS_CHECK(cudaMemsetAsync(total_out, 0, sizeof(long long), st));
        return 0;
    }
    if (n <= SCAN_SMALL_MAX) {
        scan_small<<<1, SCAN_THREADS, 0, st>>>(in, out, n, write_total ? 1 : 0, total_out);
        PIGS_LAUNCH_CHECK();
        return 0;
    }
    int64_t nb, span;
    scan_shape(n, nb, span);
    PIGS_CHECK(cudaMemsetAsync(ws, 0, (size_t)(nb + 1) * sizeof(unsigned long long), st));
    scan_lookback_kernel<uint32_t, false><<<(unsigned)nb, SCAN_THREADS, 0, st>>>(in, out, n, write_total ? 1 : 0,
                                                                                 (unsigned long long*)ws, total_out, span);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int exclusive_scan_u64(const unsigned long long* in, unsigned long long* out, int64_t n, bool write_total, void* ws,
                       cudaStream_t st)
{
    if (n <= 0) {
        if (write_total) PIGS_CHECK(cudaMemsetAsync(out, 0, sizeof(unsigned long long), st));
        return 0;
    }
    int64_t nb, span;
    scan_shape(n, nb, span);
    PIGS_CHECK(cudaMemsetAsync(ws, 0, (size_t)(nb + 1) * sizeof(unsigned long long), st));
    scan_lookback_kernel<unsigned long long, false><<<(unsigned)nb, SCAN_THREADS, 0, st>>>(
        in, out, n, write_total ? 1 : 0, (unsigned long long*)ws, nullptr, span);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int inclusive_scan_i64(const int64_t* in, int64_t* out, int64_t n, void* ws, cudaStream_t st, long long* total_out)
{
    if (n <= 0) {
        if (total_out) PIGS_CHECK(cudaMemsetAsync(total_out, 0, sizeof(long long), st));
        return 0;
    }
    int64_t nb, span;
    scan_shape(n, nb, span);
    PIGS_CHECK(cudaMemsetAsync(ws, 0, (size_t)(nb + 1) * sizeof(unsigned long long), st));
    scan_lookback_kernel<int64_t, true><<<(unsigned)nb, SCAN_THREADS, 0, st>>>(in, out, n, 0, (unsigned long long*)ws, total_out, span);
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
static_assert(SORT_TILE == SORT_TILE_ELEMS, "key-producing kernels histogram whole sort tiles");
static_assert(SORT_THREADS == 256, "radix_scatter phase 2: one thread per digit (RADIX below is 256)");
constexpr int RADIX = 256;

// lanes of the warp whose (valid) element has the same digit as mine — `nbits` ballots instead of MATCH.ANY, whose cost
// grows with the number of distinct values (a 4x4 tile grid has 16+): the radix sort's ranking primitive
__device__ __forceinline__ uint32_t match_digit(uint32_t d, int nbits, bool ok)
{
    uint32_t peers = __ballot_sync(0xffffffffu, ok);
    for (int b = 0; b < nbits; b++) {
        const bool bit = (d >> b) & 1u;
        const uint32_t m = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? m : ~m;
    }
    return peers;
}

__global__ void __launch_bounds__(SORT_THREADS) radix_hist(const uint32_t* __restrict__ keys, int64_t n, int shift,
                                                           uint32_t mask, uint32_t* __restrict__ ghist, int64_t nb, int nbits)
{
    __shared__ uint32_t hist[RADIX];
    for (int i = threadIdx.x; i < RADIX; i += SORT_THREADS) hist[i] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * SORT_TILE;
    // tile keys have a handful of distinct digits: one shared-memory atomic per distinct digit per warp (match_any)
    // instead of 32 colliding ones
    const int lane = threadIdx.x & 31;
#pragma unroll 4
    for (int i = 0; i < SORT_TILE / SORT_THREADS; i++) {
        int64_t idx = base + i * SORT_THREADS + threadIdx.x;
        const bool ok = idx < n;
        const uint32_t d = ok ? ((keys[idx] >> shift) & mask) : 0u;
        const uint32_t peers = match_digit(d, nbits, ok);
        if (ok && lane == (__ffs(peers) - 1)) atomicAdd(&hist[d], (uint32_t)__popc(peers));
    }
    __syncthreads();
    for (int i = threadIdx.x; i < RADIX; i += SORT_THREADS) ghist[(int64_t)i * nb + blockIdx.x] = hist[i];
}

__global__ void __launch_bounds__(SORT_THREADS) radix_scatter(const uint32_t* __restrict__ keys,
                                                              const uint32_t* __restrict__ vals, int64_t n, int shift,
                                                              uint32_t mask, const uint32_t* __restrict__ gbase,
                                                              int64_t nb, uint32_t* __restrict__ keys_out,
                                                              uint32_t* __restrict__ vals_out, uint2* __restrict__ ranges,
                                                              int max_slot, int nbits)
{
    // Ranked scatter THROUGH shared memory: the tile's pairs are first placed in digit order (stable) in shared memory,
    // then copied out so that consecutive threads write consecutive addresses of the same digit's output run.  With the
    // few distinct digits of a tile key (16 tiles) a direct scatter writes 8 useful bytes per 32-byte sector and the L2
    // store path, not HBM, sets the pace (43 us for 4M pairs).
    __shared__ uint32_t wcount[SORT_WARPS][RADIX];  // per-warp digit counts, then per-warp running LOCAL offsets
    __shared__ uint32_t dstart[RADIX + 1];          // local start of every digit's run inside the tile
    __shared__ uint32_t dglob[RADIX];               // global base of (digit, this CTA)
    __shared__ uint32_t s_k[SORT_TILE], s_v[SORT_TILE];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (ranges && blockIdx.x == 0) {
        // single-pass sort whose digit IS the tile key: the scanned histogram holds every tile's range
        // (identifyTileRanges, sampler_impl.cu:134-151: [first, last + 1) of present tiles, (0, 0) otherwise)
        for (int t = threadIdx.x; t <= max_slot; t += SORT_THREADS) {
            const uint32_t b0 = gbase[(int64_t)t * nb];
            const uint32_t b1 = t + 1 < RADIX ? gbase[(int64_t)(t + 1) * nb] : (uint32_t)n;
            ranges[t] = b1 > b0 ? make_uint2(b0, b1) : make_uint2(0u, 0u);
        }
    }
    for (int i = threadIdx.x; i < SORT_WARPS * RADIX; i += SORT_THREADS) (&wcount[0][0])[i] = 0;
    __syncthreads();

    const int64_t tbase = (int64_t)blockIdx.x * SORT_TILE;
    const int64_t wbase = tbase + (int64_t)w * SORT_WARP_ITEMS;
    const int cnt = (int)((n - tbase) < (int64_t)SORT_TILE ? (n - tbase) : (int64_t)SORT_TILE);
    uint32_t k[SORT_ROUNDS], v[SORT_ROUNDS], pm[SORT_ROUNDS];
    // phase 1: load everything first (independent loads in flight), then per-warp digit counts; the peer masks are
    // kept for the placement phase
#pragma unroll
    for (int r = 0; r < SORT_ROUNDS; r++) {
        const int64_t idx = wbase + r * 32 + lane;
        const bool ok = idx < n;
        k[r] = ok ? keys[idx] : 0u;
        v[r] = ok ? vals[idx] : 0u;
    }
#pragma unroll
    for (int r = 0; r < SORT_ROUNDS; r++) {
        const bool ok = wbase + r * 32 + lane < n;
        const uint32_t d = (k[r] >> shift) & mask;
        pm[r] = match_digit(d, nbits, ok);
        if (ok && lane == (__ffs(pm[r]) - 1)) wcount[w][d] += __popc(pm[r]);
        __syncwarp();
    }
    __syncthreads();
    // phase 2: digit totals -> local digit starts; per digit an exclusive scan over the warps
    {
        const int d = threadIdx.x;  // SORT_THREADS == RADIX
        uint32_t tot = 0;
#pragma unroll
        for (int ww = 0; ww < SORT_WARPS; ww++) tot += wcount[ww][d];
        // exclusive scan of tot over the 256 digits (warp scan + warp totals)
        uint32_t inc = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) dstart[RADIX - SORT_WARPS + w] = inc;  // scratch: warp totals parked in the tail (rewritten below)
        __syncthreads();
        uint32_t woff = 0;
#pragma unroll
        for (int ww = 0; ww < SORT_WARPS; ww++)
            if (ww < w) woff += dstart[RADIX - SORT_WARPS + ww];
        __syncthreads();
        const uint32_t start = woff + inc - tot;
        dstart[d] = start;
        if (d == RADIX - 1) dstart[RADIX] = start + tot;
        dglob[d] = gbase[(int64_t)d * nb + blockIdx.x];
        uint32_t run = start;
#pragma unroll
        for (int ww = 0; ww < SORT_WARPS; ww++) {
            const uint32_t c = wcount[ww][d];
            wcount[ww][d] = run;
            run += c;
        }
    }
    __syncthreads();
    // phase 3: in-order walk, stable placement in shared memory
#pragma unroll
    for (int r = 0; r < SORT_ROUNDS; r++) {
        const int64_t idx = wbase + r * 32 + lane;
        const bool ok = idx < n;
        const uint32_t d = (k[r] >> shift) & mask;
        const uint32_t peers = pm[r];
        if (ok) {
            const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
            const uint32_t pos = wcount[w][d] + rank;
            s_k[pos] = k[r];
            s_v[pos] = v[r];
        }
        __syncwarp();
        if (ok && lane == (__ffs(peers) - 1)) wcount[w][d] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    // phase 4: coalesced copy-out — slot i of the tile belongs to digit d = its key's digit, at offset i - dstart[d]
    for (int i = threadIdx.x; i < cnt; i += SORT_THREADS) {
        const uint32_t kk = s_k[i];
        const uint32_t d = (kk >> shift) & mask;
        const uint32_t pos = dglob[d] + ((uint32_t)i - dstart[d]);
        keys_out[pos] = kk;
        vals_out[pos] = s_v[i];
    }
}

size_t sort_ws_bytes(int64_t n)
{
    int64_t nb = (n + SORT_TILE - 1) / SORT_TILE;
    if (nb < 1) nb = 1;
    size_t hist = align_up((size_t)(RADIX * nb + 1) * sizeof(uint32_t));
    return hist + scan_ws_bytes(RADIX * nb) + 2 * align_up((size_t)(n > 0 ? n : 1) * sizeof(uint32_t));
}

uint32_t* sort_ghist(void* ws) { return (uint32_t*)ws; }
int64_t sort_tiles(int64_t n) { return (n + SORT_TILE - 1) / SORT_TILE; }

int stable_sort_pairs_u32(uint32_t* keys_in, uint32_t* vals_in, uint32_t* keys_out, uint32_t* vals_out, int64_t n,
                          int bits, void* ws, cudaStream_t st, const SortExtras* ex)
{
    if (n <= 0) return 0;
    const int64_t nb = (n + SORT_TILE - 1) / SORT_TILE;
    Carver c(ws);
    uint32_t* ghist = c.take<uint32_t>(RADIX * nb + 1);  // sort_ghist(ws)
    char* scan_ws = c.take<char>(scan_ws_bytes(RADIX * nb));
    uint32_t* tmp_k = c.take<uint32_t>(n);
    uint32_t* tmp_v = c.take<uint32_t>(n);
    int passes = (bits + 7) / 8;
    if (passes < 1) passes = 1;
    if (ex && ex->ranges_written) *ex->ranges_written = false;
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
        if (!(p == 0 && ex && ex->hist_done)) {  // the key-producing kernel may have left the first histogram behind
            radix_hist<<<(unsigned)nb, SORT_THREADS, 0, st>>>(src_k, n, shift, mask, ghist, nb, pb);
            PIGS_LAUNCH_CHECK();
            PIGS_STAGE(st, "radix histogram");
        }
        int rc = exclusive_scan_u32(ghist, ghist, RADIX * nb, false, scan_ws, st);
        if (rc) return rc;
        PIGS_STAGE(st, "radix scan");
        uint2* rg = nullptr;
        int max_slot = 0;
        if (passes == 1 && ex && ex->ranges && ex->max_slot < RADIX) {
            rg = ex->ranges;
            max_slot = ex->max_slot;
            if (ex->ranges_written) *ex->ranges_written = true;
        }
        radix_scatter<<<(unsigned)nb, SORT_THREADS, 0, st>>>(src_k, src_v, n, shift, mask, ghist, nb, dst_k, dst_v, rg, max_slot, pb);
        PIGS_LAUNCH_CHECK();
        PIGS_STAGE(st, "radix scatter (+ ranges)");
        src_k = dst_k;
        src_v = dst_v;
    }
    return 0;
}

}  // namespace pigs

extern "C" {

void pigs_set_stage_timing(int on)
{
    pigs::g_stage_timing = on != 0;
    pigs::g_nmarks = 0;
}

// Synchronises on the recorded marks; writes up to `cap` (name, milliseconds since the previous mark) pairs and returns
// how many; the first mark is the origin (0 ms).  Clears the list.
int pigs_stage_timing_read(const char** names, float* ms, int cap)
{
    using namespace pigs;
    int n = 0;
    for (int i = 0; i < g_nmarks && n < cap; i++) {
        float t = 0.f;
        if (cudaEventSynchronize(g_marks[i].ev) != cudaSuccess) break;
        if (i > 0) cudaEventElapsedTime(&t, g_marks[i - 1].ev, g_marks[i].ev);
        names[n] = g_marks[i].name;
        ms[n] = t;
        n++;
    }
    g_nmarks = 0;
    return n;
}

}  // extern "C"

// ---- allocation-free variant of pigs_neighbors_ranges ---------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).  The look-back scan's scratch — [ticket u64]
// [descriptor u64 x tiles] — lives in a file-scope device array instead of a caller workspace.  Sized like
// scan_ws_bytes() for the stated maximum of PIGS_RANGES_STATIC_MAX_P rows (one descriptor per 4096-element tile + the
// ticket): 32 KB + 256 B that every GPU which loads the library holds until the process exits.
namespace pigs {

constexpr int64_t RANGES_STATIC_MAX_P = PIGS_RANGES_STATIC_MAX_P;
constexpr int64_t RANGES_STATIC_WORDS = RANGES_STATIC_MAX_P / LB_TILE + 2;
static __device__ __align__(256) unsigned long long g_ranges_static_scratch[RANGES_STATIC_WORDS];

// In host code `g_ranges_static_scratch` is a host-side stand-in, NOT a device address: the device address is resolved
// once per device (with that device current) through cudaGetSymbolAddress and kept here.
constexpr int RANGES_STATIC_MAX_DEVICES = 64;
static unsigned long long* g_ranges_static_ptr[RANGES_STATIC_MAX_DEVICES] = {};

static int ranges_static_scratch(unsigned long long** out)
{
    int dev = -1;
    PIGS_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= RANGES_STATIC_MAX_DEVICES)
        return set_error(-1, "pigs_neighbors_ranges_static: device ordinal beyond the 64 this entry point tracks", __FILE__, __LINE__);
    unsigned long long* p = __atomic_load_n(&g_ranges_static_ptr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        void* sym = nullptr;
        PIGS_CHECK(cudaGetSymbolAddress(&sym, g_ranges_static_scratch));
        if (!sym) return set_error(-1, "pigs_neighbors_ranges_static: cudaGetSymbolAddress returned a null address", __FILE__, __LINE__);
        p = (unsigned long long*)sym;  // every thread resolves the same address for one device: a racing store is benign
        __atomic_store_n(&g_ranges_static_ptr[dev], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return 0;
}

}  // namespace pigs

extern "C" int pigs_neighbors_ranges_static(int64_t P, const int64_t* counts, int64_t* ranges, int64_t* total, void* stream)
{
    using namespace pigs;
    PIGS_ARG(P > 0 && counts && ranges && total);
    if (P > RANGES_STATIC_MAX_P) {
        static thread_local char msg[160];
        snprintf(msg, sizeof(msg), "pigs_neighbors_ranges_static: P = %lld exceeds the static scratch's maximum of %lld rows "
                 "(use pigs_neighbors_ranges with a caller workspace)", (long long)P, (long long)RANGES_STATIC_MAX_P);
        return set_error(-1, msg, __FILE__, __LINE__);
    }
    unsigned long long* ws = nullptr;
    int rc = ranges_static_scratch(&ws);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    int64_t nb, span;
    scan_shape(P, nb, span);
    if (nb + 1 > RANGES_STATIC_WORDS)
        return set_error(-1, "pigs_neighbors_ranges_static: more scan spans than the static scratch holds", __FILE__, __LINE__);
    // the array keeps the previous call's ticket and descriptors: clear what this launch reads before it runs
    PIGS_CHECK(cudaMemsetAsync(ws, 0, (size_t)(nb + 1) * sizeof(unsigned long long), st));
    scan_lookback_kernel<int64_t, true><<<(unsigned)nb, SCAN_THREADS, 0, st>>>(counts, ranges, P, 0, ws, (long long*)total, span);
    PIGS_LAUNCH_CHECK();
    return 0;
}
