// pack.cu — staging for the evaluation kernels.
// Once per preprocess (stage_classes; depends on the binning only, reused by every forward / backward of the cycle):
//   * classify every (tile, Gaussian) list entry as FAST, SHIFT or GENERAL and scan the class flags (one 64-bit scan),
//   * build the device-side work lists (forward with 8 / 4 samples per thread, backward).
// Per call:
//   * stable-partition each tile's list into [fast | shift | general] and write the Gaussian parameters as packed
//     float4 records in that order (what the evaluation kernels stream through shared memory),
//   * gather the sorted samples' coordinates and upstream gradients into float4 records (backward).
// All HBM-bound, a few bytes per list entry; no host synchronisation.
#include "eval_common.cuh"
#include "layouts.cuh"
#include "worklist.cuh"

#include <stdlib.h>

#include <mutex>

namespace pigs {

// Classification of every (tile, Gaussian) list entry (classes: eval_common.cuh).
// fl(mu - x) is monotone in x, so testing the two ends of the tile's sample bounding box is exact:
//   FAST  : conic safely positive definite and |mu - x| <= 1 in every dimension for every sample of the tile
//           (the reference's torus wrap, forward.cu:149-157, is the identity and power <= 0)
//   SHIFT : conic safely PD and per dimension EITHER |X| <= 1 for all samples, OR 1 < X < 2 for all (wrap = X - 2
//           exactly, fmod(X,2) == X), OR -2 < X < -1 for all (wrap = X + 2)
//   GENERAL otherwise.
template <int D>
__global__ void __launch_bounds__(256) flag_kernel(int64_t R, const uint32_t* __restrict__ list,
                                                   const uint32_t* __restrict__ tiles, const float* __restrict__ means,
                                                   const float* __restrict__ conics, const float* __restrict__ bbox,
                                                   int all_general, unsigned long long* __restrict__ cls,
                                                   uint32_t* __restrict__ scode)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= R) return;
    const uint32_t id = list[e], t = tiles[e];
    bool pd;
    if constexpr (D == 1) {
        pd = conics[id] > 0.0f;
    } else {
        const float c0 = conics[3 * (size_t)id], c1 = conics[3 * (size_t)id + 1], c2 = conics[3 * (size_t)id + 2];
        const float cc = c0 * c2;
        pd = (c0 > 0.0f) && (c2 > 0.0f) && (cc - c1 * c1 >= 1e-4f * cc) && (cc < 3.0e38f);
    }
    bool ok = pd && !all_general;
    uint32_t code = 0;
#pragma unroll
    for (int d = 0; d < D; d++) {
        const float m = means[(size_t)id * D + d];
        const float xlo = bbox[(size_t)t * 2 * D + d], xhi = bbox[(size_t)t * 2 * D + D + d];
        const float Xa = m - xhi, Xb = m - xlo;  // Xa <= X <= Xb for every sample of the tile
        uint32_t k = 3;
        if (Xa >= -1.0f && Xb <= 1.0f)
            k = 0;
        else if (Xa > 1.0f && Xb < 2.0f)
            k = 1;
        else if (Xa > -2.0f && Xb < -1.0f)
            k = 2;
        if (k == 3) ok = false;
        code |= (k & 3u) << (2 * d);
    }
    const bool general = !ok;
    cls[e] = general ? CLS_GENERAL : (code != 0u ? CLS_SHIFT : 0ull);
    scode[e] = general ? 0u : code;
}

// Small problems / single-class mode: every instance GENERAL, so the EXCLUSIVE SCAN of the flags is known in closed form
// (cls[e] = e * CLS_GENERAL, total at [R]) and is written directly: no flag pass, no scan launch.
__global__ void __launch_bounds__(256) flag_all_general_kernel(int64_t R, unsigned long long* __restrict__ cls,
                                                               uint32_t* __restrict__ scode)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e > R) return;
    cls[e] = (unsigned long long)e * CLS_GENERAL;
    if (e < R) scode[e] = 0u;
}

// cls = exclusive scan of the packed class flags (with the total at [R]).  Writes the records in partitioned order
// [fast | shift | general] within each tile (stable within each class); block 0 also turns the work-list template into
// this call's WorkHeader (dynamic counter at zero).
template <int D, int C, bool FWD>
__global__ void __launch_bounds__(256) pack_gauss_kernel(int64_t R, const uint32_t* __restrict__ list,
                                                         const uint32_t* __restrict__ tiles,
                                                         const unsigned long long* __restrict__ cls,
                                                         const uint32_t* __restrict__ scode,
                                                         const uint2* __restrict__ ranges, int blocks,
                                                         const float* __restrict__ means, const float* __restrict__ values,
                                                         const float* __restrict__ conics, int Ctot, int ch0,
                                                         float4* __restrict__ recs, uint32_t* __restrict__ rec_id,
                                                         const WorkTemplate* __restrict__ wt, WorkHeader* __restrict__ wh)
{
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        for (int k = 0; k < 4; k++) {
            wh->counter[k] = 0;
            wh->items[k] = wt->items[k];
        }
        wh->pairs = wt->pairs;
        wh->neg_zero = 0x80000000u;
    }
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= R) return;
    const uint32_t id = list[e], t = tiles[e];
    if ((int)t >= blocks) return;  // alias slots are never evaluated (the reference launches `blocks` CTAs)
    const uint2 r = ranges[t];
    const unsigned long long ce = cls[e], c0 = cls[r.x], c1 = cls[r.y], cn = cls[e + 1];
    const uint32_t gen_before = (uint32_t)((ce & CLS_MASK) - (c0 & CLS_MASK));
    const uint32_t gen_total = (uint32_t)((c1 & CLS_MASK) - (c0 & CLS_MASK));
    const uint32_t sh_before = (uint32_t)((ce >> 31) - (c0 >> 31));
    const uint32_t sh_total = (uint32_t)((c1 >> 31) - (c0 >> 31));
    const uint32_t nf = (r.y - r.x) - gen_total - sh_total;
    const bool general = (cn & CLS_MASK) != (ce & CLS_MASK);
    const bool shift = (cn >> 31) != (ce >> 31);
    const uint32_t before = (uint32_t)(e - r.x);
    const uint32_t pos = general ? r.x + nf + sh_total + gen_before
                                 : (shift ? r.x + nf + sh_before : r.x + before - gen_before - sh_before);
    constexpr int NV = FWD ? frec_v4(D, C) : grec_v4(D, C);
    constexpr int NC = nconic(D);
    float f[NV * 4];
#pragma unroll
    for (int i = 0; i < NV * 4; i++) f[i] = 0.0f;
#pragma unroll
    for (int d = 0; d < D; d++) f[d] = means[(size_t)id * D + d];
    if constexpr (FWD) {
        // forward layout (eval_common.cuh frec_floats): K = KH*conic, M = KH*K
#pragma unroll
        for (int k = 0; k < NC; k++) {
            const float kk = __fmul_rn(KH, conics[(size_t)id * NC + k]);
            f[D + k] = kk;
            f[D + NC + k] = __fmul_rn(KH, kk);
        }
#pragma unroll
        for (int ch = 0; ch < C; ch++) f[D + 2 * NC + ch] = values[(size_t)id * Ctot + ch0 + ch];
        f[D + 2 * NC + C] = __int_as_float((int)scode[e]);
    } else {
#pragma unroll
        for (int k = 0; k < NC; k++) f[D + k] = conics[(size_t)id * NC + k];
#pragma unroll
        for (int ch = 0; ch < C; ch++) f[D + NC + ch] = values[(size_t)id * Ctot + ch0 + ch];
        f[D + NC + C] = __int_as_float((int)scode[e]);
    }
#pragma unroll
    for (int i = 0; i < NV; i++) recs[(size_t)pos * NV + i] = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
    rec_id[pos] = id;
}

// sample records for the backward pass: [x(D)] then per channel the upstream weights of the unique
// components, symmetric duplicates summed (backward.cu:244-247,269: dL_dGxy + dL_dGyx, etc.)
template <int D, int C, unsigned MASK>
__global__ void __launch_bounds__(256) pack_samples_kernel(int64_t N, const uint32_t* __restrict__ slist,
                                                           const float* __restrict__ samples,
                                                           const float* __restrict__ gV, const float* __restrict__ gD,
                                                           const float* __restrict__ gH, const float* __restrict__ gT,
                                                           int Ctot, int ch0, int lap, float4* __restrict__ srecs)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    const size_t sid = slist[i];
    constexpr int SV = srec_v4(D, C, MASK);
    constexpr int NA = nacc(D, MASK);
    float f[SV * 4];
#pragma unroll
    for (int k = 0; k < SV * 4; k++) f[k] = 0.0f;
#pragma unroll
    for (int d = 0; d < D; d++) f[d] = samples[sid * D + d];
#pragma unroll
    for (int ch = 0; ch < C; ch++) {
        int k = D + ch * NA;
        const int c = ch0 + ch;
        if constexpr ((MASK & PIGS_FN_VALUE) != 0) {
            f[k++] = gV ? gV[sid * Ctot + c] : 0.0f;
        }
        if constexpr ((MASK & PIGS_FN_DERIV) != 0) {
#pragma unroll
            for (int d = 0; d < D; d++) f[k++] = gD ? gD[(sid * D + d) * Ctot + c] : 0.0f;
        }
        if constexpr ((MASK & PIGS_FN_HESS) != 0) {
            if constexpr (D == 1) {
                f[k++] = gH ? gH[sid * Ctot + c] : 0.0f;
            } else {
                if (lap) {  // d lap / d H = identity: the Laplacian's upstream gradient on the diagonal
                    const float gl = gH ? gH[sid * Ctot + c] : 0.0f;
                    f[k++] = gl;
                    f[k++] = 0.0f;
                    f[k++] = gl;
                } else {
                    const float* g = gH + sid * 4 * Ctot + c;
                    f[k++] = gH ? g[0] : 0.0f;
                    f[k++] = gH ? g[Ctot] + g[2 * Ctot] : 0.0f;
                    f[k++] = gH ? g[3 * Ctot] : 0.0f;
                }
            }
        }
        if constexpr ((MASK & PIGS_FN_LAPL) != 0) {
            f[k++] = gH ? gH[sid * Ctot + c] : 0.0f;  // gH is the Laplacian's [N, C] upstream gradient here
        }
        if constexpr ((MASK & PIGS_FN_THIRD) != 0) {
            if constexpr (D == 1) {
                f[k++] = gT ? gT[sid * Ctot + c] : 0.0f;
            } else {
                const float* g = gT + sid * 8 * Ctot + c;
                f[k++] = gT ? g[0] : 0.0f;
                f[k++] = gT ? g[Ctot] + g[2 * Ctot] + g[4 * Ctot] : 0.0f;
                f[k++] = gT ? g[3 * Ctot] + g[5 * Ctot] + g[6 * Ctot] : 0.0f;
                f[k++] = gT ? g[7 * Ctot] : 0.0f;
            }
        }
    }
#pragma unroll
    for (int k = 0; k < SV; k++) srecs[(size_t)i * SV + k] = make_float4(f[4 * k], f[4 * k + 1], f[4 * k + 2], f[4 * k + 3]);
}

__global__ void __launch_bounds__(256) worklist_kernel(int blocks, const uint2* __restrict__ ranges,
                                                       const uint2* __restrict__ sample_ranges,
                                                       const unsigned long long* __restrict__ cls, WorkListCfg cfg,
                                                       TileSeg* __restrict__ segs_all /* [WL_COUNT][EV_NSEG][blocks] */,
                                                       WorkTemplate* __restrict__ wt_all)
{
    worklist_body(blockIdx.x, blocks, ranges, sample_ranges, cls, cfg, segs_all, wt_all);
}

// Work items per SM the lists aim for.  The tail of a persistent launch is about one item long, and the backward's
// items (1024 Gaussians x a slice of the tile's samples) are the heavier ones: measured on B200 at C3 the backward goes
// 696.5 -> 649 ms from 48 to 768 items per SM, the forward is flat to slightly worse (profiles/r01n_split_target.log).
static int split_target_per_sm(bool backward)
{
    static int v[2] = {0, 0};
    if (v[backward] == 0) {
        const char* e = getenv(backward ? "PIGS_SPLIT_TARGET_BWD" : "PIGS_SPLIT_TARGET_FWD");  // tuning hook
        v[backward] = e ? atoi(e) : (backward ? 768 : 48);
        if (v[backward] < 1) v[backward] = 48;
    }
    return v[backward];
}

static uint32_t split_hint(int64_t stationary, uint32_t chunk, int blocks, bool backward)
{
    const int64_t target = (int64_t)num_sms() * split_target_per_sm(backward);
    const int64_t base = stationary / chunk + blocks;
    int64_t s = (target + base - 1) / base;
    if (s < 1) s = 1;
    if (s > 4096) s = 4096;
    return (uint32_t)s;
}

int g_deterministic = 0;
int g_class_concurrency = 1;

namespace {
struct ForkState {
    bool ready = false;
    cudaStream_t side[2];
    cudaEvent_t forked, joined[2];
};
ForkState g_fork[64];  // per device
}  // namespace

int64_t g_small_pairs = -1;  // -1: not read yet (env PIGS_SMALL_PAIRS, default 20 M pairs)

int64_t small_pairs_threshold()
{
    if (g_small_pairs < 0) {
        const char* e = getenv("PIGS_SMALL_PAIRS");
        g_small_pairs = e ? atoll(e) : 20000000ll;
        if (g_small_pairs < 0) g_small_pairs = 0;
    }
    return g_small_pairs;
}

bool small_problem(int64_t R, int64_t N, int blocks)
{
    const int64_t thr = small_pairs_threshold();
    if (thr <= 0 || R <= 0 || blocks <= 0) return false;
    // estimated pairs = instances x average samples per tile (exact when the tiles are equally populated)
    const double est = (double)R * ((double)N / (double)blocks);
    return est <= (double)thr;
}

namespace {
struct SmallNote {
    const void* ptr;
    int64_t R, N;
    int blocks;
    bool single;
};
constexpr int SMALL_NOTES = 256;
SmallNote g_notes[SMALL_NOTES];
int g_note_next = 0;
std::mutex g_note_mu;
}  // namespace

void small_binning_note(const void* ptr, int64_t R, int64_t N, int blocks, bool single)
{
    std::lock_guard<std::mutex> lk(g_note_mu);
    for (int i = 0; i < SMALL_NOTES; i++) {
        if (g_notes[i].ptr == ptr) {  // the buffer was refilled: its old note is void
            g_notes[i] = SmallNote{ptr, R, N, blocks, single};
            return;
        }
    }
    if (!single) return;  // unknown buffers already mean "three classes"
    g_notes[g_note_next] = SmallNote{ptr, R, N, blocks, single};
    g_note_next = (g_note_next + 1) % SMALL_NOTES;
}

bool small_binning_lookup(const void* ptr, int64_t R, int64_t N, int blocks)
{
    std::lock_guard<std::mutex> lk(g_note_mu);
    for (int i = 0; i < SMALL_NOTES; i++) {
        const SmallNote& n = g_notes[i];
        if (n.ptr == ptr) return n.single && n.R == R && n.N == N && n.blocks == blocks;
    }
    return false;
}

int ClassFork::fork(cudaStream_t st)
{
    main = st;
    forked = false;
    if (!g_class_concurrency || g_deterministic || g_debug_sync) return 0;
    int dev = 0;
    PIGS_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return 0;
    ForkState& f = g_fork[dev];
    if (!f.ready) {
        for (int k = 0; k < 2; k++) {
            PIGS_CHECK(cudaStreamCreateWithFlags(&f.side[k], cudaStreamNonBlocking));
            PIGS_CHECK(cudaEventCreateWithFlags(&f.joined[k], cudaEventDisableTiming));
        }
        PIGS_CHECK(cudaEventCreateWithFlags(&f.forked, cudaEventDisableTiming));
        f.ready = true;
    }
    PIGS_CHECK(cudaEventRecord(f.forked, st));
    for (int k = 0; k < 2; k++) PIGS_CHECK(cudaStreamWaitEvent(f.side[k], f.forked, 0));
    forked = true;
    return 0;
}

cudaStream_t ClassFork::stream(int k) const
{
    if (!forked || k == 0) return main;
    int dev = 0;
    cudaGetDevice(&dev);
    return g_fork[dev].side[k - 1];
}

int ClassFork::join()
{
    if (!forked) return 0;
    int dev = 0;
    PIGS_CHECK(cudaGetDevice(&dev));
    ForkState& f = g_fork[dev];
    for (int k = 0; k < 2; k++) {
        PIGS_CHECK(cudaEventRecord(f.joined[k], f.side[k]));
        PIGS_CHECK(cudaStreamWaitEvent(main, f.joined[k], 0));
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------
// host-side staging entry points
// ---------------------------------------------------------------------------------------------
static uint32_t small_min_stream(bool backward)
{
    static int v[2] = {0, 0};
    if (v[backward] == 0) {
        const char* e = getenv(backward ? "PIGS_SMALL_MIN_STREAM_BWD" : "PIGS_SMALL_MIN_STREAM_FWD");  // tuning hook
        v[backward] = e ? atoi(e) : (backward ? 32 : 16);
        if (v[backward] < 1) v[backward] = 32;
    }
    return (uint32_t)v[backward];
}

WorkListCfg make_worklist_cfg(int64_t R, int64_t N, int blocks, bool single_class)
{
    WorkListCfg cfg;
    const uint32_t chunks[WL_COUNT] = {EV_THREADS * (PIGS_FWD_KS16 ? 16u : 8u), EV_THREADS * 4u, EV_THREADS * (uint32_t)bwd_kg(2, 1, 7u)};
    for (int wl = 0; wl < WL_COUNT; wl++) {
        const bool bwd = wl == WL_BWD;
        cfg.chunk[wl] = chunks[wl];
        cfg.split_hint[wl] = (g_deterministic && !bwd) ? 1u : split_hint(bwd ? R : N, chunks[wl], blocks, bwd);
        cfg.min_stream[wl] = bwd ? 128u : 256u;
        // small problems are bound by the LONGEST work item, not by throughput (a few dozen items on 296 persistent CTAs):
        // cut the streamed operand finer, so that every item is short and all of them run at once
        if (single_class) cfg.min_stream[wl] = small_min_stream(bwd);
        cfg.stream_is_gauss[wl] = bwd ? 0 : 1;
    }
    return cfg;
}

int stage_classes(int D, int blocks, int64_t R, int64_t N, const float* means, const float* conics, const float* bbox,
                  const uint2* ranges, const uint2* sample_ranges, const BinLayout& b, cudaStream_t st, bool single_class)
{
    const unsigned nb = (unsigned)((R + 255) / 256);
    // deterministic mode: every instance GENERAL (exact wrap + skip per pair), so a tile's list stays in ascending
    // Gaussian id and — with the forward's Gaussians unsplit — every output element is written by exactly one item
    if (single_class && R < (int64_t)CLS_MASK) {
        flag_all_general_kernel<<<(unsigned)((R + 256) / 256), 256, 0, st>>>(R, b.cls, b.scode);
        PIGS_LAUNCH_CHECK();
        PIGS_STAGE(st, "class flags (all general, scanned)");
    } else {
        if (D == 2)
            flag_kernel<2><<<nb, 256, 0, st>>>(R, b.list, b.tiles, means, conics, bbox, g_deterministic || single_class, b.cls, b.scode);
        else
            flag_kernel<1><<<nb, 256, 0, st>>>(R, b.list, b.tiles, means, conics, bbox, g_deterministic || single_class, b.cls, b.scode);
        PIGS_LAUNCH_CHECK();
        PIGS_STAGE(st, "class flags");
        int rc = exclusive_scan_u64(b.cls, b.cls, R, true, b.scan_ws, st);
        if (rc) return rc;
        PIGS_STAGE(st, "class scan (u64)");
    }
    const WorkListCfg cfg = make_worklist_cfg(R, N, blocks, single_class);
    worklist_kernel<<<WL_COUNT, 256, 0, st>>>(blocks, ranges, sample_ranges, b.cls, cfg, b.segs, b.wt);
    PIGS_LAUNCH_CHECK();
    PIGS_STAGE(st, "work lists");
    return 0;
}

int stage_gaussians(const EvalCall& c, int C, int ch0, bool fwd_layout, int wl, const EvalWs& ws, cudaStream_t st)
{
    const int64_t R = c.R;
    const unsigned nb = (unsigned)((R + 255) / 256);
#define PACK1(D_, C_, F_)                                                                                            \
    pack_gauss_kernel<D_, C_, F_><<<nb, 256, 0, st>>>(R, c.list, c.tiles, c.cls, c.scode, c.ranges, c.blocks, c.means, \
                                                      c.values, c.conics, c.C, ch0, ws.recs, ws.rec_id, c.wt + wl, ws.wh)
#define PACK(D_, C_)       \
    do {                   \
        if (fwd_layout)    \
            PACK1(D_, C_, true);  \
        else               \
            PACK1(D_, C_, false); \
    } while (0)
    if (c.D == 2 && C == 1)
        PACK(2, 1);
    else if (c.D == 2 && C == 2)
        PACK(2, 2);
    else if (c.D == 1 && C == 1)
        PACK(1, 1);
    else
        PACK(1, 2);
#undef PACK
#undef PACK1
    PIGS_LAUNCH_CHECK();
    return 0;
}

template <int D, int C, unsigned MASK>
static int stage_samples_t(const EvalCall& c, int ch0, const float* gV, const float* gD, const float* gH, const float* gT,
                           bool lap, const EvalWs& ws, cudaStream_t st)
{
    pack_samples_kernel<D, C, MASK><<<(unsigned)((c.N + 255) / 256), 256, 0, st>>>(c.N, c.slist, c.samples, gV, gD, gH, gT,
                                                                                    c.C, ch0, lap ? 1 : 0, ws.srecs);
    PIGS_LAUNCH_CHECK();
    return 0;
}

int stage_samples(const EvalCall& c, int C, unsigned mask, int ch0, const float* gV, const float* gD, const float* gH,
                  const float* gT, bool lap, const EvalWs& ws, cudaStream_t st)
{
#define SS(D_, C_, M_) \
    if (c.D == D_ && C == C_ && mask == M_) return stage_samples_t<D_, C_, M_>(c, ch0, gV, gD, gH, gT, lap, ws, st);
#define SS_ALL(D_, C_) SS(D_, C_, 1u) SS(D_, C_, 2u) SS(D_, C_, 4u) SS(D_, C_, 7u) SS(D_, C_, 8u)
    SS_ALL(2, 1) SS_ALL(2, 2) SS_ALL(1, 1) SS_ALL(1, 2)
    SS(2, 1, 16u) SS(2, 1, 19u)
#undef SS_ALL
#undef SS
    return set_error(-1, "unsupported (D, C, mask)", __FILE__, __LINE__);
}

}  // namespace pigs
