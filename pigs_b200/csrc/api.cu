// api.cu — extern "C" entry points of the evaluation path (see include/pigs_b200.h) and the roofline
// microbenchmarks.  Host logic only: resolve the opaque buffers, split the request into natively
// compiled (mask, channel-group) kernels, stage, launch.  No host synchronisation anywhere.
#include "layouts.cuh"

#include <vector>

using namespace pigs;

namespace {

// ---- optional device timing of the evaluation kernels (bench.py roofline) ------------------------
struct TimedLaunch {
    cudaEvent_t a, b;
    int which;  // 0 forward, 1 backward
};
bool g_timing = false;
std::vector<TimedLaunch> g_timed;
std::vector<cudaEvent_t> g_event_pool;

cudaEvent_t take_event()
{
    cudaEvent_t e = nullptr;
    if (!g_event_pool.empty()) {
        e = g_event_pool.back();
        g_event_pool.pop_back();
    } else {
        cudaEventCreate(&e);
    }
    return e;
}

struct ScopedKernelTimer {
    bool on;
    TimedLaunch t;
    cudaStream_t st;
    ScopedKernelTimer(int which, cudaStream_t s) : on(g_timing), st(s)
    {
        if (!on) return;
        t.a = take_event();
        t.b = take_event();
        t.which = which;
        cudaEventRecord(t.a, st);
    }
    ~ScopedKernelTimer()
    {
        if (!on) return;
        cudaEventRecord(t.b, st);
        g_timed.push_back(t);
    }
};

// function masks with a dedicated kernel; anything else inside {V,D,H} runs the fused 7-kernel with
// the unwanted outputs/upstreams left NULL.
unsigned native_low_mask(unsigned m)
{
    if (m & PIGS_FN_LAPL) return (m & 3u) ? 19u : 16u;  // trace-only kernels (2-D, one channel): Laplacian (+ value + gradient)
    m &= 7u;
    if (m == 0u || m == 1u || m == 2u || m == 4u) return m;
    return 7u;
}

int max_sv4(int D, int C)
{
    int m = 0;
    const unsigned masks[7] = {1u, 2u, 4u, 7u, 8u, 16u, 19u};
    for (unsigned k : masks) {
        int v = srec_v4(D, C, k);
        if (v > m) m = v;
    }
    return m;
}

int resolve(EvalCall& c, int64_t P, int D, int64_t N, int C, int blocks, int64_t R, const float* means, const float* values,
            const float* conics, const float* samples, const void* binning_buffer, const void* sample_binning_buffer,
            const void* ranges, const void* sample_ranges)
{
    PIGS_ARG(D == 1 || D == 2);
    PIGS_ARG(C >= 1);
    PIGS_ARG(P > 0 && N > 0 && blocks > 0 && R >= 0);
    PIGS_ARG(means && values && conics && samples && binning_buffer && sample_binning_buffer && ranges && sample_ranges);
    c.D = D;
    c.C = C;
    c.blocks = blocks;
    c.P = P;
    c.N = N;
    c.R = R;
    c.means = means;
    c.values = values;
    c.conics = conics;
    c.samples = samples;
    BinLayout b(const_cast<void*>(binning_buffer), R, blocks);
    SampleBinLayout s(const_cast<void*>(sample_binning_buffer), N, blocks, D);
    c.list = b.list;
    c.tiles = b.tiles;
    c.cls = b.cls;
    c.scode = b.scode;
    c.segs = b.segs;
    c.wt = b.wt;
    c.slist = s.list;
    c.bbox = s.bbox;
    c.ranges = (const uint2*)ranges;
    c.sample_ranges = (const uint2*)sample_ranges;
    c.single_class = small_binning_lookup(binning_buffer, R, N, blocks);
    return 0;
}

}  // namespace

extern "C" {

int pigs_version(void) { return 100; }
const char* pigs_last_error(void) { return g_err; }
void pigs_set_grid_recip_mode(int recip) { g_recip_mode = recip ? 1 : 0; }
int64_t pigs_launch_count(void) { return g_launches; }
void pigs_set_debug(int on) { g_debug_sync = on ? 1 : 0; }
void pigs_set_deterministic(int on) { g_deterministic = on ? 1 : 0; }
int pigs_get_deterministic(void) { return g_deterministic; }
void pigs_set_class_concurrency(int on) { g_class_concurrency = on ? 1 : 0; }
void pigs_set_small_problem_pairs(int64_t pairs) { g_small_pairs = pairs < 0 ? 0 : pairs; }
int pigs_get_debug(void) { return g_debug_sync; }

void pigs_set_kernel_timing(int on)
{
    g_timing = on != 0;
}

int pigs_kernel_timing_read(int which, float* ms_sum, int* launches)
{
    PIGS_ARG(ms_sum && launches);
    float sum = 0.f;
    int n = 0;
    std::vector<TimedLaunch> keep;
    for (const TimedLaunch& t : g_timed) {
        if (t.which != which) {
            keep.push_back(t);
            continue;
        }
        PIGS_CHECK(cudaEventSynchronize(t.b));
        float ms = 0.f;
        PIGS_CHECK(cudaEventElapsedTime(&ms, t.a, t.b));
        sum += ms;
        n++;
        g_event_pool.push_back(t.a);
        g_event_pool.push_back(t.b);
    }
    g_timed.swap(keep);
    *ms_sum = sum;
    *launches = n;
    return 0;
}

size_t pigs_eval_workspace_bytes(int64_t num_rendered, int64_t N, int blocks, int D, int C, unsigned fn_mask)
{
    (void)fn_mask;
    const int Cg = C >= 2 ? 2 : 1;
    return EvalWs(nullptr, num_rendered, N, blocks, frec_v4(D, Cg), max_sv4(D, Cg)).bytes;
}

int pigs_sample_forward(unsigned fn_mask, int64_t P, int D, int64_t N, int C, int blocks, int64_t num_rendered,
                        const float* means, const float* values, const float* conics, const float* samples,
                        const void* binning_buffer, const void* sample_binning_buffer, const void* ranges,
                        const void* sample_ranges, float* out_u, float* out_du, float* out_ddu, float* out_dddu,
                        void* workspace, size_t workspace_bytes, void* stream)
{
    PIGS_ARG(fn_mask != 0 && (fn_mask & ~31u) == 0);
    PIGS_ARG(!((fn_mask & PIGS_FN_HESS) && (fn_mask & PIGS_FN_LAPL)));  // they share out_ddu
    const int lap = (fn_mask & PIGS_FN_LAPL) ? 1 : 0;
    if (lap) fn_mask = (fn_mask & ~PIGS_FN_LAPL) | PIGS_FN_HESS;  // same kernels; the epilogue writes the trace
    EvalCall c;
    int rc = resolve(c, P, D, N, C, blocks, num_rendered, means, values, conics, samples, binning_buffer,
                     sample_binning_buffer, ranges, sample_ranges);
    if (rc) return rc;
    PIGS_ARG(workspace && workspace_bytes >= pigs_eval_workspace_bytes(num_rendered, N, blocks, D, C, fn_mask));
    PIGS_ARG(!(fn_mask & PIGS_FN_VALUE) || out_u);
    PIGS_ARG(!(fn_mask & PIGS_FN_DERIV) || out_du);
    PIGS_ARG(!(fn_mask & PIGS_FN_HESS) || out_ddu);
    PIGS_ARG(!(fn_mask & PIGS_FN_THIRD) || out_dddu);
    if (num_rendered == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const int Cg = C >= 2 ? 2 : 1;
    EvalWs ws(workspace, num_rendered, N, blocks, frec_v4(D, Cg), max_sv4(D, Cg));

    for (int ch0 = 0; ch0 < C; ch0 += 2) {
        const int cw = (C - ch0) >= 2 ? 2 : 1;
        // Laplacian: 2-D one-channel groups run the trace-only kernels (mask 16 / 19); other shapes run the Hessian
        // kernels and form the trace in the epilogue
        const bool lap_native = lap && D == 2 && cw == 1;
        unsigned todo[2] = {native_low_mask(lap_native ? ((fn_mask & 3u) | PIGS_FN_LAPL) : fn_mask), fn_mask & PIGS_FN_THIRD};
        for (unsigned m : todo) {
            if (!m) continue;
            const int wl = fwd_ks(D, cw, m) >= 8 ? WL_FWD8 : WL_FWD4;
            rc = stage_gaussians(c, cw, ch0, true, wl, ws, st);  // packed records + this launch's work header
            if (rc) return rc;
            FwdOut o;
            o.u = (fn_mask & PIGS_FN_VALUE) ? out_u : nullptr;
            o.du = (fn_mask & PIGS_FN_DERIV) ? out_du : nullptr;
            o.ddu = (fn_mask & PIGS_FN_HESS) ? out_ddu : nullptr;
            o.dddu = (fn_mask & PIGS_FN_THIRD) ? out_dddu : nullptr;
            o.lap = (lap && !lap_native) ? 1 : 0;
            {
                ScopedKernelTimer tm(0, st);
                rc = launch_forward(c, cw, m, wl, ch0, ws, o, st);
                if (rc) return rc;
            }
        }
    }
    return 0;
}

int pigs_sample_backward(unsigned fn_mask, int64_t P, int D, int64_t N, int C, int blocks, int64_t num_rendered,
                         const float* means, const float* values, const float* conics, const float* samples,
                         const void* binning_buffer, const void* sample_binning_buffer, const void* ranges,
                         const void* sample_ranges, const float* dL_du, const float* dL_ddu, const float* dL_dddu,
                         const float* dL_ddddu, float* dL_dmeans, float* dL_dvalues, float* dL_dconics, void* workspace,
                         size_t workspace_bytes, void* stream)
{
    PIGS_ARG(fn_mask != 0 && (fn_mask & ~31u) == 0);
    PIGS_ARG(!((fn_mask & PIGS_FN_HESS) && (fn_mask & PIGS_FN_LAPL)));  // they share dL_dddu
    bool lap = (fn_mask & PIGS_FN_LAPL) != 0;
    if (lap) fn_mask = (fn_mask & ~PIGS_FN_LAPL) | PIGS_FN_HESS;  // d lap / d H = I (trace-only kernels, or applied while packing)
    EvalCall c;
    int rc = resolve(c, P, D, N, C, blocks, num_rendered, means, values, conics, samples, binning_buffer,
                     sample_binning_buffer, ranges, sample_ranges);
    if (rc) return rc;
    PIGS_ARG(workspace && workspace_bytes >= pigs_eval_workspace_bytes(num_rendered, N, blocks, D, C, fn_mask));
    PIGS_ARG(dL_dmeans && dL_dvalues && dL_dconics);
    // a NULL upstream gradient is an all-zero one
    if (!dL_du) fn_mask &= ~PIGS_FN_VALUE;
    if (!dL_ddu) fn_mask &= ~PIGS_FN_DERIV;
    if (!dL_dddu) fn_mask &= ~PIGS_FN_HESS;
    if (!dL_ddddu) fn_mask &= ~PIGS_FN_THIRD;
    lap = lap && (fn_mask & PIGS_FN_HESS);
    if (num_rendered == 0 || fn_mask == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const int Cg = C >= 2 ? 2 : 1;
    EvalWs ws(workspace, num_rendered, N, blocks, frec_v4(D, Cg), max_sv4(D, Cg));

    for (int ch0 = 0; ch0 < C; ch0 += 2) {
        const int cw = (C - ch0) >= 2 ? 2 : 1;
        const bool lap_native = lap && D == 2 && cw == 1;
        unsigned todo[2] = {native_low_mask(lap_native ? ((fn_mask & 3u) | PIGS_FN_LAPL) : fn_mask), fn_mask & PIGS_FN_THIRD};
        for (unsigned m : todo) {
            if (!m) continue;
            rc = stage_gaussians(c, cw, ch0, false, WL_BWD, ws, st);
            if (rc) return rc;
            rc = stage_samples(c, cw, m, ch0, (fn_mask & PIGS_FN_VALUE) ? dL_du : nullptr,
                               (fn_mask & PIGS_FN_DERIV) ? dL_ddu : nullptr, (fn_mask & PIGS_FN_HESS) ? dL_dddu : nullptr,
                               (fn_mask & PIGS_FN_THIRD) ? dL_ddddu : nullptr, lap && !lap_native, ws, st);
            if (rc) return rc;
            BwdOut o{dL_dmeans, dL_dvalues, dL_dconics};
            {
                ScopedKernelTimer tm(1, st);
                rc = launch_backward(c, cw, m, ch0, ws, o, st);
                if (rc) return rc;
            }
        }
    }
    return 0;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// roofline microbenchmarks: register-resident FFMA and MUFU.EX2 loops (bench.py measures the peaks
// the evaluation kernels are compared against)
// ------------------------------------------------------------------------------------------------
namespace pigs {

constexpr int MB_ILP = 16;

__global__ void __launch_bounds__(256) mb_ffma_kernel(int iters, float* sink)
{
    float a[MB_ILP];
    const float b = 1.0000001f + 1e-9f * threadIdx.x, c = 1e-7f;
#pragma unroll
    for (int i = 0; i < MB_ILP; i++) a[i] = 1.0f + i * 1e-3f + threadIdx.x * 1e-6f;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int i = 0; i < MB_ILP; i++) a[i] = fmaf(a[i], b, c);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < MB_ILP; i++) s += a[i];
    if (s == 123.456f) sink[0] = s;
}

__global__ void __launch_bounds__(256) mb_ex2_kernel(int iters, float* sink)
{
    float a[MB_ILP];
#pragma unroll
    for (int i = 0; i < MB_ILP; i++) a[i] = -0.5f - i * 1e-3f - threadIdx.x * 1e-6f;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int i = 0; i < MB_ILP; i++) a[i] = ex2_approx(a[i]) - 1.5f;  // 1 MUFU + 1 FADD
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < MB_ILP; i++) s += a[i];
    if (s == 123.456f) sink[0] = s;
}

// packed FP32x2 (Blackwell FFMA2): 2 FMAs per lane per instruction
__device__ __forceinline__ unsigned long long mb_pack(float a, float b)
{
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ unsigned long long mb_fma2(unsigned long long a, unsigned long long b, unsigned long long c)
{
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

// MIX = 0: FFMA2 only.  MIX = 1: per 16 FFMA2 (32 lane-FMAs) two MUFU.EX2 — the ratio of the fused forward.
template <int MIX>
__global__ void __launch_bounds__(256) mb_ffma2_kernel(int iters, float* sink)
{
    unsigned long long a[MB_ILP];
    float m0 = -0.5f - threadIdx.x * 1e-6f, m1 = -0.25f - threadIdx.x * 1e-6f;
#pragma unroll
    for (int i = 0; i < MB_ILP; i++) a[i] = mb_pack(1.0f + i * 1e-3f + threadIdx.x * 1e-6f, 0.5f + i * 1e-3f);
    // distinct (non-uniform) multiplicands so operands come from the vector register file like in the real kernels
    const unsigned long long b = mb_pack(1.0000001f + 1e-9f * threadIdx.x, 0.9999999f), c = mb_pack(1e-7f * threadIdx.x, 2e-7f);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int i = 0; i < MB_ILP; i++) a[i] = mb_fma2(a[i], b, c);
            if (MIX) {
                m0 = ex2_approx(m0) - 1.5f;
                m1 = ex2_approx(m1) - 1.5f;
            }
        }
    }
    float s = m0 + m1;
#pragma unroll
    for (int i = 0; i < MB_ILP; i++) {
        float lo, hi;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a[i]));
        s += lo + hi;
    }
    if (s == 123.456f) sink[0] = s;
}

}  // namespace pigs

extern "C" {

int64_t pigs_microbench_ffma2(int ctas, int iters, int mix, float* sink, void* stream)
{
    if (mix)
        mb_ffma2_kernel<1><<<ctas, 256, 0, (cudaStream_t)stream>>>(iters, sink);
    else
        mb_ffma2_kernel<0><<<ctas, 256, 0, (cudaStream_t)stream>>>(iters, sink);
    g_launches++;
    return (int64_t)ctas * 256 * (int64_t)iters * 4 * MB_ILP * 2;  // lane-FMAs
}

int64_t pigs_microbench_ffma(int ctas, int iters, float* sink, void* stream)
{
    mb_ffma_kernel<<<ctas, 256, 0, (cudaStream_t)stream>>>(iters, sink);
    g_launches++;
    return (int64_t)ctas * 256 * (int64_t)iters * 4 * MB_ILP;
}
int64_t pigs_microbench_ex2(int ctas, int iters, float* sink, void* stream)
{
    mb_ex2_kernel<<<ctas, 256, 0, (cudaStream_t)stream>>>(iters, sink);
    g_launches++;
    return (int64_t)ctas * 256 * (int64_t)iters * 4 * MB_ILP;
}

}  // extern "C"
