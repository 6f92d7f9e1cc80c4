// worklist.cuh — device-side work lists of the evaluation kernels (built once per preprocess).  Shared by pack.cu
// (worklist_kernel: one CTA per list) and binning.cu (small_bin_kernel: the whole binning of a small problem in one CTA).
#pragma once
#include "layouts.cuh"

namespace pigs {

// One CTA of 256 threads per work list (wl = WL_FWD8 | WL_FWD4 | WL_BWD).  For every tile: the FAST / SHIFT / GENERAL segments
// and how many work items each yields.
//   forward : stationary = samples in chunks of `chunk`, streamed = Gaussians split `split_hint` ways
//   backward: stationary = Gaussians in chunks of `chunk`, streamed = samples split `split_hint` ways
struct WorkListCfg {
    uint32_t chunk[WL_COUNT], split_hint[WL_COUNT], min_stream[WL_COUNT];
    int stream_is_gauss[WL_COUNT];
};

// the lists' configuration for one binning (host; pack.cu): chunk sizes, split hints, finest streamed slice
WorkListCfg make_worklist_cfg(int64_t R, int64_t N, int blocks, bool single_class);

// barrier over the 256 threads that build a list: a NAMED barrier, so that the first 256 threads of a larger CTA
// (small_bin_kernel, 1024 threads) can run the body while the others wait at the next __syncthreads()
__device__ __forceinline__ void wl_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// Must be called by exactly the threads 0..255 of the CTA, all of them (barriers inside; grids of up to 32 tiles are
// handled by the first warp alone, without barriers).
__device__ __forceinline__ void worklist_body(int wl, int blocks, const uint2* __restrict__ ranges,
                                              const uint2* __restrict__ sample_ranges,
                                              const unsigned long long* __restrict__ cls, const WorkListCfg& cfg,
                                              TileSeg* __restrict__ segs_all /* [WL_COUNT][EV_NSEG][blocks] */,
                                              WorkTemplate* __restrict__ wt_all)
{
    const uint32_t chunk = cfg.chunk[wl], split_hint = cfg.split_hint[wl], min_stream = cfg.min_stream[wl];
    const int stream_is_gauss = cfg.stream_is_gauss[wl];
    TileSeg* segs = segs_all + (size_t)wl * EV_NSEG * blocks;
    if (blocks <= 32) {
        // the usual 4 x 4 grid (any grid of up to 32 tiles): ONE warp, one tile per lane, prefix sums by shuffle — none of
        // the block-wide scan's 72 barrier rounds per list (they were a quarter of small_bin_kernel's time)
        if (threadIdx.x >= 32) return;
        const int t = threadIdx.x;
        TileSeg sg[EV_NSEG];
        uint32_t items[EV_NSEG];
        unsigned long long pairs = 0;
#pragma unroll
        for (int k = 0; k < EV_NSEG; k++) items[k] = 0;
        if (t < blocks) {
            const uint2 gr = ranges[t], sr = sample_ranges[t];
            const uint32_t G = gr.y - gr.x, S = sr.y - sr.x;
            uint32_t ng = 0, ns = 0;
            if (G > 0) {
                const unsigned long long c0 = cls[gr.x], c1 = cls[gr.y];
                ng = (uint32_t)((c1 & CLS_MASK) - (c0 & CLS_MASK));
                ns = (uint32_t)((c1 >> 31) - (c0 >> 31));
            }
            const uint32_t nf = G - ng - ns;
            pairs = (unsigned long long)G * S;
            const uint32_t begin[EV_NSEG] = {gr.x, gr.x + nf, gr.x + nf + ns};
            const uint32_t count[EV_NSEG] = {nf, ns, ng};
#pragma unroll
            for (int k = 0; k < EV_NSEG; k++) {
                sg[k].g_begin = begin[k];
                sg[k].g_count = count[k];
                sg[k].s_begin = sr.x;
                sg[k].s_count = S;
                const uint32_t stat_n = stream_is_gauss ? sg[k].s_count : sg[k].g_count;
                const uint32_t strm_n = stream_is_gauss ? sg[k].g_count : sg[k].s_count;
                uint32_t split = 1;
                if (stat_n > 0 && strm_n > 0) {
                    split = min(split_hint, (strm_n + min_stream - 1) / min_stream);
                    if (split < 1) split = 1;
                    items[k] = ((stat_n + chunk - 1) / chunk) * split;
                }
                sg[k].split = split;
            }
        }
        WorkTemplate w;
#pragma unroll
        for (int k = 0; k < EV_NSEG; k++) {
            uint32_t inc = items[k];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
                if (t >= o) inc += v;
            }
            if (t < blocks) {
                sg[k].item_begin = inc - items[k];
                segs[(size_t)k * blocks + t] = sg[k];
            }
            w.items[k] = __shfl_sync(0xffffffffu, inc, 31);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) pairs += __shfl_xor_sync(0xffffffffu, pairs, o);
        if (t == 0) {
            w.items[3] = 0;
            w.pairs = pairs;
            w.pad[0] = w.pad[1] = 0;
            wt_all[wl] = w;
        }
        return;
    }
    __shared__ uint32_t s_scan[256];
    __shared__ uint32_t s_carry[EV_NSEG];
    __shared__ unsigned long long s_pairs;
    if (threadIdx.x == 0) {
        for (int k = 0; k < EV_NSEG; k++) s_carry[k] = 0;
        s_pairs = 0ull;
    }
    wl_sync();
    unsigned long long my_pairs = 0;
    for (int base = 0; base < blocks; base += 256) {
        const int t = base + threadIdx.x;
        TileSeg sg[EV_NSEG];
        uint32_t items[EV_NSEG];
#pragma unroll
        for (int k = 0; k < EV_NSEG; k++) items[k] = 0;
        if (t < blocks) {
            const uint2 gr = ranges[t], sr = sample_ranges[t];
            const uint32_t G = gr.y - gr.x, S = sr.y - sr.x;
            uint32_t ng = 0, ns = 0;
            if (G > 0) {
                const unsigned long long c0 = cls[gr.x], c1 = cls[gr.y];
                ng = (uint32_t)((c1 & CLS_MASK) - (c0 & CLS_MASK));
                ns = (uint32_t)((c1 >> 31) - (c0 >> 31));
            }
            const uint32_t nf = G - ng - ns;
            my_pairs += (unsigned long long)G * S;
            const uint32_t begin[EV_NSEG] = {gr.x, gr.x + nf, gr.x + nf + ns};
            const uint32_t count[EV_NSEG] = {nf, ns, ng};
#pragma unroll
            for (int k = 0; k < EV_NSEG; k++) {
                sg[k].g_begin = begin[k];
                sg[k].g_count = count[k];
                sg[k].s_begin = sr.x;
                sg[k].s_count = S;
                const uint32_t stat_n = stream_is_gauss ? sg[k].s_count : sg[k].g_count;
                const uint32_t strm_n = stream_is_gauss ? sg[k].g_count : sg[k].s_count;
                uint32_t split = 1;
                if (stat_n > 0 && strm_n > 0) {
                    split = min(split_hint, (strm_n + min_stream - 1) / min_stream);
                    if (split < 1) split = 1;
                    items[k] = ((stat_n + chunk - 1) / chunk) * split;
                }
                sg[k].split = split;
            }
        }
#pragma unroll
        for (int k = 0; k < EV_NSEG; k++) {
            // block exclusive scan of items[k]
            s_scan[threadIdx.x] = items[k];
            wl_sync();
            for (int o = 1; o < 256; o <<= 1) {
                uint32_t v = threadIdx.x >= o ? s_scan[threadIdx.x - o] : 0;
                wl_sync();
                s_scan[threadIdx.x] += v;
                wl_sync();
            }
            const uint32_t incl = s_scan[threadIdx.x];
            const uint32_t carry = s_carry[k];
            if (t < blocks) {
                sg[k].item_begin = carry + incl - items[k];
                segs[(size_t)k * blocks + t] = sg[k];
            }
            wl_sync();
            if (threadIdx.x == 255) s_carry[k] = carry + incl;
            wl_sync();
        }
    }
    atomicAdd(&s_pairs, my_pairs);
    wl_sync();
    if (threadIdx.x == 0) {
        WorkTemplate w;
        for (int k = 0; k < EV_NSEG; k++) w.items[k] = s_carry[k];
        w.items[3] = 0;
        w.pairs = s_pairs;
        w.pad[0] = w.pad[1] = 0;
        wt_all[wl] = w;
    }
}


}  // namespace pigs
