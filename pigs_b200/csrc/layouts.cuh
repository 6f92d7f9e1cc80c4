// layouts.cuh — byte-buffer layouts (all torch-owned) and the argument bundles of the evaluation path.
#pragma once
#include "eval_common.cuh"

namespace pigs {

struct GeomLayout {
    pigs_header* hdr;
    uint32_t* touched;  // [P+1]; becomes exclusive offsets (+ total) after the scan
    char* scan_ws;
    float* cull_r;  // [P] culled-mode radius r_c (written only by pigs_preprocess_count_culled with subdiv > 1)
    size_t bytes;
    GeomLayout(void* base, int64_t P)
    {
        Carver c(base);
        hdr = c.take<pigs_header>(1);
        touched = c.take<uint32_t>((size_t)P + 1);
        scan_ws = c.take<char>(scan_ws_bytes(P));
        cull_r = c.take<float>((size_t)(P > 0 ? P : 1));
        bytes = c.total();
    }
};

// work lists of the evaluation kernels, built ONCE per preprocess (they depend on the binning only): forward with
// 8 or 4 samples per thread, backward.  A WorkTemplate is copied into the per-call WorkHeader by the packing kernel.
constexpr int WL_FWD8 = 0, WL_FWD4 = 1, WL_BWD = 2, WL_COUNT = 3;
struct WorkTemplate {
    uint32_t items[4];  // items per class (FAST, SHIFT, GENERAL)
    unsigned long long pairs;
    uint32_t pad[2];
};
// class prefix word: GENERAL count in the low 31 bits, SHIFT count in the next 31 (one 64-bit scan does both)
constexpr unsigned long long CLS_GENERAL = 1ull, CLS_SHIFT = 1ull << 31, CLS_MASK = (1ull << 31) - 1ull;

struct BinLayout {
    pigs_header* hdr;
    uint32_t* list;   // [L] sorted ids
    uint32_t* tiles;  // [L] sorted tile keys
    // evaluation staging that depends on the binning only (pack.cu stage_classes), so every forward / backward call of
    // the preprocess cycle reuses it instead of rebuilding it:
    unsigned long long* cls;  // [L+1] exclusive prefix of the packed class flags (+ total)
    uint32_t* scode;          // [L]   wrap-shift code per instance
    TileSeg* segs;            // [WL_COUNT][EV_NSEG][blocks]
    WorkTemplate* wt;         // [WL_COUNT]
    char* scan_ws;
    size_t bytes;
    BinLayout(void* base, int64_t L, int blocks = 1)
    {
        Carver c(base);
        const size_t n = (size_t)(L > 0 ? L : 1);
        hdr = c.take<pigs_header>(1);
        list = c.take<uint32_t>(n);
        tiles = c.take<uint32_t>(n);
        cls = c.take<unsigned long long>(n + 1);
        scode = c.take<uint32_t>(n);
        segs = c.take<TileSeg>((size_t)WL_COUNT * EV_NSEG * (blocks > 0 ? blocks : 1));
        wt = c.take<WorkTemplate>(WL_COUNT);
        scan_ws = c.take<char>(scan_ws_bytes(L));
        bytes = c.total();
    }
};

struct SampleBinLayout {
    pigs_header* hdr;
    uint32_t* list;
    uint32_t* tiles;
    float* bbox;  // [nslots][2*D]
    size_t bytes;
    SampleBinLayout(void* base, int64_t N, int blocks, int D)
    {
        Carver c(base);
        hdr = c.take<pigs_header>(1);
        list = c.take<uint32_t>((size_t)(N > 0 ? N : 1));
        tiles = c.take<uint32_t>((size_t)(N > 0 ? N : 1));
        // nslots = blocks + grid[0] + 2 <= 2*blocks + 2
        bbox = c.take<float>((size_t)(2 * blocks + 4) * 2 * D);
        bytes = c.total();
    }
};


// everything an evaluation call needs, resolved from the opaque buffers
struct EvalCall {
    int D, C, blocks;
    int64_t P, N, R;
    const float *means, *values, *conics, *samples;
    const uint32_t *list, *tiles;   // sorted Gaussian ids / tile keys  [R]
    const uint32_t* slist;          // sorted sample ids               [N]
    const float* bbox;              // per-tile sample bounding boxes
    const uint2 *ranges, *sample_ranges;
    const unsigned long long* cls;  // class prefix (BinLayout)
    const uint32_t* scode;
    const TileSeg* segs;            // [WL_COUNT][EV_NSEG][blocks]
    const WorkTemplate* wt;         // [WL_COUNT]
    bool single_class = false;      // the binning put every instance into the GENERAL class (small_binning_*): one kernel
};

// per-call workspace (torch-owned, sized by pigs_eval_workspace_bytes)
struct EvalWs {
    WorkHeader* wh;
    float4* recs;      // [R][frec_v4] (forward layout) or [R][grec_v4] (backward layout)
    uint32_t* rec_id;  // [R]
    float4* srecs;     // [N][srec_v4]   (backward only)
    size_t bytes;
    EvalWs(void* base, int64_t R, int64_t N, int blocks, int gv4, int sv4)
    {
        (void)blocks;
        Carver c(base);
        wh = c.take<WorkHeader>(1);
        recs = c.take<float4>((size_t)(R > 0 ? R : 1) * gv4);
        rec_id = c.take<uint32_t>((size_t)(R > 0 ? R : 1));
        srecs = c.take<float4>((size_t)(N > 0 ? N : 1) * sv4);
        bytes = c.total();
    }
};

// once per preprocess (called from pigs_preprocess_bin): classify every (tile, Gaussian) instance, scan the class flags,
// build the three work lists
int stage_classes(int D, int blocks, int64_t R, int64_t N, const float* means, const float* conics, const float* bbox,
                  const uint2* ranges, const uint2* sample_ranges, const BinLayout& b, cudaStream_t st, bool single_class = false);

// Small problems (the reference's own drivers: 1 K Gaussians x 10 K points) are launch-bound: the three class kernels of a
// call plus the fork / join of their streams cost more than the arithmetic.  Below g_small_pairs estimated pairs
// (instances x samples per tile) stage_classes puts EVERY instance into the GENERAL class (exact per-pair wrap and skip:
// the reference's own arithmetic, so results stay inside the parity tolerance of the three-class path) and writes the
// scanned class prefix directly (no scan launch); forward / backward then launch the GENERAL kernel alone.  What a call
// may assume about a binning buffer is recorded on the host when pigs_preprocess_bin fills it (small_binning_note) and
// looked up by (pointer, R, N, blocks) at evaluation time (small_binning_lookup): an unknown buffer — a copy, another
// process' — gets the three launches, which are correct for either kind of buffer.
extern int64_t g_small_pairs;  // pigs_set_small_problem_pairs; 0 = never; -1 = not read from the environment yet
int64_t small_pairs_threshold();
bool small_problem(int64_t R, int64_t N, int blocks);
void small_binning_note(const void* binning_buffer, int64_t R, int64_t N, int blocks, bool single_class);
bool small_binning_lookup(const void* binning_buffer, int64_t R, int64_t N, int blocks);
// per call: packed Gaussian records in class order + the call's WorkHeader (from work list `wl`)
int stage_gaussians(const EvalCall& c, int C, int ch0, bool fwd_layout, int wl, const EvalWs& ws, cudaStream_t st);
// lap: gH is the upstream gradient of the Laplacian [N, C] (PIGS_FN_LAPL) instead of the Hessian's [N, D, D, C]
int stage_samples(const EvalCall& c, int C, unsigned mask, int ch0, const float* gV, const float* gD, const float* gH,
                  const float* gT, bool lap, const EvalWs& ws, cudaStream_t st);

// Fork / join of the caller's stream into the three class streams (library-owned, created once per device, non-blocking;
// events without timing).  stream(0) is the caller's stream itself.  Legal inside CUDA-graph capture (the side streams
// join back before the call returns).  pigs_set_class_concurrency(0), debug mode and deterministic mode run the three
// kernels in order on the caller's stream instead.
struct ClassFork {
    cudaStream_t main = nullptr;
    bool forked = false;
    int fork(cudaStream_t st);
    cudaStream_t stream(int k) const;
    int join();
};
extern int g_class_concurrency;

// evaluation launchers (eval_fwd.cu / eval_bwd.cu); C is the channel-group width (1 or 2)
struct FwdOut {
    float *u, *du, *ddu, *dddu;
    int lap;  // 1: ddu is the Laplacian [N, C] (PIGS_FN_LAPL), 0: the Hessian [N, D, D, C]
};
int launch_forward(const EvalCall& c, int C, unsigned mask, int wl, int ch0, const EvalWs& ws, FwdOut out, cudaStream_t st);
struct BwdOut {
    float *dmeans, *dvalues, *dconics;
};
int launch_backward(const EvalCall& c, int C, unsigned mask, int ch0, const EvalWs& ws, BwdOut out, cudaStream_t st);

}  // namespace pigs
