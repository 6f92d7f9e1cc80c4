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

struct BinLayout {
    pigs_header* hdr;
    uint32_t* list;   // [L] sorted ids
    uint32_t* tiles;  // [L] sorted tile keys
    size_t bytes;
    BinLayout(void* base, int64_t L)
    {
        Carver c(base);
        hdr = c.take<pigs_header>(1);
        list = c.take<uint32_t>((size_t)(L > 0 ? L : 1));
        tiles = c.take<uint32_t>((size_t)(L > 0 ? L : 1));
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
};

// per-call workspace (torch-owned, sized by pigs_eval_workspace_bytes)
struct EvalWs {
    WorkHeader* wh;
    uint32_t* flags;   // [R+2] GENERAL flags -> exclusive prefix
    uint32_t* flags1;  // [R+2] SHIFT flags   -> exclusive prefix
    uint32_t* scode;   // [R]   wrap-shift code per instance
    char* scan_ws;
    uint32_t* nfast;   // [2][blocks+1]: number of FAST, then of SHIFT instances per tile
    TileSeg* segs;     // [EV_NSEG][blocks]
    float4* recs;      // [R][frec_v4] (forward layout) or [R][grec_v4] (backward layout)
    uint32_t* rec_id;  // [R]
    float4* srecs;     // [N][srec_v4]   (backward only)
    size_t bytes;
    EvalWs(void* base, int64_t R, int64_t N, int blocks, int gv4, int sv4)
    {
        Carver c(base);
        wh = c.take<WorkHeader>(1);
        flags = c.take<uint32_t>((size_t)R + 2);
        flags1 = c.take<uint32_t>((size_t)R + 2);
        scode = c.take<uint32_t>((size_t)(R > 0 ? R : 1));
        scan_ws = c.take<char>(scan_ws_bytes(R));
        nfast = c.take<uint32_t>((size_t)2 * (blocks + 1));
        segs = c.take<TileSeg>((size_t)EV_NSEG * blocks);
        recs = c.take<float4>((size_t)(R > 0 ? R : 1) * gv4);
        rec_id = c.take<uint32_t>((size_t)(R > 0 ? R : 1));
        srecs = c.take<float4>((size_t)(N > 0 ? N : 1) * sv4);
        bytes = c.total();
    }
};

int stage_gaussians(const EvalCall& c, int C, int ch0, bool fwd_layout, const EvalWs& ws, cudaStream_t st);
int stage_worklist(const EvalCall& c, const EvalWs& ws, uint32_t chunk, uint32_t split_hint, uint32_t min_stream,
                   bool stream_is_gauss, cudaStream_t st);
// lap: gH is the upstream gradient of the Laplacian [N, C] (PIGS_FN_LAPL) instead of the Hessian's [N, D, D, C]
int stage_samples(const EvalCall& c, int C, unsigned mask, int ch0, const float* gV, const float* gD, const float* gH,
                  const float* gT, bool lap, const EvalWs& ws, cudaStream_t st);

// evaluation launchers (eval_fwd.cu / eval_bwd.cu); C is the channel-group width (1 or 2)
struct FwdOut {
    float *u, *du, *ddu, *dddu;
    int lap;  // 1: ddu is the Laplacian [N, C] (PIGS_FN_LAPL), 0: the Hessian [N, D, D, C]
};
int launch_forward(const EvalCall& c, int C, unsigned mask, int mode, int ch0, const EvalWs& ws, FwdOut out,
                   cudaStream_t st);
struct BwdOut {
    float *dmeans, *dvalues, *dconics;
};
int launch_backward(const EvalCall& c, int C, unsigned mask, int mode, int ch0, const EvalWs& ws, BwdOut out,
                    cudaStream_t st);

}  // namespace pigs
