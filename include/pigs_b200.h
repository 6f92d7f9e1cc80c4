/*
 * pigs_b200.h — C ABI of libpigs_b200.so: the B200-native (sm_100a) implementation of the
 * differentiable Gaussian-mixture sampling path of kr4b/pigs (`diff-gaussian-sampling`).
 *
 * This is the drop-in boundary.  Every entry point takes plain device pointers, sizes and a
 * cudaStream_t (as void*), returns an int status (0 = ok, >0 = cudaError_t, <0 = argument
 * error; pigs_last_error() gives the text), never allocates, never frees, never retains a
 * pointer past the call, and never synchronises the host unless documented.  The caller
 * (torch, through pigs_b200/_C.py) owns every allocation.
 *
 * Reference interfaces replaced (paths relative to /root/reference/diff-gaussian-sampling):
 *   ext.cpp:20-31                  the 12 pybind entry points
 *   sample_points.h:20-131         PreprocessCUDA / SampleGaussians*CUDA / *BackwardCUDA
 *   aggregate_neighbors.h:11-47    AggregateNeighbors{Preprocess,,Backward}CUDA
 *   cuda_sampler/sampler.h:26-80   CudaSampler::Sampler::{preprocess,forward,backward}
 *
 * All float data is fp32, ids are u32, tile ranges are uint2 (start,end) pairs, the aggregate
 * CSR is int64 — the same dtypes the reference exposes.
 */
#ifndef PIGS_B200_H
#define PIGS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PIGS_TILE_WIDTH 0.51f /* cuda_sampler/config.h:18 BLOCK_SIZE */

/* function bits (cuda_sampler/sampler.h:22 enum Function {gaussian, derivative, laplacian, third}) */
#define PIGS_FN_VALUE 1u /* u        -> [N, C]          */
#define PIGS_FN_DERIV 2u /* grad u   -> [N, D, C]       */
#define PIGS_FN_HESS 4u  /* Hessian  -> [N, D, D, C]    */
#define PIGS_FN_THIRD 8u /* 3rd der. -> [N, D, D, D, C] */
/* Extension: the Laplacian = trace of the Hessian -> [N, C], what pde_rhs consumes
 * (model_pn.py:613-625 `uxx[:,0,0] + uxx[:,1,1]`).  Mutually exclusive with PIGS_FN_HESS; it uses the Hessian's
 * argument slots: out_ddu / dL_dddu then hold [N, C] instead of [N, D, D, C].  Same pair arithmetic as the Hessian
 * (sum formed once per sample in the epilogue), so lap == H_00 + H_11 of the PIGS_FN_HESS result up to the order of
 * the per-item atomic adds; the [N,D,D,C] tensor and its upstream gradient never touch HBM. */
#define PIGS_FN_LAPL 16u

/* Device-resident header written by pigs_preprocess_count at the start of `geom_ws` and copied to
 * the start of `binning_buffer` by pigs_preprocess_bin.  The host reads it back ONCE per preprocess
 * (one pinned-memory copy) to size the binning buffers — the reference needs two host syncs there
 * (sample_points.cu:74 `.item()`, sampler_impl.cu:257 cudaMemcpy) and one more in every
 * forward/backward call (sample_points.cu:124,174), which this ABI does not need at all. */
typedef struct pigs_header {
    uint32_t magic;        /* 'PIGS' */
    int32_t D;             /* 1 or 2 */
    int32_t grid[2];       /* tiles per dimension (sample_points.cu:73) */
    int32_t blocks;        /* prod(grid) */
    int32_t nslots;        /* tile-id slots incl. alias slots: blocks + grid[0] + 2 */
    float min_b[2];        /* grid offset = per-dimension min over samples (sample_points.cu:70) */
    float max_b[2];
    int64_t P, N;
    int64_t num_rendered;  /* Gaussian/tile instances (sampler_impl.cu:256-257) */
    /* culled mode only (pigs_preprocess_count_culled); zero / reference values otherwise */
    float tile_width;      /* width of the tiles the lists are built on: 0.51f, or 0.51f / subdiv */
    float cull_sigmas;     /* 0 = reference pair set */
    int32_t subdiv;        /* k: every reference tile is cut k x k; 0 or 1 = reference tiles */
    int32_t ref_grid[2];   /* the reference's grid (S1); grid[] above is then k * ref_grid[] */
    uint32_t reserved[43]; /* pad to 256 bytes */
} pigs_header;

int pigs_version(void);
const char *pigs_last_error(void);

/* Library-level option.  recip=1 (default): tile grid = ceil((max-min+1e-6f) * fl(1/0.51f)), the
 * arithmetic of ATen's CUDA division-by-scalar which the reference calls at sample_points.cu:73;
 * recip=0: IEEE division. */
void pigs_set_grid_recip_mode(int recip);

/* ---- S1: sample bounds (replaces samples.min(0)/max(0), sample_points.cu:70-71) ------------------
 * d_minmax: device float[2*D] = {min[0..D), max[0..D)}.  Multi-GPU callers all-reduce it
 * (MIN on the first D, MAX on the last D) before pigs_preprocess_count. */
int pigs_sample_bounds(const float *samples, int64_t N, int D, float *d_minmax, void *stream);

/* ---- preprocess stage A: grid, radii, tiles_touched, offsets (forward.cu:24-83, sampler_impl.cu:253) */
size_t pigs_geom_workspace_bytes(int64_t P);
int pigs_preprocess_count(int64_t P, int D, int64_t N, const float *means, const float *covariances,
                          const float *d_minmax, float *radii /* [P] out */, void *geom_ws,
                          void *stream);
/* after this call the first sizeof(pigs_header) bytes of geom_ws hold the header (device memory) */

/* Opt-in CULLED mode — NOT the reference's pair set, a subset of it.  The reference evaluates
 * every (Gaussian, sample) pair whose 0.51-wide tile is in the Gaussian's tile rectangle (S8): at C3 about 128 K
 * Gaussians per sample, of which < 10 K lie within 6 sigma.  Here every reference tile is cut into subdiv x subdiv fine
 * tiles and a (Gaussian, fine tile) instance is kept iff the reference emits the parent tile AND the fine tile lies
 * within r_c = cull_sigmas * sqrt(lambda_max(Sigma)) of the mean in every dimension (period-2 wrap, no 1e-6 floor).
 * Every dropped pair is farther than cull_sigmas standard deviations and contributes < exp(-cull_sigmas^2 / 2) of its
 * value.  `radii` keeps the reference radii.  Everything downstream (pigs_preprocess_bin, forward, backward) reads the
 * parameters from the header and is otherwise unchanged; blocks / ranges / lists are the fine ones.  bench and tests
 * report this mode separately, with its own pair count and its measured deviation from the exact mode. */
int pigs_preprocess_count_culled(int64_t P, int D, int64_t N, const float *means, const float *covariances,
                                 const float *d_minmax, float *radii, void *geom_ws, int subdiv,
                                 float cull_sigmas, void *stream);

/* Bounds + stage A in one call (what the bindings call).  have_bounds = 0: the sample bounds are computed from
 * `samples` and left in d_minmax (device float[2*D]); have_bounds = 1: d_minmax is an input (multi-GPU callers all-reduce
 * it first) and `samples` may be NULL.  subdiv = 1 / cull_sigmas = 0 for the reference pair set.  Small problems
 * (P <= 16 384 and N <= 16 384, the reference's own drivers; pigs_set_small_problem_pairs(0) switches it off) run ONE
 * single-CTA launch instead of three dependent ones; larger ones run pigs_sample_bounds + pigs_preprocess_count_culled.
 * Same results either way, bit for bit. */
int pigs_preprocess_count_fused(int64_t P, int D, int64_t N, const float *means, const float *covariances,
                                const float *samples, float *d_minmax, int have_bounds, float *radii, void *geom_ws,
                                int subdiv, float cull_sigmas, void *stream);

/* ---- preprocess stage B: key emission, stable sort by tile, ranges; sample binning ---------------
 * (sampler_impl.cu:54-189, 259-327).  Sizes are functions of the header values read back by the host.
 * Also builds, once per preprocess, the evaluation staging that depends on the binning only and is reused by every
 * forward / backward call of the cycle (kept inside the opaque binning_buffer): the class of every (tile, Gaussian)
 * instance (no torus wrap possible / constant wrap / general — decided from `conics`, the ones PreprocessCUDA receives,
 * sample_points.cu:41, and the tile's sample bounding box), its scan, and the kernels' device-side work lists. */
size_t pigs_binning_bytes(int64_t num_rendered, int blocks);
size_t pigs_sample_binning_bytes(int64_t N, int blocks, int D);
size_t pigs_sort_workspace_bytes(int64_t num_rendered, int64_t N);
int pigs_preprocess_bin(int64_t P, int D, int64_t N, int blocks, int64_t num_rendered, const float *means,
                        const float *conics, const float *samples, const float *radii, const void *geom_ws,
                        void *binning_buffer, void *sample_binning_buffer,
                        void *ranges /* uint2[blocks+1] out */, void *sample_ranges /* uint2[blocks+1] out */,
                        void *sort_ws, void *stream);

/* views into the opaque buffers (device pointers; for parity tests and debugging) */
const uint32_t *pigs_point_list(const void *binning_buffer);           /* [num_rendered] (tile asc, id asc) */
const uint32_t *pigs_point_tiles(const void *binning_buffer, int64_t num_rendered);
const uint32_t *pigs_sample_list(const void *sample_binning_buffer);   /* [N] (tile asc, id asc) */

/* ---- evaluation (forward.cu:88-343, backward.cu:27-499) -----------------------------------------
 * fn_mask: any non-empty subset of PIGS_FN_*; outputs whose bit is clear may be NULL.
 * Outputs MUST be zero-initialised by the caller (as sample_points.cu:215,240,265,290 / :165-167 do);
 * results are accumulated with atomic adds.  No host synchronisation. */
size_t pigs_eval_workspace_bytes(int64_t num_rendered, int64_t N, int blocks, int D, int C, unsigned fn_mask);
int pigs_sample_forward(unsigned fn_mask, int64_t P, int D, int64_t N, int C, int blocks, int64_t num_rendered,
                        const float *means, const float *values, const float *conics, const float *samples,
                        const void *binning_buffer, const void *sample_binning_buffer, const void *ranges,
                        const void *sample_ranges, float *out_u, float *out_du, float *out_ddu,
                        float *out_dddu, void *workspace, size_t workspace_bytes, void *stream);
int pigs_sample_backward(unsigned fn_mask, int64_t P, int D, int64_t N, int C, int blocks, int64_t num_rendered,
                         const float *means, const float *values, const float *conics, const float *samples,
                         const void *binning_buffer, const void *sample_binning_buffer, const void *ranges,
                         const void *sample_ranges, const float *dL_du, const float *dL_ddu,
                         const float *dL_dddu, const float *dL_ddddu, float *dL_dmeans, float *dL_dvalues,
                         float *dL_dconics, void *workspace, size_t workspace_bytes, void *stream);

/* number of kernels launched by this library since load (bench.py's gpu_launches) and the pair count
 * of the last forward/backward work list (device scalar, int64) */
int64_t pigs_launch_count(void);

/* Debug mode = the reference's `debug` flag (cuda_sampler/auxiliary.h:33-40 CHECK_CUDA: cudaDeviceSynchronize + error
 * check after EACH kernel launch, std::runtime_error naming file:line).  While on, every launch of this library is followed
 * by a device synchronisation and a failing kernel makes the entry point return its cudaError_t with pigs_last_error()
 * naming the launch site.  The Python binding sets it from the `debug` argument of every call. */
void pigs_set_debug(int on);
int pigs_get_debug(void);

/* Deterministic forward (default off; read by pigs_preprocess_bin).  The reference forward gives every sample to one
 * thread and adds its Gaussians in ascending id (forward.cu:100-166), so its outputs are bitwise reproducible; the default
 * mode here splits a tile's Gaussians over several work items and classes that each add their partial sum with one
 * RED.ADD, so u / grad u / Hessian vary at the 1e-6 level from run to run.  With the option on, every instance takes the
 * GENERAL class (exact per-pair wrap and skip; a tile's list stays in ascending id) and the forward's Gaussians are not
 * split: each output element is written by exactly one work item, in the reference's order -> run-to-run bit equality
 * of pigs_sample_forward, at the GENERAL loop's pace.  The backward was never deterministic in the reference either
 * (atomics per pair, backward.cu:131-147) and is not made so here. */
void pigs_set_deterministic(int on);
int pigs_get_deterministic(void);

/* Library-level option (default 1): the FAST / SHIFT / GENERAL kernels of one forward / backward call run concurrently on
 * three streams forked from and joined back into the caller's stream (one launch tail instead of three).  0 launches them
 * in order on the caller's stream.  Results are the same up to the order of the RED.ADDs. */
void pigs_set_class_concurrency(int on);

/* Library-level option (default 20 000 000; environment PIGS_SMALL_PAIRS; read by pigs_preprocess_bin).  Problems whose
 * estimated pair count — num_rendered x N / blocks — is at most `pairs` are launch-bound (the reference's own drivers:
 * 1 K Gaussians x 10 K points), so the binning puts every (tile, Gaussian) instance into the GENERAL class (the exact
 * per-pair torus wrap and `power > 0` skip of forward.cu:149-171: the reference's own arithmetic) and every forward /
 * backward call of that binning launches ONE evaluation kernel instead of three on forked streams.  Results stay within
 * the parity tolerance of the three-class path (tests compare the two).  0 disables it.  The library remembers on the
 * host which binning buffers were built this way (by address and sizes); a buffer it does not know — e.g. a copy — gets
 * the three launches, which are correct for both kinds.  (Corollary: a binning buffer is written by pigs_preprocess_bin
 * only — copying another binning's bytes INTO a buffer the library has noted is not supported.) */
void pigs_set_small_problem_pairs(int64_t pairs);

/* Optional per-stage device timing of the preprocess chain (development: tools/chain_timing.py).  While on, the
 * library records a cudaEvent behind each stage of pigs_sample_bounds / pigs_preprocess_count / pigs_preprocess_bin;
 * pigs_stage_timing_read synchronises on them and returns (stage name, ms since the previous mark) pairs. */
void pigs_set_stage_timing(int on);
int pigs_stage_timing_read(const char **names, float *ms, int cap);

/* Optional device timing of the evaluation kernels (bench.py roofline).  While on, the forward /
 * backward evaluation launches of every pigs_sample_forward / pigs_sample_backward call are
 * bracketed by cudaEvents on the call's stream.  pigs_kernel_timing_read(which = 0 forward |
 * 1 backward) synchronises on those events, returns their summed duration and count and clears them. */
void pigs_set_kernel_timing(int on);
int pigs_kernel_timing_read(int which, float *ms_sum, int *launches);

/* ---- neighbour lists + aggregation (aggregate_neighbors.cu:18-475) ------------------------------ */
size_t pigs_neighbor_workspace_bytes(int64_t P, int D);
/* stage A: per-Gaussian neighbour counts (findCollisions semantics) -> counts int64[P] */
int pigs_neighbors_count(int64_t P, int D, const float *means, const float *radii, int64_t *counts,
                         void *workspace, void *stream);
/* stage B: caller computes ranges = inclusive cumsum(counts) (int64[P]) and allocates the CSR arrays (no pre-fill
 * needed: every slot is written, skipped neighbours as index -1 / density 0); this fills them (preprocess, :57-127). */
int pigs_neighbors_fill(int64_t P, int D, const float *means, const float *conics, const float *radii,
                        const int64_t *ranges, int64_t *indices, float *dists, float *densities,
                        float *inv_total_densities, void *workspace, void *stream);
/* between the stages: ranges = inclusive prefix sum of counts (the library's single-pass scan; the reference uses
 * cub::DeviceScan / torch.cumsum, aggregate_neighbors.cu:343-350) and *total (device int64) = the CSR length M, which
 * the host reads back once to size the arrays.  workspace: pigs_neighbors_scan_workspace_bytes(P). */
size_t pigs_neighbors_scan_workspace_bytes(int64_t P);
int pigs_neighbors_ranges(int64_t P, const int64_t *counts, int64_t *ranges, int64_t *total, void *workspace,
                          void *stream);
/* Allocation-free variant of pigs_neighbors_ranges.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same inputs, outputs and results as pigs_neighbors_ranges, but the scan's scratch (ticket +
 * one 64-bit descriptor per 4096-row tile) is a file-scope `static __device__` array inside the library (32 KB + 256 B
 * per GPU that loads the library, held until the process exits) whose device address is resolved once per device with
 * cudaGetSymbolAddress; the call itself performs no cudaMalloc / cudaFree / allocating library routine.
 * Limit: P <= PIGS_RANGES_STATIC_MAX_P rows; a larger P returns -1 with a pigs_last_error() text naming the limit.
 * Every call on a device shares the one array (the call clears the words it uses before its launch): calls that OVERLAP
 * on the GPU — different streams, or graph branches without an ordering edge — race; serialise them.  The first call on a
 * device resolves the symbol address; make it outside graph capture. */
#define PIGS_RANGES_STATIC_MAX_P 16777216 /* 2^24 rows (16x the P = 1M of BASELINE config 5) */
int pigs_neighbors_ranges_static(int64_t P, const int64_t *counts, int64_t *ranges, int64_t *total, void *stream);
/* The same two stages for the rows (query Gaussians) [row_lo, row_hi) only — multi-GPU aggregation shards the CSR by
 * query Gaussian.  Every Gaussian stays a candidate neighbour; rows outside the window get count 0 and
 * the no-neighbour total 1/(0 + 1e-6), so the aggregation kernels run unchanged on the sharded CSR. */
int pigs_neighbors_count_rows(int64_t P, int D, const float *means, const float *radii, int64_t row_lo, int64_t row_hi,
                              int64_t *counts, void *workspace, void *stream);
int pigs_neighbors_fill_rows(int64_t P, int D, const float *means, const float *conics, const float *radii,
                             int64_t row_lo, int64_t row_hi, const int64_t *ranges, int64_t *indices, float *dists,
                             float *densities, float *inv_total_densities, void *workspace, void *stream);
int pigs_aggregate_forward(int64_t P, int D, int L, int K, int E, const float *features,
                           const float *transform, const float *queries, const float *keys,
                           const float *frequencies, const float *distance_transform,
                           const int64_t *indices, const int64_t *ranges, const float *dists,
                           const float *densities, const float *inv_total_densities, float *weights,
                           float *embeddings, float *factors, float *neighbor_features, void *stream);
/* Library-level option (default 1): sizes L = K = 16, E = 25 (model_pn.py:44-49) run kernels specialised for them
 * (aggregate16.cu); 0 forces the generic kernels (same results within rounding; tests compare the two). */
void pigs_set_aggregate_specialised(int on);
/* Library-level option of the neighbour search: cell_cap = candidates per grid cell the cell-cooperative kernels hold in
 * shared memory (default and maximum 4096; 0 sends every row through the per-row kernels), cell_div = grid cells per
 * largest effective radius (default 2).  Results do not depend on either; tests compare the paths. */
void pigs_set_neighbor_search(int cell_cap, float cell_div);
/* Library-level option (default 4096; environment PIGS_NB_SMALL_P): searches over at most max_P Gaussians — the reference's
 * own models hold 400 .. a few thousand (main_pn.py) — test every pair with one warp per row walking the candidates in
 * ascending id (the reference's loop order): one launch per stage instead of the grid path's dozen.  Same tests, same
 * per-slot arithmetic, same lists.  0 sends every search through the grid path. */
void pigs_set_neighbor_small(int64_t max_P);
size_t pigs_aggregate_backward_workspace_bytes(int64_t P, int L, int K, int E);
int pigs_aggregate_backward(int64_t P, int D, int L, int K, int E, int64_t M, const float *features,
                            const float *transform, const float *queries, const float *keys,
                            const float *frequencies, const float *distance_transform,
                            const int64_t *indices, const int64_t *ranges, const float *dists,
                            const float *densities, const float *weights, const float *embeddings,
                            const float *factors, const float *inv_total_densities,
                            const float *dL_dneighbor_features, float *dL_dfeatures, float *dL_dtransform,
                            float *dL_dqueries, float *dL_dkeys, float *dL_dfrequencies,
                            float *dL_ddistance_transform, void *workspace, void *stream);

/* Row-windowed aggregation (multi-GPU, SURVEY.md 8e: the CSR is sharded by query Gaussian).  Same kernels and arithmetic
 * as pigs_aggregate_forward / _backward (aggregate_neighbors.cu:129-321), restricted to the rows [row_lo, row_hi): only
 * those rows of the ROW-indexed arrays (queries, ranges, inv_total_densities, neighbor_features, dL_dneighbor_features,
 * dL_dqueries) are read or written — all addressed with the GLOBAL row index, so a caller that holds just its window passes
 * the window's base pointer displaced by -row_lo rows.  features / keys / dL_dfeatures / dL_dkeys are addressed by the
 * neighbour index j and must cover every j the window's CSR slots name.  ranges is the inclusive cumsum over all P rows
 * (rows outside the window empty), as pigs_neighbors_fill_rows produces it. */
int pigs_aggregate_forward_rows(int64_t P, int D, int L, int K, int E, int64_t row_lo, int64_t row_hi,
                                const float *features, const float *transform, const float *queries, const float *keys,
                                const float *frequencies, const float *distance_transform, const int64_t *indices,
                                const int64_t *ranges, const float *dists, const float *densities,
                                const float *inv_total_densities, float *weights, float *embeddings, float *factors,
                                float *neighbor_features, void *stream);
int pigs_aggregate_backward_rows(int64_t P, int D, int L, int K, int E, int64_t M, int64_t row_lo, int64_t row_hi,
                                 const float *features, const float *transform, const float *queries, const float *keys,
                                 const float *frequencies, const float *distance_transform, const int64_t *indices,
                                 const int64_t *ranges, const float *dists, const float *densities, const float *weights,
                                 const float *embeddings, const float *factors, const float *inv_total_densities,
                                 const float *dL_dneighbor_features, float *dL_dfeatures, float *dL_dtransform,
                                 float *dL_dqueries, float *dL_dkeys, float *dL_dfrequencies,
                                 float *dL_ddistance_transform, void *workspace, void *stream);

/* float64 instantiation of the (generic) aggregation kernels = the reference built with `typedef double FLOAT`
 * (cuda_sampler/config.h:20), which its one live test needs: torch.autograd.gradcheck of aggregate_neighbors on double
 * tensors (test_neighbor_aggregation.py:94-101).  Same arguments as pigs_aggregate_forward / _backward with every float
 * array as double (the CSR's dists / densities / inv_total_densities converted by the caller); all P rows. */
int pigs_aggregate_forward_f64(int64_t P, int D, int L, int K, int E, const double *features, const double *transform,
                               const double *queries, const double *keys, const double *frequencies,
                               const double *distance_transform, const int64_t *indices, const int64_t *ranges,
                               const double *dists, const double *densities, const double *inv_total_densities,
                               double *weights, double *embeddings, double *factors, double *neighbor_features,
                               void *stream);
int pigs_aggregate_backward_f64(int64_t P, int D, int L, int K, int E, const double *features, const double *transform,
                                const double *queries, const double *keys, const double *frequencies,
                                const double *distance_transform, const int64_t *indices, const int64_t *ranges,
                                const double *dists, const double *densities, const double *weights,
                                const double *embeddings, const double *factors, const double *inv_total_densities,
                                const double *dL_dneighbor_features, double *dL_dfeatures, double *dL_dtransform,
                                double *dL_dqueries, double *dL_dkeys, double *dL_dfrequencies,
                                double *dL_ddistance_transform, void *stream);

/* ---- covariance / conic construction ---------------------------------------
 * Replaces the torch chain of gaussians.py:163-193 `build_covariances(s, t)` (tanh, sqrt(prod), diag_embed, indexed
 * assignment, batched torch.inverse, flatten to [xx, xy, yy]) that feeds GaussianSampler.preprocess in every driver.
 * scaling [P, D], transform [P, 1] (ignored for D = 1) -> covariances, conics [P, D(D+1)/2].
 * Backward: upstream gradients of the two flattened outputs (either may be NULL = zero) -> dL/dscaling [P, D],
 * dL/dtransform [P, 1]; equals torch.autograd through the reference chain. */
int pigs_build_covariances_forward(int64_t P, int D, const float *scaling, const float *transform, float *covariances,
                                   float *conics, void *stream);
int pigs_build_covariances_backward(int64_t P, int D, const float *scaling, const float *transform,
                                    const float *dL_dcovariances, const float *dL_dconics, float *dL_dscaling,
                                    float *dL_dtransform, void *stream);

/* ---- PDE right-hand side + residual loss -------------------------------------
 * Replaces the torch chain behind the sampler in the PDE drivers: model_pn.py:612-625 `pde_rhs` and :822-834
 * `rhs = dt * pde_rhs(...)`, `pde_loss += mean((ut - rhs)**2)`.  Inputs are the sampler's outputs u [N, C],
 * ux [N, D, C], lap [N, C] (PIGS_FN_LAPL: the trace the reference takes of its Hessian) and ut [N, C].
 *   DIFFUSION rhs = lap ; BURGERS rhs = nu*lap - u*ux[:,0] ; WAVE (C == 2) rhs = (u1, 10*lap0 - 0.1*u1)
 *   r = ut - dt*rhs ; loss = mean(r^2)  (WAVE: 0.01*mean(r0^2) + mean(r1^2))
 * forward: writes the device scalar `loss` (deterministic two-stage reduction) and, if non-NULL, residual [N, C].
 * backward: dL_dloss is a device scalar; every output pointer may be NULL (not wanted); dL_dux is fully written
 * for BURGERS (components other than [:,0] are zero) and untouched otherwise. No host synchronisation. */
#define PIGS_PDE_DIFFUSION 1
#define PIGS_PDE_BURGERS 3
#define PIGS_PDE_WAVE 4 /* numbering = model_pn.py:13-19 `Problem` enum.auto() values */
size_t pigs_pde_residual_workspace_bytes(int64_t N, int C);
int pigs_pde_residual_forward(int problem, int64_t N, int D, int C, float nu, float dt, const float *u, const float *ux,
                              const float *lap, const float *ut, float *residual, float *loss, void *workspace,
                              void *stream);
int pigs_pde_residual_backward(int problem, int64_t N, int D, int C, float nu, float dt, const float *u, const float *ux,
                               const float *lap, const float *ut, const float *dL_dloss, float *dL_du, float *dL_dux,
                               float *dL_dlap, float *dL_dut, void *stream);

/* NAVIER_STOKES in vorticity form: model_pn.py:629-631 `pde_rhs`, :818 `wt`, :843-845 loss.
 *   rhs  = nu*(wxx[:,0,0] + wxx[:,1,1]) - (u[:,0]*wx[:,0] + u[:,1]*wx[:,1])
 *   loss = mean((ux[:,0,0] + ux[:,1,1])^2) + mean((wt - dt*rhs)^2)
 * u [N,2], ux [N,2,2] = [n,d,c], wx [N,2], wxx [N,2,2], wt [N] (all 16-byte aligned, contiguous).  forward writes the
 * device scalar `loss` and, if non-NULL, residual [N] = wt - dt*rhs and divergence [N].  backward: every output
 * pointer may be NULL; dL_dux / dL_dwxx are fully written (off-diagonal entries zero).  Workspace: the
 * pigs_pde_residual_workspace_bytes query above.
 * pigs_vorticity_*: model_pn.py:779-781 `w = ux[:,0,1] - ux[:,1,0]`, `wx = uxx[...,0,1] - uxx[...,1,0]`,
 * `wxx = uxxx[...,0,1] - uxxx[...,1,0]` — over the trailing (spatial axis e, channel c) 2x2 group of the sampler's
 * C = 2 outputs: out[g] = in[g, 0, 1] - in[g, 1, 0]; `groups` = N, 2N, 4N for ux, uxx, uxxx. */
int pigs_pde_ns_residual_forward(int64_t N, float nu, float dt, const float *u, const float *ux, const float *wx,
                                 const float *wxx, const float *wt, float *residual, float *divergence, float *loss,
                                 void *workspace, void *stream);
int pigs_pde_ns_residual_backward(int64_t N, float nu, float dt, const float *u, const float *ux, const float *wx,
                                  const float *wxx, const float *wt, const float *dL_dloss, float *dL_du, float *dL_dux,
                                  float *dL_dwx, float *dL_dwxx, float *dL_dwt, void *stream);
/* Time integration fused into the residual (model_pn.py:794-811 TRAPEZOID / FORWARD / BACKWARD blends, :818-820 ut):
 * the two time levels [-1] (suffix 1) and [-2] (suffix 0) of the sampler's outputs go in as they are and the kernels form
 *   x = ts*x1 + (1 - ts)*x0  per sample (ts [N]; ts == NULL: the constant ts_const, 0 = FORWARD rule, 1 = BACKWARD rule)
 *   ut = u1 - u0  (NAVIER_STOKES: wt = w1 - w0)
 * themselves, fp32 operation by operation like the torch lines.  lap1 / lap0 are the traces (PIGS_FN_LAPL) of the two
 * Hessians (blending commutes with the trace up to rounding).  backward: one gradient per time level and input; any
 * output pointer may be NULL. */
int pigs_pde_residual_timelevels_forward(int problem, int64_t N, int D, int C, float nu, float dt, const float *ts,
                                         float ts_const, const float *u1, const float *u0, const float *ux1,
                                         const float *ux0, const float *lap1, const float *lap0, float *residual,
                                         float *loss, void *workspace, void *stream);
int pigs_pde_residual_timelevels_backward(int problem, int64_t N, int D, int C, float nu, float dt, const float *ts,
                                          float ts_const, const float *u1, const float *u0, const float *ux1,
                                          const float *ux0, const float *lap1, const float *lap0, const float *dL_dloss,
                                          float *dL_du1, float *dL_du0, float *dL_dux1, float *dL_dux0, float *dL_dlap1,
                                          float *dL_dlap0, void *stream);
int pigs_pde_ns_residual_timelevels_forward(int64_t N, float nu, float dt, const float *ts, float ts_const, const float *u1,
                                            const float *u0, const float *ux1, const float *ux0, const float *wx1,
                                            const float *wx0, const float *wxx1, const float *wxx0, const float *w1,
                                            const float *w0, float *residual, float *divergence, float *loss,
                                            void *workspace, void *stream);
int pigs_pde_ns_residual_timelevels_backward(int64_t N, float nu, float dt, const float *ts, float ts_const, const float *u1,
                                             const float *u0, const float *ux1, const float *ux0, const float *wx1,
                                             const float *wx0, const float *wxx1, const float *wxx0, const float *w1,
                                             const float *w0, const float *dL_dloss, float *dL_du1, float *dL_du0,
                                             float *dL_dux1, float *dL_dux0, float *dL_dwx1, float *dL_dwx0, float *dL_dwxx1,
                                             float *dL_dwxx0, float *dL_dw1, float *dL_dw0, void *stream);
int pigs_vorticity_forward(int64_t groups, const float *in, float *out, void *stream);
int pigs_vorticity_backward(int64_t groups, const float *dL_dout, float *dL_din, void *stream);

/* ---- dense all-pairs path -----------------------------------------------------
 * gaussians.py:48-58 `sample_gaussians`, :89-101 `gaussian_derivative`, :103-116 `gaussian_derivative2`: every Gaussian
 * against every sample, x = sample - mean, FULL conics [n, D, D] (not assumed symmetric), no tile cutoff, no torus
 * wrap — the pure-PyTorch path of the reference's CPU drivers, without its [N, n, d, ...] broadcast tensors.
 * fn_mask: PIGS_FN_VALUE | PIGS_FN_DERIV | PIGS_FN_HESS (any subset, one launch).  Outputs u [N,C], ux [N,D,C],
 * uxx [N,D,D,C]; D in {1,2}, C <= 4.  backward = torch.autograd of those functions: dL_dmeans [n,D], dL_dconics
 * [n,D,D], dL_dvalues [n,C] (fully written; the dL_d<out> of functions not in fn_mask may be NULL). */
/* Test knob: the kernels hold 1-4 samples (forward) / 1-2 Gaussians (backward) per thread, chosen from the problem size;
 * 1 forces one, > 1 forces the most the channel count allows, 0 (default) restores the choice. */
void pigs_set_dense_slots(int samples_per_thread, int gaussians_per_thread);
int pigs_dense_forward(unsigned fn_mask, int64_t n, int D, int C, int64_t N, const float *means, const float *conics,
                       const float *values, const float *samples, float *out_u, float *out_ux, float *out_uxx,
                       void *stream);
int pigs_dense_backward(unsigned fn_mask, int64_t n, int D, int C, int64_t N, const float *means, const float *conics,
                        const float *values, const float *samples, const float *dL_du, const float *dL_dux,
                        const float *dL_duxx, float *dL_dmeans, float *dL_dconics, float *dL_dvalues, void *stream);

/* ---- split / prune as a device-side compaction --------------------------------
 * Replaces the torch chains that change P between steps: model_pn.py:578-610 `Model.split`, :703-716 (prune by
 * keep_indices), test_no_mlp_1d.py:203-247 (keep + displaced clones, Adam moments of the clones zeroed).
 * Output rows = [rows with keep != 0, in order | m appended rows per row with extend != 0, in order, child 0..m-1].
 * keep == NULL keeps every row, extend == NULL appends nothing.
 *   pigs_compact_count : flags + two prefix sums; the first 16 bytes of `workspace` then hold int64 {n_keep, n_extend}
 *                        (device memory; the caller reads them back once to size the outputs)
 *   pigs_compact_fill  : src[Pout] = parent row of every output row, Pout = n_keep + m * n_extend
 *   pigs_gather_rows   : out[i, :] = in[src[i], :] * (i < n_keep ? 1 : appended_scale)   (W floats per row;
 *                        appended_scale 0.5 = `u / 2` of Model.split, 0 = fresh Adam moments, 1 = plain copy)
 *   pigs_split_means   : Model.split's means: kept rows copied, the two children of a split row at
 *                        mu -/+ lambda_max * v_max of the parent's covariance (flattened [xx, xy, yy]; 1-D: [s]).
 *                        The reference's order of the two children follows LAPACK's eigenvector sign; here child 0 is
 *                        the minus side of v = normalised (lambda - c2, c1) | (c1, lambda - c0). */
size_t pigs_compact_workspace_bytes(int64_t P);
int pigs_compact_count(int64_t P, const uint8_t *keep, const uint8_t *extend, void *workspace, void *stream);
int pigs_compact_fill(int64_t P, const uint8_t *keep, const uint8_t *extend, int m, const void *workspace,
                      int64_t n_keep, int64_t *src, void *stream);
int pigs_gather_rows(int64_t Pout, int W, int64_t n_keep, const int64_t *src, const float *in, float *out,
                     float appended_scale, void *stream);
int pigs_split_means(int64_t Pout, int D, int64_t n_keep, const int64_t *src, const float *means,
                     const float *covariances, float *out_means, void *stream);

/* ---- roofline microbenchmarks (bench.py): register-resident FFMA / MUFU.EX2 loops ---------------
 * Each launches `ctas` CTAs x 256 threads running `iters` iterations of 64 independent ops per thread
 * and returns the number of lane-ops executed. Timing is done by the caller with CUDA events. */
int64_t pigs_microbench_ffma(int ctas, int iters, float *sink, void *stream);
int64_t pigs_microbench_ex2(int ctas, int iters, float *sink, void *stream);
/* packed fma.rn.f32x2 (FFMA2) loop; mix != 0 adds 2 MUFU.EX2 per 32 lane-FMAs.  Returns lane-FMAs executed. */
int64_t pigs_microbench_ffma2(int ctas, int iters, int mix, float *sink, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* PIGS_B200_H */
