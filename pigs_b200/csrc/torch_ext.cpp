// torch_ext.cpp — the compiled torch binding of libpigs_b200.so: the reference's 12 pybind entry points
// (diff-gaussian-sampling/ext.cpp:19-32; signatures sample_points.h:20-131, aggregate_neighbors.h:11-47), same names,
// argument order and return values, plus the fused forms the Python layer uses (sample_forward / sample_backward with a
// function mask, preprocess with global bounds / culling).  It is GLUE ONLY: tensor checks, torch allocations, pointer
// extraction, one C-ABI call (include/pigs_b200.h) per kernel group on torch's current CUDA stream.  No kernel lives here
// and no torch type crosses into the library.
//
// Against the ctypes binding (pigs_b200/_C.py, kept as the torch-free-build alternative; PIGS_B200_BINDING=ctypes selects
// it) this saves the per-call Python marshalling — about 20 pointer conversions and a ctypes transition per entry point —
// and keeps the per-call scratch (geometry / sort / evaluation workspaces) in grow-only buffers per (device, stream)
// instead of allocating them on every call.  Built by pigs_b200/_lib.py build_ext() into pigs_b200/_Cext.so.
#include <torch/extension.h>

#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <cuda_runtime_api.h>

#include <map>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/pigs_b200.h"

namespace {

using torch::Tensor;
using OptTensor = c10::optional<Tensor>;

void check(int rc, const char* what)
{
    if (rc != 0) {
        const char* msg = pigs_last_error();
        throw std::runtime_error(std::string("pigs_b200: ") + what + " failed (status " + std::to_string(rc) + "): " + (msg ? msg : ""));
    }
}

// fp32 + CUDA asserted; non-contiguous inputs are copied like the reference's `.contiguous()` (sample_points.cu:127-131)
Tensor f32c(const Tensor& t, const char* name)
{
    TORCH_CHECK(t.scalar_type() == torch::kFloat32, name, ": expected scalar type Float but found ", t.scalar_type());
    TORCH_CHECK(t.is_cuda(), name, " must be a CUDA tensor (pigs_b200 has no CPU path)");
    return t.contiguous();
}
const float* fp(const Tensor& t) { return t.data_ptr<float>(); }
float* fpw(Tensor& t) { return t.data_ptr<float>(); }
const float* fpo(const OptTensor& t) { return t.has_value() ? t->data_ptr<float>() : nullptr; }
void* cur_stream() { return (void*)at::cuda::getCurrentCUDAStream().stream(); }
Tensor new_bytes(int64_t n, const Tensor& like) { return torch::empty({n}, like.options().dtype(torch::kUInt8)); }

// grow-only scratch per (device, stream, slot).  Consecutive library calls on one stream are ordered on the device, so a
// buffer a finished call used can be handed to the next one without synchronisation; different streams get different buffers.
enum Slot { SLOT_GEOM = 0, SLOT_SORT = 1, SLOT_EVAL = 2, SLOT_NB = 3, SLOT_SCAN = 4 };
Tensor scratch(int slot, int64_t n, const Tensor& like)
{
    static std::map<std::tuple<int, void*, int>, Tensor> pool;
    auto key = std::make_tuple((int)like.get_device(), cur_stream(), slot);
    auto it = pool.find(key);
    if (it == pool.end() || it->second.numel() < n) {
        Tensor t = new_bytes(n + n / 8 + 256, like);
        pool[key] = t;
        return t;
    }
    return it->second;
}

// one 256-byte header / one int64 through pinned memory: the single host synchronisation of a preprocess call
Tensor pinned(int64_t n)
{
    static std::map<std::pair<int, void*>, Tensor> pool;
    int dev = 0;
    cudaGetDevice(&dev);
    auto key = std::make_pair(dev, cur_stream());
    auto it = pool.find(key);
    if (it == pool.end()) {
        Tensor t = torch::empty({512}, torch::TensorOptions().dtype(torch::kUInt8).pinned_memory(true));
        pool[key] = t;
        return t.narrow(0, 0, n);
    }
    return it->second.narrow(0, 0, n);
}
void read_back(void* dst_host, const void* src_dev, size_t n)
{
    auto st = at::cuda::getCurrentCUDAStream().stream();
    C10_CUDA_CHECK(cudaMemcpyAsync(dst_host, src_dev, n, cudaMemcpyDeviceToHost, st));
    C10_CUDA_CHECK(cudaStreamSynchronize(st));
}

struct Dims {
    int64_t P, N;
    int D, C;
};
Dims dims(const Tensor& means, const Tensor& values, const Tensor& samples)
{
    return Dims{means.size(0), samples.size(0), (int)means.size(-1), (int)values.size(-1)};
}
int blocks_of(const Tensor& ranges) { return (int)(ranges.numel() / 8 - 1); }

// ---- PreprocessCUDA (sample_points.cu:37-98) ----------------------------------------------------------------------
std::tuple<int64_t, Tensor, Tensor, Tensor, Tensor, Tensor> preprocess_gaussians(const Tensor& means_, const Tensor& values,
                                                                                  const Tensor& covariances, const Tensor& conics_,
                                                                                  const Tensor& samples_, bool debug,
                                                                                  const OptTensor& bounds,
                                                                                  const c10::optional<std::tuple<double, int64_t>>& cull)
{
    Tensor means = f32c(means_, "means");
    const c10::cuda::CUDAGuard guard(means.device());
    pigs_set_debug(debug ? 1 : 0);
    const Dims d = dims(means, values, samples_);
    Tensor radii = torch::zeros({d.P}, means.options());
    Tensor empty = torch::empty({0}, means.options().dtype(torch::kUInt8));
    if (d.P == 0 || d.N == 0) return std::make_tuple((int64_t)0, empty, empty.clone(), empty.clone(), empty.clone(), radii);
    TORCH_CHECK(d.D == 1 || d.D == 2, "pigs_b200 supports 1-D and 2-D mixtures (got D=", d.D, "), like the reference kernels");
    Tensor cov = f32c(covariances, "covariances");
    Tensor samples = f32c(samples_, "samples");
    Tensor conics = f32c(conics_, "conics");
    f32c(values, "values");
    TORCH_CHECK(cov.numel() == d.P * d.D * (d.D + 1) / 2, "covariances must hold P*D*(D+1)/2 elements");
    void* st = cur_stream();
    Tensor mm = bounds.has_value() ? f32c(*bounds, "bounds") : torch::empty({2 * d.D}, means.options());
    Tensor geom = scratch(SLOT_GEOM, (int64_t)pigs_geom_workspace_bytes(d.P), means);
    // bounds + tile grid + radii / tiles touched + scan: one single-CTA launch for small problems, the grid-wide chain else
    check(pigs_preprocess_count_fused(d.P, d.D, d.N, fp(means), fp(cov), fp(samples), fpw(mm), bounds.has_value() ? 1 : 0,
                                      fpw(radii), geom.data_ptr(), cull.has_value() ? (int)std::get<1>(*cull) : 1,
                                      cull.has_value() ? (float)std::get<0>(*cull) : 0.0f, st),
          "pigs_preprocess_count_fused");
    Tensor host = pinned(sizeof(pigs_header));
    read_back(host.data_ptr(), geom.data_ptr(), sizeof(pigs_header));
    const pigs_header* h = (const pigs_header*)host.data_ptr();
    const int blocks = h->blocks;
    const int64_t R = h->num_rendered;
    Tensor binning = new_bytes((int64_t)pigs_binning_bytes(R, blocks), means);
    Tensor sbinning = new_bytes((int64_t)pigs_sample_binning_bytes(d.N, blocks, d.D), means);
    Tensor ranges = new_bytes(blocks * 8 + 8, means);
    Tensor sample_ranges = new_bytes(blocks * 8 + 8, means);
    Tensor sort_ws = scratch(SLOT_SORT, (int64_t)pigs_sort_workspace_bytes(R, d.N), means);
    check(pigs_preprocess_bin(d.P, d.D, d.N, blocks, R, fp(means), fp(conics), fp(samples), fp(radii), geom.data_ptr(),
                              binning.data_ptr(), sbinning.data_ptr(), ranges.data_ptr(), sample_ranges.data_ptr(),
                              sort_ws.data_ptr(), st),
          "pigs_preprocess_bin");
    return std::make_tuple(R, binning, sbinning, ranges, sample_ranges, radii);
}

// ---- fused forward / backward (SampleGaussians*CUDA, sample_points.cu:100-372, any subset in one call) ------------
// Zero-initialised tensors of the given shapes (an empty shape = not wanted).  Small problems are launch-bound, and every
// torch::zeros is a fill kernel: up to 1 MB in total the tensors are views of ONE zeroed buffer (segments 256-byte
// aligned), beyond that separate allocations — a caller that keeps only one of several large outputs must not pin the
// others' memory.
void zeros_many(const std::vector<std::vector<int64_t>>& shapes, const Tensor& like, std::vector<OptTensor>& out)
{
    constexpr int64_t ALIGN = 64, FLAT_MAX = 1 << 18;  // floats
    int64_t total = 0, wanted = 0;
    std::vector<int64_t> off(shapes.size(), 0), cnt(shapes.size(), 0);
    for (size_t k = 0; k < shapes.size(); k++) {
        if (shapes[k].empty()) continue;
        int64_t n = 1;
        for (int64_t v : shapes[k]) n *= v;
        off[k] = total;
        cnt[k] = n;
        total += (n + ALIGN - 1) / ALIGN * ALIGN;
        wanted++;
    }
    const bool flat = wanted > 1 && total <= FLAT_MAX;
    Tensor base;
    if (flat) base = torch::zeros({total}, like.options());
    for (size_t k = 0; k < shapes.size(); k++) {
        if (shapes[k].empty()) continue;
        out[k] = flat ? base.narrow(0, off[k], cnt[k]).view(shapes[k]) : torch::zeros(shapes[k], like.options());
    }
}

std::vector<OptTensor> sample_forward(int64_t fn_mask, const Tensor& means_, const Tensor& values_, const Tensor& conics_,
                                      const Tensor& samples_, int64_t num_rendered, const Tensor& binning_buffer,
                                      const Tensor& sample_binning_buffer, const Tensor& ranges, const Tensor& sample_ranges,
                                      bool debug)
{
    Tensor means = f32c(means_, "means"), values = f32c(values_, "values"), conics = f32c(conics_, "conics"),
           samples = f32c(samples_, "samples");
    const c10::cuda::CUDAGuard guard(means.device());
    pigs_set_debug(debug ? 1 : 0);
    const Dims d = dims(means, values, samples);
    TORCH_CHECK(!((fn_mask & PIGS_FN_LAPL) && (fn_mask & PIGS_FN_HESS)), "FN_LAPL and FN_HESS are mutually exclusive (they share an output slot)");
    std::vector<OptTensor> outs(4);
    std::vector<std::vector<int64_t>> shapes(4);
    for (int k = 0; k < 4; k++) {
        if (!(fn_mask & (1 << k))) continue;
        shapes[k] = {d.N};
        for (int i = 0; i < k; i++) shapes[k].push_back(d.D);
        shapes[k].push_back(d.C);
    }
    if (fn_mask & PIGS_FN_LAPL) shapes[2] = {d.N, (int64_t)d.C};
    zeros_many(shapes, means, outs);
    if (d.P != 0 && d.N != 0) {
        const int blocks = blocks_of(ranges);
        const size_t wsb = pigs_eval_workspace_bytes(num_rendered, d.N, blocks, d.D, d.C, (unsigned)fn_mask);
        Tensor ws = scratch(SLOT_EVAL, (int64_t)wsb, means);
        auto op = [](OptTensor& t) -> float* { return t.has_value() ? t->data_ptr<float>() : nullptr; };
        check(pigs_sample_forward((unsigned)fn_mask, d.P, d.D, d.N, d.C, blocks, num_rendered, fp(means), fp(values), fp(conics),
                                  fp(samples), binning_buffer.data_ptr(), sample_binning_buffer.data_ptr(), ranges.data_ptr(),
                                  sample_ranges.data_ptr(), op(outs[0]), op(outs[1]), op(outs[2]), op(outs[3]), ws.data_ptr(), wsb,
                                  cur_stream()),
              "pigs_sample_forward");
    }
    return outs;
}

std::tuple<Tensor, Tensor, Tensor> sample_backward(int64_t fn_mask, const Tensor& means_, const Tensor& values_,
                                                   const Tensor& conics_, const Tensor& samples_, int64_t num_rendered,
                                                   const std::vector<OptTensor>& grads, const Tensor& binning_buffer,
                                                   const Tensor& sample_binning_buffer, const Tensor& ranges,
                                                   const Tensor& sample_ranges, bool debug)
{
    Tensor means = f32c(means_, "means"), values = f32c(values_, "values"), conics = f32c(conics_, "conics"),
           samples = f32c(samples_, "samples");
    const c10::cuda::CUDAGuard guard(means.device());
    pigs_set_debug(debug ? 1 : 0);
    const Dims d = dims(means, values, samples);
    TORCH_CHECK(grads.size() == 4, "grads must be a 4-tuple (value, derivative, Hessian | Laplacian, third)");
    // one flat buffer, three contiguous views: a multi-GPU caller all-reduces `flat` with one collective
    const int64_t K = d.D * (d.D + 1) / 2;
    Tensor flat = torch::zeros({d.P * (d.D + d.C + K)}, means.options());
    Tensor dm = flat.narrow(0, 0, d.P * d.D).view({d.P, (int64_t)d.D});
    Tensor dv = flat.narrow(0, d.P * d.D, d.P * d.C).view({d.P, (int64_t)d.C});
    Tensor dc = flat.narrow(0, d.P * (d.D + d.C), d.P * K).view({d.P, K});
    std::vector<OptTensor> g(4);
    for (int k = 0; k < 4; k++)
        if (grads[k].has_value()) g[k] = f32c(*grads[k], "dL_dout");
    if (d.P != 0 && d.N != 0) {
        const int blocks = blocks_of(ranges);
        const size_t wsb = pigs_eval_workspace_bytes(num_rendered, d.N, blocks, d.D, d.C, (unsigned)fn_mask);
        Tensor ws = scratch(SLOT_EVAL, (int64_t)wsb, means);
        check(pigs_sample_backward((unsigned)fn_mask, d.P, d.D, d.N, d.C, blocks, num_rendered, fp(means), fp(values), fp(conics),
                                   fp(samples), binning_buffer.data_ptr(), sample_binning_buffer.data_ptr(), ranges.data_ptr(),
                                   sample_ranges.data_ptr(), fpo(g[0]), fpo(g[1]), fpo(g[2]), fpo(g[3]), fpw(dm), fpw(dv), fpw(dc),
                                   ws.data_ptr(), wsb, cur_stream()),
              "pigs_sample_backward");
    }
    return std::make_tuple(dm, dv, dc);
}

// the reference's eight single-function entry points (ext.cpp:21-28)
template <int BIT, int SLOT>
Tensor single_forward(const Tensor& means, const Tensor& values, const Tensor& conics, const Tensor& samples, int64_t num_rendered,
                      const Tensor& bb, const Tensor& sbb, const Tensor& ranges, const Tensor& sample_ranges, bool debug)
{
    return *sample_forward(BIT, means, values, conics, samples, num_rendered, bb, sbb, ranges, sample_ranges, debug)[SLOT];
}
template <int BIT, int SLOT>
std::tuple<Tensor, Tensor, Tensor> single_backward(const Tensor& means, const Tensor& values, const Tensor& conics,
                                                   const Tensor& samples, int64_t num_rendered, const Tensor& dL_dout,
                                                   const Tensor& bb, const Tensor& sbb, const Tensor& ranges,
                                                   const Tensor& sample_ranges, bool debug)
{
    std::vector<OptTensor> g(4);
    g[SLOT] = dL_dout;
    return sample_backward(BIT, means, values, conics, samples, num_rendered, g, bb, sbb, ranges, sample_ranges, debug);
}

// ---- neighbour aggregation (aggregate_neighbors.cu:323-475) --------------------------------------------------------
std::tuple<Tensor, Tensor, Tensor, Tensor, Tensor> preprocess_aggregate(const Tensor& means_, const Tensor& conics_,
                                                                         const Tensor& radii_, bool debug,
                                                                         const c10::optional<std::tuple<int64_t, int64_t>>& rows)
{
    Tensor means = f32c(means_, "means"), conics = f32c(conics_, "conics"), radii = f32c(radii_, "radii");
    const c10::cuda::CUDAGuard guard(means.device());
    pigs_set_debug(debug ? 1 : 0);
    const int64_t P = means.size(0);
    const int D = (int)means.size(-1);
    const int64_t lo = rows.has_value() ? std::get<0>(*rows) : 0, hi = rows.has_value() ? std::get<1>(*rows) : P;
    void* st = cur_stream();
    auto i64 = means.options().dtype(torch::kInt64);
    Tensor counts = torch::zeros({P}, i64), ranges = torch::empty({P}, i64), total = torch::zeros({1}, i64);
    int64_t length = 0;
    Tensor ws = scratch(SLOT_NB, std::max<int64_t>((int64_t)pigs_neighbor_workspace_bytes(P, D), 16), means);
    if (P > 0) {
        check(pigs_neighbors_count_rows(P, D, fp(means), fp(radii), lo, hi, counts.data_ptr<int64_t>(), ws.data_ptr(), st),
              "pigs_neighbors_count_rows");
        Tensor sws = scratch(SLOT_SCAN, std::max<int64_t>((int64_t)pigs_neighbors_scan_workspace_bytes(P), 16), means);
        check(pigs_neighbors_ranges(P, counts.data_ptr<int64_t>(), ranges.data_ptr<int64_t>(), total.data_ptr<int64_t>(),
                                    sws.data_ptr(), st),
              "pigs_neighbors_ranges");
        Tensor host = pinned(8);
        read_back(host.data_ptr(), total.data_ptr(), 8);
        length = *(const int64_t*)host.data_ptr();
    }
    // every CSR slot is written by the fill kernels (skipped neighbours get index -1 / density 0 there)
    Tensor indices = torch::empty({length}, i64);
    Tensor dists = torch::empty({length, (int64_t)D}, means.options());
    Tensor densities = torch::empty({length}, means.options());
    Tensor inv_total = torch::zeros({P}, means.options());
    if (P > 0)
        check(pigs_neighbors_fill_rows(P, D, fp(means), fp(conics), fp(radii), lo, hi, ranges.data_ptr<int64_t>(),
                                       indices.data_ptr<int64_t>(), fpw(dists), fpw(densities), fpw(inv_total), ws.data_ptr(), st),
              "pigs_neighbors_fill_rows");
    return std::make_tuple(indices, ranges, dists, densities, inv_total);
}

std::tuple<Tensor, Tensor, Tensor, Tensor> aggregate_neighbors(const Tensor& features_, const Tensor& transform_,
                                                               const Tensor& queries_, const Tensor& keys_,
                                                               const Tensor& frequencies_, const Tensor& distance_transform_,
                                                               const Tensor& indices_, const Tensor& ranges_, const Tensor& dists_,
                                                               const Tensor& densities_, const Tensor& inv_total_, bool debug)
{
    Tensor features = f32c(features_, "features"), transform = f32c(transform_, "transform"), queries = f32c(queries_, "queries"),
           keys = f32c(keys_, "keys"), freqs = f32c(frequencies_, "frequencies"), dt = f32c(distance_transform_, "distance_transform"),
           dists = f32c(dists_, "dists"), densities = f32c(densities_, "densities"), inv_total = f32c(inv_total_, "inv_total_densities");
    const c10::cuda::CUDAGuard guard(features.device());
    pigs_set_debug(debug ? 1 : 0);
    Tensor indices = indices_.contiguous(), ranges = ranges_.contiguous();
    const int64_t P = features.size(0);
    const int D = (int)dists.size(-1), L = (int)features.size(-1), K = (int)queries.size(-1), E = (int)(dt.size(-1) / 2);
    // no zero fills: the kernels write every slot of every row (skipped neighbours as zeros) and every output row
    Tensor weights = torch::empty_like(densities), embeddings = torch::empty_like(densities), factors = torch::empty_like(densities);
    Tensor out = torch::empty({P, (int64_t)L}, features.options());
    if (P > 0)
        check(pigs_aggregate_forward(P, D, L, K, E, fp(features), fp(transform), fp(queries), fp(keys), fp(freqs), fp(dt),
                                     indices.data_ptr<int64_t>(), ranges.data_ptr<int64_t>(), fp(dists), fp(densities), fp(inv_total),
                                     fpw(weights), fpw(embeddings), fpw(factors), fpw(out), cur_stream()),
              "pigs_aggregate_forward");
    return std::make_tuple(weights, embeddings, factors, out);
}

std::tuple<Tensor, Tensor, Tensor, Tensor, Tensor, Tensor> aggregate_neighbors_backward(
    const Tensor& features_, const Tensor& transform_, const Tensor& queries_, const Tensor& keys_, const Tensor& frequencies_,
    const Tensor& distance_transform_, const Tensor& indices_, const Tensor& ranges_, const Tensor& dists_, const Tensor& densities_,
    const Tensor& weights_, const Tensor& embeddings_, const Tensor& factors_, const Tensor& inv_total_, const Tensor& dL_dout_,
    bool debug)
{
    Tensor features = f32c(features_, "features"), transform = f32c(transform_, "transform"), queries = f32c(queries_, "queries"),
           keys = f32c(keys_, "keys"), freqs = f32c(frequencies_, "frequencies"), dt = f32c(distance_transform_, "distance_transform"),
           dists = f32c(dists_, "dists"), densities = f32c(densities_, "densities"), weights = f32c(weights_, "weights"),
           embeddings = f32c(embeddings_, "embeddings"), factors = f32c(factors_, "factors"),
           inv_total = f32c(inv_total_, "inv_total_densities"), g = f32c(dL_dout_, "dL_dneighbor_features");
    const c10::cuda::CUDAGuard guard(features.device());
    pigs_set_debug(debug ? 1 : 0);
    Tensor indices = indices_.contiguous(), ranges = ranges_.contiguous();
    const int64_t P = features.size(0);
    const int D = (int)dists.size(-1), L = (int)features.size(-1), K = (int)queries.size(-1), E = (int)(dt.size(-1) / 2);
    std::vector<OptTensor> z(6);
    zeros_many({features.sizes().vec(), transform.sizes().vec(), queries.sizes().vec(), keys.sizes().vec(), freqs.sizes().vec(),
                dt.sizes().vec()}, features, z);
    Tensor gf = *z[0], gt = *z[1], gq = *z[2], gk = *z[3], gfr = *z[4], gdt = *z[5];
    if (P > 0)
        check(pigs_aggregate_backward(P, D, L, K, E, densities.numel(), fp(features), fp(transform), fp(queries), fp(keys), fp(freqs),
                                      fp(dt), indices.data_ptr<int64_t>(), ranges.data_ptr<int64_t>(), fp(dists), fp(densities),
                                      fp(weights), fp(embeddings), fp(factors), fp(inv_total), fp(g), fpw(gf), fpw(gt), fpw(gq),
                                      fpw(gk), fpw(gfr), fpw(gdt), nullptr, cur_stream()),
              "pigs_aggregate_backward");
    return std::make_tuple(gf, gt, gq, gk, gfr, gdt);
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m)
{
    namespace py = pybind11;
    m.doc() = "compiled torch binding of libpigs_b200.so (reference entry points: diff-gaussian-sampling/ext.cpp:19-32)";
    // the reference's 12 (ext.cpp:20-31); the two extension keywords default to the reference behaviour
    m.def("preprocess_gaussians", &preprocess_gaussians, py::arg("means"), py::arg("values"), py::arg("covariances"), py::arg("conics"),
          py::arg("samples"), py::arg("debug"), py::arg("bounds") = py::none(), py::arg("cull") = py::none());
    m.def("sample_gaussians", &single_forward<PIGS_FN_VALUE, 0>);
    m.def("sample_gaussians_backward", &single_backward<PIGS_FN_VALUE, 0>);
    m.def("sample_gaussians_derivative", &single_forward<PIGS_FN_DERIV, 1>);
    m.def("sample_gaussians_derivative_backward", &single_backward<PIGS_FN_DERIV, 1>);
    m.def("sample_gaussians_laplacian", &single_forward<PIGS_FN_HESS, 2>);
    m.def("sample_gaussians_laplacian_backward", &single_backward<PIGS_FN_HESS, 2>);
    m.def("sample_gaussians_third_derivative", &single_forward<PIGS_FN_THIRD, 3>);
    m.def("sample_gaussians_third_derivative_backward", &single_backward<PIGS_FN_THIRD, 3>);
    m.def("preprocess_aggregate", &preprocess_aggregate, py::arg("means"), py::arg("conics"), py::arg("radii"), py::arg("debug"),
          py::arg("rows") = py::none());
    m.def("aggregate_neighbors", &aggregate_neighbors);
    m.def("aggregate_neighbors_backward", &aggregate_neighbors_backward);
    // fused forms (function mask): what GaussianSampler's autograd node calls
    m.def("sample_forward", &sample_forward);
    m.def("sample_backward", &sample_backward);
}
