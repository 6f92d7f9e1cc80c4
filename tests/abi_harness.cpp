// TEST INFRASTRUCTURE — a compiled, torch-free caller of the C ABI (include/pigs_b200.h): the host side a maintainer
// of the reference's ext.cpp / sample_points.cu would write reduced to cudaMalloc + the library
// calls.  Reads one binary problem file, runs preprocess -> fused forward (u, grad u, Hessian) -> fused backward, writes
// every result to a binary file; tests/test_abi_harness.py compares that file with the CPU oracle.
//   in : int64 P, D, N, C ; f32 means[P*D] cov[P*T] conics[P*T] values[P*C] samples[N*D] gu[N*C] gdu[N*D*C] gddu[N*D*D*C]
//   out: int64 num_rendered, blocks ; f32 radii[P] ; u32 ranges[2*(blocks+1)] sample_ranges[2*(blocks+1)]
//        point_list[R] sample_list[N] ; f32 u[N*C] du[N*D*C] ddu[N*D*D*C] dmeans[P*D] dvalues[P*C] dconics[P*T]
// (T = D(D+1)/2).  usage: abi_harness <in> <out>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "pigs_b200.h"

#define CK(call)                                                                           \
    do {                                                                                   \
        int _rc = (call);                                                                  \
        if (_rc != 0) {                                                                    \
            std::fprintf(stderr, "%s -> %d: %s\n", #call, _rc, pigs_last_error());         \
            return 2;                                                                      \
        }                                                                                  \
    } while (0)
#define CU(call)                                                                           \
    do {                                                                                   \
        cudaError_t _e = (call);                                                           \
        if (_e != cudaSuccess) {                                                           \
            std::fprintf(stderr, "%s -> %s\n", #call, cudaGetErrorString(_e));             \
            return 3;                                                                      \
        }                                                                                  \
    } while (0)

template <class T>
static T* dev(size_t n)
{
    void* p = nullptr;
    if (cudaMalloc(&p, (n ? n : 1) * sizeof(T)) != cudaSuccess) std::abort();
    cudaMemset(p, 0, (n ? n : 1) * sizeof(T));
    return (T*)p;
}

template <class T>
static T* up(const std::vector<T>& h)
{
    T* d = dev<T>(h.size());
    cudaMemcpy(d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice);
    return d;
}

static std::vector<float> rd(FILE* f, size_t n)
{
    std::vector<float> v(n);
    if (n && std::fread(v.data(), sizeof(float), n, f) != n) std::abort();
    return v;
}

template <class T>
static void wr(FILE* f, const T* d, size_t n)
{
    std::vector<T> h(n);
    cudaMemcpy(h.data(), d, n * sizeof(T), cudaMemcpyDeviceToHost);
    std::fwrite(h.data(), sizeof(T), n, f);
}

int main(int argc, char** argv)
{
    if (argc != 3) return 1;
    FILE* fi = std::fopen(argv[1], "rb");
    if (!fi) return 1;
    int64_t dims[4];
    if (std::fread(dims, sizeof(int64_t), 4, fi) != 4) return 1;
    const int64_t P = dims[0], N = dims[2];
    const int D = (int)dims[1], C = (int)dims[3], T = D * (D + 1) / 2;
    float* means = up(rd(fi, P * D));
    float* cov = up(rd(fi, P * T));
    float* conics = up(rd(fi, P * T));
    float* values = up(rd(fi, P * C));
    float* samples = up(rd(fi, N * D));
    float* gu = up(rd(fi, N * C));
    float* gdu = up(rd(fi, N * D * C));
    float* gddu = up(rd(fi, N * D * D * C));
    std::fclose(fi);
    if (pigs_version() < 100) return 1;
    cudaStream_t st;
    CU(cudaStreamCreate(&st));

    // ---- preprocess (sample_points.cu:37-98): bounds -> count -> ONE header readback -> bin
    float* mm = dev<float>(2 * D);
    CK(pigs_sample_bounds(samples, N, D, mm, st));
    float* radii = dev<float>(P);
    void* geom = dev<uint8_t>(pigs_geom_workspace_bytes(P));
    CK(pigs_preprocess_count(P, D, N, means, cov, mm, radii, geom, st));
    pigs_header h;
    CU(cudaMemcpyAsync(&h, geom, sizeof(h), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const int64_t R = h.num_rendered;
    const int blocks = h.blocks;
    void* binning = dev<uint8_t>(pigs_binning_bytes(R, blocks));
    void* sbinning = dev<uint8_t>(pigs_sample_binning_bytes(N, blocks, D));
    uint32_t* ranges = dev<uint32_t>(2 * (blocks + 1));
    uint32_t* sranges = dev<uint32_t>(2 * (blocks + 1));
    void* sort_ws = dev<uint8_t>(pigs_sort_workspace_bytes(R, N));
    CK(pigs_preprocess_bin(P, D, N, blocks, R, means, conics, samples, radii, geom, binning, sbinning, ranges, sranges, sort_ws, st));

    // ---- fused forward + backward (sample_points.cu:100-372 as one call each way); outputs zero-initialised
    const unsigned mask = PIGS_FN_VALUE | PIGS_FN_DERIV | PIGS_FN_HESS;
    const size_t wsb = pigs_eval_workspace_bytes(R, N, blocks, D, C, mask);
    void* ws = dev<uint8_t>(wsb);
    float *u = dev<float>(N * C), *du = dev<float>(N * D * C), *ddu = dev<float>(N * D * D * C);
    CK(pigs_sample_forward(mask, P, D, N, C, blocks, R, means, values, conics, samples, binning, sbinning, ranges, sranges, u,
                           du, ddu, nullptr, ws, wsb, st));
    float *dm = dev<float>(P * D), *dv = dev<float>(P * C), *dc = dev<float>(P * T);
    CK(pigs_sample_backward(mask, P, D, N, C, blocks, R, means, values, conics, samples, binning, sbinning, ranges, sranges,
                            gu, gdu, gddu, nullptr, dm, dv, dc, ws, wsb, st));
    CU(cudaStreamSynchronize(st));
    if (pigs_launch_count() <= 0) return 4;

    FILE* fo = std::fopen(argv[2], "wb");
    if (!fo) return 1;
    const int64_t head[2] = {R, blocks};
    std::fwrite(head, sizeof(int64_t), 2, fo);
    wr(fo, radii, P);
    wr(fo, ranges, 2 * (blocks + 1));
    wr(fo, sranges, 2 * (blocks + 1));
    wr(fo, pigs_point_list(binning), R);
    wr(fo, pigs_sample_list(sbinning), N);
    wr(fo, u, N * C);
    wr(fo, du, N * D * C);
    wr(fo, ddu, N * D * D * C);
    wr(fo, dm, P * D);
    wr(fo, dv, P * C);
    wr(fo, dc, P * T);
    std::fclose(fo);
    std::printf("abi_harness ok: P=%lld N=%lld num_rendered=%lld blocks=%d launches=%lld\n", (long long)P, (long long)N,
                (long long)R, blocks, (long long)pigs_launch_count());
    return 0;
}
