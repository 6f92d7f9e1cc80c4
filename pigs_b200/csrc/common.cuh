// common.cuh — shared device/host helpers for libpigs_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/pigs_b200.h"

#define PIGS_MAGIC 0x53474950u  // 'PIGS'
#define PIGS_ALIGN 128
static_assert(sizeof(pigs_header) == 256, "pigs_header must be 256 bytes");

namespace pigs {

extern thread_local char g_err[512];
extern int64_t g_launches;
extern int g_recip_mode;
extern int g_deterministic;  // pigs_set_deterministic (pack.cu)
extern int g_debug_sync;  // pigs_set_debug: synchronise + check after EVERY launch (reference auxiliary.h:33-40)

int set_error(int code, const char* what, const char* file, int line);

#define PIGS_CHECK(expr)                                                           \
    do {                                                                           \
        cudaError_t _e = (expr);                                                   \
        if (_e != cudaSuccess) return pigs::set_error((int)_e, #expr, __FILE__, __LINE__); \
    } while (0)

// Every kernel launch of the library goes through this.  With pigs_set_debug(1) it reproduces the reference's
// CHECK_CUDA(A, debug) (cuda_sampler/auxiliary.h:33-40): cudaDeviceSynchronize + error check after EACH launch, so a
// faulting kernel is reported at its own launch site (file:line) instead of at some later call.
#define PIGS_LAUNCH_CHECK()                                                        \
    do {                                                                           \
        pigs::g_launches++;                                                        \
        cudaError_t _e = cudaGetLastError();                                       \
        if (_e != cudaSuccess) return pigs::set_error((int)_e, "kernel launch", __FILE__, __LINE__); \
        if (pigs::g_debug_sync) {                                                  \
            _e = cudaDeviceSynchronize();                                          \
            if (_e != cudaSuccess) return pigs::set_error((int)_e, "kernel execution (debug: synchronised after this launch)", __FILE__, __LINE__); \
        }                                                                          \
    } while (0)

#define PIGS_ARG(cond)                                                             \
    do {                                                                           \
        if (!(cond)) return pigs::set_error(-1, "bad argument: " #cond, __FILE__, __LINE__); \
    } while (0)

static inline size_t align_up(size_t x, size_t a = PIGS_ALIGN) { return (x + a - 1) / a * a; }

// carve typed sub-arrays out of a byte buffer (128-B aligned, like the reference's `obtain`,
// cuda_sampler/sampler_impl.h:25-31, but with our own layouts)
struct Carver {
    char* p;
    size_t used;
    explicit Carver(void* base) : p((char*)base), used(0) {}
    template <typename T>
    T* take(size_t count)
    {
        used = align_up(used);
        T* r = p ? (T*)(p + used) : nullptr;
        used += count * sizeof(T);
        return r;
    }
    size_t total() const { return align_up(used); }
};

int num_sms();

// Optional per-stage device timing of the preprocess chain (tools/chain_timing.py): while on, PIGS_STAGE(st, "name")
// records a cudaEvent on the stream behind the launches issued so far; pigs_stage_timing_read returns the gaps.
extern bool g_stage_timing;
void stage_mark(cudaStream_t st, const char* name);
#define PIGS_STAGE(st, name)                          \
    do {                                              \
        if (pigs::g_stage_timing) pigs::stage_mark(st, name); \
    } while (0)

// ---------------------------------------------------------------------------------------------
// exclusive scan of u32: ONE launch for any n (single-pass, decoupled look-back; one CTA for small n).  out may alias
// in.  out[n] = total if write_total; *total_out (device int64, optional) = total.  ws: scan_ws_bytes(n).
size_t scan_ws_bytes(int64_t n);
int exclusive_scan_u32(const uint32_t* in, uint32_t* out, int64_t n, bool write_total, void* ws, cudaStream_t st,
                       long long* total_out = nullptr);
// exclusive scan of u64 words (two packed 31-bit class counters, pack.cu), same kernel
int exclusive_scan_u64(const unsigned long long* in, unsigned long long* out, int64_t n, bool write_total, void* ws,
                       cudaStream_t st);
// inclusive scan of int64 (neighbour counts -> CSR ranges), same kernel
int inclusive_scan_i64(const int64_t* in, int64_t* out, int64_t n, void* ws, cudaStream_t st, long long* total_out = nullptr);

// stable LSD radix sort of (key u32, value u32) pairs on the low `bits` bits of the key.
// keys_in/vals_in are clobbered when more than one pass is needed (ping-pong); result lands in
// keys_out/vals_out.  ws: sort_ws_bytes(n).
size_t sort_ws_bytes(int64_t n);
// hist_done: the caller's key-producing kernel already wrote the first pass's per-tile digit histogram
//   sort_ghist(ws)[digit * sort_tiles(n) + tile] (tiles of SORT_TILE_ELEMS consecutive keys, digit = key & 0xff)
// ranges / max_slot: when ONE pass sorts the whole key (bits <= 8) the scatter kernel also writes the tile ranges
//   [first, last + 1) of the keys 0..max_slot from the scanned histogram (*ranges_written tells whether it did).
struct SortExtras {
    bool hist_done = false;
    uint2* ranges = nullptr;
    int max_slot = 0;
    bool* ranges_written = nullptr;
};
constexpr int SORT_TILE_ELEMS = 4096;
uint32_t* sort_ghist(void* ws);
int64_t sort_tiles(int64_t n);
int stable_sort_pairs_u32(uint32_t* keys_in, uint32_t* vals_in, uint32_t* keys_out, uint32_t* vals_out, int64_t n,
                          int bits, void* ws, cudaStream_t st, const SortExtras* ex = nullptr);

// ---------------------------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ void atomic_min_float(float* addr, float v)
{
    // works on the raw bit pattern for any mix of signs as long as *addr was initialised to +inf
    if (v >= 0.0f)
        atomicMin((int*)addr, __float_as_int(v));
    else
        atomicMax((unsigned int*)addr, __float_as_uint(v));
}
__device__ __forceinline__ void atomic_max_float(float* addr, float v)
{
    if (v >= 0.0f)
        atomicMax((int*)addr, __float_as_int(v));
    else
        atomicMin((unsigned int*)addr, __float_as_uint(v));
}
// ---- packed FP32x2 (Blackwell FFMA2 / FMUL2 / FADD2): two IEEE fp32 operations per lane per instruction on a
// 64-bit register pair.  ptxas folds pk(s, s) into a scalar-broadcast operand, negations into operand modifiers
// and pk(imm, imm) into an immediate, so none of these helpers costs an instruction by itself.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi)
{
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
// pair of two scalars that live in unrelated registers, materialised ONCE in an adjacent register pair: ptxas treats
// a plain mov.b64 as a free copy and re-packs loop-invariant pairs from their halves in every loop iteration.
// Adding -0.0f read from memory at run time (x + (-0) == x bit-exactly for every x, signed zeros included) gives the
// pair a defining arithmetic instruction that cannot be folded away, so it stays where it is written.
__device__ __forceinline__ f32x2 pk2_once(float lo, float hi, float neg_zero)
{
    f32x2 r;
    asm volatile("{\n\t.reg .b64 t, z;\n\tmov.b64 t, {%1, %2};\n\tmov.b64 z, {%3, %3};\n\tadd.rn.f32x2 %0, t, z;\n\t}"
                 : "=l"(r)
                 : "f"(lo), "f"(hi), "f"(neg_zero));
    return r;
}
__device__ __forceinline__ f32x2 bc2(float s) { return pk2(s, s); }
__device__ __forceinline__ void upk2(f32x2 v, float& lo, float& hi)
{
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 neg2(f32x2 v)
{
    float lo, hi;
    upk2(v, lo, hi);
    return pk2(-lo, -hi);
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b)
{
    f32x2 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b)
{
    f32x2 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ float ex2_approx(float x)
{
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
#endif

}  // namespace pigs
