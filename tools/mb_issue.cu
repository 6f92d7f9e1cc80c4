// mb_issue.cu — dev microbenchmarks for the issue/dispatch cost of the instruction mixes the evaluation
// kernels use (FP32 with three distinct register operands, MUFU.EX2 among FFMAs, LDS.128 broadcasts).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/mb_issue tools/mb_issue.cu
// Prints issue slots per SMSP-cycle for each mix (1.0 = one warp instruction per scheduler per clock).
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ float ex2(float x)
{
    float y;
    asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

constexpr int N = 16;

typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b)
{
    u64 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ float lo_of(u64 v)
{
    float a, b;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
    return a;
}
__device__ __forceinline__ float hi_of(u64 v)
{
    float a, b;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
    return b;
}
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c)
{
    u64 d;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ u64 mul2(u64 a, u64 b)
{
    u64 d;
    asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ u64 add2(u64 a, u64 b)
{
    u64 d;
    asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// packed-FP32 modes (16 packed instructions = 32 lane-ops per inner pass):
// PM 0: A = fma2(X, Y, A)  three distinct 64-bit operands
// PM 1: A = fma2(bcast(g), Y, A)  scalar broadcast operand (.F32) + two distinct pairs
// PM 2: PM 0 + 2 MUFU per 16 FFMA2 (1 MUFU per 16 lane-FMAs)
// PM 3: PM 0 + 8 scalar FADD per 16 FFMA2
// PM 4: PM 0 + 8 IADD3/LOP3 per 16 FFMA2
// PM 5: A = mul2(X, A)   PM 6: A = add2(X, A)
// PM 8 / 9 / 10: one pair operand shared by consecutive FFMA2s (does the operand-reuse cache save the third cycle?)
// PM 7: PM 0 + 4 MUFU per 16 FFMA2 + 4 scalar FMUL (the forward's real side-work per 2x... )
template <int PM>
__global__ void __launch_bounds__(256) kp(int iters, float* sink, const float* src)
{
    u64 a[N], x[N], y[N];
    float m[4], f[8];
    int q[8];
    const float g = 1.0f + src[threadIdx.x & 511];
#pragma unroll
    for (int i = 0; i < N; i++) {
        a[i] = pk(1.0f + i * 1e-3f + threadIdx.x * 1e-6f, 0.5f + i * 1e-3f);
        x[i] = pk(1.0f + src[(threadIdx.x + i) & 511] * 1e-6f, 1.0f + src[(threadIdx.x + 64 + i) & 511] * 1e-6f);
        y[i] = pk(1e-6f * src[(threadIdx.x + 128 + i) & 511] + 1e-7f, 1e-6f * src[(threadIdx.x + 192 + i) & 511]);
    }
#pragma unroll
    for (int i = 0; i < 4; i++) m[i] = -0.5f - src[(threadIdx.x + i + 99) & 511];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        f[i] = src[(threadIdx.x + i + 199) & 511];
        q[i] = threadIdx.x + i;
    }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int i = 0; i < N; i++) {
                if (PM == 0 || PM == 2 || PM == 3 || PM == 4 || PM == 7) a[i] = fma2(x[i], y[i], a[i]);
                if (PM == 1) a[i] = fma2(pk(g, g), y[i], a[i]);
                if (PM == 8) a[i] = fma2(x[0], y[i], a[i]);            // ONE pair shared by all 16 (slot A; .reuse)
                if (PM == 9) a[i] = fma2(x[i / 6], y[i], a[i]);        // groups of 6 share a pair (the backward's accumulate phase)
                if (PM == 10) a[i] = fma2(x[i & 1], y[i], a[i]);       // two shared pairs alternating (interleaved k = 0 / 1)
                if (PM == 5) a[i] = mul2(x[i], a[i]);
                if (PM == 6) a[i] = add2(x[i], a[i]);
                if (PM == 3 && (i & 1)) f[i >> 1] = f[i >> 1] + g;
                if (PM == 4 && (i & 1)) q[i >> 1] = (q[i >> 1] + it) ^ r;
                if (PM == 7 && (i & 3) == 0) m[i >> 2] = ex2(m[i >> 2]);
                if (PM == 7 && (i & 3) == 2) f[i >> 2] = f[i >> 2] * g;
            }
            if (PM == 2) {
                m[r] = ex2(m[r]);
                m[(r + 1) & 3] = ex2(m[(r + 1) & 3]);
            }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < N; i++) s += lo_of(a[i]) + hi_of(a[i]);
#pragma unroll
    for (int i = 0; i < 4; i++) s += m[i];
#pragma unroll
    for (int i = 0; i < 8; i++) s += f[i] + q[i];
    if (s == 123.456f) sink[0] = s;
}

// MODE 0: a = fma(a, b, c)        (b, c loop-invariant -> reuse cache friendly)
// MODE 1: a[i] = fma(x[i], y[i], a[i])   3 distinct registers per FFMA
// MODE 2: a[i] = fma(w, y[i], a[i])      one shared operand
// MODE 3: MODE 1 + 1 MUFU per 16 FFMA
// MODE 4: MODE 1 + 2 MUFU per 16 FFMA
// MODE 5: a[i] = x[i] + a[i] (FADD, 2 operands)
// MODE 6: a[i] = x[i] * a[i] (FMUL)
// MODE 7: MODE 0 + 1 MUFU per 16
// MODE 8: MODE 1 + 1 LDS.128 (broadcast) per 16 FFMA
// MODE 9: MODE 1 + 4 MUFU per 16 FFMA
template <int MODE>
__global__ void __launch_bounds__(256) k(int iters, float* sink, const float* src)
{
    __shared__ float4 sm[64];
    if (threadIdx.x < 64) sm[threadIdx.x] = make_float4(src[threadIdx.x], 1.f, 2.f, 3.f);
    __syncthreads();
    float a[N], x[N], y[N], z[N], m[4];
    const float b = 1.0000001f + 1e-9f * threadIdx.x, c = 1e-7f + src[1];
#pragma unroll
    for (int i = 0; i < N; i++) {
        a[i] = 1.0f + i * 1e-3f + threadIdx.x * 1e-6f;
        x[i] = 1.0f + src[(threadIdx.x + i) & 511] * 1e-6f;
        y[i] = 1e-6f * src[(threadIdx.x + 37 * i + N) & 511] + 1e-7f;
        z[i] = 1.0f + 1e-6f * src[(threadIdx.x + 41 * i + 3) & 511];
    }
#pragma unroll
    for (int i = 0; i < 4; i++) m[i] = -0.5f - src[(threadIdx.x + i + 99) & 511];
    float4 acc4 = make_float4(0, 0, 0, 0);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int i = 0; i < N; i++) {
                if (MODE == 0 || MODE == 7) a[i] = fmaf(a[i], b, c);
                if (MODE == 1 || MODE == 3 || MODE == 4 || MODE == 8 || MODE == 9) a[i] = fmaf(x[i], y[i], a[i]);
                if (MODE == 2) a[i] = fmaf(b, y[i], a[i]);
                if (MODE == 10) {  // shared-operand FFMA alternating with an unrelated 2-operand FMUL
                    a[i] = fmaf(b, y[i], a[i]);
                    x[i] = x[i] * z[i];
                }
                if (MODE == 5) a[i] = x[i] + a[i];
                if (MODE == 6) a[i] = x[i] * a[i];
                if ((MODE == 9) && (i % 4 == 0)) m[(i / 4) & 3] = ex2(m[(i / 4) & 3]);
            }
            if (MODE == 3 || MODE == 7) m[r] = ex2(m[r]);
            if (MODE == 4) {
                m[r] = ex2(m[r]);
                m[(r + 1) & 3] = ex2(m[(r + 1) & 3]);
            }
            if (MODE == 8) {
                const float4 t = sm[(it + r) & 63];
                acc4.x += t.x;
            }
        }
    }
    float s = acc4.x;
#pragma unroll
    for (int i = 0; i < N; i++) s += a[i] + x[i];
#pragma unroll
    for (int i = 0; i < 4; i++) s += m[i];
    if (s == 123.456f) sink[0] = s;
}

template <int PM>
static void runp(const char* name, double packed_per_iter, double other_per_iter, float* sink, const float* src, int sms, double mhz)
{
    for (int cps = 2; cps <= 4; cps *= 2) {
        const int iters = 20000;
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        kp<PM><<<sms * cps, 256>>>(200, sink, src);
        float best = 1e30f;
        for (int rep = 0; rep < 3; rep++) {
            cudaEventRecord(e0);
            kp<PM><<<sms * cps, 256>>>(iters, sink, src);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
        }
        const double warps_per_smsp = cps * 8 / 4.0;
        const double cycles = best * 1e-3 * mhz * 1e6 / (iters * warps_per_smsp);
        printf("%-40s ctas/sm=%d  %.3f ms  cycles per (%.0f packed + %.0f other)=%.2f  -> %.3f cycles per packed instr, lane-FMA rate %.3f of peak\n",
               name, cps, best, packed_per_iter, other_per_iter, cycles, cycles / packed_per_iter, 2.0 * packed_per_iter / cycles);
    }
}

template <int MODE>
static void run(const char* name, double fp_per_iter, double other_per_iter, float* sink, const float* src, int sms, double mhz)
{
    for (int cps = 2; cps <= 8; cps *= 2) {
        const int iters = 20000;
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        k<MODE><<<sms * cps, 256>>>(200, sink, src);
        float best = 1e30f;
        for (int rep = 0; rep < 3; rep++) {
            cudaEventRecord(e0);
            k<MODE><<<sms * cps, 256>>>(iters, sink, src);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
        }
        const double warps_per_smsp = cps * 8 / 4.0;
        const double inst = (fp_per_iter + other_per_iter) * iters * warps_per_smsp;  // warp instr per SMSP
        const double cycles = best * 1e-3 * mhz * 1e6;
        printf("%-34s ctas/sm=%d  %.3f ms  issue/clk/SMSP=%.3f  cycles per (%.0f fp + %.0f other)=%.2f\n", name, cps, best,
               inst / cycles, fp_per_iter, other_per_iter, cycles / (iters * warps_per_smsp));
    }
}

int main()
{
    int sms = 148, khz = 1965000;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double mhz = khz / 1000.0;
    printf("SMs %d, clock %.0f MHz\n", sms, mhz);
    float *sink, *src;
    cudaMalloc(&sink, 64);
    cudaMalloc(&src, 4096);
    cudaMemset(src, 0, 4096);
    run<0>("ffma reuse (a=a*b+c)", 64, 0, sink, src, sms, mhz);
    run<1>("ffma 3 distinct regs", 64, 0, sink, src, sms, mhz);
    run<2>("ffma shared A operand", 64, 0, sink, src, sms, mhz);
    run<10>("ffma shared A alternating w/ fmul", 128, 0, sink, src, sms, mhz);
    run<5>("fadd 2 regs", 64, 0, sink, src, sms, mhz);
    run<6>("fmul 2 regs", 64, 0, sink, src, sms, mhz);
    run<7>("ffma reuse + 1 mufu/16", 64, 4, sink, src, sms, mhz);
    run<3>("ffma distinct + 1 mufu/16", 64, 4, sink, src, sms, mhz);
    run<4>("ffma distinct + 2 mufu/16", 64, 8, sink, src, sms, mhz);
    run<9>("ffma distinct + 4 mufu/16", 64, 16, sink, src, sms, mhz);
    run<8>("ffma distinct + 1 lds128/16", 64, 4, sink, src, sms, mhz);
    runp<0>("ffma2 3 distinct pairs", 64, 0, sink, src, sms, mhz);
    runp<1>("ffma2 bcast .F32 + 2 pairs", 64, 0, sink, src, sms, mhz);
    runp<8>("ffma2 one shared pair (reuse) + 2 pairs", 64, 0, sink, src, sms, mhz);
    runp<9>("ffma2 shared pair per group of 6", 64, 0, sink, src, sms, mhz);
    runp<10>("ffma2 two shared pairs alternating", 64, 0, sink, src, sms, mhz);
    runp<5>("fmul2", 64, 0, sink, src, sms, mhz);
    runp<6>("fadd2", 64, 0, sink, src, sms, mhz);
    runp<2>("ffma2 distinct + 2 mufu/16", 64, 8, sink, src, sms, mhz);
    runp<7>("ffma2 distinct + 4 mufu + 4 fmul /16", 64, 32, sink, src, sms, mhz);
    runp<3>("ffma2 distinct + 8 scalar fadd/16", 64, 32, sink, src, sms, mhz);
    runp<4>("ffma2 distinct + 8 int ops/16", 64, 32, sink, src, sms, mhz);
    return 0;
}
