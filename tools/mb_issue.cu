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
    float a[N], x[N], y[N], m[4];
    const float b = 1.0000001f + 1e-9f * threadIdx.x, c = 1e-7f + src[1];
#pragma unroll
    for (int i = 0; i < N; i++) {
        a[i] = 1.0f + i * 1e-3f + threadIdx.x * 1e-6f;
        x[i] = 1.0f + src[(threadIdx.x + i) & 511] * 1e-6f;
        y[i] = 1e-6f * src[(threadIdx.x + 37 * i + N) & 511] + 1e-7f;
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
    for (int i = 0; i < N; i++) s += a[i];
#pragma unroll
    for (int i = 0; i < 4; i++) s += m[i];
    if (s == 123.456f) sink[0] = s;
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
    run<5>("fadd 2 regs", 64, 0, sink, src, sms, mhz);
    run<6>("fmul 2 regs", 64, 0, sink, src, sms, mhz);
    run<7>("ffma reuse + 1 mufu/16", 64, 4, sink, src, sms, mhz);
    run<3>("ffma distinct + 1 mufu/16", 64, 4, sink, src, sms, mhz);
    run<4>("ffma distinct + 2 mufu/16", 64, 8, sink, src, sms, mhz);
    run<9>("ffma distinct + 4 mufu/16", 64, 16, sink, src, sms, mhz);
    run<8>("ffma distinct + 1 lds128/16", 64, 4, sink, src, sms, mhz);
    return 0;
}
