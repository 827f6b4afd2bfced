// pipebench.cu — per-SM issue rates of the CUDA-core instructions the rollout kernels are made of, measured on the
// box (MEASURED_PEAKS.json has HBM and tensor figures only).  Each test runs NCHAIN independent dependency chains per
// thread, 1024 threads x (8 x #SM) blocks, so the number printed is the saturated rate in thread-ops / clk / SM
// (128 = one warp instruction per clock per SM sub-partition).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/pipebench tools/pipebench.cu && tools/pipebench
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define NCHAIN 8
#define ITERS 2048

#define BODY_KERNEL(name, T, INIT, OPS)                                          \
    __global__ void __launch_bounds__(1024) name(T *sink, T seed) {              \
        T a[NCHAIN];                                                             \
        _Pragma("unroll") for (int i = 0; i < NCHAIN; ++i) a[i] = INIT;          \
        const T m = seed, c = seed * (T)0.5;                                     \
        (void)c;                                                                 \
        for (int it = 0; it < ITERS; ++it) {                                     \
            _Pragma("unroll") for (int i = 0; i < NCHAIN; ++i) { OPS; }          \
        }                                                                        \
        T s = a[0];                                                              \
        _Pragma("unroll") for (int i = 1; i < NCHAIN; ++i) s += a[i];            \
        if (s == (T)123456.789) sink[0] = s;                                     \
    }

#define FINIT ((float)(threadIdx.x + i) * 1e-3f + 0.5f)
#define DINIT ((double)(threadIdx.x + i) * 1e-3 + 0.5)
#define IINIT ((int)(threadIdx.x + i))

BODY_KERNEL(k_ffma, float, FINIT, asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(m), "f"(c)))
BODY_KERNEL(k_fadd, float, FINIT, asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(c)))
BODY_KERNEL(k_fmul, float, FINIT, asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(m)))
BODY_KERNEL(k_fmnmx, float, FINIT, asm volatile("min.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(m)))
BODY_KERNEL(k_frnd, float, FINIT, asm volatile("cvt.rmi.f32.f32 %0, %0;" : "+f"(a[i])))
BODY_KERNEL(k_rcp, float, FINIT, asm volatile("rcp.approx.ftz.f32 %0, %0;\n\tadd.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(c)))
BODY_KERNEL(k_sin, float, FINIT, asm volatile("sin.approx.ftz.f32 %0, %0;" : "+f"(a[i])))
BODY_KERNEL(k_ex2, float, FINIT, asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i])))
BODY_KERNEL(k_dadd, double, DINIT, asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(a[i]) : "d"(c)))
BODY_KERNEL(k_dfma, double, DINIT, asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(a[i]) : "d"(m), "d"(c)))
BODY_KERNEL(k_imad, int, IINIT, asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(m), "r"(c)))
BODY_KERNEL(k_lop, int, IINIT, asm volatile("xor.b32 %0, %0, %1;" : "+r"(a[i]) : "r"(m)))
BODY_KERNEL(k_iadd, int, IINIT, asm volatile("add.s32 %0, %0, %1;" : "+r"(a[i]) : "r"(m)))

// select: setp + selp pair on a float (FSETP + FSEL)
BODY_KERNEL(k_fsel, float, FINIT,
            asm volatile("{ .reg .pred p; setp.gt.f32 p, %0, %1; selp.f32 %0, %2, %0, p; }" : "+f"(a[i]) : "f"(m), "f"(c)))

// conversions: f32 -> f64 -> f32 round trip (F2F.F64.F32 + F2F.F32.F64), f32 -> s32 -> f32 (F2I + I2F)
BODY_KERNEL(k_f2f64, float, FINIT,
            asm volatile("{ .reg .f64 d; cvt.f64.f32 d, %0; cvt.rn.f32.f64 %0, d; }" : "+f"(a[i])))
BODY_KERNEL(k_f2i2f, float, FINIT,
            asm volatile("{ .reg .s32 r; cvt.rni.s32.f32 r, %0; cvt.rn.f32.s32 %0, r; }" : "+f"(a[i])))

// packed fp32: two FMAs per lane per instruction (FFMA2 / FADD2 / FMUL2)
__global__ void __launch_bounds__(1024) k_ffma2(float *sink, float seed) {
    unsigned long long a[NCHAIN];
#pragma unroll
    for (int i = 0; i < NCHAIN; ++i) {
        float lo = (threadIdx.x + i) * 1e-3f + 0.5f, hi = lo + 0.25f;
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(a[i]) : "f"(lo), "f"(hi));
    }
    unsigned long long m, c;
    asm volatile("mov.b64 %0, {%1, %1};" : "=l"(m) : "f"(seed));
    asm volatile("mov.b64 %0, {%1, %1};" : "=l"(c) : "f"(seed * 0.5f));
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < NCHAIN; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(a[i]) : "l"(m), "l"(c));
    }
    unsigned long long s = a[0];
#pragma unroll
    for (int i = 1; i < NCHAIN; ++i) s ^= a[i];
    if (s == 0x123456789abcdefull) sink[0] = 1.f;
}

// mixed issue: one FFMA2 (FMA pipe, 2 cycles) interleaved with one ALU-pipe op (FMNMX): do the two overlap?
__global__ void __launch_bounds__(1024) k_ffma2_plus_alu(float *sink, float seed) {
    unsigned long long a[NCHAIN];
    float b[NCHAIN];
#pragma unroll
    for (int i = 0; i < NCHAIN; ++i) {
        float lo = (threadIdx.x + i) * 1e-3f + 0.5f, hi = lo + 0.25f;
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(a[i]) : "f"(lo), "f"(hi));
        b[i] = lo;
    }
    unsigned long long m, c;
    asm volatile("mov.b64 %0, {%1, %1};" : "=l"(m) : "f"(seed));
    asm volatile("mov.b64 %0, {%1, %1};" : "=l"(c) : "f"(seed * 0.5f));
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < NCHAIN; ++i) {
            asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(a[i]) : "l"(m), "l"(c));
            asm volatile("min.f32 %0, %0, %1;" : "+f"(b[i]) : "f"(seed));
        }
    }
    unsigned long long s = a[0];
    float t = b[0];
#pragma unroll
    for (int i = 1; i < NCHAIN; ++i) {
        s ^= a[i];
        t += b[i];
    }
    if (s == 0x123456789abcdefull || t == 123456.789f) sink[0] = 1.f;
}

// FFMA + FMNMX alternating (scalar): issue-bound at 128/clk/SM total if both count against the same issue port
__global__ void __launch_bounds__(1024) k_ffma_plus_alu(float *sink, float seed) {
    float a[NCHAIN], b[NCHAIN];
#pragma unroll
    for (int i = 0; i < NCHAIN; ++i) {
        a[i] = (threadIdx.x + i) * 1e-3f + 0.5f;
        b[i] = a[i] + 0.25f;
    }
    const float m = seed, c = seed * 0.5f;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < NCHAIN; ++i) {
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(m), "f"(c));
            asm volatile("min.f32 %0, %0, %1;" : "+f"(b[i]) : "f"(seed));
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < NCHAIN; ++i) s += a[i] + b[i];
    if (s == 123456.789f) sink[0] = s;
}

template <typename K, typename T>
static void run(const char *name, K kern, T seed, double ops_per_iter, int sms, double clk_hz) {
    T *sink;
    cudaMalloc(&sink, 64);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int blocks = sms * 8;
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        kern<<<blocks, 1024>>>(sink, seed);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0 && ms < best) best = ms;
    }
    const double ops = (double)blocks * 1024.0 * NCHAIN * ITERS * ops_per_iter;
    const double per_s = ops / (best * 1e-3);
    printf("%-18s %8.3f ms  %9.3e thread-ops/s  %7.2f /clk/SM (at %.0f MHz)\n", name, best, per_s,
           per_s / sms / clk_hz, clk_hz / 1e6);
    cudaFree(sink);
}

int main() {
    int sms = 0, khz = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double clk = khz * 1e3;
    printf("SMs %d, max SM clock %.0f MHz; rates assume the max clock\n", sms, clk / 1e6);
    run("FFMA", k_ffma, 0.999f, 1, sms, clk);
    run("FADD", k_fadd, 0.999f, 1, sms, clk);
    run("FMUL", k_fmul, 0.999f, 1, sms, clk);
    run("FFMA2 (x2 lanes)", k_ffma2, 0.999f, 2, sms, clk);
    run("FFMA2+FMNMX", k_ffma2_plus_alu, 0.999f, 3, sms, clk);
    run("FFMA+FMNMX", k_ffma_plus_alu, 0.999f, 2, sms, clk);
    run("FMNMX", k_fmnmx, 0.999f, 1, sms, clk);
    run("FSETP+FSEL", k_fsel, 0.999f, 2, sms, clk);
    run("FRND.FLOOR", k_frnd, 0.999f, 1, sms, clk);
    run("F2I+I2F", k_f2i2f, 0.999f, 2, sms, clk);
    run("F2F 32->64->32", k_f2f64, 0.999f, 2, sms, clk);
    run("MUFU.RCP", k_rcp, 0.999f, 1, sms, clk);
    run("MUFU.SIN", k_sin, 0.999f, 1, sms, clk);
    run("MUFU.EX2", k_ex2, 0.999f, 1, sms, clk);
    run("DADD", k_dadd, 0.999, 1, sms, clk);
    run("DFMA", k_dfma, 0.999, 1, sms, clk);
    run("IMAD", k_imad, 3, 1, sms, clk);
    run("LOP3", k_lop, 3, 1, sms, clk);
    run("IADD3", k_iadd, 3, 1, sms, clk);
    return 0;
}
