// Measured FP32-FMA, MUFU.EX2 and shared-memory-load peaks of the B200 this runs on: the roofline denominators of the
// compositor kernels (MEASURED_PEAKS.json only holds HBM and bf16 tensor figures). VERDICT r1 "missing" item 9.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/peak_fp32 tools/peak_fp32.cu && tools/peak_fp32
//
// Prints one JSON object: burst = best of 10 launches of ~0.5 ms; sustained = mean over a 2 s loop of launches.
// Each kernel keeps 8 independent dependency chains per thread, 1024 threads per CTA, 2 CTAs per SM.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

constexpr int kIters = 4096;
constexpr int kChains = 8;

__global__ void __launch_bounds__(1024, 2) ffma_kernel(float *out, float a, float b) {
    float x[kChains];
#pragma unroll
    for (int k = 0; k < kChains; ++k) x[k] = (float)(threadIdx.x + k);
    for (int i = 0; i < kIters; ++i) {
#pragma unroll
        for (int k = 0; k < kChains; ++k) x[k] = fmaf(x[k], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < kChains; ++k) s += x[k];
    if (s == 123.456f) out[0] = s;
}

// packed FP32 (Blackwell FFMA2: two FMAs per issued instruction, `fma.rn.f32x2`)
__global__ void __launch_bounds__(1024, 2) ffma2_kernel(float *out, float a, float b) {
    unsigned long long x[kChains], pa, pb;
    asm("mov.b64 %0, {%1, %1};" : "=l"(pa) : "f"(a));
    asm("mov.b64 %0, {%1, %1};" : "=l"(pb) : "f"(b));
#pragma unroll
    for (int k = 0; k < kChains; ++k) {
        const float v = (float)(threadIdx.x + k);
        asm("mov.b64 %0, {%1, %1};" : "=l"(x[k]) : "f"(v));
    }
    for (int i = 0; i < kIters; ++i) {
#pragma unroll
        for (int k = 0; k < kChains; ++k) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(x[k]) : "l"(pa), "l"(pb));
    }
    unsigned long long s = 0;
#pragma unroll
    for (int k = 0; k < kChains; ++k) s ^= x[k];
    if (s == 123456ull) out[0] = 1.f;
}

// issue-slot economy: the same FMA work as `ffma` / `ffma2` with one integer ALU instruction per FMA pair riding along
template <bool PACKED>
__global__ void __launch_bounds__(1024, 2) fma_alu_kernel(float *out, float a, float b, int c) {
    unsigned long long x[kChains], pa, pb;
    int y[kChains];
    asm("mov.b64 %0, {%1, %1};" : "=l"(pa) : "f"(a));
    asm("mov.b64 %0, {%1, %1};" : "=l"(pb) : "f"(b));
#pragma unroll
    for (int k = 0; k < kChains; ++k) {
        const float v = (float)(threadIdx.x + k);
        asm("mov.b64 %0, {%1, %1};" : "=l"(x[k]) : "f"(v));
        y[k] = threadIdx.x + k;
    }
    for (int i = 0; i < kIters; ++i) {
#pragma unroll
        for (int k = 0; k < kChains; ++k) {
            if (PACKED) {
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(x[k]) : "l"(pa), "l"(pb));
            } else {
                float lo, hi;
                asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(x[k]));
                lo = fmaf(lo, a, b);
                hi = fmaf(hi, a, b);
                asm("mov.b64 %0, {%1, %2};" : "=l"(x[k]) : "f"(lo), "f"(hi));
            }
            asm volatile("lop3.b32 %0, %0, %1, %0, 0x96;" : "+r"(y[k]) : "r"(c));
        }
    }
    unsigned long long s = 0;
#pragma unroll
    for (int k = 0; k < kChains; ++k) s ^= x[k] + y[k];
    if (s == 123456ull) out[0] = 1.f;
}

__global__ void __launch_bounds__(1024, 2) ex2_kernel(float *out, float a) {
    float x[kChains];
#pragma unroll
    for (int k = 0; k < kChains; ++k) x[k] = -(float)(threadIdx.x + k) * a;
    for (int i = 0; i < kIters; ++i) {
#pragma unroll
        for (int k = 0; k < kChains; ++k) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[k]));
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < kChains; ++k) s += x[k];
    if (s == 123.456f) out[0] = s;
}

// the compositor's inner-loop mix: 1 MUFU.EX2 per `F` FFMAs (forward: ~14, backward phase 1: ~20)
template <int F>
__global__ void __launch_bounds__(1024, 2) mix_kernel(float *out, float a, float b) {
    float x[kChains], y[kChains];
#pragma unroll
    for (int k = 0; k < kChains; ++k) x[k] = (float)(threadIdx.x + k) * 1e-3f, y[k] = x[k];
    for (int i = 0; i < kIters / 8; ++i) {
#pragma unroll
        for (int k = 0; k < kChains; ++k) {
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(y[k]));
#pragma unroll
            for (int f = 0; f < F; ++f) x[k] = fmaf(x[k], a, b);
            y[k] = y[k] * a - 1.0f;
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < kChains; ++k) s += x[k] + y[k];
    if (s == 123.456f) out[0] = s;
}

__global__ void __launch_bounds__(1024, 2) lds_kernel(float *out) {
    __shared__ float4 buf[1024];
    buf[threadIdx.x] = make_float4(threadIdx.x, 1.f, 2.f, 3.f);
    __syncthreads();
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    int idx = threadIdx.x;
    for (int i = 0; i < kIters; ++i) {
        const float4 v = buf[idx];  // conflict-free LDS.128
        acc.x += v.x;
        acc.y += v.y;
        acc.z += v.z;
        acc.w += v.w;
        idx = (idx + 33) & 1023;
    }
    if (acc.x + acc.y + acc.z + acc.w == 123.456f) out[0] = acc.x;
}

template <typename F>
static void measure(F launch, double work_per_launch, const char *name, const char *unit, bool last) {
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    for (int i = 0; i < 3; ++i) launch();
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int i = 0; i < 10; ++i) {
        cudaEventRecord(a);
        launch();
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
    }
    int n = (int)(2000.0f / best) + 1;  // ~2 s back to back
    cudaEventRecord(a);
    for (int i = 0; i < n; ++i) launch();
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float tot;
    cudaEventElapsedTime(&tot, a, b);
    printf(" \"%s\": {\"burst\": %.2f, \"sustained\": %.2f, \"unit\": \"%s\", \"ms_per_launch\": %.4f}%s\n", name,
           work_per_launch / (best * 1e-3) / 1e12, work_per_launch * n / (tot * 1e-3) / 1e12, unit, best,
           last ? "" : ",");
}

int main() {
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    int sm_clock_khz = 0;
    cudaDeviceGetAttribute(&sm_clock_khz, cudaDevAttrClockRate, 0);
    float *out;
    cudaMalloc(&out, 4);
    const int grid = prop.multiProcessorCount * 2 * 8;
    const double threads = (double)grid * 1024;
    printf("{\n \"gpu\": \"%s\", \"sms\": %d, \"sm_clock_mhz_max\": %.0f,\n", prop.name, prop.multiProcessorCount,
           sm_clock_khz / 1e3);
    printf(" \"arithmetic_fp32_tflops_at_max_clock\": %.2f, \"arithmetic_mufu_tops_at_max_clock\": %.3f,\n",
           prop.multiProcessorCount * 128 * 2.0 * sm_clock_khz * 1e3 / 1e12,
           prop.multiProcessorCount * 16.0 * sm_clock_khz * 1e3 / 1e12);
    measure([&] { ffma_kernel<<<grid, 1024>>>(out, 1.0001f, 0.5f); }, threads * kIters * kChains * 2.0, "fp32_fma",
            "TFLOP/s", false);
    measure([&] { ffma2_kernel<<<grid, 1024>>>(out, 1.0001f, 0.5f); }, threads * kIters * kChains * 4.0,
            "fp32_fma_packed_f32x2", "TFLOP/s", false);
    measure([&] { fma_alu_kernel<false><<<grid, 1024>>>(out, 1.0001f, 0.5f, 3); }, threads * kIters * kChains * 4.0,
            "fp32_2ffma_plus_1alu", "TFLOP/s", false);
    measure([&] { fma_alu_kernel<true><<<grid, 1024>>>(out, 1.0001f, 0.5f, 3); }, threads * kIters * kChains * 4.0,
            "fp32_1ffma2_plus_1alu", "TFLOP/s", false);
    measure([&] { ex2_kernel<<<grid, 1024>>>(out, 1e-3f); }, threads * kIters * kChains, "mufu_ex2", "Tops/s", false);
    measure([&] { mix_kernel<14><<<grid, 1024>>>(out, 1.0001f, 0.5f); }, threads * (kIters / 8) * kChains * (14 + 1) * 2.0,
            "mix_14ffma_1ex2_fp32", "TFLOP/s", false);
    measure([&] { mix_kernel<6><<<grid, 1024>>>(out, 1.0001f, 0.5f); }, threads * (kIters / 8) * kChains * (6 + 1) * 2.0,
            "mix_6ffma_1ex2_fp32", "TFLOP/s", false);
    measure([&] { lds_kernel<<<grid, 1024>>>(out); }, threads * kIters * 16.0, "lds128_bytes", "TB/s", true);
    printf("}\n");
    cudaFree(out);
    return 0;
}
