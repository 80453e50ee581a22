// fp32_peak.cu — measured SIMT peaks of the B200 the blend kernels are bound by:
// FP32 FMA issue (register and immediate operand forms), MUFU.EX2, shared-memory 128-bit loads.
// MEASURED_PEAKS.json has HBM and tensor-core figures only; the blend rooflines need these.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/tools/fp32_peak tools/fp32_peak.cu && build/tools/fp32_peak
// Prints one JSON object.  CUDA events, 3 warm-up launches, best and median of 10.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <vector>

constexpr int kThreads = 256;
constexpr int kIters = 2048;
constexpr int kChains = 8;

// x = fma(x, a, b), operands in registers (a, b come from memory so nothing folds)
__global__ void __launch_bounds__(kThreads) ffma_reg(const float* __restrict__ in, float* out) {
    const float a = in[0], b = in[1];
    float x[kChains];
#pragma unroll
    for (int c = 0; c < kChains; ++c) x[c] = in[2 + c] + threadIdx.x;
#pragma unroll 1
    for (int i = 0; i < kIters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int c = 0; c < kChains; ++c) x[c] = fmaf(x[c], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kChains; ++c) s += x[c];
    if (s == 12345.678f) out[0] = s;
}

// x = fma(x, imm, b): one operand is an immediate (the FFMA form with the highest issue rate)
__global__ void __launch_bounds__(kThreads) ffma_imm(const float* __restrict__ in, float* out) {
    const float b = in[1];
    float x[kChains];
#pragma unroll
    for (int c = 0; c < kChains; ++c) x[c] = in[2 + c] + threadIdx.x;
#pragma unroll 1
    for (int i = 0; i < kIters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int c = 0; c < kChains; ++c) x[c] = fmaf(x[c], 0.99993f, b);
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kChains; ++c) s += x[c];
    if (s == 12345.678f) out[0] = s;
}

// x = fma(y, z, x) with three DISTINCT source registers per instruction (the shape of the blend inner loops)
__global__ void __launch_bounds__(kThreads) ffma_3src(const float* __restrict__ in, float* out) {
    float x[kChains], y[kChains];
#pragma unroll
    for (int c = 0; c < kChains; ++c) { x[c] = in[2 + c] + threadIdx.x; y[c] = in[10 + c]; }
#pragma unroll 1
    for (int i = 0; i < kIters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int c = 0; c < kChains; ++c) x[c] = fmaf(y[c], y[(c + 3) & 7], x[c]);
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kChains; ++c) s += x[c];
    if (s == 12345.678f) out[0] = s;
}

__global__ void __launch_bounds__(kThreads) mufu_ex2(const float* __restrict__ in, float* out) {
    float x[kChains];
#pragma unroll
    for (int c = 0; c < kChains; ++c) x[c] = in[2 + c] * 1e-3f;
#pragma unroll 1
    for (int i = 0; i < kIters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int c = 0; c < kChains; ++c) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[c]));
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kChains; ++c) s += x[c];
    if (s == 12345.678f) out[0] = s;
}

// conflict-free 128-bit shared loads (each lane its own 16-byte slot), 8 in flight
__global__ void __launch_bounds__(kThreads) lds128(const float* __restrict__ in, float* out) {
    __shared__ float4 s[kThreads * 2];
    s[threadIdx.x] = make_float4(in[0], in[1], in[2], in[3]);
    s[threadIdx.x + kThreads] = s[threadIdx.x];
    __syncthreads();
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const volatile float4* p = s;
#pragma unroll 1
    for (int i = 0; i < kIters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            const int j = (threadIdx.x + 32 * u) & (2 * kThreads - 1);
            acc.x += p[j].x; acc.y += p[j].y; acc.z += p[j].z; acc.w += p[j].w;
        }
    }
    if (acc.x + acc.y + acc.z + acc.w == 12345.678f) out[0] = acc.x;
}

template <typename F>
static void time_kernel(const char* name, F launch, double work_per_launch, const char* unit, double per_clk_div,
                        bool last) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    for (int i = 0; i < 3; ++i) launch();
    cudaDeviceSynchronize();
    std::vector<float> ms(10);
    for (int i = 0; i < 10; ++i) {
        cudaEventRecord(e0);
        launch();
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms[i], e0, e1);
    }
    std::sort(ms.begin(), ms.end());
    const double best = work_per_launch / (ms[0] * 1e-3), med = work_per_launch / (ms[5] * 1e-3);
    printf("  \"%s\": {\"best\": %.4g, \"median\": %.4g, \"unit\": \"%s\", \"ms_best\": %.4f, \"per_sm_per_clk_at_max_clock\": %.2f}%s\n",
           name, best, med, unit, ms[0], best / per_clk_div, last ? "" : ",");
}

int main() {
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    int clk_khz = 0;
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const int sms = prop.multiProcessorCount;
    const int blocks = sms * 8;
    float h[32];
    for (int i = 0; i < 32; ++i) h[i] = 0.5f + 0.01f * i;
    float *d_in, *d_out;
    cudaMalloc(&d_in, sizeof(h));
    cudaMalloc(&d_out, 64);
    cudaMemcpy(d_in, h, sizeof(h), cudaMemcpyHostToDevice);
    const double lanes = (double)blocks * kThreads;
    const double n_op = lanes * kIters * 8.0 * kChains;          // lane-instructions per launch
    const double clk = clk_khz * 1e3;
    printf("{\n  \"gpu\": \"%s\", \"sms\": %d, \"clock_rate_khz\": %d,\n", prop.name, sms, clk_khz);
    time_kernel("ffma_reg_tflops", [&] { ffma_reg<<<blocks, kThreads>>>(d_in, d_out); }, 2.0 * n_op / 1e12, "TFLOP/s",
                2.0 * sms * clk / 1e12, false);
    time_kernel("ffma_imm_tflops", [&] { ffma_imm<<<blocks, kThreads>>>(d_in, d_out); }, 2.0 * n_op / 1e12, "TFLOP/s",
                2.0 * sms * clk / 1e12, false);
    time_kernel("ffma_3src_tflops", [&] { ffma_3src<<<blocks, kThreads>>>(d_in, d_out); }, 2.0 * n_op / 1e12, "TFLOP/s",
                2.0 * sms * clk / 1e12, false);
    time_kernel("mufu_ex2_tops", [&] { mufu_ex2<<<blocks, kThreads>>>(d_in, d_out); }, n_op / 1e12, "Tex2/s",
                sms * clk / 1e12, false);
    time_kernel("lds128_tbs", [&] { lds128<<<blocks, kThreads>>>(d_in, d_out); }, lanes * kIters * 16.0 * 16.0 / 1e12,
                "TB/s", sms * clk / 1e12, true);
    printf("}\n");
    const cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { fprintf(stderr, "CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
