// bulk_probe.cu — how fast does cp.async.bulk (UBLKCP) stream HBM into shared memory as a function of the size of
// one request?  Persistent CTAs, one mbarrier per buffer, double-buffered; the consumer only touches one word per
// row so the copy engine is what is measured.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/tools/bulk_probe tools/bulk_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t n) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(n) : "memory");
}
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* b) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(dst)),
                 "l"(src), "r"(bytes), "r"(s32(b))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    asm volatile(
        "{\n .reg .pred p;\n W: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra D;\n bra W;\n D:\n}" ::"r"(s32(b)),
        "r"(parity)
        : "memory");
}

constexpr int kRows = 128, kRowBytes = 192, kStride = 208;
// mode 0: one request per tile (dense); 1: one per row (padded dst, every thread issues its own); 2: one per 32 rows
template <int kMode>
__global__ void __launch_bounds__(128) probe(const char* __restrict__ src, int ntiles, float* out) {
    extern __shared__ __align__(128) char smem[];
    char* buf[2] = {smem, smem + kRows * kStride};
    __shared__ __align__(8) uint64_t bar[2];
    if (threadIdx.x == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    auto issue = [&](int tile, int b) {
        const char* g = src + (size_t)tile * kRows * kRowBytes;
        if (kMode == 0) {
            if (threadIdx.x == 0) { mbar_expect(&bar[b], kRows * kRowBytes); bulk_g2s(buf[b], g, kRows * kRowBytes, &bar[b]); }
        } else if (kMode == 1) {
            if (threadIdx.x == 0) mbar_expect(&bar[b], kRows * kRowBytes);
            bulk_g2s(buf[b] + threadIdx.x * kStride, g + threadIdx.x * kRowBytes, kRowBytes, &bar[b]);
        } else {
            if (threadIdx.x == 0) mbar_expect(&bar[b], kRows * kRowBytes);
            if ((threadIdx.x & 31) == 0)
                bulk_g2s(buf[b] + threadIdx.x * kRowBytes, g + threadIdx.x * kRowBytes, 32 * kRowBytes, &bar[b]);
        }
    };
    float acc = 0.f;
    uint32_t ph[2] = {0, 0};
    int it = 0;
    if ((int)blockIdx.x < ntiles) issue(blockIdx.x, 0);
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int b = it & 1;
        const int next = tile + gridDim.x;
        if (next < ntiles) issue(next, b ^ 1);
        mbar_wait(&bar[b], ph[b]);
        ph[b] ^= 1;
        acc += *reinterpret_cast<const float*>(buf[b] + threadIdx.x * (kMode == 1 ? kStride : kRowBytes));
        __syncthreads();   // buffer b is free for the request after next
    }
    if (acc == 123.456f) out[0] = acc;
}

template <int kMode>
void run(const char* name, const char* d, int ntiles, float* out, int ctas_per_sm) {
    const int smem = 2 * kRows * kStride;
    cudaFuncSetAttribute(probe<kMode>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e9f;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0);
        probe<kMode><<<148 * ctas_per_sm, 128, smem>>>(d, ntiles, out);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double bytes = (double)ntiles * kRows * kRowBytes;
    printf("{\"mode\": \"%s\", \"ctas_per_sm\": %d, \"ms\": %.4f, \"GBps\": %.1f, \"requests_per_us_per_sm\": %.2f, \"err\": \"%s\"}\n", name,
           ctas_per_sm, best, bytes / best / 1e6,
           (kMode == 0 ? 1.0 : kMode == 1 ? 128.0 : 4.0) * ntiles / 148.0 / (best * 1e3), cudaGetErrorString(cudaGetLastError()));
}

int main() {
    const int ntiles = 3000000 / kRows;   // 576 MB: the SH rows of bench config 3
    char* d; float* out;
    cudaMalloc(&d, (size_t)ntiles * kRows * kRowBytes);
    cudaMemset(d, 0, (size_t)ntiles * kRows * kRowBytes);
    cudaMalloc(&out, 4);
    for (int c : {2, 4}) {
        run<0>("one 24576-byte request per 128 rows", d, ntiles, out, c);
        run<2>("one 6144-byte request per 32 rows", d, ntiles, out, c);
        run<1>("one 192-byte request per row (208-byte destination stride)", d, ntiles, out, c);
    }
    return 0;
}
