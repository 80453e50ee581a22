// gsr_loss.cu — the step between forward and backward: the photometric loss of the
// training loop, L = (1 - lambda) * mean|x - y| + lambda * (1 - mean(SSIM map)), and its gradient w.r.t.
// the rendered image, in two launches instead of five grouped 11x11 convolutions, ~15 elementwise
// kernels and their autograd.
//
// Reference: utils/loss_utils.py:17-18 (l1_loss), :23-31 (window: binary32 1-D Gaussian sigma 1.5 divided by
// its sum, 2-D window = outer product), :39-63 (ssim: mu, E[x^2], E[y^2], E[xy] by zero-padded grouped
// conv2d, C1 = 0.01^2, C2 = 0.03^2, mean of the map), train.py:99-101 (the combination).
//
// Forward (loss_forward_kernel): one CTA per 32x32 tile of one channel.  The 42x42 halo tiles of x and y
// are staged in shared memory once; the window is separable, so the five moments take a horizontal
// 11-tap pass into shared memory and a vertical 11-tap pass in registers.  Per pixel the kernel forms the
// SSIM map value and the three partial derivatives of the map w.r.t. the convolution outputs that depend
// on x (mu_x, E[x^2], E[xy]) and stores those (12 B/pixel/channel) for the backward launch.  |x - y| and
// the map are summed per CTA and added to two binary64 accumulators; the last CTA to finish writes the
// three scalars (loss, l1, ssim).
//
// Backward (loss_backward_kernel): the adjoint of a correlation with a symmetric window is the same
// correlation, so dSSIM/dx = conv(d_mu) + 2x * conv(d_exx) + y * conv(d_exy); the kernel convolves the
// three stored maps the same separable way and adds the L1 term's sign(x - y).
//
// Algorithmic bytes per pixel-channel: forward 8 in + 12 out, backward 20 in + 4 out = 44 B; HBM-bound.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <mutex>

#include "gsr_internal.cuh"

namespace gsr {

namespace {

constexpr int kWin = 11;
constexpr int kHalo = kWin / 2;
constexpr int kTile = 32;
constexpr int kIn = kTile + 2 * kHalo;  // 42
constexpr int kLossThreads = 256;
constexpr int kRowsPerThread = kTile * kTile / kLossThreads;  // 4

struct LossArgs {
    int C, H, W;
    int vec4;        // l1-only kernels: all pointers 16-byte aligned
    const float* x;  // rendered image [C,H,W]
    const float* y;  // ground truth   [C,H,W]
    float* maps;     // [3][C][H][W]: d map / d mu_x, d map / d E[x^2], d map / d E[xy]
    double* sums;    // [0] sum |x-y|, [1] sum of the SSIM map; ((unsigned*)sums)[4] = CTAs finished
    float* out3;     // loss, l1, ssim
    double a0, a_l1, a_ssim;   // loss = a0 + a_l1 * l1 + a_ssim * ssim
    float coef_l1, coef_ssim;  // backward: dL/dx = up * (coef_l1 * sign(x-y) + coef_ssim * dSSIMsum/dx) (both / N)
    const float* upstream;     // device scalar dL/dloss, or null = 1
    float* dL_dx;
    float w[kWin];
};

// Stages the zero-padded 42x42 halo tile of one channel plane.
__device__ __forceinline__ void stage_tile(float (*dst)[kIn], const float* __restrict__ plane, int H, int W, int x0,
                                           int y0) {
    for (int i = threadIdx.x; i < kIn * kIn; i += kLossThreads) {
        const int r = i / kIn, c = i - r * kIn;
        const int gy = y0 + r - kHalo, gx = x0 + c - kHalo;
        dst[r][c] = (gy >= 0 && gy < H && gx >= 0 && gx < W) ? __ldg(plane + (size_t)gy * W + gx) : 0.f;
    }
}

// CTA sums (fixed order inside the CTA), two binary64 atomics per CTA; the last CTA to arrive writes the scalars.
__device__ __forceinline__ void finish_sums(const LossArgs& a, float l1_sum, float ssim_sum,
                                            float (*s_part)[kLossThreads / 32], double N) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        l1_sum += __shfl_xor_sync(0xffffffffu, l1_sum, o);
        ssim_sum += __shfl_xor_sync(0xffffffffu, ssim_sum, o);
    }
    if ((threadIdx.x & 31) == 0) {
        s_part[0][threadIdx.x >> 5] = l1_sum;
        s_part[1][threadIdx.x >> 5] = ssim_sum;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double l1 = 0.0, ss = 0.0;
        for (int k = 0; k < kLossThreads / 32; ++k) {
            l1 += (double)s_part[0][k];
            ss += (double)s_part[1][k];
        }
        atomicAdd(a.sums + 0, l1);
        atomicAdd(a.sums + 1, ss);
        __threadfence();
        unsigned* done = reinterpret_cast<unsigned*>(a.sums + 2);
        const unsigned total = gridDim.x * gridDim.y * gridDim.z;
        if (atomicAdd(done, 1u) + 1u == total) {
            __threadfence();
            const double l1m = *reinterpret_cast<volatile double*>(a.sums + 0) / N;
            const double ssm = *reinterpret_cast<volatile double*>(a.sums + 1) / N;
            a.out3[0] = (float)(a.a0 + a.a_l1 * l1m + a.a_ssim * ssm);
            a.out3[1] = (float)l1m;
            a.out3[2] = (float)ssm;
        }
    }
}

// a_ssim == 0 (l1_loss alone): no moments, no maps — one streaming pass each way.
__global__ void __launch_bounds__(kLossThreads) l1_forward_kernel(const LossArgs a) {
    __shared__ float s_part[2][kLossThreads / 32];
    const long long n = (long long)a.C * a.H * a.W, n4 = a.vec4 ? n >> 2 : 0;
    const long long stride = (long long)gridDim.x * kLossThreads;
    float sum = 0.f;
    for (long long i = (long long)blockIdx.x * kLossThreads + threadIdx.x; i < n4; i += stride) {
        const float4 x = __ldg(reinterpret_cast<const float4*>(a.x) + i), y = __ldg(reinterpret_cast<const float4*>(a.y) + i);
        sum += (fabsf(x.x - y.x) + fabsf(x.y - y.y)) + (fabsf(x.z - y.z) + fabsf(x.w - y.w));
    }
    if (blockIdx.x == 0)
        for (long long i = 4 * n4 + threadIdx.x; i < n; i += kLossThreads) sum += fabsf(a.x[i] - a.y[i]);
    finish_sums(a, sum, 0.f, s_part, (double)n);
}

__global__ void __launch_bounds__(kLossThreads) l1_backward_kernel(const LossArgs a) {
    const long long n = (long long)a.C * a.H * a.W, n4 = a.vec4 ? n >> 2 : 0;
    const long long stride = (long long)gridDim.x * kLossThreads;
    const float k = (a.upstream ? __ldg(a.upstream) : 1.f) * a.coef_l1;
    auto sgn = [k](float d) { return d > 0.f ? k : (d < 0.f ? -k : 0.f); };
    for (long long i = (long long)blockIdx.x * kLossThreads + threadIdx.x; i < n4; i += stride) {
        const float4 x = __ldg(reinterpret_cast<const float4*>(a.x) + i), y = __ldg(reinterpret_cast<const float4*>(a.y) + i);
        reinterpret_cast<float4*>(a.dL_dx)[i] = make_float4(sgn(x.x - y.x), sgn(x.y - y.y), sgn(x.z - y.z), sgn(x.w - y.w));
    }
    if (blockIdx.x == 0)
        for (long long i = 4 * n4 + threadIdx.x; i < n; i += kLossThreads) a.dL_dx[i] = sgn(a.x[i] - a.y[i]);
}

__global__ void __launch_bounds__(kLossThreads) loss_forward_kernel(const LossArgs a) {
    __shared__ float s_x[kIn][kIn], s_y[kIn][kIn];
    __shared__ float s_h[5][kIn][kTile];
    __shared__ float s_part[2][kLossThreads / 32];
    const int x0 = blockIdx.x * kTile, y0 = blockIdx.y * kTile, ch = blockIdx.z;
    const size_t plane = (size_t)a.H * a.W;
    stage_tile(s_x, a.x + ch * plane, a.H, a.W, x0, y0);
    stage_tile(s_y, a.y + ch * plane, a.H, a.W, x0, y0);
    __syncthreads();
    // horizontal pass: 42 rows x 32 columns, five moments each
    for (int i = threadIdx.x; i < kIn * kTile; i += kLossThreads) {
        const int r = i >> 5, c = i & 31;
        float hx = 0.f, hy = 0.f, hxx = 0.f, hyy = 0.f, hxy = 0.f;
#pragma unroll
        for (int t = 0; t < kWin; ++t) {
            const float xv = s_x[r][c + t], yv = s_y[r][c + t], wt = a.w[t];
            const float wx = wt * xv, wy = wt * yv;
            hx += wx;
            hy += wy;
            hxx = fmaf(wx, xv, hxx);
            hyy = fmaf(wy, yv, hyy);
            hxy = fmaf(wx, yv, hxy);
        }
        s_h[0][r][c] = hx;
        s_h[1][r][c] = hy;
        s_h[2][r][c] = hxx;
        s_h[3][r][c] = hyy;
        s_h[4][r][c] = hxy;
    }
    __syncthreads();
    // vertical pass: each thread owns 4 consecutive rows of one column, so the 14 rows its windows span are
    // read once into registers (5 x 14 shared loads for 4 pixels instead of 5 x 44)
    const int c = threadIdx.x & 31, r0 = (threadIdx.x >> 5) * kRowsPerThread;
    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
    float l1_sum = 0.f, ssim_sum = 0.f;
    float mom[5][kRowsPerThread];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        float col[kRowsPerThread + kWin - 1];
#pragma unroll
        for (int i = 0; i < kRowsPerThread + kWin - 1; ++i) col[i] = s_h[k][r0 + i][c];
#pragma unroll
        for (int j = 0; j < kRowsPerThread; ++j) {
            float acc = 0.f;
#pragma unroll
            for (int t = 0; t < kWin; ++t) acc = fmaf(a.w[t], col[j + t], acc);
            mom[k][j] = acc;
        }
    }
#pragma unroll
    for (int j = 0; j < kRowsPerThread; ++j) {
        const int r = r0 + j;
        const int gy = y0 + r, gx = x0 + c;
        if (gy >= a.H || gx >= a.W) continue;
        const float mu1 = mom[0][j], mu2 = mom[1][j], e11 = mom[2][j], e22 = mom[3][j], e12 = mom[4][j];
        const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu1_mu2 = mu1 * mu2;
        const float sigma1_sq = e11 - mu1_sq, sigma2_sq = e22 - mu2_sq, sigma12 = e12 - mu1_mu2;
        const float A1 = 2.f * mu1_mu2 + C1, A2 = 2.f * sigma12 + C2;
        const float B1 = mu1_sq + mu2_sq + C1, B2 = sigma1_sq + sigma2_sq + C2;
        const float inv_B1 = 1.f / B1, inv_B2 = 1.f / B2;
        const float inv_B = inv_B1 * inv_B2;
        const float S = A1 * A2 * inv_B;
        ssim_sum += S;
        l1_sum += fabsf(s_x[r + kHalo][c + kHalo] - s_y[r + kHalo][c + kHalo]);
        if (a.maps) {
            const size_t n = (size_t)a.C * plane, at = ch * plane + (size_t)gy * a.W + gx;
            a.maps[at] = 2.f * mu2 * (A2 - A1) * inv_B - 2.f * mu1 * S * (inv_B1 - inv_B2);
            a.maps[n + at] = -S * inv_B2;
            a.maps[2 * n + at] = 2.f * A1 * inv_B;
        }
    }
    finish_sums(a, l1_sum, ssim_sum, s_part, (double)a.C * (double)plane);
}

__global__ void __launch_bounds__(kLossThreads) loss_backward_kernel(const LossArgs a) {
    __shared__ float s_m[3][kIn][kIn];
    __shared__ float s_h[3][kIn][kTile];
    const int x0 = blockIdx.x * kTile, y0 = blockIdx.y * kTile, ch = blockIdx.z;
    const size_t plane = (size_t)a.H * a.W, n = (size_t)a.C * plane;
#pragma unroll
    for (int k = 0; k < 3; ++k) stage_tile(s_m[k], a.maps + k * n + ch * plane, a.H, a.W, x0, y0);
    __syncthreads();
    for (int i = threadIdx.x; i < kIn * kTile; i += kLossThreads) {
        const int r = i >> 5, c = i & 31;
        float h0 = 0.f, h1 = 0.f, h2 = 0.f;
#pragma unroll
        for (int t = 0; t < kWin; ++t) {
            const float wt = a.w[t];
            h0 = fmaf(wt, s_m[0][r][c + t], h0);
            h1 = fmaf(wt, s_m[1][r][c + t], h1);
            h2 = fmaf(wt, s_m[2][r][c + t], h2);
        }
        s_h[0][r][c] = h0;
        s_h[1][r][c] = h1;
        s_h[2][r][c] = h2;
    }
    __syncthreads();
    const int c = threadIdx.x & 31, r0 = (threadIdx.x >> 5) * kRowsPerThread;
    const float up = a.upstream ? __ldg(a.upstream) : 1.f;
    const float k_l1 = up * a.coef_l1, k_ssim = up * a.coef_ssim;
    float g[3][kRowsPerThread];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float col[kRowsPerThread + kWin - 1];
#pragma unroll
        for (int i = 0; i < kRowsPerThread + kWin - 1; ++i) col[i] = s_h[k][r0 + i][c];
#pragma unroll
        for (int j = 0; j < kRowsPerThread; ++j) {
            float acc = 0.f;
#pragma unroll
            for (int t = 0; t < kWin; ++t) acc = fmaf(a.w[t], col[j + t], acc);
            g[k][j] = acc;
        }
    }
#pragma unroll
    for (int j = 0; j < kRowsPerThread; ++j) {
        const int gy = y0 + r0 + j, gx = x0 + c;
        if (gy >= a.H || gx >= a.W) continue;
        const size_t at = ch * plane + (size_t)gy * a.W + gx;
        const float xv = __ldg(a.x + at), yv = __ldg(a.y + at);
        const float d = xv - yv;
        const float sgn = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
        a.dL_dx[at] = fmaf(k_ssim, g[0][j] + 2.f * xv * g[1][j] + yv * g[2][j], k_l1 * sgn);
    }
}

void fill_window(float* w) {
    // utils/loss_utils.py:23-25: exp() in binary64, stored as binary32, divided by the binary32 sum
    float g[kWin], sum = 0.f;
    for (int x = 0; x < kWin; ++x) {
        const double d = (double)(x - kWin / 2);
        g[x] = (float)std::exp(-(d * d) / (2.0 * 1.5 * 1.5));
        sum += g[x];
    }
    for (int x = 0; x < kWin; ++x) w[x] = g[x] / sum;
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

unsigned l1_grid(int C, int H, int W) {
    const long long n4 = ((long long)C * H * W) >> 2;
    return (unsigned)std::max(1ll, std::min((n4 + kLossThreads - 1) / kLossThreads, 148ll * 8));
}

dim3 loss_grid(int C, int H, int W) { return dim3((W + kTile - 1) / kTile, (H + kTile - 1) / kTile, C); }

}  // namespace

size_t loss_workspace_bytes(int C, int H, int W, bool save_for_backward) {
    return kLossHeaderBytes + (save_for_backward ? (size_t)3 * C * H * W * sizeof(float) : 0);
}

int launch_loss_forward(int C, int H, int W, const float* image, const float* gt, double a0, double a_l1,
                        double a_ssim, void* workspace, bool save_for_backward, float* out3, cudaStream_t s) {
    LossArgs a{};
    a.C = C; a.H = H; a.W = W;
    a.x = image; a.y = gt;
    a.sums = static_cast<double*>(workspace);
    a.maps = save_for_backward ? reinterpret_cast<float*>(static_cast<char*>(workspace) + kLossHeaderBytes) : nullptr;
    a.out3 = out3;
    a.a0 = a0; a.a_l1 = a_l1; a.a_ssim = a_ssim;
    fill_window(a.w);
    if (cudaMemsetAsync(workspace, 0, kLossHeaderBytes, s) != cudaSuccess) return GSR_ERR_CUDA;
    if (a_ssim == 0.0) {
        a.maps = nullptr;
        a.vec4 = aligned16(image) && aligned16(gt);
        l1_forward_kernel<<<l1_grid(C, H, W), kLossThreads, 0, s>>>(a);
    } else {
        loss_forward_kernel<<<loss_grid(C, H, W), kLossThreads, 0, s>>>(a);
    }
    GSR_AFTER_LAUNCH(s, 0);
    return GSR_OK;
}

int launch_loss_backward(int C, int H, int W, const float* image, const float* gt, double a_l1, double a_ssim,
                         const void* workspace, const float* upstream, float* dL_dimage, cudaStream_t s) {
    LossArgs a{};
    a.C = C; a.H = H; a.W = W;
    a.x = image; a.y = gt;
    a.maps = reinterpret_cast<float*>(static_cast<char*>(const_cast<void*>(workspace)) + kLossHeaderBytes);
    const double N = (double)C * H * W;
    a.coef_l1 = (float)(a_l1 / N);
    a.coef_ssim = (float)(a_ssim / N);
    a.upstream = upstream;
    a.dL_dx = dL_dimage;
    fill_window(a.w);
    if (a_ssim == 0.0) {
        a.vec4 = aligned16(image) && aligned16(gt) && aligned16(dL_dimage);
        l1_backward_kernel<<<l1_grid(C, H, W), kLossThreads, 0, s>>>(a);   // never touches the maps
    } else {
        loss_backward_kernel<<<loss_grid(C, H, W), kLossThreads, 0, s>>>(a);
    }
    GSR_AFTER_LAUNCH(s, 0);
    return GSR_OK;
}

}  // namespace gsr

// ---- gsr_loss_value_and_grad_static: the loss and its image gradient without a caller-provided workspace ----
//
// The two kernels above need scratch in global memory: the 256-byte header (two binary64 sums and the
// finished-CTA counter) and the three per-pixel derivative maps the forward launch leaves for the backward one
// (12 bytes per image element).  gsr_loss_forward / gsr_loss_backward take that scratch from the caller.  This
// entry point is for hosts that have nothing to give: it keeps the scratch in the array below, which the CUDA
// runtime places in device memory when this file's module is loaded on a GPU and frees at process exit, so the
// call itself never asks anybody for memory.
namespace {

constexpr size_t kStaticScratchBytes = gsr::kLossHeaderBytes + (size_t)3 * GSR_LOSS_STATIC_MAX_ELEMS * sizeof(float);

// One copy per GPU that loads the library (about 71 MiB each).  Host code must not hand this NAME to a
// kernel: on the host it is only a handle for cudaGetSymbolAddress, not a device address.
static __device__ __align__(256) unsigned char g_loss_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_static_mutex;
void* g_static_ptr[kMaxDevices] = {};   // device address of the array on each device, filled on first use

// Looks the array's device address up once per device, with that device current (the address differs from
// GPU to GPU).  Everything after the first call on a device is a table read under a mutex.
int static_scratch_on_current_device(void** out) {
    int dev = -1;
    GSR_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return GSR_ERR_UNSUPPORTED;
    std::lock_guard<std::mutex> hold(g_static_mutex);
    if (!g_static_ptr[dev]) {
        void* p = nullptr;
        GSR_CUDA(cudaGetSymbolAddress(&p, g_loss_static_scratch));
        if (!p || (reinterpret_cast<uintptr_t>(p) & 255)) return GSR_ERR_WORKSPACE;
        g_static_ptr[dev] = p;
    }
    *out = g_static_ptr[dev];
    return GSR_OK;
}

}  // namespace

extern "C" GSR_API int gsr_loss_value_and_grad_static(int C, int H, int W, const float* image, const float* gt,
                                                      double a0, double a_l1, double a_ssim,
                                                      const float* upstream, float* out3, float* dL_dimage,
                                                      void* stream) {
    if (C <= 0 || H <= 0 || W <= 0 || !image || !gt || !out3 || !dL_dimage) return GSR_ERR_INVALID_ARGUMENT;
    if (C > 65535 || (H + 31) / 32 > 65535) return GSR_ERR_UNSUPPORTED;
    if ((long long)C * H * W > (long long)GSR_LOSS_STATIC_MAX_ELEMS) return GSR_ERR_SCRATCH_LIMIT;
    void* scratch = nullptr;
    int st = static_scratch_on_current_device(&scratch);
    if (st != GSR_OK) return st;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    // The forward launch zeroes the header (the sums and the counter still hold the previous call's values)
    // and overwrites every map element the backward launch will read; nothing else carries over.
    st = gsr::launch_loss_forward(C, H, W, image, gt, a0, a_l1, a_ssim, scratch, /*save_for_backward=*/true, out3, s);
    if (st != GSR_OK) return st;
    return gsr::launch_loss_backward(C, H, W, image, gt, a_l1, a_ssim, scratch, upstream, dL_dimage, s);
}
