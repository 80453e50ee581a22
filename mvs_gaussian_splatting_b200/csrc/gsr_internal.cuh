// gsr_internal.cuh — shared declarations of the sm_100a rasterizer library (private).
//
// Numerics policy.  Everything that decides an INTEGER result (radii, tile rects, depth key bits,
// cull decisions) is written with explicit round-to-nearest intrinsics (__fmul_rn / __fadd_rn /
// __fmaf_rn / __fdiv_rn / __fsqrt_rn), which nvcc never contracts or reassociates, in exactly the
// operation order oracle/gs_oracle.c documents.  The result is therefore independent of -fmad and
// bit-identical to the oracle's -ffp-contract=off build.  Tolerance-level arithmetic (blending,
// all gradients) is written naturally and left to the compiler.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/gsrast_b200.h"

namespace gsr {

constexpr int kTile = 16;          // pixels per tile edge (16x16, as the reference bins)
constexpr int kTilePixels = 256;
constexpr float kNearPlane = 0.2f; // view-space z cull threshold

// ---- launch accounting + error plumbing -------------------------------------------------
void count_launch(int n = 1);
int cuda_fail(cudaError_t e);  // records e for gsr_last_cuda_error(), returns GSR_ERR_CUDA

#define GSR_CUDA(expr)                                    \
    do {                                                  \
        cudaError_t _e = (expr);                          \
        if (_e != cudaSuccess) return ::gsr::cuda_fail(_e); \
    } while (0)

// After a launch: always catch launch-configuration errors; in debug mode also synchronise so a
// faulting kernel is reported by the call that launched it (the reference's `debug` behaviour).
#define GSR_AFTER_LAUNCH(stream, debug)                                  \
    do {                                                                 \
        ::gsr::count_launch();                                           \
        cudaError_t _e = cudaGetLastError();                             \
        if (_e == cudaSuccess && (debug)) _e = cudaStreamSynchronize(stream); \
        if (_e != cudaSuccess) return ::gsr::cuda_fail(_e);              \
    } while (0)

// ---- optional per-stage device timing (gsr_stage_timing_*; off by default) -----------------
enum Stage {
    kStagePreprocess = 0, kStageDepthSort, kStageEmitKeys, kStageTileRanges, kStageSort, kStageBlendFwd,
    kStageBlendBwd, kStagePreprocessBwd, kStageCount
};
bool stage_timing_on();
void stage_mark(int stage, cudaStream_t s, bool begin);
// RAII: brackets the launches of one stage with a CUDA event pair on `s` when timing is enabled.
struct StageScope {
    int stage; cudaStream_t s; bool on;
    StageScope(int stage_, cudaStream_t s_) : stage(stage_), s(s_), on(stage_timing_on()) {
        if (on) stage_mark(stage, s, true);
    }
    ~StageScope() { if (on) stage_mark(stage, s, false); }
};

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Runs `setup` once per device (per-device bit in `mask`, for the CURRENT device): function attributes such as
// the dynamic shared-memory opt-in apply only to the device that is current when they are set, and one process
// may render on several devices.  The mutex is held across `setup`, so a second thread making its first call on
// the same device cannot launch before the attributes exist; the bit is set only when `setup` succeeded.
cudaError_t once_per_device(unsigned long long* mask, cudaError_t (*setup)());
// Multiprocessors of the current device (cached per device): persistent grids are sized from it.
int sm_count();

struct Grid {
    int W, H, gx, gy, tiles;
};
inline Grid make_grid(int W, int H) {
    Grid g;
    g.W = W; g.H = H;
    g.gx = (W + kTile - 1) / kTile;
    g.gy = (H + kTile - 1) / kTile;
    g.tiles = g.gx * g.gy;
    return g;
}

// Typed views of the three workspaces.
struct GeomView {
    float* depths;
    float2* means2D;
    float4* conic_opacity;
    float4* rgb;  // .w carries the clamped bits
    uint32_t* tiles_touched;
    gsr_forward_info* info;
    float* grad_accum;  // [P,12]
    uint32_t* order_keys;      // [num_visible] depth bits of the visible Gaussians, ascending (stable)
    uint32_t* order;           // [num_visible] Gaussian index of depth-order slot k
    uint32_t* order_keys_alt;  // ping-pong partners of the per-Gaussian sort
    uint32_t* order_alt;
    uint32_t* sort_ctl;        // histograms, tickets, look-back status words
};
struct BinView {           // tile keys are u16 up to 65536 tiles, else u32 (tile_key_bytes)
    void* keys_unsorted;   // tile id of every instance, emitted in DEPTH order
    uint32_t* vals_unsorted;
    void* keys;            // sorted by tile (stable)
    uint32_t* point_list;
    void* keys_alt;        // radix-sort ping-pong partner of (keys, point_list)
    uint32_t* vals_alt;
    uint32_t* sort_ctl;   // histograms, tickets, look-back status words
};
struct ImageView {
    uint2* ranges;
    float* final_T;
    uint32_t* n_contrib;
};
// View of rows [r, ...) of the per-Gaussian arrays (slot addressing of the sharded entries); the
// sort/scan arrays and the info block are not per-slot and stay where they are.
inline GeomView offset_rows(GeomView v, size_t r) {
    v.depths += r; v.means2D += r; v.conic_opacity += r; v.rgb += r; v.tiles_touched += r;
    v.grad_accum += 12 * r;
    v.info = nullptr;  // the counts are taken by the depth-order stage of the rank that renders the view
    return v;
}
void compute_layout(int P, int W, int H, int64_t cap, gsr_layout* L);
GeomView geom_view(void* ws, const gsr_layout& L);
BinView bin_view(void* ws, const gsr_layout& L);
ImageView image_view(void* ws, const gsr_layout& L);

// u32 entries of the control regions of the per-Gaussian depth sort (+ prefix sum) and of the
// per-instance tile sort.
size_t depth_sort_ctl_entries(size_t P);
size_t tile_sort_ctl_entries(size_t cap, size_t P, size_t tiles);
int tile_key_bytes(int tiles);   // 2 up to 65536 tiles, else 4

struct RawMode;
// ---- stage launchers (each returns a gsr_status) ------------------------------------------
int launch_preprocess(int P, int D, int M, const Grid& g, const float* means3D, const float* shs,
                      const float* colors_precomp, const float* opacities, const float* scales,
                      float scale_modifier, const float* rotations, const float* cov3D_precomp,
                      const float* view, const float* proj, const float* campos, float tanfovx,
                      float tanfovy, int32_t* radii, GeomView gv, cudaStream_t s, int debug,
                      const struct RawMode& rm, bool* depth_counts_done = nullptr);

// Zeroes the control region of the depth sort (histograms, tickets, look-back words); must precede the kernels
// that fill it: launch_preprocess when it takes the counts itself (depth_counts_done), launch_depth_order always.
int launch_depth_ctl_reset(int P, GeomView gv, cudaStream_t s);
// Counts (info->num_visible, info->num_rendered, on the device; skipped when the preprocess kernel already took
// them: counts_done) and the depth-order sort of the visible (depth bits, index) pairs.
int launch_depth_order(int P, GeomView gv, cudaStream_t s, int debug, bool counts_done = false);

// The reference's 64-bit (tile << 32 | depth bits) key of every list entry, unsorted (emission order) or sorted,
// rebuilt from the stored tile ids and depths[index] (tests / debugging; nothing on the render path reads it).
int launch_materialize_keys(int P, const Grid& g, GeomView gv, BinView bv, int64_t cap, int sorted, uint64_t* out,
                            cudaStream_t s);

int launch_binning(int P, const Grid& g, const int32_t* radii, GeomView gv, BinView bv,
                   int64_t cap, ImageView iv, cudaStream_t s, int debug);

int launch_blend_forward(const Grid& g, const float* bg, GeomView gv, BinView bv, ImageView iv,
                         float* out_color, cudaStream_t s, int debug);

// Instrumented forward (no outputs): counters[0] = pairs the reference algorithm evaluates,
// counters[1] = pairs it blends.  Tools / benchmarks only.
int launch_blend_pair_counts(const Grid& g, GeomView gv, BinView bv, ImageView iv,
                             unsigned long long* counters, cudaStream_t s);

int launch_blend_backward(int P, const Grid& g, const float* bg, GeomView gv, BinView bv,
                          ImageView iv, const float* dL_dpix, cudaStream_t s, int debug);

int launch_preprocess_backward(int P, int D, int M, const Grid& g, const float* means3D,
                               const float* shs, int colors_precomp, const float* scales,
                               float scale_modifier, const float* rotations,
                               const float* cov3D_precomp, const float* view, const float* proj,
                               const float* campos, float tanfovx, float tanfovy,
                               const int32_t* radii, GeomView gv, float* dL_dmeans2D,
                               float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D,
                               float* dL_dcov3D, float* dL_dsh, float* dL_dscales, float* dL_drots,
                               int accumulate, cudaStream_t s, int debug, const struct RawMode& rm);

// Raw-parameter mode: the per-Gaussian inputs are the reference's PRE-activation
// parameters (scene/gaussian_model.py:152-183): scales = log-scale, rotations = unnormalised quaternion,
// opacities = logit, shs = _features_dc [P,1,3] and sh_rest = _features_rest [P,M-1,3].  The kernels
// apply exp / normalize / sigmoid / the dc|rest concatenation in registers and shared memory, and the
// backward applies their chain rules and writes dL/dsh split the same way.
struct RawMode {
    int raw = 0;
    const float* sh_rest = nullptr;
    float* dL_dsh_rest = nullptr;
};

struct ForwardViewArgs {
    const float* view; const float* proj; const float* campos;
    float tanfovx, tanfovy;
    int32_t* radii;
    GeomView gv;
    int32_t* radii_local = nullptr;   // optional local copies (gsr_view_slot): radius | clamped << 28, conic + opacity
    float4* conic_local = nullptr;
};
int launch_preprocess_views(int P, int D, int M, const Grid& g, const float* means3D, const float* shs,
                            const float* colors_precomp, const float* opacities, const float* scales,
                            float scale_modifier, const float* rotations, const float* cov3D_precomp, int n_views,
                            const ForwardViewArgs* views, cudaStream_t s, int debug, const RawMode& rm = RawMode(),
                            bool* depth_counts_done = nullptr);

// Multi-view form: the Gaussians' parameters are read once, `views` (host array, n_views <= GSR_MAX_VIEWS)
// says per view where the camera, the forward rows (conic_opacity, rgb), radii and blend accumulators of
// THESE P Gaussians live and where that view's dL/dmeans2D goes; parameter gradients are summed over views.
struct BackwardViewArgs {
    const float* view; const float* proj; const float* campos;
    float tanfovx, tanfovy;
    const int32_t* radii;
    GeomView gv;
    float* dL_dmeans2D;
    const int32_t* radii_local = nullptr;   // when set: read instead of radii / gv.conic_opacity / gv.rgb
    const float4* conic_local = nullptr;
};
int launch_preprocess_backward_views(int P, int D, int M, const Grid& g, const float* means3D, const float* shs,
                                     int colors_precomp, const float* scales, float scale_modifier,
                                     const float* rotations, const float* cov3D_precomp, int n_views,
                                     const BackwardViewArgs* views, float* dL_dcolors, float* dL_dopacity,
                                     float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscales,
                                     float* dL_drots, int accumulate, cudaStream_t s, int debug,
                                     const RawMode& rm = RawMode());

// The activations alone (the unfused form of RawMode, same device functions): tests and tools.
int launch_activate(int P, int M, const float* features_dc, const float* features_rest, const float* opacity_raw,
                    const float* scaling_raw, const float* rotation_raw, float* shs, float* opacities, float* scales,
                    float* rotations, cudaStream_t s);

// fused Adam over all parameter groups + densification statistics (gsr_optim.cu).
int launch_adam(int n_groups, const gsr_adam_group* groups, double beta1, double beta2, double eps, int step,
                cudaStream_t s);
int launch_densification_stats(int P, const int32_t* radii, const float* dL_dmeans2D, float* max_radii2D,
                               float* xyz_gradient_accum, float* denom, cudaStream_t s);

constexpr size_t kLossHeaderBytes = 256;  // two binary64 sums + the finished-CTA counter, maps follow
size_t loss_workspace_bytes(int C, int H, int W, bool save_for_backward);
int launch_loss_forward(int C, int H, int W, const float* image, const float* gt, double a0, double a_l1,
                        double a_ssim, void* workspace, bool save_for_backward, float* out3, cudaStream_t s);
int launch_loss_backward(int C, int H, int W, const float* image, const float* gt, double a_l1, double a_ssim,
                         const void* workspace, const float* upstream, float* dL_dimage, cudaStream_t s);

int launch_ply(bool pack, int64_t P, const gsr_ply_columns* c, float* rows, float* xyz, float* f_dc, float* f_rest,
               float* opacity, float* scale, float* rot, cudaStream_t s);

int launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present,
                        cudaStream_t s);

// ---- device helpers shared by several translation units ------------------------------------
#ifdef __CUDACC__

// p0 + p1 + p2 with the oracle's contraction policy: fma(a2,b2, fma(a0,b0, rn(a1*b1))).
__device__ __forceinline__ float dot3_rn(float a0, float b0, float a1, float b1, float a2, float b2) {
    return __fmaf_rn(a2, b2, __fmaf_rn(a0, b0, __fmul_rn(a1, b1)));
}
__device__ __forceinline__ float dot2_rn(float a0, float b0, float a1, float b1) {
    return __fmaf_rn(a0, b0, __fmul_rn(a1, b1));
}
// Column-major 4x4 (flat 16 floats) times point; row r = m[r], m[4+r], m[8+r], m[12+r].
__device__ __forceinline__ float xform_row_rn(const float* m, int r, float x, float y, float z) {
    return __fadd_rn(dot3_rn(m[r], x, m[4 + r], y, m[8 + r], z), m[12 + r]);
}

// Tile rectangle [min,max) of a splat, clamped to the grid (bit-exact with the oracle).
__device__ __forceinline__ void tile_rect(float px, float py, int radius, int gx, int gy,
                                          int& x0, int& y0, int& x1, int& y1) {
    const float rf = (float)radius;
    x0 = min(gx, max(0, __float2int_rz(__fdiv_rn(__fadd_rn(px, -rf), 16.0f))));
    y0 = min(gy, max(0, __float2int_rz(__fdiv_rn(__fadd_rn(py, -rf), 16.0f))));
    x1 = min(gx, max(0, __float2int_rz(__fdiv_rn(__fadd_rn(__fadd_rn(__fadd_rn(px, rf), 16.0f), -1.0f), 16.0f))));
    y1 = min(gy, max(0, __float2int_rz(__fdiv_rn(__fadd_rn(__fadd_rn(__fadd_rn(py, rf), 16.0f), -1.0f), 16.0f))));
}

// ---- activations of the raw-parameter mode: the formulas of torch.exp / torch.sigmoid /
// torch.nn.functional.normalize(eps = 1e-12) that scene/gaussian_model.py:28-45 installs ----
__device__ __forceinline__ float act_scale(float r) { return expf(r); }
__device__ __forceinline__ float act_opacity(float r) { return __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-r))); }
__device__ __forceinline__ float quat_norm(float4 q) {  // max(||q||, eps)
    const float n2 = __fmaf_rn(q.w, q.w, __fmaf_rn(q.z, q.z, __fmaf_rn(q.y, q.y, __fmul_rn(q.x, q.x))));
    return fmaxf(__fsqrt_rn(n2), 1e-12f);
}
__device__ __forceinline__ float4 act_rotation(float4 q) {
    const float n = quat_norm(q);
    return make_float4(__fdiv_rn(q.x, n), __fdiv_rn(q.y, n), __fdiv_rn(q.z, n), __fdiv_rn(q.w, n));
}

// Warp-cooperative copy between a contiguous global block of `nrows` rows x `rr` floats (16-byte
// aligned start) and columns [c0, c0 + rr) of the warp's shared tile (row r at tile + r*stride), moved as
// 16-byte pieces that may straddle two rows.  Fetch: only rows in `want` and the first `nr` floats of a
// row are loaded / written.  Used for the split SH layout of RawMode (dc: rr = 3; rest: rr = 3(M-1)).
__device__ __forceinline__ void fetch_rows_f4(const float* __restrict__ base, int nrows, int rr, int nr, int c0,
                                              unsigned want, float* tile, int stride, unsigned lane) {
    const int total = nrows * rr, nvec = total >> 2;
    for (int jb = 0; jb < nvec; jb += 128) {
        float4 val[4];
        int r0[4], k0[4];
        bool ld[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int j = jb + (int)lane + 32 * u;
            ld[u] = false;
            r0[u] = k0[u] = 0;
            if (j < nvec) {
                r0[u] = (4 * j) / rr;
                k0[u] = 4 * j - r0[u] * rr;
                const bool first = ((want >> r0[u]) & 1u) && k0[u] < nr;
                const bool second = k0[u] + 3 >= rr && r0[u] + 1 < 32 && ((want >> (r0[u] + 1)) & 1u);
                ld[u] = first || second;
                if (ld[u]) val[u] = __ldg(reinterpret_cast<const float4*>(base) + j);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (!ld[u]) continue;
            const float e[4] = {val[u].x, val[u].y, val[u].z, val[u].w};
            int r = r0[u], k = k0[u];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                if (k >= rr) { k -= rr; ++r; }
                if (r < 32 && ((want >> r) & 1u) && k < nr) tile[r * stride + c0 + k] = e[i];
                ++k;
            }
        }
    }
    for (int f = 4 * nvec + (int)lane; f < total; f += 32) {
        const int r = f / rr, k = f - r * rr;
        if (((want >> r) & 1u) && k < nr) tile[r * stride + c0 + k] = __ldg(base + f);
    }
}

// The reverse: columns [c0, c0 + rr) of all `nrows` tile rows -> the contiguous global block (streaming).
__device__ __forceinline__ void store_rows_f4(float* __restrict__ base, int nrows, int rr, int c0, const float* tile,
                                              int stride, unsigned lane) {
    const int total = nrows * rr, nvec = total >> 2;
#pragma unroll 4
    for (int j = lane; j < nvec; j += 32) {
        int r = (4 * j) / rr, k = 4 * j - r * rr;
        float e[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (k >= rr) { k -= rr; ++r; }
            e[i] = tile[r * stride + c0 + k];
            ++k;
        }
        __stcs(reinterpret_cast<float4*>(base) + j, make_float4(e[0], e[1], e[2], e[3]));
    }
    for (int f = 4 * nvec + (int)lane; f < total; f += 32) {
        const int r = f / rr, k = f - r * rr;
        base[f] = tile[r * stride + c0 + k];
    }
}

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31u; }

// ---- bulk asynchronous copies (the TMA engine's descriptor-less form: cp.async.bulk, SASS UBLKCP) + mbarrier ----
// Sizes and both addresses are multiples of 16 bytes.  Measured on the B200 (tools/bulk_probe.cu,
// profiles/r02j_bulk_probe.json): one 24 KB request per 128 SH rows streams at 6.6 TB/s, one 192-byte request per
// row (the form that lets culled rows be skipped and rows land on a padded, conflict-free stride) at 5.8 TB/s.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(arrivals) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// The one arrival of a phase, announcing how many bytes the copies of that phase will deliver.
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n .reg .pred p;\n GSR_W: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra GSR_D;\n bra GSR_W;\n GSR_D:\n}" ::"r"(
            smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// shared -> global, tracked by the issuing thread's bulk group
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src_smem, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src_smem)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// generic-proxy writes to shared memory become visible to the copy engine
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

#endif  // __CUDACC__

}  // namespace gsr
