// gsr_api.cu — the C ABI declared in include/gsrast_b200.h: argument validation, workspace layout,
// stage sequencing.  No device memory is allocated here and the host never waits on the device
// (except in debug mode, where every launch is followed by a stream synchronise, as the
// reference's `debug` flag does).
#include <atomic>
#include <cstring>
#include <mutex>
#include <vector>

#include "gsr_internal.cuh"

namespace gsr {

static std::atomic<int64_t> g_launches{0};
static thread_local cudaError_t t_last_error = cudaSuccess;

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
int cuda_fail(cudaError_t e) {
    t_last_error = e;
    cudaGetLastError();  // clear the sticky-less error state
    return GSR_ERR_CUDA;
}

cudaError_t once_per_device(unsigned long long* mask, cudaError_t (*setup)()) {
    static std::mutex mu;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return setup();  // unknown: set the attributes again
    std::lock_guard<std::mutex> lock(mu);   // held across setup(): a second thread's first launch waits for the opt-in
    const unsigned long long bit = 1ull << dev;
    if (*mask & bit) return cudaSuccess;
    const cudaError_t e = setup();
    if (e == cudaSuccess) *mask |= bit;     // a failed set-up is retried by the next call
    return e;
}

int sm_count() {
    static std::atomic<int> cached[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    int n = cached[dev].load(std::memory_order_relaxed);
    if (n == 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev].store(n, std::memory_order_relaxed);
    }
    return n;
}

// ---- per-stage timing: event pairs appended while enabled, summed by gsr_stage_timing_read ----
static std::atomic<bool> g_timing{false};
static std::mutex g_timing_mu;
struct StageMark { int stage; cudaEvent_t ev[2]; };
static std::vector<StageMark> g_marks;       // in use since the last read
static std::vector<cudaEvent_t> g_event_pool; // recycled events

bool stage_timing_on() { return g_timing.load(std::memory_order_relaxed); }

void stage_mark(int stage, cudaStream_t s, bool begin) {
    std::lock_guard<std::mutex> lock(g_timing_mu);
    auto get_event = []() {
        cudaEvent_t e = nullptr;
        if (!g_event_pool.empty()) { e = g_event_pool.back(); g_event_pool.pop_back(); }
        else cudaEventCreate(&e);
        return e;
    };
    if (begin) {
        StageMark m;
        m.stage = stage;
        m.ev[0] = get_event();
        m.ev[1] = get_event();
        cudaEventRecord(m.ev[0], s);
        g_marks.push_back(m);
    } else {
        for (size_t i = g_marks.size(); i-- > 0;)
            if (g_marks[i].stage == stage) { cudaEventRecord(g_marks[i].ev[1], s); break; }
    }
}

void compute_layout(int P, int W, int H, int64_t cap, gsr_layout* L) {
    std::memset(L, 0, sizeof(*L));
    const size_t p = (size_t)(P > 0 ? P : 1);
    const size_t c = (size_t)(cap > 0 ? cap : 1);
    const Grid g = make_grid(W, H);
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes, 256); return at; };
    L->g_depths = take(4 * p);
    L->g_means2D = take(8 * p);
    L->g_conic_opacity = take(16 * p);
    L->g_rgb = take(16 * p);
    L->g_tiles_touched = take(4 * p);
    L->g_info = take(sizeof(gsr_forward_info));
    L->g_grad_accum = take(48 * p);
    L->g_order_keys = take(4 * p);
    L->g_order = take(4 * p);
    L->g_order_keys_alt = take(4 * p);
    L->g_order_alt = take(4 * p);
    L->g_sort_ctl = take(4 * depth_sort_ctl_entries(p));
    L->geom_bytes = o;
    o = 0;
    const size_t kb = (size_t)tile_key_bytes(g.tiles);
    L->b_keys_unsorted = take(kb * c);
    L->b_vals_unsorted = take(4 * c);
    L->b_keys = take(kb * c);
    L->b_point_list = take(4 * c);
    L->b_keys_alt = take(kb * c);
    L->b_vals_alt = take(4 * c);
    L->b_sort_ctl = take(4 * tile_sort_ctl_entries(c, p, (size_t)g.tiles));
    L->binning_bytes = o;
    L->b_key_bytes = kb;
    o = 0;
    L->i_ranges = take(8 * (size_t)g.tiles);
    L->i_final_T = take(4 * (size_t)W * H);
    L->i_n_contrib = take(4 * (size_t)W * H);
    L->image_bytes = o;
}

template <typename T>
static T* at(void* ws, size_t off) { return reinterpret_cast<T*>(static_cast<char*>(ws) + off); }

GeomView geom_view(void* ws, const gsr_layout& L) {
    GeomView v;
    v.depths = at<float>(ws, L.g_depths);
    v.means2D = at<float2>(ws, L.g_means2D);
    v.conic_opacity = at<float4>(ws, L.g_conic_opacity);
    v.rgb = at<float4>(ws, L.g_rgb);
    v.tiles_touched = at<uint32_t>(ws, L.g_tiles_touched);
    v.info = at<gsr_forward_info>(ws, L.g_info);
    v.grad_accum = at<float>(ws, L.g_grad_accum);
    v.order_keys = at<uint32_t>(ws, L.g_order_keys);
    v.order = at<uint32_t>(ws, L.g_order);
    v.order_keys_alt = at<uint32_t>(ws, L.g_order_keys_alt);
    v.order_alt = at<uint32_t>(ws, L.g_order_alt);
    v.sort_ctl = at<uint32_t>(ws, L.g_sort_ctl);
    return v;
}
BinView bin_view(void* ws, const gsr_layout& L) {
    BinView v;
    v.keys_unsorted = at<char>(ws, L.b_keys_unsorted);
    v.vals_unsorted = at<uint32_t>(ws, L.b_vals_unsorted);
    v.keys = at<char>(ws, L.b_keys);
    v.point_list = at<uint32_t>(ws, L.b_point_list);
    v.keys_alt = at<char>(ws, L.b_keys_alt);
    v.vals_alt = at<uint32_t>(ws, L.b_vals_alt);
    v.sort_ctl = at<uint32_t>(ws, L.b_sort_ctl);
    return v;
}
ImageView image_view(void* ws, const gsr_layout& L) {
    ImageView v;
    v.ranges = at<uint2>(ws, L.i_ranges);
    v.final_T = at<float>(ws, L.i_final_T);
    v.n_contrib = at<uint32_t>(ws, L.i_n_contrib);
    return v;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

static int check_dims(int P, int W, int H, int64_t cap) {
    if (P < 0 || W <= 0 || H <= 0 || cap < 0) return GSR_ERR_INVALID_ARGUMENT;
    if (cap >= (1ll << 30) || P >= (1 << 30)) return GSR_ERR_UNSUPPORTED;  // 30-bit look-back counts
    const Grid g = make_grid(W, H);
    if (g.gx > 65535 || g.gy > 65535 || (int64_t)g.gx * g.gy > (1ll << 30)) return GSR_ERR_UNSUPPORTED;
    return GSR_OK;
}

}  // namespace gsr

using namespace gsr;

extern "C" {

int gsr_abi_version(void) { return GSR_ABI_VERSION; }

const char* gsr_status_string(int status) {
    switch (status) {
        case GSR_OK: return "ok";
        case GSR_ERR_INVALID_ARGUMENT: return "invalid argument";
        case GSR_ERR_CUDA: return "CUDA error";
        case GSR_ERR_WORKSPACE: return "workspace pointer null or misaligned";
        case GSR_ERR_UNSUPPORTED: return "unsupported size or configuration";
        case GSR_ERR_SCRATCH_LIMIT:
            return "input larger than the built-in scratch of this *_static entry point "
                   "(C*H*W > GSR_LOSS_STATIC_MAX_ELEMS): use the workspace form";
        default: return "unknown status";
    }
}

const char* gsr_last_cuda_error(void) {
    return t_last_error == cudaSuccess ? "" : cudaGetErrorString(t_last_error);
}

int64_t gsr_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int gsr_stage_count(void) { return kStageCount; }

const char* gsr_stage_name(int stage) {
    static const char* names[kStageCount] = {"preprocess", "depth_sort", "emit_keys", "tile_ranges", "radix_sort",
                                             "blend_forward", "blend_backward", "preprocess_backward"};
    return (stage >= 0 && stage < kStageCount) ? names[stage] : "";
}

int gsr_stage_timing_enable(int enable) {
    g_timing.store(enable != 0, std::memory_order_relaxed);
    return GSR_OK;
}

int gsr_stage_timing_read(double* ms_sum, int64_t* calls, int n) {
    if (!ms_sum || !calls || n < kStageCount) return GSR_ERR_INVALID_ARGUMENT;
    std::lock_guard<std::mutex> lock(g_timing_mu);
    for (int i = 0; i < n; ++i) { ms_sum[i] = 0.0; calls[i] = 0; }
    int st = GSR_OK;
    for (const StageMark& m : g_marks) {
        float ms = 0.f;
        cudaError_t e = cudaEventSynchronize(m.ev[1]);
        if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, m.ev[0], m.ev[1]);
        if (e == cudaSuccess) { ms_sum[m.stage] += ms; calls[m.stage] += 1; }
        else st = cuda_fail(e);
        g_event_pool.push_back(m.ev[0]);
        g_event_pool.push_back(m.ev[1]);
    }
    g_marks.clear();
    return st;
}

int gsr_workspace_layout(int P, int width, int height, int64_t instance_capacity, gsr_layout* out) {
    if (!out) return GSR_ERR_INVALID_ARGUMENT;
    int st = check_dims(P, width, height, instance_capacity);
    if (st != GSR_OK) return st;
    compute_layout(P, width, height, instance_capacity, out);
    return GSR_OK;
}

int gsr_workspace_bytes(int P, int width, int height, int64_t instance_capacity, size_t* geom_bytes,
                        size_t* binning_bytes, size_t* image_bytes) {
    gsr_layout L;
    int st = gsr_workspace_layout(P, width, height, instance_capacity, &L);
    if (st != GSR_OK) return st;
    if (geom_bytes) *geom_bytes = L.geom_bytes;
    if (binning_bytes) *binning_bytes = L.binning_bytes;
    if (image_bytes) *image_bytes = L.image_bytes;
    return GSR_OK;
}

static int forward_geometry_impl(const RawMode& rm, int P, int D, int M, int width, int height, const float* means3D, const float* shs,
                         const float* colors_precomp, const float* opacities, const float* scales,
                         float scale_modifier, const float* rotations, const float* cov3D_precomp,
                         const float* viewmatrix, const float* projmatrix, const float* campos, float tanfovx,
                         float tanfovy, int prefiltered, int32_t* radii, void* geom_ws, gsr_forward_info* h_info,
                         void* stream, int debug) {
    (void)prefiltered;  // accepted for API compatibility; the cull is always evaluated
    int st = check_dims(P, width, height, 0);
    if (st != GSR_OK) return st;
    if (!geom_ws || !aligned16(geom_ws)) return GSR_ERR_WORKSPACE;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    gsr_layout L;
    compute_layout(P, width, height, 0, &L);
    GeomView gv = geom_view(geom_ws, L);
    GSR_CUDA(cudaMemsetAsync(gv.info, 0, sizeof(gsr_forward_info), s));
    if (P > 0) {
        if (!means3D || !opacities || !radii || !viewmatrix || !projmatrix || !campos) return GSR_ERR_INVALID_ARGUMENT;
        if ((shs == nullptr) == (colors_precomp == nullptr)) return GSR_ERR_INVALID_ARGUMENT;
        const bool has_sr = scales != nullptr && rotations != nullptr;
        if (has_sr == (cov3D_precomp != nullptr)) return GSR_ERR_INVALID_ARGUMENT;
        if (!has_sr && (scales || rotations)) return GSR_ERR_INVALID_ARGUMENT;
        if (shs) {
            if (D < 0 || D > 3 || M < (D + 1) * (D + 1)) return GSR_ERR_INVALID_ARGUMENT;
            if (M > 16) return GSR_ERR_UNSUPPORTED;
        }
        if (!aligned16(means3D) || !aligned16(opacities) || !aligned16(shs) || !aligned16(scales) || !aligned16(rotations) ||
            !aligned16(cov3D_precomp) || !aligned16(colors_precomp))
            return GSR_ERR_INVALID_ARGUMENT;
        const Grid g = make_grid(width, height);
        st = launch_depth_ctl_reset(P, gv, s);
        if (st != GSR_OK) return st;
        bool counts_done = false;   // the persistent preprocess kernel takes the depth histograms and counts itself
        st = launch_preprocess(P, D, M, g, means3D, shs, colors_precomp, opacities, scales, scale_modifier,
                               rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tanfovx, tanfovy, radii,
                               gv, s, debug, rm, &counts_done);
        if (st != GSR_OK) return st;
        st = launch_depth_order(P, gv, s, debug, counts_done);
        if (st != GSR_OK) return st;
    }
    if (h_info) GSR_CUDA(cudaMemcpyAsync(h_info, gv.info, sizeof(gsr_forward_info), cudaMemcpyDeviceToHost, s));
    return GSR_OK;
}

int gsr_forward_render(int P, int width, int height, const float* bg, const int32_t* radii, void* geom_ws,
                       void* binning_ws, int64_t instance_capacity, void* image_ws, float* out_color,
                       gsr_forward_info* h_info, void* stream, int debug) {
    int st = check_dims(P, width, height, instance_capacity);
    if (st != GSR_OK) return st;
    if (!geom_ws || !image_ws || !aligned16(geom_ws) || !aligned16(image_ws)) return GSR_ERR_WORKSPACE;
    if (instance_capacity > 0 && (!binning_ws || !aligned16(binning_ws))) return GSR_ERR_WORKSPACE;
    if (!bg || !out_color || (P > 0 && !radii)) return GSR_ERR_INVALID_ARGUMENT;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    gsr_layout L;
    compute_layout(P, width, height, instance_capacity, &L);
    GeomView gv = geom_view(geom_ws, L);
    BinView bv = bin_view(binning_ws, L);
    ImageView iv = image_view(image_ws, L);
    const Grid g = make_grid(width, height);
    if (P > 0) {
        st = launch_binning(P, g, radii, gv, bv, instance_capacity, iv, s, debug);
        if (st != GSR_OK) return st;
    } else {
        GSR_CUDA(cudaMemsetAsync(iv.ranges, 0, sizeof(uint2) * (size_t)g.tiles, s));
    }
    st = launch_blend_forward(g, bg, gv, bv, iv, out_color, s, debug);
    if (st != GSR_OK) return st;
    if (h_info) GSR_CUDA(cudaMemcpyAsync(h_info, gv.info, sizeof(gsr_forward_info), cudaMemcpyDeviceToHost, s));
    return GSR_OK;
}

static int backward_impl(const RawMode& rm, int P, int D, int M, int width, int height, const float* bg, const float* means3D,
                 const float* shs, const float* colors_precomp, const float* scales, float scale_modifier,
                 const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                 const float* projmatrix, const float* campos, float tanfovx, float tanfovy, const int32_t* radii,
                 void* geom_ws, void* binning_ws, int64_t instance_capacity, void* image_ws, const float* dL_dpix,
                 float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh,
                 float* dL_dscales, float* dL_drots, void* stream, int debug) {
    int st = check_dims(P, width, height, instance_capacity);
    if (st != GSR_OK) return st;
    if (P == 0) return GSR_OK;
    if (!geom_ws || !image_ws || !binning_ws || !aligned16(geom_ws) || !aligned16(image_ws) || !aligned16(binning_ws))
        return GSR_ERR_WORKSPACE;
    if (!bg || !means3D || !viewmatrix || !projmatrix || !campos || !radii || !dL_dpix || !dL_dmeans2D ||
        !dL_dopacity || !dL_dmeans3D)
        return GSR_ERR_INVALID_ARGUMENT;
    if ((shs == nullptr) == (colors_precomp == nullptr)) return GSR_ERR_INVALID_ARGUMENT;
    const bool has_sr = scales != nullptr && rotations != nullptr;
    if (has_sr == (cov3D_precomp != nullptr)) return GSR_ERR_INVALID_ARGUMENT;
    if (shs && (D < 0 || D > 3 || M < (D + 1) * (D + 1) || M > 16)) return GSR_ERR_INVALID_ARGUMENT;
    if (has_sr && (!dL_dscales || !dL_drots)) return GSR_ERR_INVALID_ARGUMENT;
    if (!aligned16(dL_dsh) || !aligned16(dL_drots) || !aligned16(dL_dcov3D)) return GSR_ERR_INVALID_ARGUMENT;
    // the persistent backward kernels fetch these with bulk copies (16-byte aligned source addresses)
    if (!aligned16(means3D) || !aligned16(shs) || !aligned16(scales) || !aligned16(rotations))
        return GSR_ERR_INVALID_ARGUMENT;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    gsr_layout L;
    compute_layout(P, width, height, instance_capacity, &L);
    GeomView gv = geom_view(geom_ws, L);
    BinView bv = bin_view(binning_ws, L);
    ImageView iv = image_view(image_ws, L);
    const Grid g = make_grid(width, height);
    st = launch_blend_backward(P, g, bg, gv, bv, iv, dL_dpix, s, debug);
    if (st != GSR_OK) return st;
    return launch_preprocess_backward(P, D, M, g, means3D, shs, colors_precomp != nullptr, scales, scale_modifier,
                                      rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tanfovx, tanfovy,
                                      radii, gv, dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D,
                                      dL_dsh, dL_dscales, dL_drots, 0, s, debug, rm);
}

int gsr_forward_geometry(int P, int D, int M, int width, int height, const float* means3D, const float* shs,
                         const float* colors_precomp, const float* opacities, const float* scales,
                         float scale_modifier, const float* rotations, const float* cov3D_precomp,
                         const float* viewmatrix, const float* projmatrix, const float* campos, float tanfovx,
                         float tanfovy, int prefiltered, int32_t* radii, void* geom_ws, gsr_forward_info* h_info,
                         void* stream, int debug) {
    return forward_geometry_impl(RawMode(), P, D, M, width, height, means3D, shs, colors_precomp, opacities, scales,
                                 scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tanfovx,
                                 tanfovy, prefiltered, radii, geom_ws, h_info, stream, debug);
}

int gsr_backward(int P, int D, int M, int width, int height, const float* bg, const float* means3D,
                 const float* shs, const float* colors_precomp, const float* scales, float scale_modifier,
                 const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                 const float* projmatrix, const float* campos, float tanfovx, float tanfovy, const int32_t* radii,
                 void* geom_ws, void* binning_ws, int64_t instance_capacity, void* image_ws, const float* dL_dpix,
                 float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D,
                 float* dL_dsh, float* dL_dscales, float* dL_drots, void* stream, int debug) {
    return backward_impl(RawMode(), P, D, M, width, height, bg, means3D, shs, colors_precomp, scales, scale_modifier,
                         rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tanfovx, tanfovy, radii, geom_ws,
                         binning_ws, instance_capacity, image_ws, dL_dpix, dL_dmeans2D, dL_dcolors, dL_dopacity,
                         dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drots, stream, debug);
}

// ---- raw-parameter entries -----------------------------------------------------
static bool raw_args_ok(int M, const float* features_dc, const float* features_rest, const float* opacity_raw,
                        const float* scaling_raw, const float* rotation_raw) {
    if (!features_dc || !opacity_raw || !scaling_raw || !rotation_raw) return false;
    if (M < 1 || M > 16 || (M > 1 && !features_rest)) return false;
    return aligned16(features_dc) && aligned16(features_rest) && aligned16(opacity_raw) && aligned16(scaling_raw) &&
           aligned16(rotation_raw);
}

int gsr_forward_geometry_raw(int P, int D, int M, int width, int height, const float* means3D,
                             const float* features_dc, const float* features_rest, const float* opacity_raw,
                             const float* scaling_raw, float scale_modifier, const float* rotation_raw,
                             const float* viewmatrix, const float* projmatrix, const float* campos, float tanfovx,
                             float tanfovy, int32_t* radii, void* geom_ws, gsr_forward_info* h_info, void* stream,
                             int debug) {
    if (P > 0 && !raw_args_ok(M, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw))
        return GSR_ERR_INVALID_ARGUMENT;
    RawMode rm;
    rm.raw = 1;
    rm.sh_rest = M > 1 ? features_rest : features_dc;  // non-null selects the split fetch; unused when M == 1
    return forward_geometry_impl(rm, P, D, M, width, height, means3D, features_dc, nullptr, opacity_raw, scaling_raw,
                                 scale_modifier, rotation_raw, nullptr, viewmatrix, projmatrix, campos, tanfovx,
                                 tanfovy, 0, radii, geom_ws, h_info, stream, debug);
}

int gsr_backward_raw(int P, int D, int M, int width, int height, const float* bg, const float* means3D,
                     const float* features_dc, const float* features_rest, const float* scaling_raw,
                     float scale_modifier, const float* rotation_raw, const float* viewmatrix,
                     const float* projmatrix, const float* campos, float tanfovx, float tanfovy, const int32_t* radii,
                     void* geom_ws, void* binning_ws, int64_t instance_capacity, void* image_ws, const float* dL_dpix,
                     float* dL_dmeans2D, float* dL_dopacity_raw, float* dL_dmeans3D, float* dL_dfeatures_dc,
                     float* dL_dfeatures_rest, float* dL_dscaling_raw, float* dL_drotation_raw, void* stream,
                     int debug) {
    if (P > 0 && (!features_dc || !scaling_raw || !rotation_raw || (M > 1 && !features_rest) || !dL_dfeatures_dc ||
                  (M > 1 && !dL_dfeatures_rest)))
        return GSR_ERR_INVALID_ARGUMENT;
    RawMode rm;
    rm.raw = 1;
    rm.sh_rest = M > 1 ? features_rest : features_dc;
    rm.dL_dsh_rest = M > 1 ? dL_dfeatures_rest : dL_dfeatures_dc;
    return backward_impl(rm, P, D, M, width, height, bg, means3D, features_dc, nullptr, scaling_raw, scale_modifier,
                         rotation_raw, nullptr, viewmatrix, projmatrix, campos, tanfovx, tanfovy, radii, geom_ws,
                         binning_ws, instance_capacity, image_ws, dL_dpix, dL_dmeans2D, nullptr, dL_dopacity_raw,
                         dL_dmeans3D, nullptr, dL_dfeatures_dc, dL_dscaling_raw, dL_drotation_raw, stream, debug);
}

int gsr_activate(int P, int M, const float* features_dc, const float* features_rest, const float* opacity_raw,
                 const float* scaling_raw, const float* rotation_raw, float* shs, float* opacities, float* scales,
                 float* rotations, void* stream) {
    if (P < 0) return GSR_ERR_INVALID_ARGUMENT;
    if (P == 0) return GSR_OK;
    if (!raw_args_ok(M, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw) || !shs || !opacities ||
        !scales || !rotations)
        return GSR_ERR_INVALID_ARGUMENT;
    return launch_activate(P, M, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw, shs, opacities,
                           scales, rotations, static_cast<cudaStream_t>(stream));
}

// ---- Gaussian-sharded multi-view entries ---------------------------------------------------
static int check_slots(int P_ws, int slot_offset, int Ps) {
    if (P_ws < 0 || Ps < 0 || slot_offset < 0 || (int64_t)slot_offset + Ps > P_ws) return GSR_ERR_INVALID_ARGUMENT;
    if (Ps == 0) return GSR_OK;  // an empty trailing slice may start at an unaligned P: nothing is addressed
    if (slot_offset & 3) return GSR_ERR_INVALID_ARGUMENT;  // keeps the 16-byte row alignment of every array
    return GSR_OK;
}

int gsr_preprocess_slots(int P_ws, int slot_offset, int Ps, int D, int M, int width, int height,
                         const float* means3D, const float* shs, const float* colors_precomp,
                         const float* opacities, const float* scales, float scale_modifier,
                         const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                         const float* projmatrix, const float* campos, float tanfovx, float tanfovy,
                         int32_t* radii_slots, void* geom_ws, void* stream, int debug) {
    int st = check_dims(P_ws, width, height, 0);
    if (st == GSR_OK) st = check_slots(P_ws, slot_offset, Ps);
    if (st != GSR_OK) return st;
    if (Ps == 0) return GSR_OK;
    if (!geom_ws || !aligned16(geom_ws)) return GSR_ERR_WORKSPACE;
    if (!means3D || !opacities || !radii_slots || !viewmatrix || !projmatrix || !campos) return GSR_ERR_INVALID_ARGUMENT;
    if ((shs == nullptr) == (colors_precomp == nullptr)) return GSR_ERR_INVALID_ARGUMENT;
    const bool has_sr = scales != nullptr && rotations != nullptr;
    if (has_sr == (cov3D_precomp != nullptr)) return GSR_ERR_INVALID_ARGUMENT;
    if (shs && (D < 0 || D > 3 || M < (D + 1) * (D + 1) || M > 16)) return GSR_ERR_INVALID_ARGUMENT;
    if (!aligned16(means3D) || !aligned16(opacities) || !aligned16(shs) || !aligned16(scales) || !aligned16(rotations) ||
        !aligned16(cov3D_precomp) || !aligned16(colors_precomp))
        return GSR_ERR_INVALID_ARGUMENT;
    gsr_layout L;
    compute_layout(P_ws, width, height, 0, &L);
    const GeomView gv = offset_rows(geom_view(geom_ws, L), (size_t)slot_offset);
    return launch_preprocess(Ps, D, M, make_grid(width, height), means3D, shs, colors_precomp, opacities, scales,
                             scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tanfovx,
                             tanfovy, radii_slots, gv, static_cast<cudaStream_t>(stream), debug, RawMode());
}

int gsr_depth_order(int P, int width, int height, void* geom_ws, gsr_forward_info* h_info, void* stream,
                    int debug) {
    int st = check_dims(P, width, height, 0);
    if (st != GSR_OK) return st;
    if (!geom_ws || !aligned16(geom_ws)) return GSR_ERR_WORKSPACE;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    gsr_layout L;
    compute_layout(P, width, height, 0, &L);
    GeomView gv = geom_view(geom_ws, L);
    GSR_CUDA(cudaMemsetAsync(gv.info, 0, sizeof(gsr_forward_info), s));
    if (P > 0) {
        st = launch_depth_ctl_reset(P, gv, s);
        if (st == GSR_OK) st = launch_depth_order(P, gv, s, debug);
        if (st != GSR_OK) return st;
    }
    if (h_info) GSR_CUDA(cudaMemcpyAsync(h_info, gv.info, sizeof(gsr_forward_info), cudaMemcpyDeviceToHost, s));
    return GSR_OK;
}

int gsr_blend_backward(int P, int width, int height, const float* bg, void* geom_ws, void* binning_ws,
                       int64_t instance_capacity, void* image_ws, const float* dL_dpix, void* stream, int debug) {
    int st = check_dims(P, width, height, instance_capacity);
    if (st != GSR_OK) return st;
    if (P == 0) return GSR_OK;
    if (!geom_ws || !image_ws || !binning_ws || !aligned16(geom_ws) || !aligned16(image_ws) || !aligned16(binning_ws))
        return GSR_ERR_WORKSPACE;
    if (!bg || !dL_dpix) return GSR_ERR_INVALID_ARGUMENT;
    gsr_layout L;
    compute_layout(P, width, height, instance_capacity, &L);
    return launch_blend_backward(P, make_grid(width, height), bg, geom_view(geom_ws, L), bin_view(binning_ws, L),
                                 image_view(image_ws, L), dL_dpix, static_cast<cudaStream_t>(stream), debug);
}

int gsr_preprocess_backward_slots(int P_ws, int slot_offset, int Ps, int D, int M, int width, int height,
                                  const float* means3D, const float* shs, const float* colors_precomp,
                                  const float* scales, float scale_modifier, const float* rotations,
                                  const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                                  const float* campos, float tanfovx, float tanfovy, const int32_t* radii_slots,
                                  void* geom_ws, float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity,
                                  float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscales,
                                  float* dL_drots, int accumulate, void* stream, int debug) {
    int st = check_dims(P_ws, width, height, 0);
    if (st == GSR_OK) st = check_slots(P_ws, slot_offset, Ps);
    if (st != GSR_OK) return st;
    if (Ps == 0) return GSR_OK;
    if (!geom_ws || !aligned16(geom_ws)) return GSR_ERR_WORKSPACE;
    if (!means3D || !viewmatrix || !projmatrix || !campos || !radii_slots || !dL_dmeans2D || !dL_dopacity || !dL_dmeans3D)
        return GSR_ERR_INVALID_ARGUMENT;
    if ((shs == nullptr) == (colors_precomp == nullptr)) return GSR_ERR_INVALID_ARGUMENT;
    const bool has_sr = scales != nullptr && rotations != nullptr;
    if (has_sr == (cov3D_precomp != nullptr)) return GSR_ERR_INVALID_ARGUMENT;
    if (shs && (D < 0 || D > 3 || M < (D + 1) * (D + 1) || M > 16)) return GSR_ERR_INVALID_ARGUMENT;
    if (has_sr && (!dL_dscales || !dL_drots)) return GSR_ERR_INVALID_ARGUMENT;
    if (!aligned16(dL_dsh) || !aligned16(dL_drots) || !aligned16(dL_dcov3D)) return GSR_ERR_INVALID_ARGUMENT;
    // the persistent backward kernels fetch these with bulk copies (16-byte aligned source addresses)
    if (!aligned16(means3D) || !aligned16(shs) || !aligned16(scales) || !aligned16(rotations))
        return GSR_ERR_INVALID_ARGUMENT;
    gsr_layout L;
    compute_layout(P_ws, width, height, 0, &L);
    const GeomView gv = offset_rows(geom_view(geom_ws, L), (size_t)slot_offset);
    return launch_preprocess_backward(Ps, D, M, make_grid(width, height), means3D, shs, colors_precomp != nullptr,
                                      scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix,
                                      campos, tanfovx, tanfovy, radii_slots, gv, dL_dmeans2D, dL_dcolors, dL_dopacity,
                                      dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drots, accumulate != 0,
                                      static_cast<cudaStream_t>(stream), debug, RawMode());
}

int gsr_preprocess_views(int P_ws, int slot_offset, int Ps, int D, int M, int width, int height,
                         const float* means3D, const float* shs, const float* colors_precomp,
                         const float* opacities, const float* scales, float scale_modifier, const float* rotations,
                         const float* cov3D_precomp, int n_views, const gsr_view_slot* views, void* stream,
                         int debug) {
    if (!views || n_views < 1 || n_views > GSR_MAX_VIEWS) return GSR_ERR_INVALID_ARGUMENT;
    int st = check_dims(P_ws, width, height, 0);
    if (st == GSR_OK) st = check_slots(P_ws, slot_offset, Ps);
    if (st != GSR_OK) return st;
    if (Ps == 0) return GSR_OK;
    if (!means3D || !opacities) return GSR_ERR_INVALID_ARGUMENT;
    if ((shs == nullptr) == (colors_precomp == nullptr)) return GSR_ERR_INVALID_ARGUMENT;
    const bool has_sr = scales != nullptr && rotations != nullptr;
    if (has_sr == (cov3D_precomp != nullptr)) return GSR_ERR_INVALID_ARGUMENT;
    if (shs && (D < 0 || D > 3 || M < (D + 1) * (D + 1) || M > 16)) return GSR_ERR_INVALID_ARGUMENT;
    if (!aligned16(means3D) || !aligned16(opacities) || !aligned16(shs) || !aligned16(scales) || !aligned16(rotations) ||
        !aligned16(cov3D_precomp) || !aligned16(colors_precomp))
        return GSR_ERR_INVALID_ARGUMENT;
    gsr_layout L;
    compute_layout(P_ws, width, height, 0, &L);
    ForwardViewArgs args[GSR_MAX_VIEWS];
    for (int v = 0; v < n_views; ++v) {
        const gsr_view_slot& a = views[v];
        if (!a.viewmatrix || !a.projmatrix || !a.campos || !a.geom_ws || !a.radii_slots) return GSR_ERR_INVALID_ARGUMENT;
        if (!aligned16(a.geom_ws)) return GSR_ERR_WORKSPACE;
        args[v].view = a.viewmatrix; args[v].proj = a.projmatrix; args[v].campos = a.campos;
        args[v].tanfovx = a.tanfovx; args[v].tanfovy = a.tanfovy;
        args[v].radii = a.radii_slots;
        args[v].gv = offset_rows(geom_view(a.geom_ws, L), (size_t)slot_offset);
        if ((a.radii_local == nullptr) != (a.conic_local == nullptr) || !aligned16(a.conic_local))
            return GSR_ERR_INVALID_ARGUMENT;
        args[v].radii_local = a.radii_local;
        args[v].conic_local = reinterpret_cast<float4*>(a.conic_local);
    }
    return launch_preprocess_views(Ps, D, M, make_grid(width, height), means3D, shs, colors_precomp, opacities, scales,
                                   scale_modifier, rotations, cov3D_precomp, n_views, args,
                                   static_cast<cudaStream_t>(stream), debug);
}

int gsr_preprocess_backward_views(int P_ws, int slot_offset, int Ps, int D, int M, int width, int height,
                                  const float* means3D, const float* shs, const float* colors_precomp,
                                  const float* scales, float scale_modifier, const float* rotations,
                                  const float* cov3D_precomp, int n_views, const gsr_view_slot* views,
                                  float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D,
                                  float* dL_dsh, float* dL_dscales, float* dL_drots, int accumulate, void* stream,
                                  int debug) {
    if (!views || n_views < 1 || n_views > GSR_MAX_VIEWS) return GSR_ERR_INVALID_ARGUMENT;
    int st = check_dims(P_ws, width, height, 0);
    if (st == GSR_OK) st = check_slots(P_ws, slot_offset, Ps);
    if (st != GSR_OK) return st;
    if (Ps == 0) return GSR_OK;
    if (!means3D || !dL_dopacity || !dL_dmeans3D) return GSR_ERR_INVALID_ARGUMENT;
    if ((shs == nullptr) == (colors_precomp == nullptr)) return GSR_ERR_INVALID_ARGUMENT;
    const bool has_sr = scales != nullptr && rotations != nullptr;
    if (has_sr == (cov3D_precomp != nullptr)) return GSR_ERR_INVALID_ARGUMENT;
    if (shs && (D < 0 || D > 3 || M < (D + 1) * (D + 1) || M > 16)) return GSR_ERR_INVALID_ARGUMENT;
    if (has_sr && (!dL_dscales || !dL_drots)) return GSR_ERR_INVALID_ARGUMENT;
    if (!aligned16(dL_dsh) || !aligned16(dL_drots) || !aligned16(dL_dcov3D)) return GSR_ERR_INVALID_ARGUMENT;
    // the persistent backward kernels fetch these with bulk copies (16-byte aligned source addresses)
    if (!aligned16(means3D) || !aligned16(shs) || !aligned16(scales) || !aligned16(rotations))
        return GSR_ERR_INVALID_ARGUMENT;
    gsr_layout L;
    compute_layout(P_ws, width, height, 0, &L);
    BackwardViewArgs args[GSR_MAX_VIEWS];
    for (int v = 0; v < n_views; ++v) {
        const gsr_view_slot& a = views[v];
        if (!a.viewmatrix || !a.projmatrix || !a.campos || !a.geom_ws || !a.radii_slots || !a.dL_dmeans2D)
            return GSR_ERR_INVALID_ARGUMENT;
        if (!aligned16(a.geom_ws)) return GSR_ERR_WORKSPACE;
        args[v].view = a.viewmatrix; args[v].proj = a.projmatrix; args[v].campos = a.campos;
        args[v].tanfovx = a.tanfovx; args[v].tanfovy = a.tanfovy;
        args[v].radii = a.radii_slots;
        args[v].gv = offset_rows(geom_view(a.geom_ws, L), (size_t)slot_offset);
        args[v].dL_dmeans2D = a.dL_dmeans2D;
        if ((a.radii_local == nullptr) != (a.conic_local == nullptr) || !aligned16(a.conic_local))
            return GSR_ERR_INVALID_ARGUMENT;
        args[v].radii_local = a.radii_local;
        args[v].conic_local = reinterpret_cast<const float4*>(a.conic_local);
    }
    return launch_preprocess_backward_views(Ps, D, M, make_grid(width, height), means3D, shs, colors_precomp != nullptr,
                                            scales, scale_modifier, rotations, cov3D_precomp, n_views, args, dL_dcolors,
                                            dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drots,
                                            accumulate, static_cast<cudaStream_t>(stream), debug);
}

int gsr_adam_step(int n_groups, const gsr_adam_group* groups, double beta1, double beta2, double eps, int step,
                  void* stream) {
    if (!groups) return GSR_ERR_INVALID_ARGUMENT;
    return launch_adam(n_groups, groups, beta1, beta2, eps, step, static_cast<cudaStream_t>(stream));
}

int gsr_densification_stats(int P, const int32_t* radii, const float* dL_dmeans2D, float* max_radii2D,
                            float* xyz_gradient_accum, float* denom, void* stream) {
    if (P < 0) return GSR_ERR_INVALID_ARGUMENT;
    if (P == 0) return GSR_OK;
    if (!radii || !dL_dmeans2D || !max_radii2D || !xyz_gradient_accum || !denom) return GSR_ERR_INVALID_ARGUMENT;
    return launch_densification_stats(P, radii, dL_dmeans2D, max_radii2D, xyz_gradient_accum, denom,
                                      static_cast<cudaStream_t>(stream));
}

static int check_loss_dims(int C, int H, int W) {
    if (C <= 0 || H <= 0 || W <= 0) return GSR_ERR_INVALID_ARGUMENT;
    if ((long long)C * H * W >= (1ll << 31) || C > 65535 || (H + 31) / 32 > 65535) return GSR_ERR_UNSUPPORTED;
    return GSR_OK;
}

int gsr_loss_workspace_bytes(int C, int H, int W, int save_for_backward, size_t* bytes) {
    if (!bytes) return GSR_ERR_INVALID_ARGUMENT;
    int st = check_loss_dims(C, H, W);
    if (st != GSR_OK) return st;
    *bytes = loss_workspace_bytes(C, H, W, save_for_backward != 0);
    return GSR_OK;
}

int gsr_loss_forward(int C, int H, int W, const float* image, const float* gt, double a0, double a_l1, double a_ssim,
                     void* workspace, int save_for_backward, float* out3, void* stream) {
    int st = check_loss_dims(C, H, W);
    if (st != GSR_OK) return st;
    if (!image || !gt || !out3) return GSR_ERR_INVALID_ARGUMENT;
    if (!workspace || (reinterpret_cast<uintptr_t>(workspace) & 255)) return GSR_ERR_WORKSPACE;
    return launch_loss_forward(C, H, W, image, gt, a0, a_l1, a_ssim, workspace, save_for_backward != 0, out3,
                               static_cast<cudaStream_t>(stream));
}

int gsr_loss_backward(int C, int H, int W, const float* image, const float* gt, double a_l1, double a_ssim,
                      const void* workspace, const float* upstream, float* dL_dimage, void* stream) {
    int st = check_loss_dims(C, H, W);
    if (st != GSR_OK) return st;
    if (!image || !gt || !dL_dimage) return GSR_ERR_INVALID_ARGUMENT;
    if (!workspace || (reinterpret_cast<uintptr_t>(workspace) & 255)) return GSR_ERR_WORKSPACE;
    return launch_loss_backward(C, H, W, image, gt, a_l1, a_ssim, workspace, upstream, dL_dimage,
                                static_cast<cudaStream_t>(stream));
}

static int check_ply(int64_t P, const gsr_ply_columns* c, const void* rows) {
    if (P < 0 || !c) return GSR_ERR_INVALID_ARGUMENT;
    if (c->n_props < 14 || c->n_props > GSR_PLY_MAX_PROPS) return GSR_ERR_UNSUPPORTED;
    if (c->n_rest != 0 && c->n_rest != 9 && c->n_rest != 24 && c->n_rest != 45) return GSR_ERR_INVALID_ARGUMENT;
    if (P > 0 && (!rows || (reinterpret_cast<uintptr_t>(rows) & 15))) return GSR_ERR_INVALID_ARGUMENT;
    if (P >= (1ll << 30)) return GSR_ERR_UNSUPPORTED;
    return GSR_OK;
}

int gsr_ply_unpack(int64_t P, const gsr_ply_columns* cols, const float* rows, float* xyz, float* f_dc, float* f_rest,
                   float* opacity, float* scale, float* rot, void* stream) {
    int st = check_ply(P, cols, rows);
    if (st != GSR_OK || P == 0) return st;
    return launch_ply(false, P, cols, const_cast<float*>(rows), xyz, f_dc, f_rest, opacity, scale, rot,
                      static_cast<cudaStream_t>(stream));
}

int gsr_ply_pack(int64_t P, const gsr_ply_columns* cols, const float* xyz, const float* f_dc, const float* f_rest,
                 const float* opacity, const float* scale, const float* rot, float* rows, void* stream) {
    int st = check_ply(P, cols, rows);
    if (st != GSR_OK || P == 0) return st;
    return launch_ply(true, P, cols, rows, const_cast<float*>(xyz), const_cast<float*>(f_dc),
                      const_cast<float*>(f_rest), const_cast<float*>(opacity), const_cast<float*>(scale),
                      const_cast<float*>(rot), static_cast<cudaStream_t>(stream));
}

int gsr_blend_pair_counts(int P, int width, int height, void* geom_ws, void* binning_ws,
                           int64_t instance_capacity, void* image_ws, uint64_t* d_counters, void* stream) {
    int st = check_dims(P, width, height, instance_capacity);
    if (st != GSR_OK) return st;
    if (!geom_ws || !image_ws || !binning_ws || !d_counters) return GSR_ERR_WORKSPACE;
    gsr_layout L;
    compute_layout(P, width, height, instance_capacity, &L);
    return launch_blend_pair_counts(make_grid(width, height), geom_view(geom_ws, L), bin_view(binning_ws, L),
                                    image_view(image_ws, L), reinterpret_cast<unsigned long long*>(d_counters),
                                    static_cast<cudaStream_t>(stream));
}

int gsr_materialize_keys(int P, int width, int height, void* geom_ws, void* binning_ws, int64_t instance_capacity,
                         int sorted, uint64_t* d_keys, void* stream) {
    int st = check_dims(P, width, height, instance_capacity);
    if (st != GSR_OK) return st;
    if (!geom_ws || !binning_ws || !aligned16(geom_ws) || !aligned16(binning_ws)) return GSR_ERR_WORKSPACE;
    if (!d_keys) return GSR_ERR_INVALID_ARGUMENT;
    gsr_layout L;
    compute_layout(P, width, height, instance_capacity, &L);
    return launch_materialize_keys(P, make_grid(width, height), geom_view(geom_ws, L), bin_view(binning_ws, L),
                                   instance_capacity, sorted != 0, d_keys, static_cast<cudaStream_t>(stream));
}

int gsr_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* stream) {
    (void)projmatrix;  // the reference's test uses only the view-space depth
    if (P < 0) return GSR_ERR_INVALID_ARGUMENT;
    if (P == 0) return GSR_OK;
    if (!means3D || !viewmatrix || !present) return GSR_ERR_INVALID_ARGUMENT;
    return launch_mark_visible(P, means3D, viewmatrix, present, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
