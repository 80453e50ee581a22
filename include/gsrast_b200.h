/*
 * gsrast_b200.h — C ABI of the B200-native differentiable Gaussian-splatting rasterizer.
 *
 * This is the drop-in boundary for the ONE hot path of ChikaYan/mvs_gaussian_splatting:
 * the rasterizer that /root/reference/gaussian_renderer/__init__.py:15 imports
 * (`from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer`),
 * constructs at :42-57 and calls at :257-265.  In the reference that module is a pybind11
 * extension `_C` taking torch::Tensor (un-vendored submodule, .gitmodules:4-6); its three
 * entry points `_C.rasterize_gaussians`, `_C.rasterize_gaussians_backward`, `_C.mark_visible`
 * are what the functions below replace.  Plain `extern "C"`: raw DEVICE pointers, explicit
 * sizes, an explicit cudaStream_t (passed as void*), an integer status return, no exceptions,
 * no global state, no torch types.
 *
 * Ownership: the caller owns every byte.  The library never allocates or frees device memory;
 * the three workspaces (geometry / binning / image — the reference's geomBuffer /
 * binningBuffer / imgBuffer) are sized with gsr_workspace_bytes() and must stay alive, untouched,
 * from a forward call until the matching backward call has been enqueued.
 *
 * Conventions (same as the reference call site):
 *   means3D [P,3] f32 | shs [P,M,3] f32 (coefficient-major, RGB innermost) | colors_precomp [P,3]
 *   opacities [P] f32 | scales [P,3] | rotations [P,4] (w,x,y,z; used as given) | cov3D_precomp [P,6]
 *   viewmatrix / projmatrix: 16 f32, the reference's TRANSPOSED torch matrices read flat
 *   (scene/cameras.py:54-56), i.e. column-major maths matrices | campos [3] | bg [3]
 *   out_color [3,H,W] planar f32 | radii [P] i32 (0 = culled)
 * Exactly one of (shs, colors_precomp) and one of ((scales,rotations), cov3D_precomp) is non-NULL.
 * All pointers must be 16-byte aligned and dense.
 */
#ifndef GSRAST_B200_H_
#define GSRAST_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSR_ABI_VERSION 9
#define GSR_MAX_VIEWS 16   /* views one multi-view call can take */

#if defined(__GNUC__)
#define GSR_API __attribute__((visibility("default")))
#else
#define GSR_API
#endif

enum gsr_status {
    GSR_OK = 0,
    GSR_ERR_INVALID_ARGUMENT = 1, /* NULL / inconsistent pointers, negative sizes, bad degree */
    GSR_ERR_CUDA = 2,             /* a CUDA runtime call or (debug mode) a kernel failed     */
    GSR_ERR_WORKSPACE = 3,        /* workspace pointer NULL or misaligned                    */
    GSR_ERR_UNSUPPORTED = 4,      /* e.g. more than 2^16 tiles per axis, M > 16 coefficients */
    GSR_ERR_SCRATCH_LIMIT = 5     /* a *_static entry point was given more elements than its built-in
                                     scratch holds (GSR_LOSS_STATIC_MAX_ELEMS)                */
};

/* Host-visible result block of a forward call.  `h_info` arguments point at PINNED host memory
 * (or NULL); the library enqueues an async D2H copy of the device copy into it, so the values
 * are valid once the stream has reached that point (no host sync inside the library). */
typedef struct gsr_forward_info {
    int32_t num_rendered;  /* R = sum(tiles_touched): tile instances the view needs            */
    int32_t overflow;      /* 1 if R > instance_capacity: nothing was binned/blended; re-run
                              gsr_forward_render with a larger binning workspace               */
    int32_t num_visible;   /* Gaussians with radii > 0                                         */
    int32_t reserved;
} gsr_forward_info;

/* Byte offsets of the arrays inside the three workspaces (for tests, debugging and tools; the
 * layout is otherwise private).  Obtained from gsr_workspace_layout(). */
typedef struct gsr_layout {
    /* geometry workspace, per Gaussian */
    size_t g_depths;        /* f32[P]   view-space z; culled Gaussians hold the bits 0xFFFFFFFF  */
    size_t g_means2D;       /* f32[P,2] pixel centre                                           */
    size_t g_conic_opacity; /* f32[P,4] conic A,B,C + opacity                                  */
    size_t g_rgb;           /* f32[P,4] r,g,b + clamped bits (as u32 in .w, bit c = channel c) */
    size_t g_tiles_touched; /* u32[P]                                                          */
    size_t g_info;          /* gsr_forward_info (device copy)                                  */
    size_t g_grad_accum;    /* f32[P,12] backward blend accumulators                           */
    size_t g_order_keys;    /* u32[P]   depth bits of the num_visible visible Gaussians, ascending (stable) */
    size_t g_order;         /* u32[P]   Gaussian index of depth-order slot k                     */
    size_t g_order_keys_alt;/* u32[P]   sort ping-pong                                           */
    size_t g_order_alt;     /* u32[P]                                                            */
    size_t g_sort_ctl;      /* u32[...] histograms, tickets, look-back status                    */
    size_t geom_bytes;
    /* binning workspace, per tile instance (capacity entries each) */
    size_t b_keys_unsorted; /* K[C]    tile id of every instance, emitted in DEPTH order; K = u16 up to
                               65536 tiles, else u32 (b_key_bytes).  The reference's 64-bit key
                               (tile << 32) | depth bits is gsr_materialize_keys(): its low half is
                               depths[index], which no stage of the render path needs to move         */
    size_t b_vals_unsorted; /* u32[C]  Gaussian index                                          */
    size_t b_keys;          /* K[C]    sorted by tile (stable)                                 */
    size_t b_point_list;    /* u32[C]  sorted                                                  */
    size_t b_keys_alt;      /* K[C]    radix-sort ping-pong buffer                             */
    size_t b_vals_alt;      /* u32[C]                                                          */
    size_t b_sort_ctl;      /* u32[...] histograms, tickets, per-tile counts, look-back status   */
    size_t binning_bytes;
    size_t b_key_bytes;     /* sizeof(K): 2 or 4                                               */
    /* image workspace */
    size_t i_ranges;        /* u32[tiles,2] [start,end) into point_list                        */
    size_t i_final_T;       /* f32[H*W]                                                        */
    size_t i_n_contrib;     /* u32[H*W] 1-based list position of the last blended entry        */
    size_t image_bytes;
} gsr_layout;

GSR_API int gsr_abi_version(void);

/* Replaces: the exception text the reference's CHECK_CUDA macro raises. Never NULL. */
GSR_API const char* gsr_status_string(int status);
/* cudaGetErrorString of the last CUDA error this THREAD saw inside the library ("" if none). */
GSR_API const char* gsr_last_cuda_error(void);

/* Replaces: the three resize lambdas (`resizeFunctional`) the reference's glue hands to
 * CudaRasterizer::Rasterizer::forward — here the caller sizes and owns the buffers up front. */
GSR_API int gsr_workspace_bytes(int P, int width, int height, int64_t instance_capacity,
                        size_t* geom_bytes, size_t* binning_bytes, size_t* image_bytes);
GSR_API int gsr_workspace_layout(int P, int width, int height, int64_t instance_capacity, gsr_layout* out);

/* Forward, stage 1 (replaces the first half of `_C.rasterize_gaussians`: preprocess, the instance count the
 * reference takes with its InclusiveSum, and the per-Gaussian depth order the instances are emitted in).
 * Writes radii[P] and the geometry workspace; enqueues a copy of gsr_forward_info (num_rendered,
 * num_visible) into h_info if non-NULL.  D = active SH degree 0..3, M = coefficients per channel
 * in `shs` (>= (D+1)^2, <= 16). */
GSR_API int gsr_forward_geometry(int P, int D, int M, int width, int height, const float* means3D,
                         const float* shs, const float* colors_precomp, const float* opacities,
                         const float* scales, float scale_modifier, const float* rotations,
                         const float* cov3D_precomp, const float* viewmatrix,
                         const float* projmatrix, const float* campos, float tanfovx,
                         float tanfovy, int prefiltered, int32_t* radii, void* geom_ws,
                         gsr_forward_info* h_info, void* stream, int debug);

/* Forward, stage 2 (replaces the second half: duplicateWithKeys, radix sort, identifyTileRanges,
 * render).  Never syncs the host: R is read on the device; if R > instance_capacity the overflow
 * flag is raised, nothing is binned (out_color then holds the background only), and the caller
 * re-runs this stage with a larger binning workspace (the geometry workspace stays valid). */
GSR_API int gsr_forward_render(int P, int width, int height, const float* bg, const int32_t* radii,
                       void* geom_ws, void* binning_ws, int64_t instance_capacity, void* image_ws,
                       float* out_color, gsr_forward_info* h_info, void* stream, int debug);

/* Backward (replaces `_C.rasterize_gaussians_backward`).  Needs the inputs and the three
 * workspaces of the matching forward.  EVERY output array is fully written by the library
 * (zeros where the reference would leave its zero-initialised tensors untouched); the caller
 * passes uninitialised memory.  Outputs that do not apply may be NULL:
 *   dL_dsh (colors_precomp path), dL_dscales/dL_drots (cov3D_precomp path).
 * dL_dmeans2D is [P,3] (z = 0) in the reference's NDC-scaled convention
 * (scene/gaussian_model.py:775-777 reads [:, :2]).  dL_dcolors [P,3], dL_dopacity [P],
 * dL_dmeans3D [P,3], dL_dcov3D [P,6], dL_dsh [P,M,3], dL_dscales [P,3], dL_drots [P,4]. */
GSR_API int gsr_backward(int P, int D, int M, int width, int height, const float* bg,
                 const float* means3D, const float* shs, const float* colors_precomp,
                 const float* scales, float scale_modifier, const float* rotations,
                 const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                 const float* campos, float tanfovx, float tanfovy, const int32_t* radii,
                 void* geom_ws, void* binning_ws, int64_t instance_capacity, void* image_ws,
                 const float* dL_dpix, float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D,
                 float* dL_dcov3D, float* dL_dsh, float* dL_dscales, float* dL_drots, void* stream,
                 int debug);

/* ---- Raw-parameter entries (no counterpart in `_C`).  The reference activates its
 * parameters in PyTorch before every rasterizer call (scene/gaussian_model.py:152-183:
 * scaling = exp(_scaling), rotation = normalize(_rotation), opacity = sigmoid(_opacity),
 * features = cat(_features_dc, _features_rest)) and autograd walks the same ops backwards: > 1.5 GB of
 * extra traffic per step at 3 M Gaussians.  These entries take the PRE-activation tensors, apply the
 * same formulas inside preprocess / preprocess_backward, and return gradients w.r.t. the raw tensors.
 * features_dc is [P,1,3], features_rest [P,M-1,3] (NULL when M == 1), M = SH coefficients per channel.
 * Everything else (workspaces, gsr_forward_render, error behaviour) is as for the standard entries.
 * gsr_activate runs the activations alone with the same device functions (shs [P,M,3], opacities [P],
 * scales [P,3], rotations [P,4]): feeding its outputs to the standard entries reproduces the raw
 * entries' results bit for bit (tests/test_gpu_raw.py). */
GSR_API int gsr_forward_geometry_raw(int P, int D, int M, int width, int height, const float* means3D,
                         const float* features_dc, const float* features_rest, const float* opacity_raw,
                         const float* scaling_raw, float scale_modifier, const float* rotation_raw,
                         const float* viewmatrix, const float* projmatrix, const float* campos,
                         float tanfovx, float tanfovy, int32_t* radii, void* geom_ws,
                         gsr_forward_info* h_info, void* stream, int debug);

GSR_API int gsr_backward_raw(int P, int D, int M, int width, int height, const float* bg,
                         const float* means3D, const float* features_dc, const float* features_rest,
                         const float* scaling_raw, float scale_modifier, const float* rotation_raw,
                         const float* viewmatrix, const float* projmatrix, const float* campos,
                         float tanfovx, float tanfovy, const int32_t* radii, void* geom_ws,
                         void* binning_ws, int64_t instance_capacity, void* image_ws, const float* dL_dpix,
                         float* dL_dmeans2D, float* dL_dopacity_raw, float* dL_dmeans3D,
                         float* dL_dfeatures_dc, float* dL_dfeatures_rest, float* dL_dscaling_raw,
                         float* dL_drotation_raw, void* stream, int debug);

GSR_API int gsr_activate(int P, int M, const float* features_dc, const float* features_rest,
                         const float* opacity_raw, const float* scaling_raw, const float* rotation_raw,
                         float* shs, float* opacities, float* scales, float* rotations, void* stream);

/* ---- The step after backward (no counterpart in `_C`: the reference runs
 * torch.optim.Adam(l, lr=0.0, eps=1e-15) over six parameter groups, scene/gaussian_model.py:245-268,
 * train.py:140, and keeps per-Gaussian densification statistics with indexed PyTorch ops,
 * train.py:130-131, scene/gaussian_model.py:775-777).  gsr_adam_step applies ONE Adam step (torch's
 * formula: no amsgrad, no weight decay) to every group in one launch; `step` is the 1-based step count
 * used for the bias corrections; all pointers 16-byte aligned device pointers of n floats.
 * gsr_densification_stats: for radii[i] > 0: max_radii2D[i] = max(., radii[i]),
 * xyz_gradient_accum[i] += ||dL_dmeans2D[i,:2]||, denom[i] += 1. */
#define GSR_MAX_ADAM_GROUPS 16
typedef struct gsr_adam_group {
    float* param;
    const float* grad;
    float* exp_avg;
    float* exp_avg_sq;
    int64_t n;
    double lr;   /* as Python holds it: lr / (1 - beta1^step) is formed in binary64, as torch does */
} gsr_adam_group;

GSR_API int gsr_adam_step(int n_groups, const gsr_adam_group* groups, double beta1, double beta2, double eps,
                          int step, void* stream);

GSR_API int gsr_densification_stats(int P, const int32_t* radii, const float* dL_dmeans2D,
                                    float* max_radii2D, float* xyz_gradient_accum, float* denom,
                                    void* stream);

/* ---- The step between forward and backward (no counterpart in `_C`: the reference forms
 * the photometric loss with PyTorch ops, utils/loss_utils.py:17-63 — l1_loss, and ssim as five zero-padded
 * grouped 11x11 Gaussian-window conv2d calls plus ~15 elementwise ops — combined at train.py:99-101 as
 * (1 - lambda) * l1 + lambda * (1 - ssim), then autograd).  image / gt: [C,H,W] contiguous float device
 * pointers.  gsr_loss_forward writes out3 = { a0 + a_l1 * l1 + a_ssim * ssim, l1, ssim } (device floats;
 * the training loss is a0 = lambda, a_l1 = 1 - lambda, a_ssim = -lambda) and, when save_for_backward != 0,
 * the three per-pixel partial-derivative maps gsr_loss_backward needs into the workspace
 * (gsr_loss_workspace_bytes; 256-byte aligned).  gsr_loss_backward writes
 * dL_dimage = upstream * d(a_l1 * l1 + a_ssim * ssim)/d image for every pixel; `upstream` is a device
 * scalar (autograd's grad_output) or NULL for 1.  Two launches in all; no host synchronisation.
 * With a_ssim == 0 (l1_loss alone) both calls take a streaming L1-only path: no maps are written or read
 * (save_for_backward may be 0) and out3[2] is 0. */
GSR_API int gsr_loss_workspace_bytes(int C, int H, int W, int save_for_backward, size_t* bytes);
GSR_API int gsr_loss_forward(int C, int H, int W, const float* image, const float* gt, double a0, double a_l1,
                             double a_ssim, void* workspace, int save_for_backward, float* out3, void* stream);
GSR_API int gsr_loss_backward(int C, int H, int W, const float* image, const float* gt, double a_l1,
                              double a_ssim, const void* workspace, const float* upstream, float* dL_dimage,
                              void* stream);

/* ---- The same loss and its gradient for callers that cannot allocate (no counterpart in the reference:
 * there the loss is PyTorch ops and torch's allocator is always at hand, utils/loss_utils.py).
 * Who this is for: a host that binds this C ABI without an allocator of its own and must not call
 * cudaMalloc where it calls us — a viewer or evaluation plugin whose render callback is forbidden to
 * allocate, a service whose device memory was handed out in full at start-up, or a loop that shares the GPU
 * with other streams (cudaMalloc and cudaFree wait for the whole device, so one call would stall them all).
 * Such a caller cannot provide the workspace gsr_loss_forward / gsr_loss_backward need.
 * This one call does both launches — out3 = { a0 + a_l1 * l1 + a_ssim * ssim, l1, ssim } and
 * dL_dimage = upstream * d(a_l1 * l1 + a_ssim * ssim)/d image (upstream: device scalar or NULL for 1) —
 * with the scratch (256-byte header + 12 bytes per image element) in an array built into the library.
 * The caller still owns and passes image, gt, out3 and dL_dimage.  The call makes no cudaMalloc / cudaFree
 * and uses no allocating library routine.
 *   Limit: C * H * W <= GSR_LOSS_STATIC_MAX_ELEMS (a 3 x 1080 x 1920 image); larger inputs are refused with
 *     GSR_ERR_SCRATCH_LIMIT before anything is launched.  Use the workspace form for those.
 *   Cost: the array (about 71 MiB) occupies device memory on every GPU that loads the library's loss module,
 *     from that load until the process exits, whether or not this entry point is ever called.
 *   First call on each device: looks up the array's device address (cudaGetSymbolAddress, with the calling
 *     thread's current device — make the stream's device current, as for every entry point here).  That
 *     first call should be made before any stream capture is begun.
 *   Sharing: every call on a device uses the same array.  It still holds the previous call's contents on
 *     entry; the call clears the header itself and overwrites every map element it later reads, so nothing
 *     has to be cleared by the caller.  Two calls that can run on the GPU at the same time (different streams
 *     without an ordering between them, or two host threads) race on it and give wrong results: order them. */
#define GSR_LOSS_STATIC_MAX_ELEMS (3 * 1080 * 1920)
GSR_API int gsr_loss_value_and_grad_static(int C, int H, int W, const float* image, const float* gt, double a0,
                                           double a_l1, double a_ssim, const float* upstream, float* out3,
                                           float* dL_dimage, void* stream);

/* ---- The on-disk format either side of the render path (no counterpart in `_C`: the
 * reference reads and writes Gaussian .ply files with per-property Python loops over `plyfile`,
 * scene/gaussian_model.py:279-358).  `rows` is the file body on the device: row-major [P, n_props] binary32.
 * gsr_ply_columns gives, for every element of a row of each parameter array, its column in the file row
 * (the caller resolves property names; f_rest is stored transposed in the file: column of
 * features_rest[p, k, c] = column of "f_rest_{c*(M-1)+k}").  n_rest = 3*(M-1) in {0, 9, 24, 45}.
 * gsr_ply_unpack: rows -> xyz [P,3], f_dc [P,1,3], f_rest [P,M-1,3], opacity [P,1], scale [P,3], rot [P,4].
 * gsr_ply_pack: the reverse; columns no array maps to (nx ny nz) are written as zeros, as save_ply does.
 * Every column may be named at most once.  `rows` must be 16-byte aligned. */
#define GSR_PLY_MAX_PROPS 96
typedef struct gsr_ply_columns {
    int32_t n_props;
    int32_t n_rest;
    int16_t xyz[3];
    int16_t f_dc[3];
    int16_t f_rest[45];
    int16_t opacity;
    int16_t scale[3];
    int16_t rot[4];
} gsr_ply_columns;

GSR_API int gsr_ply_unpack(int64_t P, const gsr_ply_columns* cols, const float* rows, float* xyz, float* f_dc,
                           float* f_rest, float* opacity, float* scale, float* rot, void* stream);
GSR_API int gsr_ply_pack(int64_t P, const gsr_ply_columns* cols, const float* xyz, const float* f_dc,
                         const float* f_rest, const float* opacity, const float* scale, const float* rot,
                         float* rows, void* stream);

/* ---- Gaussian-sharded multi-view entries (no reference equivalent: the reference is single-GPU,
 * utils/general_utils.py:133).  With N ranks, rank r owns the parameter rows of slice r and renders
 * view r.  Forward: every rank runs gsr_preprocess_slots for each of the N cameras over ITS slice,
 * the per-Gaussian geometry rows (depths, means2D, conic_opacity, rgb, tiles_touched, radii) are
 * exchanged all-to-all (52 B per Gaussian and view instead of 236 B of parameter gradients), then
 * gsr_depth_order + gsr_forward_render run on the assembled full-P geometry workspace.  Backward:
 * gsr_blend_backward fills grad_accum[P,12] of the rendered view, its slices are exchanged back
 * all-to-all (48 B per Gaussian), and gsr_preprocess_backward_slots turns each view's accumulators
 * of the slice into parameter gradients, summed over the views.  The arithmetic per (view, Gaussian)
 * is exactly that of gsr_forward_geometry / gsr_backward.
 *
 * `geom_ws` is a geometry workspace laid out for `P_ws` rows (gsr_workspace_layout(P_ws, ...));
 * the call touches rows [slot_offset, slot_offset + Ps) of its per-Gaussian arrays.  Parameter
 * pointers address the Ps rows of the slice; slot_offset must be a multiple of 4. */
GSR_API int gsr_preprocess_slots(int P_ws, int slot_offset, int Ps, int D, int M, int width, int height,
                         const float* means3D, const float* shs, const float* colors_precomp,
                         const float* opacities, const float* scales, float scale_modifier,
                         const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                         const float* projmatrix, const float* campos, float tanfovx, float tanfovy,
                         int32_t* radii_slots, void* geom_ws, void* stream, int debug);

/* Counts (num_visible, num_rendered) + depth-order sort over a geometry workspace whose per-Gaussian arrays are already
 * filled (second half of gsr_forward_geometry). */
GSR_API int gsr_depth_order(int P, int width, int height, void* geom_ws, gsr_forward_info* h_info,
                            void* stream, int debug);

/* First half of gsr_backward: dL/dpixel -> grad_accum[P,12] of the geometry workspace. */
GSR_API int gsr_blend_backward(int P, int width, int height, const float* bg, void* geom_ws,
                               void* binning_ws, int64_t instance_capacity, void* image_ws,
                               const float* dL_dpix, void* stream, int debug);

/* Second half of gsr_backward over rows [slot_offset, slot_offset + Ps) of a P_ws-row geometry
 * workspace (conic_opacity, rgb, grad_accum) and `radii_slots` (Ps rows).  dL_dmeans2D (Ps rows,
 * per view) is always overwritten; with accumulate != 0 the parameter gradients are ADDED to the
 * output arrays (rows the view does not see are left untouched), otherwise overwritten in full. */
GSR_API int gsr_preprocess_backward_slots(int P_ws, int slot_offset, int Ps, int D, int M, int width,
                         int height, const float* means3D, const float* shs, const float* colors_precomp,
                         const float* scales, float scale_modifier, const float* rotations,
                         const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                         const float* campos, float tanfovx, float tanfovy, const int32_t* radii_slots,
                         void* geom_ws, float* dL_dmeans2D, float* dL_dcolors, float* dL_dopacity,
                         float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscales,
                         float* dL_drots, int accumulate, void* stream, int debug);

/* Multi-view forms of the two slot entries: ONE call (and, for the backward, ONE kernel that reads the
 * slice's parameters and SH rows once and loops over the views, accumulating in registers / shared
 * memory) for all n_views <= GSR_MAX_VIEWS views of the step.  `views` is a HOST array; its pointers are
 * device pointers.  Every view uses the same P_ws / slot_offset / Ps. */
typedef struct gsr_view_slot {
    const float* viewmatrix;   /* [16] */
    const float* projmatrix;   /* [16] */
    const float* campos;       /* [3]  */
    float tanfovx, tanfovy;
    void* geom_ws;             /* P_ws-row geometry workspace of this view (may be a peer GPU's) */
    int32_t* radii_slots;      /* radii rows [slot_offset, slot_offset + Ps) of this view */
    float* dL_dmeans2D;        /* backward only: [Ps,3] of this view, always overwritten */
    /* Optional LOCAL copies (both or neither; [Ps] rows, private to the slice owner) of what the forward produced
     * for this view: the packed word  radius | clamped_bits << 28  and the conic + opacity.  gsr_preprocess_views
     * writes them; gsr_preprocess_backward_views then reads THEM instead of radii_slots / conic_opacity / rgb of
     * geom_ws, so that only the 48-byte blend accumulators are fetched from a peer's workspace.  The radii of
     * the densification statistics (train.py:130) are  word & GSR_LOCAL_RADIUS_MASK. */
    int32_t* radii_local;
    float* conic_local;        /* [Ps,4] */
} gsr_view_slot;
#define GSR_LOCAL_RADIUS_MASK 0x0FFFFFFF

GSR_API int gsr_preprocess_views(int P_ws, int slot_offset, int Ps, int D, int M, int width, int height,
                         const float* means3D, const float* shs, const float* colors_precomp,
                         const float* opacities, const float* scales, float scale_modifier,
                         const float* rotations, const float* cov3D_precomp, int n_views,
                         const gsr_view_slot* views, void* stream, int debug);

/* `accumulate` of gsr_preprocess_backward_views is a set of flags: */
#define GSR_BWD_ACCUMULATE 1         /* add the parameter gradients to the outputs (a second chunk of views)  */
#define GSR_BWD_PEER_ACCUMULATORS 2  /* the views' workspaces are peer-GPU memory: fetch the 48-byte accumulator
                                        rows by bulk asynchronous copies, four views ahead, instead of one
                                        dependent NVLink load per view                                        */
GSR_API int gsr_preprocess_backward_views(int P_ws, int slot_offset, int Ps, int D, int M, int width,
                         int height, const float* means3D, const float* shs, const float* colors_precomp,
                         const float* scales, float scale_modifier, const float* rotations,
                         const float* cov3D_precomp, int n_views, const gsr_view_slot* views,
                         float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D,
                         float* dL_dsh, float* dL_dscales, float* dL_drots, int accumulate, void* stream,
                         int debug);

/* The reference's sorted / unsorted 64-bit key list (binningState.point_list_keys[_unsorted]: (tile << 32) |
 * depth bits) of the last gsr_forward_render on these workspaces, written to d_keys[0 .. num_rendered) (device
 * memory, instance_capacity entries).  sorted = 0: emission order (Gaussians in depth order, each one's tiles
 * row-major — a per-Gaussian-block permutation of the reference's index-order list); sorted = 1: the sorted list.
 * For parity tests and debugging: the render path stores only tile ids and indices. */
GSR_API int gsr_materialize_keys(int P, int width, int height, void* geom_ws, void* binning_ws,
                         int64_t instance_capacity, int sorted, uint64_t* d_keys, void* stream);

/* Replaces `_C.mark_visible`: present[i] = 1 iff Gaussian i passes the near-plane test. */
GSR_API int gsr_mark_visible(int P, const float* means3D, const float* viewmatrix,
                     const float* projmatrix, uint8_t* present, void* stream);

/* Number of kernel launches the library has issued in this process (all threads); bench.py
 * reports the difference across its timed region as `gpu_launches`. */
GSR_API int64_t gsr_launch_count(void);

/* Instrumented re-run of the blend forward over the workspaces of a finished forward call (no
 * reference equivalent; SURVEY 8d "debug counters in a separate instrumented build, not the timed
 * one").  Writes FOUR u64 to DEVICE memory: d_counters[0] = (pixel, splat) pairs the reference
 * algorithm evaluates (each pixel walks its tile list until it saturates), d_counters[1] = pairs
 * that are blended, d_counters[2] = pairs THIS library's kernel evaluates after its two culling
 * levels, d_counters[3] = (warp, splat) loop iterations it issues.  bench.py derives the blend
 * kernels' algorithmic FLOPs from [0] and [1]. */
GSR_API int gsr_blend_pair_counts(int P, int width, int height, void* geom_ws, void* binning_ws,
                                  int64_t instance_capacity, void* image_ws, uint64_t* d_counters,
                                  void* stream);

/* Optional per-stage device timing for benchmarks (no reference equivalent; the reference's only
 * timer is the CUDA-event pair at train.py:47-48,70,109).  While enabled, the launches of each
 * stage are bracketed by a CUDA event pair ON THE CALLER'S STREAM; gsr_stage_timing_read waits for
 * the recorded events, returns per-stage summed milliseconds and bracket counts since the last
 * read (arrays of at least gsr_stage_count() entries) and recycles the events.  Process-wide. */
GSR_API int gsr_stage_count(void);
GSR_API const char* gsr_stage_name(int stage);
GSR_API int gsr_stage_timing_enable(int enable);
GSR_API int gsr_stage_timing_read(double* ms_sum, int64_t* calls, int n);

#ifdef __cplusplus
}
#endif
#endif /* GSRAST_B200_H_ */
