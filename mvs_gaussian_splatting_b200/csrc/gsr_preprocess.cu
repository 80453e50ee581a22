// gsr_preprocess.cu — forward per-Gaussian stage (the reference's `preprocessCUDA`, SURVEY §2.1 K1,
// §8a rows a8-a12) and markVisible (K11), written for sm_100a.
//
// One thread per Gaussian, 128-thread CTAs.  HBM-bound: 44 B/Gaussian always + the active SH
// prefix (up to 192 B) only for Gaussians that survive the cull.  means/scales are fetched as
// coalesced float4 through shared memory; SH rows are fetched warp-cooperatively as float4 (lane j
// of a warp fetches 16-byte piece j of the warp's surviving rows) into an odd-stride shared tile,
// so every 32-byte sector requested from HBM is fully used and culled rows cost nothing.
//
// All arithmetic that decides radii / tile rects / depth bits follows oracle/gs_oracle.c
// operation for operation with non-contractable _rn intrinsics (see gsr_internal.cuh).
#include <algorithm>
#include <cstdlib>

#include "gsr_internal.cuh"

namespace gsr {

namespace {

constexpr int kPreThreads = 128;
constexpr int kPreWarps = kPreThreads / 32;
constexpr int kShStride = 49;  // 48 floats (16 coeffs x RGB) + 1 pad -> conflict-free column reads

__constant__ float c_SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                 -1.0925484305920792f, 0.5462742152960396f};
__constant__ float c_SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f,  -0.4570457994644658f,
                                 0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                                 -0.5900435899266435f};

constexpr int kMaxViews = GSR_MAX_VIEWS;

// What differs from view to view: the camera and where the rows of that view go (possibly a peer
// GPU's workspace, written with plain remote stores).
struct PreView {
    const float* view;
    const float* proj;
    const float* campos;
    float tanfovx, tanfovy, focal_x, focal_y;
    int32_t* radii;
    GeomView gv;
    int32_t* radii_local;   // optional local copies for the slice owner's backward (gsr_view_slot)
    float4* conic_local;
};

struct PreParams {
    int P, D, M;
    int W, H, gx, gy;
    int n_views;
    const float* means3D;
    const float* shs;
    const float* colors_precomp;
    const float* opacities;
    const float* scales;
    float scale_modifier;
    const float* rotations;
    const float* cov3D_precomp;
    int raw;               // RawMode: pre-activation scales / rotations / opacities
    const float* sh_rest;  // RawMode: shs = dc [P,1,3], sh_rest = [P,M-1,3]
    uint32_t* depth_hist;  // bulk-copy kernel: digit histograms of the depth sort [4][256] (zeroed), or null
    PreView v[kMaxViews];
};

// Stage a [n,3] f32 array slice for this CTA into shared memory with float4 loads.
__device__ __forceinline__ void stage_xyz(const float* __restrict__ src, int P, int base, float* s) {
    const int nfloat_total = 3 * P;
    const int f0 = 3 * base;                                 // multiple of 384 floats -> 16B aligned
    const int nfl = min(3 * kPreThreads, nfloat_total - f0); // floats this CTA needs
    const int nvec = nfl >> 2;
    const float4* src4 = reinterpret_cast<const float4*>(src + f0);
    for (int j = threadIdx.x; j < nvec; j += kPreThreads) {
        float4 v = __ldg(src4 + j);
        s[4 * j + 0] = v.x; s[4 * j + 1] = v.y; s[4 * j + 2] = v.z; s[4 * j + 3] = v.w;
    }
    for (int j = 4 * nvec + threadIdx.x; j < nfl; j += kPreThreads) s[j] = __ldg(src + f0 + j);
}

// Sigma = R diag(s)^2 R^T (6 upper-tri floats), quaternion used as given.
__device__ __forceinline__ void cov3d_from_scale_rot(float sx, float sy, float sz, float mod, float4 q,
                                                     float* c6) {
    const float s0 = __fmul_rn(mod, sx), s1 = __fmul_rn(mod, sy), s2 = __fmul_rn(mod, sz);
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    const float R00 = __fadd_rn(1.0f, -__fmul_rn(2.0f, dot2_rn(y, y, z, z)));
    const float R01 = __fmul_rn(2.0f, __fmaf_rn(x, y, -__fmul_rn(r, z)));
    const float R02 = __fmul_rn(2.0f, dot2_rn(x, z, r, y));
    const float R10 = __fmul_rn(2.0f, dot2_rn(x, y, r, z));
    const float R11 = __fadd_rn(1.0f, -__fmul_rn(2.0f, dot2_rn(x, x, z, z)));
    const float R12 = __fmul_rn(2.0f, __fmaf_rn(y, z, -__fmul_rn(r, x)));
    const float R20 = __fmul_rn(2.0f, __fmaf_rn(x, z, -__fmul_rn(r, y)));
    const float R21 = __fmul_rn(2.0f, dot2_rn(y, z, r, x));
    const float R22 = __fadd_rn(1.0f, -__fmul_rn(2.0f, dot2_rn(x, x, y, y)));
    const float M00 = __fmul_rn(s0, R00), M01 = __fmul_rn(s0, R10), M02 = __fmul_rn(s0, R20);
    const float M10 = __fmul_rn(s1, R01), M11 = __fmul_rn(s1, R11), M12 = __fmul_rn(s1, R21);
    const float M20 = __fmul_rn(s2, R02), M21 = __fmul_rn(s2, R12), M22 = __fmul_rn(s2, R22);
    c6[0] = dot3_rn(M00, M00, M10, M10, M20, M20);
    c6[1] = dot3_rn(M00, M01, M10, M11, M20, M21);
    c6[2] = dot3_rn(M00, M02, M10, M12, M20, M22);
    c6[3] = dot3_rn(M01, M01, M11, M11, M21, M21);
    c6[4] = dot3_rn(M01, M02, M11, M12, M21, M22);
    c6[5] = dot3_rn(M02, M02, M12, M12, M22, M22);
}

// EWA projection of Sigma to the screen: (a,b,c) incl. the +0.3 dilation.  t = view-space mean.
__device__ __forceinline__ void cov2d_project(float tx, float ty, float tz, float fx, float fy,
                                              float tanx, float tany, const float* c6,
                                              const float* __restrict__ v, float& a, float& b, float& c) {
    const float limx = __fmul_rn(1.3f, tanx), limy = __fmul_rn(1.3f, tany);
    const float txtz = __fdiv_rn(tx, tz), tytz = __fdiv_rn(ty, tz);
    tx = __fmul_rn(fminf(limx, fmaxf(-limx, txtz)), tz);
    ty = __fmul_rn(fminf(limy, fmaxf(-limy, tytz)), tz);
    const float j00 = __fdiv_rn(fx, tz);
    const float j11 = __fdiv_rn(fy, tz);
    const float tz2 = __fmul_rn(tz, tz);
    const float j02 = __fdiv_rn(-__fmul_rn(fx, tx), tz2);
    const float j12 = __fdiv_rn(-__fmul_rn(fy, ty), tz2);
    float M0[3], M1[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        M0[k] = __fmaf_rn(v[4 * k + 2], j02, __fmul_rn(v[4 * k + 0], j00));
        M1[k] = __fmaf_rn(v[4 * k + 2], j12, __fmul_rn(v[4 * k + 1], j11));
    }
    const float S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
    float A0[3], A1[3];
#pragma unroll
    for (int cc = 0; cc < 3; ++cc) {
        A0[cc] = dot3_rn(M0[0], S[0][cc], M0[1], S[1][cc], M0[2], S[2][cc]);
        A1[cc] = dot3_rn(M1[0], S[0][cc], M1[1], S[1][cc], M1[2], S[2][cc]);
    }
    a = __fadd_rn(dot3_rn(A0[0], M0[0], A0[1], M0[1], A0[2], M0[2]), 0.3f);
    b = dot3_rn(A1[0], M0[0], A1[1], M0[1], A1[2], M0[2]);
    c = __fadd_rn(dot3_rn(A1[0], M1[0], A1[1], M1[1], A1[2], M1[2]), 0.3f);
}

// Degree 0..3 real SH -> RGB (+0.5, clamp >= 0, clamped bits).  sh = this Gaussian's row in the
// shared tile: sh[3*k + c].  Same evaluation order as the oracle's sh_to_rgb.
__device__ __forceinline__ void sh_to_rgb(int deg, float dx, float dy, float dz, const float* sh,
                                          float* rgb, uint32_t& clamped) {
    const float len = __fsqrt_rn(dot3_rn(dx, dx, dy, dy, dz, dz));
    const float x = __fdiv_rn(dx, len), y = __fdiv_rn(dy, len), z = __fdiv_rn(dz, len);
    const float SH_C0 = 0.28209479177387814f, SH_C1 = 0.4886025119029199f;
    clamped = 0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
#define SH(k) sh[3 * (k) + c]
        float res = __fmul_rn(SH_C0, SH(0));
        if (deg > 0) {
            res = __fmaf_rn(-__fmul_rn(SH_C1, y), SH(1), res);
            res = __fmaf_rn(__fmul_rn(SH_C1, z), SH(2), res);
            res = __fmaf_rn(-__fmul_rn(SH_C1, x), SH(3), res);
            if (deg > 1) {
                const float xx = __fmul_rn(x, x), yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
                const float xy = __fmul_rn(x, y), yz = __fmul_rn(y, z), xz = __fmul_rn(x, z);
                res = __fmaf_rn(__fmul_rn(c_SH_C2[0], xy), SH(4), res);
                res = __fmaf_rn(__fmul_rn(c_SH_C2[1], yz), SH(5), res);
                res = __fmaf_rn(__fmul_rn(c_SH_C2[2], __fadd_rn(__fmaf_rn(2.0f, zz, -xx), -yy)), SH(6), res);
                res = __fmaf_rn(__fmul_rn(c_SH_C2[3], xz), SH(7), res);
                res = __fmaf_rn(__fmul_rn(c_SH_C2[4], __fadd_rn(xx, -yy)), SH(8), res);
                if (deg > 2) {
                    res = __fmaf_rn(__fmul_rn(__fmul_rn(c_SH_C3[0], y), __fmaf_rn(3.0f, xx, -yy)), SH(9), res);
                    res = __fmaf_rn(__fmul_rn(__fmul_rn(c_SH_C3[1], xy), z), SH(10), res);
                    res = __fmaf_rn(__fmul_rn(__fmul_rn(c_SH_C3[2], y), __fadd_rn(__fmaf_rn(4.0f, zz, -xx), -yy)), SH(11), res);
                    res = __fmaf_rn(__fmul_rn(__fmul_rn(c_SH_C3[3], z),
                                              __fmaf_rn(-3.0f, yy, __fmaf_rn(2.0f, zz, -__fmul_rn(3.0f, xx)))),
                                    SH(12), res);
                    res = __fmaf_rn(__fmul_rn(__fmul_rn(c_SH_C3[4], x), __fadd_rn(__fmaf_rn(4.0f, zz, -xx), -yy)), SH(13), res);
                    res = __fmaf_rn(__fmul_rn(__fmul_rn(c_SH_C3[5], z), __fadd_rn(xx, -yy)), SH(14), res);
                    res = __fmaf_rn(__fmul_rn(__fmul_rn(c_SH_C3[6], x), __fmaf_rn(-3.0f, yy, xx)), SH(15), res);
                }
            }
        }
#undef SH
        res = __fadd_rn(res, 0.5f);
        if (res < 0.0f) clamped |= (1u << c);
        rgb[c] = fmaxf(res, 0.0f);
    }
}

// kMulti = true: means / scales / rotation / opacity and the SH rows of the CTA are read ONCE and the
// views are looped over (SH rows are fetched when the first view that sees the Gaussian needs them).
template <bool kMulti, bool kRaw>
__global__ void __launch_bounds__(kPreThreads, kMulti ? 5 : 8) preprocess_kernel(const PreParams p) {
    __shared__ float s_mean[3 * kPreThreads];
    __shared__ float s_scale[3 * kPreThreads];
    __shared__ float s_sh[kPreWarps][32 * kShStride];
    __shared__ float s_cam[35];  // view[16] proj[16] campos[3]

    const int base = blockIdx.x * kPreThreads;
    const int idx = base + threadIdx.x;
    const unsigned lane = lane_id();
    const int warp = threadIdx.x >> 5;

    stage_xyz(p.means3D, p.P, base, s_mean);
    if (p.scales) stage_xyz(p.scales, p.P, base, s_scale);
    const float* view = s_cam;
    const float* proj = s_cam + 16;
    unsigned fetched = 0;      // SH rows of this warp already in the shared tile
    bool have_c6 = false;      // 3D covariance (view-independent) already formed
    float c6[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};

  for (int vi = 0; vi < p.n_views; ++vi) {
    const PreView& V = p.v[vi];
    if (kMulti && vi > 0) __syncthreads();  // the previous view's camera is no longer read
    if (threadIdx.x < 16) s_cam[threadIdx.x] = __ldg(V.view + threadIdx.x);
    else if (threadIdx.x < 32) s_cam[threadIdx.x] = __ldg(V.proj + threadIdx.x - 16);
    else if (threadIdx.x < 35) s_cam[threadIdx.x] = __ldg(V.campos + threadIdx.x - 32);
    __syncthreads();

    bool alive = idx < p.P;
    float mx = 0, my = 0, mz = 0, depth = 0, px = 0, py = 0, ca = 0, cb = 0, cc = 0, det_inv = 0;
    int rad = 0, x0 = 0, y0 = 0, x1 = 0, y1 = 0;

    if (alive) {
        mx = s_mean[3 * threadIdx.x]; my = s_mean[3 * threadIdx.x + 1]; mz = s_mean[3 * threadIdx.x + 2];
        const float tx = xform_row_rn(view, 0, mx, my, mz);
        const float ty = xform_row_rn(view, 1, mx, my, mz);
        depth = xform_row_rn(view, 2, mx, my, mz);
        alive = depth > kNearPlane;  // near-plane cull only; NaN fails too
        if (alive) {
            const float hx = xform_row_rn(proj, 0, mx, my, mz);
            const float hy = xform_row_rn(proj, 1, mx, my, mz);
            const float hw = xform_row_rn(proj, 3, mx, my, mz);
            const float pw = __fdiv_rn(1.0f, __fadd_rn(hw, 0.0000001f));
            const float ndcx = __fmul_rn(hx, pw), ndcy = __fmul_rn(hy, pw);
            if (!have_c6) {
                if (p.cov3D_precomp) {
                    const float2* cp = reinterpret_cast<const float2*>(p.cov3D_precomp) + 3 * (size_t)idx;
                    float2 u0 = __ldg(cp), u1 = __ldg(cp + 1), u2 = __ldg(cp + 2);
                    c6[0] = u0.x; c6[1] = u0.y; c6[2] = u1.x; c6[3] = u1.y; c6[4] = u2.x; c6[5] = u2.y;
                } else {
                    float4 q = __ldg(reinterpret_cast<const float4*>(p.rotations) + idx);
                    float sx = s_scale[3 * threadIdx.x], sy = s_scale[3 * threadIdx.x + 1], sz = s_scale[3 * threadIdx.x + 2];
                    if (kRaw) {
                        q = act_rotation(q);
                        sx = act_scale(sx); sy = act_scale(sy); sz = act_scale(sz);
                    }
                    cov3d_from_scale_rot(sx, sy, sz, p.scale_modifier, q, c6);
                }
                have_c6 = kMulti;  // single view: nothing to keep
            }
            cov2d_project(tx, ty, depth, V.focal_x, V.focal_y, V.tanfovx, V.tanfovy, c6, view, ca, cb, cc);
            const float det = __fmaf_rn(ca, cc, -__fmul_rn(cb, cb));
            alive = det != 0.0f;
            if (alive) {
                det_inv = __fdiv_rn(1.0f, det);
                const float mid = __fmul_rn(0.5f, __fadd_rn(ca, cc));
                const float sq = __fsqrt_rn(fmaxf(0.1f, __fmaf_rn(mid, mid, -det)));
                const float lam1 = __fadd_rn(mid, sq), lam2 = __fadd_rn(mid, -sq);
                rad = __float2int_rz(ceilf(__fmul_rn(3.0f, __fsqrt_rn(fmaxf(lam1, lam2)))));
                px = __fmul_rn(__fmaf_rn(__fadd_rn(ndcx, 1.0f), (float)p.W, -1.0f), 0.5f);
                py = __fmul_rn(__fmaf_rn(__fadd_rn(ndcy, 1.0f), (float)p.H, -1.0f), 0.5f);
                tile_rect(px, py, rad, p.gx, p.gy, x0, y0, x1, y1);
                alive = (x1 - x0) * (y1 - y0) != 0;
            }
        }
    }

    // ---- colour: SH rows of surviving Gaussians, fetched warp-cooperatively ----
    const unsigned vis = __ballot_sync(0xffffffffu, alive);
    float rgb[3] = {0.f, 0.f, 0.f};
    uint32_t clamped = 0;
    if (p.colors_precomp) {
        if (alive) {
            rgb[0] = __ldg(p.colors_precomp + 3 * (size_t)idx);
            rgb[1] = __ldg(p.colors_precomp + 3 * (size_t)idx + 1);
            rgb[2] = __ldg(p.colors_precomp + 3 * (size_t)idx + 2);
        }
    } else if (vis) {
        float* tile = s_sh[warp];
        const int row = 3 * p.M;                      // floats per Gaussian in global memory
        const int need = 3 * (p.D + 1) * (p.D + 1);   // floats of the active degree
        const int warp_base = base + warp * 32;
        const unsigned want = vis & ~fetched;   // rows this view needs that are not in the tile yet
        fetched |= kMulti ? vis : 0u;
        const int nvis = __popc(want);
        if (kRaw) {  // split layout: dc [P,1,3] | rest [P,M-1,3]
            const int nrows = min(32, p.P - warp_base);
            fetch_rows_f4(p.shs + 3 * (size_t)warp_base, nrows, 3, 3, 0, want, tile, kShStride, lane);
            if (need > 3)
                fetch_rows_f4(p.sh_rest + (size_t)warp_base * (row - 3), nrows, row - 3, need - 3, 3, want, tile,
                              kShStride, lane);
        } else if ((row & 3) == 0) {
            const int vpr = (need + 3) >> 2;  // float4 pieces per row (row start is 16B aligned)
            const int total = nvis * vpr;
            const int my_src = __fns(want, 0, lane + 1);  // lane of the (lane+1)-th row to fetch
            // four independent 16-byte loads in flight per lane before the first shared store
            for (int jb = 0; jb < total; jb += 128) {  // warp-uniform trip count (shuffles inside)
                const int j0 = jb + (int)lane;
                float4 val[4];
                int dst[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int j = j0 + 32 * u;
                    const int v = min(j, total - 1) / vpr, part = min(j, total - 1) - v * vpr;
                    const int src_lane = __shfl_sync(0xffffffffu, my_src, v);
                    dst[u] = src_lane * kShStride + 4 * part;
                    if (j < total)
                        val[u] = __ldg(reinterpret_cast<const float4*>(p.shs + (size_t)(warp_base + src_lane) * row) + part);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (j0 + 32 * u < total) {
                        float* d = tile + dst[u];
                        d[0] = val[u].x; d[1] = val[u].y; d[2] = val[u].z; d[3] = val[u].w;
                    }
                }
            }
        } else {
            const int total = nvis * need;
            for (int j = lane; j < total; j += 32) {
                const int v = j / need, part = j - v * need;
                const int src_lane = __fns(want, 0, v + 1);
                tile[src_lane * kShStride + part] = __ldg(p.shs + (size_t)(warp_base + src_lane) * row + part);
            }
        }
        __syncwarp();
        if (alive) {
            sh_to_rgb(p.D, __fadd_rn(mx, -view[32]), __fadd_rn(my, -view[33]), __fadd_rn(mz, -view[34]),
                      tile + lane * kShStride, rgb, clamped);
        }
    }

    if (idx < p.P) {
        V.radii[idx] = alive ? rad : 0;
        V.gv.tiles_touched[idx] = alive ? (uint32_t)((x1 - x0) * (y1 - y0)) : 0u;
        // the depth array doubles as the key of the per-Gaussian depth sort: culled rows sink to the end
        V.gv.depths[idx] = alive ? depth : __uint_as_float(0xFFFFFFFFu);
        float4 conic = make_float4(0.f, 0.f, 0.f, 0.f);
        if (alive) {
            V.gv.means2D[idx] = make_float2(px, py);
            conic = make_float4(__fmul_rn(cc, det_inv), __fmul_rn(-cb, det_inv), __fmul_rn(ca, det_inv),
                                kRaw ? act_opacity(__ldg(p.opacities + idx)) : __ldg(p.opacities + idx));
            V.gv.conic_opacity[idx] = conic;
            V.gv.rgb[idx] = make_float4(rgb[0], rgb[1], rgb[2], __uint_as_float(clamped));
        }
        if (V.radii_local) {   // what the slice owner's backward needs again, kept on this GPU
            V.radii_local[idx] = alive ? (int32_t)((uint32_t)rad | (clamped << 28)) : 0;
            if (alive) V.conic_local[idx] = conic;
        }
    }
  }  // views
}

// ------------------------------------------------------------------------------------------------------
// Persistent form of the single-view kernel with the next tile's bytes in flight (cp.async.bulk + mbarrier):
// the per-Gaussian kernels are latency-bound, not bandwidth-bound (ncu r02a: 47 % of HBM, 56 % issue, three
// dependent round trips per CTA: means -> rotation -> SH rows), so each CTA walks over tiles of 128 Gaussians and
//   * the operand block of tile i+1 (xyz, scale, rotation, opacity: four contiguous runs, 5.6 KB) is requested as
//     soon as tile i's geometry phase has read its own, and lands while tile i-1's colour phase runs;
//   * the SH rows of tile i's SURVIVORS — known at the end of its geometry phase — are requested one 16-byte-
//     multiple run per row (culled rows cost nothing) onto a 208-byte stride (13 sixteen-byte pieces: every
//     quarter-warp of a 128-bit shared load touches eight different bank groups), and are consumed one iteration
//     later, behind tile i+1's geometry phase;
// so a tile's two dependent memory phases overlap the arithmetic of its neighbours inside one CTA, on top of what
// the other resident CTAs hide.  Raw mode fetches its two contiguous blocks (dc | rest) per tile instead.
// Arithmetic and its order are exactly those of preprocess_kernel (bit-identical outputs, tests/test_gpu_*).
constexpr int kShRow4 = 13;                 // float4 pieces per staged SH row (12 used)
struct PreTmaSmem {
    alignas(16) float xyz[3 * kPreThreads];
    alignas(16) float scale[3 * kPreThreads];
    alignas(16) float4 rot[kPreThreads];
    alignas(16) float opac[kPreThreads];
    alignas(16) float4 sh[kPreThreads * kShRow4];   // standard: row t at sh + 13 t; raw: dc block, then rest block
    alignas(8) uint64_t bar_opnd, bar_sh;
    float cam[36];
    uint32_t hist[4][256];   // digit histograms of the visible depth keys (what depth_hist_kernel would count)
    uint32_t sums[2];        // visible Gaussians, tile instances
};

template <bool kRaw>
__global__ void __launch_bounds__(kPreThreads, 6) preprocess_tma_kernel(const PreParams p) {
    __shared__ PreTmaSmem sm;
    const PreView& V = p.v[0];
    const int ntiles = (p.P + kPreThreads - 1) / kPreThreads;
    const int t = threadIdx.x;
    if (t < 16) sm.cam[t] = __ldg(V.view + t);
    else if (t < 32) sm.cam[t] = __ldg(V.proj + t - 16);
    else if (t < 35) sm.cam[t] = __ldg(V.campos + t - 32);
    if (t == 0) { mbar_init(&sm.bar_opnd, 1); mbar_init(&sm.bar_sh, 1); mbar_init_fence(); }
    for (int i = t; i < 4 * 256; i += kPreThreads) (&sm.hist[0][0])[i] = 0;
    if (t < 2) sm.sums[t] = 0;
    __syncthreads();
    uint32_t my_vis = 0, my_inst = 0;
    const float* view = sm.cam;
    const float* proj = sm.cam + 16;
    const int row = 3 * p.M;                       // floats per SH row in global memory (a multiple of 4 here)
    const int need = 3 * (p.D + 1) * (p.D + 1);    // floats of the active degree
    const uint32_t need_bytes = (uint32_t)((need * 4 + 15) & ~15);

    // operands of tile `tile` -> shared (full tiles by the copy engine; the ragged last tile by plain loads)
    auto request_operands = [&](int tile) {
        const int base = tile * kPreThreads;
        const int n = min(kPreThreads, p.P - base);
        if (n == kPreThreads) {
            if (t == 0) {
                mbar_arrive_expect_tx(&sm.bar_opnd, 44u * kPreThreads);
                bulk_g2s(sm.xyz, p.means3D + 3 * (size_t)base, 12u * kPreThreads, &sm.bar_opnd);
                bulk_g2s(sm.scale, p.scales + 3 * (size_t)base, 12u * kPreThreads, &sm.bar_opnd);
                bulk_g2s(sm.rot, p.rotations + 4 * (size_t)base, 16u * kPreThreads, &sm.bar_opnd);
                bulk_g2s(sm.opac, p.opacities + (size_t)base, 4u * kPreThreads, &sm.bar_opnd);
            }
        } else if (t < n) {
            const size_t i = (size_t)base + t;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                sm.xyz[3 * t + k] = __ldg(p.means3D + 3 * i + k);
                sm.scale[3 * t + k] = __ldg(p.scales + 3 * i + k);
            }
            sm.rot[t] = __ldg(reinterpret_cast<const float4*>(p.rotations) + i);
            sm.opac[t] = __ldg(p.opacities + i);
        }
    };

    uint32_t ph_opnd = 0, ph_sh = 0;
    int tile = blockIdx.x;
    if (tile < ntiles) request_operands(tile);
    // state carried from a tile's geometry phase to its colour phase one iteration later
    bool prev_any = false, prev_alive = false;
    int prev_idx = 0;
    float pmx = 0.f, pmy = 0.f, pmz = 0.f;

    for (; tile < ntiles || prev_any; tile += gridDim.x) {
        const bool cur = tile < ntiles;
        const int base = tile * kPreThreads;
        const int idx = base + t;
        bool alive = false;
        float mx = 0.f, my = 0.f, mz = 0.f;
        if (cur) {
            if (p.P - base >= kPreThreads) { mbar_wait(&sm.bar_opnd, ph_opnd); ph_opnd ^= 1u; }
            alive = idx < p.P;
            float depth = 0, px = 0, py = 0, ca = 0, cb = 0, cc = 0, det_inv = 0;
            int rad = 0, x0 = 0, y0 = 0, x1 = 0, y1 = 0;
            if (alive) {
                mx = sm.xyz[3 * t]; my = sm.xyz[3 * t + 1]; mz = sm.xyz[3 * t + 2];
                const float tx = xform_row_rn(view, 0, mx, my, mz);
                const float ty = xform_row_rn(view, 1, mx, my, mz);
                depth = xform_row_rn(view, 2, mx, my, mz);
                alive = depth > kNearPlane;  // near-plane cull only; NaN fails too
                if (alive) {
                    const float hx = xform_row_rn(proj, 0, mx, my, mz);
                    const float hy = xform_row_rn(proj, 1, mx, my, mz);
                    const float hw = xform_row_rn(proj, 3, mx, my, mz);
                    const float pw = __fdiv_rn(1.0f, __fadd_rn(hw, 0.0000001f));
                    const float ndcx = __fmul_rn(hx, pw), ndcy = __fmul_rn(hy, pw);
                    float c6[6];
                    float4 q = sm.rot[t];
                    float sx = sm.scale[3 * t], sy = sm.scale[3 * t + 1], sz = sm.scale[3 * t + 2];
                    if (kRaw) {
                        q = act_rotation(q);
                        sx = act_scale(sx); sy = act_scale(sy); sz = act_scale(sz);
                    }
                    cov3d_from_scale_rot(sx, sy, sz, p.scale_modifier, q, c6);
                    cov2d_project(tx, ty, depth, V.focal_x, V.focal_y, V.tanfovx, V.tanfovy, c6, view, ca, cb, cc);
                    const float det = __fmaf_rn(ca, cc, -__fmul_rn(cb, cb));
                    alive = det != 0.0f;
                    if (alive) {
                        det_inv = __fdiv_rn(1.0f, det);
                        const float mid = __fmul_rn(0.5f, __fadd_rn(ca, cc));
                        const float sq = __fsqrt_rn(fmaxf(0.1f, __fmaf_rn(mid, mid, -det)));
                        const float lam1 = __fadd_rn(mid, sq), lam2 = __fadd_rn(mid, -sq);
                        rad = __float2int_rz(ceilf(__fmul_rn(3.0f, __fsqrt_rn(fmaxf(lam1, lam2)))));
                        px = __fmul_rn(__fmaf_rn(__fadd_rn(ndcx, 1.0f), (float)p.W, -1.0f), 0.5f);
                        py = __fmul_rn(__fmaf_rn(__fadd_rn(ndcy, 1.0f), (float)p.H, -1.0f), 0.5f);
                        tile_rect(px, py, rad, p.gx, p.gy, x0, y0, x1, y1);
                        alive = (x1 - x0) * (y1 - y0) != 0;
                    }
                }
            }
            if (p.depth_hist) {   // the depth sort's histograms and the two counts, while the key is in a register
                const uint32_t key = __float_as_uint(depth);
                const bool full = __all_sync(0xffffffffu, alive);
                if (alive) { ++my_vis; my_inst += (uint32_t)((x1 - x0) * (y1 - y0)); }
#pragma unroll
                for (int pass = 0; pass < 4; ++pass) {
                    const uint32_t d = (key >> (8 * pass)) & 255u;
                    const uint32_t d0 = __shfl_sync(0xffffffffu, d, 0);
                    if (full && __all_sync(0xffffffffu, d == d0)) {   // constant digit (the high depth bytes)
                        if ((t & 31) == 0) atomicAdd(&sm.hist[pass][d0], 32u);
                    } else if (alive) {
                        atomicAdd(&sm.hist[pass][d], 1u);
                    }
                }
            }
            if (idx < p.P) {   // everything but the colour goes out now
                V.radii[idx] = alive ? rad : 0;
                V.gv.tiles_touched[idx] = alive ? (uint32_t)((x1 - x0) * (y1 - y0)) : 0u;
                V.gv.depths[idx] = alive ? depth : __uint_as_float(0xFFFFFFFFu);
                if (alive) {
                    V.gv.means2D[idx] = make_float2(px, py);
                    const float op = kRaw ? act_opacity(sm.opac[t]) : sm.opac[t];
                    V.gv.conic_opacity[idx] = make_float4(__fmul_rn(cc, det_inv), __fmul_rn(-cb, det_inv),
                                                          __fmul_rn(ca, det_inv), op);
                }
            }
        }
        __syncthreads();   // the operand block has been read by everyone
        if (cur && tile + (int)gridDim.x < ntiles) request_operands(tile + gridDim.x);

        // ---- colour phase of the PREVIOUS tile: its SH rows were requested one iteration ago ----
        if (prev_any) {
            mbar_wait(&sm.bar_sh, ph_sh);
            ph_sh ^= 1u;
            if (prev_alive) {
                float shr[48];
                if (kRaw) {
                    const float* dc = reinterpret_cast<const float*>(sm.sh);
                    const float* rest = dc + 3 * kPreThreads;
                    const int rr = row - 3;
#pragma unroll
                    for (int k = 0; k < 3; ++k) shr[k] = dc[3 * t + k];
#pragma unroll
                    for (int k = 3; k < 48; ++k) shr[k] = k < need ? rest[rr * t + k - 3] : 0.f;
                } else {
                    const float4* r4 = sm.sh + kShRow4 * t;
#pragma unroll
                    for (int j = 0; j < 12; ++j) {
                        const float4 v = 4 * j < need ? r4[j] : make_float4(0.f, 0.f, 0.f, 0.f);
                        shr[4 * j] = v.x; shr[4 * j + 1] = v.y; shr[4 * j + 2] = v.z; shr[4 * j + 3] = v.w;
                    }
                }
                float rgb[3];
                uint32_t clamped;
                sh_to_rgb(p.D, __fadd_rn(pmx, -view[32]), __fadd_rn(pmy, -view[33]), __fadd_rn(pmz, -view[34]), shr, rgb,
                          clamped);
                V.gv.rgb[prev_idx] = make_float4(rgb[0], rgb[1], rgb[2], __uint_as_float(clamped));
            }
        }
        const int nalive = __syncthreads_count(cur && alive);   // also: the SH tile has been read by everyone
        if (cur) {
            if (kRaw) {   // two contiguous blocks of the tile's rows (dc | rest), whoever survives
                const int n = min(kPreThreads, p.P - base);
                const uint32_t b_dc = 12u * (uint32_t)n, b_rest = 4u * (uint32_t)(row - 3) * (uint32_t)n;
                float* dc = reinterpret_cast<float*>(sm.sh);
                if (((b_dc | b_rest) & 15u) == 0) {
                    if (t == 0) {
                        mbar_arrive_expect_tx(&sm.bar_sh, b_dc + b_rest);
                        bulk_g2s(dc, p.shs + 3 * (size_t)base, b_dc, &sm.bar_sh);
                        if (b_rest) bulk_g2s(dc + 3 * kPreThreads, p.sh_rest + (size_t)base * (row - 3), b_rest, &sm.bar_sh);
                    }
                } else {   // ragged last tile: plain loads, then an empty phase
                    for (int j = t; j < 3 * n; j += kPreThreads) dc[j] = __ldg(p.shs + 3 * (size_t)base + j);
                    for (int j = t; j < (row - 3) * n; j += kPreThreads)
                        dc[3 * kPreThreads + j] = __ldg(p.sh_rest + (size_t)base * (row - 3) + j);
                    __syncthreads();
                    if (t == 0) mbar_arrive_expect_tx(&sm.bar_sh, 0u);
                }
            } else {
                if (t == 0) mbar_arrive_expect_tx(&sm.bar_sh, (uint32_t)nalive * need_bytes);
                if (alive) bulk_g2s(sm.sh + kShRow4 * t, p.shs + (size_t)idx * row, need_bytes, &sm.bar_sh);
            }
        }
        prev_any = cur;
        prev_alive = alive;
        prev_idx = idx;
        pmx = mx; pmy = my; pmz = mz;
    }
    if (p.depth_hist) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            my_vis += __shfl_xor_sync(0xffffffffu, my_vis, d);
            my_inst += __shfl_xor_sync(0xffffffffu, my_inst, d);
        }
        if ((t & 31) == 0 && my_vis) { atomicAdd(&sm.sums[0], my_vis); atomicAdd(&sm.sums[1], my_inst); }
        __syncthreads();
        for (int i = t; i < 4 * 256; i += kPreThreads) {
            const uint32_t c = (&sm.hist[0][0])[i];
            if (c) atomicAdd(p.depth_hist + i, c);
        }
        if (t == 0 && sm.sums[0] && V.gv.info) {
            atomicAdd(&V.gv.info->num_visible, (int32_t)sm.sums[0]);
            atomicAdd(&V.gv.info->num_rendered, (int32_t)sm.sums[1]);
        }
    }
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D,
                                    const float* __restrict__ view, uint8_t* __restrict__ present) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float x = __ldg(means3D + 3 * (size_t)idx), y = __ldg(means3D + 3 * (size_t)idx + 1),
                z = __ldg(means3D + 3 * (size_t)idx + 2);
    float m[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) m[k] = __ldg(view + k);
    present[idx] = xform_row_rn(m, 2, x, y, z) > kNearPlane ? 1 : 0;
}

__global__ void activate_kernel(int P, int M, const float* __restrict__ dc, const float* __restrict__ rest,
                                const float* __restrict__ opacity_raw, const float* __restrict__ scaling_raw,
                                const float* __restrict__ rotation_raw, float* __restrict__ shs,
                                float* __restrict__ opacities, float* __restrict__ scales, float* __restrict__ rotations) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const size_t i = (size_t)idx;
    opacities[i] = act_opacity(opacity_raw[i]);
    for (int k = 0; k < 3; ++k) scales[3 * i + k] = act_scale(scaling_raw[3 * i + k]);
    reinterpret_cast<float4*>(rotations)[i] = act_rotation(reinterpret_cast<const float4*>(rotation_raw)[i]);
    float* row = shs + 3 * (size_t)M * i;
    for (int k = 0; k < 3; ++k) row[k] = dc[3 * i + k];
    for (int k = 0; k < 3 * (M - 1); ++k) row[3 + k] = rest[3 * (size_t)(M - 1) * i + k];
}

}  // namespace

int launch_activate(int P, int M, const float* features_dc, const float* features_rest, const float* opacity_raw,
                    const float* scaling_raw, const float* rotation_raw, float* shs, float* opacities, float* scales,
                    float* rotations, cudaStream_t s) {
    activate_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, M, features_dc, features_rest, opacity_raw, scaling_raw,
                                                    rotation_raw, shs, opacities, scales, rotations);
    GSR_AFTER_LAUNCH(s, 0);
    return GSR_OK;
}

int launch_preprocess_views(int P, int D, int M, const Grid& g, const float* means3D, const float* shs,
                            const float* colors_precomp, const float* opacities, const float* scales,
                            float scale_modifier, const float* rotations, const float* cov3D_precomp, int n_views,
                            const ForwardViewArgs* views, cudaStream_t s, int debug, const RawMode& rm,
                            bool* depth_counts_done) {
    if (n_views < 1 || n_views > kMaxViews) return GSR_ERR_INVALID_ARGUMENT;
    if (depth_counts_done) *depth_counts_done = false;
    PreParams p;
    p.depth_hist = nullptr;
    p.raw = rm.raw; p.sh_rest = rm.sh_rest;
    p.P = P; p.D = D; p.M = M; p.W = g.W; p.H = g.H; p.gx = g.gx; p.gy = g.gy;
    p.n_views = n_views;
    p.means3D = means3D; p.shs = shs; p.colors_precomp = colors_precomp; p.opacities = opacities;
    p.scales = scales; p.scale_modifier = scale_modifier; p.rotations = rotations;
    p.cov3D_precomp = cov3D_precomp;
    for (int v = 0; v < n_views; ++v) {
        const ForwardViewArgs& a = views[v];
        PreView& o = p.v[v];
        o.view = a.view; o.proj = a.proj; o.campos = a.campos;
        o.tanfovx = a.tanfovx; o.tanfovy = a.tanfovy;
        o.focal_x = (float)g.W / (2.0f * a.tanfovx);
        o.focal_y = (float)g.H / (2.0f * a.tanfovy);
        o.radii = a.radii; o.gv = a.gv;
        o.radii_local = a.radii_local; o.conic_local = a.conic_local;
    }
    const int blocks = (P + kPreThreads - 1) / kPreThreads;
    StageScope scope(kStagePreprocess, s);
    // single view, SH colours, scale + rotation inputs, 16-byte-multiple SH rows: the persistent bulk-copy kernel
    if (n_views == 1 && shs && !colors_precomp && scales && rotations && !cov3D_precomp && (3 * M) % 4 == 0 &&
        !views[0].radii_local) {
        const int grid = std::min(blocks, sm_count() * 6);   // 6 CTAs per multiprocessor stay resident
        if (depth_counts_done && views[0].gv.info && views[0].gv.sort_ctl) {   // the caller zeroed the control region
            p.depth_hist = views[0].gv.sort_ctl;
            *depth_counts_done = true;
        }
        if (rm.raw) preprocess_tma_kernel<true><<<grid, kPreThreads, 0, s>>>(p);
        else preprocess_tma_kernel<false><<<grid, kPreThreads, 0, s>>>(p);
        GSR_AFTER_LAUNCH(s, debug);
        return GSR_OK;
    }
    if (rm.raw) {
        if (n_views == 1) preprocess_kernel<false, true><<<blocks, kPreThreads, 0, s>>>(p);
        else preprocess_kernel<true, true><<<blocks, kPreThreads, 0, s>>>(p);
    } else {
        if (n_views == 1) preprocess_kernel<false, false><<<blocks, kPreThreads, 0, s>>>(p);
        else preprocess_kernel<true, false><<<blocks, kPreThreads, 0, s>>>(p);
    }
    GSR_AFTER_LAUNCH(s, debug);
    return GSR_OK;
}

int launch_preprocess(int P, int D, int M, const Grid& g, const float* means3D, const float* shs,
                      const float* colors_precomp, const float* opacities, const float* scales,
                      float scale_modifier, const float* rotations, const float* cov3D_precomp,
                      const float* view, const float* proj, const float* campos, float tanfovx,
                      float tanfovy, int32_t* radii, GeomView gv, cudaStream_t s, int debug, const RawMode& rm,
                      bool* depth_counts_done) {
    ForwardViewArgs a;
    a.view = view; a.proj = proj; a.campos = campos; a.tanfovx = tanfovx; a.tanfovy = tanfovy;
    a.radii = radii; a.gv = gv;
    return launch_preprocess_views(P, D, M, g, means3D, shs, colors_precomp, opacities, scales, scale_modifier,
                                   rotations, cov3D_precomp, 1, &a, s, debug, rm, depth_counts_done);
}

int launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present,
                        cudaStream_t s) {
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, view, present);
    GSR_AFTER_LAUNCH(s, 0);
    return GSR_OK;
}

}  // namespace gsr
