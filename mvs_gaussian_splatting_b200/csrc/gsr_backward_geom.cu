// gsr_backward_geom.cu — per-Gaussian backward stage for sm_100a: the reference's
// `computeCov2DCUDA` + backward `preprocessCUDA` fused into one
// kernel.  HBM-bound: reads the 12-float blend accumulators, the Gaussian's parameters and (for
// visible Gaussians only) its SH row; writes every gradient array in full, so the caller never
// pre-zeroes ~300 B/Gaussian as the reference's glue does.  SH rows move warp-cooperatively as
// float4 through an odd-stride shared tile in both directions (coalesced loads and stores).
//
// The blend kernel accumulates RAW sums per Gaussian (see gsr_blend.cu):
//   acc[0..1] = sum h*dx, sum h*dy          acc[2..4] = sum h*dx*dx, h*dx*dy, h*dy*dy
//   acc[5]    = dL/dopacity                 acc[6..8] = dL/dcolor
// with h = dL/dG * G and d = mean2D - pixel; the conic/mean chain-rule factors are applied here.
#include <algorithm>
#include <cstdlib>

#include "gsr_internal.cuh"

namespace gsr {

namespace {

constexpr int kBwdThreads = 128;
constexpr int kBwdWarps = kBwdThreads / 32;
constexpr int kShStride = 49;

constexpr int kMaxViews = GSR_MAX_VIEWS;

// What differs from view to view: the camera, and where that view's forward rows and blend
// accumulators of these Gaussians live (possibly a peer GPU's workspace).
struct BwdView {
    const float* view;
    const float* proj;
    const float* campos;
    float tanfovx, tanfovy, focal_x, focal_y;
    const int32_t* radii;
    const float* grad_accum;
    const float4* conic_opacity;
    const float4* rgb;
    float* dL_dmeans2D;
    const int32_t* radii_local;   // optional local copies written by the multi-view forward (gsr_view_slot):
    const float4* conic_local;    // radius | clamped << 28, conic + opacity; read instead of radii / conic_opacity / rgb
};

struct BwdParams {
    int P, D, M, W, H;
    int accumulate;  // parameter gradients are added to the outputs (rows no view sees: untouched)
    int n_views;
    const float* means3D;
    const float* shs;
    int colors_precomp;
    const float* scales;
    float scale_modifier;
    const float* rotations;
    const float* cov3D_precomp;
    float *dL_dcolors, *dL_dopacity, *dL_dmeans3D, *dL_dcov3D, *dL_dsh, *dL_dscales, *dL_drots;
    int raw;               // RawMode: inputs are pre-activation, gradients are w.r.t. them
    const float* sh_rest;  // RawMode: shs = dc [P,1,3], sh_rest = [P,M-1,3]
    float* dL_dsh_rest;    // RawMode: dL_dsh = [P,1,3], dL_dsh_rest = [P,M-1,3]
    BwdView v[kMaxViews];
};

__device__ __forceinline__ void cov3d(float s0, float s1, float s2, float4 q, float R[3][3], float* c6) {
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    R[0][0] = 1.f - 2.f * (y * y + z * z); R[0][1] = 2.f * (x * y - r * z); R[0][2] = 2.f * (x * z + r * y);
    R[1][0] = 2.f * (x * y + r * z); R[1][1] = 1.f - 2.f * (x * x + z * z); R[1][2] = 2.f * (y * z - r * x);
    R[2][0] = 2.f * (x * z - r * y); R[2][1] = 2.f * (y * z + r * x); R[2][2] = 1.f - 2.f * (x * x + y * y);
    float Mm[3][3];  // M[k][j] = s_k * R[j][k]
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        Mm[0][j] = s0 * R[j][0]; Mm[1][j] = s1 * R[j][1]; Mm[2][j] = s2 * R[j][2];
    }
    c6[0] = Mm[0][0] * Mm[0][0] + Mm[1][0] * Mm[1][0] + Mm[2][0] * Mm[2][0];
    c6[1] = Mm[0][0] * Mm[0][1] + Mm[1][0] * Mm[1][1] + Mm[2][0] * Mm[2][1];
    c6[2] = Mm[0][0] * Mm[0][2] + Mm[1][0] * Mm[1][2] + Mm[2][0] * Mm[2][2];
    c6[3] = Mm[0][1] * Mm[0][1] + Mm[1][1] * Mm[1][1] + Mm[2][1] * Mm[2][1];
    c6[4] = Mm[0][1] * Mm[0][2] + Mm[1][1] * Mm[1][2] + Mm[2][1] * Mm[2][2];
    c6[5] = Mm[0][2] * Mm[0][2] + Mm[1][2] * Mm[1][2] + Mm[2][2] * Mm[2][2];
}

// SH backward for one Gaussian: reads the SH values from `row`, writes dL/dsh to `out` and returns
// dL/d(direction).  kAcc = false: out may alias row (in place), coefficients above the active degree
// are zeroed.  kAcc = true (multi-view): out is a separate, pre-zeroed row that is added to.
template <bool kAcc>
__device__ __forceinline__ void sh_backward(int deg, int M, float x, float y, float z, const float* row, float* out,
                                            const float* g3, float* gdir) {
    const float C0 = 0.28209479177387814f, C1 = 0.4886025119029199f;
    const float C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f,
                         0.5462742152960396f};
    const float C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                         -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};
    gdir[0] = gdir[1] = gdir[2] = 0.f;
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        const float g = g3[ch];
        float ddx = 0.f, ddy = 0.f, ddz = 0.f;
#define ROW(k) row[3 * (k) + ch]
#define PUT(k, val) do { float& o_ = out[3 * (k) + ch]; o_ = kAcc ? o_ + (val) : (val); } while (0)
        PUT(0, C0 * g);
        if (deg > 0) {
            const float s1 = ROW(1), s2 = ROW(2), s3 = ROW(3);
            PUT(1, -C1 * y * g); PUT(2, C1 * z * g); PUT(3, -C1 * x * g);
            ddx = -C1 * s3; ddy = -C1 * s1; ddz = C1 * s2;
            if (deg > 1) {
                const float s4 = ROW(4), s5 = ROW(5), s6 = ROW(6), s7 = ROW(7), s8 = ROW(8);
                PUT(4, C2[0] * xy * g); PUT(5, C2[1] * yz * g); PUT(6, C2[2] * (2.f * zz - xx - yy) * g);
                PUT(7, C2[3] * xz * g); PUT(8, C2[4] * (xx - yy) * g);
                ddx += C2[0] * y * s4 + C2[2] * 2.f * -x * s6 + C2[3] * z * s7 + C2[4] * 2.f * x * s8;
                ddy += C2[0] * x * s4 + C2[1] * z * s5 + C2[2] * 2.f * -y * s6 + C2[4] * 2.f * -y * s8;
                ddz += C2[1] * y * s5 + C2[2] * 4.f * z * s6 + C2[3] * x * s7;
                if (deg > 2) {
                    const float s9 = ROW(9), s10 = ROW(10), s11 = ROW(11), s12 = ROW(12), s13 = ROW(13),
                                s14 = ROW(14), s15 = ROW(15);
                    PUT(9, C3[0] * y * (3.f * xx - yy) * g);
                    PUT(10, C3[1] * xy * z * g);
                    PUT(11, C3[2] * y * (4.f * zz - xx - yy) * g);
                    PUT(12, C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy) * g);
                    PUT(13, C3[4] * x * (4.f * zz - xx - yy) * g);
                    PUT(14, C3[5] * z * (xx - yy) * g);
                    PUT(15, C3[6] * x * (xx - 3.f * yy) * g);
                    ddx += C3[0] * s9 * 6.f * xy + C3[1] * s10 * yz + C3[2] * s11 * -2.f * xy +
                           C3[3] * s12 * -6.f * xz + C3[4] * s13 * (-3.f * xx + 4.f * zz - yy) +
                           C3[5] * s14 * 2.f * xz + C3[6] * s15 * 3.f * (xx - yy);
                    ddy += C3[0] * s9 * 3.f * (xx - yy) + C3[1] * s10 * xz +
                           C3[2] * s11 * (-3.f * yy + 4.f * zz - xx) + C3[3] * s12 * -6.f * yz +
                           C3[4] * s13 * -2.f * xy + C3[5] * s14 * -2.f * yz + C3[6] * s15 * -6.f * xy;
                    ddz += C3[1] * s10 * xy + C3[2] * s11 * 8.f * yz + C3[3] * s12 * 3.f * (2.f * zz - xx - yy) +
                           C3[4] * s13 * 8.f * xz + C3[5] * s14 * (xx - yy);
                }
            }
        }
        const int used = (deg + 1) * (deg + 1);
        if (!kAcc) for (int k = used; k < M; ++k) out[3 * k + ch] = 0.f;
#undef PUT
#undef ROW
        gdir[0] += ddx * g; gdir[1] += ddy * g; gdir[2] += ddz * g;
    }
}

// kMulti = false: one view, the dL/dsh row is formed in place over the SH values (one shared tile).
// kMulti = true: the Gaussian's parameters and SH row are read ONCE and the views are looped over;
// scalar gradients accumulate in registers, dL/dsh in a second shared tile; outputs written once.
// Multi-view form with PEER accumulators (kRing; the sharded step on more than one GPU): the 48-byte blend
// accumulators of a Gaussian live in the workspace of the rank that rendered the view.  A dependent remote load per
// view costs an NVLink round trip (~2 us) eight times in a row per thread; instead every WARP requests the 32 rows of
// its Gaussians (1536 contiguous bytes) of up to kAccRing views ahead with cp.async.bulk into a shared ring (one
// mbarrier per ring slot, one arrival per warp) and re-arms a slot for view v + kAccRing as soon as it has consumed
// view v.  (One request per ROW was measured first: 36 M 48-byte requests run into the copy engine's request rate,
// ~200 per microsecond and SM — 2.66 -> 4.06 ms on one GPU.)  Costs the fourth CTA per SM, so local accumulators
// keep the plain loads.
#ifndef GSR_ACC_RING
#define GSR_ACC_RING 1   // slots: 1 keeps four CTAs per SM (56.5 KB of shared memory); 4 slots cost the fourth CTA
#endif
constexpr int kAccRing = GSR_ACC_RING;
constexpr size_t kAccRingBytes = (size_t)kAccRing * kBwdThreads * 48;

template <bool kMulti, bool kRaw, bool kRing = false>
__global__ void __launch_bounds__(kBwdThreads, kMulti ? ((kRing && kAccRing > 1) ? 3 : 4) : 6) preprocess_backward_kernel(const BwdParams p) {
    extern __shared__ __align__(16) float s_dyn[];
    float* s_cam = s_dyn;                                   // view[16] proj[16] campos[3]
    float* tile_in = s_dyn + 36 + (threadIdx.x >> 5) * (32 * kShStride);
    float* tile = kMulti ? tile_in + kBwdWarps * 32 * kShStride : tile_in;   // dL/dsh rows of this warp
    // kMulti: ring of accumulator rows [slot][thread][3 x float4] + its mbarriers, behind the two SH tiles
    float4* s_acc = reinterpret_cast<float4*>(s_dyn + 36 + 2 * kBwdWarps * 32 * kShStride);
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(reinterpret_cast<unsigned char*>(s_acc) + kAccRingBytes);

    // camera of the first view: staged before anything else so no CTA barrier follows the SH fetch
    if (threadIdx.x < 16) s_cam[threadIdx.x] = __ldg(p.v[0].view + threadIdx.x);
    else if (threadIdx.x < 32) s_cam[threadIdx.x] = __ldg(p.v[0].proj + threadIdx.x - 16);
    else if (threadIdx.x < 35) s_cam[threadIdx.x] = __ldg(p.v[0].campos + threadIdx.x - 32);
    if (kRing && threadIdx.x == 0) {
        for (int k = 0; k < kAccRing; ++k) mbar_init(&s_bar[k], kBwdWarps);
        mbar_init_fence();
    }
    __syncthreads();

    const int idx = blockIdx.x * kBwdThreads + threadIdx.x;
    const unsigned lane = lane_id();
    const int warp = threadIdx.x >> 5;
    const int warp_base = blockIdx.x * kBwdThreads + warp * 32;
    const bool valid = idx < p.P;
    uint32_t visbits = 0;  // bit v: view v sees this Gaussian
    for (int v = 0; v < p.n_views; ++v) {
        if (!valid) break;
        const int r = (kMulti && p.v[v].radii_local) ? (__ldg(p.v[v].radii_local + idx) & GSR_LOCAL_RADIUS_MASK)
                                                     : p.v[v].radii[idx];
        if (r > 0) visbits |= 1u << v;
    }
    const bool vis_any = visbits != 0;
    const unsigned vmask = __ballot_sync(0xffffffffu, vis_any);
    // request the accumulator rows of this warp's Gaussians for view v into ring slot v % kAccRing (one arrival per warp)
    auto request_acc = [&](int v) {
        const unsigned any = __ballot_sync(0xffffffffu, (visbits >> v) & 1u);
        if (lane == 0) {
            uint64_t* bar = &s_bar[v % kAccRing];
            const int nrows = min(32, p.P - warp_base);
            if (any && nrows > 0) {
                mbar_arrive_expect_tx(bar, 48u * (uint32_t)nrows);
                bulk_g2s(s_acc + ((v % kAccRing) * kBwdThreads + warp * 32) * 3, p.v[v].grad_accum + (size_t)warp_base * 12,
                         48u * (uint32_t)nrows, bar);
            } else {
                asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
            }
        }
    };
    if (kRing)
        for (int v = 0; v < min(kAccRing, p.n_views); ++v) request_acc(v);
    const bool do_sh = (p.dL_dsh != nullptr) && !p.colors_precomp;
    const int row = 3 * p.M;
    const int need = 3 * (p.D + 1) * (p.D + 1);

    // ---- per-Gaussian operands: issued before the SH fetch so their latency overlaps it ----
    float4 q = make_float4(1.f, 0.f, 0.f, 0.f), q_raw = q;
    float opacity = 0.f;  // activated opacity (RawMode chain rule), from any view that sees the Gaussian
    float mx = 0.f, my = 0.f, mz = 0.f, rs0 = 0.f, rs1 = 0.f, rs2 = 0.f;
    float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0, a2 = a0, con = a0;
    uint32_t clamped_bits = 0;
    if (vis_any) {
        mx = p.means3D[3 * (size_t)idx]; my = p.means3D[3 * (size_t)idx + 1]; mz = p.means3D[3 * (size_t)idx + 2];
        if (!p.cov3D_precomp) {
            q = __ldg(reinterpret_cast<const float4*>(p.rotations) + idx);
            rs0 = p.scales[3 * (size_t)idx]; rs1 = p.scales[3 * (size_t)idx + 1]; rs2 = p.scales[3 * (size_t)idx + 2];
            if (kRaw) {
                q_raw = q;
                q = act_rotation(q);
                rs0 = act_scale(rs0); rs1 = act_scale(rs1); rs2 = act_scale(rs2);
            }
        }
        if (!kMulti) {  // the single view's accumulators travel with the operands
            const float4* accp = reinterpret_cast<const float4*>(p.v[0].grad_accum + (size_t)idx * 12);
            a0 = accp[0]; a1 = accp[1]; a2 = accp[2];
            con = p.v[0].conic_opacity[idx];
            if (do_sh) clamped_bits = __float_as_uint(p.v[0].rgb[idx].w);
        }
    }

    // ---- fetch the SH rows of the visible Gaussians of this warp ----
    if (do_sh && vmask) {
        const int nvis = __popc(vmask);
        if (kRaw) {  // split layout: dc [P,1,3] | rest [P,M-1,3]
            const int nr_ = min(32, p.P - warp_base);
            fetch_rows_f4(p.shs + 3 * (size_t)warp_base, nr_, 3, 3, 0, vmask, tile_in, kShStride, lane);
            if (need > 3)
                fetch_rows_f4(p.sh_rest + (size_t)warp_base * (row - 3), nr_, row - 3, need - 3, 3, vmask, tile_in,
                              kShStride, lane);
        } else if ((row & 3) == 0) {
            const int vpr = (need + 3) >> 2;
            const int total = nvis * vpr;
            const int my_src = __fns(vmask, 0, lane + 1);  // lane of the (lane+1)-th visible Gaussian
            for (int jb = 0; jb < total; jb += 128) {  // warp-uniform trip count (shuffles inside)
                float4 val[4];
                int dsto[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int j = jb + (int)lane + 32 * u;
                    const int jc = min(j, total - 1);
                    const int v = jc / vpr, part = jc - v * vpr;
                    const int src = __shfl_sync(0xffffffffu, my_src, v);
                    dsto[u] = src * kShStride + 4 * part;
                    if (j < total)
                        val[u] = __ldg(reinterpret_cast<const float4*>(p.shs + (size_t)(warp_base + src) * row) + part);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (jb + (int)lane + 32 * u < total) {
                        float* d = tile_in + dsto[u];
                        d[0] = val[u].x; d[1] = val[u].y; d[2] = val[u].z; d[3] = val[u].w;
                    }
                }
            }
        } else {
            for (int j = lane; j < nvis * need; j += 32) {
                const int v = j / need, part = j - v * need;
                const int src = __fns(vmask, 0, v + 1);
                tile_in[src * kShStride + part] = __ldg(p.shs + (size_t)(warp_base + src) * row + part);
            }
        }
    }
    if (kMulti && do_sh)
        for (int j = lane; j < 32 * kShStride; j += 32) tile[j] = 0.f;
    __syncwarp();

    float gm[3] = {0.f, 0.f, 0.f}, gc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float gs[3] = {0.f, 0.f, 0.f}, gq[4] = {0.f, 0.f, 0.f, 0.f};
    float gcol[3] = {0.f, 0.f, 0.f}, gop = 0.f;

    // view-independent: 3D covariance from scale / rotation
    float c6[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, R[3][3] = {{0.f}}, s[3] = {0.f, 0.f, 0.f};
    if (vis_any) {
        if (p.cov3D_precomp) {
#pragma unroll
            for (int k = 0; k < 6; ++k) c6[k] = p.cov3D_precomp[6 * (size_t)idx + k];
        } else {
            s[0] = p.scale_modifier * rs0;
            s[1] = p.scale_modifier * rs1;
            s[2] = p.scale_modifier * rs2;
            cov3d(s[0], s[1], s[2], q, R, c6);
        }
    }

    for (int vi = 0; vi < p.n_views; ++vi) {
        const BwdView& V = p.v[vi];
        if (vi > 0) {
            __syncthreads();  // the previous view's camera is no longer read
            if (threadIdx.x < 16) s_cam[threadIdx.x] = __ldg(V.view + threadIdx.x);
            else if (threadIdx.x < 32) s_cam[threadIdx.x] = __ldg(V.proj + threadIdx.x - 16);
            else if (threadIdx.x < 35) s_cam[threadIdx.x] = __ldg(V.campos + threadIdx.x - 32);
            __syncthreads();
        }
        const float* view = s_cam;
        const float* proj = s_cam + 16;
        const float* campos = s_cam + 32;
        const bool vis = (visbits >> vi) & 1u;
        float g2d[2] = {0.f, 0.f};

        if (kRing) mbar_wait(&s_bar[vi % kAccRing], (uint32_t)(vi / kAccRing) & 1u);   // this view's rows have landed
        if (vis) {
            if (kMulti) {
                const float4* accp = kRing ? s_acc + ((vi % kAccRing) * kBwdThreads + threadIdx.x) * 3
                                           : reinterpret_cast<const float4*>(V.grad_accum + (size_t)idx * 12);
                a0 = accp[0]; a1 = accp[1]; a2 = accp[2];
            }
        }
        if (kRing && vi + kAccRing < p.n_views) {   // the warp's rows of this slot are in registers: re-arm it for view vi + kAccRing
            __syncwarp();
            fence_async_smem();
            request_acc(vi + kAccRing);
        }
        if (vis) {
            if (kMulti) {
                if (V.radii_local) {   // forward rows kept locally: only the accumulators come from the view's owner
                    con = __ldg(V.conic_local + idx);
                    clamped_bits = (uint32_t)__ldg(V.radii_local + idx) >> 28;
                } else {
                    con = V.conic_opacity[idx];
                    if (do_sh) clamped_bits = __float_as_uint(V.rgb[idx].w);
                }
            }
            if (kRaw) opacity = con.w;
            const float S1 = a0.x, S2 = a0.y;
            // true partials of the conic (A,B,C) and the NDC-scaled mean2D gradient
            const float gA = -0.5f * a0.z, gB = -a0.w, gC = -0.5f * a1.x;
            g2d[0] = (-con.x * S1 - con.y * S2) * (0.5f * (float)p.W);
            g2d[1] = (-con.z * S2 - con.y * S1) * (0.5f * (float)p.H);
            gop += a1.y;
            const float vcol[3] = {a1.z, a1.w, a2.x};
            gcol[0] += vcol[0]; gcol[1] += vcol[1]; gcol[2] += vcol[2];

            // ---- EWA projection (forward values) ----
            float t[3];
            t[0] = view[0] * mx + view[4] * my + view[8] * mz + view[12];
            t[1] = view[1] * mx + view[5] * my + view[9] * mz + view[13];
            t[2] = view[2] * mx + view[6] * my + view[10] * mz + view[14];
            const float limx = 1.3f * V.tanfovx, limy = 1.3f * V.tanfovy;
            const float txtz = t[0] / t[2], tytz = t[1] / t[2];
            const float gmulx = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
            const float gmuly = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
            t[0] = fminf(limx, fmaxf(-limx, txtz)) * t[2];
            t[1] = fminf(limy, fmaxf(-limy, tytz)) * t[2];
            const float fx = V.focal_x, fy = V.focal_y;
            const float j00 = fx / t[2], j11 = fy / t[2];
            const float j02 = -(fx * t[0]) / (t[2] * t[2]), j12 = -(fy * t[1]) / (t[2] * t[2]);
            float M0[3], M1[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                M0[k] = view[4 * k + 0] * j00 + view[4 * k + 2] * j02;
                M1[k] = view[4 * k + 1] * j11 + view[4 * k + 2] * j12;
            }
            const float S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
            float SM0[3], SM1[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                SM0[k] = S[k][0] * M0[0] + S[k][1] * M0[1] + S[k][2] * M0[2];
                SM1[k] = S[k][0] * M1[0] + S[k][1] * M1[1] + S[k][2] * M1[2];
            }
            const float a = M0[0] * SM0[0] + M0[1] * SM0[1] + M0[2] * SM0[2] + 0.3f;
            const float b = M1[0] * SM0[0] + M1[1] * SM0[1] + M1[2] * SM0[2];
            const float c = M1[0] * SM1[0] + M1[1] * SM1[1] + M1[2] * SM1[2] + 0.3f;
            const float det = a * c - b * b;
            const float d2inv = 1.f / (det * det + 0.0000001f);
            float dLa = 0.f, dLb = 0.f, dLc = 0.f;
            if (d2inv != 0.f) {
                dLa = d2inv * (-c * c * gA + b * c * gB + (det - a * c) * gC);
                dLc = d2inv * (-a * a * gC + a * b * gB + (det - a * c) * gA);
                dLb = d2inv * (2.f * b * c * gA - (det + 2.f * b * b) * gB + 2.f * a * b * gC);
                gc[0] += M0[0] * M0[0] * dLa + M0[0] * M1[0] * dLb + M1[0] * M1[0] * dLc;
                gc[3] += M0[1] * M0[1] * dLa + M0[1] * M1[1] * dLb + M1[1] * M1[1] * dLc;
                gc[5] += M0[2] * M0[2] * dLa + M0[2] * M1[2] * dLb + M1[2] * M1[2] * dLc;
                gc[1] += 2.f * M0[0] * M0[1] * dLa + (M0[0] * M1[1] + M0[1] * M1[0]) * dLb + 2.f * M1[0] * M1[1] * dLc;
                gc[2] += 2.f * M0[0] * M0[2] * dLa + (M0[0] * M1[2] + M0[2] * M1[0]) * dLb + 2.f * M1[0] * M1[2] * dLc;
                gc[4] += 2.f * M0[2] * M0[1] * dLa + (M0[1] * M1[2] + M0[2] * M1[1]) * dLb + 2.f * M1[1] * M1[2] * dLc;
            }
            float vm[3];  // this view's dL/dmean3D
            {
                float dM0[3], dM1[3];
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    dM0[k] = 2.f * SM0[k] * dLa + SM1[k] * dLb;
                    dM1[k] = 2.f * SM1[k] * dLc + SM0[k] * dLb;
                }
                float dj00 = 0.f, dj02 = 0.f, dj11 = 0.f, dj12 = 0.f;
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    dj00 += view[4 * k + 0] * dM0[k];
                    dj02 += view[4 * k + 2] * dM0[k];
                    dj11 += view[4 * k + 1] * dM1[k];
                    dj12 += view[4 * k + 2] * dM1[k];
                }
                const float tz = 1.f / t[2], tz2 = tz * tz, tz3 = tz2 * tz;
                const float dtx = gmulx * -fx * tz2 * dj02;
                const float dty = gmuly * -fy * tz2 * dj12;
                const float dtz = -fx * tz2 * dj00 - fy * tz2 * dj11 + (2.f * fx * t[0]) * tz3 * dj02 +
                                  (2.f * fy * t[1]) * tz3 * dj12;
                vm[0] = view[0] * dtx + view[1] * dty + view[2] * dtz;
                vm[1] = view[4] * dtx + view[5] * dty + view[6] * dtz;
                vm[2] = view[8] * dtx + view[9] * dty + view[10] * dtz;
            }
            // ---- mean2D -> mean3D through the perspective divide ----
            {
                const float hx = proj[0] * mx + proj[4] * my + proj[8] * mz + proj[12];
                const float hy = proj[1] * mx + proj[5] * my + proj[9] * mz + proj[13];
                const float hw = proj[3] * mx + proj[7] * my + proj[11] * mz + proj[15];
                const float mw = 1.f / (hw + 0.0000001f);
                const float mul1 = hx * mw * mw, mul2 = hy * mw * mw;
                vm[0] += (proj[0] * mw - proj[3] * mul1) * g2d[0] + (proj[1] * mw - proj[3] * mul2) * g2d[1];
                vm[1] += (proj[4] * mw - proj[7] * mul1) * g2d[0] + (proj[5] * mw - proj[7] * mul2) * g2d[1];
                vm[2] += (proj[8] * mw - proj[11] * mul1) * g2d[0] + (proj[9] * mw - proj[11] * mul2) * g2d[1];
            }
            // ---- colour -> SH + view direction ----
            if (do_sh) {
                float g3[3];
#pragma unroll
                for (int ch = 0; ch < 3; ++ch) g3[ch] = (clamped_bits >> ch) & 1u ? 0.f : vcol[ch];
                const float d0 = mx - campos[0], d1 = my - campos[1], d2 = mz - campos[2];
                const float s2 = d0 * d0 + d1 * d1 + d2 * d2;
                const float len = sqrtf(s2);
                float gdir[3];
                sh_backward<kMulti>(p.D, p.M, d0 / len, d1 / len, d2 / len, tile_in + lane * kShStride,
                                    tile + lane * kShStride, g3, gdir);
                const float inv32 = 1.f / sqrtf(s2 * s2 * s2);
                const float vd = d0 * gdir[0] + d1 * gdir[1] + d2 * gdir[2];
                vm[0] += (s2 * gdir[0] - d0 * vd) * inv32;
                vm[1] += (s2 * gdir[1] - d1 * vd) * inv32;
                vm[2] += (s2 * gdir[2] - d2 * vd) * inv32;
            }
            gm[0] += vm[0]; gm[1] += vm[1]; gm[2] += vm[2];
        }
        if (valid) {  // per-view output: always overwritten
            const size_t i = (size_t)idx;
            V.dL_dmeans2D[3 * i] = g2d[0]; V.dL_dmeans2D[3 * i + 1] = g2d[1]; V.dL_dmeans2D[3 * i + 2] = 0.f;
        }
    }

    // ---- cov3D -> scale, quaternion: linear in dL/dcov3D and view-independent, so applied once ----
    if (vis_any && !p.cov3D_precomp) {
        const float Gs[3][3] = {{gc[0], 0.5f * gc[1], 0.5f * gc[2]},
                                {0.5f * gc[1], gc[3], 0.5f * gc[4]},
                                {0.5f * gc[2], 0.5f * gc[4], gc[5]}};
        float GR[3][3], dR[3][3];
#pragma unroll
        for (int a2 = 0; a2 < 3; ++a2)
#pragma unroll
            for (int k = 0; k < 3; ++k) GR[a2][k] = Gs[a2][0] * R[0][k] + Gs[a2][1] * R[1][k] + Gs[a2][2] * R[2][k];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float rgr = R[0][k] * GR[0][k] + R[1][k] * GR[1][k] + R[2][k] * GR[2][k];
            gs[k] = 2.f * s[k] * rgr;
#pragma unroll
            for (int a2 = 0; a2 < 3; ++a2) dR[a2][k] = 2.f * GR[a2][k] * s[k] * s[k];
        }
        const float r = q.x, x = q.y, y = q.z, z = q.w;
        gq[0] = 2.f * (z * (dR[1][0] - dR[0][1]) + y * (dR[0][2] - dR[2][0]) + x * (dR[2][1] - dR[1][2]));
        gq[1] = 2.f * (y * (dR[0][1] + dR[1][0]) + z * (dR[0][2] + dR[2][0]) + r * (dR[2][1] - dR[1][2])) -
                4.f * x * (dR[1][1] + dR[2][2]);
        gq[2] = 2.f * (x * (dR[0][1] + dR[1][0]) + r * (dR[0][2] - dR[2][0]) + z * (dR[1][2] + dR[2][1])) -
                4.f * y * (dR[0][0] + dR[2][2]);
        gq[3] = 2.f * (r * (dR[1][0] - dR[0][1]) + x * (dR[0][2] + dR[2][0]) + y * (dR[1][2] + dR[2][1])) -
                4.f * z * (dR[0][0] + dR[1][1]);
    }
    if (!kMulti && do_sh && valid && !vis_any) {  // in-place tile: culled rows carry zeros
        float* rowp = tile + lane * kShStride;
        for (int k = 0; k < row; ++k) rowp[k] = 0.f;
    }

    // ---- RawMode: chain rules of exp / normalize / sigmoid (scene/gaussian_model.py:28-45) ----
    if (kRaw && vis_any) {
        gop *= opacity * (1.0f - opacity);
        if (!p.cov3D_precomp) {
            gs[0] *= rs0; gs[1] *= rs1; gs[2] *= rs2;
            const float n = quat_norm(q_raw);
            const float dotp = q.x * gq[0] + q.y * gq[1] + q.z * gq[2] + q.w * gq[3];
            const float inv = 1.0f / n;
            gq[0] = (gq[0] - q.x * dotp) * inv; gq[1] = (gq[1] - q.y * dotp) * inv;
            gq[2] = (gq[2] - q.z * dotp) * inv; gq[3] = (gq[3] - q.w * dotp) * inv;
        }
    }

    // ---- outputs (every row written; in accumulate mode rows no view sees are skipped) ----
    const bool acc = p.accumulate != 0;
    if (valid && (!acc || vis_any)) {
        const size_t i = (size_t)idx;
        auto put = [acc](float* a, float v) { *a = acc ? *a + v : v; };
        put(p.dL_dopacity + i, gop);
        put(p.dL_dmeans3D + 3 * i, gm[0]); put(p.dL_dmeans3D + 3 * i + 1, gm[1]); put(p.dL_dmeans3D + 3 * i + 2, gm[2]);
        if (p.dL_dcolors) {
            put(p.dL_dcolors + 3 * i, gcol[0]); put(p.dL_dcolors + 3 * i + 1, gcol[1]); put(p.dL_dcolors + 3 * i + 2, gcol[2]);
        }
        if (p.dL_dcov3D) {
#pragma unroll
            for (int k = 0; k < 6; ++k) put(p.dL_dcov3D + 6 * i + k, gc[k]);
        }
        if (p.dL_dscales) {
            put(p.dL_dscales + 3 * i, gs[0]); put(p.dL_dscales + 3 * i + 1, gs[1]); put(p.dL_dscales + 3 * i + 2, gs[2]);
        }
        if (p.dL_drots) {
            float4* o = reinterpret_cast<float4*>(p.dL_drots) + i;
            float4 v = make_float4(gq[0], gq[1], gq[2], gq[3]);
            if (acc) { const float4 old = *o; v.x += old.x; v.y += old.y; v.z += old.z; v.w += old.w; }
            *o = v;
        }
    }
    if (p.dL_dsh) {
        if (p.colors_precomp) {
            // SH not in use: the gradient is identically zero
            if (valid && !acc) for (int k = 0; k < row; ++k) p.dL_dsh[(size_t)idx * row + k] = 0.f;
        } else {
            __syncwarp();
            const int nrows = min(32, p.P - warp_base);
            if (kRaw && nrows > 0) {  // RawMode: dc rows -> dL_dsh [P,1,3], the rest -> dL_dsh_rest [P,M-1,3]
                const int rr = row - 3;
                float* ddc = p.dL_dsh + 3 * (size_t)warp_base;
                float* drest = p.dL_dsh_rest + (size_t)warp_base * rr;
                if (!acc) {
                    store_rows_f4(ddc, nrows, 3, 0, tile, kShStride, lane);
                    if (rr > 0) store_rows_f4(drest, nrows, rr, 3, tile, kShStride, lane);
                } else {
                    for (int j = lane; j < 3 * nrows; j += 32) {
                        const int r = j / 3, k = j - 3 * r;
                        if ((vmask >> r) & 1u) ddc[j] += tile[r * kShStride + k];
                    }
                    for (int j = lane; j < nrows * rr; j += 32) {
                        const int r = j / rr, k = j - r * rr;
                        if ((vmask >> r) & 1u) drest[j] += tile[r * kShStride + 3 + k];
                    }
                }
            } else if (nrows > 0) {
                float* dst = p.dL_dsh + (size_t)warp_base * row;
                if ((row & 3) == 0) {
                    const int vpr = row >> 2;
                    const int total = nrows * vpr;
                    if (!acc) {  // the common case: pure streaming stores, unrolled
#pragma unroll 4
                        for (int j = lane; j < total; j += 32) {
                            const int r = j / vpr, part = j - r * vpr;
                            const float* sp = tile + r * kShStride + 4 * part;
                            __stcs(reinterpret_cast<float4*>(dst) + j, make_float4(sp[0], sp[1], sp[2], sp[3]));
                        }
                    } else {
                        for (int j = lane; j < total; j += 32) {
                            const int r = j / vpr, part = j - r * vpr;
                            if (!((vmask >> r) & 1u)) continue;
                            const float* sp = tile + r * kShStride + 4 * part;
                            float4* o = reinterpret_cast<float4*>(dst) + j;
                            const float4 old = *o;
                            __stcs(o, make_float4(sp[0] + old.x, sp[1] + old.y, sp[2] + old.z, sp[3] + old.w));
                        }
                    }
                } else {
                    for (int j = lane; j < nrows * row; j += 32) {
                        const int r = j / row, part = j - r * row;
                        if (acc) {
                            if ((vmask >> r) & 1u) dst[j] += tile[r * kShStride + part];
                        } else {
                            dst[j] = tile[r * kShStride + part];
                        }
                    }
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------------
// Persistent single-view form with the next tile's bytes in flight (cp.async.bulk + mbarrier), the backward
// counterpart of preprocess_tma_kernel.  Visibility is known from the forward (radii), so per tile of 128 Gaussians:
//   * the operand block (xyz, scale, rotation, the 48-byte blend accumulators, conic + opacity, colour/clamp rows:
//     six contiguous runs, 15 KB) is requested as soon as the previous tile's geometry phase has read its own;
//   * the SH rows of the VISIBLE Gaussians are requested one run per row onto a 208-byte stride (128-bit shared
//     accesses touch eight bank groups per quarter-warp) and consumed one iteration later;
//   * phase A (EWA / covariance / mean chain rules, needs the operands) of tile i and phase B (SH backward: dL/dsh
//     formed IN PLACE over the staged row, the view-direction term of dL/dmean3D) of tile i-1 alternate, so either
//     tile's memory phase overlaps the other's arithmetic;
//   * dL/dsh leaves as one bulk store per row straight from the staged tile (zeros for culled rows: every row of
//     every gradient array is written, the caller passes uninitialised memory).
// Raw-parameter mode (kRaw): the activations and their chain rules are applied in registers, the SH values arrive as
// the tile's two contiguous blocks (dc | rest) and dL/d(dc), dL/d(rest) leave as two bulk stores per tile.
// SH backward is evaluated coefficient-major here (t_k = g . sh_k, gdir += dY_k/ddir t_k, dL/dsh_k = Y_k g), the
// same sums as sh_backward in a different association — gradients are a tolerance item (tests/helpers.py).
constexpr int kBwdShRow4 = 13;
struct BwdTmaSmem {
    alignas(16) float xyz[3 * kBwdThreads];
    alignas(16) float scale[3 * kBwdThreads];
    alignas(16) float4 rot[kBwdThreads];
    alignas(16) float4 acc[3 * kBwdThreads];
    alignas(16) float4 con[kBwdThreads];
    alignas(16) float4 rgb[kBwdThreads];
    alignas(16) float4 sh[kBwdThreads * kBwdShRow4];
    alignas(8) uint64_t bar_opnd, bar_sh;
    float cam[36];
};

// Real SH basis function k (with its constant) and its partial derivatives w.r.t. the unit direction.
__device__ __forceinline__ void sh_basis(int k, float x, float y, float z, float& B, float& Dx, float& Dy, float& Dz) {
    const float C0 = 0.28209479177387814f, C1 = 0.4886025119029199f;
    const float C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f,
                         0.5462742152960396f};
    const float C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                         -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
    Dx = Dy = Dz = 0.f;
    switch (k) {
        case 0: B = C0; break;
        case 1: B = -C1 * y; Dy = -C1; break;
        case 2: B = C1 * z; Dz = C1; break;
        case 3: B = -C1 * x; Dx = -C1; break;
        case 4: B = C2[0] * xy; Dx = C2[0] * y; Dy = C2[0] * x; break;
        case 5: B = C2[1] * yz; Dy = C2[1] * z; Dz = C2[1] * y; break;
        case 6: B = C2[2] * (2.f * zz - xx - yy); Dx = C2[2] * 2.f * -x; Dy = C2[2] * 2.f * -y; Dz = C2[2] * 4.f * z; break;
        case 7: B = C2[3] * xz; Dx = C2[3] * z; Dz = C2[3] * x; break;
        case 8: B = C2[4] * (xx - yy); Dx = C2[4] * 2.f * x; Dy = C2[4] * 2.f * -y; break;
        case 9: B = C3[0] * y * (3.f * xx - yy); Dx = C3[0] * 6.f * xy; Dy = C3[0] * 3.f * (xx - yy); break;
        case 10: B = C3[1] * xy * z; Dx = C3[1] * yz; Dy = C3[1] * xz; Dz = C3[1] * xy; break;
        case 11: B = C3[2] * y * (4.f * zz - xx - yy); Dx = C3[2] * -2.f * xy; Dy = C3[2] * (-3.f * yy + 4.f * zz - xx);
                 Dz = C3[2] * 8.f * yz; break;
        case 12: B = C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy); Dx = C3[3] * -6.f * xz; Dy = C3[3] * -6.f * yz;
                 Dz = C3[3] * 3.f * (2.f * zz - xx - yy); break;
        case 13: B = C3[4] * x * (4.f * zz - xx - yy); Dx = C3[4] * (-3.f * xx + 4.f * zz - yy); Dy = C3[4] * -2.f * xy;
                 Dz = C3[4] * 8.f * xz; break;
        case 14: B = C3[5] * z * (xx - yy); Dx = C3[5] * 2.f * xz; Dy = C3[5] * -2.f * yz; Dz = C3[5] * (xx - yy); break;
        default: B = C3[6] * x * (xx - 3.f * yy); Dx = C3[6] * 3.f * (xx - yy); Dy = C3[6] * -6.f * xy; break;
    }
}

// SH backward of one Gaussian, in place on its staged row (12 floats = 4 coefficients = three 16-byte pieces at a
// time): row values in, dL/dsh out (zeros above the active degree), returns dL/d(direction).
__device__ __forceinline__ void sh_backward_row(int deg, int row_floats, float x, float y, float z, float4* r4,
                                                const float* g3, float* gdir) {
    const int used = (deg + 1) * (deg + 1);
    gdir[0] = gdir[1] = gdir[2] = 0.f;
#pragma unroll
    for (int grp = 0; grp < 4; ++grp) {
        if (12 * grp >= row_floats) break;
        float sv[12], o[12];
        if (4 * grp < used) {
            const float4 a = r4[3 * grp], b = r4[3 * grp + 1], c = r4[3 * grp + 2];
            sv[0] = a.x; sv[1] = a.y; sv[2] = a.z; sv[3] = a.w; sv[4] = b.x; sv[5] = b.y; sv[6] = b.z; sv[7] = b.w;
            sv[8] = c.x; sv[9] = c.y; sv[10] = c.z; sv[11] = c.w;
        }
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            const int k = 4 * grp + kk;
            if (k < used) {
                float B, Dx, Dy, Dz;
                sh_basis(k, x, y, z, B, Dx, Dy, Dz);
                const float tk = g3[0] * sv[3 * kk] + g3[1] * sv[3 * kk + 1] + g3[2] * sv[3 * kk + 2];
                gdir[0] += Dx * tk; gdir[1] += Dy * tk; gdir[2] += Dz * tk;
                o[3 * kk] = B * g3[0]; o[3 * kk + 1] = B * g3[1]; o[3 * kk + 2] = B * g3[2];
            } else {
                o[3 * kk] = o[3 * kk + 1] = o[3 * kk + 2] = 0.f;
            }
        }
        r4[3 * grp] = make_float4(o[0], o[1], o[2], o[3]);
        r4[3 * grp + 1] = make_float4(o[4], o[5], o[6], o[7]);
        r4[3 * grp + 2] = make_float4(o[8], o[9], o[10], o[11]);
    }
}

// The same, coefficient by coefficient, on a row whose coefficient 0 and coefficients 1.. live in two dense blocks
// (the raw-parameter layout: dc [.,1,3] and rest [.,M-1,3]; scalar accesses, odd strides: conflict-free).
__device__ __forceinline__ void sh_backward_split(int deg, int M, float x, float y, float z, float* dc, float* rest,
                                                  const float* g3, float* gdir) {
    const int used = (deg + 1) * (deg + 1);
    gdir[0] = gdir[1] = gdir[2] = 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        if (k >= M) break;
        float* c = k == 0 ? dc : rest + 3 * (k - 1);
        if (k < used) {
            float B, Dx, Dy, Dz;
            sh_basis(k, x, y, z, B, Dx, Dy, Dz);
            const float tk = g3[0] * c[0] + g3[1] * c[1] + g3[2] * c[2];
            gdir[0] += Dx * tk; gdir[1] += Dy * tk; gdir[2] += Dz * tk;
            c[0] = B * g3[0]; c[1] = B * g3[1]; c[2] = B * g3[2];
        } else {
            c[0] = c[1] = c[2] = 0.f;
        }
    }
}

template <bool kRaw>
__global__ void __launch_bounds__(kBwdThreads, 5) preprocess_backward_tma_kernel(const BwdParams p) {
    extern __shared__ __align__(16) unsigned char s_tma_raw[];
    BwdTmaSmem& sm = *reinterpret_cast<BwdTmaSmem*>(s_tma_raw);
    const BwdView& V = p.v[0];
    const int ntiles = (p.P + kBwdThreads - 1) / kBwdThreads;
    const int t = threadIdx.x;
    if (t < 16) sm.cam[t] = __ldg(V.view + t);
    else if (t < 32) sm.cam[t] = __ldg(V.proj + t - 16);
    else if (t < 35) sm.cam[t] = __ldg(V.campos + t - 32);
    if (t == 0) { mbar_init(&sm.bar_opnd, 1); mbar_init(&sm.bar_sh, 1); mbar_init_fence(); }
    __syncthreads();
    const float* view = sm.cam;
    const float* proj = sm.cam + 16;
    const float* campos = sm.cam + 32;
    const int row = 3 * p.M;                       // a multiple of 12 here
    const int need = 3 * (p.D + 1) * (p.D + 1);
    const uint32_t need_bytes = (uint32_t)((need * 4 + 15) & ~15);

    auto request_operands = [&](int tile) {
        const int base = tile * kBwdThreads;
        const int n = min(kBwdThreads, p.P - base);
        if (n == kBwdThreads) {
            if (t == 0) {
                mbar_arrive_expect_tx(&sm.bar_opnd, 120u * kBwdThreads);
                bulk_g2s(sm.xyz, p.means3D + 3 * (size_t)base, 12u * kBwdThreads, &sm.bar_opnd);
                bulk_g2s(sm.scale, p.scales + 3 * (size_t)base, 12u * kBwdThreads, &sm.bar_opnd);
                bulk_g2s(sm.rot, p.rotations + 4 * (size_t)base, 16u * kBwdThreads, &sm.bar_opnd);
                bulk_g2s(sm.acc, V.grad_accum + 12 * (size_t)base, 48u * kBwdThreads, &sm.bar_opnd);
                bulk_g2s(sm.con, V.conic_opacity + base, 16u * kBwdThreads, &sm.bar_opnd);
                bulk_g2s(sm.rgb, V.rgb + base, 16u * kBwdThreads, &sm.bar_opnd);
            }
        } else if (t < n) {
            const size_t i = (size_t)base + t;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                sm.xyz[3 * t + k] = __ldg(p.means3D + 3 * i + k);
                sm.scale[3 * t + k] = __ldg(p.scales + 3 * i + k);
                sm.acc[3 * t + k] = reinterpret_cast<const float4*>(V.grad_accum)[3 * i + k];
            }
            sm.rot[t] = __ldg(reinterpret_cast<const float4*>(p.rotations) + i);
            sm.con[t] = V.conic_opacity[i];
            sm.rgb[t] = V.rgb[i];
        }
    };
    auto radius_of = [&](int tile) {
        const int i = tile * kBwdThreads + t;
        return (tile < ntiles && i < p.P) ? V.radii[i] : 0;
    };

    uint32_t ph_opnd = 0, ph_sh = 0;
    int tile = blockIdx.x;
    if (tile < ntiles) request_operands(tile);
    int r_cur = radius_of(tile), r_next = radius_of(tile + (int)gridDim.x);   // radii travel one tile ahead in registers
    // state carried from a tile's phase A to its phase B one iteration later
    bool prev_any = false, prev_vis = false;
    int prev_idx = 0, prev_base = 0;
    float pm[3] = {0.f, 0.f, 0.f}, pvm[3] = {0.f, 0.f, 0.f}, pg3[3] = {0.f, 0.f, 0.f};

    for (; tile < ntiles || prev_any; tile += gridDim.x) {
        const bool cur = tile < ntiles;
        const int base = tile * kBwdThreads;
        const int idx = base + t;
        const bool valid = cur && idx < p.P;
        const bool vis = valid && r_cur > 0;
        float mx = 0.f, my = 0.f, mz = 0.f;
        float vm[3] = {0.f, 0.f, 0.f}, g3[3] = {0.f, 0.f, 0.f};
        if (cur) {
            if (p.P - base >= kBwdThreads) { mbar_wait(&sm.bar_opnd, ph_opnd); ph_opnd ^= 1u; }
            float g2d[2] = {0.f, 0.f}, gop = 0.f, gs[3] = {0.f, 0.f, 0.f}, gq[4] = {0.f, 0.f, 0.f, 0.f};
            if (vis) {
                mx = sm.xyz[3 * t]; my = sm.xyz[3 * t + 1]; mz = sm.xyz[3 * t + 2];
                float4 q = sm.rot[t];
                const float4 q_raw = q;
                float rs0 = sm.scale[3 * t], rs1 = sm.scale[3 * t + 1], rs2 = sm.scale[3 * t + 2];
                if (kRaw) {   // pre-activation inputs (scene/gaussian_model.py:28-45)
                    q = act_rotation(q);
                    rs0 = act_scale(rs0); rs1 = act_scale(rs1); rs2 = act_scale(rs2);
                }
                float s[3], c6[6], R[3][3];
                s[0] = p.scale_modifier * rs0;
                s[1] = p.scale_modifier * rs1;
                s[2] = p.scale_modifier * rs2;
                cov3d(s[0], s[1], s[2], q, R, c6);
                const float4 a0 = sm.acc[3 * t], a1 = sm.acc[3 * t + 1], a2 = sm.acc[3 * t + 2], con = sm.con[t];
                const uint32_t clamped_bits = __float_as_uint(sm.rgb[t].w);
                const float S1 = a0.x, S2 = a0.y;
                const float gA = -0.5f * a0.z, gB = -a0.w, gC = -0.5f * a1.x;
                g2d[0] = (-con.x * S1 - con.y * S2) * (0.5f * (float)p.W);
                g2d[1] = (-con.z * S2 - con.y * S1) * (0.5f * (float)p.H);
                gop = a1.y;
                const float vcol[3] = {a1.z, a1.w, a2.x};
#pragma unroll
                for (int ch = 0; ch < 3; ++ch) g3[ch] = (clamped_bits >> ch) & 1u ? 0.f : vcol[ch];

                // ---- EWA projection (forward values) ----
                float tt[3];
                tt[0] = view[0] * mx + view[4] * my + view[8] * mz + view[12];
                tt[1] = view[1] * mx + view[5] * my + view[9] * mz + view[13];
                tt[2] = view[2] * mx + view[6] * my + view[10] * mz + view[14];
                const float limx = 1.3f * V.tanfovx, limy = 1.3f * V.tanfovy;
                const float txtz = tt[0] / tt[2], tytz = tt[1] / tt[2];
                const float gmulx = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
                const float gmuly = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
                tt[0] = fminf(limx, fmaxf(-limx, txtz)) * tt[2];
                tt[1] = fminf(limy, fmaxf(-limy, tytz)) * tt[2];
                const float fx = V.focal_x, fy = V.focal_y;
                const float j00 = fx / tt[2], j11 = fy / tt[2];
                const float j02 = -(fx * tt[0]) / (tt[2] * tt[2]), j12 = -(fy * tt[1]) / (tt[2] * tt[2]);
                float M0[3], M1[3];
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    M0[k] = view[4 * k + 0] * j00 + view[4 * k + 2] * j02;
                    M1[k] = view[4 * k + 1] * j11 + view[4 * k + 2] * j12;
                }
                const float S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
                float SM0[3], SM1[3];
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    SM0[k] = S[k][0] * M0[0] + S[k][1] * M0[1] + S[k][2] * M0[2];
                    SM1[k] = S[k][0] * M1[0] + S[k][1] * M1[1] + S[k][2] * M1[2];
                }
                const float a = M0[0] * SM0[0] + M0[1] * SM0[1] + M0[2] * SM0[2] + 0.3f;
                const float b = M1[0] * SM0[0] + M1[1] * SM0[1] + M1[2] * SM0[2];
                const float c = M1[0] * SM1[0] + M1[1] * SM1[1] + M1[2] * SM1[2] + 0.3f;
                const float det = a * c - b * b;
                const float d2inv = 1.f / (det * det + 0.0000001f);
                float dLa = 0.f, dLb = 0.f, dLc = 0.f;
                float gc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                if (d2inv != 0.f) {
                    dLa = d2inv * (-c * c * gA + b * c * gB + (det - a * c) * gC);
                    dLc = d2inv * (-a * a * gC + a * b * gB + (det - a * c) * gA);
                    dLb = d2inv * (2.f * b * c * gA - (det + 2.f * b * b) * gB + 2.f * a * b * gC);
                    gc[0] = M0[0] * M0[0] * dLa + M0[0] * M1[0] * dLb + M1[0] * M1[0] * dLc;
                    gc[3] = M0[1] * M0[1] * dLa + M0[1] * M1[1] * dLb + M1[1] * M1[1] * dLc;
                    gc[5] = M0[2] * M0[2] * dLa + M0[2] * M1[2] * dLb + M1[2] * M1[2] * dLc;
                    gc[1] = 2.f * M0[0] * M0[1] * dLa + (M0[0] * M1[1] + M0[1] * M1[0]) * dLb + 2.f * M1[0] * M1[1] * dLc;
                    gc[2] = 2.f * M0[0] * M0[2] * dLa + (M0[0] * M1[2] + M0[2] * M1[0]) * dLb + 2.f * M1[0] * M1[2] * dLc;
                    gc[4] = 2.f * M0[2] * M0[1] * dLa + (M0[1] * M1[2] + M0[2] * M1[1]) * dLb + 2.f * M1[1] * M1[2] * dLc;
                }
                {
                    float dM0[3], dM1[3];
#pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        dM0[k] = 2.f * SM0[k] * dLa + SM1[k] * dLb;
                        dM1[k] = 2.f * SM1[k] * dLc + SM0[k] * dLb;
                    }
                    float dj00 = 0.f, dj02 = 0.f, dj11 = 0.f, dj12 = 0.f;
#pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        dj00 += view[4 * k + 0] * dM0[k];
                        dj02 += view[4 * k + 2] * dM0[k];
                        dj11 += view[4 * k + 1] * dM1[k];
                        dj12 += view[4 * k + 2] * dM1[k];
                    }
                    const float tz = 1.f / tt[2], tz2 = tz * tz, tz3 = tz2 * tz;
                    const float dtx = gmulx * -fx * tz2 * dj02;
                    const float dty = gmuly * -fy * tz2 * dj12;
                    const float dtz = -fx * tz2 * dj00 - fy * tz2 * dj11 + (2.f * fx * tt[0]) * tz3 * dj02 +
                                      (2.f * fy * tt[1]) * tz3 * dj12;
                    vm[0] = view[0] * dtx + view[1] * dty + view[2] * dtz;
                    vm[1] = view[4] * dtx + view[5] * dty + view[6] * dtz;
                    vm[2] = view[8] * dtx + view[9] * dty + view[10] * dtz;
                }
                {   // mean2D -> mean3D through the perspective divide
                    const float hx = proj[0] * mx + proj[4] * my + proj[8] * mz + proj[12];
                    const float hy = proj[1] * mx + proj[5] * my + proj[9] * mz + proj[13];
                    const float hw = proj[3] * mx + proj[7] * my + proj[11] * mz + proj[15];
                    const float mw = 1.f / (hw + 0.0000001f);
                    const float mul1 = hx * mw * mw, mul2 = hy * mw * mw;
                    vm[0] += (proj[0] * mw - proj[3] * mul1) * g2d[0] + (proj[1] * mw - proj[3] * mul2) * g2d[1];
                    vm[1] += (proj[4] * mw - proj[7] * mul1) * g2d[0] + (proj[5] * mw - proj[7] * mul2) * g2d[1];
                    vm[2] += (proj[8] * mw - proj[11] * mul1) * g2d[0] + (proj[9] * mw - proj[11] * mul2) * g2d[1];
                }
                {   // cov3D -> scale, quaternion
                    const float Gs[3][3] = {{gc[0], 0.5f * gc[1], 0.5f * gc[2]},
                                            {0.5f * gc[1], gc[3], 0.5f * gc[4]},
                                            {0.5f * gc[2], 0.5f * gc[4], gc[5]}};
                    float GR[3][3], dR[3][3];
#pragma unroll
                    for (int a2 = 0; a2 < 3; ++a2)
#pragma unroll
                        for (int k = 0; k < 3; ++k)
                            GR[a2][k] = Gs[a2][0] * R[0][k] + Gs[a2][1] * R[1][k] + Gs[a2][2] * R[2][k];
#pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        const float rgr = R[0][k] * GR[0][k] + R[1][k] * GR[1][k] + R[2][k] * GR[2][k];
                        gs[k] = 2.f * s[k] * rgr;
#pragma unroll
                        for (int a2 = 0; a2 < 3; ++a2) dR[a2][k] = 2.f * GR[a2][k] * s[k] * s[k];
                    }
                    const float r = q.x, x = q.y, y = q.z, z = q.w;
                    gq[0] = 2.f * (z * (dR[1][0] - dR[0][1]) + y * (dR[0][2] - dR[2][0]) + x * (dR[2][1] - dR[1][2]));
                    gq[1] = 2.f * (y * (dR[0][1] + dR[1][0]) + z * (dR[0][2] + dR[2][0]) + r * (dR[2][1] - dR[1][2])) -
                            4.f * x * (dR[1][1] + dR[2][2]);
                    gq[2] = 2.f * (x * (dR[0][1] + dR[1][0]) + r * (dR[0][2] - dR[2][0]) + z * (dR[1][2] + dR[2][1])) -
                            4.f * y * (dR[0][0] + dR[2][2]);
                    gq[3] = 2.f * (r * (dR[1][0] - dR[0][1]) + x * (dR[0][2] + dR[2][0]) + y * (dR[1][2] + dR[2][1])) -
                            4.f * z * (dR[0][0] + dR[1][1]);
                }
                if (kRaw) {   // chain rules of sigmoid / exp / normalize
                    const float o = con.w;
                    gop *= o * (1.0f - o);
                    gs[0] *= rs0; gs[1] *= rs1; gs[2] *= rs2;
                    const float n = quat_norm(q_raw);
                    const float dotp = q.x * gq[0] + q.y * gq[1] + q.z * gq[2] + q.w * gq[3];
                    const float inv = 1.0f / n;
                    gq[0] = (gq[0] - q.x * dotp) * inv; gq[1] = (gq[1] - q.y * dotp) * inv;
                    gq[2] = (gq[2] - q.z * dotp) * inv; gq[3] = (gq[3] - q.w * dotp) * inv;
                }
            }
            if (valid) {   // everything that does not need the SH row goes out now (zeros for culled rows)
                const size_t i = (size_t)idx;
                V.dL_dmeans2D[3 * i] = g2d[0]; V.dL_dmeans2D[3 * i + 1] = g2d[1]; V.dL_dmeans2D[3 * i + 2] = 0.f;
                p.dL_dopacity[i] = gop;
                p.dL_dscales[3 * i] = gs[0]; p.dL_dscales[3 * i + 1] = gs[1]; p.dL_dscales[3 * i + 2] = gs[2];
                reinterpret_cast<float4*>(p.dL_drots)[i] = make_float4(gq[0], gq[1], gq[2], gq[3]);
            }
        }
        __syncthreads();   // the operand block has been read by everyone
        if (cur && tile + (int)gridDim.x < ntiles) request_operands(tile + gridDim.x);

        // ---- phase B of the PREVIOUS tile: SH backward in place on the staged rows, rows out by bulk store ----
        if (prev_any) {
            mbar_wait(&sm.bar_sh, ph_sh);
            ph_sh ^= 1u;
            float4* r4 = sm.sh + kBwdShRow4 * t;
            // raw layout: the tile's dc block [128][3], then its rest block [128][3(M-1)], both dense
            float* dc_row = reinterpret_cast<float*>(sm.sh) + 3 * t;
            float* rest_row = reinterpret_cast<float*>(sm.sh) + 3 * kBwdThreads + (row - 3) * t;
            float gm[3] = {pvm[0], pvm[1], pvm[2]};
            if (prev_vis) {
                const float d0 = pm[0] - campos[0], d1 = pm[1] - campos[1], d2 = pm[2] - campos[2];
                const float s2 = d0 * d0 + d1 * d1 + d2 * d2;
                const float len = sqrtf(s2);
                float gdir[3];
                if (kRaw) sh_backward_split(p.D, p.M, d0 / len, d1 / len, d2 / len, dc_row, rest_row, pg3, gdir);
                else sh_backward_row(p.D, row, d0 / len, d1 / len, d2 / len, r4, pg3, gdir);
                const float inv32 = 1.f / sqrtf(s2 * s2 * s2);
                const float vd = d0 * gdir[0] + d1 * gdir[1] + d2 * gdir[2];
                gm[0] += (s2 * gdir[0] - d0 * vd) * inv32;
                gm[1] += (s2 * gdir[1] - d1 * vd) * inv32;
                gm[2] += (s2 * gdir[2] - d2 * vd) * inv32;
            } else if (kRaw) {
                dc_row[0] = dc_row[1] = dc_row[2] = 0.f;
                for (int j = 0; j < row - 3; ++j) rest_row[j] = 0.f;
            } else {
                for (int j = 0; j < (row >> 2); ++j) r4[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
            if (prev_idx < p.P) {
                const size_t i = (size_t)prev_idx;
                p.dL_dmeans3D[3 * i] = gm[0]; p.dL_dmeans3D[3 * i + 1] = gm[1]; p.dL_dmeans3D[3 * i + 2] = gm[2];
                if (!kRaw) {
                    fence_async_smem();   // my row, written through the generic proxy, is read by the copy engine next
                    bulk_s2g(p.dL_dsh + i * row, r4, 4u * (uint32_t)row);
                    bulk_commit();
                    bulk_wait_read();     // the staged row may be overwritten once the engine has read it
                }
            }
            if (kRaw) {   // the two gradient blocks of the whole tile leave together: dL/d(dc) and dL/d(rest)
                const int pbase = prev_base;
                const int n = min(kBwdThreads, p.P - pbase);
                fence_async_smem();
                __syncthreads();   // every row of the tile is written
                float* dcb = reinterpret_cast<float*>(sm.sh);
                if (n == kBwdThreads) {
                    if (t == 0) {
                        bulk_s2g(p.dL_dsh + 3 * (size_t)pbase, dcb, 12u * kBwdThreads);
                        if (row > 3)
                            bulk_s2g(p.dL_dsh_rest + (size_t)pbase * (row - 3), dcb + 3 * kBwdThreads,
                                     4u * (uint32_t)(row - 3) * kBwdThreads);
                        bulk_commit();
                        bulk_wait_read();
                    }
                } else {   // ragged last tile: plain stores
                    for (int j = t; j < 3 * n; j += kBwdThreads) p.dL_dsh[3 * (size_t)pbase + j] = dcb[j];
                    for (int j = t; j < (row - 3) * n; j += kBwdThreads)
                        p.dL_dsh_rest[(size_t)pbase * (row - 3) + j] = dcb[3 * kBwdThreads + j];
                }
            }
        }
        const int nvis = __syncthreads_count(vis);   // also: the SH tile is free
        if (cur) {
            if (kRaw) {   // two contiguous blocks of the tile's rows (dc | rest), whoever is visible
                const int n = min(kBwdThreads, p.P - base);
                float* dcb = reinterpret_cast<float*>(sm.sh);
                if (n == kBwdThreads) {
                    if (t == 0) {
                        const uint32_t b_dc = 12u * kBwdThreads, b_rest = 4u * (uint32_t)(row - 3) * kBwdThreads;
                        mbar_arrive_expect_tx(&sm.bar_sh, b_dc + b_rest);
                        bulk_g2s(dcb, p.shs + 3 * (size_t)base, b_dc, &sm.bar_sh);
                        if (b_rest) bulk_g2s(dcb + 3 * kBwdThreads, p.sh_rest + (size_t)base * (row - 3), b_rest, &sm.bar_sh);
                    }
                } else {   // ragged last tile: plain loads, then an empty phase
                    for (int j = t; j < 3 * n; j += kBwdThreads) dcb[j] = __ldg(p.shs + 3 * (size_t)base + j);
                    for (int j = t; j < (row - 3) * n; j += kBwdThreads)
                        dcb[3 * kBwdThreads + j] = __ldg(p.sh_rest + (size_t)base * (row - 3) + j);
                    __syncthreads();
                    if (t == 0) mbar_arrive_expect_tx(&sm.bar_sh, 0u);
                }
            } else {
                if (t == 0) mbar_arrive_expect_tx(&sm.bar_sh, (uint32_t)nvis * need_bytes);
                if (vis) bulk_g2s(sm.sh + kBwdShRow4 * t, p.shs + (size_t)idx * row, need_bytes, &sm.bar_sh);
            }
        }
        prev_base = base;
        prev_any = cur;
        prev_vis = vis;
        prev_idx = cur ? idx : p.P;
        pm[0] = mx; pm[1] = my; pm[2] = mz;
        pvm[0] = vm[0]; pvm[1] = vm[1]; pvm[2] = vm[2];
        pg3[0] = g3[0]; pg3[1] = g3[1]; pg3[2] = g3[2];
        r_cur = r_next;
        r_next = radius_of(tile + 2 * (int)gridDim.x);
    }
}

constexpr size_t bwd_smem_bytes(bool multi, bool ring = false) {
    return sizeof(float) * (36 + (size_t)(multi ? 2 : 1) * kBwdWarps * 32 * kShStride) +
           (ring ? kAccRingBytes + sizeof(uint64_t) * kAccRing : 0);
}

}  // namespace

int launch_preprocess_backward_views(int P, int D, int M, const Grid& g, const float* means3D, const float* shs,
                                     int colors_precomp, const float* scales, float scale_modifier,
                                     const float* rotations, const float* cov3D_precomp, int n_views,
                                     const BackwardViewArgs* views, float* dL_dcolors, float* dL_dopacity,
                                     float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscales,
                                     float* dL_drots, int accumulate, cudaStream_t s, int debug, const RawMode& rm) {
    if (n_views < 1 || n_views > kMaxViews) return GSR_ERR_INVALID_ARGUMENT;
    static unsigned long long attr_devices = 0;
    GSR_CUDA(once_per_device(&attr_devices, []() -> cudaError_t {
        const auto attr = cudaFuncAttributeMaxDynamicSharedMemorySize;
        cudaError_t e = cudaFuncSetAttribute(preprocess_backward_kernel<true, false>, attr, (int)bwd_smem_bytes(true));
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(preprocess_backward_kernel<true, true>, attr, (int)bwd_smem_bytes(true));
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(preprocess_backward_kernel<true, false, true>, attr, (int)bwd_smem_bytes(true, true));
        return e;
    }));
    BwdParams p;
    p.raw = rm.raw; p.sh_rest = rm.sh_rest; p.dL_dsh_rest = rm.dL_dsh_rest;
    p.accumulate = accumulate & GSR_BWD_ACCUMULATE;
    p.n_views = n_views;
    p.P = P; p.D = D; p.M = M; p.W = g.W; p.H = g.H;
    p.means3D = means3D; p.shs = shs; p.colors_precomp = colors_precomp; p.scales = scales;
    p.scale_modifier = scale_modifier; p.rotations = rotations; p.cov3D_precomp = cov3D_precomp;
    p.dL_dcolors = dL_dcolors; p.dL_dopacity = dL_dopacity; p.dL_dmeans3D = dL_dmeans3D; p.dL_dcov3D = dL_dcov3D;
    p.dL_dsh = dL_dsh; p.dL_dscales = dL_dscales; p.dL_drots = dL_drots;
    for (int v = 0; v < n_views; ++v) {
        const BackwardViewArgs& a = views[v];
        BwdView& o = p.v[v];
        o.view = a.view; o.proj = a.proj; o.campos = a.campos;
        o.tanfovx = a.tanfovx; o.tanfovy = a.tanfovy;
        o.focal_x = (float)g.W / (2.0f * a.tanfovx);
        o.focal_y = (float)g.H / (2.0f * a.tanfovy);
        o.radii = a.radii; o.grad_accum = a.gv.grad_accum; o.conic_opacity = a.gv.conic_opacity; o.rgb = a.gv.rgb;
        o.dL_dmeans2D = a.dL_dmeans2D;
        o.radii_local = a.radii_local; o.conic_local = a.conic_local;
    }
    StageScope scope(kStagePreprocessBwd, s);
    const unsigned blocks = (unsigned)((P + kBwdThreads - 1) / kBwdThreads);
    // single view, SH colours, scale + rotation inputs, 16-byte-multiple SH rows, every output requested: the
    // persistent bulk-copy kernel
    const bool raw_ok = rm.raw && (M == 1 || (rm.sh_rest && rm.dL_dsh_rest)) &&
                        ((reinterpret_cast<uintptr_t>(rm.sh_rest) | reinterpret_cast<uintptr_t>(rm.dL_dsh_rest)) & 15) == 0;
    if ((!rm.raw || raw_ok) && n_views == 1 && !(accumulate & GSR_BWD_ACCUMULATE) && shs && !colors_precomp &&
        scales && rotations && !cov3D_precomp && (rm.raw || M % 4 == 0) && dL_dsh && dL_dscales && dL_drots &&
        !dL_dcolors && !dL_dcov3D && !views[0].radii_local) {
        static unsigned long long tma_devices = 0;
        GSR_CUDA(once_per_device(&tma_devices, []() -> cudaError_t {
            const auto attr = cudaFuncAttributeMaxDynamicSharedMemorySize;
            cudaError_t e = cudaFuncSetAttribute(preprocess_backward_tma_kernel<false>, attr, (int)sizeof(BwdTmaSmem));
            if (e == cudaSuccess)
                e = cudaFuncSetAttribute(preprocess_backward_tma_kernel<true>, attr, (int)sizeof(BwdTmaSmem));
            return e;
        }));
        const unsigned grid = std::min(blocks, (unsigned)sm_count() * 5u);   // 5 CTAs per multiprocessor stay resident
        if (rm.raw) preprocess_backward_tma_kernel<true><<<grid, kBwdThreads, sizeof(BwdTmaSmem), s>>>(p);
        else preprocess_backward_tma_kernel<false><<<grid, kBwdThreads, sizeof(BwdTmaSmem), s>>>(p);
        GSR_AFTER_LAUNCH(s, debug);
        return GSR_OK;
    }
    if (rm.raw) {
        if (n_views == 1) preprocess_backward_kernel<false, true><<<blocks, kBwdThreads, bwd_smem_bytes(false), s>>>(p);
        else preprocess_backward_kernel<true, true><<<blocks, kBwdThreads, bwd_smem_bytes(true), s>>>(p);
    } else {
        if (n_views == 1) preprocess_backward_kernel<false, false><<<blocks, kBwdThreads, bwd_smem_bytes(false), s>>>(p);
        else if (accumulate & GSR_BWD_PEER_ACCUMULATORS)
            preprocess_backward_kernel<true, false, true><<<blocks, kBwdThreads, bwd_smem_bytes(true, true), s>>>(p);
        else preprocess_backward_kernel<true, false><<<blocks, kBwdThreads, bwd_smem_bytes(true), s>>>(p);
    }
    GSR_AFTER_LAUNCH(s, debug);
    return GSR_OK;
}

int launch_preprocess_backward(int P, int D, int M, const Grid& g, const float* means3D, const float* shs,
                               int colors_precomp, const float* scales, float scale_modifier,
                               const float* rotations, const float* cov3D_precomp, const float* view,
                               const float* proj, const float* campos, float tanfovx, float tanfovy,
                               const int32_t* radii, GeomView gv, float* dL_dmeans2D, float* dL_dcolors,
                               float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh,
                               float* dL_dscales, float* dL_drots, int accumulate, cudaStream_t s, int debug,
                               const RawMode& rm) {
    BackwardViewArgs a;
    a.view = view; a.proj = proj; a.campos = campos; a.tanfovx = tanfovx; a.tanfovy = tanfovy;
    a.radii = radii; a.gv = gv; a.dL_dmeans2D = dL_dmeans2D;
    return launch_preprocess_backward_views(P, D, M, g, means3D, shs, colors_precomp, scales, scale_modifier, rotations,
                                            cov3D_precomp, 1, &a, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D,
                                            dL_dsh, dL_dscales, dL_drots, accumulate, s, debug, rm);
}

}  // namespace gsr
