// gsr_blend.cu — per-16x16-tile alpha compositing, forward and backward, for sm_100a
// (the reference's two `renderCUDA` kernels).
//
// One CTA per tile, 256 threads, one pixel per thread; warp w owns an 8x4 pixel block so that the
// lanes of a warp are spatially compact.  Splats are staged through shared memory in batches of
// 256 list entries.  B200-first differences from the reference algorithm, none of which changes
// a result:
//   * staging applies a CONSERVATIVE exact-ellipse-vs-tile test (minimum of the conic's quadratic
//     form over the tile rectangle against ln(255*opacity)) and compacts the batch in order with a
//     ballot, so entries that cannot reach alpha >= 1/255 anywhere in the tile are never evaluated
//     per pixel.  Such entries are `continue`d by every pixel in the reference, so dropping them
//     is exact; the original list position travels with each entry so n_contrib keeps its meaning.
//   * the conic is pre-scaled by -log2(e)/2 at staging, so the per-pair exponent is 5 FP32 ops and
//     one MUFU.EX2 (ex2.approx), not an expf() call.
//   * a warp leaves the inner loop as soon as all 32 of its pixels are saturated (vote), the CTA
//     leaves the batch loop when all 256 are.
//   * backward: per-pair gradients are reduced across the warp with a 12-shuffle transposing
//     butterfly (9 sums land in 9 different lanes), per-warp partials are combined through shared
//     memory WITHOUT shared float atomics (those are CAS loops on sm_100), and each (tile, splat)
//     issues three 128-bit vector reductions (REDG.E.ADD.F32x4) to L2 — versus 9 scalar global
//     atomics per contributing (pixel, splat) pair in the reference.
#include "gsr_internal.cuh"

namespace gsr {

namespace {

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kAlphaMin = 1.0f / 255.0f;
constexpr float kTmin = 0.0001f;
constexpr int kBlendThreads = 256;
constexpr int kBlendWarps = 8;

// Can the splat reach alpha >= 1/255 at some pixel centre of [rx0,rx1]x[ry0,ry1]?  Conservative:
// false only when certainly not.  q(d) = (A dx^2 + C dy^2)/2 + B dx dy is convex; when the centre
// is outside the rectangle its minimum over the rectangle lies on an edge facing the centre.
__device__ __forceinline__ bool splat_may_touch(float x, float y, float A, float B, float C, float op,
                                                float rx0, float ry0, float rx1, float ry1) {
    const float o255 = op * 255.0f;
    if (o255 < 1.0f) return false;  // alpha <= opacity < 1/255 everywhere
    const bool in_x = (x >= rx0) && (x <= rx1), in_y = (y >= ry0) && (y <= ry1);
    if (in_x && in_y) return true;
    if (!(A > 0.0f) || !(C > 0.0f)) return true;  // not a proper conic: leave it to the pixels
    const float tau = __logf(o255);
    const float dxe = fminf(fmaxf(x, rx0), rx1) - x, dye = fminf(fmaxf(y, ry0), ry1) - y;
    float qmin = 3.0e38f;
    if (!in_x) {
        const float py = fminf(fmaxf(y - __fdividef(B * dxe, C), ry0), ry1);
        const float dy = py - y;
        qmin = 0.5f * (A * dxe * dxe + C * dy * dy) + B * dxe * dy;
    }
    if (!in_y) {
        const float px = fminf(fmaxf(x - __fdividef(B * dye, A), rx0), rx1);
        const float dx = px - x;
        qmin = fminf(qmin, 0.5f * (A * dx * dx + C * dye * dye) + B * dx * dye);
    }
    const float dxm = fmaxf(fabsf(rx0 - x), fabsf(rx1 - x)), dym = fmaxf(fabsf(ry0 - y), fabsf(ry1 - y));
    const float smax = 0.5f * (A * dxm * dxm + C * dym * dym) + fabsf(B) * dxm * dym;
    return !(qmin > tau + 0.02f + 3.0e-6f * smax);  // NaN -> keep
}

struct TileCtx {
    int tile, tx, ty;
    int px, py;  // this thread's pixel
    bool inside;
    float rx0, ry0, rx1, ry1;
};

__device__ __forceinline__ TileCtx tile_ctx(int gx, int W, int H) {
    TileCtx c;
    c.tile = blockIdx.x;
    c.ty = c.tile / gx;
    c.tx = c.tile - c.ty * gx;
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    c.px = c.tx * kTile + (w & 1) * 8 + (l & 7);
    c.py = c.ty * kTile + (w >> 1) * 4 + (l >> 3);
    c.inside = c.px < W && c.py < H;
    c.rx0 = (float)(c.tx * kTile);
    c.ry0 = (float)(c.ty * kTile);
    c.rx1 = (float)min(c.tx * kTile + kTile - 1, W - 1);
    c.ry1 = (float)min(c.ty * kTile + kTile - 1, H - 1);
    return c;
}

// Ordered CTA-wide compaction slot for `keep`; returns slot (valid if keep) and the total.
__device__ __forceinline__ int compact_slot(bool keep, int* s_wcount, int& total) {
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    const int w = threadIdx.x >> 5;
    if (lane_id() == 0) s_wcount[w] = __popc(bal);
    __syncthreads();
    int off = 0;
    total = 0;
#pragma unroll
    for (int i = 0; i < kBlendWarps; ++i) {
        const int c = s_wcount[i];
        off += (i < w) ? c : 0;
        total += c;
    }
    return off + __popc(bal & ((1u << lane_id()) - 1u));
}

// ------------------------------------ forward ----------------------------------------------
// kCount = true is the INSTRUMENTED variant (gsr_blend_pair_counts): same decisions, no image
// outputs, counts the (pixel, splat) pairs the reference algorithm evaluates and blends.
template <bool kCount>
__global__ void __launch_bounds__(kBlendThreads) blend_forward_kernel(
    int gx, int W, int H, const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
    const float2* __restrict__ means2D, const float4* __restrict__ conic_opacity, const float4* __restrict__ rgb,
    const float* __restrict__ bg, float* __restrict__ out_color, float* __restrict__ final_T,
    uint32_t* __restrict__ n_contrib, unsigned long long* __restrict__ counters) {
    __shared__ float4 s_g[kBlendThreads];    // x, y, a2, b2
    __shared__ float2 s_c[kBlendThreads];    // c2, opacity
    __shared__ float4 s_col[kBlendThreads];  // r, g, b, list position (1-based, int bits)
    __shared__ int s_wcount[kBlendWarps];

    const TileCtx c = tile_ctx(gx, W, H);
    const uint2 range = ranges[c.tile];
    const int n = (int)(range.y - range.x);
    const float fx = (float)c.px, fy = (float)c.py;

    float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f;
    uint32_t last = 0;
    bool done = !c.inside;
    uint32_t n_blended = 0, stop_pos = 0;  // kCount only

    for (int base = 0; base < n; base += kBlendThreads) {
        if (__syncthreads_count(done) == kBlendThreads) break;
        const int p = base + threadIdx.x;
        bool keep = false;
        uint32_t gid = 0;
        float2 m = make_float2(0.f, 0.f);
        float4 co = make_float4(0.f, 0.f, 0.f, 0.f);
        if (p < n) {
            gid = __ldg(point_list + range.x + p);
            m = __ldg(means2D + gid);
            co = __ldg(conic_opacity + gid);
            keep = splat_may_touch(m.x, m.y, co.x, co.y, co.z, co.w, c.rx0, c.ry0, c.rx1, c.ry1);
        }
        int total;
        const int slot = compact_slot(keep, s_wcount, total);
        if (keep) {
            const float4 col = __ldg(rgb + gid);
            s_g[slot] = make_float4(m.x, m.y, -0.5f * kLog2e * co.x, -kLog2e * co.y);
            s_c[slot] = make_float2(-0.5f * kLog2e * co.z, co.w);
            s_col[slot] = make_float4(col.x, col.y, col.z, __int_as_float(p + 1));
        }
        __syncthreads();

        for (int j = 0; j < total; ++j) {
            if (__all_sync(0xffffffffu, done)) break;
            const float4 g = s_g[j];
            const float2 cc = s_c[j];
            const float dx = g.x - fx, dy = g.y - fy;
            const float pw = fmaf(cc.x * dy, dy, fmaf(g.w, dy, g.z * dx) * dx);
            const float alpha = fminf(0.99f, cc.y * ex2_approx(pw));
            if (done || pw > 0.0f || alpha < kAlphaMin) continue;
            const float test_T = T * (1.0f - alpha);
            if (test_T < kTmin) {
                done = true;
                if (kCount) stop_pos = (uint32_t)__float_as_int(s_col[j].w);
                continue;
            }
            if (kCount) ++n_blended;
            const float4 col = s_col[j];
            const float wgt = alpha * T;
            C0 = fmaf(col.x, wgt, C0);
            C1 = fmaf(col.y, wgt, C1);
            C2 = fmaf(col.z, wgt, C2);
            T = test_T;
            last = (uint32_t)__float_as_int(col.w);
        }
    }

    if (kCount) {
        // the reference walks its tile list until the pixel is done: stop position, else all n
        unsigned long long ev = c.inside ? (stop_pos ? stop_pos : (uint32_t)n) : 0u;
        unsigned long long bl = n_blended;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            ev += __shfl_xor_sync(0xffffffffu, ev, d);
            bl += __shfl_xor_sync(0xffffffffu, bl, d);
        }
        if (lane_id() == 0) {
            atomicAdd(counters, ev);
            atomicAdd(counters + 1, bl);
        }
        return;
    }
    if (c.inside) {
        const size_t pix = (size_t)c.py * W + c.px;
        const size_t plane = (size_t)W * H;
        final_T[pix] = T;
        n_contrib[pix] = last;
        out_color[pix] = fmaf(T, __ldg(bg + 0), C0);
        out_color[plane + pix] = fmaf(T, __ldg(bg + 1), C1);
        out_color[2 * plane + pix] = fmaf(T, __ldg(bg + 2), C2);
    }
}

// ------------------------------------ backward ---------------------------------------------
// Transposing butterfly: reduces 9 per-lane values over the warp in 12 shuffles.  On return,
// lanes with an even lane id and slot >= 0 hold the warp total of value `slot` (slots 9..11 are
// exact zeros used as float4 padding).
__device__ __forceinline__ int reduce_slot_of_lane(unsigned lane) {
    const int b4 = (lane >> 4) & 1, k = (lane >> 1) & 7;  // k = b3 b2 b1
    // within a 5-group: 000->0 001->1 010->2 011->pad 100->3 101->4 110->pad 111->pad
    // lower half-warp: values 0..4, pads become the zero slots 9,10,11; upper: values 5..8.
    const uint32_t packed = b4 ? 0xFFF8F765u : 0xBA439210u;
    const int nib = (int)((packed >> (4 * k)) & 15u);
    return nib == 15 ? -1 : nib;
}

__device__ __forceinline__ float warp_reduce9(const float* v, unsigned lane) {
    const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4, h2 = lane & 2;
    float u[5], t[3], s[2];
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        const float hi_v = (i < 4) ? v[5 + i] : 0.0f;
        const float send = h16 ? v[i] : hi_v;
        const float keep = h16 ? hi_v : v[i];
        u[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const float hi_v = (i < 2) ? u[3 + i] : 0.0f;
        const float send = h8 ? u[i] : hi_v;
        const float keep = h8 ? hi_v : u[i];
        t[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const float hi_v = (i < 1) ? t[2] : 0.0f;
        const float send = h4 ? t[i] : hi_v;
        const float keep = h4 ? hi_v : t[i];
        s[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    float r;
    {
        const float send = h2 ? s[0] : s[1];
        const float keep = h2 ? s[1] : s[0];
        r = keep + __shfl_xor_sync(0xffffffffu, send, 2);
    }
    r += __shfl_xor_sync(0xffffffffu, r, 1);
    return r;
}

constexpr int kSub = 32;  // splats per flush group

__global__ void __launch_bounds__(kBlendThreads) blend_backward_kernel(
    int gx, int W, int H, const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
    const float2* __restrict__ means2D, const float4* __restrict__ conic_opacity, const float4* __restrict__ rgb,
    const float* __restrict__ bg, const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib,
    const float* __restrict__ dL_dpix, float* __restrict__ grad_accum) {
    __shared__ float4 s_g[kBlendThreads];    // x, y, a2, b2
    __shared__ float4 s_c[kBlendThreads];    // c2, opacity, list position (1-based), gaussian id
    __shared__ float4 s_col[kBlendThreads];  // r, g, b, -
    __shared__ __align__(16) float s_part[kBlendWarps][kSub][12];
    __shared__ int s_wcount[kBlendWarps];
    __shared__ uint32_t s_wmax[kBlendWarps];

    const TileCtx c = tile_ctx(gx, W, H);
    const uint2 range = ranges[c.tile];
    const unsigned lane = lane_id();
    const int warp = threadIdx.x >> 5;
    const float fx = (float)c.px, fy = (float)c.py;
    const int my_slot = (lane & 1) ? -1 : reduce_slot_of_lane(lane);

    float Tfin = 0.f, g0 = 0.f, g1 = 0.f, g2 = 0.f;
    uint32_t last = 0;
    if (c.inside) {
        const size_t pix = (size_t)c.py * W + c.px;
        const size_t plane = (size_t)W * H;
        Tfin = final_T[pix];
        last = n_contrib[pix];
        g0 = __ldg(dL_dpix + pix);
        g1 = __ldg(dL_dpix + plane + pix);
        g2 = __ldg(dL_dpix + 2 * plane + pix);
    }
    const float bgdot = __ldg(bg) * g0 + __ldg(bg + 1) * g1 + __ldg(bg + 2) * g2;

    const uint32_t wmax = __reduce_max_sync(0xffffffffu, last);
    if (lane == 0) s_wmax[warp] = wmax;
    __syncthreads();
    uint32_t m = 0;
#pragma unroll
    for (int i = 0; i < kBlendWarps; ++i) m = max(m, s_wmax[i]);

    float T = Tfin, last_alpha = 0.f;
    float lc0 = 0.f, lc1 = 0.f, lc2 = 0.f, bh0 = 0.f, bh1 = 0.f, bh2 = 0.f;

    for (int hi = (int)m; hi > 0; hi -= kBlendThreads) {
        __syncthreads();
        const int p = hi - 1 - (int)threadIdx.x;  // descending list position
        bool keep = false;
        uint32_t gid = 0;
        float2 mm = make_float2(0.f, 0.f);
        float4 co = make_float4(0.f, 0.f, 0.f, 0.f);
        if (p >= 0) {
            gid = __ldg(point_list + range.x + p);
            mm = __ldg(means2D + gid);
            co = __ldg(conic_opacity + gid);
            keep = splat_may_touch(mm.x, mm.y, co.x, co.y, co.z, co.w, c.rx0, c.ry0, c.rx1, c.ry1);
        }
        int total;
        const int slot = compact_slot(keep, s_wcount, total);
        if (keep) {
            const float4 col = __ldg(rgb + gid);
            s_g[slot] = make_float4(mm.x, mm.y, -0.5f * kLog2e * co.x, -kLog2e * co.y);
            s_c[slot] = make_float4(-0.5f * kLog2e * co.z, co.w, __int_as_float(p + 1), __uint_as_float(gid));
            s_col[slot] = col;
        }
        __syncthreads();

        for (int sb = 0; sb < total; sb += kSub) {
            const int nsub = min(kSub, total - sb);
            for (int jj = 0; jj < nsub; ++jj) {
                const int j = sb + jj;
                const float4 g = s_g[j];
                const float4 cc = s_c[j];
                const float dx = g.x - fx, dy = g.y - fy;
                const float pw = fmaf(cc.x * dy, dy, fmaf(g.w, dy, g.z * dx) * dx);
                const float G = ex2_approx(pw);
                const float alpha = fminf(0.99f, cc.y * G);
                const bool ok = ((uint32_t)__float_as_int(cc.z) <= last) && !(pw > 0.0f) && !(alpha < kAlphaMin);
                float v[9];
#pragma unroll
                for (int k = 0; k < 9; ++k) v[k] = 0.0f;
                if (ok) {
                    const float4 col = s_col[j];
                    const float ra = rcp_approx(1.0f - alpha);
                    T *= ra;
                    const float wgt = alpha * T;
                    bh0 = last_alpha * lc0 + (1.0f - last_alpha) * bh0;
                    bh1 = last_alpha * lc1 + (1.0f - last_alpha) * bh1;
                    bh2 = last_alpha * lc2 + (1.0f - last_alpha) * bh2;
                    lc0 = col.x; lc1 = col.y; lc2 = col.z;
                    float dL_dalpha = (col.x - bh0) * g0 + (col.y - bh1) * g1 + (col.z - bh2) * g2;
                    dL_dalpha *= T;
                    last_alpha = alpha;
                    dL_dalpha -= Tfin * ra * bgdot;
                    const float h = cc.y * dL_dalpha * G;  // dL/dG * G
                    const float hx = h * dx, hy = h * dy;
                    v[0] = hx; v[1] = hy; v[2] = hx * dx; v[3] = hx * dy; v[4] = hy * dy;
                    v[5] = G * dL_dalpha;
                    v[6] = wgt * g0; v[7] = wgt * g1; v[8] = wgt * g2;
                }
                if (__any_sync(0xffffffffu, ok)) {
                    const float r = warp_reduce9(v, lane);
                    if (my_slot >= 0) s_part[warp][jj][my_slot] = r;
                } else if (lane < 12) {
                    s_part[warp][jj][lane] = 0.0f;
                }
            }
            __syncthreads();
            // flush: one thread per (splat, float4 quad)
            if ((int)threadIdx.x < nsub * 3) {
                const int jj = threadIdx.x / 3, quad = threadIdx.x - jj * 3;
                float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int w = 0; w < kBlendWarps; ++w) {
                    const float4 pv = *reinterpret_cast<const float4*>(&s_part[w][jj][quad * 4]);
                    acc.x += pv.x; acc.y += pv.y; acc.z += pv.z; acc.w += pv.w;
                }
                if (acc.x != 0.f || acc.y != 0.f || acc.z != 0.f || acc.w != 0.f) {
                    const uint32_t gid2 = __float_as_uint(s_c[sb + jj].w);
                    atomicAdd(reinterpret_cast<float4*>(grad_accum + (size_t)gid2 * 12 + quad * 4), acc);
                }
            }
            __syncthreads();
        }
    }
}

}  // namespace

int launch_blend_forward(const Grid& g, const float* bg, GeomView gv, BinView bv, ImageView iv,
                         float* out_color, cudaStream_t s, int debug) {
    StageScope scope(kStageBlendFwd, s);
    blend_forward_kernel<false><<<g.tiles, kBlendThreads, 0, s>>>(g.gx, g.W, g.H, iv.ranges, bv.point_list,
                                                                  gv.means2D, gv.conic_opacity, gv.rgb, bg, out_color,
                                                                  iv.final_T, iv.n_contrib, nullptr);
    GSR_AFTER_LAUNCH(s, debug);
    return GSR_OK;
}

int launch_blend_pair_counts(const Grid& g, GeomView gv, BinView bv, ImageView iv, unsigned long long* counters,
                             cudaStream_t s) {
    GSR_CUDA(cudaMemsetAsync(counters, 0, 2 * sizeof(unsigned long long), s));
    blend_forward_kernel<true><<<g.tiles, kBlendThreads, 0, s>>>(g.gx, g.W, g.H, iv.ranges, bv.point_list,
                                                                 gv.means2D, gv.conic_opacity, gv.rgb, nullptr, nullptr,
                                                                 nullptr, nullptr, counters);
    GSR_AFTER_LAUNCH(s, 0);
    return GSR_OK;
}

int launch_blend_backward(int P, const Grid& g, const float* bg, GeomView gv, BinView bv, ImageView iv,
                          const float* dL_dpix, cudaStream_t s, int debug) {
    StageScope scope(kStageBlendBwd, s);
    GSR_CUDA(cudaMemsetAsync(gv.grad_accum, 0, sizeof(float) * 12 * (size_t)P, s));
    blend_backward_kernel<<<g.tiles, kBlendThreads, 0, s>>>(g.gx, g.W, g.H, iv.ranges, bv.point_list, gv.means2D,
                                                            gv.conic_opacity, gv.rgb, bg, iv.final_T, iv.n_contrib,
                                                            dL_dpix, gv.grad_accum);
    GSR_AFTER_LAUNCH(s, debug);
    return GSR_OK;
}

}  // namespace gsr
