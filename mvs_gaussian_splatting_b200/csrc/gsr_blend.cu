// gsr_blend.cu — per-16x16-tile alpha compositing, forward and backward, for sm_100a
// (the reference's two `renderCUDA` kernels).
//
// One CTA per tile.  Forward: 256 threads, one pixel per thread, warp w owns an 8x4 pixel block so that the
// lanes of a warp are spatially compact; splats are staged through shared memory in batches of 256 list
// entries.  Backward (blend_backward2_kernel, the default): 128 threads, TWO pixels per lane, warp w owns an
// 8x8 block, batches of 128 — the per-(warp, splat) reduction is paid once per 64 pixels (0.750 -> 0.604 ms);
// the round-1 one-pixel-per-lane backward is gone (A/B builds go through GSRAST_B200_LIB, never an environment
// switch inside the library).  B200-first differences from the reference algorithm, none of which changes a result:
//   * two-level exact culling.  Staging applies a CONSERVATIVE ellipse-vs-rectangle test (maximum of
//     the conic's exponent over the rectangle against -log2(255*opacity)) for the TILE and compacts
//     the batch in order with a ballot; then every WARP tests the compacted entries, 32 at a time
//     (one per lane), against its own 8x4 pixel block and walks only the set bits of the ballot.
//     Entries that cannot reach alpha >= 1/255 anywhere in the block are `continue`d by every one of
//     its pixels in the reference, so dropping them is exact; the original list position travels
//     with each entry so n_contrib keeps its meaning.  Small splats touch 1-3 of the 8 blocks of a
//     tile, so most (warp, splat) iterations of the reference never issue.
//   * the conic is pre-scaled by -log2(e)/2 at staging, so the per-pair exponent is 7 FP32 ops and
//     one MUFU.EX2 (ex2.approx), not an expf() call.
//   * a warp leaves the loop as soon as all 32 of its pixels are saturated (vote), the CTA leaves
//     the batch loop when all 256 are.
//   * backward: the 9 per-pair gradient terms of a warp's 32 pixels are TRANSPOSED through a per-warp
//     shared tile (9 conflict-free STS per splat) and summed three splats at a time by 27 lanes that
//     each add up one 32-entry row (8 LDS.128 + 31 FADD) — a third of the issue slots of a shuffle
//     butterfly — then flushed as 128-bit vector reductions (RED.E.ADD.F32x4) to L2: three per
//     (8x4 block, splat) — versus 9 scalar global atomics per contributing (pixel, splat) pair in
//     the reference.  No CTA-wide barrier inside the walk.
#include "gsr_internal.cuh"

namespace gsr {

namespace {

// Staging gathers (list entry -> mean + conic) of the NEXT batch issued before the walk of the current one.  A/B on
// the B200 (config 3): backward 0.605 -> 0.592 ms at 77 registers / 6 CTAs per SM (0.620 at 5 CTAs); forward
// 0.313 -> 0.347 ms (the seven extra registers cost the fifth CTA per SM) — so backward only.
#ifndef GSR_BLEND_PREFETCH_FWD
#define GSR_BLEND_PREFETCH_FWD 0
#endif
#ifndef GSR_BLEND_PREFETCH_BWD
#define GSR_BLEND_PREFETCH_BWD 1
#endif
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kAlphaMin = 1.0f / 255.0f;
constexpr float kTmin = 0.0001f;
constexpr int kBlendThreads = 256;
constexpr int kBlendWarps = 8;

struct Rect { float x0, y0, x1, y1; };

// Can the splat reach alpha >= 1/255 at some pixel centre of the rectangle?  Conservative: false
// only when certainly not.  Works on the pre-scaled conic: e(d) = a2 dx^2 + b2 dx dy + c2 dy^2 is
// the base-2 exponent (concave, <= 0), alpha = op * 2^e, so the question is whether max e over the
// rectangle reaches -tau2 with tau2 = log2(255 * op).  When the centre is outside the rectangle the
// maximum lies on an edge facing the centre.
__device__ __forceinline__ bool splat_may_reach(float x, float y, float a2, float b2, float c2, float tau2,
                                                const Rect& r) {
    const bool in_x = (x >= r.x0) && (x <= r.x1), in_y = (y >= r.y0) && (y <= r.y1);
    if (in_x && in_y) return true;
    if (!(a2 < 0.0f) || !(c2 < 0.0f)) return true;  // not a proper conic: leave it to the pixels
    const float dxe = fminf(fmaxf(x, r.x0), r.x1) - x, dye = fminf(fmaxf(y, r.y0), r.y1) - y;
    float emax = -3.0e38f;
    if (!in_x) {
        const float py = fminf(fmaxf(y - __fdividef(0.5f * b2 * dxe, c2), r.y0), r.y1);
        const float dy = py - y;
        emax = a2 * dxe * dxe + c2 * dy * dy + b2 * dxe * dy;
    }
    if (!in_y) {
        const float px = fminf(fmaxf(x - __fdividef(0.5f * b2 * dye, a2), r.x0), r.x1);
        const float dx = px - x;
        emax = fmaxf(emax, a2 * dx * dx + c2 * dye * dye + b2 * dx * dye);
    }
    const float dxm = fmaxf(fabsf(r.x0 - x), fabsf(r.x1 - x)), dym = fmaxf(fabsf(r.y0 - y), fabsf(r.y1 - y));
    const float smax = fabsf(b2) * dxm * dym - a2 * dxm * dxm - c2 * dym * dym;
    return !(emax < -(tau2 + 0.03f + 3.0e-6f * smax));  // NaN -> keep
}

struct TileCtx {
    int tile, tx, ty;
    int px, py;  // this thread's pixel
    bool inside;
    Rect tile_rect;  // pixel centres of the tile, clipped to the image
    Rect warp_rect;  // pixel centres of this warp's 8x4 block, clipped to the image
    bool warp_inside;  // the block has at least one pixel inside the image
};

__device__ __forceinline__ TileCtx tile_ctx(int gx, int W, int H) {
    TileCtx c;
    c.tile = blockIdx.x;
    c.ty = c.tile / gx;
    c.tx = c.tile - c.ty * gx;
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    const int bx = c.tx * kTile + (w & 1) * 8, by = c.ty * kTile + (w >> 1) * 4;
    c.px = bx + (l & 7);
    c.py = by + (l >> 3);
    c.inside = c.px < W && c.py < H;
    c.tile_rect.x0 = (float)(c.tx * kTile);
    c.tile_rect.y0 = (float)(c.ty * kTile);
    c.tile_rect.x1 = (float)min(c.tx * kTile + kTile - 1, W - 1);
    c.tile_rect.y1 = (float)min(c.ty * kTile + kTile - 1, H - 1);
    c.warp_inside = bx < W && by < H;
    c.warp_rect.x0 = (float)bx;
    c.warp_rect.y0 = (float)by;
    c.warp_rect.x1 = (float)min(bx + 7, W - 1);
    c.warp_rect.y1 = (float)min(by + 3, H - 1);
    return c;
}

// Ordered CTA-wide compaction slot for `keep`; returns slot (valid if keep) and the total.
template <int kWarps = kBlendWarps>
__device__ __forceinline__ int compact_slot(bool keep, int* s_wcount, int& total) {
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    const int w = threadIdx.x >> 5;
    if (lane_id() == 0) s_wcount[w] = __popc(bal);
    __syncthreads();
    int off = 0;
    total = 0;
#pragma unroll
    for (int i = 0; i < kWarps; ++i) {
        const int c = s_wcount[i];
        off += (i < w) ? c : 0;
        total += c;
    }
    return off + __popc(bal & ((1u << lane_id()) - 1u));
}

// Tile-level staging test of one list entry; on keep, returns the pre-scaled conic and tau2.
__device__ __forceinline__ bool stage_test(float2 m, float4 co, const Rect& tile_rect, float& a2, float& b2,
                                           float& c2, float& tau2) {
    const float o255 = co.w * 255.0f;
    a2 = -0.5f * kLog2e * co.x;
    b2 = -kLog2e * co.y;
    c2 = -0.5f * kLog2e * co.z;
    tau2 = __log2f(o255);
    if (o255 < 1.0f) return false;  // alpha <= opacity < 1/255 everywhere
    return splat_may_reach(m.x, m.y, a2, b2, c2, tau2, tile_rect);
}

// ------------------------------------ forward ----------------------------------------------
// kCount = true is the INSTRUMENTED variant (gsr_blend_pair_counts): same decisions, no image
// outputs, counts the (pixel, splat) pairs the reference algorithm evaluates and blends, plus the
// pairs this kernel actually evaluates after both culling levels.
template <bool kCount>
__global__ void __launch_bounds__(kBlendThreads) blend_forward_kernel(
    int gx, int W, int H, const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
    const float2* __restrict__ means2D, const float4* __restrict__ conic_opacity, const float4* __restrict__ rgb,
    const float* __restrict__ bg, float* __restrict__ out_color, float* __restrict__ final_T,
    uint32_t* __restrict__ n_contrib, unsigned long long* __restrict__ counters) {
    // staged entry j = s_e[3j..3j+2]: (x, y, a2, b2) (c2, opacity, tau2, list position 1-based) (r, g, b, -)
    // one base address per entry; 48-byte stride keeps the lane-per-entry 16-byte reads conflict-free
    __shared__ float4 s_e[3 * kBlendThreads];
    __shared__ int s_wcount[kBlendWarps];

    const TileCtx c = tile_ctx(gx, W, H);
    const uint2 range = ranges[c.tile];
    const int n = (int)(range.y - range.x);
    const float fx = (float)c.px, fy = (float)c.py;
    const unsigned lane = lane_id();

    float T = 1.0f, C0 = 0.f, C1 = 0.f, C2 = 0.f;
    uint32_t last = 0;
    bool done = !c.inside;
    uint32_t n_blended = 0, stop_pos = 0, n_eval = 0, n_iter = 0;  // kCount only

#if GSR_BLEND_PREFETCH_FWD
    uint32_t n_gid = 0;
    float2 n_m = make_float2(0.f, 0.f);
    float4 n_co = make_float4(0.f, 0.f, 0.f, 0.f);
    if ((int)threadIdx.x < n) {
        n_gid = __ldg(point_list + range.x + threadIdx.x);
        n_m = __ldg(means2D + n_gid);
        n_co = __ldg(conic_opacity + n_gid);
    }
#endif
    for (int base = 0; base < n; base += kBlendThreads) {
        if (__syncthreads_count(done) == kBlendThreads) break;
        const int p = base + threadIdx.x;
        bool keep = false;
        uint32_t gid = 0;
        float2 m = make_float2(0.f, 0.f);
        float a2 = 0.f, b2 = 0.f, c2 = 0.f, tau2 = 0.f, op = 0.f;
        if (p < n) {
#if GSR_BLEND_PREFETCH_FWD
            gid = n_gid;
            m = n_m;
            const float4 co = n_co;
#else
            gid = __ldg(point_list + range.x + p);
            m = __ldg(means2D + gid);
            const float4 co = __ldg(conic_opacity + gid);
#endif
            op = co.w;
            keep = stage_test(m, co, c.tile_rect, a2, b2, c2, tau2);
        }
#if GSR_BLEND_PREFETCH_FWD
        if (p + kBlendThreads < n) {   // the next batch's gathers fly during this batch's walk
            n_gid = __ldg(point_list + range.x + p + kBlendThreads);
            n_m = __ldg(means2D + n_gid);
            n_co = __ldg(conic_opacity + n_gid);
        }
#endif
        int total;
        const int slot = compact_slot(keep, s_wcount, total);
        if (keep) {
            const float4 col = __ldg(rgb + gid);
            s_e[3 * slot] = make_float4(m.x, m.y, a2, b2);
            s_e[3 * slot + 1] = make_float4(c2, op, tau2, __int_as_float(p + 1));
            s_e[3 * slot + 2] = col;
        }
        __syncthreads();

        for (int cb = 0; cb < total; cb += 32) {
            if (__all_sync(0xffffffffu, done)) break;
            const int jt = cb + (int)lane;
            bool hit = false;
            if (jt < total) {
                const float4 ta = s_e[3 * jt];
                const float4 tb = s_e[3 * jt + 1];
                hit = splat_may_reach(ta.x, ta.y, ta.z, ta.w, tb.x, tb.z, c.warp_rect);
            }
            unsigned mask = __ballot_sync(0xffffffffu, hit);
            while (mask) {
                const int j = cb + __ffs(mask) - 1;
                mask &= mask - 1;
                const float4 g = s_e[3 * j];
                const float4 cc = s_e[3 * j + 1];
                const float dx = g.x - fx, dy = g.y - fy;
                const float pw = fmaf(cc.x * dy, dy, fmaf(g.w, dy, g.z * dx) * dx);
                const float alpha = fminf(0.99f, cc.y * ex2_approx(pw));
                if (kCount) { if (!done) ++n_eval; ++n_iter; }
                if (done || pw > 0.0f || alpha < kAlphaMin) continue;
                const float test_T = T * (1.0f - alpha);
                if (test_T < kTmin) {
                    done = true;
                    if (kCount) stop_pos = (uint32_t)__float_as_int(cc.w);
                    continue;
                }
                if (kCount) ++n_blended;
                const float4 col = s_e[3 * j + 2];
                const float wgt = alpha * T;
                C0 = fmaf(col.x, wgt, C0);
                C1 = fmaf(col.y, wgt, C1);
                C2 = fmaf(col.z, wgt, C2);
                T = test_T;
                last = (uint32_t)__float_as_int(cc.w);
            }
        }
    }

    if (kCount) {
        // the reference walks its tile list until the pixel is done: stop position, else all n
        unsigned long long ev = c.inside ? (stop_pos ? stop_pos : (uint32_t)n) : 0u;
        unsigned long long bl = n_blended, ne = n_eval;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            ev += __shfl_xor_sync(0xffffffffu, ev, d);
            bl += __shfl_xor_sync(0xffffffffu, bl, d);
            ne += __shfl_xor_sync(0xffffffffu, ne, d);
        }
        if (lane == 0) {
            atomicAdd(counters, ev);
            atomicAdd(counters + 1, bl);
            atomicAdd(counters + 2, ne);
            atomicAdd(counters + 3, (unsigned long long)n_iter);  // (warp, splat) iterations issued
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
constexpr int kRedSplats = 3;  // splats per transposed-reduction batch: 27 rows <= 32 lanes

// Sums the parked rows of `nb` splats (row r = splat r/9, term r%9; 32 per-pixel entries each) and
// adds the totals to grad_accum as three 16-byte vector reductions per splat.  Lane r reads its row
// in 16-byte pieces starting at piece r%8, so the four lanes that share a bank group are the four
// wavefronts a 512-byte warp read needs anyway.
__device__ __forceinline__ void flush_parked(int nb, unsigned lane, const float* red, float* tot,
                                             const uint32_t* gid, float* __restrict__ grad_accum) {
    __syncwarp();
    if ((int)lane < 9 * nb) {
        const float4* row = reinterpret_cast<const float4*>(red) + lane * 8;
        float4 a = row[lane & 7];
#pragma unroll
        for (int j = 1; j < 8; ++j) {
            const float4 b = row[(lane + j) & 7];
            a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        }
        const int sp = (int)lane / 9;
        tot[sp * 12 + ((int)lane - sp * 9)] = (a.x + a.y) + (a.z + a.w);
    }
    __syncwarp();
    if ((int)lane < 3 * nb) {
        const int e = (int)lane / 3, quad = (int)lane - e * 3;
        const float4 acc = *reinterpret_cast<const float4*>(tot + e * 12 + quad * 4);
        if (acc.x != 0.f || acc.y != 0.f || acc.z != 0.f || acc.w != 0.f)
            atomicAdd(reinterpret_cast<float4*>(grad_accum + (size_t)gid[e] * 12 + quad * 4), acc);
    }
    __syncwarp();
}

// Two pixels per lane: 128 threads per tile, warp w owns an 8x8 pixel block, lane l the pixels
// (l & 7, l >> 3) and (l & 7, (l >> 3) + 4) of it.  Same decisions and the same per-pair arithmetic as a
// one-pixel-per-lane walk; what changes is that the per-(warp, splat) costs — the walk, the three shared
// loads of the staged entry, and above all the transposed shared-memory reduction (9 stores + the row
// sums, the pipe that bounds a one-pixel kernel) — are paid once per 64 pixels instead of once per 32:
// each lane adds its two pixels' nine terms in registers before parking them.
constexpr int kB2Threads = 128;
constexpr int kB2Warps = 4;

#ifndef GSR_B2_MINBLOCKS
#define GSR_B2_MINBLOCKS 6   // 80 registers, no spills once the per-pixel state is packed: 0.604 ms (5 CTAs/SM: 0.643 ms)
#endif
__global__ void __launch_bounds__(kB2Threads, GSR_B2_MINBLOCKS) blend_backward2_kernel(
    int gx, int W, int H, const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
    const float2* __restrict__ means2D, const float4* __restrict__ conic_opacity, const float4* __restrict__ rgb,
    const float* __restrict__ bg, const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib,
    const float* __restrict__ dL_dpix, float* __restrict__ grad_accum) {
    __shared__ float4 s_e[3 * kB2Threads];
    __shared__ __align__(16) float s_red[kB2Warps][kRedSplats * 9][32];
    __shared__ __align__(16) float s_tot[kB2Warps][kRedSplats][12];
    __shared__ uint32_t s_gid[kB2Warps][kRedSplats + 1];
    __shared__ int s_wcount[kB2Warps];
    __shared__ uint32_t s_wmax[kB2Warps];

    const int tile = blockIdx.x;
    const int ty = tile / gx, tx = tile - ty * gx;
    const unsigned lane = lane_id();
    const int warp = threadIdx.x >> 5;
    const int bx = tx * kTile + (warp & 1) * 8, by = ty * kTile + (warp >> 1) * 8;
    const int px = bx + (int)(lane & 7), py0 = by + (int)(lane >> 3);
    Rect tile_rect, warp_rect;
    tile_rect.x0 = (float)(tx * kTile);
    tile_rect.y0 = (float)(ty * kTile);
    tile_rect.x1 = (float)min(tx * kTile + kTile - 1, W - 1);
    tile_rect.y1 = (float)min(ty * kTile + kTile - 1, H - 1);
    warp_rect.x0 = (float)bx;
    warp_rect.y0 = (float)by;
    warp_rect.x1 = (float)min(bx + 7, W - 1);
    warp_rect.y1 = (float)min(by + 7, H - 1);
    const uint2 range = ranges[tile];
    const float fx = (float)px;
    if (lane < kRedSplats * 3) s_tot[warp][lane / 3][9 + lane % 3] = 0.0f;
    float* const my_red = &s_red[warp][0][lane];
    int nh = 0;

    // per-pixel state, packed for registers: row py0 + 4k is recomputed, final_T only survives as the product
    // final_T * (bg . dL/dpixel), and the running "colour behind" bh is advanced right after a splat is used
    // (alpha * colour + (1 - alpha) * bh — the expression the reference evaluates at the start of the NEXT
    // contributing splat from the saved alpha and colour) so neither has to be kept.
    float Tb[2], T[2], g0[2], g1[2], g2[2];
    uint32_t last[2];
    const size_t plane = (size_t)W * H;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const int py = py0 + 4 * k;
        T[k] = 0.f; g0[k] = 0.f; g1[k] = 0.f; g2[k] = 0.f; last[k] = 0;
        if (px < W && py < H) {
            const size_t pix = (size_t)py * W + px;
            T[k] = final_T[pix];
            last[k] = n_contrib[pix];
            g0[k] = __ldg(dL_dpix + pix);
            g1[k] = __ldg(dL_dpix + plane + pix);
            g2[k] = __ldg(dL_dpix + 2 * plane + pix);
        }
        Tb[k] = T[k] * (__ldg(bg) * g0[k] + __ldg(bg + 1) * g1[k] + __ldg(bg + 2) * g2[k]);
    }
    const float fy0 = (float)py0;
    const uint32_t wmax = __reduce_max_sync(0xffffffffu, max(last[0], last[1]));
    if (lane == 0) s_wmax[warp] = wmax;
    __syncthreads();
    uint32_t m = 0;
#pragma unroll
    for (int i = 0; i < kB2Warps; ++i) m = max(m, s_wmax[i]);

    float bh0[2] = {0.f, 0.f}, bh1[2] = {0.f, 0.f}, bh2[2] = {0.f, 0.f};

#if GSR_BLEND_PREFETCH_BWD
    // staging gathers (list entry -> mean + conic) of the NEXT batch are issued before the walk of the current one
    uint32_t n_gid = 0;
    float2 n_mm = make_float2(0.f, 0.f);
    float4 n_co = make_float4(0.f, 0.f, 0.f, 0.f);
    {
        const int p0 = (int)m - 1 - (int)threadIdx.x;
        if (p0 >= 0) {
            n_gid = __ldg(point_list + range.x + p0);
            n_mm = __ldg(means2D + n_gid);
            n_co = __ldg(conic_opacity + n_gid);
        }
    }
#endif
    for (int hi = (int)m; hi > 0; hi -= kB2Threads) {
        __syncthreads();
        const int p = hi - 1 - (int)threadIdx.x;
        bool keep = false;
        uint32_t gid = 0;
        float2 mm = make_float2(0.f, 0.f);
        float a2 = 0.f, b2 = 0.f, c2 = 0.f, tau2 = 0.f, op = 0.f;
        if (p >= 0) {
#if GSR_BLEND_PREFETCH_BWD
            gid = n_gid;
            mm = n_mm;
            const float4 co = n_co;
#else
            gid = __ldg(point_list + range.x + p);
            mm = __ldg(means2D + gid);
            const float4 co = __ldg(conic_opacity + gid);
#endif
            op = co.w;
            keep = stage_test(mm, co, tile_rect, a2, b2, c2, tau2);
        }
#if GSR_BLEND_PREFETCH_BWD
        {
            const int pn = p - kB2Threads;
            if (pn >= 0) {
                n_gid = __ldg(point_list + range.x + pn);
                n_mm = __ldg(means2D + n_gid);
                n_co = __ldg(conic_opacity + n_gid);
            }
        }
#endif
        int total;
        const int slot = compact_slot<kB2Warps>(keep, s_wcount, total);
        if (keep) {
            const float4 col = __ldg(rgb + gid);
            s_e[3 * slot] = make_float4(mm.x, mm.y, a2, b2);
            s_e[3 * slot + 1] = make_float4(c2, op, tau2, __int_as_float(p + 1));
            s_e[3 * slot + 2] = make_float4(col.x, col.y, col.z, __uint_as_float(gid));
        }
        __syncthreads();
        if (wmax == 0) continue;

        for (int cb = 0; cb < total; cb += 32) {
            const int jt = cb + (int)lane;
            bool hit = false;
            if (jt < total) {
                const float4 ta = s_e[3 * jt];
                const float4 tb = s_e[3 * jt + 1];
                hit = ((uint32_t)__float_as_int(tb.w) <= wmax) &&
                      splat_may_reach(ta.x, ta.y, ta.z, ta.w, tb.x, tb.z, warp_rect);
            }
            unsigned mask = __ballot_sync(0xffffffffu, hit);
            while (mask) {
                const int j = cb + __ffs(mask) - 1;
                mask &= mask - 1;
                const float4 g = s_e[3 * j];
                const float4 cc = s_e[3 * j + 1];
                const float dx = g.x - fx;
                const float lin = g.z * dx;  // a2 * dx
                float dy[2], G[2], alpha[2], pw[2];
                bool ok[2];
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    dy[k] = g.y - (fy0 + (float)(4 * k));
                    pw[k] = fmaf(cc.x * dy[k], dy[k], fmaf(g.w, dy[k], lin) * dx);
                    G[k] = ex2_approx(pw[k]);
                    alpha[k] = fminf(0.99f, cc.y * G[k]);
                    ok[k] = ((uint32_t)__float_as_int(cc.w) <= last[k]) && !(pw[k] > 0.0f) && !(alpha[k] < kAlphaMin);
                }
                if (!__any_sync(0xffffffffu, ok[0] || ok[1])) continue;
                const float4 col = s_e[3 * j + 2];
                float t0 = 0.f, t1 = 0.f, t2 = 0.f, t3 = 0.f, t4 = 0.f, t5 = 0.f, t6 = 0.f, t7 = 0.f, t8 = 0.f;
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    if (ok[k]) {
                        const float ra = rcp_approx(1.0f - alpha[k]);
                        T[k] *= ra;
                        const float wgt = alpha[k] * T[k];
                        float dL_dalpha = (col.x - bh0[k]) * g0[k] + (col.y - bh1[k]) * g1[k] + (col.z - bh2[k]) * g2[k];
                        dL_dalpha *= T[k];
                        dL_dalpha -= Tb[k] * ra;
                        bh0[k] = alpha[k] * col.x + (1.0f - alpha[k]) * bh0[k];
                        bh1[k] = alpha[k] * col.y + (1.0f - alpha[k]) * bh1[k];
                        bh2[k] = alpha[k] * col.z + (1.0f - alpha[k]) * bh2[k];
                        const float gdl = G[k] * dL_dalpha;
                        const float h = cc.y * gdl;
                        const float hx = h * dx, hy = h * dy[k];
                        t0 += hx;
                        t1 += hy;
                        t2 = fmaf(hx, dx, t2);
                        t3 = fmaf(hx, dy[k], t3);
                        t4 = fmaf(hy, dy[k], t4);
                        t5 += gdl;
                        t6 = fmaf(wgt, g0[k], t6);
                        t7 = fmaf(wgt, g1[k], t7);
                        t8 = fmaf(wgt, g2[k], t8);
                    }
                }
                float* dst = my_red + nh * (9 * 32);
                dst[0 * 32] = t0;
                dst[1 * 32] = t1;
                dst[2 * 32] = t2;
                dst[3 * 32] = t3;
                dst[4 * 32] = t4;
                dst[5 * 32] = t5;
                dst[6 * 32] = t6;
                dst[7 * 32] = t7;
                dst[8 * 32] = t8;
                if (lane == 0) s_gid[warp][nh] = __float_as_uint(col.w);
                if (++nh == kRedSplats) {
                    flush_parked(nh, lane, &s_red[warp][0][0], &s_tot[warp][0][0], s_gid[warp], grad_accum);
                    nh = 0;
                }
            }
        }
    }
    if (nh) flush_parked(nh, lane, &s_red[warp][0][0], &s_tot[warp][0][0], s_gid[warp], grad_accum);
}

}  // namespace

int launch_blend_forward(const Grid& g, const float* bg, GeomView gv, BinView bv, ImageView iv,
                         float* out_color, cudaStream_t s, int debug) {
    StageScope scope(kStageBlendFwd, s);
    // (a two-pixels-per-lane forward was measured too: 0.319 ms against 0.312 ms — the forward has no reduction to
    // amortise and the coarser 8x8 block test lets more pairs through — so the forward keeps one pixel per lane)
    blend_forward_kernel<false><<<g.tiles, kBlendThreads, 0, s>>>(g.gx, g.W, g.H, iv.ranges, bv.point_list,
                                                                  gv.means2D, gv.conic_opacity, gv.rgb, bg, out_color,
                                                                  iv.final_T, iv.n_contrib, nullptr);
    GSR_AFTER_LAUNCH(s, debug);
    return GSR_OK;
}

int launch_blend_pair_counts(const Grid& g, GeomView gv, BinView bv, ImageView iv, unsigned long long* counters,
                             cudaStream_t s) {
    GSR_CUDA(cudaMemsetAsync(counters, 0, 4 * sizeof(unsigned long long), s));
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
    blend_backward2_kernel<<<g.tiles, kB2Threads, 0, s>>>(g.gx, g.W, g.H, iv.ranges, bv.point_list, gv.means2D,
                                                          gv.conic_opacity, gv.rgb, bg, iv.final_T, iv.n_contrib,
                                                          dL_dpix, gv.grad_accum);
    GSR_AFTER_LAUNCH(s, debug);
    return GSR_OK;
}

}  // namespace gsr
