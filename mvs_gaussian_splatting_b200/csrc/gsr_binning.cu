// gsr_binning.cu — tile binning for sm_100a: prefix sum of tiles_touched, (tile|depth) key emission,
// a hand-written stable LSD radix sort of the 64-bit keys with their 32-bit payload, and per-tile
// range identification (the reference's K2, K4, K5, K6: cub::DeviceScan::InclusiveSum,
// duplicateWithKeys, cub::DeviceRadixSort::SortPairs, identifyTileRanges — SURVEY §2.1, §8a a13-a16).
//
// Everything is integer work bound by HBM.  The instance count R lives on the DEVICE (last entry
// of point_offsets); kernels are launched for the caller's instance_capacity and bound themselves
// by R, so the host never waits for R (the reference stalls on a 4-byte D2H copy here, K3).
// Results are bit-exact with the reference algorithm: same key format, stable sort (equal keys
// keep Gaussian-index order), same [start,end) ranges.
#include "gsr_internal.cuh"

namespace gsr {

namespace {

// =========================== generic inclusive scan (u32) ================================
constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;  // per thread (two uint4)
constexpr int kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane_id() >= (unsigned)d) v += t;
    }
    return v;
}

// Block-wide exclusive scan of one value per thread (256 threads); returns exclusive prefix and
// the block total through `total`.
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* s_warp, uint32_t& total) {
    const uint32_t incl = warp_incl_scan(v);
    const int warp = threadIdx.x >> 5;
    if (lane_id() == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = lane_id() < kScanThreads / 32 ? s_warp[lane_id()] : 0u;
        uint32_t wi = warp_incl_scan(w);
        if (lane_id() < kScanThreads / 32) s_warp[lane_id()] = wi - w;
        if (lane_id() == kScanThreads / 32 - 1) s_warp[kScanThreads / 32] = wi;
    }
    __syncthreads();
    total = s_warp[kScanThreads / 32];
    const uint32_t r = s_warp[warp] + incl - v;
    __syncthreads();
    return r;
}

__device__ __forceinline__ void scan_load_tile(const uint32_t* __restrict__ in, size_t n, size_t base,
                                               uint32_t* v) {
    const size_t i0 = base + (size_t)threadIdx.x * kScanItems;
    if (i0 + kScanItems <= n && ((reinterpret_cast<uintptr_t>(in + i0) & 15) == 0)) {
        const uint4 a = *reinterpret_cast<const uint4*>(in + i0);
        const uint4 b = *reinterpret_cast<const uint4*>(in + i0 + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int k = 0; k < kScanItems; ++k) v[k] = (i0 + k < n) ? in[i0 + k] : 0u;
    }
}

__global__ void __launch_bounds__(kScanThreads) scan_reduce_kernel(const uint32_t* __restrict__ in, size_t n,
                                                                   uint32_t* __restrict__ block_sums) {
    __shared__ uint32_t s_warp[kScanThreads / 32 + 1];
    uint32_t v[kScanItems];
    scan_load_tile(in, n, (size_t)blockIdx.x * kScanTile, v);
    uint32_t sum = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) sum += v[k];
    uint32_t total;
    block_excl_scan(sum, s_warp, total);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

// Single CTA: in-place exclusive scan of the block sums.
__global__ void __launch_bounds__(kScanThreads) scan_sums_kernel(uint32_t* __restrict__ sums, size_t nb) {
    __shared__ uint32_t s_warp[kScanThreads / 32 + 1];
    uint32_t carry = 0;
    for (size_t base = 0; base < nb; base += kScanThreads) {
        const size_t i = base + threadIdx.x;
        const uint32_t v = i < nb ? sums[i] : 0u;
        uint32_t total;
        const uint32_t ex = block_excl_scan(v, s_warp, total);
        if (i < nb) sums[i] = carry + ex;
        carry += total;
    }
}

template <bool kExclusive>
__global__ void __launch_bounds__(kScanThreads) scan_apply_kernel(const uint32_t* __restrict__ in,
                                                                  uint32_t* __restrict__ out, size_t n,
                                                                  const uint32_t* __restrict__ block_offsets) {
    __shared__ uint32_t s_warp[kScanThreads / 32 + 1];
    uint32_t v[kScanItems];
    const size_t base = (size_t)blockIdx.x * kScanTile;
    scan_load_tile(in, n, base, v);
    uint32_t sum = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) sum += v[k];
    uint32_t total;
    uint32_t run = block_excl_scan(sum, s_warp, total) + block_offsets[blockIdx.x];
    const size_t i0 = base + (size_t)threadIdx.x * kScanItems;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
        const uint32_t ex = run;
        run += v[k];
        if (i0 + k < n) out[i0 + k] = kExclusive ? ex : run;
    }
}

int scan_u32(const uint32_t* in, uint32_t* out, size_t n, uint32_t* tmp, bool exclusive,
             cudaStream_t s, int debug) {
    if (n == 0) return GSR_OK;
    const size_t nb = (n + kScanTile - 1) / kScanTile;
    scan_reduce_kernel<<<(unsigned)nb, kScanThreads, 0, s>>>(in, n, tmp);
    GSR_AFTER_LAUNCH(s, debug);
    scan_sums_kernel<<<1, kScanThreads, 0, s>>>(tmp, nb);
    GSR_AFTER_LAUNCH(s, debug);
    if (exclusive) scan_apply_kernel<true><<<(unsigned)nb, kScanThreads, 0, s>>>(in, out, n, tmp);
    else scan_apply_kernel<false><<<(unsigned)nb, kScanThreads, 0, s>>>(in, out, n, tmp);
    GSR_AFTER_LAUNCH(s, debug);
    return GSR_OK;
}

// Effective instance count on the device: R if it fits the capacity, else 0 (overflow).
__device__ __forceinline__ uint32_t effective_n(const uint32_t* __restrict__ r_ptr, uint32_t cap) {
    const uint32_t r = __ldg(r_ptr);
    return r <= cap ? r : 0u;
}

// Publishes num_rendered / overflow for the host and clears the tile ranges' guard.
__global__ void publish_info_kernel(const uint32_t* __restrict__ r_ptr, uint32_t cap, int has_cap,
                                    gsr_forward_info* info) {
    const uint32_t r = *r_ptr;
    info->num_rendered = (int32_t)r;
    if (has_cap) info->overflow = r > cap ? 1 : 0;
}

// =========================== key emission (duplicateWithKeys) =============================
// One lane per Gaussian; rects of up to kSmallRect tiles are written by the owning lane, larger
// ones cooperatively by the whole warp (coalesced, and a screen-filling splat no longer serialises
// thousands of stores on one thread).  Entry t of a rect (row-major, y outer) goes to
// offsets[idx-1] + t, so the unsorted list is identical to a serial emission in Gaussian order.
constexpr int kSmallRect = 4;

__global__ void __launch_bounds__(256) emit_keys_kernel(int P, int gx, int gy, const int32_t* __restrict__ radii,
                                                        const float2* __restrict__ means2D,
                                                        const float* __restrict__ depths,
                                                        const uint32_t* __restrict__ offsets, uint32_t cap,
                                                        uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (effective_n(offsets + (P - 1), cap) == 0u) return;
    int x0 = 0, y0 = 0, w = 0, count = 0;
    uint32_t off = 0, dbits = 0;
    if (idx < P) {
        const int r = radii[idx];
        if (r > 0) {
            const float2 m = means2D[idx];
            int x1, y1;
            tile_rect(m.x, m.y, r, gx, gy, x0, y0, x1, y1);
            w = x1 - x0;
            count = w * (y1 - y0);
            off = idx == 0 ? 0u : offsets[idx - 1];
            dbits = __float_as_uint(depths[idx]);
        }
    }
    if (count > 0 && count <= kSmallRect) {
        for (int t = 0; t < count; ++t) {
            const int y = y0 + t / w, x = x0 + t - (t / w) * w;
            keys[off + t] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | dbits;
            vals[off + t] = (uint32_t)idx;
        }
    }
    unsigned big = __ballot_sync(0xffffffffu, count > kSmallRect);
    while (big) {
        const int src = __ffs(big) - 1;
        big &= big - 1;
        const int bx0 = __shfl_sync(0xffffffffu, x0, src), by0 = __shfl_sync(0xffffffffu, y0, src);
        const int bw = __shfl_sync(0xffffffffu, w, src), bcount = __shfl_sync(0xffffffffu, count, src);
        const uint32_t boff = __shfl_sync(0xffffffffu, off, src), bd = __shfl_sync(0xffffffffu, dbits, src);
        const uint32_t bidx = (uint32_t)__shfl_sync(0xffffffffu, idx, src);
        for (int t = lane_id(); t < bcount; t += 32) {
            const int yy = t / bw;
            const int y = by0 + yy, x = bx0 + t - yy * bw;
            keys[boff + t] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | bd;
            vals[boff + t] = bidx;
        }
    }
}

// =========================== stable LSD radix sort ========================================
// Per pass: (1) per-CTA digit histogram, (2) exclusive scan of the [digit][cta] table, (3) stable
// scatter.  A CTA owns kSortTile consecutive items; warp w owns a consecutive slice of those, read
// 32 at a time in order, so warp-match ranking + warp order + CTA order = global stable order.
constexpr int kSortThreads = 512;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortItems = 16;  // per thread
constexpr int kSortTile = kSortThreads * kSortItems;
constexpr int kSortWarpSpan = 32 * kSortItems;

__global__ void __launch_bounds__(kSortThreads) sort_hist_kernel(const uint64_t* __restrict__ keys,
                                                                 const uint32_t* __restrict__ r_ptr, uint32_t cap,
                                                                 int shift, int bits, uint32_t* __restrict__ table,
                                                                 int nblk) {
    extern __shared__ uint32_t s_hist[];
    const int bins = 1 << bits;
    const uint32_t n = effective_n(r_ptr, cap);
    for (int d = threadIdx.x; d < bins; d += kSortThreads) s_hist[d] = 0;
    __syncthreads();
    const size_t start = (size_t)blockIdx.x * kSortTile;
    if (start < n) {
        const uint32_t mask = (uint32_t)bins - 1u;
#pragma unroll 4
        for (int k = 0; k < kSortItems; ++k) {
            const size_t i = start + (size_t)k * kSortThreads + threadIdx.x;
            if (i < n) atomicAdd(&s_hist[(uint32_t)(keys[i] >> shift) & mask], 1u);
        }
    }
    __syncthreads();
    for (int d = threadIdx.x; d < bins; d += kSortThreads) table[(size_t)d * nblk + blockIdx.x] = s_hist[d];
}

__global__ void __launch_bounds__(kSortThreads) sort_scatter_kernel(
    const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in, uint64_t* __restrict__ keys_out,
    uint32_t* __restrict__ vals_out, const uint32_t* __restrict__ r_ptr, uint32_t cap, int shift, int bits,
    const uint32_t* __restrict__ table, int nblk) {
    extern __shared__ uint32_t s_cnt[];  // [kSortWarps][bins]
    const int bins = 1 << bits;
    const uint32_t n = effective_n(r_ptr, cap);
    const size_t start = (size_t)blockIdx.x * kSortTile;
    if (start >= n) return;
    for (int d = threadIdx.x; d < bins * kSortWarps; d += kSortThreads) s_cnt[d] = 0;
    __syncthreads();

    const int warp = threadIdx.x >> 5;
    const unsigned lane = lane_id();
    const unsigned lt = (1u << lane) - 1u;
    const uint32_t mask = (uint32_t)bins - 1u;
    uint32_t* my = s_cnt + warp * bins;
    const size_t wbase = start + (size_t)warp * kSortWarpSpan;

    uint64_t key[kSortItems];
    uint32_t rank[kSortItems];
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const size_t i = wbase + (size_t)k * 32 + lane;
        key[k] = i < n ? keys_in[i] : ~0ull;
    }
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const size_t i = wbase + (size_t)k * 32 + lane;
        const bool valid = i < n;
        const uint32_t d = valid ? ((uint32_t)(key[k] >> shift) & mask) : 0xffffffffu;
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        const uint32_t r = __popc(peers & lt);
        uint32_t prior = 0;
        if (valid) prior = my[d];
        __syncwarp();
        if (valid && r == 0) my[d] = prior + __popc(peers);
        __syncwarp();
        rank[k] = prior + r;
    }
    __syncthreads();
    // per digit: turn the per-warp counts into per-warp global base offsets
    for (int d = threadIdx.x; d < bins; d += kSortThreads) {
        uint32_t run = table[(size_t)d * nblk + blockIdx.x];
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w) {
            const uint32_t c = s_cnt[w * bins + d];
            s_cnt[w * bins + d] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const size_t i = wbase + (size_t)k * 32 + lane;
        if (i < n) {
            const uint32_t d = (uint32_t)(key[k] >> shift) & mask;
            const uint32_t dst = my[d] + rank[k];
            keys_out[dst] = key[k];
            vals_out[dst] = vals_in[i];
        }
    }
}

// =========================== tile ranges ==================================================
__global__ void __launch_bounds__(256) tile_ranges_kernel(const uint64_t* __restrict__ keys,
                                                          const uint32_t* __restrict__ r_ptr, uint32_t cap,
                                                          uint2* __restrict__ ranges) {
    const uint32_t n = effective_n(r_ptr, cap);
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t t = (uint32_t)(keys[i] >> 32);
    if (i == 0) ranges[t].x = 0;
    else {
        const uint32_t tp = (uint32_t)(keys[i - 1] >> 32);
        if (tp != t) {
            ranges[tp].y = i;
            ranges[t].x = i;
        }
    }
    if (i == n - 1) ranges[t].y = n;
}

int higher_msb(uint32_t n) {  // number of bits needed to represent values < n (at least 1)
    int b = 1;
    while (b < 32 && (1u << b) < n) ++b;
    return b;
}

}  // namespace

size_t scan_tmp_entries(size_t n) { return (n + kScanTile - 1) / kScanTile + 1; }

// Radix digit width: the smallest that sorts `total_bits` in an ODD number of passes, so that the
// ping-pong ends in the (keys, point_list) buffers while the unsorted lists stay intact.
static void sort_plan(int total_bits, int& passes, int& bits) {
    passes = 5;
    bits = (total_bits + passes - 1) / passes;
    if (bits > 10) { passes = 7; bits = (total_bits + passes - 1) / passes; }
    if (bits < 4) bits = 4;
}

size_t sort_tmp_entries(size_t cap) {
    const size_t nblk = (cap + kSortTile - 1) / kSortTile;
    const size_t table = nblk * 1024;  // worst-case 10-bit digits
    return table + scan_tmp_entries(table) + 16;
}

int launch_inclusive_scan(const uint32_t* in, uint32_t* out, size_t n, uint32_t* tmp, cudaStream_t s,
                          int debug) {
    StageScope scope(kStageScan, s);
    return scan_u32(in, out, n, tmp, false, s, debug);
}

int launch_binning(int P, const Grid& g, const int32_t* radii, GeomView gv, BinView bv, int64_t cap64,
                   ImageView iv, cudaStream_t s, int debug) {
    const uint32_t cap = (uint32_t)cap64;
    const uint32_t* r_ptr = gv.point_offsets + (P - 1);
    publish_info_kernel<<<1, 1, 0, s>>>(r_ptr, cap, 1, gv.info);
    GSR_AFTER_LAUNCH(s, debug);
    GSR_CUDA(cudaMemsetAsync(iv.ranges, 0, sizeof(uint2) * (size_t)g.tiles, s));
    if (cap == 0) return GSR_OK;

    {
    StageScope scope(kStageEmitKeys, s);
    emit_keys_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, g.gx, g.gy, radii, gv.means2D, gv.depths,
                                                     gv.point_offsets, cap, bv.keys_unsorted, bv.vals_unsorted);
    GSR_AFTER_LAUNCH(s, debug);
    }

    int passes, bits;
    sort_plan(32 + higher_msb((uint32_t)g.tiles), passes, bits);
    const int nblk = (int)((cap + kSortTile - 1) / kSortTile);
    const size_t table_n = (size_t)nblk << bits;
    uint32_t* table = bv.sort_tmp;
    uint32_t* scan_tmp = bv.sort_tmp + (size_t)nblk * 1024;
    const size_t hist_smem = sizeof(uint32_t) << bits;
    const size_t scat_smem = hist_smem * kSortWarps;
    static bool attr_set = false;  // idempotent; racing threads set the same value
    if (!attr_set) {
        GSR_CUDA(cudaFuncSetAttribute(sort_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)(sizeof(uint32_t) * 1024 * kSortWarps)));
        attr_set = true;
    }
    // pass 0 reads the unsorted lists; later passes ping-pong between (keys, point_list) and the
    // alternate pair, so the unsorted lists stay intact and an odd pass count ends in point_list.
    const uint64_t* kin = bv.keys_unsorted;
    const uint32_t* vin = bv.vals_unsorted;
    uint64_t* kout = bv.keys;
    uint32_t* vout = bv.point_list;
    {
    StageScope scope(kStageSort, s);
    for (int pass = 0; pass < passes; ++pass) {
        const int shift = pass * bits;
        sort_hist_kernel<<<nblk, kSortThreads, hist_smem, s>>>(kin, r_ptr, cap, shift, bits, table, nblk);
        GSR_AFTER_LAUNCH(s, debug);
        int st = scan_u32(table, table, table_n, scan_tmp, true, s, debug);
        if (st != GSR_OK) return st;
        sort_scatter_kernel<<<nblk, kSortThreads, scat_smem, s>>>(kin, vin, kout, vout, r_ptr, cap, shift, bits,
                                                                  table, nblk);
        GSR_AFTER_LAUNCH(s, debug);
        kin = kout;
        vin = vout;
        const bool to_alt = (kout == bv.keys);
        kout = to_alt ? bv.keys_alt : bv.keys;
        vout = to_alt ? bv.vals_alt : bv.point_list;
    }
    }
    StageScope scope(kStageTileRanges, s);
    tile_ranges_kernel<<<(cap + 255) / 256, 256, 0, s>>>(bv.keys, r_ptr, cap, iv.ranges);
    GSR_AFTER_LAUNCH(s, debug);
    return GSR_OK;
}

}  // namespace gsr
