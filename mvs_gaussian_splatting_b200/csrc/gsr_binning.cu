// gsr_binning.cu — tile binning for sm_100a (the reference's K2-K6: cub::DeviceScan::InclusiveSum,
// duplicateWithKeys, cub::DeviceRadixSort::SortPairs, identifyTileRanges — SURVEY §2.1, §8a a13-a16).
//
// The reference emits one 64-bit (tile << 32 | depth bits) key per tile instance in Gaussian order
// and sorts all R of them over 32 + log2(tiles) bits.  The sorted result is fully determined by
// the (tile, depth bits, Gaussian index) order, which factorises; this file produces the SAME sorted
// key list / point_list / ranges with far less traffic:
//   1. stable LSD radix sort of the P per-Gaussian (depth bits -> index) pairs, 4 x 8 bits
//      (culled Gaussians carry key 0xFFFFFFFF and sink to the end; P-sized, mostly L2-resident);
//   2. single-pass prefix sum of tiles_touched in that order;
//   3. key emission in (depth, index) order — the unsorted list is a permutation of the reference's;
//   4. stable radix sort of the R instances on the TILE bits only (2 passes for <= 65536 tiles).
// Every radix pass is ONE kernel: chunk histograms are chained with decoupled look-back (ticketed
// chunks, 2 flag bits + 30 count bits per (chunk, digit) status word), keys are ranked per warp with
// one ballot per digit bit, reordered through shared memory and written as coalesced per-digit runs.  Digit
// histograms of all passes of a sort come from one read of the keys.
//
// Everything is integer work bound by HBM / L2.  The instance count R lives on the DEVICE
// (info->num_rendered); kernels are launched for the caller's instance_capacity and bound themselves
// by R, so the host never waits for R (the reference stalls on a 4-byte D2H copy here, K3).
#include "gsr_internal.cuh"

namespace gsr {

namespace {

constexpr uint32_t kInvalidDepthKey = 0xFFFFFFFFu;  // written by preprocess for culled Gaussians

__device__ __forceinline__ uint32_t ld_relaxed(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane_id() >= (unsigned)d) v += t;
    }
    return v;
}

// Block-wide exclusive scan of one value per thread (256 threads); returns the exclusive prefix and
// the block total through `total`.  s_warp: 9 words.
constexpr int kBlk = 256;
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* s_warp, uint32_t& total) {
    const uint32_t incl = warp_incl_scan(v);
    const int warp = threadIdx.x >> 5;
    if (lane_id() == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = lane_id() < kBlk / 32 ? s_warp[lane_id()] : 0u;
        uint32_t wi = warp_incl_scan(w);
        if (lane_id() < kBlk / 32) s_warp[lane_id()] = wi - w;
        if (lane_id() == kBlk / 32 - 1) s_warp[kBlk / 32] = wi;
    }
    __syncthreads();
    total = s_warp[kBlk / 32];
    const uint32_t r = s_warp[warp] + incl - v;
    __syncthreads();
    return r;
}

// Effective instance count on the device: R if it fits the capacity, else 0 (overflow).
// (R was written by an earlier kernel of the stream, so the read-only path is safe.)
__device__ __forceinline__ uint32_t effective_n(const uint32_t* __restrict__ r_ptr, uint32_t cap) {
    const uint32_t r = __ldg(r_ptr);
    return r <= cap ? r : 0u;
}

// Publishes the overflow flag for the host.
__global__ void publish_info_kernel(uint32_t cap, gsr_forward_info* info) {
    info->overflow = (uint32_t)info->num_rendered > cap ? 1 : 0;
}

// =========================== prefix sum in depth order =====================================
// out[k] = sum_{j <= k} tiles_touched[order[j]]; single pass, decoupled look-back over ticketed
// chunks (status word: flag << 32 | value; flag 1 = chunk aggregate, 2 = inclusive prefix).
constexpr int kScanItems = 8;
constexpr int kScanChunk = kBlk * kScanItems;

__global__ void __launch_bounds__(kBlk) scan_sorted_tiles_kernel(int P, const uint32_t* __restrict__ order,
                                                                 const uint32_t* __restrict__ tiles_touched,
                                                                 uint32_t* __restrict__ out,
                                                                 unsigned long long* status, uint32_t* ticket,
                                                                 gsr_forward_info* info) {
    __shared__ uint32_t s_warp[kBlk / 32 + 1];
    __shared__ uint32_t s_b, s_prefix;
    if (threadIdx.x == 0) s_b = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t b = s_b;
    const size_t i0 = (size_t)b * kScanChunk + (size_t)threadIdx.x * kScanItems;
    uint32_t v[kScanItems];
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) v[k] = 0;
    if (i0 + kScanItems <= (size_t)P) {
        const uint4 a = *reinterpret_cast<const uint4*>(order + i0);
        const uint4 c = *reinterpret_cast<const uint4*>(order + i0 + 4);
        const uint32_t id[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
#pragma unroll
        for (int k = 0; k < kScanItems; ++k) v[k] = __ldg(tiles_touched + id[k]);
    } else {
#pragma unroll
        for (int k = 0; k < kScanItems; ++k)
            if (i0 + k < (size_t)P) v[k] = __ldg(tiles_touched + order[i0 + k]);
    }
    uint32_t sum = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) sum += v[k];
    uint32_t total;
    const uint32_t ex = block_excl_scan(sum, s_warp, total);

    if (threadIdx.x < 32) {  // warp 0: publish, then look back 32 chunks at a time
        const unsigned lane = lane_id();
        uint32_t prefix = 0;
        if (b == 0) {
            if (lane == 0) st_relaxed64(status, (2ull << 32) | total);
        } else {
            if (lane == 0) st_relaxed64(status + b, (1ull << 32) | total);
            int base = (int)b - 1;
            while (true) {
                const int j = base - (int)lane;
                unsigned long long s = 2ull << 32;  // before the first chunk: inclusive 0
                if (j >= 0) {
                    do { s = ld_relaxed64(status + j); } while ((s >> 32) == 0);
                }
                const unsigned incl = __ballot_sync(0xffffffffu, (s >> 32) == 2);
                const int first = __ffs(incl) - 1;  // nearest chunk holding an inclusive prefix, or -1
                uint32_t val = (first < 0 || (int)lane <= first) ? (uint32_t)s : 0u;
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) val += __shfl_xor_sync(0xffffffffu, val, d);
                prefix += val;
                if (first >= 0) break;
                base -= 32;
            }
            if (lane == 0) st_relaxed64(status + b, (2ull << 32) | (prefix + total));
        }
        if (lane == 0) s_prefix = prefix;
    }
    __syncthreads();
    uint32_t run = s_prefix + ex;
    uint32_t o[kScanItems];
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) { run += v[k]; o[k] = run; }
    if (i0 + kScanItems <= (size_t)P) {
        *reinterpret_cast<uint4*>(out + i0) = make_uint4(o[0], o[1], o[2], o[3]);
        *reinterpret_cast<uint4*>(out + i0 + 4) = make_uint4(o[4], o[5], o[6], o[7]);
    } else {
#pragma unroll
        for (int k = 0; k < kScanItems; ++k)
            if (i0 + k < (size_t)P) out[i0 + k] = o[k];
    }
    if ((size_t)b * kScanChunk + kScanChunk >= (size_t)P && threadIdx.x == kBlk - 1)
        info->num_rendered = (int32_t)run;  // last chunk, last thread: grand total R
}

// =========================== stable LSD radix sort, one kernel per pass ====================
#ifndef GSR_SORT_ITEMS
#define GSR_SORT_ITEMS 24   // A/B on B200 (config 3): 8/12/16/20/24/32 items -> 24 at 2 CTAs/SM is fastest and spill-free
#endif
#ifndef GSR_SORT_MINBLOCKS
#define GSR_SORT_MINBLOCKS 2
#endif
constexpr int kSortItems = GSR_SORT_ITEMS;      // per thread
constexpr int kSortChunk = kBlk * kSortItems;   // 6144 items per ticket
constexpr int kSortWarps = kBlk / 32;
constexpr int kSortWarpSpan = 32 * kSortItems;  // consecutive items owned by one warp
constexpr int kMaxBins = 256;
constexpr int kMaxPasses = 4;
constexpr uint32_t kFlagAgg = 1u << 30, kFlagIncl = 2u << 30, kValMask = (1u << 30) - 1u;

// Digit histograms of ALL passes of one sort from a single read of the keys: hist[pass][digit].
// Shared atomics, with a vote so that warps whose digit is constant (culled keys, high depth bytes)
// issue one add instead of 32 serialised ones.
template <typename K>
__global__ void __launch_bounds__(kBlk) radix_hist_kernel(const K* __restrict__ keys, const uint32_t* __restrict__ n_ptr,
                                                          uint32_t n_host, uint32_t cap, int shift0, int bits,
                                                          int npass, uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_h[kMaxPasses][kMaxBins];
    for (int i = threadIdx.x; i < kMaxPasses * kMaxBins; i += kBlk) (&s_h[0][0])[i] = 0;
    __syncthreads();
    const uint32_t n = n_ptr ? effective_n(n_ptr, cap) : n_host;
    const uint32_t mask = (1u << bits) - 1u;
    const unsigned lane = lane_id();
    // whole warps iterate together (the uniformity vote needs converged lanes): round n up to a warp multiple
    const size_t stride = (size_t)gridDim.x * kBlk;
    const size_t n_up = ((size_t)n + 31) & ~(size_t)31;
    for (size_t i = (size_t)blockIdx.x * kBlk + threadIdx.x; i < n_up; i += stride) {
        const bool valid = i < n;
        const K key = valid ? keys[i] : (K)0;
        const bool full = __all_sync(0xffffffffu, valid);
        for (int p = 0; p < npass; ++p) {
            const uint32_t d = (uint32_t)(key >> (shift0 + p * bits)) & mask;
            const uint32_t d0 = __shfl_sync(0xffffffffu, d, 0);
            if (full && __all_sync(0xffffffffu, d == d0)) {  // constant digit (culled keys, high depth bytes)
                if (lane == 0) atomicAdd(&s_h[p][d0], 32u);
            } else if (valid) {
                atomicAdd(&s_h[p][d], 1u);
            }
        }
    }
    __syncthreads();
    const int bins = 1 << bits;
    for (int i = threadIdx.x; i < npass * bins; i += kBlk) {
        const int p = i / bins, d = i - p * bins;
        const uint32_t c = s_h[p][d];
        if (c) atomicAdd(hist + p * kMaxBins + d, c);
    }
}

// One pass.  kIdentityVal: the payload of item i is i itself (first pass of the depth sort), so no
// payload array is read.  status: [chunks][bins] words, zeroed before the launch; ticket: zeroed.
template <typename K, bool kIdentityVal, int kBits>
__global__ void __launch_bounds__(kBlk, GSR_SORT_MINBLOCKS) radix_pass_kernel(const K* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                          K* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                                                          const uint32_t* __restrict__ n_ptr, uint32_t n_host, uint32_t cap,
                                                          int shift, const uint32_t* __restrict__ hist,
                                                          uint32_t* status, uint32_t* ticket) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    K* s_keys = reinterpret_cast<K*>(s_raw);                                          // [kSortChunk]
    uint32_t* s_vals = reinterpret_cast<uint32_t*>(s_raw + sizeof(K) * kSortChunk);   // [kSortChunk]
    uint32_t* s_cnt = s_vals + kSortChunk;                                            // [kSortWarps][bins]
    __shared__ uint32_t s_binstart[kMaxBins];  // exclusive start of each digit inside this chunk
    __shared__ uint32_t s_gbase[kMaxBins];     // global position of sorted-chunk item i of digit d: s_gbase[d] + i
    __shared__ uint32_t s_warp[kBlk / 32 + 1];
    __shared__ uint32_t s_b;

    constexpr int bins = 1 << kBits;
    const uint32_t n = n_ptr ? effective_n(n_ptr, cap) : n_host;
    if (threadIdx.x == 0) s_b = atomicAdd(ticket, 1u);
    for (int d = threadIdx.x; d < bins * kSortWarps; d += kBlk) s_cnt[d] = 0;
    __syncthreads();
    const uint32_t b = s_b;
    const size_t start = (size_t)b * kSortChunk;
    if (start >= n) return;
    const uint32_t count = (uint32_t)min((size_t)kSortChunk, (size_t)n - start);

    const int warp = threadIdx.x >> 5;
    const unsigned lane = lane_id();
    const unsigned lt = (1u << lane) - 1u;
    const uint32_t mask = (uint32_t)bins - 1u;
    uint32_t* my = s_cnt + warp * bins;
    const size_t wbase = start + (size_t)warp * kSortWarpSpan;

    K key[kSortItems];
    uint32_t rank[kSortItems];
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const size_t i = wbase + (size_t)k * 32 + lane;
        key[k] = i < n ? keys_in[i] : (K)0;
    }
    // Rank inside the warp: lanes holding the same digit are found with one ballot per digit bit
    // (pipelined votes; MATCH.ANY costs ~150 cycles a piece on sm_100), the lowest such lane bumps the
    // warp's private counter with one shared atomic and broadcasts the old value.
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const size_t i = wbase + (size_t)k * 32 + lane;
        const bool valid = i < n;
        const uint32_t d = (uint32_t)(key[k] >> shift) & mask;
        unsigned peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
        for (int bit = 0; bit < kBits; ++bit) {
            const bool one = (d >> bit) & 1u;
            const unsigned bal = __ballot_sync(0xffffffffu, one);
            peers &= one ? bal : ~bal;
        }
        const int leader = __ffs(peers) - 1;
        uint32_t prior = 0;
        if (valid && (int)lane == leader) prior = atomicAdd(&my[d], (uint32_t)__popc(peers));
        prior = __shfl_sync(0xffffffffu, prior, leader < 0 ? 0 : leader);
        rank[k] = prior + __popc(peers & lt);
    }
    __syncthreads();

    // thread d owns digit d: per-warp counts -> exclusive offsets, chunk count, look-back
    uint32_t h = 0;
    if ((int)threadIdx.x < bins) {
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w) {
            const uint32_t c = s_cnt[w * bins + threadIdx.x];
            s_cnt[w * bins + threadIdx.x] = h;
            h += c;
        }
        if (b != 0) st_relaxed(status + (size_t)b * bins + threadIdx.x, kFlagAgg | h);
    }
    uint32_t total;
    const uint32_t binstart = block_excl_scan(h, s_warp, total);  // position of digit d in the sorted chunk
    const uint32_t gh = ((int)threadIdx.x < bins) ? __ldg(hist + threadIdx.x) : 0u;
    const uint32_t gstart = block_excl_scan(gh, s_warp, total);   // global start of digit d
    if ((int)threadIdx.x < bins) {
        // decoupled look-back, kWin predecessors per round trip
        constexpr int kWin = 16;
        uint32_t excl = 0;
        int prev = (int)b - 1;
        bool done = (b == 0);
        const uint32_t* col = status + threadIdx.x;
        while (!done) {
            uint32_t sw[kWin];
#pragma unroll
            for (int j = 0; j < kWin; ++j) {
                const int idx = prev - j;
                sw[j] = idx >= 0 ? ld_relaxed(col + (size_t)idx * bins) : kFlagIncl;
            }
            bool halt = false;
#pragma unroll
            for (int j = 0; j < kWin; ++j) {
                const uint32_t f = sw[j] >> 30;
                if (!halt && !done) {
                    if (f == 0) halt = true;  // not published yet: retry from here
                    else {
                        excl += sw[j] & kValMask;
                        --prev;
                        if (f == 2) done = true;
                    }
                }
            }
        }
        st_relaxed(status + (size_t)b * bins + threadIdx.x, kFlagIncl | (excl + h));
        s_binstart[threadIdx.x] = binstart;
        s_gbase[threadIdx.x] = gstart + excl - binstart;
    }
    __syncthreads();

    // reorder through shared memory: chunk-sorted position = binstart[d] + warp offset + rank
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const size_t i = wbase + (size_t)k * 32 + lane;
        if (i < n) {
            const uint32_t d = (uint32_t)(key[k] >> shift) & mask;
            const uint32_t pos = s_binstart[d] + my[d] + rank[k];
            s_keys[pos] = key[k];
            s_vals[pos] = kIdentityVal ? (uint32_t)i : __ldg(vals_in + i);
        }
    }
    __syncthreads();
    // coalesced runs: consecutive threads hold consecutive items of the same digit
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const uint32_t i = (uint32_t)k * kBlk + threadIdx.x;
        if (i < count) {
            const K kk = s_keys[i];
            const uint32_t d = (uint32_t)(kk >> shift) & mask;
            const uint32_t dst = s_gbase[d] + i;
            keys_out[dst] = kk;
            vals_out[dst] = s_vals[i];
        }
    }
}

template <typename K>
constexpr size_t radix_pass_smem(int bits) {
    return (sizeof(K) + 4) * (size_t)kSortChunk + sizeof(uint32_t) * kSortWarps * ((size_t)1 << bits);
}

// =========================== key emission (duplicateWithKeys) =============================
// One lane per Gaussian IN DEPTH ORDER (slot k of the per-Gaussian sort); rects of up to kSmallRect
// tiles are written by the owning lane, larger ones cooperatively by the whole warp (coalesced, and
// a screen-filling splat no longer serialises thousands of stores on one thread).  Entry t of a
// rect (row-major, y outer) goes to offsets[k-1] + t.
constexpr int kSmallRect = 4;

__global__ void __launch_bounds__(256) emit_keys_kernel(int P, int gx, int gy, const uint32_t* __restrict__ order_keys,
                                                        const uint32_t* __restrict__ order,
                                                        const int32_t* __restrict__ radii,
                                                        const float2* __restrict__ means2D,
                                                        const uint32_t* __restrict__ offsets,
                                                        const gsr_forward_info* __restrict__ info, uint32_t cap,
                                                        uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (effective_n(reinterpret_cast<const uint32_t*>(&info->num_rendered), cap) == 0u) return;
    int x0 = 0, y0 = 0, w = 0, count = 0;
    uint32_t off = 0, dbits = 0, idx = 0;
    if (k < P) {
        dbits = order_keys[k];
        if (dbits != kInvalidDepthKey) {
            idx = order[k];
            const int r = radii[idx];
            const float2 m = means2D[idx];
            int x1, y1;
            tile_rect(m.x, m.y, r, gx, gy, x0, y0, x1, y1);
            w = x1 - x0;
            count = w * (y1 - y0);
            off = k == 0 ? 0u : offsets[k - 1];
        }
    }
    if (count > 0 && count <= kSmallRect) {
        for (int t = 0; t < count; ++t) {
            const int y = y0 + t / w, x = x0 + t - (t / w) * w;
            keys[off + t] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | dbits;
            vals[off + t] = idx;
        }
    }
    unsigned big = __ballot_sync(0xffffffffu, count > kSmallRect);
    while (big) {
        const int src = __ffs(big) - 1;
        big &= big - 1;
        const int bx0 = __shfl_sync(0xffffffffu, x0, src), by0 = __shfl_sync(0xffffffffu, y0, src);
        const int bw = __shfl_sync(0xffffffffu, w, src), bcount = __shfl_sync(0xffffffffu, count, src);
        const uint32_t boff = __shfl_sync(0xffffffffu, off, src), bd = __shfl_sync(0xffffffffu, dbits, src);
        const uint32_t bidx = __shfl_sync(0xffffffffu, idx, src);
        for (int t = lane_id(); t < bcount; t += 32) {
            const int yy = t / bw;
            const int y = by0 + yy, x = bx0 + t - yy * bw;
            keys[boff + t] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | bd;
            vals[boff + t] = bidx;
        }
    }
}

// =========================== tile ranges ==================================================
// Two sorted keys per thread (one 16-byte load); the predecessor of the pair comes from the
// neighbouring lane, lane 0 fetches it.
__global__ void __launch_bounds__(256) tile_ranges_kernel(const uint64_t* __restrict__ keys,
                                                          const uint32_t* __restrict__ r_ptr, uint32_t cap,
                                                          uint2* __restrict__ ranges) {
    const uint32_t n = effective_n(r_ptr, cap);
    const uint32_t i = 2u * (blockIdx.x * blockDim.x + threadIdx.x);
    uint32_t t0 = 0xffffffffu, t1 = 0xffffffffu;
    if (i + 1 < n) {
        const ulonglong2 kk = *reinterpret_cast<const ulonglong2*>(keys + i);
        t0 = (uint32_t)(kk.x >> 32);
        t1 = (uint32_t)(kk.y >> 32);
    } else if (i < n) {
        t0 = (uint32_t)(keys[i] >> 32);
    }
    uint32_t tp = __shfl_up_sync(0xffffffffu, t1, 1);
    if (lane_id() == 0 && i > 0 && i < n) tp = (uint32_t)(keys[i - 1] >> 32);
    if (i < n) {
        if (i == 0) ranges[t0].x = 0;
        else if (tp != t0) { ranges[tp].y = i; ranges[t0].x = i; }
        if (i == n - 1) ranges[t0].y = n;
    }
    if (i + 1 < n) {
        if (t0 != t1) { ranges[t0].y = i + 1; ranges[t1].x = i + 1; }
        if (i + 1 == n - 1) ranges[t1].y = n;
    }
}

int higher_msb(uint32_t n) {  // number of bits needed to represent values < n (at least 1)
    int b = 1;
    while (b < 32 && (1u << b) < n) ++b;
    return b;
}

// Tile-bit sort plan: fewest passes of at most 8 bits, evenly split.
void tile_sort_plan(int tile_bits, int& passes, int& bits) {
    passes = (tile_bits + 7) / 8;
    bits = (tile_bits + passes - 1) / passes;
    if (bits < 4) bits = 4;
}

size_t chunks_of(size_t n) { return (n + kSortChunk - 1) / kSortChunk; }

// Tile-bit passes are instantiated for 4..8-bit digits (the digit width is a compile-time constant so
// the ballot ranking unrolls without predicate spills).
template <int kBits>
void launch_tile_pass_bits(unsigned grid, size_t smem, cudaStream_t s, const uint64_t* kin, const uint32_t* vin,
                           uint64_t* kout, uint32_t* vout, const uint32_t* r_ptr, uint32_t cap, int shift,
                           const uint32_t* hist, uint32_t* status, uint32_t* ticket) {
    radix_pass_kernel<uint64_t, false, kBits><<<grid, kBlk, smem, s>>>(kin, vin, kout, vout, r_ptr, 0u, cap, shift, hist,
                                                                      status, ticket);
}
void launch_tile_pass(int bits, unsigned grid, size_t smem, cudaStream_t s, const uint64_t* kin, const uint32_t* vin,
                      uint64_t* kout, uint32_t* vout, const uint32_t* r_ptr, uint32_t cap, int shift,
                      const uint32_t* hist, uint32_t* status, uint32_t* ticket) {
    switch (bits) {
        case 4: launch_tile_pass_bits<4>(grid, smem, s, kin, vin, kout, vout, r_ptr, cap, shift, hist, status, ticket); break;
        case 5: launch_tile_pass_bits<5>(grid, smem, s, kin, vin, kout, vout, r_ptr, cap, shift, hist, status, ticket); break;
        case 6: launch_tile_pass_bits<6>(grid, smem, s, kin, vin, kout, vout, r_ptr, cap, shift, hist, status, ticket); break;
        case 7: launch_tile_pass_bits<7>(grid, smem, s, kin, vin, kout, vout, r_ptr, cap, shift, hist, status, ticket); break;
        default: launch_tile_pass_bits<8>(grid, smem, s, kin, vin, kout, vout, r_ptr, cap, shift, hist, status, ticket); break;
    }
}
cudaError_t set_tile_pass_attributes() {
    cudaError_t e = cudaSuccess;
    const auto attr = cudaFuncAttributeMaxDynamicSharedMemorySize;
    if (e == cudaSuccess) e = cudaFuncSetAttribute(radix_pass_kernel<uint32_t, true, 8>, attr, (int)radix_pass_smem<uint32_t>(8));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(radix_pass_kernel<uint32_t, false, 8>, attr, (int)radix_pass_smem<uint32_t>(8));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(radix_pass_kernel<uint64_t, false, 4>, attr, (int)radix_pass_smem<uint64_t>(4));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(radix_pass_kernel<uint64_t, false, 5>, attr, (int)radix_pass_smem<uint64_t>(5));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(radix_pass_kernel<uint64_t, false, 6>, attr, (int)radix_pass_smem<uint64_t>(6));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(radix_pass_kernel<uint64_t, false, 7>, attr, (int)radix_pass_smem<uint64_t>(7));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(radix_pass_kernel<uint64_t, false, 8>, attr, (int)radix_pass_smem<uint64_t>(8));
    return e;
}

}  // namespace

// ---- control regions (u32 words): [hist kMaxPasses*kMaxBins][tickets 8][status ...] ----------
constexpr size_t kCtlHeader = (size_t)kMaxPasses * kMaxBins + 8;

size_t depth_sort_ctl_entries(size_t P) {
    const size_t scan_status = 2 * ((P + kScanChunk - 1) / kScanChunk + 1);  // u64 words
    return kCtlHeader + 4 * chunks_of(P) * kMaxBins + scan_status + 2;
}
size_t tile_sort_ctl_entries(size_t cap) { return kCtlHeader + (size_t)kMaxPasses * chunks_of(cap) * kMaxBins; }

// Per-Gaussian stage after preprocess: depth-order sort of all P (key, index) pairs, then the prefix
// sum of tiles_touched in that order (writes info->num_rendered).
int launch_depth_order(int P, GeomView gv, cudaStream_t s, int debug) {
    static unsigned long long attr_devices = 0;
    if (first_use_on_current_device(&attr_devices)) GSR_CUDA(set_tile_pass_attributes());
    const size_t nchunk = chunks_of((size_t)P);
    uint32_t* ctl = gv.sort_ctl;
    uint32_t* hist = ctl;
    uint32_t* tickets = ctl + kMaxPasses * kMaxBins;
    uint32_t* status = ctl + kCtlHeader;
    uint32_t* scan_status = status + 4 * nchunk * kMaxBins;
    scan_status += ((reinterpret_cast<uintptr_t>(scan_status) & 7) ? 1 : 0);
    GSR_CUDA(cudaMemsetAsync(ctl, 0, sizeof(uint32_t) * depth_sort_ctl_entries((size_t)P), s));
    {
        StageScope scope(kStageDepthSort, s);
        const int hblocks = (int)min((size_t)(148 * 8), ((size_t)P + kBlk - 1) / kBlk);
        radix_hist_kernel<uint32_t><<<hblocks, kBlk, 0, s>>>(reinterpret_cast<const uint32_t*>(gv.depths), nullptr,
                                                            (uint32_t)P, 0u, 0, 8, 4, hist);
        GSR_AFTER_LAUNCH(s, debug);
        const size_t smem = radix_pass_smem<uint32_t>(8);
        // depths (identity payload) -> A -> B -> A -> B ... 4 passes end in (order_keys, order)
        const uint32_t* kin = reinterpret_cast<const uint32_t*>(gv.depths);
        const uint32_t* vin = nullptr;
        uint32_t* kout = gv.order_keys_alt;
        uint32_t* vout = gv.order_alt;
        for (int pass = 0; pass < 4; ++pass) {
            if (pass == 0)
                radix_pass_kernel<uint32_t, true, 8><<<(unsigned)nchunk, kBlk, smem, s>>>(
                    kin, vin, kout, vout, nullptr, (uint32_t)P, 0u, 8 * pass, hist + pass * kMaxBins,
                    status + (size_t)pass * nchunk * kMaxBins, tickets + pass);
            else
                radix_pass_kernel<uint32_t, false, 8><<<(unsigned)nchunk, kBlk, smem, s>>>(
                    kin, vin, kout, vout, nullptr, (uint32_t)P, 0u, 8 * pass, hist + pass * kMaxBins,
                    status + (size_t)pass * nchunk * kMaxBins, tickets + pass);
            GSR_AFTER_LAUNCH(s, debug);
            kin = kout;
            vin = vout;
            const bool to_main = (kout == gv.order_keys_alt);
            kout = to_main ? gv.order_keys : gv.order_keys_alt;
            vout = to_main ? gv.order : gv.order_alt;
        }
    }
    {
        StageScope scope(kStageScan, s);
        const unsigned blocks = (unsigned)(((size_t)P + kScanChunk - 1) / kScanChunk);
        scan_sorted_tiles_kernel<<<blocks, kBlk, 0, s>>>(P, gv.order, gv.tiles_touched, gv.point_offsets,
                                                        reinterpret_cast<unsigned long long*>(scan_status), tickets + 4,
                                                        gv.info);
        GSR_AFTER_LAUNCH(s, debug);
    }
    return GSR_OK;
}

int launch_binning(int P, const Grid& g, const int32_t* radii, GeomView gv, BinView bv, int64_t cap64,
                   ImageView iv, cudaStream_t s, int debug) {
    const uint32_t cap = (uint32_t)cap64;
    const uint32_t* r_ptr = reinterpret_cast<const uint32_t*>(&gv.info->num_rendered);
    publish_info_kernel<<<1, 1, 0, s>>>(cap, gv.info);
    GSR_AFTER_LAUNCH(s, debug);
    GSR_CUDA(cudaMemsetAsync(iv.ranges, 0, sizeof(uint2) * (size_t)g.tiles, s));
    if (cap == 0) return GSR_OK;

    {
        StageScope scope(kStageEmitKeys, s);
        emit_keys_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, g.gx, g.gy, gv.order_keys, gv.order, radii, gv.means2D,
                                                         gv.point_offsets, gv.info, cap, bv.keys_unsorted,
                                                         bv.vals_unsorted);
        GSR_AFTER_LAUNCH(s, debug);
    }

    int passes, bits;
    tile_sort_plan(higher_msb((uint32_t)g.tiles), passes, bits);
    const size_t nchunk = chunks_of((size_t)cap);
    uint32_t* ctl = bv.sort_ctl;
    uint32_t* hist = ctl;
    uint32_t* tickets = ctl + kMaxPasses * kMaxBins;
    uint32_t* status = ctl + kCtlHeader;
    GSR_CUDA(cudaMemsetAsync(ctl, 0, sizeof(uint32_t) * (kCtlHeader + (size_t)passes * nchunk * kMaxBins), s));
    {
        StageScope scope(kStageSort, s);
        const int hblocks = (int)min((size_t)(148 * 8), ((size_t)cap + kBlk - 1) / kBlk);
        radix_hist_kernel<uint64_t><<<hblocks, kBlk, 0, s>>>(bv.keys_unsorted, r_ptr, 0u, cap, 32, bits, passes, hist);
        GSR_AFTER_LAUNCH(s, debug);
        const size_t smem = radix_pass_smem<uint64_t>(bits);
        // the unsorted lists stay intact; the last pass lands in (keys, point_list)
        const uint64_t* kin = bv.keys_unsorted;
        const uint32_t* vin = bv.vals_unsorted;
        bool to_main = (passes & 1) != 0;
        for (int pass = 0; pass < passes; ++pass) {
            uint64_t* kout = to_main ? bv.keys : bv.keys_alt;
            uint32_t* vout = to_main ? bv.point_list : bv.vals_alt;
            launch_tile_pass(bits, (unsigned)nchunk, smem, s, kin, vin, kout, vout, r_ptr, cap, 32 + pass * bits,
                             hist + pass * kMaxBins, status + (size_t)pass * nchunk * kMaxBins, tickets + pass);
            GSR_AFTER_LAUNCH(s, debug);
            kin = kout;
            vin = vout;
            to_main = !to_main;
        }
    }
    StageScope scope(kStageTileRanges, s);
    tile_ranges_kernel<<<(cap + 511) / 512, 256, 0, s>>>(bv.keys, r_ptr, cap, iv.ranges);
    GSR_AFTER_LAUNCH(s, debug);
    return GSR_OK;
}

}  // namespace gsr
