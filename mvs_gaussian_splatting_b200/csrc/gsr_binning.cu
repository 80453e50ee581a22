// gsr_binning.cu — tile binning for sm_100a (the reference's K2-K6: cub::DeviceScan::InclusiveSum,
// duplicateWithKeys, cub::DeviceRadixSort::SortPairs, identifyTileRanges — SURVEY §2.1, §8a a13-a16).
//
// The reference emits one 64-bit (tile << 32 | depth bits) key per tile instance in Gaussian order
// and sorts all R of them over 32 + log2(tiles) bits.  The sorted result is fully determined by
// the (tile, depth bits, Gaussian index) order, which factorises; this file produces the SAME sorted
// point_list / ranges (and, on request, the same 64-bit key lists) with far less traffic:
//   1. one read of the depths: digit histograms of the VISIBLE Gaussians, their count and R = sum(tiles_touched);
//   2. stable LSD radix sort of the per-Gaussian (depth bits -> index) pairs, 4 x 8 bits; the first pass
//      drops the culled Gaussians, so passes 2-4 move P_visible pairs (P-sized, L2-resident);
//   3. ONE kernel for the prefix sum over the depth-ordered tile counts AND the key emission: chunk offsets by
//      decoupled look-back, entries written with a load-balanced search (every thread writes consecutive list
//      positions: coalesced), per-tile instance counts by RED.  An instance is stored as its TILE id only
//      (u16 up to 65536 tiles, else u32) + the Gaussian index: 6 bytes, not 12 — the depth bits of the
//      reference's key are a function of the index (depths[idx]) and the tile sort below never looks at them;
//   4. exclusive scan of the per-tile counts = the tile ranges (identifyTileRanges without reading a key)
//      and, folded over the digits, the histograms of the tile-bit passes (no histogram pass over the instances);
//   5. stable radix sort of the R instances on the tile bits only (2 passes up to 65536 tiles).
// The reference's 64-bit key lists are materialised from (tile, depths[idx]) by gsr_materialize_keys for the
// parity tests and debugging; nothing on the render path reads them.
// Every radix pass is ONE kernel: chunk histograms are chained with decoupled look-back (ticketed
// chunks, 2 flag bits + 30 count bits per (chunk, digit) status word), keys are ranked per warp with
// one ballot per digit bit, reordered through shared memory and written as coalesced per-digit runs.
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

// Block-wide exclusive scan of one value per thread (kThreads threads); returns the exclusive prefix and
// the block total through `total`.  s_warp: kThreads / 32 + 1 words.
constexpr int kBlk = 256;
template <int kThreads = kBlk>
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* s_warp, uint32_t& total) {
    constexpr int kW = kThreads / 32;
    const uint32_t incl = warp_incl_scan(v);
    const int warp = threadIdx.x >> 5;
    if (lane_id() == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = lane_id() < kW ? s_warp[lane_id()] : 0u;
        uint32_t wi = warp_incl_scan(w);
        if (lane_id() < kW) s_warp[lane_id()] = wi - w;
        if (lane_id() == kW - 1) s_warp[kW] = wi;
    }
    __syncthreads();
    total = s_warp[kW];
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

// Publishes the overflow flag for the host (zero-capacity path; otherwise tile_offsets_kernel does it).
__global__ void publish_info_kernel(uint32_t cap, gsr_forward_info* info) {
    info->overflow = (uint32_t)info->num_rendered > cap ? 1 : 0;
}

// =========================== stable LSD radix sort, one kernel per pass ====================
// A/B on B200 (config 3, tools/gpu_ab.sh): 256 threads x 8/12/16/20/24/32 items, 384 x 16, 512 x 8/12/16/20, 768 x 12,
// 1024 x 8, at 1-4 CTAs/SM.  The passes are latency-bound (dependent ballot / shared-atomic chains per key, three
// CTA barriers and a look-back per chunk), so twice the warps on half the keys per thread wins: 512 x 16 at 2 CTAs/SM.
#ifndef GSR_SORT_ITEMS
#define GSR_SORT_ITEMS 16
#endif
#ifndef GSR_SORT_MINBLOCKS
#define GSR_SORT_MINBLOCKS 2
#endif
#ifndef GSR_TILE_SORT_MINBLOCKS
#define GSR_TILE_SORT_MINBLOCKS 2
#endif
constexpr int kMaxBinsC = 256;
#ifndef GSR_SORT_THREADS
#define GSR_SORT_THREADS 512
#endif
constexpr int kSortThreads = GSR_SORT_THREADS;
constexpr int kSortItems = GSR_SORT_ITEMS;      // per thread
constexpr int kSortChunk = kSortThreads * kSortItems;   // 8192 items per ticket
constexpr int kSortWarps = kSortThreads / 32;
static_assert(kSortThreads >= kMaxBinsC && kSortThreads <= 1024 && kSortThreads % 32 == 0, "one thread per digit");
constexpr int kSortWarpSpan = 32 * kSortItems;  // consecutive items owned by one warp
constexpr int kMaxBins = 256;
constexpr int kMaxPasses = 4;
constexpr uint32_t kFlagAgg = 1u << 30, kFlagIncl = 2u << 30, kValMask = (1u << 30) - 1u;

// One read of the per-Gaussian depths (+ tiles_touched): digit histograms of all four passes of the depth
// sort over the VISIBLE Gaussians (culled ones carry the key 0xFFFFFFFF and are dropped by the first pass),
// their count -> info->num_visible and R = sum(tiles_touched) -> info->num_rendered.  Shared atomics, with a
// vote so that warps whose digit is constant (the high depth bytes) issue one add instead of 32 serialised ones.
__global__ void __launch_bounds__(kBlk) depth_hist_kernel(const uint32_t* __restrict__ keys,
                                                          const uint32_t* __restrict__ tiles_touched, uint32_t n,
                                                          uint32_t* __restrict__ hist, gsr_forward_info* info) {
    __shared__ uint32_t s_h[kMaxPasses][kMaxBins];
    __shared__ uint32_t s_sum[2];
    for (int i = threadIdx.x; i < kMaxPasses * kMaxBins; i += kBlk) (&s_h[0][0])[i] = 0;
    if (threadIdx.x < 2) s_sum[threadIdx.x] = 0;
    __syncthreads();
    const unsigned lane = lane_id();
    uint32_t nvis = 0, nren = 0;
    // whole warps iterate together (the uniformity vote needs converged lanes): round n up to a warp multiple
    const size_t stride = (size_t)gridDim.x * kBlk;
    const size_t n_up = ((size_t)n + 31) & ~(size_t)31;
    for (size_t i = (size_t)blockIdx.x * kBlk + threadIdx.x; i < n_up; i += stride) {
        const uint32_t key = i < n ? keys[i] : kInvalidDepthKey;
        const bool valid = key != kInvalidDepthKey;
        if (valid) { ++nvis; nren += __ldg(tiles_touched + i); }
        const bool full = __all_sync(0xffffffffu, valid);
#pragma unroll
        for (int p = 0; p < kMaxPasses; ++p) {
            const uint32_t d = (key >> (8 * p)) & 255u;
            const uint32_t d0 = __shfl_sync(0xffffffffu, d, 0);
            if (full && __all_sync(0xffffffffu, d == d0)) {
                if (lane == 0) atomicAdd(&s_h[p][d0], 32u);
            } else if (valid) {
                atomicAdd(&s_h[p][d], 1u);
            }
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        nvis += __shfl_xor_sync(0xffffffffu, nvis, d);
        nren += __shfl_xor_sync(0xffffffffu, nren, d);
    }
    if (lane == 0 && nvis) { atomicAdd(&s_sum[0], nvis); atomicAdd(&s_sum[1], nren); }
    __syncthreads();
    for (int i = threadIdx.x; i < kMaxPasses * kMaxBins; i += kBlk) {
        const uint32_t c = (&s_h[0][0])[i];
        if (c) atomicAdd(hist + i, c);
    }
    if (threadIdx.x == 0 && s_sum[0]) {
        atomicAdd(&info->num_visible, (int32_t)s_sum[0]);
        atomicAdd(&info->num_rendered, (int32_t)s_sum[1]);
    }
}

// One pass.  kFirst: first pass of the depth sort — the payload of item i is i itself (no payload array is
// read) and items whose key is 0xFFFFFFFF (culled Gaussians) are dropped, so the output holds the visible
// ones only.  status: [chunks][bins] words, zeroed before the launch; ticket: zeroed.  hist: global digit
// counts of the items this pass keeps.
template <typename K, bool kFirst, int kBits, int kMinBlocks>
__global__ void __launch_bounds__(kSortThreads, kMinBlocks) radix_pass_kernel(const K* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                                          K* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                                                          const uint32_t* __restrict__ n_ptr, uint32_t n_host, uint32_t cap,
                                                          int shift, const uint32_t* __restrict__ hist,
                                                          uint32_t* status, uint32_t* ticket) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    uint32_t* s_vals = reinterpret_cast<uint32_t*>(s_raw);                            // [kSortChunk]
    uint32_t* s_cnt = s_vals + kSortChunk;                                            // [kSortWarps][bins]
    K* s_keys = reinterpret_cast<K*>(s_cnt + kSortWarps * (1 << kBits));              // [kSortChunk]
    __shared__ uint32_t s_binstart[kMaxBins];  // exclusive start of each digit inside this chunk
    __shared__ uint32_t s_gbase[kMaxBins];     // global position of sorted-chunk item i of digit d: s_gbase[d] + i
    __shared__ uint32_t s_warp[kSortThreads / 32 + 1];
    __shared__ uint32_t s_b;

    constexpr int bins = 1 << kBits;
    const uint32_t n = n_ptr ? effective_n(n_ptr, cap) : n_host;
    if (threadIdx.x == 0) s_b = atomicAdd(ticket, 1u);
    for (int d = threadIdx.x; d < bins * kSortWarps; d += kSortThreads) s_cnt[d] = 0;
    __syncthreads();
    const uint32_t b = s_b;
    const size_t start = (size_t)b * kSortChunk;
    if (start >= n) return;

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
        key[k] = i < n ? keys_in[i] : (K)(kFirst ? kInvalidDepthKey : 0);
    }
    // Rank inside the warp: lanes holding the same digit are found with one ballot per digit bit
    // (pipelined votes; MATCH.ANY costs ~150 cycles a piece on sm_100), the lowest such lane bumps the
    // warp's private counter with one shared atomic and broadcasts the old value.
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const size_t i = wbase + (size_t)k * 32 + lane;
        const bool valid = kFirst ? (uint32_t)key[k] != kInvalidDepthKey : i < n;
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
    uint32_t count, total;
    const uint32_t binstart = block_excl_scan<kSortThreads>(h, s_warp, count);  // position of digit d in the sorted chunk; items kept
    const uint32_t gh = ((int)threadIdx.x < bins) ? __ldg(hist + threadIdx.x) : 0u;
    const uint32_t gstart = block_excl_scan<kSortThreads>(gh, s_warp, total);   // global start of digit d
    if ((int)threadIdx.x < bins) s_binstart[threadIdx.x] = binstart;
    __syncthreads();
    // reorder through shared memory: chunk-sorted position = binstart[d] + warp offset + rank
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const size_t i = wbase + (size_t)k * 32 + lane;
        const bool valid = kFirst ? (uint32_t)key[k] != kInvalidDepthKey : i < n;
        if (valid) {
            const uint32_t d = (uint32_t)(key[k] >> shift) & mask;
            const uint32_t pos = s_binstart[d] + my[d] + rank[k];
            s_keys[pos] = key[k];
            s_vals[pos] = kFirst ? (uint32_t)i : __ldg(vals_in + i);
        }
    }
    if ((int)threadIdx.x < bins) {
        // decoupled look-back, kWin predecessors per round trip.  It runs AFTER the shared-memory reorder above has been
        // issued (measured both ways on the B200: depth sort 0.171 -> 0.134 ms with the reorder first — the stores
        // cover the look-back's round trips; the other order was dropped)
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
        s_gbase[threadIdx.x] = gstart + excl - binstart;
    }
    __syncthreads();
    // coalesced runs: consecutive threads hold consecutive items of the same digit
#pragma unroll
    for (int k = 0; k < kSortItems; ++k) {
        const uint32_t i = (uint32_t)k * kSortThreads + threadIdx.x;
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

// =========================== prefix sum + key emission (duplicateWithKeys) ==================
// One chunk = kEmitSlots consecutive slots of the depth order per thread (one 16-byte load of `order`).  Per slot:
// the splat's tile rect from (radii, means2D) — its area IS tiles_touched.  The chunk's first list position comes
// from a decoupled look-back over the chunk totals (the reference's InclusiveSum, never stored): the chunk
// publishes its total as soon as it is known and only asks for its prefix AFTER it has expanded its rects into a
// shared staging window — the expansion needs no global position, so the predecessors' totals have landed by then.
// Entry t of a rect (row-major, y outer) is tile (y0 + t / w) * gx + x0 + t % w; rects of up to kEmitSmall tiles are
// expanded by the owning lane, larger ones by the whole warp (a screen-filling splat no longer serialises thousands
// of stores on one thread).  The staged window then goes out with coalesced stores, one RED per entry counting the
// instances of its tile.
constexpr int kEmitSlots = 4;
constexpr int kEmitChunk = kBlk * kEmitSlots;
#ifndef GSR_EMIT_SMALL
#define GSR_EMIT_SMALL 12
#endif
constexpr int kEmitSmall = GSR_EMIT_SMALL;
// The per-tile instance counters are kept in kTileReplicas (8) copies (chunk b adds to copy b mod kTileReplicas) and
// summed by tile_offsets_kernel: 8 M REDs into the ~250 cache lines of one copy saturate the L2 slices that hold the
// busy screen-centre tiles (ncu: hottest slice at 88 % while the average sits at 28 %).
#ifndef GSR_TILE_REPLICAS
#define GSR_TILE_REPLICAS 8   // A/B: 1 copy 0.124 ms emission; 8 / 16 / 32 copies 0.086 ms, tile_offsets 0.015 / 0.019 / 0.023 ms
#endif
constexpr int kTileReplicas = GSR_TILE_REPLICAS;

__device__ __forceinline__ void publish_aggregate(unsigned long long* status, uint32_t b, uint32_t total) {
    st_relaxed64(status + b, ((b == 0 ? 2ull : 1ull) << 32) | total);
}
// Exclusive prefix of chunk b (warp 0, after publish_aggregate); upgrades the chunk's word to an inclusive prefix.
__device__ __forceinline__ uint32_t lookback_prefix(unsigned long long* status, uint32_t b, uint32_t total) {
    if (b == 0) return 0;
    const unsigned lane = lane_id();
    uint32_t prefix = 0;
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
    return prefix;
}

// Writes the entries [t0, t1) of one rect into the staging window at s_* + pos0 (pos0 = slot of entry t0).
template <typename TK>
__device__ __forceinline__ void stage_run(TK* s_tile, uint32_t* s_val, uint32_t pos0, uint32_t t0, uint32_t t1,
                                          uint32_t step, uint32_t x0, uint32_t y0, uint32_t w, uint32_t gx, uint32_t idx) {
    for (uint32_t t = t0; t < t1; t += step) {
        const uint32_t yy = t / w;
        s_tile[pos0 + (t - t0)] = (TK)((y0 + yy) * gx + x0 + (t - yy * w));
        s_val[pos0 + (t - t0)] = idx;
    }
}

#ifndef GSR_EMIT_MINBLOCKS
#define GSR_EMIT_MINBLOCKS 1
#endif
template <typename TK>
__global__ void __launch_bounds__(kBlk, GSR_EMIT_MINBLOCKS) emit_scan_kernel(int gx, int gy, const uint32_t* __restrict__ order,
                                                         const int32_t* __restrict__ radii,
                                                         const float2* __restrict__ means2D,
                                                         const gsr_forward_info* __restrict__ info, uint32_t cap,
                                                         unsigned long long* status, uint32_t* ticket,
                                                         TK* __restrict__ keys, uint32_t* __restrict__ vals,
                                                         uint32_t* __restrict__ tile_count, uint32_t replica_stride) {
    constexpr int kStage = sizeof(TK) == 2 ? 6144 : 4096;   // staged entries: 36 KB / 32 KB of shared memory
    __shared__ uint32_t s_val[kStage];
    __shared__ TK s_tile[kStage];
    __shared__ uint32_t s_warp[kBlk / 32 + 1];
    __shared__ uint32_t s_b, s_prefix;
    if (effective_n(reinterpret_cast<const uint32_t*>(&info->num_rendered), cap) == 0u) return;  // empty or overflow
    const uint32_t nvis = (uint32_t)__ldg(&info->num_visible);
    if (threadIdx.x == 0) s_b = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t b = s_b;
    const uint32_t k0 = b * kEmitChunk + threadIdx.x * kEmitSlots;
    if (b * kEmitChunk >= nvis) return;
    const unsigned lane = lane_id();

    uint32_t id[kEmitSlots];
    if (k0 + kEmitSlots <= nvis) {
        const uint4 a = *reinterpret_cast<const uint4*>(order + k0);
        id[0] = a.x; id[1] = a.y; id[2] = a.z; id[3] = a.w;
    } else {
#pragma unroll
        for (int j = 0; j < kEmitSlots; ++j) id[j] = k0 + j < nvis ? order[k0 + j] : 0xFFFFFFFFu;
    }
    int rr[kEmitSlots];
    float2 mm[kEmitSlots];
#pragma unroll
    for (int j = 0; j < kEmitSlots; ++j) {
        const bool on = id[j] != 0xFFFFFFFFu;
        rr[j] = on ? __ldg(radii + id[j]) : 0;
        mm[j] = on ? __ldg(means2D + id[j]) : make_float2(0.f, 0.f);
    }
    uint32_t xy[kEmitSlots], ww[kEmitSlots], cnt[kEmitSlots], ex[kEmitSlots];
    uint32_t sum = 0;
#pragma unroll
    for (int j = 0; j < kEmitSlots; ++j) {
        int x0 = 0, y0 = 0, x1 = 0, y1 = 0;
        if (rr[j] > 0) tile_rect(mm[j].x, mm[j].y, rr[j], gx, gy, x0, y0, x1, y1);
        const uint32_t w = (uint32_t)(x1 - x0);
        cnt[j] = w * (uint32_t)(y1 - y0);
        xy[j] = (uint32_t)x0 | ((uint32_t)y0 << 16);
        ww[j] = w ? w : 1u;
        sum += cnt[j];
    }
    uint32_t total;
    uint32_t run = block_excl_scan(sum, s_warp, total);
    if (threadIdx.x == 0) publish_aggregate(status, b, total);   // successors can add it up from now on
#pragma unroll
    for (int j = 0; j < kEmitSlots; ++j) { ex[j] = run; run += cnt[j]; }

    for (uint32_t w0 = 0; w0 < total; w0 += kStage) {   // one window for all but pathological chunks
        const uint32_t w1 = min(w0 + (uint32_t)kStage, total);
        if (w0) __syncthreads();   // the previous window has been copied out
#pragma unroll
        for (int j = 0; j < kEmitSlots; ++j) {
            // the part [t0, t1) of this slot's entries that falls into the window
            const uint32_t lo = max(ex[j], w0), hi = min(ex[j] + cnt[j], w1);
            const bool any = lo < hi;
            const bool small = any && cnt[j] <= (uint32_t)kEmitSmall;
            if (small)
                stage_run(s_tile, s_val, lo - w0, lo - ex[j], hi - ex[j], 1u, xy[j] & 0xFFFFu, xy[j] >> 16, ww[j],
                          (uint32_t)gx, id[j]);
            unsigned big = __ballot_sync(0xffffffffu, any && !small);
            while (big) {
                const int src = __ffs(big) - 1;
                big &= big - 1;
                const uint32_t bxy = __shfl_sync(0xffffffffu, xy[j], src), bw = __shfl_sync(0xffffffffu, ww[j], src);
                const uint32_t bid = __shfl_sync(0xffffffffu, id[j], src);
                const uint32_t bex = __shfl_sync(0xffffffffu, ex[j], src);
                const uint32_t blo = __shfl_sync(0xffffffffu, lo, src), bhi = __shfl_sync(0xffffffffu, hi, src);
                stage_run(s_tile, s_val, blo - w0 + lane, blo - bex + lane, bhi - bex, 32u, bxy & 0xFFFFu, bxy >> 16, bw,
                          (uint32_t)gx, bid);
            }
        }
        if (w0 == 0 && threadIdx.x < 32) {
            const uint32_t prefix = lookback_prefix(status, b, total);
            if (threadIdx.x == 0) s_prefix = prefix;
        }
        __syncthreads();
        const uint32_t goff = s_prefix + w0;
        uint32_t* my_count = tile_count + (size_t)(b & (kTileReplicas - 1)) * replica_stride;
        for (uint32_t i = threadIdx.x; i < w1 - w0; i += kBlk) {
            const TK tile = s_tile[i];
            keys[goff + i] = tile;
            vals[goff + i] = s_val[i];
            atomicAdd(my_count + (uint32_t)tile, 1u);
        }
    }
}

// =========================== tile ranges + tile-pass histograms =============================
// Exclusive scan of the per-tile instance counts: ranges[t] = [start, end) into the sorted list (what the
// reference's identifyTileRanges finds by comparing neighbouring keys), and the digit histograms of the tile-bit
// radix passes folded out of the same counts.  One CTA (tiles is a few thousand to a few hundred thousand).
constexpr int kOffsThreads = 1024;
constexpr int kOffsPer = 8;   // rows of 1024 tiles per batch (one batch covers 1080p; 4K takes four)
__global__ void __launch_bounds__(kOffsThreads) tile_offsets_kernel(int tiles, const uint32_t* __restrict__ tile_count,
                                                                    uint32_t replica_stride, uint2* __restrict__ ranges,
                                                                    int npass, int bits, uint32_t* __restrict__ hist,
                                                                    uint32_t cap, gsr_forward_info* info) {
    __shared__ uint32_t s_h[kMaxPasses][kMaxBins];
    __shared__ uint32_t s_wsum[kOffsThreads / 32 + 1];
    for (int i = threadIdx.x; i < kMaxPasses * kMaxBins; i += kOffsThreads) (&s_h[0][0])[i] = 0;
    if (threadIdx.x == 0) info->overflow = (uint32_t)info->num_rendered > cap ? 1 : 0;
    __syncthreads();
    const uint32_t mask = (1u << bits) - 1u;
    const unsigned lane = lane_id();
    const int warp = threadIdx.x >> 5;
    uint32_t carry = 0;  // instances of the tiles before the current row (kept identical in every thread)
    for (int base = 0; base < tiles; base += kOffsThreads * kOffsPer) {
        // row k of the batch = tiles base + k * 1024 + thread: coalesced, and every load of the batch is independent
        // of the others, so the whole batch costs one memory round trip
        uint32_t c[kOffsPer];
#pragma unroll
        for (int k = 0; k < kOffsPer; ++k) c[k] = 0;
#pragma unroll
        for (int r = 0; r < kTileReplicas; ++r) {
#pragma unroll
            for (int k = 0; k < kOffsPer; ++k) {
                const int t = base + k * kOffsThreads + (int)threadIdx.x;
                if (t < tiles) c[k] += tile_count[(size_t)r * replica_stride + t];
            }
        }
#pragma unroll
        for (int k = 0; k < kOffsPer; ++k) {
            const int t = base + k * kOffsThreads + (int)threadIdx.x;
            if (base + k * kOffsThreads >= tiles) break;   // uniform
            const uint32_t incl = warp_incl_scan(c[k]);
            if (lane == 31) s_wsum[warp] = incl;
            __syncthreads();
            if (warp == 0) {
                const uint32_t w = s_wsum[lane];
                const uint32_t wi = warp_incl_scan(w);
                s_wsum[lane] = wi - w;
                if (lane == 31) s_wsum[kOffsThreads / 32] = wi;
            }
            __syncthreads();
            const uint32_t begin = carry + s_wsum[warp] + incl - c[k];
            carry += s_wsum[kOffsThreads / 32];
            if (t < tiles)   // empty tiles: (0, 0) as the reference leaves them
                ranges[t] = c[k] ? make_uint2(begin, begin + c[k]) : make_uint2(0u, 0u);
            // digit histograms: a warp holds 32 consecutive tiles, so every digit above the lowest is the same for
            // all its lanes — one add of the warp's sum instead of 32 serialised adds to one shared word
            const uint32_t wsum = __reduce_add_sync(0xffffffffu, c[k]);
            for (int p = 0; p < npass; ++p) {
                const uint32_t d = ((uint32_t)t >> (p * bits)) & mask;
                const uint32_t d0 = __shfl_sync(0xffffffffu, d, 0);
                if (__all_sync(0xffffffffu, d == d0)) {
                    if (lane == 0 && wsum) atomicAdd(&s_h[p][d0], wsum);
                } else if (c[k]) {
                    atomicAdd(&s_h[p][d], c[k]);
                }
            }
            __syncthreads();
        }
    }
    for (int i = threadIdx.x; i < npass * kMaxBins; i += kOffsThreads) hist[i] = (&s_h[0][0])[i];
}

// The reference's 64-bit key of list entry i: (tile << 32) | depth bits of its Gaussian (tests / debugging).
template <typename TK>
__global__ void __launch_bounds__(256) materialize_keys_kernel(const TK* __restrict__ tile_keys,
                                                               const uint32_t* __restrict__ vals,
                                                               const uint32_t* __restrict__ depth_bits,
                                                               const uint32_t* __restrict__ r_ptr, uint32_t cap,
                                                               uint64_t* __restrict__ out) {
    const uint32_t n = effective_n(r_ptr, cap);
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = ((uint64_t)tile_keys[i] << 32) | __ldg(depth_bits + vals[i]);
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
size_t emit_chunks_of(size_t n) { return (n + kEmitChunk - 1) / kEmitChunk; }

constexpr int kDepthMin = GSR_SORT_MINBLOCKS, kTileMin = GSR_TILE_SORT_MINBLOCKS;

// Tile-bit passes are instantiated for 4..8-bit digits (the digit width is a compile-time constant so
// the ballot ranking unrolls without predicate spills) and for 16- / 32-bit tile keys.
template <typename TK, int kBits>
cudaError_t tile_pass_bits(unsigned grid, cudaStream_t s, const TK* kin, const uint32_t* vin, TK* kout, uint32_t* vout,
                           const uint32_t* r_ptr, uint32_t cap, int shift, const uint32_t* hist, uint32_t* status,
                           uint32_t* ticket, bool attr_only) {
    auto* fn = radix_pass_kernel<TK, false, kBits, kTileMin>;
    const size_t smem = radix_pass_smem<TK>(kBits);
    if (attr_only) return cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    fn<<<grid, kSortThreads, smem, s>>>(kin, vin, kout, vout, r_ptr, 0u, cap, shift, hist, status, ticket);
    return cudaSuccess;
}
template <typename TK>
cudaError_t tile_pass(int bits, unsigned grid, cudaStream_t s, const TK* kin, const uint32_t* vin, TK* kout,
                      uint32_t* vout, const uint32_t* r_ptr, uint32_t cap, int shift, const uint32_t* hist,
                      uint32_t* status, uint32_t* ticket, bool attr_only = false) {
    switch (bits) {
        case 4: return tile_pass_bits<TK, 4>(grid, s, kin, vin, kout, vout, r_ptr, cap, shift, hist, status, ticket, attr_only);
        case 5: return tile_pass_bits<TK, 5>(grid, s, kin, vin, kout, vout, r_ptr, cap, shift, hist, status, ticket, attr_only);
        case 6: return tile_pass_bits<TK, 6>(grid, s, kin, vin, kout, vout, r_ptr, cap, shift, hist, status, ticket, attr_only);
        case 7: return tile_pass_bits<TK, 7>(grid, s, kin, vin, kout, vout, r_ptr, cap, shift, hist, status, ticket, attr_only);
        default: return tile_pass_bits<TK, 8>(grid, s, kin, vin, kout, vout, r_ptr, cap, shift, hist, status, ticket, attr_only);
    }
}
cudaError_t set_pass_attributes() {
    cudaError_t e = cudaSuccess;
    const auto attr = cudaFuncAttributeMaxDynamicSharedMemorySize;
    if (e == cudaSuccess) e = cudaFuncSetAttribute(radix_pass_kernel<uint32_t, true, 8, kDepthMin>, attr, (int)radix_pass_smem<uint32_t>(8));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(radix_pass_kernel<uint32_t, false, 8, kDepthMin>, attr, (int)radix_pass_smem<uint32_t>(8));
    for (int bits = 4; bits <= 8 && e == cudaSuccess; ++bits) {
        e = tile_pass<uint16_t>(bits, 0, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, 0, nullptr, nullptr, nullptr, true);
        if (e == cudaSuccess)
            e = tile_pass<uint32_t>(bits, 0, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, 0, nullptr, nullptr, nullptr, true);
    }
    return e;
}
cudaError_t ensure_pass_attributes() {
    static unsigned long long attr_devices = 0;
    return once_per_device(&attr_devices, set_pass_attributes);
}

// ---- control regions (u32 words) --------------------------------------------------------------------
// geometry:  [hist kMaxPasses*kMaxBins][tickets 8][pass status 4 x chunks(P) x kMaxBins]
// binning:   [hist kMaxPasses*kMaxBins][tickets 8][tile_count tiles, padded][emit status 2 x emit_chunks(P)]
//            [pass status passes x chunks(cap) x kMaxBins]      (tile_count: kTileReplicas copies)
constexpr size_t kCtlHeader = (size_t)kMaxPasses * kMaxBins + 8;
size_t pad4(size_t n) { return (n + 3) & ~(size_t)3; }

struct TileCtl {
    uint32_t *hist, *tickets, *tile_count, *status;
    unsigned long long* emit_status;
    size_t head_words;  // everything before the pass status words
};
TileCtl tile_ctl(uint32_t* ctl, size_t P, size_t tiles) {
    TileCtl c;
    c.hist = ctl;
    c.tickets = ctl + kMaxPasses * kMaxBins;
    c.tile_count = ctl + kCtlHeader;
    size_t o = kCtlHeader + kTileReplicas * pad4(tiles);
    c.emit_status = reinterpret_cast<unsigned long long*>(ctl + o);
    o += 2 * (emit_chunks_of(P) + 1);
    c.status = ctl + o;
    c.head_words = o;
    return c;
}

template <typename TK>
int binning_impl(int P, const Grid& g, const int32_t* radii, GeomView gv, BinView bv, uint32_t cap, ImageView iv,
                 cudaStream_t s, int debug) {
    const uint32_t* r_ptr = reinterpret_cast<const uint32_t*>(&gv.info->num_rendered);
    int passes, bits;
    tile_sort_plan(higher_msb((uint32_t)g.tiles), passes, bits);
    const size_t nchunk = chunks_of((size_t)cap);
    const TileCtl c = tile_ctl(bv.sort_ctl, (size_t)P, (size_t)g.tiles);
    GSR_CUDA(cudaMemsetAsync(bv.sort_ctl, 0, sizeof(uint32_t) * (c.head_words + (size_t)passes * nchunk * kMaxBins), s));
    TK* k_unsorted = reinterpret_cast<TK*>(bv.keys_unsorted);
    TK* k_main = reinterpret_cast<TK*>(bv.keys);
    TK* k_alt = reinterpret_cast<TK*>(bv.keys_alt);
    {
        StageScope scope(kStageEmitKeys, s);
        emit_scan_kernel<TK><<<(unsigned)emit_chunks_of((size_t)P), kBlk, 0, s>>>(
            g.gx, g.gy, gv.order, radii, gv.means2D, gv.info, cap, c.emit_status, c.tickets + 7, k_unsorted,
            bv.vals_unsorted, c.tile_count, (uint32_t)pad4((size_t)g.tiles));
        GSR_AFTER_LAUNCH(s, debug);
    }
    {
        StageScope scope(kStageTileRanges, s);
        tile_offsets_kernel<<<1, kOffsThreads, 0, s>>>(g.tiles, c.tile_count, (uint32_t)pad4((size_t)g.tiles), iv.ranges, passes, bits, c.hist, cap, gv.info);
        GSR_AFTER_LAUNCH(s, debug);
    }
    StageScope scope(kStageSort, s);
    // the unsorted lists stay intact; the last pass lands in (keys, point_list)
    const TK* kin = k_unsorted;
    const uint32_t* vin = bv.vals_unsorted;
    bool to_main = (passes & 1) != 0;
    for (int pass = 0; pass < passes; ++pass) {
        TK* kout = to_main ? k_main : k_alt;
        uint32_t* vout = to_main ? bv.point_list : bv.vals_alt;
        GSR_CUDA(tile_pass<TK>(bits, (unsigned)nchunk, s, kin, vin, kout, vout, r_ptr, cap, pass * bits,
                               c.hist + pass * kMaxBins, c.status + (size_t)pass * nchunk * kMaxBins, c.tickets + pass));
        GSR_AFTER_LAUNCH(s, debug);
        kin = kout;
        vin = vout;
        to_main = !to_main;
    }
    return GSR_OK;
}

}  // namespace

size_t depth_sort_ctl_entries(size_t P) { return kCtlHeader + 4 * chunks_of(P) * kMaxBins; }
size_t tile_sort_ctl_entries(size_t cap, size_t P, size_t tiles) {
    return kCtlHeader + kTileReplicas * pad4(tiles) + 2 * (emit_chunks_of(P) + 1) + (size_t)kMaxPasses * chunks_of(cap) * kMaxBins;
}
int tile_key_bytes(int tiles) { return tiles <= 65536 ? 2 : 4; }

// Per-Gaussian stage after preprocess: counts (num_visible, num_rendered) and the depth-order sort of the
// visible (key, index) pairs; (order_keys, order)[0 .. num_visible) on return.
int launch_depth_ctl_reset(int P, GeomView gv, cudaStream_t s) {
    GSR_CUDA(cudaMemsetAsync(gv.sort_ctl, 0, sizeof(uint32_t) * depth_sort_ctl_entries((size_t)P), s));
    return GSR_OK;
}

int launch_depth_order(int P, GeomView gv, cudaStream_t s, int debug, bool counts_done) {
    GSR_CUDA(ensure_pass_attributes());
    const size_t nchunk = chunks_of((size_t)P);
    uint32_t* ctl = gv.sort_ctl;
    uint32_t* hist = ctl;
    uint32_t* tickets = ctl + kMaxPasses * kMaxBins;
    uint32_t* status = ctl + kCtlHeader;
    StageScope scope(kStageDepthSort, s);
    if (!counts_done) {
        const int hblocks = (int)min((size_t)sm_count() * 8, ((size_t)P + kBlk - 1) / kBlk);
        depth_hist_kernel<<<hblocks, kBlk, 0, s>>>(reinterpret_cast<const uint32_t*>(gv.depths), gv.tiles_touched,
                                                   (uint32_t)P, hist, gv.info);
        GSR_AFTER_LAUNCH(s, debug);
    }
    const size_t smem = radix_pass_smem<uint32_t>(8);
    const uint32_t* nvis_ptr = reinterpret_cast<const uint32_t*>(&gv.info->num_visible);
    // depths (identity payload, culled dropped) -> A -> B -> A -> B: 4 passes end in (order_keys, order)
    const uint32_t* kin = reinterpret_cast<const uint32_t*>(gv.depths);
    const uint32_t* vin = nullptr;
    uint32_t* kout = gv.order_keys_alt;
    uint32_t* vout = gv.order_alt;
    for (int pass = 0; pass < 4; ++pass) {
        if (pass == 0)
            radix_pass_kernel<uint32_t, true, 8, kDepthMin><<<(unsigned)nchunk, kSortThreads, smem, s>>>(
                kin, vin, kout, vout, nullptr, (uint32_t)P, 0u, 0, hist, status, tickets);
        else
            radix_pass_kernel<uint32_t, false, 8, kDepthMin><<<(unsigned)nchunk, kSortThreads, smem, s>>>(
                kin, vin, kout, vout, nvis_ptr, 0u, (uint32_t)P, 8 * pass, hist + pass * kMaxBins,
                status + (size_t)pass * nchunk * kMaxBins, tickets + pass);
        GSR_AFTER_LAUNCH(s, debug);
        kin = kout;
        vin = vout;
        const bool to_main = (kout == gv.order_keys_alt);
        kout = to_main ? gv.order_keys : gv.order_keys_alt;
        vout = to_main ? gv.order : gv.order_alt;
    }
    return GSR_OK;
}

int launch_binning(int P, const Grid& g, const int32_t* radii, GeomView gv, BinView bv, int64_t cap64,
                   ImageView iv, cudaStream_t s, int debug) {
    const uint32_t cap = (uint32_t)cap64;
    if (cap == 0) {
        publish_info_kernel<<<1, 1, 0, s>>>(cap, gv.info);
        GSR_AFTER_LAUNCH(s, debug);
        GSR_CUDA(cudaMemsetAsync(iv.ranges, 0, sizeof(uint2) * (size_t)g.tiles, s));
        return GSR_OK;
    }
    GSR_CUDA(ensure_pass_attributes());
    return tile_key_bytes(g.tiles) == 2 ? binning_impl<uint16_t>(P, g, radii, gv, bv, cap, iv, s, debug)
                                        : binning_impl<uint32_t>(P, g, radii, gv, bv, cap, iv, s, debug);
}

int launch_materialize_keys(int P, const Grid& g, GeomView gv, BinView bv, int64_t cap64, int sorted, uint64_t* out,
                            cudaStream_t s) {
    (void)P;
    const uint32_t cap = (uint32_t)cap64;
    if (cap == 0) return GSR_OK;
    const uint32_t* r_ptr = reinterpret_cast<const uint32_t*>(&gv.info->num_rendered);
    const uint32_t* dbits = reinterpret_cast<const uint32_t*>(gv.depths);
    const void* tk = sorted ? (const void*)bv.keys : (const void*)bv.keys_unsorted;
    const uint32_t* vals = sorted ? bv.point_list : bv.vals_unsorted;
    const unsigned grid = (cap + 255) / 256;
    if (tile_key_bytes(g.tiles) == 2)
        materialize_keys_kernel<uint16_t><<<grid, 256, 0, s>>>(static_cast<const uint16_t*>(tk), vals, dbits, r_ptr, cap, out);
    else
        materialize_keys_kernel<uint32_t><<<grid, 256, 0, s>>>(static_cast<const uint32_t*>(tk), vals, dbits, r_ptr, cap, out);
    GSR_AFTER_LAUNCH(s, 0);
    return GSR_OK;
}

}  // namespace gsr
