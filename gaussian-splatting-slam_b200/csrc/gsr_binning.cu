// gsr_binning.cu -- depth ordering of Gaussians, tile-count scan, instance emission (balanced by output), tile table
// (ranges + digit histograms), tile sort. Replaces rasterizer_impl.cu:70-138 (duplicateWithKeys, identifyTileRanges)
// and the CUB calls at :277 (InclusiveSum) and :303 (SortPairs). HBM-bound integer work.
//
// Order of the final point list (must equal the reference's stable 64-bit sort):
//   by tile id, then by depth bits, then by Gaussian id.
// We get it as: (1) stable sort of the VISIBLE Gaussians' ids by depth bits (culled ones carry the key 0xFFFFFFFF and
//                   are dropped by the sort's first pass, so 3 of the 4 passes, the scan and the emission run over
//                   the visible count only -- half of P on the bench scene; that count stays on the device);
//               (2) emit instances walking the Gaussians in that order (offsets from a scan taken in
//                   the same order);  (3) stable sort of instances by tile id only.
// The tile ranges (identifyTileRanges) and the digit histograms of (3) need no pass over the instances: every
// rectangle adds +1 / -1 at its four corners of a difference array, whose 2-D prefix sum is the list length per tile.
#include "gsr_common.cuh"
#include <algorithm>

namespace gsr {

// ------------------------------------------------------------------ scan + emission (A6 + A7)
constexpr int kScanThreads = 256;
constexpr int kScanItems = GSR_SCAN_TILE / kScanThreads;   // 8, blocked per thread
static_assert(kScanItems % 4 == 0, "scan_offsets_kernel loads its ids with 16-byte requests");

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* warp_sums, uint32_t& total) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
    }
    if (lane == 31) warp_sums[w] = incl;
    __syncthreads();
    uint32_t woff = 0, tot = 0;
#pragma unroll
    for (int k = 0; k < kScanThreads / 32; k++) {
        const uint32_t s = warp_sums[k];
        if (k < w) woff += s;
        tot += s;
    }
    total = tot;
    __syncthreads();
    return woff + incl - v;
}

__device__ __forceinline__ void rect_of(float px, float py, int max_radius, int gx, int gy, int& x0, int& y0,
                                        int& x1, int& y1) {
    // auxiliary.h:46-56 (same float expressions as the forward preprocess)
    x0 = min(gx, max((int)0, (int)((px - max_radius) / GSR_TILE)));
    y0 = min(gy, max((int)0, (int)((py - max_radius) / GSR_TILE)));
    x1 = min(gx, max((int)0, (int)((px + max_radius + GSR_TILE - 1) / GSR_TILE)));
    y1 = min(gy, max((int)0, (int)((py + max_radius + GSR_TILE - 1) / GSR_TILE)));
}

// Replaces cub::DeviceScan::InclusiveSum (rasterizer_impl.cu:277): single-pass chained scan (decoupled look-back) of
// the tile counts in DEPTH order. CTA b takes the b-th GSR_SCAN_TILE visible Gaussians, gathers their packed rectangles through
// `order` (L2 hits: 4 B x P), scans locally, publishes its sum and looks back with ALL its threads (1024 predecessors per round trip). A tile's
// status word carries flag (bits 30-31: 1 = the tile's own sum, 2 = inclusive prefix) and value (30 bits, saturating)
// together, so one 32-bit store publishes both. status[0..m) = tile words (all inclusive prefixes once the kernel has
// finished: the emission uses them as the first level of its search), status[m_slots] = grand total,
// status[m_slots + 2] = the tile table's arrival counter; all zeroed before the launch. The rectangles are also written out in depth order.
constexpr uint32_t kScanOwn = 1u << 30, kScanIncl = 2u << 30, kScanVal = (1u << 30) - 1u;
template <bool PACKED>
__global__ void __launch_bounds__(kScanThreads) scan_offsets_kernel(const uint32_t* __restrict__ n_dev,
                                                                     const uint32_t* __restrict__ order,
                                                                     const uint32_t* __restrict__ counts, int m_slots,
                                                                     uint32_t* __restrict__ status,
                                                                     uint32_t* __restrict__ point_offsets,
                                                                     uint32_t* __restrict__ rect_sorted,
                                                                     uint32_t* __restrict__ host_total) {
    // n (visible Gaussians, in depth order) lives on the device; the grid covers the host's upper bound and m_slots only
    // fixes where the total and the ticket sit. CTAs whose ticket lies beyond the last tile leave at once.
    __shared__ uint32_t warp_sums[kScanThreads / 32];
    __shared__ unsigned long long s_part[kScanThreads / 32];
    __shared__ int s_found[kScanThreads / 32];
    // CTA b scans tile b: it only ever waits for lower-numbered tiles, whose CTAs were dispatched before it. (A ticket
    // per tile costs a same-address atomic, and those serialise in L2 at ~20 ns each: 2930 CTAs -> the whole kernel.)
    const int tile = (int)blockIdx.x;
    const int n = (int)__ldg(n_dev);
    const int m = (n + GSR_SCAN_TILE - 1) / GSR_SCAN_TILE;
    if (tile >= m) {
        // nothing visible at all: the count (0) still has to reach the host, which polls host_total (mapped pinned memory)
        if (m == 0 && tile == 0 && threadIdx.x == 0 && host_total != nullptr) { *host_total = 0u; __threadfence_system(); }
        return;
    }
    const int base = tile * GSR_SCAN_TILE + threadIdx.x * kScanItems;
    uint32_t g[kScanItems], c[kScanItems], v = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; i++) g[i] = 0u;
    if (base + kScanItems <= n) {
#pragma unroll
        for (int q = 0; q < kScanItems / 4; q++) {      // base is a multiple of 4, order 128-byte aligned
            const uint4 o4 = __ldg(reinterpret_cast<const uint4*>(order + base) + q);
            g[4 * q] = o4.x; g[4 * q + 1] = o4.y; g[4 * q + 2] = o4.z; g[4 * q + 3] = o4.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < kScanItems; i++)
            if (base + i < n) g[i] = __ldg(order + base + i);
    }
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        c[i] = (base + i < n) ? __ldg(counts + g[i]) : 0u;
        if (PACKED) {
            if (base + i < n) rect_sorted[base + i] = c[i];
            c[i] = ((c[i] >> 16) & 255u) * (c[i] >> 24);
        }
        v += c[i];
    }
    uint32_t total;
    const uint32_t ex = block_exclusive_scan(v, warp_sums, total);
    volatile uint32_t* vs = status;
    if (threadIdx.x == 0) vs[tile] = (tile == 0 ? kScanIncl : kScanOwn) | total;
    // look-back with the whole CTA, 1024 predecessors per round trip: warp w reads tiles j - 128w - 4*lane - q (nearest
    // first). Saturating sums: a status word has 30 value bits, so a running prefix that reaches kScanVal = 2^30 - 1 stays
    // there (sums are formed in 64 bits and clamped) instead of spilling into the flag bits. A total equal to kScanVal
    // therefore means "num_rendered >= 2^30 - 1": gsr_forward answers GSR_E_TOOLARGE, and every consumer of the offsets
    // clamps its writes to the capacity of the binning blob, so nothing is written out of bounds.
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    unsigned long long prefix = 0;
    for (int j = tile - 1; j >= 0; j -= 4 * kScanThreads) {
        unsigned long long part = 0;                       // sum of this lane's words up to (and including) its first inclusive one
        bool has_incl = false;
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const int idx = j - 128 * w - 4 * lane - q;
            uint32_t st = kScanIncl;                       // before the first tile: inclusive prefix 0
            if (idx >= 0) {
                st = vs[idx];
                while ((st >> 30) == 0u) { __nanosleep(20); st = vs[idx]; }
            }
            if (!has_incl) part += st & kScanVal;
            has_incl = has_incl || (st >> 30) == 2u;
        }
        const unsigned incl = __ballot_sync(0xffffffffu, has_incl);
        const int first = incl ? __ffs(incl) - 1 : 31;
        unsigned long long val = lane <= first ? part : 0ull;          // <= 4 * 2^30 per lane, 2^37 per warp
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
        if (lane == 0) { s_part[w] = val; s_found[w] = incl != 0u; }
        __syncthreads();
        bool done = false;
#pragma unroll
        for (int k = 0; k < kScanThreads / 32; k++) {
            if (!done) prefix += s_part[k];
            done = done || s_found[k] != 0;
        }
        __syncthreads();
        if (done) break;
    }
    if (threadIdx.x == 0) {
        const unsigned long long incl_sum = prefix + total;
        const uint32_t sat = incl_sum >= kScanVal ? kScanVal : (uint32_t)incl_sum;
        if (tile > 0) vs[tile] = kScanIncl | sat;
        if (tile == m - 1) {
            status[m_slots] = sat;
            // the count goes home from here: a store into mapped pinned memory, no copy operation behind the sort passes
            if (host_total != nullptr) { *host_total = sat; __threadfence_system(); }
        }
    }
    uint32_t run = ex + (prefix >= kScanVal ? kScanVal : (uint32_t)prefix);
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        run += c[i];
        if (base + i < n) point_offsets[base + i] = run;   // inclusive, as rasterizer_impl.cu:277 (in depth order)
    }
}

// The same scan in two launches, for clouds of more than kChainedMaxTiles scan tiles: the chained kernel above is bound by
// the latency of its look-back chain (ncu at C3: 42 us at 7 % issue activity, barrier + long-scoreboard stalls), two
// dependent launches without any waiting inside are shorter. Pass 1: every tile gathers its counts through `order`, parks
// the rectangles in depth order and publishes its sum. Pass 2: every tile adds up the sums in front of it (all threads,
// 64-bit, saturating as above), scans its own counts -- now a dense read of the parked rectangles -- and writes the same
// outputs as the chained kernel: offsets, the tiles' inclusive prefixes in status[0..m), the total, the host's copy.
constexpr int kChainedMaxTiles = 16;
template <bool PACKED>
__global__ void __launch_bounds__(kScanThreads) scan_tile_sums_kernel(const uint32_t* __restrict__ n_dev,
                                                                       const uint32_t* __restrict__ order,
                                                                       const uint32_t* __restrict__ counts,
                                                                       uint32_t* __restrict__ sums,
                                                                       uint32_t* __restrict__ rect_sorted) {
    __shared__ uint32_t warp_sums[kScanThreads / 32];
    const int tile = (int)blockIdx.x;
    const int n = (int)__ldg(n_dev);
    if (tile * GSR_SCAN_TILE >= n) return;
    const int base = tile * GSR_SCAN_TILE + threadIdx.x * kScanItems;
    uint32_t g[kScanItems], v = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; i++) g[i] = 0u;
    if (base + kScanItems <= n) {
#pragma unroll
        for (int q = 0; q < kScanItems / 4; q++) {
            const uint4 o4 = __ldg(reinterpret_cast<const uint4*>(order + base) + q);
            g[4 * q] = o4.x; g[4 * q + 1] = o4.y; g[4 * q + 2] = o4.z; g[4 * q + 3] = o4.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < kScanItems; i++)
            if (base + i < n) g[i] = __ldg(order + base + i);
    }
    uint32_t c[kScanItems];
#pragma unroll
    for (int i = 0; i < kScanItems; i++) c[i] = (base + i < n) ? __ldg(counts + g[i]) : 0u;
    if (PACKED) {
        if (base + kScanItems <= n) {
#pragma unroll
            for (int q = 0; q < kScanItems / 4; q++)
                reinterpret_cast<uint4*>(rect_sorted + base)[q] = make_uint4(c[4 * q], c[4 * q + 1], c[4 * q + 2], c[4 * q + 3]);
        } else {
#pragma unroll
            for (int i = 0; i < kScanItems; i++)
                if (base + i < n) rect_sorted[base + i] = c[i];
        }
    }
#pragma unroll
    for (int i = 0; i < kScanItems; i++) v += PACKED ? ((c[i] >> 16) & 255u) * (c[i] >> 24) : c[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t tot = 0;
#pragma unroll
        for (int k = 0; k < kScanThreads / 32; k++) tot += warp_sums[k];
        sums[tile] = tot;                                  // <= GSR_SCAN_TILE * 255 * 255: fits
    }
}

template <bool PACKED>
__global__ void __launch_bounds__(kScanThreads) scan_apply_sums_kernel(const uint32_t* __restrict__ n_dev,
                                                                        const uint32_t* __restrict__ order,
                                                                        const uint32_t* __restrict__ counts,
                                                                        const uint32_t* __restrict__ sums,
                                                                        const uint32_t* __restrict__ rect_sorted, int m_slots,
                                                                        uint32_t* __restrict__ status,
                                                                        uint32_t* __restrict__ point_offsets,
                                                                        uint32_t* __restrict__ host_total) {
    __shared__ uint32_t warp_sums[kScanThreads / 32];
    __shared__ unsigned long long s_part[kScanThreads / 32];
    const int tile = (int)blockIdx.x;
    const int n = (int)__ldg(n_dev);
    const int m = (n + GSR_SCAN_TILE - 1) / GSR_SCAN_TILE;
    if (tile >= m) {
        if (m == 0 && tile == 0 && threadIdx.x == 0 && host_total != nullptr) { *host_total = 0u; __threadfence_system(); }
        return;
    }
    unsigned long long part = 0;
    for (int i = threadIdx.x; i < tile; i += kScanThreads) part += __ldg(sums + i);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = part;
    const int base = tile * GSR_SCAN_TILE + threadIdx.x * kScanItems;
    uint32_t c[kScanItems], v = 0;
    if (PACKED) {
        if (base + kScanItems <= n) {
#pragma unroll
            for (int q = 0; q < kScanItems / 4; q++) {
                const uint4 r4 = __ldg(reinterpret_cast<const uint4*>(rect_sorted + base) + q);
                c[4 * q] = r4.x; c[4 * q + 1] = r4.y; c[4 * q + 2] = r4.z; c[4 * q + 3] = r4.w;
            }
        } else {
#pragma unroll
            for (int i = 0; i < kScanItems; i++) c[i] = (base + i < n) ? __ldg(rect_sorted + base + i) : 0u;
        }
#pragma unroll
        for (int i = 0; i < kScanItems; i++) c[i] = ((c[i] >> 16) & 255u) * (c[i] >> 24);
    } else {
#pragma unroll
        for (int i = 0; i < kScanItems; i++) c[i] = (base + i < n) ? __ldg(counts + __ldg(order + base + i)) : 0u;
    }
#pragma unroll
    for (int i = 0; i < kScanItems; i++) v += c[i];
    uint32_t total;
    const uint32_t ex = block_exclusive_scan(v, warp_sums, total);          // syncs: s_part is visible behind it
    unsigned long long prefix = 0;
#pragma unroll
    for (int k = 0; k < kScanThreads / 32; k++) prefix += s_part[k];
    if (threadIdx.x == 0) {
        const unsigned long long incl_sum = prefix + total;
        const uint32_t sat = incl_sum >= kScanVal ? kScanVal : (uint32_t)incl_sum;
        status[tile] = kScanIncl | sat;
        if (tile == m - 1) {
            status[m_slots] = sat;
            if (host_total != nullptr) { *host_total = sat; __threadfence_system(); }
        }
    }
    uint32_t run = ex + (prefix >= kScanVal ? kScanVal : (uint32_t)prefix);
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        run += c[i];
        if (base + i < n) point_offsets[base + i] = run;
    }
}

// Replaces duplicateWithKeys (rasterizer_impl.cu:70-111), balanced by OUTPUT: a CTA writes kEmitTile (2048) consecutive instance
// slots, whatever number of Gaussians they belong to. (Emitting per Gaussian -- the reference's mapping, one thread per
// Gaussian -- leaves the near Gaussians, which cover hundreds of tiles each and come first in depth order, to a few
// threads while the rest of the GPU idles.) The CTA finds the Gaussians that own its slots in two cooperative steps,
// both counts over sorted arrays: the scan's per-tile inclusive prefixes give the first and last scan tile, the offsets
// of those two tiles the first and last Gaussian. Their offsets / ids / rectangles (<= kEmitTile + 1 of each) are staged in shared
// memory; a slot's owner is the number of range ends at or before it: the ends are strictly increasing, so they are
// distinct bits of a kEmitTile-bit map, and the owner is a prefix popcount. Stores are coalesced, writes clamped to `cap`.
constexpr int kEmitThreads = 256;
#ifndef GSR_EMIT_TILE
#define GSR_EMIT_TILE 2048
#endif
constexpr int kEmitTile = GSR_EMIT_TILE;                    // slots per CTA (a multiple of 1024). Measured at C3: 4096 -> 0.0597 ms,
                                                            // 2048 -> 0.0570 ms (twice the CTAs resident hide the search latency), 1024 -> 0.0782 ms
constexpr int kEmitMax = kEmitTile + 8;                     // Gaussians a CTA can need (4097), padded
template <bool PACKED>
__global__ void __launch_bounds__(kEmitThreads) emit_instances_kernel(const uint32_t* __restrict__ n_dev,
                                                                       const uint32_t* __restrict__ status, int m_slots,
                                                                       const uint32_t* __restrict__ offsets,
                                                                       const uint32_t* __restrict__ order,
                                                                       const uint32_t* __restrict__ rect_sorted,
                                                                       const float4* __restrict__ rec,
                                                                       const int* __restrict__ radii, int gx, int gy,
                                                                       uint32_t cap, uint32_t* __restrict__ tile_keys,
                                                                       uint32_t* __restrict__ vals) {
    extern __shared__ __align__(16) uint32_t emit_smem[];
    uint32_t* s_end = emit_smem;                            // [kEmitMax] inclusive end offset of the k-th Gaussian of the CTA
    uint32_t* s_gid = s_end + kEmitMax;
    uint32_t* s_rect = s_gid + kEmitMax;                    // PACKED: x0 | y0 << 8 | w << 16 | h << 24;  else x0 | y0 << 16
    uint32_t* s_bw = s_rect + kEmitMax;                     // !PACKED only: rectangle width
    __shared__ uint32_t s_bits[kEmitTile / 32];             // bit r of the map: some Gaussian's range ends at slot base + r
    __shared__ uint32_t s_wpre[kEmitTile / 32];             // ends in the words before this one
    __shared__ int s_cnt[4];
    const int t = threadIdx.x;
    const uint32_t total = min(__ldg(status + m_slots), cap);
    const uint32_t base = blockIdx.x * (uint32_t)kEmitTile;
    if (base >= total) return;
    const uint32_t stop = min(base + (uint32_t)kEmitTile, total);        // slots [base, stop)
    const int n = (int)__ldg(n_dev);
    const int m = (n + GSR_SCAN_TILE - 1) / GSR_SCAN_TILE;
    if (t < 4) s_cnt[t] = 0;
    for (int i = t; i < kEmitTile / 32; i += kEmitThreads) s_bits[i] = 0u;
    __syncthreads();
    // level 1: scan tiles. t_lo = tiles whose inclusive prefix is <= base (they end before our first slot);
    //          t_hi = tiles whose inclusive prefix is < stop (the tile after them holds the Gaussian that owns slot stop-1)
    {
        int lo = 0, hi = 0;
        for (int i = t; i < m; i += kEmitThreads) {
            const uint32_t incl = __ldg(status + i) & kScanVal;
            lo += incl <= base;
            hi += incl < stop;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { lo += __shfl_xor_sync(0xffffffffu, lo, o); hi += __shfl_xor_sync(0xffffffffu, hi, o); }
        if ((t & 31) == 0) { atomicAdd(&s_cnt[0], lo); atomicAdd(&s_cnt[1], hi); }
    }
    __syncthreads();
    const int t_lo = s_cnt[0], t_hi = min(s_cnt[1], m - 1);
    // level 2: Gaussians inside those two tiles. j_lo = first Gaussian whose end is > base; j_hi = first whose end is >= stop
    {
        int lo = 0, hi = 0;
        for (int i = t; i < GSR_SCAN_TILE; i += kEmitThreads) {
            const int ja = t_lo * GSR_SCAN_TILE + i, jb = t_hi * GSR_SCAN_TILE + i;
            if (ja < n) lo += __ldg(offsets + ja) <= base;
            if (jb < n) hi += __ldg(offsets + jb) < stop;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { lo += __shfl_xor_sync(0xffffffffu, lo, o); hi += __shfl_xor_sync(0xffffffffu, hi, o); }
        if ((t & 31) == 0) { atomicAdd(&s_cnt[2], lo); atomicAdd(&s_cnt[3], hi); }
    }
    __syncthreads();
    const int j_lo = t_lo * GSR_SCAN_TILE + s_cnt[2];
    const int j_hi = min(t_hi * GSR_SCAN_TILE + s_cnt[3], n - 1);
    const int cnt = min(j_hi - j_lo + 1, kEmitMax - 1);    // 1 .. 4097 (ends are strictly increasing: every visible Gaussian has >= 1 tile)
    const uint32_t first_start = j_lo > 0 ? __ldg(offsets + j_lo - 1) : 0u;
    for (int k = t; k < cnt; k += kEmitThreads) {
        const int j = j_lo + k;
        const uint32_t e = __ldg(offsets + j), gid = __ldg(order + j);
        s_end[k] = e;
        s_gid[k] = gid;
        if (PACKED) {
            s_rect[k] = __ldg(rect_sorted + j);
        } else {                                            // grids above 255 x 255 tiles: rectangle from record + radius
            int x0, y0, x1, y1;
            const float4 r0 = __ldg(rec + 2 * (size_t)gid);
            rect_of(r0.x, r0.y, __ldg(radii + gid), gx, gy, x0, y0, x1, y1);
            s_rect[k] = (uint32_t)x0 | ((uint32_t)y0 << 16);
            s_bw[k] = (uint32_t)max(1, x1 - x0);
        }
        const uint32_t rel = e - base;                      // e > base for every k (j >= j_lo)
        if (rel < (uint32_t)kEmitTile) atomicOr(&s_bits[rel >> 5], 1u << (rel & 31u));
    }
    __syncthreads();
    if (t < 32) {                                           // exclusive prefix of the words' popcounts (kEmitTile / 1024 words per lane)
        constexpr int kW = kEmitTile / 32 / 32;
        uint32_t c4[kW], sum = 0;
#pragma unroll
        for (int q = 0; q < kW; q++) { c4[q] = __popc(s_bits[kW * t + q]); sum += c4[q]; }
        uint32_t incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
            if (t >= o) incl += y;
        }
        uint32_t run = incl - sum;
#pragma unroll
        for (int q = 0; q < kW; q++) { s_wpre[kW * t + q] = run; run += c4[q]; }
    }
    __syncthreads();
#pragma unroll 4
    for (int i = 0; i < kEmitTile / kEmitThreads; i++) {
        const uint32_t r = (uint32_t)(i * kEmitThreads + t);           // slot base + r
        if (base + r >= stop) break;
        // owner = Gaussians of the CTA's list whose range ends at or before this slot, i.e. end - base <= r
        const uint32_t wd = r >> 5, bit = r & 31u;
        const int k = (int)(s_wpre[wd] + __popc(s_bits[wd] & (0xffffffffu >> (31u - bit))));
        const uint32_t start = k > 0 ? s_end[k - 1] : first_start;
        const uint32_t q = base + r - start;                            // row-major index inside the rectangle
        const uint32_t rc = s_rect[k];
        int x0, y0, bw;
        if (PACKED) { x0 = (int)(rc & 255u); y0 = (int)((rc >> 8) & 255u); bw = max(1, (int)((rc >> 16) & 255u)); }
        else { x0 = (int)(rc & 0xffffu); y0 = (int)(rc >> 16); bw = (int)s_bw[k]; }
        const int row = (int)(((float)q + 0.5f) * __frcp_rn((float)bw));   // exact: the quotient stays >= 0.5 / 4096 away
        const int col = (int)q - row * bw;                                  //        from an integer, the product errs by < 1e-3 of that
        tile_keys[base + r] = (uint32_t)((y0 + row) * gx + x0 + col);
        vals[base + r] = s_gid[k];
    }
}

// ------------------------------------------------------------------ instance count without binning (first call of a shape)
template <bool PACKED>
__global__ void __launch_bounds__(256) count_instances_kernel(int P, const uint32_t* __restrict__ counts,
                                                               unsigned long long* __restrict__ out) {
    unsigned long long sum = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
        const uint32_t c = __ldg(counts + i);
        sum += PACKED ? ((c >> 16) & 255u) * (c >> 24) : c;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if ((threadIdx.x & 31) == 0 && sum) atomicAdd(out, sum);
}

int launch_count_instances(const gsr_view& v, GeomState& g, cudaStream_t s) {
    int sms = 148;
    { const int e = sm_count(&sms); if (e) return e; }
    const int blocks = min((v.P + 255) / 256, sms * 8);
    if (rect_packing_ok(v.W, v.H))
        GSR_LAUNCH("count_instances", v.debug, s, (count_instances_kernel<true><<<blocks, 256, 0, s>>>(v.P, g.rect_packed, g.count64)));
    else
        GSR_LAUNCH("count_instances", v.debug, s, (count_instances_kernel<false><<<blocks, 256, 0, s>>>(v.P, g.tiles_touched, g.count64)));
    return 0;
}

// the depth sort's tables and the instance counter are carved next to each other: one region to clear
void depth_sort_zero_region(GeomState& g, void** begin, size_t* bytes) {
    *begin = g.sort_scratch;
    *bytes = (size_t)(reinterpret_cast<const char*>(g.count64 + 2) - reinterpret_cast<const char*>(g.sort_scratch));
}

int launch_depth_order(const gsr_view& v, GeomState& g, cudaStream_t s, bool scratch_zeroed) {
    const int P = v.P;
    bool in_a = true;
    if (!scratch_zeroed) {
        void* zb; size_t zn;
        depth_sort_zero_region(g, &zb, &zn);
        GSR_MEMSET("memset_depth_sort", s, zb, 0, zn);
    }
    uint32_t* n_vis = nullptr;
    int e = radix_sort_u32(g.dkey, g.order, g.dkey_tmp, g.order_tmp, P, 32, g.sort_scratch, &in_a, v.debug != 0, s, &n_vis,
                           /*scratch_zeroed=*/true);
    if (e) return e;
    // 4 passes: the result is back in (dkey, order); its first *n_vis entries are the visible Gaussians
    if (n_vis != g.n_visible) return GSR_E_BADARG;              // layout contract between the two files
    return 0;
}

// ------------------------------------------------------------------ tile table (A9 + the histograms of A8)
// Every visible Gaussian adds +1 / -1 at the four corners of its tile rectangle to a difference array; the array's 2-D
// prefix sum is the number of instances per tile. Their exclusive scan in tile order gives the ranges (identifyTileRanges,
// rasterizer_impl.cu:116-138: empty tiles keep (0, 0), :310), and the counts binned by the digits of the tile id are the
// histograms the instance sort's passes start from -- no pass over the instances for either.
//
// Straight global REDs for the corners are out: they pile up on the cells around the image centre, and same-address
// reductions serialise in L2 at ~30 ns each (measured: 5.9M REDs on 8349 cells cost 0.15 ms). So a few persistent CTAs
// accumulate their share of the Gaussians in shared memory (the rectangles stream in coalesced, 16 bytes per thread, in id
// order -- no depth order needed) and flush their non-zero cells with one RED each; the LAST CTA to finish (a counter in
// global memory) then turns the array into ranges and histograms, so the table costs one launch.
constexpr int kTableThreads = 1024;
constexpr int kTableSmemMax = 200 * 1024;

__device__ __forceinline__ void diff_add_rect(int32_t* acc, int stride, int gx, int gy, int x0, int y0, int bw, int bh) {
    // corners on the far border (x = gx or y = gy) are never read by the prefix sums: skipped
    const int xe = x0 + bw, ye = y0 + bh;
    atomicAdd(acc + y0 * stride + x0, 1);
    if (xe < gx) atomicAdd(acc + y0 * stride + xe, -1);
    if (ye < gy) atomicAdd(acc + ye * stride + x0, -1);
    if (xe < gx && ye < gy) atomicAdd(acc + ye * stride + xe, 1);
}

// the table itself, run by one CTA of kTableThreads threads. `work` holds the difference array (shared memory copy, or
// the global array itself for grids too large for shared memory: slow path, correct).
__device__ void tile_table_block(int32_t* work, int gx, int gy, int passes, int total_bits, uint2* __restrict__ ranges,
                                 uint32_t* __restrict__ ghist, uint32_t (*hist)[GSR_RADIX], uint32_t* warp_sums) {
    const int t = threadIdx.x, lane = t & 31, w = t >> 5, stride = gx + 1;
    for (int i = t; i < 4 * GSR_RADIX; i += kTableThreads) (&hist[0][0])[i] = 0u;
    // prefix along every row, one warp per row, 32 columns per step with a carried sum
    for (int y = w; y < gy; y += kTableThreads / 32) {
        int32_t carry = 0;
        for (int x0 = 0; x0 < gx; x0 += 32) {
            const int x = x0 + lane;
            int32_t v = x < gx ? work[y * stride + x] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int32_t u = __shfl_up_sync(0xffffffffu, v, o);
                if (lane >= o) v += u;
            }
            v += carry;
            if (x < gx) work[y * stride + x] = v;
            carry = __shfl_sync(0xffffffffu, v, 31);
        }
    }
    __syncthreads();
    for (int x = t; x < gx; x += kTableThreads) {          // prefix down every column: work[y][x] = instances of tile (x, y)
        int32_t run = 0;
        for (int y = 0; y < gy; y++) { run += work[y * stride + x]; work[y * stride + x] = run; }
    }
    __syncthreads();
    const int tiles = gx * gy;
    const int per = (tiles + kTableThreads - 1) / kTableThreads;
    const int lo = min(t * per, tiles), hi = min(lo + per, tiles);
    uint32_t local = 0;
    {
        int y = lo / gx, x = lo - y * gx;
        for (int k = lo; k < hi; k++) {
            local += (uint32_t)work[y * stride + x];
            if (++x == gx) { x = 0; y++; }
        }
    }
    uint32_t incl = local;                                  // exclusive scan of `local` over the 1024 threads
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += u;
    }
    if (lane == 31) warp_sums[w] = incl;
    __syncthreads();
    uint32_t ofs = incl - local;
    for (int k = 0; k < w; k++) ofs += warp_sums[k];
    {
        int y = lo / gx, x = lo - y * gx;
        for (int k = lo; k < hi; k++) {
            const uint32_t c = (uint32_t)work[y * stride + x];
            ranges[k] = c ? make_uint2(ofs, ofs + c) : make_uint2(0u, 0u);
            ofs += c;
            if (c) {
                for (int p = 0; p < passes; p++) {
                    const int bits = min(8, total_bits - 8 * p);
                    atomicAdd(&hist[p][((uint32_t)k >> (8 * p)) & ((1u << bits) - 1u)], c);
                }
            }
            if (++x == gx) { x = 0; y++; }
        }
    }
    __syncthreads();
    for (int i = t; i < passes * GSR_RADIX; i += kTableThreads) ghist[i] = (&hist[0][0])[i];
}

template <bool PACKED>
__global__ void __launch_bounds__(kTableThreads) tile_table_kernel(int P, const uint32_t* __restrict__ counts,
                                                                    const float4* __restrict__ rec,
                                                                    const int* __restrict__ radii, int gx, int gy,
                                                                    int32_t* __restrict__ tile_diff, int use_smem,
                                                                    uint32_t* __restrict__ done_counter, int passes,
                                                                    int total_bits, uint2* __restrict__ ranges,
                                                                    uint32_t* __restrict__ ghist) {
    extern __shared__ __align__(16) int32_t table_smem[];
    __shared__ uint32_t hist[4][GSR_RADIX];
    __shared__ uint32_t warp_sums[kTableThreads / 32];
    __shared__ int s_last;
    const int t = threadIdx.x, stride = gx + 1, cells = stride * (gy + 1);
    int32_t* acc = use_smem ? table_smem : tile_diff;
    if (use_smem) {
        for (int i = t; i < cells; i += kTableThreads) acc[i] = 0;
        __syncthreads();
    }
    // this CTA's share of the Gaussians: four per thread and step (16-byte loads; `counts` is 128-byte aligned)
    const int nvec = P >> 2;
    const uint4* c4 = reinterpret_cast<const uint4*>(counts);
    for (int i = blockIdx.x * kTableThreads + t; i < nvec + 1; i += gridDim.x * kTableThreads) {
        uint32_t c[4] = {0u, 0u, 0u, 0u};
        if (i < nvec) { const uint4 v = __ldg(c4 + i); c[0] = v.x; c[1] = v.y; c[2] = v.z; c[3] = v.w; }
        else { for (int q = 0; q < (P & 3); q++) c[q] = __ldg(counts + 4 * nvec + q); }
#pragma unroll
        for (int q = 0; q < 4; q++) {
            if (c[q] == 0u) continue;                               // culled
            int x0, y0, bw, bh;
            if (PACKED) {
                x0 = (int)(c[q] & 255u); y0 = (int)((c[q] >> 8) & 255u); bw = (int)((c[q] >> 16) & 255u); bh = (int)(c[q] >> 24);
            } else {
                const int g = 4 * i + q;
                int x1, y1;
                const float4 r0 = __ldg(rec + 2 * (size_t)g);
                rect_of(r0.x, r0.y, __ldg(radii + g), gx, gy, x0, y0, x1, y1);
                bw = x1 - x0; bh = y1 - y0;
            }
            diff_add_rect(acc, stride, gx, gy, x0, y0, bw, bh);
        }
    }
    if (use_smem) {
        __syncthreads();
        if (gridDim.x > 1) {
            for (int i = t; i < cells; i += kTableThreads) {
                const int32_t v = acc[i];
                if (v != 0) atomicAdd(tile_diff + i, v);
            }
        }
    }
    if (gridDim.x > 1) {
        // the last CTA to arrive builds the table from the complete array
        __threadfence();
        __syncthreads();
        if (t == 0) s_last = atomicAdd(done_counter, 1u) == gridDim.x - 1 ? 1 : 0;
        __syncthreads();
        if (!s_last) return;
        __threadfence();
        if (use_smem)
            for (int i = t; i < cells; i += kTableThreads) acc[i] = __ldcg(tile_diff + i);
    }
    __syncthreads();
    tile_table_block(acc, gx, gy, passes, total_bits, ranges, ghist, hist, warp_sums);
}

static int bits_for(uint32_t count) {   // bits needed to represent ids 0..count-1
    int b = 0;
    while (b < 32 && (1ull << b) < count) b++;
    return b;
}

// Clears the binning blob's tables and builds the tile table (ranges + the instance sort's histograms). Depends on the
// preprocess only, so gsr_forward runs it on a side stream under the depth sort.
int launch_tile_table(const gsr_view& v, GeomState& g, const int* radii, BinningState& b, ImageState& img, cudaStream_t s,
                      bool scratch_zeroed) {
    const int gx = (v.W + GSR_TILE - 1) / GSR_TILE, gy = (v.H + GSR_TILE - 1) / GSR_TILE;
    const int tiles = gx * gy;
    const int tbits = bits_for((uint32_t)tiles);
    const int passes = radix_passes(tbits);
    if (passes > 4) return GSR_E_TOOLARGE;
    const int m = (int)BinningState::scan_tiles((size_t)v.P);
    if (!scratch_zeroed) GSR_MEMSET("memset_binning", s, b.zero_begin, 0, (size_t)(b.zero_end - b.zero_begin));
    {
        const size_t table_bytes = BinningState::diff_cells(v.W, v.H) * sizeof(int32_t);
        const int use_smem = table_bytes <= (size_t)kTableSmemMax ? 1 : 0;
        int sms = 148;
        { const int e = sm_count(&sms); if (e) return e; }
        // one CTA per ~32K Gaussians, at most one per SM: a single CTA (small scenes) needs no flush and no hand-over
        const int blocks = std::max(1, std::min((v.P + 32767) / 32768, sms));
        const size_t dyn = use_smem ? table_bytes : 0;
        uint32_t* done_counter = b.scan_sums + m + 2;           // zeroed with the scan words
        static DeviceOnce once_p, once_u;
        if (rect_packing_ok(v.W, v.H)) {
            { const int e = ensure_dynamic_smem(tile_table_kernel<true>, kTableSmemMax, once_p); if (e) return e; }
            GSR_LAUNCH("tile_table", v.debug, s, (tile_table_kernel<true><<<blocks, kTableThreads, dyn, s>>>(
                           v.P, g.rect_packed, g.rec, radii, gx, gy, b.tile_diff, use_smem, done_counter, passes, tbits, img.ranges,
                           b.sort_scratch)));
        } else {
            { const int e = ensure_dynamic_smem(tile_table_kernel<false>, kTableSmemMax, once_u); if (e) return e; }
            GSR_LAUNCH("tile_table", v.debug, s, (tile_table_kernel<false><<<blocks, kTableThreads, dyn, s>>>(
                           v.P, g.tiles_touched, g.rec, radii, gx, gy, b.tile_diff, use_smem, done_counter, passes, tbits, img.ranges,
                           b.sort_scratch)));
        }
    }
    return 0;
}

static int g_scan_two_pass = [] { const char* e = getenv("GSR_SCAN_TWO_PASS"); return e ? atoi(e) : 1; }();   // A/B switch

int launch_binning(const gsr_view& v, GeomState& g, const int* radii, BinningState& b, ImageState& img, uint32_t cap,
                   uint32_t** total_dev, uint32_t* host_total, cudaStream_t s) {
    (void)img;
    const int gx = (v.W + GSR_TILE - 1) / GSR_TILE, gy = (v.H + GSR_TILE - 1) / GSR_TILE;
    const int tiles = gx * gy;
    const int tbits = bits_for((uint32_t)tiles);
    const int passes = radix_passes(tbits);
    if (passes > 4) return GSR_E_TOOLARGE;
    const int m = (int)BinningState::scan_tiles((size_t)v.P);
    *total_dev = b.scan_sums + m;
    // make the sorted result land in point_list
    uint32_t *ka = b.tile_keys, *va = b.point_list, *kb = b.tile_keys_tmp, *vb = b.vals_tmp;
    if (passes % 2) { ka = b.tile_keys_tmp; va = b.vals_tmp; kb = b.tile_keys; vb = b.point_list; }
    if (cap > 0) {
        const int emit_blocks = (int)((cap + (uint32_t)kEmitTile - 1) / (uint32_t)kEmitTile);
        static DeviceOnce once_ep, once_eu;
        if (rect_packing_ok(v.W, v.H)) {
            const size_t smem = sizeof(uint32_t) * 3 * kEmitMax;
            { const int e = ensure_dynamic_smem(emit_instances_kernel<true>, (int)smem, once_ep); if (e) return e; }
            if (m > kChainedMaxTiles && g_scan_two_pass) {
                uint32_t* sums = b.scan_sums + m + 8;
                GSR_LAUNCH("scan_tile_sums", v.debug, s,
                           (scan_tile_sums_kernel<true><<<m, kScanThreads, 0, s>>>(g.n_visible, g.order, g.rect_packed, sums, g.rect_sorted)));
                GSR_LAUNCH("scan_offsets", v.debug, s,
                           (scan_apply_sums_kernel<true><<<m, kScanThreads, 0, s>>>(g.n_visible, g.order, g.rect_packed, sums, g.rect_sorted,
                                                                                   m, b.scan_sums, g.point_offsets, host_total)));
            } else
            GSR_LAUNCH("scan_offsets", v.debug, s,
                       (scan_offsets_kernel<true><<<m, kScanThreads, 0, s>>>(g.n_visible, g.order, g.rect_packed, m, b.scan_sums,
                                                                            g.point_offsets, g.rect_sorted, host_total)));
            GSR_LAUNCH("emit_instances", v.debug, s,
                       (emit_instances_kernel<true><<<emit_blocks, kEmitThreads, smem, s>>>(
                           g.n_visible, b.scan_sums, m, g.point_offsets, g.order, g.rect_sorted, g.rec, radii, gx, gy, cap, ka, va)));
        } else {
            const size_t smem = sizeof(uint32_t) * 4 * kEmitMax;
            { const int e = ensure_dynamic_smem(emit_instances_kernel<false>, (int)smem, once_eu); if (e) return e; }
            if (m > kChainedMaxTiles && g_scan_two_pass) {
                uint32_t* sums = b.scan_sums + m + 8;
                GSR_LAUNCH("scan_tile_sums", v.debug, s,
                           (scan_tile_sums_kernel<false><<<m, kScanThreads, 0, s>>>(g.n_visible, g.order, g.tiles_touched, sums, nullptr)));
                GSR_LAUNCH("scan_offsets", v.debug, s,
                           (scan_apply_sums_kernel<false><<<m, kScanThreads, 0, s>>>(g.n_visible, g.order, g.tiles_touched, sums, nullptr,
                                                                                    m, b.scan_sums, g.point_offsets, host_total)));
            } else
            GSR_LAUNCH("scan_offsets", v.debug, s,
                       (scan_offsets_kernel<false><<<m, kScanThreads, 0, s>>>(g.n_visible, g.order, g.tiles_touched, m, b.scan_sums,
                                                                             g.point_offsets, nullptr, host_total)));
            GSR_LAUNCH("emit_instances", v.debug, s,
                       (emit_instances_kernel<false><<<emit_blocks, kEmitThreads, smem, s>>>(
                           g.n_visible, b.scan_sums, m, g.point_offsets, g.order, nullptr, g.rec, radii, gx, gy, cap, ka, va)));
        }
    }
    if (cap == 0) return 0;
    bool in_a = true;
    const SortOptions opt{/*hist_ready=*/true, /*n_dev=*/*total_dev, /*values_only_last=*/true};
    return radix_sort_u32(ka, va, kb, vb, (int)cap, tbits, b.sort_scratch, &in_a, v.debug != 0, s, nullptr, /*scratch_zeroed=*/true, &opt);
}

// ------------------------------------------------------------------ parity helper: the reference's 64-bit keys
// one CTA per tile: the instances of its range carry its id in the upper half (rasterizer_impl.cu:98-104)
__global__ void sorted_keys_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                                   const float* __restrict__ depths, uint64_t* __restrict__ keys) {
    const uint2 r = ranges[blockIdx.x];
    for (uint32_t i = r.x + threadIdx.x; i < r.y; i += blockDim.x)
        keys[i] = ((uint64_t)blockIdx.x << 32) | __float_as_uint(depths[point_list[i]]);
}

int launch_sorted_keys_debug(const gsr_view& v, const GeomState& g, const uint32_t* point_list, const ImageState& img, int R,
                             uint64_t* keys_out, cudaStream_t s) {
    if (R <= 0) return 0;
    const int gx = (v.W + GSR_TILE - 1) / GSR_TILE, gy = (v.H + GSR_TILE - 1) / GSR_TILE;
    GSR_LAUNCH("sorted_keys", v.debug, s, (sorted_keys_kernel<<<gx * gy, 256, 0, s>>>(img.ranges, point_list, g.depths, keys_out)));
    return 0;
}

}  // namespace gsr
