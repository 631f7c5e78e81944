// gsr_binning.cu -- depth ordering of Gaussians, tile-count scan, instance emission, tile sort,
// tile ranges. Replaces rasterizer_impl.cu:70-138 (duplicateWithKeys, identifyTileRanges) and the
// CUB calls at :277 (InclusiveSum) and :303 (SortPairs). HBM-bound integer work.
//
// Order of the final point list (must equal the reference's stable 64-bit sort):
//   by tile id, then by depth bits, then by Gaussian id.
// We get it as: (1) stable sort of the VISIBLE Gaussians' ids by depth bits (culled ones carry the key 0xFFFFFFFF and
//                   are dropped by the sort's first pass, so 3 of the 4 passes, the scan and the emission run over
//                   the visible count only -- half of P on the bench scene; that count stays on the device);
//               (2) emit instances walking the Gaussians in that order (offsets from a scan taken in
//                   the same order);  (3) stable sort of instances by tile id only.
#include "gsr_common.cuh"

namespace gsr {

// ------------------------------------------------------------------ scan
constexpr int kScanThreads = 256;
constexpr int kScanItems = GSR_SCAN_TILE / kScanThreads;   // 8, blocked per thread

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

// Single-pass chained scan (decoupled look-back) of tiles_touched in depth order: one launch instead of
// reduce / spine / apply. A tile's status word carries flag (bits 30-31: 1 = the tile's own sum, 2 = inclusive prefix)
// and value (30 bits: num_rendered < 2^30 is enforced by the caller) together, so one 32-bit store publishes both.
// Tiles are taken by ticket, which makes waiting on lower-numbered tiles safe whatever the CTA scheduling order.
// status[0..m) = tile words, status[m] = grand total, status[m + 1] = ticket; all zeroed before the launch.
constexpr uint32_t kScanOwn = 1u << 30, kScanIncl = 2u << 30, kScanVal = (1u << 30) - 1u;
// PACKED: `counts` holds packed tile rectangles (GeomState::rect_packed); the count is width x height and the word is
// also written out in depth order (rect_sorted), so the emission streams it instead of gathering record + radius.
template <bool PACKED>
__global__ void __launch_bounds__(kScanThreads) scan_chained_kernel(const uint32_t* __restrict__ counts,
                                                                     const uint32_t* __restrict__ order,
                                                                     const uint32_t* __restrict__ n_dev, int m_slots,
                                                                     uint32_t* __restrict__ status,
                                                                     uint32_t* __restrict__ out,
                                                                     uint32_t* __restrict__ rect_sorted) {
    // n (items, in depth order) lives on the device; the grid covers the host's upper bound and m_slots only fixes
    // where the total and the ticket sit. CTAs whose ticket lies beyond the last tile leave at once.
    __shared__ uint32_t warp_sums[kScanThreads / 32];
    __shared__ uint32_t s_tile, s_prefix;
    if (threadIdx.x == 0) s_tile = atomicAdd(status + m_slots + 1, 1u);
    __syncthreads();
    const int tile = (int)s_tile;
    const int n = (int)__ldg(n_dev);
    const int m = (n + GSR_SCAN_TILE - 1) / GSR_SCAN_TILE;
    if (tile >= m) return;
    const int base = tile * GSR_SCAN_TILE + threadIdx.x * kScanItems;
    uint32_t c[kScanItems], v = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        c[i] = (base + i < n) ? __ldg(counts + __ldg(order + base + i)) : 0u;
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
    if (threadIdx.x < 32) {
        // look-back, 128 predecessors per round trip: lane l reads tiles j - 4l .. j - 4l - 3 (nearest first). The front
        // of finished (inclusive) prefixes advances one window per L2 round trip, so a window of 32 made the last of
        // ~700 tiles wait for ~20 round trips.
        const int lane = threadIdx.x;
        uint32_t prefix = 0;
        for (int j = tile - 1; j >= 0; j -= 128) {
            uint32_t part = 0;                             // sum of this lane's words up to (and including) its first inclusive one
            bool has_incl = false;
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const int idx = j - 4 * lane - q;
                uint32_t st = kScanIncl;                   // before the first tile: inclusive prefix 0
                if (idx >= 0) {
                    st = vs[idx];
                    while ((st >> 30) == 0u) { __nanosleep(20); st = vs[idx]; }
                }
                if (!has_incl) part += st & kScanVal;
                has_incl = has_incl || (st >> 30) == 2u;
            }
            const unsigned incl = __ballot_sync(0xffffffffu, has_incl);
            const int first = incl ? __ffs(incl) - 1 : 31;
            uint32_t val = lane <= first ? part : 0u;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
            prefix += val;
            if (incl) break;
        }
        if (lane == 0) {
            if (tile > 0) vs[tile] = kScanIncl | (prefix + total);
            if (tile == m - 1) status[m_slots] = prefix + total;
            s_prefix = prefix;
        }
    }
    __syncthreads();
    uint32_t run = ex + s_prefix;
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        run += c[i];
        if (base + i < n) out[base + i] = run;   // inclusive, as rasterizer_impl.cu:277
    }
}


int launch_order_and_scan(const gsr_view& v, GeomState& g, uint32_t** total_dev, cudaStream_t s) {
    const int P = v.P;
    bool in_a = true;
    // one memset clears the scan's status words and the depth sort's tables: they are carved next to each other
    const size_t sort_words = 4 * GSR_RADIX + 64 + (size_t)radix_passes(32) * sort_tiles((size_t)P) * GSR_RADIX;
    const char* zero_end = reinterpret_cast<const char*>(g.sort_scratch + sort_words);
    GSR_MEMSET("memset_scan", s, g.scan_sums, 0, (size_t)(zero_end - reinterpret_cast<const char*>(g.scan_sums)));
    uint32_t* n_vis = nullptr;
    int e = radix_sort_u32(g.dkey, g.order, g.dkey_tmp, g.order_tmp, P, 32, g.sort_scratch, &in_a, v.debug != 0, s, &n_vis,
                           /*scratch_zeroed=*/true);
    if (e) return e;
    // 4 passes: the result is back in (dkey, order); its first *n_vis entries are the visible Gaussians
    if (n_vis != g.n_visible) return GSR_E_BADARG;              // layout contract between the two files
    const int m = (P + GSR_SCAN_TILE - 1) / GSR_SCAN_TILE;
    if (rect_packing_ok(v.W, v.H))
        GSR_LAUNCH("scan_chained", v.debug, s,
                   (scan_chained_kernel<true><<<m, kScanThreads, 0, s>>>(g.rect_packed, g.order, n_vis, m, g.scan_sums, g.point_offsets, g.rect_sorted)));
    else
        GSR_LAUNCH("scan_chained", v.debug, s,
                   (scan_chained_kernel<false><<<m, kScanThreads, 0, s>>>(g.tiles_touched, g.order, n_vis, m, g.scan_sums, g.point_offsets, nullptr)));
    *total_dev = g.scan_sums + m;
    return 0;
}

// ------------------------------------------------------------------ instance emission (A7)
__device__ __forceinline__ void rect_of(float px, float py, int max_radius, int gx, int gy, int& x0, int& y0,
                                        int& x1, int& y1) {
    // auxiliary.h:46-56 (same float expressions as the forward preprocess)
    x0 = min(gx, max((int)0, (int)((px - max_radius) / GSR_TILE)));
    y0 = min(gy, max((int)0, (int)((py - max_radius) / GSR_TILE)));
    x1 = min(gx, max((int)0, (int)((px + max_radius + GSR_TILE - 1) / GSR_TILE)));
    y1 = min(gy, max((int)0, (int)((py + max_radius + GSR_TILE - 1) / GSR_TILE)));
}

// One warp emits the instances of 32 consecutive Gaussians (depth order). Their output slots are one contiguous
// range, so the lanes walk that range 32 slots at a time -- fully coalesced stores -- and find the owner of their
// slot by a 5-step binary search over the lanes' end offsets (shuffles). Replaces the thread-serial, divergent loop of
// duplicateWithKeys (rasterizer_impl.cu:88-109): a warp-store there touches 32 different sectors.
template <bool PACKED>
__global__ void __launch_bounds__(256) emit_instances_kernel(const uint32_t* __restrict__ n_dev, const uint32_t* __restrict__ order,
                                                              const uint32_t* __restrict__ rect_sorted,
                                                              const uint32_t* __restrict__ offsets,
                                                              const float4* __restrict__ rec,
                                                              const int* __restrict__ radii, int gx, int gy,
                                                              uint32_t* __restrict__ tile_keys,
                                                              uint32_t* __restrict__ vals) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const int P = (int)__ldg(n_dev);               // visible Gaussians (depth order); the grid covers the host's bound
    if (j - lane >= P) return;
    uint32_t g = 0, n = 0, end = 0;
    int x0 = 0, y0 = 0, bw = 1;
    if (j < P) {
        g = __ldg(order + j);
        end = __ldg(offsets + j);                   // inclusive scan in depth order
        n = end - (j > 0 ? __ldg(offsets + j - 1) : 0u);   // = tiles_touched[g], without the gather
        if (PACKED) {                               // streamed, in depth order (written by the scan)
            const uint32_t r = __ldg(rect_sorted + j);
            x0 = (int)(r & 255u); y0 = (int)((r >> 8) & 255u); bw = max(1, (int)((r >> 16) & 255u));
        } else if (n > 0) {
            int x1, y1;
            const float4 r0 = __ldg(rec + 3 * (size_t)g);
            rect_of(r0.x, r0.y, __ldg(radii + g), gx, gy, x0, y0, x1, y1);
            bw = x1 - x0;
        }
    }
    if (__ballot_sync(0xffffffffu, n > 0) == 0u) return;
    // lanes beyond P: end = end of the last valid lane, so they never own a slot that is written
    const uint32_t last_end = __shfl_sync(0xffffffffu, end, min(31, P - 1 - (j - lane)));
    if (j >= P) end = last_end;
    const uint32_t start = end - n;
    const uint32_t first = __shfl_sync(0xffffffffu, start, 0);
    const uint32_t xy = (uint32_t)x0 | ((uint32_t)y0 << 16);
    // The lanes walk the warp's output range 32 slots at a time. Owner of slot s = number of Gaussians whose range ends
    // at or before s. The ends are strictly increasing, so inside a window of 32 slots they are distinct bits: ONE warp-wide
    // OR (REDUX) of those bits plus a running count of the ends before the window replaces a per-slot binary search over
    // the lanes (five dependent shuffles); four shuffles then fetch the owner's id, range start, corner and width.
    // The kernel was bound by the shuffle / LSU pipe (69 % busy) before this.
    uint32_t before = 0;                            // ends below the window's first slot (warp-uniform)
    for (uint32_t base = first; base < last_end; base += 32) {
        const uint32_t rel = end - base;            // wraps to a huge value for ends below the window
        const uint32_t mine = (rel < 32u && j < P) ? (1u << rel) : 0u;
        const uint32_t ends = __reduce_or_sync(0xffffffffu, mine);
        const uint32_t s = base + lane;
        const int owner = min((int)(before + __popc(ends & (0xffffffffu >> (31 - lane)))), 31);
        before += __popc(ends);
        const uint32_t og = __shfl_sync(0xffffffffu, g, owner);
        const uint32_t ostart = __shfl_sync(0xffffffffu, start, owner);
        const uint32_t oxy = __shfl_sync(0xffffffffu, xy, owner);
        const int obw = __shfl_sync(0xffffffffu, bw, owner);
        if (s < last_end) {
            const uint32_t k = s - ostart;                             // row-major index inside the rectangle
            const int row = (int)(((float)k + 0.5f) * __frcp_rn((float)obw));   // exact: the quotient stays >= 0.5 / 4096 away
            const int col = (int)k - row * obw;                        //        from an integer, the product errs by < 1e-3 of that
            tile_keys[s] = (uint32_t)(((int)(oxy >> 16) + row) * gx + (int)(oxy & 0xffffu) + col);
            vals[s] = og;
        }
    }
}

// ------------------------------------------------------------------ tile ranges (A9)
// four consecutive sorted keys per thread (one 16-byte load) + the key before them
__global__ void __launch_bounds__(256) tile_ranges_kernel(int L, const uint32_t* __restrict__ tile_keys,
                                                           uint2* __restrict__ ranges) {
    const int base = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (base >= L) return;
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(tile_keys) + (base >> 2));   // buffer is padded by 8 words
    const uint32_t k[4] = {v.x, v.y, v.z, v.w};
    uint32_t prev = base > 0 ? __ldg(tile_keys + base - 1) : 0u;
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const int idx = base + q;
        if (idx < L) {
            const uint32_t cur = k[q];
            if (idx == 0) ranges[cur].x = 0;
            else if (cur != prev) {
                ranges[prev].y = idx;
                ranges[cur].x = idx;
            }
            if (idx == L - 1) ranges[cur].y = L;
            prev = cur;
        }
    }
}

static int bits_for(uint32_t count) {   // bits needed to represent ids 0..count-1
    int b = 0;
    while (b < 32 && (1ull << b) < count) b++;
    return b;
}

int launch_binning(const gsr_view& v, const GeomState& g, const int* radii, BinningState& b, ImageState& img, int R,
                   cudaStream_t s) {
    const int gx = (v.W + GSR_TILE - 1) / GSR_TILE, gy = (v.H + GSR_TILE - 1) / GSR_TILE;
    const int tiles = gx * gy;
    cudaError_t ce = cudaMemsetAsync(img.ranges, 0, sizeof(uint2) * (size_t)tiles, s);   // rasterizer_impl.cu:310
    if (ce != cudaSuccess) return (int)ce;
    if (R <= 0) return 0;
    const int tbits = bits_for((uint32_t)tiles);
    const int passes = radix_passes(tbits);
    // make the sorted result land in (tile_keys, point_list)
    uint32_t *ka = b.tile_keys, *va = b.point_list, *kb = b.tile_keys_tmp, *vb = b.vals_tmp;
    if (passes % 2) { ka = b.tile_keys_tmp; va = b.vals_tmp; kb = b.tile_keys; vb = b.point_list; }
    if (rect_packing_ok(v.W, v.H))
        GSR_LAUNCH("emit_instances", v.debug, s,
                   (emit_instances_kernel<true><<<(v.P + 255) / 256, 256, 0, s>>>(g.n_visible, g.order, g.rect_sorted, g.point_offsets,
                                                                              g.rec, radii, gx, gy, ka, va)));
    else
        GSR_LAUNCH("emit_instances", v.debug, s,
                   (emit_instances_kernel<false><<<(v.P + 255) / 256, 256, 0, s>>>(g.n_visible, g.order, nullptr, g.point_offsets,
                                                                               g.rec, radii, gx, gy, ka, va)));
    bool in_a = true;
    int e = radix_sort_u32(ka, va, kb, vb, R, tbits, b.sort_scratch, &in_a, v.debug != 0, s);
    if (e) return e;
    GSR_LAUNCH("tile_ranges", v.debug, s, (tile_ranges_kernel<<<(R + 1023) / 1024, 256, 0, s>>>(R, b.tile_keys, img.ranges)));
    return 0;
}

// ------------------------------------------------------------------ parity helper: the reference's 64-bit keys
__global__ void sorted_keys_kernel(int R, const uint32_t* __restrict__ tile_keys, const uint32_t* __restrict__ point_list,
                                   const float* __restrict__ depths, uint64_t* __restrict__ keys) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= R) return;
    uint64_t k = tile_keys[i];
    k <<= 32;
    k |= __float_as_uint(depths[point_list[i]]);    // rasterizer_impl.cu:102-104
    keys[i] = k;
}

int launch_sorted_keys_debug(const gsr_view& v, const GeomState& g, const BinningState& b, int R, uint64_t* keys_out,
                             cudaStream_t s) {
    if (R <= 0) return 0;
    GSR_LAUNCH("sorted_keys", v.debug, s,
               (sorted_keys_kernel<<<(R + 255) / 256, 256, 0, s>>>(R, b.tile_keys, b.point_list, g.depths, keys_out)));
    return 0;
}

}  // namespace gsr
