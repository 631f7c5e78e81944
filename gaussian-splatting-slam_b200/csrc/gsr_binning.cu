// gsr_binning.cu -- depth ordering of Gaussians, tile-count scan + instance emission (one kernel), tile table
// (ranges + digit histograms), tile sort. Replaces rasterizer_impl.cu:70-138 (duplicateWithKeys, identifyTileRanges)
// and the CUB calls at :277 (InclusiveSum) and :303 (SortPairs). HBM-bound integer work.
//
// Order of the final point list (must equal the reference's stable 64-bit sort):
//   by tile id, then by depth bits, then by Gaussian id.
// We get it as: (1) stable sort of the VISIBLE Gaussians' ids by depth bits (culled ones carry the key 0xFFFFFFFF and
//                   are dropped by the sort's first pass, so 3 of the 4 passes, the scan and the emission run over
//                   the visible count only -- half of P on the bench scene; that count stays on the device);
//               (2) emit instances walking the Gaussians in that order (offsets from a scan taken in
//                   the same order, inside the same kernel);  (3) stable sort of instances by tile id only.
// The tile ranges (identifyTileRanges) and the digit histograms of (3) need no pass over the instances: every
// rectangle adds +1 / -1 at its four corners of a difference array, whose 2-D prefix sum is the list length per tile.
#include "gsr_common.cuh"

namespace gsr {

// ------------------------------------------------------------------ scan + emission (A6 + A7)
constexpr int kScanThreads = 256;
constexpr int kScanItems = GSR_SCAN_TILE / kScanThreads;   // 4, blocked per thread
static_assert(kScanItems == 4, "scan_emit_kernel loads its four ids with one 16-byte request");

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

// One kernel replaces cub::DeviceScan::InclusiveSum (rasterizer_impl.cu:277) AND duplicateWithKeys (:70-111).
//
// Scan: single-pass chained scan (decoupled look-back) of the tile counts in DEPTH order. A CTA takes 1024 visible
// Gaussians by ticket (which makes waiting on lower-numbered tiles safe whatever the CTA scheduling order), gathers
// their packed rectangles through `order` (L2 hits: 4 B x P), scans locally, publishes its sum and looks back. A tile's
// status word carries flag (bits 30-31: 1 = the tile's own sum, 2 = inclusive prefix) and value (30 bits, saturating)
// together, so one 32-bit store publishes both. status[0..m) = tile words, status[m_slots] = grand total,
// status[m_slots + 1] = ticket; all zeroed before the launch.
//
// Emission: the CTA then writes the instances of its own Gaussians -- it knows their offsets, nothing goes through
// memory in between. A warp takes 32 consecutive Gaussians; their output slots are one contiguous range, so the lanes
// walk that range 32 slots at a time (coalesced stores). Owner of slot s = number of Gaussians whose range ends at or
// before s; the ends are strictly increasing, so inside a window of 32 slots they are distinct bits: ONE warp-wide OR
// (REDUX) of those bits plus a running count of the ends before the window gives the owner, four shuffles fetch its
// id, range start, corner and width. Writes are clamped to `cap` (the blob's capacity): with a speculative capacity the
// count may turn out larger, and then the caller repeats the binning.
//
// Tile table input: every Gaussian adds its rectangle to the difference array (4 RED operations).
constexpr uint32_t kScanOwn = 1u << 30, kScanIncl = 2u << 30, kScanVal = (1u << 30) - 1u;
template <bool PACKED>
__global__ void __launch_bounds__(kScanThreads) scan_emit_kernel(const uint32_t* __restrict__ n_dev,
                                                                  const uint32_t* __restrict__ order,
                                                                  const uint32_t* __restrict__ counts,
                                                                  const float4* __restrict__ rec,
                                                                  const int* __restrict__ radii, int gx, int gy, int m_slots,
                                                                  uint32_t* __restrict__ status,
                                                                  uint32_t* __restrict__ point_offsets,
                                                                  int32_t* __restrict__ tile_diff, uint32_t cap,
                                                                  uint32_t* __restrict__ tile_keys,
                                                                  uint32_t* __restrict__ vals) {
    // n (visible Gaussians, in depth order) lives on the device; the grid covers the host's upper bound and m_slots only
    // fixes where the total and the ticket sit. CTAs whose ticket lies beyond the last tile leave at once.
    __shared__ uint32_t warp_sums[kScanThreads / 32];
    __shared__ uint32_t s_tile, s_prefix;
    __shared__ uint32_t s_end[GSR_SCAN_TILE], s_gid[GSR_SCAN_TILE], s_xy[GSR_SCAN_TILE], s_bw[GSR_SCAN_TILE];
    if (threadIdx.x == 0) s_tile = atomicAdd(status + m_slots + 1, 1u);
    __syncthreads();
    const int tile = (int)s_tile;
    const int n = (int)__ldg(n_dev);
    const int m = (n + GSR_SCAN_TILE - 1) / GSR_SCAN_TILE;
    if (tile >= m) return;
    const int tbase = tile * GSR_SCAN_TILE;
    const int base = tbase + threadIdx.x * kScanItems;
    uint32_t g[kScanItems] = {0u, 0u, 0u, 0u}, c[kScanItems], v = 0;
    if (base + kScanItems <= n) {
        const uint4 o4 = __ldg(reinterpret_cast<const uint4*>(order + base));     // base is a multiple of 4, order 128-byte aligned
        g[0] = o4.x; g[1] = o4.y; g[2] = o4.z; g[3] = o4.w;
    } else {
#pragma unroll
        for (int i = 0; i < kScanItems; i++)
            if (base + i < n) g[i] = __ldg(order + base + i);
    }
    const int stride = gx + 1;
#pragma unroll
    for (int i = 0; i < kScanItems; i++) c[i] = (base + i < n) ? __ldg(counts + g[i]) : 0u;
    int x0[kScanItems], y0[kScanItems], bw[kScanItems], bh[kScanItems];
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        x0[i] = 0; y0[i] = 0; bw[i] = 1; bh[i] = 0;
        if (PACKED) {
            const uint32_t r = c[i];
            x0[i] = (int)(r & 255u); y0[i] = (int)((r >> 8) & 255u); bw[i] = (int)((r >> 16) & 255u); bh[i] = (int)(r >> 24);
            c[i] = (uint32_t)(bw[i] * bh[i]);
        } else if (c[i] > 0u) {                         // grids above 255 x 255 tiles: rectangle from record + radius
            int x1, y1;
            const float4 r0 = __ldg(rec + 3 * (size_t)g[i]);
            rect_of(r0.x, r0.y, __ldg(radii + g[i]), gx, gy, x0[i], y0[i], x1, y1);
            bw[i] = x1 - x0[i]; bh[i] = y1 - y0[i];
        }
        v += c[i];
        if (c[i] > 0u) {
            // corners on the far border (x = gx or y = gy) are never read by the prefix sums: skipped
            const int xe = x0[i] + bw[i], ye = y0[i] + bh[i];
            atomicAdd(tile_diff + y0[i] * stride + x0[i], 1);
            if (xe < gx) atomicAdd(tile_diff + y0[i] * stride + xe, -1);
            if (ye < gy) atomicAdd(tile_diff + ye * stride + x0[i], -1);
            if (xe < gx && ye < gy) atomicAdd(tile_diff + ye * stride + xe, 1);
        }
    }
    uint32_t total;
    const uint32_t ex = block_exclusive_scan(v, warp_sums, total);
    volatile uint32_t* vs = status;
    if (threadIdx.x == 0) vs[tile] = (tile == 0 ? kScanIncl : kScanOwn) | total;
    if (threadIdx.x < 32) {
        // look-back, 128 predecessors per round trip: lane l reads tiles j - 4l .. j - 4l - 3 (nearest first).
        const int lane = threadIdx.x;
        // Saturating sums: a status word has 30 value bits, so a running prefix that reaches kScanVal = 2^30 - 1 stays
        // there (sums are formed in 64 bits and clamped) instead of spilling into the flag bits. A total equal to
        // kScanVal therefore means "num_rendered >= 2^30 - 1": gsr_forward answers GSR_E_TOOLARGE, and every consumer of
        // the offsets clamps its writes to the capacity of the binning blob, so nothing is written out of bounds.
        unsigned long long prefix = 0;
        for (int j = tile - 1; j >= 0; j -= 128) {
            unsigned long long part = 0;                   // sum of this lane's words up to (and including) its first inclusive one
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
            unsigned long long val = lane <= first ? part : 0ull;      // <= 4 * 2^30 per lane, 2^37 per warp
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
            prefix += val;
            if (incl) break;
        }
        if (lane == 0) {
            const unsigned long long incl_sum = prefix + total;
            const uint32_t sat = incl_sum >= kScanVal ? kScanVal : (uint32_t)incl_sum;
            if (tile > 0) vs[tile] = kScanIncl | sat;
            if (tile == m - 1) status[m_slots] = sat;
            s_prefix = prefix >= kScanVal ? kScanVal : (uint32_t)prefix;
        }
    }
    __syncthreads();
    const uint32_t tile_start = s_prefix;
    uint32_t run = ex + tile_start;
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        run += c[i];
        const int k = threadIdx.x * kScanItems + i;
        s_end[k] = run;
        s_gid[k] = g[i];
        s_xy[k] = (uint32_t)x0[i] | ((uint32_t)y0[i] << 16);
        s_bw[k] = (uint32_t)max(1, bw[i]);
        if (base + i < n) point_offsets[base + i] = run;   // inclusive, as rasterizer_impl.cu:277 (in depth order)
    }
    __syncthreads();
    if (tile_start >= cap) return;                         // everything this tile would write lies beyond the capacity
    // ---- emission: warp w takes Gaussians w*128 + r*32 + lane of the tile
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll 1
    for (int r = 0; r < GSR_SCAN_TILE / kScanThreads; r++) {
        const int k = w * (GSR_SCAN_TILE / 8) + r * 32 + lane;
        const int j = tbase + k;
        if (j - lane >= n) break;
        uint32_t gid = 0, end = 0, start = 0, xy = 0;
        int obw_mine = 1;
        if (j < n) {
            gid = s_gid[k]; end = s_end[k]; xy = s_xy[k]; obw_mine = (int)s_bw[k];
            start = k > 0 ? s_end[k - 1] : tile_start;
        }
        // lanes beyond n: end = start = end of the last valid lane, so they never own a slot that is written
        const uint32_t last_end = __shfl_sync(0xffffffffu, end, min(31, n - 1 - (j - lane)));
        if (j >= n) { end = last_end; start = last_end; }
        const uint32_t first = __shfl_sync(0xffffffffu, start, 0);
        const uint32_t stop = min(last_end, cap);
        uint32_t before = 0;                            // ends below the window's first slot (warp-uniform)
        for (uint32_t wb = first; wb < stop; wb += 32) {
            const uint32_t rel = end - wb;              // wraps to a huge value for ends below the window
            const uint32_t mine = (rel < 32u && j < n) ? (1u << rel) : 0u;
            const uint32_t ends = __reduce_or_sync(0xffffffffu, mine);
            const uint32_t s = wb + lane;
            const int owner = min((int)(before + __popc(ends & (0xffffffffu >> (31 - lane)))), 31);
            before += __popc(ends);
            const uint32_t og = __shfl_sync(0xffffffffu, gid, owner);
            const uint32_t ostart = __shfl_sync(0xffffffffu, start, owner);
            const uint32_t oxy = __shfl_sync(0xffffffffu, xy, owner);
            const int obw = __shfl_sync(0xffffffffu, obw_mine, owner);
            if (s < stop) {
                const uint32_t q = s - ostart;                             // row-major index inside the rectangle
                const int row = (int)(((float)q + 0.5f) * __frcp_rn((float)obw));   // exact: the quotient stays >= 0.5 / 4096 away
                const int col = (int)q - row * obw;                        //        from an integer, the product errs by < 1e-3 of that
                tile_keys[s] = (uint32_t)(((int)(oxy >> 16) + row) * gx + (int)(oxy & 0xffffu) + col);
                vals[s] = og;
            }
        }
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

int launch_depth_order(const gsr_view& v, GeomState& g, cudaStream_t s) {
    const int P = v.P;
    bool in_a = true;
    // one memset clears the depth sort's tables and the instance counter: they are carved next to each other
    const char* zero_end = reinterpret_cast<const char*>(g.count64 + 2);
    GSR_MEMSET("memset_depth_sort", s, g.sort_scratch, 0, (size_t)(zero_end - reinterpret_cast<const char*>(g.sort_scratch)));
    uint32_t* n_vis = nullptr;
    int e = radix_sort_u32(g.dkey, g.order, g.dkey_tmp, g.order_tmp, P, 32, g.sort_scratch, &in_a, v.debug != 0, s, &n_vis,
                           /*scratch_zeroed=*/true);
    if (e) return e;
    // 4 passes: the result is back in (dkey, order); its first *n_vis entries are the visible Gaussians
    if (n_vis != g.n_visible) return GSR_E_BADARG;              // layout contract between the two files
    return 0;
}

// ------------------------------------------------------------------ tile table (A9 + the histograms of A8)
// One CTA. The difference array's 2-D prefix sum is the number of instances per tile; their exclusive scan in tile order
// gives the ranges (identifyTileRanges, rasterizer_impl.cu:116-138: empty tiles keep (0, 0), :310), and the counts binned
// by the digits of the tile id are the histograms the instance sort's passes start from -- no pass over the instances.
constexpr int kTableThreads = 1024;
constexpr int kTableSmemMax = 200 * 1024;
__global__ void __launch_bounds__(kTableThreads) tile_table_kernel(int32_t* __restrict__ diff, int gx, int gy, int passes,
                                                                    int total_bits, uint2* __restrict__ ranges,
                                                                    uint32_t* __restrict__ ghist, int use_smem) {
    extern __shared__ __align__(16) int32_t table_smem[];
    __shared__ uint32_t hist[4][GSR_RADIX];
    __shared__ uint32_t warp_sums[kTableThreads / 32];
    const int t = threadIdx.x, stride = gx + 1, cells = stride * (gy + 1);
    int32_t* work = use_smem ? table_smem : diff;          // large grids: in place in global memory (slow path, correct)
    if (use_smem)
        for (int i = t; i < cells; i += kTableThreads) work[i] = diff[i];
    for (int i = t; i < 4 * GSR_RADIX; i += kTableThreads) (&hist[0][0])[i] = 0u;
    __syncthreads();
    for (int x = t; x < gx; x += kTableThreads) {          // prefix down every column
        int32_t run = 0;
        for (int y = 0; y < gy; y++) { run += work[y * stride + x]; work[y * stride + x] = run; }
    }
    __syncthreads();
    for (int y = t; y < gy; y += kTableThreads) {          // prefix along every row: work[y][x] = instances of tile (x, y)
        int32_t run = 0;
        for (int x = 0; x < gx; x++) { run += work[y * stride + x]; work[y * stride + x] = run; }
    }
    __syncthreads();
    const int tiles = gx * gy;
    const int per = (tiles + kTableThreads - 1) / kTableThreads;
    const int lo = min(t * per, tiles), hi = min(lo + per, tiles);
    uint32_t local = 0;
    for (int k = lo; k < hi; k++) local += (uint32_t)work[(k / gx) * stride + (k % gx)];
    // exclusive scan of `local` over the 1024 threads
    const int lane = t & 31, w = t >> 5;
    uint32_t incl = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
    }
    if (lane == 31) warp_sums[w] = incl;
    __syncthreads();
    uint32_t ofs = incl - local;
    for (int k = 0; k < w; k++) ofs += warp_sums[k];
    for (int k = lo; k < hi; k++) {
        const uint32_t c = (uint32_t)work[(k / gx) * stride + (k % gx)];
        ranges[k] = c ? make_uint2(ofs, ofs + c) : make_uint2(0u, 0u);
        ofs += c;
        if (c) {
            for (int p = 0; p < passes; p++) {
                const int bits = min(8, total_bits - 8 * p);
                atomicAdd(&hist[p][((uint32_t)k >> (8 * p)) & ((1u << bits) - 1u)], c);
            }
        }
    }
    __syncthreads();
    for (int i = t; i < passes * GSR_RADIX; i += kTableThreads) ghist[i] = (&hist[0][0])[i];
}

static int bits_for(uint32_t count) {   // bits needed to represent ids 0..count-1
    int b = 0;
    while (b < 32 && (1ull << b) < count) b++;
    return b;
}

int launch_binning(const gsr_view& v, GeomState& g, const int* radii, BinningState& b, ImageState& img, uint32_t cap,
                   uint32_t** total_dev, cudaStream_t s) {
    const int gx = (v.W + GSR_TILE - 1) / GSR_TILE, gy = (v.H + GSR_TILE - 1) / GSR_TILE;
    const int tiles = gx * gy;
    const int tbits = bits_for((uint32_t)tiles);
    const int passes = radix_passes(tbits);
    if (passes > 4) return GSR_E_TOOLARGE;
    const int m = (int)BinningState::scan_tiles((size_t)v.P);
    *total_dev = b.scan_sums + m;
    GSR_MEMSET("memset_binning", s, b.zero_begin, 0, (size_t)(b.zero_end - b.zero_begin));
    // make the sorted result land in point_list
    uint32_t *ka = b.tile_keys, *va = b.point_list, *kb = b.tile_keys_tmp, *vb = b.vals_tmp;
    if (passes % 2) { ka = b.tile_keys_tmp; va = b.vals_tmp; kb = b.tile_keys; vb = b.point_list; }
    if (cap > 0) {
        if (rect_packing_ok(v.W, v.H))
            GSR_LAUNCH("scan_emit", v.debug, s,
                       (scan_emit_kernel<true><<<m, kScanThreads, 0, s>>>(g.n_visible, g.order, g.rect_packed, g.rec, radii, gx, gy, m,
                                                                         b.scan_sums, g.point_offsets, b.tile_diff, cap, ka, va)));
        else
            GSR_LAUNCH("scan_emit", v.debug, s,
                       (scan_emit_kernel<false><<<m, kScanThreads, 0, s>>>(g.n_visible, g.order, g.tiles_touched, g.rec, radii, gx, gy, m,
                                                                          b.scan_sums, g.point_offsets, b.tile_diff, cap, ka, va)));
    }
    const size_t table_bytes = BinningState::diff_cells(v.W, v.H) * sizeof(int32_t);
    const int use_smem = table_bytes <= (size_t)kTableSmemMax ? 1 : 0;
    static DeviceOnce once;
    { const int e = ensure_dynamic_smem(tile_table_kernel, kTableSmemMax, once); if (e) return e; }
    GSR_LAUNCH("tile_table", v.debug, s,
               (tile_table_kernel<<<1, kTableThreads, use_smem ? table_bytes : 0, s>>>(b.tile_diff, gx, gy, passes, tbits, img.ranges,
                                                                                     b.sort_scratch, use_smem)));
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
