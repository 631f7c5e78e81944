// gsr_binning.cu -- depth ordering of Gaussians, tile-count scan, instance emission, tile sort,
// tile ranges. Replaces rasterizer_impl.cu:70-138 (duplicateWithKeys, identifyTileRanges) and the
// CUB calls at :277 (InclusiveSum) and :303 (SortPairs). HBM-bound integer work.
//
// Order of the final point list (must equal the reference's stable 64-bit sort):
//   by tile id, then by depth bits, then by Gaussian id.
// We get it as: (1) stable sort of Gaussian ids by depth bits (P items, 4 passes);
//               (2) emit instances walking the Gaussians in that order (offsets from a scan taken in
//                   the same order);  (3) stable sort of instances by tile id only.
#include "gsr_common.cuh"

namespace gsr {

// ------------------------------------------------------------------ scan (3 launches, no spinning)
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

__global__ void __launch_bounds__(kScanThreads) scan_reduce_kernel(const uint32_t* __restrict__ counts,
                                                                    const uint32_t* __restrict__ order, int n,
                                                                    uint32_t* __restrict__ sums) {
    __shared__ uint32_t warp_sums[kScanThreads / 32];
    const int base = blockIdx.x * GSR_SCAN_TILE + threadIdx.x * kScanItems;
    uint32_t v = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; i++)
        if (base + i < n) v += __ldg(counts + __ldg(order + base + i));
    uint32_t total;
    block_exclusive_scan(v, warp_sums, total);
    if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

// single block: exclusive scan of sums[0..m) in place; sums[m] = grand total
__global__ void __launch_bounds__(kScanThreads) scan_spine_kernel(uint32_t* __restrict__ sums, int m) {
    __shared__ uint32_t warp_sums[kScanThreads / 32];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (int c0 = 0; c0 < m; c0 += kScanThreads) {
        const int i = c0 + threadIdx.x;
        const uint32_t x = i < m ? sums[i] : 0u;
        uint32_t total;
        const uint32_t ex = block_exclusive_scan(x, warp_sums, total);
        const uint32_t carry = s_carry;
        if (i < m) sums[i] = carry + ex;
        __syncthreads();
        if (threadIdx.x == 0) s_carry = carry + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) sums[m] = s_carry;
}

__global__ void __launch_bounds__(kScanThreads) scan_apply_kernel(const uint32_t* __restrict__ counts,
                                                                   const uint32_t* __restrict__ order, int n,
                                                                   const uint32_t* __restrict__ sums,
                                                                   uint32_t* __restrict__ out) {
    __shared__ uint32_t warp_sums[kScanThreads / 32];
    const int base = blockIdx.x * GSR_SCAN_TILE + threadIdx.x * kScanItems;
    uint32_t c[kScanItems], v = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        c[i] = (base + i < n) ? __ldg(counts + __ldg(order + base + i)) : 0u;
        v += c[i];
    }
    uint32_t total;
    uint32_t run = block_exclusive_scan(v, warp_sums, total) + sums[blockIdx.x];
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        run += c[i];
        if (base + i < n) out[base + i] = run;   // inclusive, as rasterizer_impl.cu:277
    }
}

int launch_order_and_scan(const gsr_view& v, GeomState& g, uint32_t** total_dev, cudaStream_t s) {
    const int P = v.P;
    bool in_a = true;
    int e = radix_sort_u32(g.dkey, g.order, g.dkey_tmp, g.order_tmp, P, 32, g.sort_scratch, &in_a, v.debug != 0, s);
    if (e) return e;
    // 4 passes: the result is back in (dkey, order)
    const int m = (P + GSR_SCAN_TILE - 1) / GSR_SCAN_TILE;
    GSR_LAUNCH("scan_reduce", v.debug, s,
               (scan_reduce_kernel<<<m, kScanThreads, 0, s>>>(g.tiles_touched, g.order, P, g.scan_sums)));
    GSR_LAUNCH("scan_spine", v.debug, s, (scan_spine_kernel<<<1, kScanThreads, 0, s>>>(g.scan_sums, m)));
    GSR_LAUNCH("scan_apply", v.debug, s,
               (scan_apply_kernel<<<m, kScanThreads, 0, s>>>(g.tiles_touched, g.order, P, g.scan_sums, g.point_offsets)));
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

constexpr int kDupSerialMax = 12;   // rects up to this many tiles are written by their own thread

__global__ void __launch_bounds__(256) emit_instances_kernel(int P, const uint32_t* __restrict__ order,
                                                              const uint32_t* __restrict__ tiles_touched,
                                                              const uint32_t* __restrict__ offsets,
                                                              const float4* __restrict__ rec,
                                                              const int* __restrict__ radii, int gx, int gy,
                                                              uint32_t* __restrict__ tile_keys,
                                                              uint32_t* __restrict__ vals) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    uint32_t g = 0, n = 0, off = 0;
    int x0 = 0, y0 = 0, x1 = 0, y1 = 0;
    if (j < P) {
        g = __ldg(order + j);
        n = __ldg(tiles_touched + g);
        if (n > 0) {
            off = __ldg(offsets + j) - n;
            const float4 r0 = __ldg(rec + 3 * (size_t)g);
            rect_of(r0.x, r0.y, __ldg(radii + g), gx, gy, x0, y0, x1, y1);
        }
    }
    if (n > 0 && n <= kDupSerialMax) {
        for (int y = y0; y < y1; y++)
            for (int x = x0; x < x1; x++) {
                tile_keys[off] = (uint32_t)(y * gx + x);
                vals[off] = g;
                off++;
            }
    }
    // large rectangles: the whole warp writes them, one rectangle at a time
    uint32_t big = __ballot_sync(0xffffffffu, n > kDupSerialMax);
    while (big) {
        const int src = __ffs(big) - 1;
        big &= big - 1;
        const uint32_t bg = __shfl_sync(0xffffffffu, g, src);
        const uint32_t bn = __shfl_sync(0xffffffffu, n, src);
        const uint32_t boff = __shfl_sync(0xffffffffu, off, src);
        const int bx0 = __shfl_sync(0xffffffffu, x0, src), by0 = __shfl_sync(0xffffffffu, y0, src);
        const int bw = __shfl_sync(0xffffffffu, x1, src) - bx0;
        for (uint32_t k = lane; k < bn; k += 32) {
            const int y = by0 + (int)(k / (uint32_t)bw), x = bx0 + (int)(k % (uint32_t)bw);
            tile_keys[boff + k] = (uint32_t)(y * gx + x);
            vals[boff + k] = bg;
        }
    }
}

// ------------------------------------------------------------------ tile ranges (A9)
__global__ void __launch_bounds__(256) tile_ranges_kernel(int L, const uint32_t* __restrict__ tile_keys,
                                                           uint2* __restrict__ ranges) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= L) return;
    const uint32_t cur = tile_keys[idx];
    if (idx == 0) ranges[cur].x = 0;
    else {
        const uint32_t prev = tile_keys[idx - 1];
        if (cur != prev) {
            ranges[prev].y = idx;
            ranges[cur].x = idx;
        }
    }
    if (idx == L - 1) ranges[cur].y = L;
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
    GSR_LAUNCH("emit_instances", v.debug, s,
               (emit_instances_kernel<<<(v.P + 255) / 256, 256, 0, s>>>(v.P, g.order, g.tiles_touched, g.point_offsets,
                                                                        g.rec, radii, gx, gy, ka, va)));
    bool in_a = true;
    int e = radix_sort_u32(ka, va, kb, vb, R, tbits, b.sort_scratch, &in_a, v.debug != 0, s);
    if (e) return e;
    GSR_LAUNCH("tile_ranges", v.debug, s, (tile_ranges_kernel<<<(R + 255) / 256, 256, 0, s>>>(R, b.tile_keys, img.ranges)));
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
