// gsr_sort.cu -- hand-written stable LSD radix sort for (u32 key, u32 value) pairs, 8-bit digits.
//
// Replaces cub::DeviceRadixSort::SortPairs (rasterizer_impl.cu:303-308, simple_knn.cu:211-214).
// The rasterizer no longer sorts 64-bit (tile|depth) keys over all R instances (6 passes of 24 B).
// It sorts the P Gaussians by depth bits ONCE (4 passes over P items), emits the instances in that
// order, and then needs only a stable sort by tile id (2 passes at 1080p) -- same final order as
// the reference's 45-bit sort, ~4x fewer bytes moved ("Binning").
//
// Per pass, three launches, no spinning and no inter-block waits (deterministic, hang-free):
//   digit_hist   : per sort-tile (4096 keys) digit counts -> table[digit][tile], totals[digit]
//   digit_scan   : block d: base(d) = sum totals[<d]; exclusive scan of row d in place
//   digit_scatter: re-rank the tile's keys stably (warp match + per-warp counters), reorder them in
//                  shared memory so that equal digits are contiguous, write coalesced runs.
// HBM bytes per pass and item: 4 (hist) + 8 (load) + 8 (store) = 20; the last pass may skip keys.
#include "gsr_common.cuh"

namespace gsr {

constexpr int kSortThreads = 256;
constexpr int kSortItems = GSR_SORT_TILE / kSortThreads;   // 16
constexpr int kSortWarps = kSortThreads / 32;               // 8

__global__ void __launch_bounds__(kSortThreads) digit_hist_kernel(const uint32_t* __restrict__ keys, int n, int shift,
                                                                   uint32_t mask, uint32_t* __restrict__ table,
                                                                   uint32_t* __restrict__ totals, int tiles) {
    __shared__ uint32_t hist[GSR_RADIX];
    const int tile = blockIdx.x;
    hist[threadIdx.x] = 0;
    __syncthreads();
    const int base = tile * GSR_SORT_TILE;
#pragma unroll
    for (int i = 0; i < kSortItems; i++) {
        const int idx = base + i * kSortThreads + threadIdx.x;
        if (idx < n) atomicAdd(&hist[(__ldg(keys + idx) >> shift) & mask], 1u);
    }
    __syncthreads();
    const uint32_t c = hist[threadIdx.x];
    table[(size_t)threadIdx.x * tiles + tile] = c;
    if (c) atomicAdd(&totals[threadIdx.x], c);
}

// grid = 256 blocks (one per digit), 256 threads
__global__ void __launch_bounds__(256) digit_scan_kernel(uint32_t* __restrict__ table, const uint32_t* __restrict__ totals,
                                                          int tiles) {
    __shared__ uint32_t warp_sums[8];
    __shared__ uint32_t s_carry;
    const int d = blockIdx.x, t = threadIdx.x, lane = t & 31, w = t >> 5;
    // base = sum of totals of smaller digits
    uint32_t v = (t < d) ? totals[t] : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) warp_sums[w] = v;
    __syncthreads();
    if (t == 0) {
        uint32_t b = 0;
        for (int i = 0; i < 8; i++) b += warp_sums[i];
        s_carry = b;
    }
    __syncthreads();
    uint32_t* row = table + (size_t)d * tiles;
    for (int c0 = 0; c0 < tiles; c0 += 256) {
        const int i = c0 + t;
        const uint32_t x = (i < tiles) ? row[i] : 0u;
        uint32_t incl = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += y;
        }
        if (lane == 31) warp_sums[w] = incl;
        __syncthreads();
        uint32_t woff = 0;
        for (int k = 0; k < w; k++) woff += warp_sums[k];
        const uint32_t carry = s_carry;
        if (i < tiles) row[i] = carry + woff + incl - x;
        __syncthreads();
        if (t == 255) s_carry = carry + woff + incl;
        __syncthreads();
    }
}

template <bool WRITE_KEYS>
__global__ void __launch_bounds__(kSortThreads) digit_scatter_kernel(const uint32_t* __restrict__ keys_in,
                                                                      const uint32_t* __restrict__ vals_in,
                                                                      uint32_t* __restrict__ keys_out,
                                                                      uint32_t* __restrict__ vals_out, int n, int shift,
                                                                      uint32_t mask, const uint32_t* __restrict__ table,
                                                                      int tiles) {
    __shared__ uint32_t whist[kSortWarps][GSR_RADIX + 1];   // per-warp digit counters, then warp offsets
    __shared__ uint32_t dstart[GSR_RADIX + 1];              // local start of each digit's run
    __shared__ uint32_t gbase[GSR_RADIX];                   // global start of this tile's run per digit
    __shared__ uint32_t s_keys[GSR_SORT_TILE];
    __shared__ uint32_t s_vals[GSR_SORT_TILE];
    __shared__ uint32_t warp_sums[kSortWarps];
    const int tile = blockIdx.x, t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int base = tile * GSR_SORT_TILE;
    const int count = min(GSR_SORT_TILE, n - base);
    for (int i = t; i < kSortWarps * (GSR_RADIX + 1); i += kSortThreads) (&whist[0][0])[i] = 0;
    gbase[t] = table[(size_t)t * tiles + tile];
    __syncthreads();

    // warp-striped load: order index inside the tile = w*512 + i*32 + lane
    uint32_t key[kSortItems], val[kSortItems], rank[kSortItems];
    const int wbase = base + w * (32 * kSortItems);
#pragma unroll
    for (int i = 0; i < kSortItems; i++) {
        const int idx = wbase + i * 32 + lane;
        const bool ok = idx < n;
        key[i] = ok ? __ldg(keys_in + idx) : 0xFFFFFFFFu;
        val[i] = ok ? __ldg(vals_in + idx) : 0u;
    }
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < kSortItems; i++) {
        const int idx = wbase + i * 32 + lane;
        const uint32_t d = (idx < n) ? ((key[i] >> shift) & mask) : GSR_RADIX;   // 256 = "not an item"
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const uint32_t lower = __popc(peers & lt_mask);
        const uint32_t prev = whist[w][d];
        rank[i] = prev + lower;
        __syncwarp();
        if (lower == 0) whist[w][d] = prev + __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    // digit t: exclusive scan over warps; block count
    uint32_t run = 0;
#pragma unroll
    for (int k = 0; k < kSortWarps; k++) {
        const uint32_t c = whist[k][t];
        whist[k][t] = run;
        run += c;
    }
    // exclusive scan over digits -> dstart
    uint32_t incl = run;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += y;
    }
    if (lane == 31) warp_sums[w] = incl;
    __syncthreads();
    uint32_t woff = 0;
    for (int k = 0; k < w; k++) woff += warp_sums[k];
    dstart[t] = woff + incl - run;
    if (t == 255) dstart[GSR_RADIX] = woff + incl;
    __syncthreads();
    // place into shared memory at the local sorted position
#pragma unroll
    for (int i = 0; i < kSortItems; i++) {
        const int idx = wbase + i * 32 + lane;
        if (idx < n) {
            const uint32_t d = (key[i] >> shift) & mask;
            const uint32_t lp = dstart[d] + whist[w][d] + rank[i];
            s_keys[lp] = key[i];
            s_vals[lp] = val[i];
        }
    }
    __syncthreads();
    // coalesced runs out
#pragma unroll
    for (int i = 0; i < kSortItems; i++) {
        const int slot = i * kSortThreads + t;
        if (slot < count) {
            const uint32_t k = s_keys[slot];
            const uint32_t d = (k >> shift) & mask;
            const uint32_t pos = gbase[d] + (slot - dstart[d]);
            if (WRITE_KEYS) keys_out[pos] = k;
            vals_out[pos] = s_vals[slot];
        }
    }
}

int radix_sort_u32(uint32_t* ka, uint32_t* va, uint32_t* kb, uint32_t* vb, int n, int total_bits, uint32_t* scratch,
                   bool* result_in_a, bool debug, cudaStream_t s) {
    *result_in_a = true;
    if (n <= 0 || total_bits <= 0) return 0;
    const int passes = radix_passes(total_bits);
    const int tiles = (int)sort_tiles((size_t)n);
    uint32_t* totals = scratch;                 // [8][256]
    uint32_t* table = scratch + 8 * GSR_RADIX;  // [256][tiles]
    cudaError_t e = cudaMemsetAsync(totals, 0, sizeof(uint32_t) * 8 * GSR_RADIX, s);
    if (e != cudaSuccess) return (int)e;
    uint32_t *kin = ka, *vin = va, *kout = kb, *vout = vb;
    for (int p = 0; p < passes; p++) {
        const int shift = 8 * p;
        const int bits = (total_bits - shift) < 8 ? (total_bits - shift) : 8;
        const uint32_t mask = (1u << bits) - 1u;
        GSR_LAUNCH("digit_hist", debug, s,
                   (digit_hist_kernel<<<tiles, kSortThreads, 0, s>>>(kin, n, shift, mask, table, totals + p * GSR_RADIX, tiles)));
        GSR_LAUNCH("digit_scan", debug, s, (digit_scan_kernel<<<GSR_RADIX, 256, 0, s>>>(table, totals + p * GSR_RADIX, tiles)));
        GSR_LAUNCH("digit_scatter", debug, s,
                   (digit_scatter_kernel<true><<<tiles, kSortThreads, 0, s>>>(kin, vin, kout, vout, n, shift, mask, table, tiles)));
        uint32_t* tk = kin; kin = kout; kout = tk;
        uint32_t* tv = vin; vin = vout; vout = tv;
    }
    *result_in_a = (passes % 2) == 0;
    return 0;
}

}  // namespace gsr
