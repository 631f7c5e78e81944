// gsr_sort.cu -- hand-written stable LSD radix sort for (u32 key, u32 value) pairs, 8-bit digits, onesweep style.
//
// Replaces cub::DeviceRadixSort::SortPairs (rasterizer_impl.cu:303-308, simple_knn.cu:211-214).
// The rasterizer no longer sorts 64-bit (tile|depth) keys over all R instances (6 passes of 24 B).
// It sorts the P Gaussians by depth bits ONCE (4 passes over P items), emits the instances in that
// order, and then needs only a stable sort by tile id (2 passes at 1080p) -- same final order as
// the reference's 45-bit sort, ~4x fewer bytes moved ("Binning").
//
// Launches per sort: one memset (histograms, tickets, look-back status), ONE histogram kernel for the digits
// of every pass, then ONE kernel per pass:
//   * a tile of 4096 pairs per CTA, handed out by an atomic ticket (so a CTA only ever waits for tiles whose
//     CTAs are already resident: the look-back cannot dead-lock);
//   * stable ranks from per-bit ballots + per-warp shared counters (no block-wide atomics);
//   * the tile's per-digit totals are published, the pairs are reordered in shared memory, and the global
//     position of each digit run comes from a decoupled look-back over the predecessors' status words
//     (flag + value in one 32-bit word, relaxed gpu-scope accesses);
//   * runs are written out coalesced.
// HBM bytes per pair: 4 (histogram, once) + 16 per pass (read 8, write 8).
#include "gsr_common.cuh"

namespace gsr {

constexpr int kSortThreads = 256;
constexpr int kSortItems = GSR_SORT_TILE / kSortThreads;   // 16
#ifndef GSR_SORT_CTAS
#define GSR_SORT_CTAS (GSR_SORT_TILE >= 4096 ? 3 : 5)
#endif
constexpr int kSortCtasPerSm = GSR_SORT_CTAS;
constexpr int kSortWarps = kSortThreads / 32;               // 8
constexpr uint32_t kFlagAgg = 1u << 30, kFlagInc = 2u << 30, kValMask = (1u << 30) - 1u;
constexpr unsigned kAll = 0xffffffffu;
constexpr int kHistCopies = 8;

// lanes of the warp holding the same BITS-bit digit as this lane: one ballot per bit, folded with a 3-input logic op.
// (MATCH.ANY does this in one instruction but measured ~48 cycles per warp-instruction per SM on B200.)
template <int BITS>
__device__ __forceinline__ uint32_t same_digit_lanes(uint32_t d, uint32_t peers) {
#pragma unroll
    for (int b = 0; b < BITS; b++) {
        const uint32_t m = 0u - ((d >> b) & 1u);                   // all ones if my bit is set
        const uint32_t bal = __ballot_sync(kAll, m != 0u);
        peers &= ~(bal ^ m);                                       // keep lanes whose bit equals mine
    }
    return peers;
}

__device__ __forceinline__ uint32_t ld_relaxed(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// exclusive scan of one value per thread over the 256 threads of the CTA
__device__ __forceinline__ uint32_t block_excl_scan256(uint32_t v, uint32_t* warp_sums) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(kAll, incl, o);
        if (lane >= o) incl += y;
    }
    if (lane == 31) warp_sums[w] = incl;
    __syncthreads();
    uint32_t woff = 0;
#pragma unroll
    for (int k = 0; k < kSortWarps; k++)
        if (k < w) woff += warp_sums[k];
    __syncthreads();
    return woff + incl - v;
}

// ------------------------------------------------------------------ digit histograms of every pass, one read of the keys
__global__ void __launch_bounds__(kSortThreads) radix_hist_kernel(const uint32_t* __restrict__ keys, int n, int passes,
                                                                   int total_bits, uint32_t* __restrict__ ghist, int drop) {
    // kHistCopies replicas per digit, picked by lane: depth exponents put most keys of a warp into a few digits of
    // the top byte, and same-address shared atomics serialise (MATCH.ANY aggregation measured slower than this)
    __shared__ uint32_t sh[4][kHistCopies][GSR_RADIX];
    const int t = threadIdx.x, lane = t & 31, cp = lane & (kHistCopies - 1);
    for (int i = t; i < 4 * kHistCopies * GSR_RADIX; i += kSortThreads) (&sh[0][0][0])[i] = 0;
    __syncthreads();
    const int nvec = n >> 2;                                       // keys is 128-byte aligned (carved scratch)
    const uint4* k4 = reinterpret_cast<const uint4*>(keys);
    const int stride = gridDim.x * kSortThreads;
    constexpr int kU = 4;                                          // 16-byte loads in flight per thread
    for (int i0 = blockIdx.x * kSortThreads + t; i0 < nvec; i0 += stride * kU) {
        uint4 v[kU];
#pragma unroll
        for (int u = 0; u < kU; u++)
            v[u] = (i0 + u * stride < nvec) ? __ldg(k4 + i0 + u * stride) : make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
#pragma unroll
        for (int u = 0; u < kU; u++) {
            if (i0 + u * stride >= nvec) continue;
            const uint32_t kk[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
            for (int j = 0; j < 4; j++) {
                if (drop && kk[j] == 0xFFFFFFFFu) continue;        // dropped keys are in no histogram
                for (int p = 0; p < passes; p++) {
                    const int bits = min(8, total_bits - 8 * p);
                    atomicAdd(&sh[p][cp][(kk[j] >> (8 * p)) & ((1u << bits) - 1u)], 1u);
                }
            }
        }
    }
    if (blockIdx.x == 0 && t < (n & 3)) {                          // tail (< 4 keys)
        const uint32_t k = __ldg(keys + (nvec << 2) + t);
        for (int p = 0; p < (drop && k == 0xFFFFFFFFu ? 0 : passes); p++) {
            const int bits = min(8, total_bits - 8 * p);
            atomicAdd(&sh[p][0][(k >> (8 * p)) & ((1u << bits) - 1u)], 1u);
        }
    }
    __syncthreads();
    for (int p = 0; p < passes; p++) {
        uint32_t c = 0;
#pragma unroll
        for (int k = 0; k < kHistCopies; k++) c += sh[p][k][t];
        if (c) atomicAdd(&ghist[p * GSR_RADIX + t], c);
    }
}

// ------------------------------------------------------------------ one pass
// Persistent CTAs (3 per SM): a CTA draws tile tickets in a loop and cp.async-prefetches the NEXT tile's pairs into
// shared memory while it ranks / reorders / writes the current one, so global loads are in flight all the time
// (a single-tile-per-CTA version spent half its cycles waiting for its own loads: 1.5 TB/s).
struct SortSmem {
    uint32_t pf_keys[GSR_SORT_TILE];         // prefetched pairs of the tile about to be processed (cp.async)
    uint32_t pf_vals[GSR_SORT_TILE];
    uint32_t s_keys[GSR_SORT_TILE];          // reorder buffers
    uint32_t s_vals[GSR_SORT_TILE];
    uint32_t whist[kSortWarps][GSR_RADIX];   // per-warp digit counters, then local start of (warp, digit)
    uint32_t gofs[GSR_RADIX];                // global start of the digit's run minus its local start
    uint32_t warp_sums[kSortWarps];
    uint32_t tile, next;
    uint32_t n_valid, count_out;             // drop mode: keys that survive (whole sort / this tile)
};

__device__ __forceinline__ void sort_cp16(void* smem, const void* gmem) {
    const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void sort_cp4(void* smem, const void* gmem) {
    const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(gmem) : "memory");
}
// all threads: start copying tile `tile` into the prefetch buffers (no wait)
__device__ __forceinline__ void sort_prefetch(SortSmem& sm, const uint32_t* __restrict__ keys_in,
                                              const uint32_t* __restrict__ vals_in, int n, int tile) {
    const int t = threadIdx.x, base = tile * GSR_SORT_TILE;
    if (base + GSR_SORT_TILE <= n) {                      // tile bases are 16 KB apart: 16-byte aligned
#pragma unroll
        for (int i = 0; i < GSR_SORT_TILE / 4 / kSortThreads; i++) {
            const int q = (i * kSortThreads + t) * 4;
            sort_cp16(&sm.pf_keys[q], keys_in + base + q);
            sort_cp16(&sm.pf_vals[q], vals_in + base + q);
        }
    } else {
        for (int q = t; base + q < n; q += kSortThreads) {
            sort_cp4(&sm.pf_keys[q], keys_in + base + q);
            sort_cp4(&sm.pf_vals[q], vals_in + base + q);
        }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}

// ranks, publishes, reorders and writes out one tile whose pairs are already in registers
template <int BITS, bool FULL, bool DROP, bool WRITE_KEYS>
__device__ __forceinline__ void onesweep_tile(SortSmem& sm, const uint32_t (&key)[kSortItems], const uint32_t (&val)[kSortItems],
                                              uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int n,
                                              int shift, int tile, uint32_t gbase, uint32_t* __restrict__ status) {
    constexpr uint32_t mask = (1u << BITS) - 1u;
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int base = tile * GSR_SORT_TILE;
    const int count = FULL ? GSR_SORT_TILE : n - base;
    // order index inside the tile = w*512 + i*32 + lane (warp-striped)
    uint32_t rank[kSortItems];
    const int wofs = w * (32 * kSortItems) + lane;
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < kSortItems; i++) {
        const bool ok = (FULL || wofs + i * 32 < count) && (!DROP || key[i] != 0xFFFFFFFFu);
        const uint32_t d = (key[i] >> shift) & mask;
        const uint32_t peers = same_digit_lanes<BITS>(d, (FULL && !DROP) ? kAll : __ballot_sync(kAll, ok));
        const int leader = __ffs(peers) - 1;                       // invalid lanes: garbage, never used
        uint32_t prev = 0;
        if (ok && lane == leader) prev = atomicAdd(&sm.whist[w][d], (uint32_t)__popc(peers));   // same warp: program order
        prev = __shfl_sync(kAll, prev, leader & 31);
        rank[i] = prev + __popc(peers & lt_mask);
    }
    __syncthreads();
    // digit t: totals over the warps -> publish -> local start of every (warp, digit) run
    uint32_t cnt[kSortWarps], run = 0;
#pragma unroll
    for (int k = 0; k < kSortWarps; k++) { cnt[k] = sm.whist[k][t]; run += cnt[k]; }
    uint32_t* my_status = status + (size_t)tile * GSR_RADIX + t;
    st_relaxed(my_status, (tile == 0 ? kFlagInc : kFlagAgg) | run);            // publish early: successors can proceed
    const uint32_t ds = block_excl_scan256(run, sm.warp_sums);
    if (DROP && t == kSortThreads - 1) sm.count_out = ds + run;                // pairs of this tile that are kept
    uint32_t acc = ds;
#pragma unroll
    for (int k = 0; k < kSortWarps; k++) { sm.whist[k][t] = acc; acc += cnt[k]; }
    __syncthreads();
    // reorder inside shared memory
#pragma unroll
    for (int i = 0; i < kSortItems; i++) {
        if ((FULL || wofs + i * 32 < count) && (!DROP || key[i] != 0xFFFFFFFFu)) {
            const uint32_t lp = sm.whist[w][(key[i] >> shift) & mask] + rank[i];
            sm.s_keys[lp] = key[i];
            sm.s_vals[lp] = val[i];
        }
    }
    // decoupled look-back for digit t, kLookWin predecessors per round trip: the status words of a window are loaded
    // together (independent L2 reads in flight) and then consumed in order; a serial walk paid one L2 latency per
    // predecessor and was a third of the pass
    uint32_t excl = 0;
    if (tile > 0) {
        // (measured and not adopted: a window of 32 -- most tiles find an inclusive prefix among their nearest predecessors,
        // the 24 extra status loads per thread are pure L2 traffic; CTA b = tile b without tickets when the grid covers all
        // tiles, and the next ticket requested before the ranking and consumed after it -- both a few per cent slower)
        constexpr int kLookWin = 8;
        bool done = false;
        for (int j = tile - 1; !done; j -= kLookWin) {
            uint32_t sw[kLookWin];
#pragma unroll
            for (int q = 0; q < kLookWin; q++)
                sw[q] = (j - q >= 0) ? ld_relaxed(status + (size_t)(j - q) * GSR_RADIX + t) : kFlagInc;   // before tile 0: inclusive 0
#pragma unroll
            for (int q = 0; q < kLookWin; q++) {
                if (done) break;
                uint32_t sv = sw[q];
                while ((sv >> 30) == 0u) {
                    __nanosleep(32);
                    sv = ld_relaxed(status + (size_t)(j - q) * GSR_RADIX + t);
                }
                excl += sv & kValMask;
                done = (sv >> 30) == 2u;
            }
        }
        st_relaxed(my_status, kFlagInc | (excl + run));
    }
    sm.gofs[t] = gbase + excl - ds;
    __syncthreads();
    // coalesced runs out
    const int count_out = DROP ? (int)sm.count_out : count;
#pragma unroll
    for (int i = 0; i < kSortItems; i++) {
        const int slot = i * kSortThreads + t;
        if ((FULL && !DROP) || slot < count_out) {
            const uint32_t k = sm.s_keys[slot];
            const uint32_t pos = sm.gofs[(k >> shift) & mask] + (uint32_t)slot;
            if (WRITE_KEYS) keys_out[pos] = k;
            vals_out[pos] = sm.s_vals[slot];
        }
    }
}

// MODE 0: plain pass. MODE 1 ("drop"): pairs whose key is 0xFFFFFFFF are left out of the output (the histograms do not
// count them either), so the pass also compacts; the number of kept pairs goes to *n_valid_out. MODE 2: a pass after a
// drop pass -- the pair count is not known to the host, it is the total of this pass's own histogram.
// n_dev != NULL: the pair count lives on the device (the instance sort of the speculative forward: the host sizes the
// launch for the capacity `n` of the buffers and learns the real count later). A count above the capacity means the
// buffers were too small: the pass does nothing (the caller repeats the binning with larger ones).
// WRITE_KEYS = false: the last pass of the instance sort leaves the keys unwritten (nobody reads sorted tile ids: the
// tile ranges come from the tile table).
template <int BITS, int MODE, bool WRITE_KEYS>
__global__ void __launch_bounds__(kSortThreads, kSortCtasPerSm) onesweep_kernel(const uint32_t* __restrict__ keys_in,
                                                                    const uint32_t* __restrict__ vals_in,
                                                                    uint32_t* __restrict__ keys_out,
                                                                    uint32_t* __restrict__ vals_out, int n, int n_tiles,
                                                                    int shift, const uint32_t* __restrict__ ghist,
                                                                    uint32_t* __restrict__ ticket,
                                                                    uint32_t* __restrict__ status,
                                                                    uint32_t* __restrict__ n_valid_out,
                                                                    const uint32_t* __restrict__ n_dev) {
    extern __shared__ __align__(16) unsigned char sort_smem_raw[];
    SortSmem& sm = *reinterpret_cast<SortSmem*>(sort_smem_raw);
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    if (MODE == 0 && n_dev != nullptr) {
        const uint32_t nd = __ldg(n_dev);
        if (nd > (uint32_t)n) return;                      // overflow of the speculative capacity: uniform over the grid
        n = (int)nd;
        n_tiles = (n + GSR_SORT_TILE - 1) / GSR_SORT_TILE;
    }
    if (t == 0) sm.tile = atomicAdd(ticket, 1u);
    const uint32_t hcount = __ldg(ghist + t);
    const uint32_t gbase = block_excl_scan256(hcount, sm.warp_sums);   // first output slot of digit t (syncs)
    if (MODE != 0) {
        if (t == kSortThreads - 1) sm.n_valid = gbase + hcount;
        __syncthreads();
        if (MODE == 2) {
            n = (int)sm.n_valid;
            n_tiles = (n + GSR_SORT_TILE - 1) / GSR_SORT_TILE;
        }
    }
    int tile = (int)sm.tile;
    if (MODE == 1 && tile == 0 && t == 0) *n_valid_out = sm.n_valid;
    if (tile >= n_tiles) return;
    sort_prefetch(sm, keys_in, vals_in, n, tile);
    const int wofs = w * (32 * kSortItems) + lane;
    while (true) {
        if (t == 0) sm.next = atomicAdd(ticket, 1u);
#pragma unroll
        for (int k = 0; k < kSortWarps; k++) sm.whist[k][t] = 0;
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();                                   // the tile's pairs, the next ticket and the cleared counters are visible
        const int next = (int)sm.next;
        uint32_t key[kSortItems], val[kSortItems];
#pragma unroll
        for (int i = 0; i < kSortItems; i++) { key[i] = sm.pf_keys[wofs + i * 32]; val[i] = sm.pf_vals[wofs + i * 32]; }
        __syncthreads();                                   // everybody holds its pairs: the prefetch buffers are free again
        if (next < n_tiles) sort_prefetch(sm, keys_in, vals_in, n, next);
        if ((tile + 1) * GSR_SORT_TILE <= n)
            onesweep_tile<BITS, true, MODE == 1, WRITE_KEYS>(sm, key, val, keys_out, vals_out, n, shift, tile, gbase, status);
        else
            onesweep_tile<BITS, false, MODE == 1, WRITE_KEYS>(sm, key, val, keys_out, vals_out, n, shift, tile, gbase, status);
        if (next >= n_tiles) break;
        tile = next;
        __syncthreads();                                   // reorder buffers / gofs are read until here
    }
}

// ------------------------------------------------------------------ small sorts: every pass in ONE CTA
// Up to kSmallMax pairs (config 1's 10k Gaussians, a SLAM keyframe's new points): the multi-launch sort costs a histogram
// kernel plus one latency-bound launch per pass (4 x 11 us + 14 us at 10k keys, a quarter of config 1's view). Here one
// CTA of 1024 threads holds all pairs in registers (12 per thread, warp-striped order) and runs every pass through shared
// memory: per-warp digit counters by the same ballot ranking as the tiles above, digit-major / warp-minor prefix, scatter,
// read back. Slots beyond n carry the key 0xFFFFFFFF, which sorts behind everything in every pass (ties keep input order:
// real keys first), so there is no validity mask anywhere; drop mode only counts the keys that are not 0xFFFFFFFF.
constexpr int kSmallThreads = 1024, kSmallItems = 12, kSmallMax = kSmallThreads * kSmallItems;   // 12288
constexpr int kSmallWarps = kSmallThreads / 32;
struct SmallSortSmem {
    uint32_t keys[kSmallMax];
    uint32_t vals[kSmallMax];
    uint32_t whist[kSmallWarps][GSR_RADIX];
    uint32_t dsum[GSR_RADIX];
    uint32_t warp_sums[8];
    uint32_t kept[kSmallWarps];
};
__global__ void __launch_bounds__(kSmallThreads, 1) small_sort_kernel(uint32_t* __restrict__ keys, uint32_t* __restrict__ vals, int n,
                                                                      int total_bits, uint32_t* __restrict__ n_valid_out) {
    extern __shared__ __align__(16) unsigned char small_smem_raw[];
    SmallSortSmem& sm = *reinterpret_cast<SmallSortSmem*>(small_smem_raw);
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const int wofs = w * (32 * kSmallItems) + lane;            // order index of item i: wofs + 32 * i
    uint32_t key[kSmallItems], val[kSmallItems];
    uint32_t kept = 0;
#pragma unroll
    for (int i = 0; i < kSmallItems; i++) {
        const int idx = wofs + 32 * i;
        key[i] = idx < n ? keys[idx] : 0xFFFFFFFFu;
        val[i] = idx < n ? vals[idx] : 0u;
        kept += key[i] != 0xFFFFFFFFu;
    }
    if (n_valid_out != nullptr) {                               // drop mode: how many pairs are part of the result
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) kept += __shfl_xor_sync(kAll, kept, o);
        if (lane == 0) sm.kept[w] = kept;
    }
    const int passes = (total_bits + 7) / 8;
    for (int p = 0; p < passes; p++) {
        const int shift = 8 * p;
        const uint32_t mask = (total_bits - shift) >= 8 ? 255u : ((1u << (total_bits - shift)) - 1u);
        for (int i = t; i < kSmallWarps * GSR_RADIX; i += kSmallThreads) (&sm.whist[0][0])[i] = 0u;
        __syncthreads();
        uint32_t rank[kSmallItems];
#pragma unroll
        for (int i = 0; i < kSmallItems; i++) {
            const uint32_t d = (key[i] >> shift) & mask;
            const uint32_t peers = same_digit_lanes<8>(d, kAll);
            const int leader = __ffs(peers) - 1;
            uint32_t prev = 0;
            if (lane == leader) prev = atomicAdd(&sm.whist[w][d], (uint32_t)__popc(peers));   // one warp: program order
            prev = __shfl_sync(kAll, prev, leader);
            rank[i] = prev + __popc(peers & lt_mask);
        }
        __syncthreads();
        // digit t (threads 0..255): total over the warps, exclusive scan over the digits, start of every (warp, digit) run
        uint32_t run = 0;
        if (t < GSR_RADIX) {
#pragma unroll 8
            for (int k = 0; k < kSmallWarps; k++) run += sm.whist[k][t];
        }
        if (t < GSR_RADIX) {                                    // warps 0..7 are whole: shuffles are safe
            uint32_t incl = run;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(kAll, incl, o);
                if (lane >= o) incl += y;
            }
            if (lane == 31) sm.warp_sums[w] = incl;
            sm.dsum[t] = incl - run;                            // exclusive inside the warp of digits
        }
        __syncthreads();
        if (t < GSR_RADIX) {
            uint32_t acc = sm.dsum[t];
            for (int k = 0; k < w; k++) acc += sm.warp_sums[k];
#pragma unroll 8
            for (int k = 0; k < kSmallWarps; k++) { const uint32_t c = sm.whist[k][t]; sm.whist[k][t] = acc; acc += c; }
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < kSmallItems; i++) {
            const uint32_t pos = sm.whist[w][(key[i] >> shift) & mask] + rank[i];
            sm.keys[pos] = key[i];
            sm.vals[pos] = val[i];
        }
        __syncthreads();
        if (p + 1 < passes) {
#pragma unroll
            for (int i = 0; i < kSmallItems; i++) { key[i] = sm.keys[wofs + 32 * i]; val[i] = sm.vals[wofs + 32 * i]; }
        }
    }
    for (int idx = t; idx < n; idx += kSmallThreads) { keys[idx] = sm.keys[idx]; vals[idx] = sm.vals[idx]; }
    if (n_valid_out != nullptr && t == 0) {
        uint32_t tot = 0;
        for (int k = 0; k < kSmallWarps; k++) tot += sm.kept[k];
        *n_valid_out = tot;
    }
}

template <int BITS, int MODE = 0, bool WRITE_KEYS = true>
static int launch_pass(const uint32_t* kin, const uint32_t* vin, uint32_t* kout, uint32_t* vout, int n, int shift,
                       const uint32_t* ghist, uint32_t* ticket, uint32_t* status, int tiles, cudaStream_t s,
                       uint32_t* n_valid_out = nullptr, const uint32_t* n_dev = nullptr) {
    static DeviceOnce once;                                // the opt-in is per device
    int e = ensure_dynamic_smem(onesweep_kernel<BITS, MODE, WRITE_KEYS>, (int)sizeof(SortSmem), once);
    if (e) return e;
    int sms = 148;
    e = sm_count(&sms);
    if (e) return e;
    const int ctas = tiles < sms * kSortCtasPerSm ? tiles : sms * kSortCtasPerSm;
    onesweep_kernel<BITS, MODE, WRITE_KEYS><<<ctas, kSortThreads, sizeof(SortSmem), s>>>(kin, vin, kout, vout, n, tiles, shift, ghist, ticket, status, n_valid_out, n_dev);
    return 0;
}

template <bool WRITE_KEYS>
static int launch_pass_bits(int bits, const uint32_t* kin, const uint32_t* vin, uint32_t* kout, uint32_t* vout, int n, int shift,
                            const uint32_t* gh, uint32_t* ticket, uint32_t* st, int tiles, cudaStream_t s, const uint32_t* n_dev) {
    switch (bits) {
        case 1: return launch_pass<1, 0, WRITE_KEYS>(kin, vin, kout, vout, n, shift, gh, ticket, st, tiles, s, nullptr, n_dev);
        case 2: return launch_pass<2, 0, WRITE_KEYS>(kin, vin, kout, vout, n, shift, gh, ticket, st, tiles, s, nullptr, n_dev);
        case 3: return launch_pass<3, 0, WRITE_KEYS>(kin, vin, kout, vout, n, shift, gh, ticket, st, tiles, s, nullptr, n_dev);
        case 4: return launch_pass<4, 0, WRITE_KEYS>(kin, vin, kout, vout, n, shift, gh, ticket, st, tiles, s, nullptr, n_dev);
        case 5: return launch_pass<5, 0, WRITE_KEYS>(kin, vin, kout, vout, n, shift, gh, ticket, st, tiles, s, nullptr, n_dev);
        case 6: return launch_pass<6, 0, WRITE_KEYS>(kin, vin, kout, vout, n, shift, gh, ticket, st, tiles, s, nullptr, n_dev);
        case 7: return launch_pass<7, 0, WRITE_KEYS>(kin, vin, kout, vout, n, shift, gh, ticket, st, tiles, s, nullptr, n_dev);
        default: return launch_pass<8, 0, WRITE_KEYS>(kin, vin, kout, vout, n, shift, gh, ticket, st, tiles, s, nullptr, n_dev);
    }
}

static int g_small_sort = [] { const char* e = getenv("GSR_SMALL_SORT"); return e ? atoi(e) : 1; }();    // A/B switch

int radix_sort_u32(uint32_t* ka, uint32_t* va, uint32_t* kb, uint32_t* vb, int n, int total_bits, uint32_t* scratch,
                   bool* result_in_a, bool debug, cudaStream_t s, uint32_t** n_valid_dev, bool scratch_zeroed,
                   const SortOptions* opt) {
    // n_valid_dev != NULL ("drop" sort, 32-bit keys only): pairs whose key is 0xFFFFFFFF are not part of the result;
    // the first pass compacts them away, the later passes run over the kept pairs only, and *n_valid_dev points at
    // their count (device). The tail of the output arrays beyond that count is unspecified.
    // opt (instance sort): hist_ready = the digit histograms in scratch[0 .. passes*256) were filled by the caller (the
    // tile table) and no histogram kernel runs; n_dev = device word with the real pair count (n is the capacity);
    // values_only_last = the last pass writes no keys.
    const bool drop = n_valid_dev != nullptr;
    if (drop && (total_bits != 32 || opt != nullptr)) return GSR_E_BADARG;
    *result_in_a = true;
    if (drop) *n_valid_dev = scratch + 4 * GSR_RADIX + 8;          // a spare word of the ticket block (zeroed below)
    if (n <= 0 || total_bits <= 0) return 0;
    if (n <= kSmallMax && opt == nullptr && g_small_sort) {        // every pass in one CTA, in place in (ka, va)
        static DeviceOnce once_small;
        { const int e = ensure_dynamic_smem(small_sort_kernel, (int)sizeof(SmallSortSmem), once_small); if (e) return e; }
        GSR_LAUNCH("small_sort", debug, s,
                   (small_sort_kernel<<<1, kSmallThreads, sizeof(SmallSortSmem), s>>>(ka, va, n, total_bits, drop ? *n_valid_dev : nullptr)));
        return 0;
    }
    const int passes = radix_passes(total_bits);
    const int tiles = (int)sort_tiles((size_t)n);
    uint32_t* ghist = scratch;                                  // [4][256]
    uint32_t* tickets = scratch + 4 * GSR_RADIX;                // [4] (+ padding)
    uint32_t* status = scratch + 4 * GSR_RADIX + 64;            // [passes][tiles][256]
    const bool hist_ready = opt && opt->hist_ready;
    const uint32_t* n_dev = opt ? opt->n_dev : nullptr;
    if (!scratch_zeroed) {
        // with the histograms already in place only the tickets and the look-back status are cleared
        uint32_t* from = hist_ready ? tickets : scratch;
        GSR_MEMSET("memset_sort", s, from, 0, sizeof(uint32_t) * ((size_t)(scratch + 4 * GSR_RADIX + 64 - from) + (size_t)passes * tiles * GSR_RADIX));
    }
    if (!hist_ready) {
        // a CTA pays a fixed price (clearing 32 KB of counters, flushing up to 1024 of them with global atomics), so it
        // should see at least ~16K keys; 4 CTAs per SM at most
        int sms = 148;
        { const int se = sm_count(&sms); if (se) return se; }
        // few CTAs: every CTA ends with up to 1024 global atomics on the same 1024 counters, and same-address atomics
    // serialise in L2 (~20 ns each) -- at 4 CTAs per SM that flush alone cost 12 us
    const int hist_blocks = min((n + 16383) / 16384, sms);
        GSR_LAUNCH("radix_hist", debug, s,
                   (radix_hist_kernel<<<hist_blocks > 0 ? hist_blocks : 1, kSortThreads, 0, s>>>(ka, n, passes, total_bits, ghist, drop ? 1 : 0)));
    }
    uint32_t *kin = ka, *vin = va, *kout = kb, *vout = vb;
    for (int p = 0; p < passes; p++) {
        const int shift = 8 * p;
        const int bits = (total_bits - shift) < 8 ? (total_bits - shift) : 8;
        uint32_t* st = status + (size_t)p * tiles * GSR_RADIX;
        const uint32_t* gh = ghist + p * GSR_RADIX;
        const bool no_keys = opt && opt->values_only_last && p == passes - 1;
        int le = 0;
        GSR_LAUNCH(drop ? "onesweep_depth" : (opt ? "onesweep_tile" : "onesweep_pass"), debug, s, {
            if (drop && p == 0) le = launch_pass<8, 1>(kin, vin, kout, vout, n, shift, gh, tickets + p, st, tiles, s, *n_valid_dev);
            else if (drop) le = launch_pass<8, 2>(kin, vin, kout, vout, n, shift, gh, tickets + p, st, tiles, s);
            else if (no_keys) le = launch_pass_bits<false>(bits, kin, vin, kout, vout, n, shift, gh, tickets + p, st, tiles, s, n_dev);
            else le = launch_pass_bits<true>(bits, kin, vin, kout, vout, n, shift, gh, tickets + p, st, tiles, s, n_dev);
        });
        if (le) return le;
        uint32_t* tk = kin; kin = kout; kout = tk;
        uint32_t* tv = vin; vin = vout; vout = tv;
    }
    *result_in_a = (passes % 2) == 0;
    return 0;
}

}  // namespace gsr
