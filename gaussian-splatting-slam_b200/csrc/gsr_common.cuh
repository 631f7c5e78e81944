// gsr_common.cuh -- shared declarations for the sm_100a rasterizer kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include "../../include/gsr.h"

#define GSR_TILE 16          // observable contract: tile ids appear in keys / ranges (config.h:16-17)
#define GSR_ALIGN 128

namespace gsr {

// ---------------------------------------------------------------- launch accounting / checks
extern unsigned long long g_launches;
extern bool g_prof_on;                               // gsr_profile_enable(): CUDA-event pair around every launch
int check_launch(const char* what, bool debug, cudaStream_t s);
void prof_begin(const char* what, cudaStream_t s);
void prof_end(cudaStream_t s);
// NVTX: with GSR_NVTX=1 in the environment (or gsr_nvtx_enable(1)) every launch sits inside a named range, so nsys / ncu
// group the kernels by stage as the reference's iter_start / iter_end events do per iteration (train.py:60-61,87,139).
extern bool g_nvtx_on;
void nvtx_push(const char* what);
void nvtx_pop();
#define GSR_LAUNCH(what, debug, stream, ...)                 \
    do {                                                     \
        if (gsr::g_nvtx_on) gsr::nvtx_push(what);            \
        if (gsr::g_prof_on) gsr::prof_begin(what, stream);   \
        __VA_ARGS__;                                         \
        if (gsr::g_prof_on) gsr::prof_end(stream);           \
        if (gsr::g_nvtx_on) gsr::nvtx_pop();                 \
        gsr::g_launches++;                                   \
        int _e = gsr::check_launch(what, debug, stream);     \
        if (_e) return _e;                                   \
    } while (0)
// memsets: timed like launches when profiling, not counted as kernels
#define GSR_MEMSET(what, stream, ptr, val, bytes)                                  \
    do {                                                                           \
        if (gsr::g_prof_on) gsr::prof_begin(what, stream);                         \
        cudaError_t _ce = cudaMemsetAsync(ptr, val, bytes, stream);                \
        if (gsr::g_prof_on) gsr::prof_end(stream);                                 \
        if (_ce != cudaSuccess) return (int)_ce;                                   \
    } while (0)

// ---------------------------------------------------------------- per-device launch state
// Function attributes (opt-in dynamic shared memory) and the SM count belong to a DEVICE, not to the process: a
// caller may render on cuda:0 and then on cuda:1 (the binding follows the tensors' device). DeviceOnce remembers, per
// kernel, on which devices the attribute has been set; sm_count() caches cudaDevAttrMultiProcessorCount per device.
constexpr int kMaxDevices = 64;
struct DeviceOnce { bool done[kMaxDevices] = {}; };
int current_device(int* dev);                        // cudaGetDevice with error code
int sm_count(int* sms);                              // SMs of the current device (cached per device)
template <typename F>
static inline int ensure_dynamic_smem(F* kernel, int bytes, DeviceOnce& once) {
    int dev = 0;
    int e = current_device(&dev);
    if (e) return e;
    if (dev >= 0 && dev < kMaxDevices && once.done[dev]) return 0;
    cudaError_t ce = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (ce != cudaSuccess) return (int)ce;
    if (dev >= 0 && dev < kMaxDevices) once.done[dev] = true;
    return 0;
}

// ---------------------------------------------------------------- Adam update (shared by gsr_adam.cu and gsr_comm.cu)
struct AdamHyper { float omb1, b2, omb2, eps; };      // 1 - beta1, beta2, 1 - beta2 (rounded from double, as torch does), eps
#ifdef __CUDACC__
__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamHyper h, float step_size,
                                         float inv_sqrt_bc2) {
    const float eps = h.eps;
    m = m + (g - m) * h.omb1;
    v = v * h.b2 + h.omb2 * g * g;
    const float denom = sqrtf(v) * inv_sqrt_bc2 + eps;
    p = p - step_size * (m / denom);
}
#endif

// ---------------------------------------------------------------- scratch layouts (private)
template <typename T>
static inline T* carve(char*& p, size_t count) {
    uintptr_t a = (reinterpret_cast<uintptr_t>(p) + GSR_ALIGN - 1) & ~uintptr_t(GSR_ALIGN - 1);
    T* r = reinterpret_cast<T*>(a);
    p = reinterpret_cast<char*>(r + count);
    return r;
}

#ifndef GSR_SORT_TILE
#define GSR_SORT_TILE 4096    // keys per radix-sort tile (256 threads x 16)
#endif
#define GSR_RADIX 256
#ifndef GSR_SCAN_TILE
#define GSR_SCAN_TILE 1024    // items per scan tile (256 threads x 4). Measured at C3 (scan pass 2 + emission, whose second search
                              // level walks two scan tiles): 4096 -> 0.086 ms, 2048 -> 0.071 ms, 1024 -> 0.063 ms
#endif

static inline size_t sort_tiles(size_t n) { return (n + GSR_SORT_TILE - 1) / GSR_SORT_TILE; }
// scratch words for radix_sort_u32 over n items: digit histograms [4][256], tickets, and one look-back status
// word per (pass, tile, digit)
static inline size_t sort_scratch_words(size_t n) { return 4 * GSR_RADIX + 64 + 4 * GSR_RADIX * sort_tiles(n) + 64; }

// packed tile rectangles carry 8 bits per field: images up to 4080 x 4080 pixels; beyond that binning gathers record and
// radius again and preprocess keeps tiles_touched
static inline bool rect_packing_ok(int W, int H) {
    return (W + GSR_TILE - 1) / GSR_TILE <= 255 && (H + GSR_TILE - 1) / GSR_TILE <= 255;
}

struct GeomState {            // replaces GeometryState (rasterizer_impl.h:31-46)
    float* depths;            // [P] view-space z (visible Gaussians only)
    float4* rec;              // [2P] cull part of the splat records: {x,y,ex,ey} {a,b,c,opacity} = 32 bytes = one sector per
                              // Gaussian (everything the sub-tile tests of the blend kernels need)
    float4* colr;             // [P] colour part: {r,g,b,1/depth} (= rec + 2P; fetched for surviving splats only)
    uint8_t* clamped;         // [P] bit c set = channel c clamped at 0
    uint32_t* tiles_touched;  // [P]
    uint32_t* dkey;           // [P] depth bits (0xFFFFFFFF for culled Gaussians): Gaussian-level sort key
    uint32_t* dkey_tmp;       // [P]
    uint32_t* order;          // [P] Gaussian ids in (depth, id) order
    uint32_t* order_tmp;      // [P]
    uint32_t* point_offsets;  // [P] inclusive scan of tiles_touched[order[j]]  (depth order!)
    uint32_t* sort_scratch;   // radix-sort tables for the P-sized sort
    unsigned long long* count64;  // [2] instance count of the view (only written when the forward has no capacity estimate)
    uint32_t* rect_packed;    // [P] tile rectangle x0 | y0 << 8 | width << 16 | height << 24 (visible Gaussians; grids up to 255 x 255 tiles)
    uint32_t* rect_sorted;    // [P] the same words in depth order (written by the scan, read by the instance emission)
    uint32_t* n_visible;      // [1] Gaussians that survived the cull = valid prefix of order / point_offsets (device)
    static GeomState from(char*& p, size_t P) {
        GeomState g;
        g.depths = carve<float>(p, P);
        g.rec = carve<float4>(p, 3 * P);
        g.colr = g.rec + 2 * P;
        g.clamped = carve<uint8_t>(p, P);
        g.tiles_touched = carve<uint32_t>(p, P);
        g.dkey = carve<uint32_t>(p, P);
        g.dkey_tmp = carve<uint32_t>(p, P);
        g.order = carve<uint32_t>(p, P);
        g.order_tmp = carve<uint32_t>(p, P);
        g.point_offsets = carve<uint32_t>(p, P);
        g.sort_scratch = carve<uint32_t>(p, sort_scratch_words(P));
        g.count64 = carve<unsigned long long>(p, 2);
        g.rect_packed = carve<uint32_t>(p, P);
        g.rect_sorted = carve<uint32_t>(p, P);
        g.n_visible = g.sort_scratch + 4 * GSR_RADIX + 8;      // where radix_sort_u32's drop mode leaves the kept count
        return g;
    }
};

struct ImageState {           // replaces ImageState (rasterizer_impl.h:48-55)
    float* final_T;           // [N]
    uint32_t* n_contrib;      // [N]
    uint2* ranges;            // [tiles]
    static ImageState from(char*& p, size_t N, size_t tiles) {
        ImageState s;
        s.final_T = carve<float>(p, N);
        s.n_contrib = carve<uint32_t>(p, N);
        s.ranges = carve<uint2>(p, tiles);
        return s;
    }
};

struct BinningState {         // replaces BinningState (rasterizer_impl.h:57-66)
    // R below is the CAPACITY the blob was sized for (>= num_rendered: the speculative forward sizes it before the count
    // is known). point_list comes first, so consumers of a finished forward (backward, statistics) find it without
    // knowing that capacity; everything behind it is private to the forward.
    uint32_t* point_list;     // [R] Gaussian ids sorted by (tile, depth, id) (final)
    uint32_t* tile_keys;      // [R] tile id per instance (ping-pong partner of tile_keys_tmp; not kept sorted)
    uint32_t* tile_keys_tmp;  // [R]
    uint32_t* vals_tmp;       // [R]
    uint32_t* sort_scratch;   // radix-sort tables for the R-sized sort: digit histograms [4][256], tickets, look-back status
    uint32_t* scan_sums;      // [2 * scan tiles + 16]: look-back words of the tile-count scan, total (num_rendered), ticket; from
                              // [scan tiles + 8] on the plain tile sums of the two-launch scan
    int32_t* tile_diff;       // [(gx + 1) * (gy + 1)] difference array of the tile rectangles (+1 / -1 at the corners); its
                              // 2-D prefix sum is the number of instances per tile (tile table: ranges, digit histograms)
    char* zero_begin;         // [zero_begin, zero_end): what one memset clears before every binning (tickets, status words,
    char* zero_end;           // scan words, difference array); the histograms in front of it are overwritten
    static size_t diff_cells(int W, int H) {
        return (size_t)((W + GSR_TILE - 1) / GSR_TILE + 1) * (size_t)((H + GSR_TILE - 1) / GSR_TILE + 1);
    }
    static size_t scan_tiles(size_t P) { return (P + GSR_SCAN_TILE - 1) / GSR_SCAN_TILE; }
    static BinningState from(char*& p, size_t R, size_t P, int W, int H) {
        BinningState b;
        size_t Rp = R + 8;
        b.point_list = carve<uint32_t>(p, Rp);
        b.tile_keys = carve<uint32_t>(p, Rp);
        b.tile_keys_tmp = carve<uint32_t>(p, Rp);
        b.vals_tmp = carve<uint32_t>(p, Rp);
        b.sort_scratch = carve<uint32_t>(p, sort_scratch_words(R));
        b.scan_sums = carve<uint32_t>(p, 2 * scan_tiles(P) + 16);   // status words + 8 (total, counter), then the tile sums
        b.tile_diff = carve<int32_t>(p, diff_cells(W, H));
        b.zero_begin = reinterpret_cast<char*>(b.sort_scratch + 4 * GSR_RADIX);
        b.zero_end = reinterpret_cast<char*>(b.tile_diff + diff_cells(W, H));
        return b;
    }
    static const uint32_t* point_list_of(const char* blob) {      // all a finished forward's consumers need
        char* p = const_cast<char*>(blob);
        return carve<uint32_t>(p, 0);
    }
};

static inline uint32_t higher_msb(uint32_t n) {   // rasterizer_impl.cu:35-50
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

// Packed per-Gaussian gradient accumulator written by the backward blend (12 floats = 3 x float4):
//   [0] dmean2D.x  [1] dmean2D.y  [2] dconic.xx  [3] dconic.xy
//   [4] dconic.yy  [5] dopacity   [6] dcolor.r   [7] dcolor.g
//   [8] dcolor.b   [9] dinvdepth  [10],[11] unused
#define GSR_ACC_FLOATS 12

// ---------------------------------------------------------------- stage entry points (host)
// zero_a / zero_b: optional scratch regions (multiples of 4 bytes) the kernel zero-fills on the side
// geometry_only: colours / inverse depths / clamp bits are left to launch_preprocess_color (any stream, after this one)
int launch_preprocess(const gsr_view& v, GeomState& g, int* radii, cudaStream_t s, void* zero_a = nullptr,
                      size_t zero_a_bytes = 0, void* zero_b = nullptr, size_t zero_b_bytes = 0, bool geometry_only = false);
int launch_preprocess_color(const gsr_view& v, GeomState& g, const int* radii, cudaStream_t s);
int launch_mark_visible(int P, const float* means, const float* view, uint8_t* present, cudaStream_t s);
// depth-order the visible Gaussians (drop-mode radix sort; the visible count stays on the device)
int launch_depth_order(const gsr_view& v, GeomState& g, cudaStream_t s, bool scratch_zeroed = false);
void depth_sort_zero_region(GeomState& g, void** begin, size_t* bytes);
// number of instances = sum of the tile counts (only used when the forward has no capacity estimate yet); *total_dev
// points at the device word
int launch_count_instances(const gsr_view& v, GeomState& g, cudaStream_t s);
// clears the binning blob's tables and builds the tile table (ranges + digit histograms from the rectangles' difference
// array); needs the preprocess only
int launch_tile_table(const gsr_view& v, GeomState& g, const int* radii, BinningState& b, ImageState& img, cudaStream_t s,
                      bool scratch_zeroed = false);
// after launch_tile_table on the same blob: scan of the tile counts in depth order, instance emission, stable sort by tile
// id. `cap` = instances the binning blob holds; the real count lands in *total_dev (device). Safe for any count: writes
// are clamped to cap, and a count above cap turns the later stages (sort passes, forward blend) into no-ops so that the
// caller can repeat the call with a larger blob.
// host_total: NULL, or mapped pinned host memory the scan's last tile stores the count into (the host polls it).
int launch_binning(const gsr_view& v, GeomState& g, const int* radii, BinningState& b, ImageState& img,
                   uint32_t cap, uint32_t** total_dev, uint32_t* host_total, cudaStream_t s);
int launch_blend_forward(const gsr_view& v, const GeomState& g, const uint32_t* point_list, ImageState& img,
                         float* out_color, float* out_invdepth, const uint32_t* total_dev, uint32_t cap, cudaStream_t s);
int launch_blend_backward(const gsr_view& v, const GeomState& g, const uint32_t* point_list, const ImageState& img,
                          const float* dL_dpix, const float* dL_dinvdepth, float* acc, cudaStream_t s);
int launch_preprocess_backward(const gsr_view& v, const GeomState& g, const int* radii, float* acc,
                               bool have_invdepth, const gsr_grads& out, cudaStream_t s);
// Stable LSD radix sort of (u32 key, u32 value) pairs on key bits [0, total_bits). Ping-pongs between
// (ka,va) and (kb,vb); *result_in_a tells where the sorted data ended up. scratch: sort_scratch_words(n).
struct SortOptions {
    bool hist_ready;          // the digit histograms are already in scratch[0 .. passes * 256): no histogram kernel
    const uint32_t* n_dev;    // device word holding the real pair count (then n is only the capacity), or NULL
    bool values_only_last;    // the last pass writes no keys
};
int radix_sort_u32(uint32_t* ka, uint32_t* va, uint32_t* kb, uint32_t* vb, int n, int total_bits,
                   uint32_t* scratch, bool* result_in_a, bool debug, cudaStream_t s, uint32_t** n_valid_dev = nullptr,
                   bool scratch_zeroed = false, const SortOptions* opt = nullptr);
static inline int radix_passes(int total_bits) { return (total_bits + 7) / 8; }
int launch_sorted_keys_debug(const gsr_view& v, const GeomState& g, const uint32_t* point_list, const ImageState& img, int R,
                             uint64_t* keys_out, cudaStream_t s);
int launch_blend_stats(const gsr_view& v, const GeomState& g, const uint32_t* point_list, const ImageState& img,
                       unsigned long long* out, cudaStream_t s);
int launch_ssim_forward(int planes, int H, int W, float C1, float C2, const float* img1, const float* img2, float* ssim_map,
                        float* dm_dmu1, float* dm_dsigma1_sq, float* dm_dsigma12, double* sums, cudaStream_t s);
int launch_ssim_backward(int planes, int H, int W, const float* img1, const float* img2, const float* dL_dmap,
                         float map_scale, float l1_scale, const float* upstream, const float* dm_dmu1,
                         const float* dm_dsigma1_sq, const float* dm_dsigma12, float* dL_dimg1, cudaStream_t s);
size_t densify_plan_bytes(size_t P);
int launch_densify_plan(int P, const float* opacity, const float* scaling, const float* grad_accum, const float* denom,
                        const gsr_densify_rule& rule, void* blob, uint32_t* counts_out, cudaStream_t s);
int launch_prune_plan(int P, const uint8_t* prune_mask, void* blob, uint32_t* counts_out, cudaStream_t s);
int launch_densify_apply(int P, const void* blob, const gsr_densify_rule* rule, const gsr_densify_tensor* tensors, int n,
                         const float* scaling, const float* rotation, const float* noise, cudaStream_t s);
int launch_densify_stats(int P, const float* dL_dmean2D, const uint8_t* update_filter, const int* radii, float* grad_accum,
                         float* denom, float* max_radii2D, cudaStream_t s);
int set_exact_cull(int on);
int set_blend_variant(int v);
int set_comm_variant(int v);
int launch_cull_stats(const gsr_view& v, const GeomState& g, const uint32_t* point_list, const ImageState& img,
                      unsigned long long* out, cudaStream_t s);
int launch_adam(const gsr_adam_group* groups, int n, double b1, double b2, double eps, const uint8_t* visible, cudaStream_t s);
int launch_exchange_adam(const gsr_peer_group& grads, const gsr_peer_group& params, float* exp_avg, float* exp_avg_sq,
                         const gsr_adam_span* spans, int n_spans, double b1, double b2, double eps, long long numel, int max_ctas,
                         cudaStream_t s);
int launch_allreduce_sum(const gsr_peer_group& g, long long numel, int max_ctas, cudaStream_t s);
void comm_slice(long long n16, int rank, int world, long long* lo, long long* hi);
size_t comm_emulation_scratch_bytes();
int launch_allreduce_sum_emulated(const gsr_peer_group* groups, int world, long long numel, int max_ctas, void* scratch_dev,
                                  cudaStream_t s);
int launch_exchange_adam_emulated(const gsr_peer_group* grads, const gsr_peer_group* params, float* const* exp_avg,
                                  float* const* exp_avg_sq, int world, const gsr_adam_span* spans, int n_spans, double b1, double b2,
                                  double eps, long long numel, int max_ctas, void* scratch_dev, cudaStream_t s);
int launch_exp_check(unsigned long long* mismatches_dev, cudaStream_t s);
int launch_fp32_peak(int iters, float* out, int* sms_out, cudaStream_t s);
int launch_dist2(int P, const float* pts, float* out, char* temp, cudaStream_t s);
size_t dist2_temp_bytes(int P);
int launch_dist2_noalloc(int P, const float* pts, float* out, cudaStream_t s);   // scratch in a static device array

}  // namespace gsr
