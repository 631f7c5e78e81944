// gsr_api.cu -- the extern "C" boundary declared in include/gsr.h and the stage orchestration that
// replaces CudaRasterizer::Rasterizer::forward / backward (rasterizer_impl.cu:198-336, 340-434).
#include "gsr_common.cuh"
#include <nvtx3/nvToolsExt.h>
#include <cstdio>
#include <cstring>
#include <algorithm>
#include <cstdlib>
#include <map>
#include <mutex>
#include <string>
#include <vector>

namespace gsr {

unsigned long long g_launches = 0;

// ---- optional per-launch timing (bench.py's roofline leg): one CUDA-event pair per launch on the launching
// stream, resolved at report time. Off by default; costs nothing when off.
bool g_prof_on = false;
namespace {
struct ProfRec { const char* name; cudaEvent_t a, b; };
std::vector<ProfRec> g_prof_recs;
std::vector<cudaEvent_t> g_prof_pool;
cudaEvent_t prof_event() {
    cudaEvent_t e;
    if (!g_prof_pool.empty()) { e = g_prof_pool.back(); g_prof_pool.pop_back(); return e; }
    cudaEventCreate(&e);
    return e;
}
}  // namespace
static bool nvtx_env() {
    const char* e = getenv("GSR_NVTX");
    return e && e[0] == '1';
}
bool g_nvtx_on = nvtx_env();
void nvtx_push(const char* what) { nvtxRangePushA(what); }
void nvtx_pop() { nvtxRangePop(); }
void prof_begin(const char* what, cudaStream_t s) {
    ProfRec r{what, prof_event(), prof_event()};
    cudaEventRecord(r.a, s);
    g_prof_recs.push_back(r);
}
void prof_end(cudaStream_t s) { cudaEventRecord(g_prof_recs.back().b, s); }

int current_device(int* dev) { return (int)cudaGetDevice(dev); }
int sm_count(int* sms) {
    static int cache[kMaxDevices] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev >= 0 && dev < kMaxDevices && cache[dev] > 0) { *sms = cache[dev]; return 0; }
    int n = 0;
    e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return (int)e;
    if (dev >= 0 && dev < kMaxDevices) cache[dev] = n;
    *sms = n;
    return 0;
}

int check_launch(const char* what, bool debug, cudaStream_t s) {
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess && debug) {            // auxiliary.h:166-173 (CHECK_CUDA in debug mode)
        e = cudaStreamSynchronize(s);
        if (e == cudaSuccess) e = cudaGetLastError();
    }
    if (e != cudaSuccess) {
        fprintf(stderr, "\n[GSR CUDA ERROR] %s: %s\n", what, cudaGetErrorString(e));
        return (int)e;
    }
    return 0;
}

static size_t blob_size_geom(size_t P) {
    char* p = nullptr;
    GeomState::from(p, P);
    return (size_t)p + GSR_ALIGN;
}
static size_t blob_size_binning(size_t R, size_t P, int W, int H) {
    char* p = nullptr;
    BinningState::from(p, R, P, W, H);
    return (size_t)p + GSR_ALIGN;
}
static size_t blob_size_image(size_t N, size_t tiles) {
    char* p = nullptr;
    ImageState::from(p, N, tiles);
    return (size_t)p + GSR_ALIGN;
}
static inline int tiles_of(int W, int H) { return ((W + GSR_TILE - 1) / GSR_TILE) * ((H + GSR_TILE - 1) / GSR_TILE); }

static bool view_ok(const gsr_view* v) {
    if (!v || v->P < 0 || v->W <= 0 || v->H <= 0) return false;
    if (v->P == 0) return true;
    if (!v->means3D || !v->opacities || !v->viewmatrix || !v->projmatrix || !v->background) return false;
    const bool has_sh = v->shs || v->sh_dc;
    if (!has_sh && !v->colors_precomp) return false;
    if (has_sh && (!v->cam_pos || v->M < (v->D + 1) * (v->D + 1) || v->D < 0 || v->D > 3)) return false;
    if (v->sh_dc && v->M > 1 && !v->shs) return false;      /* dc given: the M-1 higher coefficients must come too */
    if (!v->cov3D_precomp && (!v->scales || !v->rotations)) return false;
    return true;
}

}  // namespace gsr

using namespace gsr;

extern "C" {

size_t gsr_required_geom(int P) { return blob_size_geom((size_t)(P > 0 ? P : 0)); }
size_t gsr_required_binning(int R, int P, int W, int H) {
    return blob_size_binning((size_t)(R > 0 ? R : 0), (size_t)(P > 0 ? P : 0), W > 0 ? W : 1, H > 0 ? H : 1);
}
size_t gsr_required_image(int W, int H) { return blob_size_image((size_t)W * H, (size_t)tiles_of(W, H)); }
int gsr_key_bits(int W, int H) { return 32 + (int)higher_msb((uint32_t)tiles_of(W, H)); }
size_t gsr_backward_scratch_bytes(int P) { return sizeof(float) * GSR_ACC_FLOATS * (size_t)(P > 0 ? P : 0) + 256; }
size_t gsr_dist2_temp_bytes(int P) { return dist2_temp_bytes(P); }
uint64_t gsr_launch_count(void) { return g_launches; }

int gsr_nvtx_enable(int on) { g_nvtx_on = on != 0; return 0; }

int gsr_profile_enable(int on) {
    if (on) {   // a new recording starts: recycle the previous log's events
        for (auto& r : g_prof_recs) { g_prof_pool.push_back(r.a); g_prof_pool.push_back(r.b); }
        g_prof_recs.clear();
    }
    g_prof_on = on != 0;
    return 0;
}

long gsr_profile_report(char* buf, size_t cap) {
    struct Agg { long n = 0; double ms = 0.0; int first = 0; };
    std::map<std::string, Agg> agg;
    int order = 0;
    for (auto& r : g_prof_recs) {
        cudaError_t e = cudaEventSynchronize(r.b);
        if (e != cudaSuccess) return -(long)e;
        float ms = 0.f;
        e = cudaEventElapsedTime(&ms, r.a, r.b);
        if (e != cudaSuccess) return -(long)e;
        Agg& a = agg[r.name];
        if (a.n == 0) a.first = order++;
        a.n++;
        a.ms += ms;
    }
    std::vector<std::pair<int, std::string>> names;
    for (auto& kv : agg) names.push_back({kv.second.first, kv.first});
    std::sort(names.begin(), names.end());
    std::string out;
    char line[256];
    for (auto& nm : names) {
        const Agg& a = agg[nm.second];
        snprintf(line, sizeof(line), "%s %ld %.6f\n", nm.second.c_str(), a.n, a.ms);
        out += line;
    }
    if (buf && cap > 0) {
        const size_t n = out.size() < cap - 1 ? out.size() : cap - 1;
        memcpy(buf, out.data(), n);
        buf[n] = 0;
    }
    return (long)out.size() + 1;
}
const char* gsr_version(void) { return "gsr 0.1 (sm_100a)"; }

const char* gsr_error_string(int code) {
    switch (code) {
        case 0: return "ok";
        case GSR_E_BADARG: return "gsr: bad argument";
        case GSR_E_TOOLARGE: return "gsr: too large (num_rendered >= 2^30 - 1, or more than 2^20 tiles)";
        case GSR_E_ALLOC: return "gsr: scratch allocation callback returned NULL";
        case GSR_E_STATIC_LIMIT:
            return "gsr: input larger than the fixed scratch of a no-allocation entry point (gsr_dist2_noalloc: at most "
                   "131072 points; use gsr_dist2 with a caller-supplied buffer)";
        default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "gsr: unknown error";
    }
}

int gsr_geom_layout(int P, size_t* off, int max_off) {
    char* p = nullptr;
    GeomState g = GeomState::from(p, (size_t)P);
    const size_t v[] = {(size_t)g.depths, (size_t)g.rec, (size_t)g.clamped, (size_t)g.tiles_touched,
                        (size_t)g.point_offsets, (size_t)g.order, (size_t)g.dkey, (size_t)g.rect_packed};
    int n = 0;
    for (; n < (int)(sizeof(v) / sizeof(v[0])) && n < max_off; n++) off[n] = v[n];
    return n;
}
int gsr_binning_layout(int R, int W, int H, size_t* off, int max_off) {
    (void)R; (void)W; (void)H;                  // point_list sits at the blob's base whatever capacity it was sized for
    if (max_off < 1) return 0;
    off[0] = (size_t)BinningState::point_list_of(nullptr);
    return 1;
}
int gsr_image_layout(int W, int H, size_t* off, int max_off) {
    char* p = nullptr;
    ImageState s = ImageState::from(p, (size_t)W * H, (size_t)tiles_of(W, H));
    const size_t v[] = {(size_t)s.final_T, (size_t)s.n_contrib, (size_t)s.ranges};
    int n = 0;
    for (; n < 3 && n < max_off; n++) off[n] = v[n];
    return n;
}

// ---- speculative capacity of the binning blob ------------------------------------------------------------------------
// The reference stops the GPU in the middle of every forward: it copies num_rendered to the host, waits, sizes the
// binning buffers and only then launches the remaining kernels (rasterizer_impl.cu:281-284). Here the host remembers,
// per (device, image size), the largest instance count seen so far (relative to the number of Gaussians), sizes the blob
// from that estimate, launches EVERYTHING, and waits for the count -- copied to pinned memory right after the emission
// kernel -- only once the whole forward is queued: the GPU never idles. Every kernel reads the count from the device and
// clamps its writes to the capacity; if the estimate was too small the later stages turn into no-ops and binning + blend
// are repeated with a blob of the right size (then the estimate is raised). The first forward of a shape has no estimate
// and takes the reference's route (count, wait, allocate). gsr_set_speculation(0) forces that route always.
namespace {
struct CapHint { int dev, W, H; double per_gaussian; double high; };
std::mutex g_hint_mu;
std::vector<CapHint> g_hints;
int g_speculate = -1;                        // -1: read GSR_SPECULATE on first use
int g_overlap = 1;                           // tile table on a side stream under the depth sort (gsr_set_speculation(2) = off)
int g_split = -1;                            // colour half of the preprocess off the critical path: OFF unless GSR_SPLIT_PRE=1
unsigned long long g_spec_hits = 0, g_spec_retries = 0, g_spec_cold = 0;

bool speculation_on() {
    if (g_speculate < 0) {
        const char* e = getenv("GSR_SPECULATE");
        g_speculate = (e && e[0] == '0') ? 0 : 1;
    }
    return g_speculate != 0;
}
bool split_on() {
    if (g_split < 0) {
        const char* e = getenv("GSR_SPLIT_PRE");
        g_split = (e && e[0] == '1') ? 1 : 0;
    }
    return g_split != 0;
}
uint32_t hint_capacity(int dev, int W, int H, int P) {
    std::lock_guard<std::mutex> lk(g_hint_mu);
    for (auto& h : g_hints)
        if (h.dev == dev && h.W == W && h.H == H) {
            // the larger of: the largest count seen, and the largest count per Gaussian seen scaled to this P (a growing
            // map raises the count with it); 1/8 headroom, rounded up to whole sort tiles
            double c = std::max(h.high, h.per_gaussian * (double)P) * 1.125 + 8192.0;
            if (c > 1.0e9) c = 1.0e9;
            return (uint32_t)(((size_t)c + 4095) / 4096 * 4096);
        }
    return 0;
}
void hint_update(int dev, int W, int H, int P, uint32_t R) {
    std::lock_guard<std::mutex> lk(g_hint_mu);
    const double per = P > 0 ? (double)R / (double)P : 0.0;
    for (auto& h : g_hints)
        if (h.dev == dev && h.W == W && h.H == H) {
            // slow decay, so that one outlier view does not pin the capacity for ever
            h.high = std::max((double)R, h.high * 0.999);
            h.per_gaussian = std::max(per, h.per_gaussian * 0.999);
            return;
        }
    if (g_hints.size() >= 64) g_hints.erase(g_hints.begin());
    g_hints.push_back(CapHint{dev, W, H, per, (double)R});
}

// pinned word + event for the count's way home (one per host thread: gsr_forward waits before it returns)
struct CountSlot {
    uint32_t* host = nullptr;                // pinned + mapped: the scan kernel stores num_rendered here
    uint32_t* host_dev = nullptr;            // ... through this device pointer
    unsigned long long* host64 = nullptr;
    cudaEvent_t ev[kMaxDevices] = {};
    bool have[kMaxDevices] = {};
    cudaStream_t side_stream[kMaxDevices] = {};
    cudaStream_t chain_stream[kMaxDevices] = {};   // highest priority: the latency-bound binning chain of the split forward
    cudaEvent_t fork_ev[kMaxDevices] = {}, join_ev[kMaxDevices] = {}, chain_ev[kMaxDevices] = {};
    int side_state[kMaxDevices] = {};      // 0 = not created, 1 = ready, -1 = creation failed (stay in-stream)
    int side(int dev, cudaStream_t* st, cudaEvent_t* fork, cudaEvent_t* join, cudaStream_t* chain = nullptr,
             cudaEvent_t* chain_done = nullptr) {
        if (dev < 0 || dev >= kMaxDevices) return GSR_E_BADARG;
        if (side_state[dev] == 0) {
            int least = 0, greatest = 0;
            cudaError_t ce = cudaDeviceGetStreamPriorityRange(&least, &greatest);
            if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&side_stream[dev], cudaStreamNonBlocking);
            if (ce == cudaSuccess) ce = cudaStreamCreateWithPriority(&chain_stream[dev], cudaStreamNonBlocking, greatest);
            if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&fork_ev[dev], cudaEventDisableTiming);
            if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&join_ev[dev], cudaEventDisableTiming);
            if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&chain_ev[dev], cudaEventDisableTiming);
            side_state[dev] = ce == cudaSuccess ? 1 : -1;
            if (ce != cudaSuccess) (void)cudaGetLastError();
        }
        if (side_state[dev] != 1) return GSR_E_BADARG;
        *st = side_stream[dev]; *fork = fork_ev[dev]; *join = join_ev[dev];
        if (chain) *chain = chain_stream[dev];
        if (chain_done) *chain_done = chain_ev[dev];
        return 0;
    }
    int get(int dev, cudaEvent_t* e) {
        if (!host) {
            void* p = nullptr;
            cudaError_t ce = cudaHostAlloc(&p, 64, cudaHostAllocPortable | cudaHostAllocMapped);
            if (ce != cudaSuccess) return (int)ce;
            void* dp = nullptr;
            ce = cudaHostGetDevicePointer(&dp, p, 0);
            if (ce != cudaSuccess) return (int)ce;
            host = reinterpret_cast<uint32_t*>(p);
            host_dev = reinterpret_cast<uint32_t*>(dp);
            host64 = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(p) + 16);
        }
        if (dev < 0 || dev >= kMaxDevices) return GSR_E_BADARG;
        if (!have[dev]) {                      // events belong to a device
            cudaError_t ce = cudaEventCreateWithFlags(&ev[dev], cudaEventDisableTiming);
            if (ce != cudaSuccess) return (int)ce;
            have[dev] = true;
        }
        *e = ev[dev];
        return 0;
    }
};
thread_local CountSlot t_slot;
}  // namespace

int gsr_debug_set_split_forward(int on) { g_split = on ? 1 : 0; return 0; }
int gsr_set_speculation(int on) { g_speculate = on ? 1 : 0; g_overlap = on == 2 ? 0 : 1; return 0; }
int gsr_reset_capacity_hints(void) {
    std::lock_guard<std::mutex> lk(g_hint_mu);
    g_hints.clear();
    return 0;
}
int gsr_speculation_stats(unsigned long long* out3) {
    if (!out3) return GSR_E_BADARG;
    out3[0] = g_spec_hits; out3[1] = g_spec_retries; out3[2] = g_spec_cold;
    return 0;
}

namespace {
struct NvtxScope {
    bool on;
    explicit NvtxScope(const char* name) : on(g_nvtx_on) { if (on) nvtx_push(name); }
    ~NvtxScope() { if (on) nvtx_pop(); }
};
}  // namespace

int gsr_forward(const gsr_view* view, gsr_alloc_fn geom_alloc, void* geom_ctx, gsr_alloc_fn binning_alloc,
                void* binning_ctx, gsr_alloc_fn image_alloc, void* image_ctx, float* out_color, int* radii,
                float* out_invdepth, int* num_rendered, void* stream) {
    NvtxScope nvtx_scope("gsr_forward");
    if (!view_ok(view) || !geom_alloc || !binning_alloc || !image_alloc || !out_color || !num_rendered) return GSR_E_BADARG;
    const gsr_view& v = *view;
    cudaStream_t s = (cudaStream_t)stream;
    *num_rendered = 0;
    const size_t N = (size_t)v.W * v.H;
    const int tiles = tiles_of(v.W, v.H);
    if (v.P == 0) {   // rasterize_points.cu:81 : outputs stay zero-filled
        cudaError_t e = cudaMemsetAsync(out_color, 0, sizeof(float) * 3 * N, s);
        if (e == cudaSuccess && out_invdepth) e = cudaMemsetAsync(out_invdepth, 0, sizeof(float) * N, s);
        return (int)e;
    }
    if (!radii) return GSR_E_BADARG;
    if ((long long)tiles > (1ll << 20)) return GSR_E_TOOLARGE;      // tile ids are sorted in at most 20 bits (16K x 16K pixels)
    char* gp = geom_alloc(geom_ctx, blob_size_geom((size_t)v.P));
    char* ip = image_alloc(image_ctx, blob_size_image(N, (size_t)tiles));
    if (!gp || !ip) return GSR_E_ALLOC;
    GeomState g = GeomState::from(gp, (size_t)v.P);
    ImageState img = ImageState::from(ip, N, (size_t)tiles);

    int dev = 0;
    int e = current_device(&dev);
    if (e) return e;
    cudaEvent_t ev;
    e = t_slot.get(dev, &ev);
    if (e) return e;
    constexpr uint32_t kLimit = (1u << 30) - 1u;                      // saturation value of the scan
    uint32_t cap = speculation_on() ? hint_capacity(dev, v.W, v.H, v.P) : 0u;
    if (cap >= kLimit) cap = kLimit - 1u;
    uint32_t R = 0;
    const bool cold = cap == 0u;
    // With a capacity estimate the binning blob exists before the first kernel: the preprocess clears the tables of the
    // depth sort and of the binning on the side (no memset launches), and the tile table -- which needs the preprocess
    // only -- runs on a side stream UNDER the depth sort (four latency-bound passes that leave most of the GPU idle).
    char* bp = nullptr;
    bool table_done = false, binning_zeroed = false;
    cudaEvent_t ev_join = nullptr;
    // Split forward -- MEASURED AND NOT ADOPTED (GSR_SPLIT_PRE=1 enables it; C3: 1.878 against 1.863 ms/view, SLAM loop 0.771
    // against 0.741 ms/keyframe): the colour kernel saturates HBM while it runs, the latency-bound chain kernels beside it
    // slow down by what the overlap was meant to save (histogram 21 -> 59 us, depth passes 95 -> 128 us, tile table 41 ->
    // 178 us) and the colour pass itself takes 135 us without the L2 prefetch distance the one-kernel preprocess gives it.
    // The design: (estimate known, side streams available) the geometry half of the preprocess runs on the caller's
    // stream; the binning chain (histogram, depth passes, scan, emission, instance sort -- latency- and issue-bound, little
    // HBM traffic) moves to a library-owned stream of the HIGHEST priority, the tile table to the side stream, and the
    // colour half of the preprocess (60 % of its bytes) stays on the caller's stream, where it runs UNDER the chain: the
    // chain's CTAs take SM slots first whenever some free up. The caller's stream joins the chain before the blend.
    cudaStream_t cs = s;                                              // stream of the binning chain
    cudaEvent_t ev_chain = nullptr;
    bool chain_forked = false;
    void* za = nullptr; size_t zan = 0;
    depth_sort_zero_region(g, &za, &zan);
    if (!cold) {
        bp = binning_alloc(binning_ctx, blob_size_binning((size_t)cap, (size_t)v.P, v.W, v.H));
        if (!bp) return GSR_E_ALLOC;
        char* bq = bp;                                                // from() advances the pointer it is given
        BinningState b0 = BinningState::from(bq, (size_t)cap, (size_t)v.P, v.W, v.H);
        cudaStream_t side = nullptr, chain = nullptr;
        cudaEvent_t ev_fork = nullptr;
        const bool have_side = g_overlap && !v.debug && t_slot.side(dev, &side, &ev_fork, &ev_join, &chain, &ev_chain) == 0;
        const bool split = have_side && split_on();
        e = launch_preprocess(v, g, radii, s, za, zan, b0.zero_begin, (size_t)(b0.zero_end - b0.zero_begin), split);
        if (e) return e;
        binning_zeroed = true;
        if (have_side) {
            cudaError_t ce = cudaEventRecord(ev_fork, s);
            if (ce == cudaSuccess) ce = cudaStreamWaitEvent(side, ev_fork, 0);
            if (ce == cudaSuccess && split) ce = cudaStreamWaitEvent(chain, ev_fork, 0);
            if (ce != cudaSuccess) return (int)ce;
            if (split) { cs = chain; chain_forked = true; }
            e = launch_tile_table(v, g, radii, b0, img, side, /*scratch_zeroed=*/true);
            if (e) return e;
            ce = cudaEventRecord(ev_join, side);
            if (ce != cudaSuccess) return (int)ce;
            table_done = true;
        }
        if (split) {
            e = launch_depth_order(v, g, cs, /*scratch_zeroed=*/true);     // chain stream first: its kernels are queued ahead
            if (e) return e;
            e = launch_preprocess_color(v, g, radii, s);
            if (e) return e;
        }
    } else {
        e = launch_preprocess(v, g, radii, s, za, zan);
        if (e) return e;
    }
    if (!chain_forked) {
        e = launch_depth_order(v, g, s, /*scratch_zeroed=*/true);
        if (e) return e;
    }
    if (table_done) {
        cudaError_t ce = cudaStreamWaitEvent(cs, ev_join, 0);
        if (ce != cudaSuccess) return (int)ce;
    }
    if (cold) {
        // no estimate for this image size yet: count first (rasterizer_impl.cu:281, the reference's one host synchronisation)
        g_spec_cold++;
        e = launch_count_instances(v, g, s);
        if (e) return e;
        cudaError_t ce = cudaMemcpyAsync(t_slot.host64, g.count64, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s);
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);
        if (ce != cudaSuccess) return (int)ce;
        if (*t_slot.host64 >= (unsigned long long)kLimit) return GSR_E_TOOLARGE;
        R = (uint32_t)*t_slot.host64;
        cap = R;
    }
    for (int attempt = 0; attempt < 2; attempt++) {
        if (!bp) {
            bp = binning_alloc(binning_ctx, blob_size_binning((size_t)cap, (size_t)v.P, v.W, v.H));
            if (!bp) return GSR_E_ALLOC;
        }
        char* bq = bp;
        BinningState b = BinningState::from(bq, (size_t)cap, (size_t)v.P, v.W, v.H);
        bp = nullptr;                                                 // a repeat allocates a new blob
        uint32_t* total_dev = nullptr;
        const cudaStream_t bs = attempt == 0 ? cs : s;                // a repeat runs on the caller's stream (already joined)
        if (!table_done) {
            e = launch_tile_table(v, g, radii, b, img, bs, binning_zeroed);
            if (e) return e;
        }
        table_done = false;
        binning_zeroed = false;                                       // a repeat works on a fresh blob
        const bool check_count = !cold || attempt > 0 || v.debug;
        constexpr uint32_t kPending = 0xFFFFFFFFu;
        if (check_count) *(volatile uint32_t*)t_slot.host = kPending;
        e = launch_binning(v, g, radii, b, img, cap, &total_dev, check_count && cap > 0 ? t_slot.host_dev : nullptr, bs);
        if (e) return e;
        if (attempt == 0 && chain_forked) {                           // the caller's stream (colours) meets the chain (lists)
            cudaError_t ce = cudaEventRecord(ev_chain, cs);
            if (ce == cudaSuccess) ce = cudaStreamWaitEvent(s, ev_chain, 0);
            if (ce != cudaSuccess) return (int)ce;
        }
        e = launch_blend_forward(v, g, b.point_list, img, out_color, out_invdepth, cold && attempt == 0 ? nullptr : total_dev, cap, s);
        if (e) return e;
        if (!check_count) break;                                      // exact capacity: nothing to check
        // The count leaves the device from inside the scan kernel (a store into mapped pinned memory): the host waits for
        // that kernel only, while the emission, the sort and the blend queued behind it keep the GPU busy.
        uint32_t got = cap > 0 ? *(volatile uint32_t*)t_slot.host : 0u;
        for (unsigned spins = 0; got == kPending; spins++) {
            if ((spins & 1023u) == 1023u) {
                const cudaError_t q = cudaStreamQuery(bs);
                if (q != cudaErrorNotReady) {                         // the stream has drained (or failed): one last look
                    got = *(volatile uint32_t*)t_slot.host;
                    if (got == kPending) return q != cudaSuccess ? (int)q : GSR_E_BADARG;
                    break;
                }
            }
            got = *(volatile uint32_t*)t_slot.host;
        }
        if (cold && attempt == 0) {                                   // debug mode: the two counts must agree
            if (got != R) return GSR_E_BADARG;
            break;
        }
        if (got >= kLimit) return GSR_E_TOOLARGE;
        R = got;
        if (R <= cap) { if (attempt == 0) g_spec_hits++; break; }
        if (attempt == 1) return GSR_E_BADARG;                        // cannot happen: the second blob has exactly R slots
        g_spec_retries++;
        cap = R;                                                      // the estimate was too small: once more, exact size
    }
    hint_update(dev, v.W, v.H, v.P, R);
    *num_rendered = (int)R;
    return 0;
}

int gsr_backward(const gsr_view* view, int R, const int* radii, const char* geom, const char* binning, const char* image,
                 const float* dL_dpix, const float* dL_dinvdepth, char* accum_scratch, const gsr_grads* out, void* stream) {
    NvtxScope nvtx_scope("gsr_backward");
    if (!view_ok(view) || !out) return GSR_E_BADARG;
    const gsr_view& v = *view;
    if (v.P == 0) return 0;
    if (!radii || !geom || !image || !dL_dpix || !accum_scratch || !out->dL_dmeans2D || !out->dL_dopacity ||
        !out->dL_dmeans3D || (R > 0 && !binning))
        return GSR_E_BADARG;
    if (v.shs && !out->dL_dsh) return GSR_E_BADARG;
    if (v.sh_dc && !out->dL_ddc) return GSR_E_BADARG;
    cudaStream_t s = (cudaStream_t)stream;
    char* gp = const_cast<char*>(geom);
    char* ip = const_cast<char*>(image);
    GeomState g = GeomState::from(gp, (size_t)v.P);
    ImageState img = ImageState::from(ip, (size_t)v.W * v.H, (size_t)tiles_of(v.W, v.H));
    char* ap = accum_scratch;
    float* acc = carve<float>(ap, (size_t)v.P * GSR_ACC_FLOATS);
    if (!out->accum_clean) GSR_MEMSET("memset_grad_acc", s, acc, 0, sizeof(float) * GSR_ACC_FLOATS * (size_t)v.P);
    int e = 0;
    if (R > 0) {
        e = launch_blend_backward(v, g, BinningState::point_list_of(binning), img, dL_dpix, dL_dinvdepth, acc, s);
        if (e) return e;
    }
    gsr_grads o = *out;
    if (!v.shs) o.dL_dsh = nullptr;
    if (!v.sh_dc) o.dL_ddc = nullptr;
    if (!v.scales) { o.dL_dscales = nullptr; o.dL_drotations = nullptr; }
    return launch_preprocess_backward(v, g, radii, acc, dL_dinvdepth != nullptr, o, s);
}

int gsr_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix, uint8_t* present,
                     void* stream) {
    (void)projmatrix;   // the reference computes p_proj but never uses it (auxiliary.h:149-151)
    if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) return GSR_E_BADARG;
    if (P == 0) return 0;
    return launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream);
}

int gsr_dist2(int P, const float* points_xyz, float* mean_d2, char* temp, void* stream) {
    if (P < 0 || (P > 0 && (!points_xyz || !mean_d2 || !temp))) return GSR_E_BADARG;
    return launch_dist2(P, points_xyz, mean_d2, temp, (cudaStream_t)stream);
}

int gsr_dist2_noalloc(int P, const float* points_xyz, float* mean_d2, void* stream) {
    if (P < 0 || (P > 0 && (!points_xyz || !mean_d2))) return GSR_E_BADARG;
    return launch_dist2_noalloc(P, points_xyz, mean_d2, (cudaStream_t)stream);
}

// Parity helper (not on the hot path): reconstruct the reference's sorted 64-bit keys
// (tile << 32 | depth bits) for the instances of the last forward, for bit-exact comparison.
int gsr_debug_sorted_keys(const gsr_view* view, int R, const char* geom, const char* binning, const char* image,
                          uint64_t* keys_out, void* stream) {
    if (!view || view->W <= 0 || view->H <= 0 || !geom || !binning || !image || !keys_out) return GSR_E_BADARG;
    char* gp = const_cast<char*>(geom);
    char* ip = const_cast<char*>(image);
    GeomState g = GeomState::from(gp, (size_t)view->P);
    ImageState img = ImageState::from(ip, (size_t)view->W * view->H, (size_t)tiles_of(view->W, view->H));
    return launch_sorted_keys_debug(*view, g, BinningState::point_list_of(binning), img, R, keys_out, (cudaStream_t)stream);
}

int gsr_ssim_forward(int B, int CH, int H, int W, float C1, float C2, const float* img1, const float* img2,
                     float* ssim_map, float* dm_dmu1, float* dm_dsigma1_sq, float* dm_dsigma12, double* sums, void* stream) {
    if (B < 0 || CH < 0 || H <= 0 || W <= 0) return GSR_E_BADARG;
    const bool train = dm_dmu1 != nullptr;
    if ((dm_dsigma1_sq != nullptr) != train || (dm_dsigma12 != nullptr) != train) return GSR_E_BADARG;
    if ((long long)B * CH > 65535) return GSR_E_TOOLARGE;
    if (B * CH == 0) return 0;
    if (!img1 || !img2 || (!ssim_map && !sums && !train)) return GSR_E_BADARG;
    (void)C1; (void)C2;
    return launch_ssim_forward(B * CH, H, W, C1, C2, img1, img2, ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12, sums,
                               (cudaStream_t)stream);
}

int gsr_ssim_backward(int B, int CH, int H, int W, float C1, float C2, const float* img1, const float* img2,
                      const float* dL_dmap, float map_scale, float l1_scale, const float* upstream,
                      const float* dm_dmu1, const float* dm_dsigma1_sq, const float* dm_dsigma12, float* dL_dimg1,
                      void* stream) {
    (void)C1; (void)C2;   // the derivative maps already carry them (ssim.cu:272-366 takes them and never reads them)
    if (B < 0 || CH < 0 || H <= 0 || W <= 0) return GSR_E_BADARG;
    if ((long long)B * CH > 65535) return GSR_E_TOOLARGE;
    if (B * CH == 0) return 0;
    if (!img1 || !img2 || !dm_dmu1 || !dm_dsigma1_sq || !dm_dsigma12 || !dL_dimg1) return GSR_E_BADARG;
    return launch_ssim_backward(B * CH, H, W, img1, img2, dL_dmap, map_scale, l1_scale, upstream, dm_dmu1, dm_dsigma1_sq, dm_dsigma12,
                                dL_dimg1, (cudaStream_t)stream);
}

int gsr_adam_step(const gsr_adam_group* groups, int n_groups, double beta1, double beta2, double eps, const uint8_t* visible,
                  void* stream) {
    if (n_groups < 0 || n_groups > GSR_ADAM_MAX_GROUPS || (n_groups > 0 && !groups)) return GSR_E_BADARG;
    for (int k = 0; k < n_groups; k++) {
        const gsr_adam_group& g = groups[k];
        if (g.numel < 0 || g.row_len <= 0) return GSR_E_BADARG;
        if (g.numel > 0 && (!g.param || !g.grad || !g.exp_avg || !g.exp_avg_sq)) return GSR_E_BADARG;
        if (g.numel > (1ll << 40)) return GSR_E_TOOLARGE;
    }
    return launch_adam(groups, n_groups, beta1, beta2, eps, visible, (cudaStream_t)stream);
}

size_t gsr_comm_signal_bytes(void) { return sizeof(uint32_t) * (GSR_COMM_MAX_CTAS * GSR_COMM_MAX_PEERS + 16); }   // + error word

int gsr_comm_slice(long long numel, int rank, int world, long long* lo, long long* hi) {
    if (numel < 0 || (numel & 3) || world < 1 || world > GSR_COMM_MAX_PEERS || rank < 0 || rank >= world || !lo || !hi)
        return GSR_E_BADARG;
    comm_slice(numel / 4, rank, world, lo, hi);
    *lo *= 4;
    *hi *= 4;
    return 0;
}

int gsr_allreduce_sum(const gsr_peer_group* group, long long numel, int max_ctas, void* stream) {
    if (!group || numel < 0 || (numel & 3) || max_ctas < 0) return GSR_E_BADARG;
    const gsr_peer_group& g = *group;
    if (g.world < 1 || g.world > GSR_COMM_MAX_PEERS || g.rank < 0 || g.rank >= g.world) return GSR_E_BADARG;
    if (numel / 4 > (1ll << 40)) return GSR_E_TOOLARGE;
    if (numel == 0) return 0;
    for (int p = 0; p < g.world; p++)
        if (!g.buf[p] || !g.signal[p] || (reinterpret_cast<uintptr_t>(g.buf[p]) & 15) || (reinterpret_cast<uintptr_t>(g.signal[p]) & 3))
            return GSR_E_BADARG;
    if (reinterpret_cast<uintptr_t>(g.multicast) & 15) return GSR_E_BADARG;
    return launch_allreduce_sum(g, numel, max_ctas, (cudaStream_t)stream);
}

static bool peer_group_ok(const gsr_peer_group& g, bool need_signal) {
    if (g.world < 1 || g.world > GSR_COMM_MAX_PEERS || g.rank < 0 || g.rank >= g.world) return false;
    for (int p = 0; p < g.world; p++) {
        if (!g.buf[p] || (reinterpret_cast<uintptr_t>(g.buf[p]) & 15)) return false;
        if (need_signal && (!g.signal[p] || (reinterpret_cast<uintptr_t>(g.signal[p]) & 3))) return false;
    }
    return (reinterpret_cast<uintptr_t>(g.multicast) & 15) == 0;
}

int gsr_exchange_adam_step(const gsr_peer_group* grads, const gsr_peer_group* params, float* exp_avg, float* exp_avg_sq,
                           const gsr_adam_span* spans, int n_spans, double beta1, double beta2, double eps, long long numel,
                           int max_ctas, void* stream) {
    if (!grads || !params || numel < 0 || (numel & 3) || max_ctas < 0 || n_spans < 0 || n_spans > GSR_ADAM_MAX_SPANS) return GSR_E_BADARG;
    if (!peer_group_ok(*grads, true) || !peer_group_ok(*params, false)) return GSR_E_BADARG;
    if (grads->rank != params->rank || grads->world != params->world) return GSR_E_BADARG;
    if ((grads->multicast == nullptr) != (params->multicast == nullptr)) return GSR_E_BADARG;   // one data path for both
    if (numel / 4 > (1ll << 40)) return GSR_E_TOOLARGE;
    if (numel == 0) return 0;
    if (n_spans > 0 && !spans) return GSR_E_BADARG;
    long long prev = 0;
    for (int k = 0; k < n_spans; k++) {
        if (spans[k].begin < prev || spans[k].end < spans[k].begin || spans[k].end > numel || (spans[k].begin & 3)) return GSR_E_BADARG;
        prev = (spans[k].end + 3) & ~3ll;
    }
    if (!exp_avg || !exp_avg_sq || (reinterpret_cast<uintptr_t>(exp_avg) & 15) || (reinterpret_cast<uintptr_t>(exp_avg_sq) & 15))
        return GSR_E_BADARG;
    return launch_exchange_adam(*grads, *params, exp_avg, exp_avg_sq, spans, n_spans, beta1, beta2, eps, numel, max_ctas,
                                (cudaStream_t)stream);
}

size_t gsr_comm_emulation_scratch_bytes(void) { return comm_emulation_scratch_bytes(); }

int gsr_allreduce_sum_emulated(const gsr_peer_group* groups, int world, long long numel, int max_ctas, void* scratch_dev,
                               void* stream) {
    if (!groups || !scratch_dev || world < 1 || world > GSR_COMM_MAX_PEERS || numel < 0 || (numel & 3) || max_ctas < 0) return GSR_E_BADARG;
    if (numel == 0) return 0;
    for (int r = 0; r < world; r++) {
        if (groups[r].rank != r || groups[r].world != world || !peer_group_ok(groups[r], true)) return GSR_E_BADARG;
    }
    return launch_allreduce_sum_emulated(groups, world, numel, max_ctas, scratch_dev, (cudaStream_t)stream);
}

int gsr_exchange_adam_step_emulated(const gsr_peer_group* grads, const gsr_peer_group* params, float* const* exp_avg,
                                    float* const* exp_avg_sq, int world, const gsr_adam_span* spans, int n_spans, double beta1,
                                    double beta2, double eps, long long numel, int max_ctas, void* scratch_dev, void* stream) {
    if (!grads || !params || !exp_avg || !exp_avg_sq || !scratch_dev || world < 1 || world > GSR_COMM_MAX_PEERS || numel < 0 ||
        (numel & 3) || max_ctas < 0 || n_spans < 0 || n_spans > GSR_ADAM_MAX_SPANS || (n_spans > 0 && !spans))
        return GSR_E_BADARG;
    if (numel == 0) return 0;
    for (int r = 0; r < world; r++) {
        if (grads[r].rank != r || grads[r].world != world || !peer_group_ok(grads[r], true) || !peer_group_ok(params[r], false) ||
            !exp_avg[r] || !exp_avg_sq[r])
            return GSR_E_BADARG;
    }
    return launch_exchange_adam_emulated(grads, params, exp_avg, exp_avg_sq, world, spans, n_spans, beta1, beta2, eps, numel,
                                         max_ctas, scratch_dev, (cudaStream_t)stream);
}

size_t gsr_densify_plan_bytes(int P) { return P < 0 ? 0 : densify_plan_bytes((size_t)P); }

int gsr_densify_plan(int P, const float* opacity, const float* scaling, const float* grad_accum, const float* denom,
                     const gsr_densify_rule* rule, void* plan, uint32_t* counts_out, void* stream) {
    if (P < 0 || !rule || !plan) return GSR_E_BADARG;
    if (P > 0 && (!opacity || !scaling || !grad_accum || !denom)) return GSR_E_BADARG;
    if (!(rule->grad_threshold > 0.0f) || rule->n_split < 1) return GSR_E_BADARG;   // clones (zero gradient) must not split
    if (P >= (1 << 29)) return GSR_E_TOOLARGE;
    return launch_densify_plan(P, opacity, scaling, grad_accum, denom, *rule, plan, counts_out, (cudaStream_t)stream);
}

int gsr_prune_plan(int P, const uint8_t* prune_mask, void* plan, uint32_t* counts_out, void* stream) {
    if (P < 0 || !plan || (P > 0 && !prune_mask)) return GSR_E_BADARG;
    if (P >= (1 << 29)) return GSR_E_TOOLARGE;
    return launch_prune_plan(P, prune_mask, plan, counts_out, (cudaStream_t)stream);
}

int gsr_densify_apply(int P, const void* plan, const gsr_densify_rule* rule, const gsr_densify_tensor* tensors, int n_tensors,
                      const float* scaling, const float* rotation, const float* noise, void* stream) {
    if (P < 0 || !plan || n_tensors < 0 || n_tensors > GSR_DENSIFY_MAX_TENSORS || (n_tensors > 0 && !tensors)) return GSR_E_BADARG;
    for (int k = 0; k < n_tensors; k++) {
        const gsr_densify_tensor& t = tensors[k];
        if (t.row_len <= 0 || (P > 0 && (!t.src || !t.dst))) return GSR_E_BADARG;
        const int have = (t.src_m != nullptr) + (t.dst_m != nullptr) + (t.src_v != nullptr) + (t.dst_v != nullptr);
        if (have != 0 && have != 4) return GSR_E_BADARG;
        if (t.role != GSR_DENSIFY_COPY && t.role != GSR_DENSIFY_XYZ && t.role != GSR_DENSIFY_SCALING) return GSR_E_BADARG;
        if (t.role != GSR_DENSIFY_COPY && !rule) return GSR_E_BADARG;
        if (t.role == GSR_DENSIFY_XYZ && (t.row_len != 3 || (P > 0 && (!scaling || !rotation)))) return GSR_E_BADARG;
        if (t.role == GSR_DENSIFY_SCALING && t.row_len != 3) return GSR_E_BADARG;
    }
    return launch_densify_apply(P, plan, rule, tensors, n_tensors, scaling, rotation, noise, (cudaStream_t)stream);
}

int gsr_densify_stats(int P, const float* dL_dmean2D, const uint8_t* update_filter, const int* radii, float* grad_accum,
                      float* denom, float* max_radii2D, void* stream) {
    if (P < 0) return GSR_E_BADARG;
    if (P > 0 && (!dL_dmean2D || !grad_accum || !denom || (!update_filter && !radii))) return GSR_E_BADARG;
    return launch_densify_stats(P, dL_dmean2D, update_filter, radii, grad_accum, denom, max_radii2D, (cudaStream_t)stream);
}

int gsr_debug_blend_stats(const gsr_view* view, int R, const char* geom, const char* binning, const char* image,
                          unsigned long long* counts_dev, void* stream) {
    if (!view || view->W <= 0 || view->H <= 0 || !geom || !image || !counts_dev) return GSR_E_BADARG;
    if (R <= 0 || view->P <= 0) return 0;
    if (!binning) return GSR_E_BADARG;
    char* gp = const_cast<char*>(geom);
    char* ip = const_cast<char*>(image);
    GeomState g = GeomState::from(gp, (size_t)view->P);
    const uint32_t* point_list = BinningState::point_list_of(binning);
    ImageState img = ImageState::from(ip, (size_t)view->W * view->H, (size_t)tiles_of(view->W, view->H));
    int e = launch_blend_stats(*view, g, point_list, img, counts_dev, (cudaStream_t)stream);
    if (e) return e;
    return launch_cull_stats(*view, g, point_list, img, counts_dev, (cudaStream_t)stream);
}

int gsr_debug_set_exact_cull(int on) { return set_exact_cull(on); }
int gsr_debug_set_blend_variant(int v) { return set_blend_variant(v); }
int gsr_debug_set_comm_variant(int v) { return set_comm_variant(v); }

// Test hook: the library's own stable radix sort of (key, value) pairs on key bits [0, total_bits), as the depth sort
// (drop != 0: keys 0xFFFFFFFF are dropped, *n_valid_out = pairs kept), the Morton sort of distCUDA2 and the instance sort
// use it. Sorted pairs end up in (keys, vals). scratch: gsr_debug_sort_scratch_bytes(n) bytes.
size_t gsr_debug_sort_scratch_bytes(int n) { return sizeof(uint32_t) * sort_scratch_words((size_t)(n > 0 ? n : 1)); }
int gsr_debug_sort_pairs(uint32_t* keys, uint32_t* vals, uint32_t* keys_tmp, uint32_t* vals_tmp, int n, int total_bits, int drop,
                         uint32_t* scratch, uint32_t* n_valid_out, void* stream) {
    if (n < 0 || total_bits < 0 || total_bits > 32 || (n > 0 && (!keys || !vals || !keys_tmp || !vals_tmp || !scratch)))
        return GSR_E_BADARG;
    if (drop && !n_valid_out) return GSR_E_BADARG;
    cudaStream_t s = (cudaStream_t)stream;
    bool in_a = true;
    uint32_t* n_dev = nullptr;
    const int e = radix_sort_u32(keys, vals, keys_tmp, vals_tmp, n, total_bits, scratch, &in_a, false, s, drop ? &n_dev : nullptr);
    if (e) return e;
    cudaError_t ce = cudaSuccess;
    if (!in_a && n > 0) {
        ce = cudaMemcpyAsync(keys, keys_tmp, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToDevice, s);
        if (ce == cudaSuccess) ce = cudaMemcpyAsync(vals, vals_tmp, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToDevice, s);
    }
    if (ce == cudaSuccess && drop && n > 0) ce = cudaMemcpyAsync(n_valid_out, n_dev, sizeof(uint32_t), cudaMemcpyDeviceToDevice, s);
    return (int)ce;
}

int gsr_debug_fp32_peak(int iters, float* out, int* sms_out, void* stream) {
    if (iters <= 0 || !out) return GSR_E_BADARG;
    return launch_fp32_peak(iters, out, sms_out, (cudaStream_t)stream);
}

int gsr_debug_exp_mismatches(unsigned long long* mismatches_dev, void* stream) {
    if (!mismatches_dev) return GSR_E_BADARG;
    return launch_exp_check(mismatches_dev, (cudaStream_t)stream);
}

}  // extern "C"
