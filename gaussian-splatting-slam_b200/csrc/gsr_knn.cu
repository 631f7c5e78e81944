// gsr_knn.cu -- distCUDA2: mean squared distance to the 3 nearest other points (K1-K4).
// Replaces SimpleKNN::knn (submodules/simple-knn/simple_knn.cu:186-222). Same algorithm family
// (Morton order, 1024-point boxes, box pruning, exact 3-NN) re-laid for coalesced access:
//   * bbox by one reduction kernel with ordered-int atomics, NO host synchronisation (the reference
//     does two blocking cudaMemcpy, simple_knn.cu:194-201); bbox always includes the origin as the
//     reference's init = {0,0,0} implies;
//   * points are COPIED into Morton order once (float4 x,y,z,id) so every later access is a
//     coalesced stream instead of points[indices[i]] gathers (simple_knn.cu:154-181);
//   * the search works on TWO box levels: the reference's 1024-point boxes and 64-point sub-boxes. A CTA of 256
//     consecutive (spatially coherent) points first finds every thread's 3 nearest among the CTA's own points (these are
//     real candidates, so the pruning radius is tight from the start -- the reference's radius comes from the +-3
//     neighbours along the Morton curve, simple_knn.cu:157-162, which can be far away), then prunes 1024-point boxes
//     and their sub-boxes against the CTA's bounding box, stages the surviving sub-boxes four at a time in shared
//     memory, and every thread scans only those whose box distance beats its current third-best.
// The result is the exact 3-NN mean, so it equals the reference up to fp32 rounding of d^2.
#include "gsr_common.cuh"
#include <mutex>
#include <cfloat>

namespace gsr {

constexpr int kBox = 1024;   // simple_knn.cu:12 BOX_SIZE
constexpr int kSub = 64;     // points per sub-box
constexpr int kSubPerBox = kBox / kSub;

struct KnnTemp {
    uint32_t* bbox;        // [8] encoded min xyz (0..2), max xyz (4..6)
    uint32_t* codes;       // [P]
    uint32_t* codes_tmp;   // [P]
    uint32_t* idx;         // [P]
    uint32_t* idx_tmp;     // [P]
    float4* sorted;        // [P] x,y,z,original id (bits)
    float4* boxes;         // [2*nb] min, max
    float4* sub;           // [2*nsub] min, max of the 64-point sub-boxes
    uint32_t* sort_scratch;
    static KnnTemp from(char*& p, size_t P) {
        KnnTemp t;
        t.bbox = carve<uint32_t>(p, 8);
        t.codes = carve<uint32_t>(p, P);
        t.codes_tmp = carve<uint32_t>(p, P);
        t.idx = carve<uint32_t>(p, P);
        t.idx_tmp = carve<uint32_t>(p, P);
        t.sorted = carve<float4>(p, P);
        t.boxes = carve<float4>(p, 2 * ((P + kBox - 1) / kBox) + 2);
        t.sub = carve<float4>(p, 2 * ((P + kSub - 1) / kSub) + 2);
        t.sort_scratch = carve<uint32_t>(p, sort_scratch_words(P));
        return t;
    }
};

size_t dist2_temp_bytes(int P) {
    char* p = nullptr;
    KnnTemp::from(p, (size_t)(P > 0 ? P : 1));
    return (size_t)p + 256;
}

__device__ __forceinline__ uint32_t enc_f(float f) {
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float dec_f(uint32_t u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__global__ void __launch_bounds__(256) bbox_kernel(int P, const float* __restrict__ pts, uint32_t* __restrict__ bbox) {
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const float v = __ldg(pts + 3 * (size_t)i + a);
            mn[a] = fminf(mn[a], v);
            mx[a] = fmaxf(mx[a], v);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; a++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[a] = fminf(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
            mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            atomicMin(&bbox[a], enc_f(mn[a]));
            atomicMax(&bbox[4 + a], enc_f(mx[a]));
        }
    }
}

__device__ __forceinline__ uint32_t spread10(uint32_t x) {   // simple_knn.cu:46-53
    x = (x | (x << 16)) & 0x030000FF;
    x = (x | (x << 8)) & 0x0300F00F;
    x = (x | (x << 4)) & 0x030C30C3;
    x = (x | (x << 2)) & 0x09249249;
    return x;
}

__global__ void __launch_bounds__(256) morton_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ bbox,
                                                      uint32_t* __restrict__ codes, uint32_t* __restrict__ idx) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    // reduction identity of the reference is {0,0,0}: the box always contains the origin
    const float mnx = fminf(0.f, dec_f(bbox[0])), mny = fminf(0.f, dec_f(bbox[1])), mnz = fminf(0.f, dec_f(bbox[2]));
    const float mxx = fmaxf(0.f, dec_f(bbox[4])), mxy = fmaxf(0.f, dec_f(bbox[5])), mxz = fmaxf(0.f, dec_f(bbox[6]));
    const float px = __ldg(pts + 3 * (size_t)i), py = __ldg(pts + 3 * (size_t)i + 1), pz = __ldg(pts + 3 * (size_t)i + 2);
    const uint32_t x = spread10((uint32_t)(((px - mnx) / (mxx - mnx)) * ((1 << 10) - 1)));   // simple_knn.cu:55-62
    const uint32_t y = spread10((uint32_t)(((py - mny) / (mxy - mny)) * ((1 << 10) - 1)));
    const uint32_t z = spread10((uint32_t)(((pz - mnz) / (mxz - mnz)) * ((1 << 10) - 1)));
    codes[i] = x | (y << 1) | (z << 2);
    idx[i] = (uint32_t)i;
}

__global__ void __launch_bounds__(256) gather_points_kernel(int P, const float* __restrict__ pts,
                                                             const uint32_t* __restrict__ idx, float4* __restrict__ sorted) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= P) return;
    const uint32_t i = __ldg(idx + j);
    sorted[j] = make_float4(__ldg(pts + 3 * (size_t)i), __ldg(pts + 3 * (size_t)i + 1), __ldg(pts + 3 * (size_t)i + 2),
                            __uint_as_float(i));
}

// one warp per 64-point sub-box (two points per lane); 8 sub-boxes per CTA
__global__ void __launch_bounds__(256) sub_minmax_kernel(int P, const float4* __restrict__ sorted, float4* __restrict__ sub) {
    const int lane = threadIdx.x & 31, sb = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int base = sb * kSub;
    if (base >= P) return;
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
#pragma unroll
    for (int e = lane; e < kSub; e += 32) {
        if (base + e < P) {
            const float4 p = __ldg(sorted + base + e);
            mn[0] = fminf(mn[0], p.x); mn[1] = fminf(mn[1], p.y); mn[2] = fminf(mn[2], p.z);
            mx[0] = fmaxf(mx[0], p.x); mx[1] = fmaxf(mx[1], p.y); mx[2] = fmaxf(mx[2], p.z);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; a++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[a] = fminf(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
            mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
        }
    }
    if (lane == 0) {
        sub[2 * sb] = make_float4(mn[0], mn[1], mn[2], 0.f);
        sub[2 * sb + 1] = make_float4(mx[0], mx[1], mx[2], 0.f);
    }
}

// 1024-point boxes (simple_knn.cu:79-118) from their 16 sub-boxes: one thread per box
__global__ void __launch_bounds__(256) box_from_sub_kernel(int nb, int nsub, const float4* __restrict__ sub, float4* __restrict__ boxes) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nb) return;
    float4 mn = make_float4(FLT_MAX, FLT_MAX, FLT_MAX, 0.f), mx = make_float4(-FLT_MAX, -FLT_MAX, -FLT_MAX, 0.f);
    for (int k = 0; k < kSubPerBox; k++) {
        const int sb = b * kSubPerBox + k;
        if (sb >= nsub) break;
        const float4 a = __ldg(sub + 2 * sb), c = __ldg(sub + 2 * sb + 1);
        mn.x = fminf(mn.x, a.x); mn.y = fminf(mn.y, a.y); mn.z = fminf(mn.z, a.z);
        mx.x = fmaxf(mx.x, c.x); mx.y = fmaxf(mx.y, c.y); mx.z = fmaxf(mx.z, c.z);
    }
    boxes[2 * b] = mn;
    boxes[2 * b + 1] = mx;
}

__device__ __forceinline__ void k_best3(const float4& ref, const float4& point, float* knn) {   // simple_knn.cu:132-146
    const float dx = point.x - ref.x, dy = point.y - ref.y, dz = point.z - ref.z;
    float dist = dx * dx + dy * dy + dz * dz;
#pragma unroll
    for (int j = 0; j < 3; j++) {
        if (knn[j] > dist) { const float t = knn[j]; knn[j] = dist; dist = t; }
    }
}

__device__ __forceinline__ float dist_box_point(const float4& bmin, const float4& bmax, const float4& p) {   // :120-130
    float dx = 0.f, dy = 0.f, dz = 0.f;
    if (p.x < bmin.x || p.x > bmax.x) dx = fminf(fabsf(p.x - bmin.x), fabsf(p.x - bmax.x));
    if (p.y < bmin.y || p.y > bmax.y) dy = fminf(fabsf(p.y - bmin.y), fabsf(p.y - bmax.y));
    if (p.z < bmin.z || p.z > bmax.z) dz = fminf(fabsf(p.z - bmin.z), fabsf(p.z - bmax.z));
    return dx * dx + dy * dy + dz * dz;
}

// squared gap between two axis-aligned boxes: a lower bound of dist_box_point for every point of the first one
__device__ __forceinline__ float box_gap2(const float* r, const float4& bmin, const float4& bmax) {
    const float gx = fmaxf(0.f, fmaxf(bmin.x - r[3], r[0] - bmax.x));
    const float gy = fmaxf(0.f, fmaxf(bmin.y - r[4], r[1] - bmax.y));
    const float gz = fmaxf(0.f, fmaxf(bmin.z - r[5], r[2] - bmax.z));
    return gx * gx + gy * gy + gz * gz;
}

// B = threads (= consecutive Morton-ordered points) per CTA: 256 for large clouds, 64 for small ones (a SLAM keyframe's
// 25k new points make 98 CTAs of 256 -- fewer than SMs, each a long serial chain of staging rounds -- but 391 CTAs of 64
// with tighter bounding boxes). kStage = B / kSub sub-boxes are staged per round (one point per thread).
template <int B>
__global__ void __launch_bounds__(B) knn_kernel(int P, const float4* __restrict__ sorted, const float4* __restrict__ boxes,
                                                   int nb, const float4* __restrict__ sub, int nsub, float* __restrict__ out) {
    constexpr int kStage = B / kSub;
    __shared__ float4 s_pts[B];
    __shared__ float4 s_sb[kStage][2];
    __shared__ int s_c1[B], s_c0[B];
    __shared__ int s_n1, s_n0;
    __shared__ float s_red[7][B / 32];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int cbase = blockIdx.x * B;
    const int pos = cbase + tid;
    const bool valid = pos < P;
    const float4 me = valid ? __ldg(sorted + pos) : make_float4(0.f, 0.f, 0.f, 0.f);
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    // ---- phase 1: the CTA's own B points (B / 64 sub-boxes), brute force. These are real candidates.
    const int own = min(B, P - cbase);
    s_pts[tid] = me;
    __syncthreads();
    if (valid) {
        for (int e = 0; e < own; e++) {
            if (e == tid) continue;
            k_best3(me, s_pts[e], best);
        }
    }
    // CTA bounding box and the largest third-best distance among its points
    float r[7] = {valid ? me.x : FLT_MAX, valid ? me.y : FLT_MAX, valid ? me.z : FLT_MAX,
                  valid ? me.x : -FLT_MAX, valid ? me.y : -FLT_MAX, valid ? me.z : -FLT_MAX, valid ? best[2] : 0.f};
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int a = 0; a < 3; a++) r[a] = fminf(r[a], __shfl_xor_sync(0xffffffffu, r[a], o));
#pragma unroll
        for (int a = 3; a < 7; a++) r[a] = fmaxf(r[a], __shfl_xor_sync(0xffffffffu, r[a], o));
    }
    if (lane == 0)
        for (int a = 0; a < 7; a++) s_red[a][w] = r[a];
    __syncthreads();
    for (int k = 0; k < B / 32; k++) {
#pragma unroll
        for (int a = 0; a < 3; a++) r[a] = fminf(r[a], s_red[a][k]);
#pragma unroll
        for (int a = 3; a < 7; a++) r[a] = fmaxf(r[a], s_red[a][k]);
    }
    const float rmax = r[6];
    const int own_sb0 = cbase / kSub, own_sb1 = own_sb0 + B / kSub;      // sub-boxes already searched in phase 1

    // ---- phase 2: 1024-point boxes -> their sub-boxes -> points
    for (int c0 = 0; c0 < nb; c0 += B) {
        __syncthreads();
        if (tid == 0) s_n1 = 0;
        __syncthreads();
        const int b = c0 + tid;
        if (b < nb) {
            const float4 bmin = __ldg(boxes + 2 * b), bmax = __ldg(boxes + 2 * b + 1);
            if (!(box_gap2(r, bmin, bmax) > rmax)) s_c1[atomicAdd(&s_n1, 1)] = b;
        }
        __syncthreads();
        const int n1 = s_n1;
        for (int g0 = 0; g0 < n1; g0 += B / kSubPerBox) {                // B / 16 boxes -> B sub-boxes, one per thread
            __syncthreads();
            if (tid == 0) s_n0 = 0;
            __syncthreads();
            const int k = g0 + tid / kSubPerBox;
            if (k < n1) {
                const int sb = s_c1[k] * kSubPerBox + (tid % kSubPerBox);
                if (sb < nsub && (sb < own_sb0 || sb >= own_sb1)) {
                    const float4 bmin = __ldg(sub + 2 * sb), bmax = __ldg(sub + 2 * sb + 1);
                    if (!(box_gap2(r, bmin, bmax) > rmax)) s_c0[atomicAdd(&s_n0, 1)] = sb;
                }
            }
            __syncthreads();
            const int n0 = s_n0;
            for (int r0 = 0; r0 < n0; r0 += kStage) {
                __syncthreads();                                         // the previous round's readers are done
                const int q = tid / kSub, e = tid % kSub;
                if (r0 + q < n0) {
                    const int sb = s_c0[r0 + q];
                    const int idx = sb * kSub + e;
                    if (idx < P) s_pts[tid] = __ldg(sorted + idx);
                    if (e < 2) s_sb[q][e] = __ldg(sub + 2 * sb + e);
                }
                __syncthreads();
                if (valid) {
                    const int nq = min(kStage, n0 - r0);
                    for (int qq = 0; qq < nq; qq++) {
                        const float d = dist_box_point(s_sb[qq][0], s_sb[qq][1], me);
                        if (d > best[2]) continue;                       // simple_knn.cu:171-173
                        const int cnt = min(kSub, P - s_c0[r0 + qq] * kSub);
                        for (int ee = 0; ee < cnt; ee++) k_best3(me, s_pts[qq * kSub + ee], best);
                    }
                }
            }
        }
    }
    if (valid) out[__float_as_uint(me.w)] = (best[0] + best[1] + best[2]) / 3.0f;   // simple_knn.cu:183
}

int launch_dist2(int P, const float* pts, float* out, char* temp, cudaStream_t s) {
    if (P <= 0) return 0;
    char* p = temp;
    KnnTemp t = KnnTemp::from(p, (size_t)P);
    cudaError_t e = cudaMemsetAsync(t.bbox, 0xFF, 16, s);        // encoded +max for the three minima
    if (e != cudaSuccess) return (int)e;
    e = cudaMemsetAsync(t.bbox + 4, 0x00, 16, s);                // encoded -max for the three maxima
    if (e != cudaSuccess) return (int)e;
    const int blocks = (P + 255) / 256;
    GSR_LAUNCH("knn_bbox", false, s, (bbox_kernel<<<min(blocks, 148 * 8), 256, 0, s>>>(P, pts, t.bbox)));
    GSR_LAUNCH("knn_morton", false, s, (morton_kernel<<<blocks, 256, 0, s>>>(P, pts, t.bbox, t.codes, t.idx)));
    bool in_a = true;
    int rc = radix_sort_u32(t.codes, t.idx, t.codes_tmp, t.idx_tmp, P, 30, t.sort_scratch, &in_a, false, s);
    if (rc) return rc;
    const uint32_t* sorted_idx = in_a ? t.idx : t.idx_tmp;
    GSR_LAUNCH("knn_gather", false, s, (gather_points_kernel<<<blocks, 256, 0, s>>>(P, pts, sorted_idx, t.sorted)));
    const int nb = (P + kBox - 1) / kBox, nsub = (P + kSub - 1) / kSub;
    GSR_LAUNCH("knn_subboxes", false, s, (sub_minmax_kernel<<<(nsub + 7) / 8, 256, 0, s>>>(P, t.sorted, t.sub)));
    GSR_LAUNCH("knn_boxes", false, s, (box_from_sub_kernel<<<(nb + 255) / 256, 256, 0, s>>>(nb, nsub, t.sub, t.boxes)));
    int sms = 148;
    { const int se = sm_count(&sms); if (se) return se; }
    if (blocks < 2 * sms)                                           // small cloud: more, smaller CTAs (see knn_kernel)
        GSR_LAUNCH("knn_search", false, s, (knn_kernel<64><<<(P + 63) / 64, 64, 0, s>>>(P, t.sorted, t.boxes, nb, t.sub, nsub, out)));
    else
        GSR_LAUNCH("knn_search", false, s, (knn_kernel<256><<<blocks, 256, 0, s>>>(P, t.sorted, t.boxes, nb, t.sub, nsub, out)));
    return 0;
}

// ------------------------------------------------------------------ no-allocation variant (gsr_dist2_noalloc)
// Same kernels as launch_dist2, with the scratch in a file-scope device array instead of a caller-supplied buffer: for
// callers that cannot obtain device memory at the call (CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc). Every GPU that loads the library holds the array (4.5 MiB) until exit. All calls on a device SHARE it:
// calls that overlap on the GPU (different streams) race. The chain clears what it needs at the start of every call
// (bounding box words, the sort's tables), so the previous call's contents do not matter.
constexpr size_t kNoAllocScratchBytes = 4608u * 1024u;       // >= dist2_temp_bytes(GSR_DIST2_NOALLOC_MAX_POINTS) = 4400256
static __device__ __align__(256) unsigned char g_dist2_static_scratch[kNoAllocScratchBytes];

int launch_dist2_noalloc(int P, const float* pts, float* out, cudaStream_t s) {
    if (P <= 0) return 0;
    if (P > GSR_DIST2_NOALLOC_MAX_POINTS || dist2_temp_bytes(P) > kNoAllocScratchBytes) return GSR_E_STATIC_LIMIT;
    int dev = 0;
    { const int e = current_device(&dev); if (e) return e; }
    if (dev < 0 || dev >= kMaxDevices) return GSR_E_BADARG;
    // In host code the array's name is only a stand-in; the device address comes from cudaGetSymbolAddress, once per
    // device, with that device current (it is: current_device() just reported it).
    static std::mutex mu;
    static char* addr[kMaxDevices] = {};
    char* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lk(mu);
        if (!addr[dev]) {
            void* p = nullptr;
            const cudaError_t e = cudaGetSymbolAddress(&p, g_dist2_static_scratch);
            if (e != cudaSuccess) return (int)e;
            if (!p) return GSR_E_BADARG;
            addr[dev] = static_cast<char*>(p);
        }
        scratch = addr[dev];
    }
    return launch_dist2(P, pts, out, scratch, s);
}

}  // namespace gsr
