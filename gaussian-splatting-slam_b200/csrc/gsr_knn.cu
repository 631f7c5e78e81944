// gsr_knn.cu -- distCUDA2: mean squared distance to the 3 nearest other points (K1-K4).
// Replaces SimpleKNN::knn (submodules/simple-knn/simple_knn.cu:186-222). Same algorithm family
// (Morton order, 1024-point boxes, box pruning, exact 3-NN) re-laid for coalesced access:
//   * bbox by one reduction kernel with ordered-int atomics, NO host synchronisation (the reference
//     does two blocking cudaMemcpy, simple_knn.cu:194-201); bbox always includes the origin as the
//     reference's init = {0,0,0} implies;
//   * points are COPIED into Morton order once (float4 x,y,z,id) so every later access is a
//     coalesced stream instead of points[indices[i]] gathers (simple_knn.cu:154-181);
//   * a CTA of 256 consecutive (spatially coherent) points first prunes boxes against the CTA's own
//     bounding box, then stages each surviving box in shared memory once for all its threads.
// The result is the exact 3-NN mean, so it equals the reference up to fp32 rounding of d^2.
#include "gsr_common.cuh"
#include <cfloat>

namespace gsr {

constexpr int kBox = 1024;   // simple_knn.cu:12 BOX_SIZE

struct KnnTemp {
    uint32_t* bbox;        // [8] encoded min xyz (0..2), max xyz (4..6)
    uint32_t* codes;       // [P]
    uint32_t* codes_tmp;   // [P]
    uint32_t* idx;         // [P]
    uint32_t* idx_tmp;     // [P]
    float4* sorted;        // [P] x,y,z,original id (bits)
    float4* boxes;         // [2*nb] min, max
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

// one CTA of 256 threads per box of 1024 sorted points (simple_knn.cu:79-118)
__global__ void __launch_bounds__(256) box_minmax_kernel(int P, const float4* __restrict__ sorted, float4* __restrict__ boxes) {
    __shared__ float red[6][8];
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    const int base = blockIdx.x * kBox;
    for (int e = threadIdx.x; e < kBox; e += 256) {
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
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) {
#pragma unroll
        for (int a = 0; a < 3; a++) { red[a][w] = mn[a]; red[3 + a][w] = mx[a]; }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 8; k++)
            for (int a = 0; a < 3; a++) { mn[a] = fminf(mn[a], red[a][k]); mx[a] = fmaxf(mx[a], red[3 + a][k]); }
        boxes[2 * blockIdx.x] = make_float4(mn[0], mn[1], mn[2], 0.f);
        boxes[2 * blockIdx.x + 1] = make_float4(mx[0], mx[1], mx[2], 0.f);
    }
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

__global__ void __launch_bounds__(256) knn_kernel(int P, const float4* __restrict__ sorted, const float4* __restrict__ boxes,
                                                   int nb, float* __restrict__ out) {
    __shared__ float4 s_pts[kBox];
    __shared__ int s_cand[256];
    __shared__ int s_ncand;
    __shared__ float s_red[7][8];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int pos = blockIdx.x * 256 + tid;
    const bool valid = pos < P;
    const float4 me = valid ? __ldg(sorted + pos) : make_float4(0.f, 0.f, 0.f, 0.f);
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    float reject = 0.f;
    if (valid) {
        for (int i = max(0, pos - 3); i <= min(P - 1, pos + 3); i++) {   // simple_knn.cu:157-162
            if (i == pos) continue;
            k_best3(me, __ldg(sorted + i), best);
        }
        reject = best[2];
        best[0] = best[1] = best[2] = FLT_MAX;
    }
    // CTA bounding box and the largest rejection radius among its points
    float r[7] = {valid ? me.x : FLT_MAX, valid ? me.y : FLT_MAX, valid ? me.z : FLT_MAX,
                  valid ? me.x : -FLT_MAX, valid ? me.y : -FLT_MAX, valid ? me.z : -FLT_MAX, valid ? reject : 0.f};
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
    for (int k = 0; k < 8; k++) {
#pragma unroll
        for (int a = 0; a < 3; a++) r[a] = fminf(r[a], s_red[a][k]);
#pragma unroll
        for (int a = 3; a < 7; a++) r[a] = fmaxf(r[a], s_red[a][k]);
    }
    const float rmax = r[6];

    for (int c0 = 0; c0 < nb; c0 += 256) {
        __syncthreads();
        if (tid == 0) s_ncand = 0;
        __syncthreads();
        const int b = c0 + tid;
        if (b < nb) {
            const float4 bmin = __ldg(boxes + 2 * b), bmax = __ldg(boxes + 2 * b + 1);
            // gap between the CTA box and box b: a lower bound of dist_box_point for every thread
            const float gx = fmaxf(0.f, fmaxf(bmin.x - r[3], r[0] - bmax.x));
            const float gy = fmaxf(0.f, fmaxf(bmin.y - r[4], r[1] - bmax.y));
            const float gz = fmaxf(0.f, fmaxf(bmin.z - r[5], r[2] - bmax.z));
            if (!(gx * gx + gy * gy + gz * gz > rmax)) s_cand[atomicAdd(&s_ncand, 1)] = b;
        }
        __syncthreads();
        const int ncand = s_ncand;
        for (int ci = 0; ci < ncand; ci++) {
            const int bb = s_cand[ci];
            const float4 bmin = __ldg(boxes + 2 * bb), bmax = __ldg(boxes + 2 * bb + 1);
            bool want = false;
            if (valid) {
                const float d = dist_box_point(bmin, bmax, me);
                want = !(d > reject || d > best[2]);                       // simple_knn.cu:171-173
            }
            if (!__syncthreads_or(want)) continue;
            const int bbase = bb * kBox, cnt = min(kBox, P - bbase);
            for (int e = tid; e < cnt; e += 256) s_pts[e] = __ldg(sorted + bbase + e);
            __syncthreads();
            if (want) {
                for (int e = 0; e < cnt; e++) {
                    if (bbase + e == pos) continue;
                    k_best3(me, s_pts[e], best);
                }
            }
            __syncthreads();
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
    const int nb = (P + kBox - 1) / kBox;
    GSR_LAUNCH("knn_boxes", false, s, (box_minmax_kernel<<<nb, 256, 0, s>>>(P, t.sorted, t.boxes)));
    GSR_LAUNCH("knn_search", false, s, (knn_kernel<<<blocks, 256, 0, s>>>(P, t.sorted, t.boxes, nb, out)));
    return 0;
}

}  // namespace gsr
