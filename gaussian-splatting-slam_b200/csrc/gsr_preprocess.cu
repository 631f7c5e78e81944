// gsr_preprocess.cu -- per-Gaussian stages: forward preprocess (A1-A5, A-AA), markVisible (A15),
// backward preprocess (A12+A13 fused, A-AA, A-DEPTH). One thread per Gaussian, HBM-bound.
//
// Bit-exactness: radii / tiles_touched / means2D / depth bits must equal the
// reference's, which is compiled with nvcc's default -fmad=true. The forward therefore keeps the
// reference's EXPRESSION TREES (operand order, association, literal zeros of the sparse matrices)
// so that the compiler's FMA contraction lands on the same operations; only the spelling is ours.
// Reference: submodules/diff-gaussian-rasterization/cuda_rasterizer/forward.cu:20-256,
// auxiliary.h:41-77,139-164, third_party/glm/glm/detail/type_mat3x3.inl:486-520 (operator*).
#include "gsr_common.cuh"
#include <cstdio>

namespace gsr {

// ---- minimal column-major 3x3 / vec3 algebra with glm's operation order -------------------
struct V3 { float x, y, z; };
__device__ __forceinline__ V3 operator*(float s, V3 v) { return {s * v.x, s * v.y, s * v.z}; }
__device__ __forceinline__ V3 operator*(V3 v, float s) { return {v.x * s, v.y * s, v.z * s}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator+(V3 a, float s) { return {a.x + s, a.y + s, a.z + s}; }
__device__ __forceinline__ V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
__device__ __forceinline__ float dot3(V3 a, V3 b) { V3 t = {a.x * b.x, a.y * b.y, a.z * b.z}; return t.x + t.y + t.z; }

struct M3 { float c[3][3]; };   // c[column][row]
__device__ __forceinline__ M3 make_m3(float x0, float y0, float z0, float x1, float y1, float z1,
                                      float x2, float y2, float z2) {
    M3 m; m.c[0][0] = x0; m.c[0][1] = y0; m.c[0][2] = z0; m.c[1][0] = x1; m.c[1][1] = y1; m.c[1][2] = z1;
    m.c[2][0] = x2; m.c[2][1] = y2; m.c[2][2] = z2; return m;
}
__device__ __forceinline__ M3 mul(const M3& a, const M3& b) {
    M3 r;
#pragma unroll
    for (int j = 0; j < 3; j++)
#pragma unroll
        for (int i = 0; i < 3; i++)
            r.c[j][i] = a.c[0][i] * b.c[j][0] + a.c[1][i] * b.c[j][1] + a.c[2][i] * b.c[j][2];
    return r;
}
__device__ __forceinline__ M3 transpose(const M3& a) {
    M3 r;
#pragma unroll
    for (int j = 0; j < 3; j++)
#pragma unroll
        for (int i = 0; i < 3; i++) r.c[j][i] = a.c[i][j];
    return r;
}

__device__ __forceinline__ float3 xform43(const float3& p, const float* m) {
    return {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
            m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]};
}
__device__ __forceinline__ float4 xform44(const float3& p, const float* m) {
    return {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
            m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14], m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15]};
}
__device__ __forceinline__ float ndc_to_pix(float v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }

__device__ __forceinline__ void tile_rect(const float2 p, int max_radius, uint2& rmin, uint2& rmax, dim3 grid) {
    rmin = {min(grid.x, max((int)0, (int)((p.x - max_radius) / GSR_TILE))),
            min(grid.y, max((int)0, (int)((p.y - max_radius) / GSR_TILE)))};
    rmax = {min(grid.x, max((int)0, (int)((p.x + max_radius + GSR_TILE - 1) / GSR_TILE))),
            min(grid.y, max((int)0, (int)((p.y + max_radius + GSR_TILE - 1) / GSR_TILE)))};
}

__device__ const float kC0 = 0.28209479177387814f;
__device__ const float kC1 = 0.4886025119029199f;
__device__ const float kC2[] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                -1.0925484305920792f, 0.5462742152960396f};
__device__ const float kC3[] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                                -0.5900435899266435f};

// cov3D from scale / rotation (forward.cu:118-152)
__device__ __forceinline__ void cov3d_of(const V3 scale, float mod, const float4 q, float* cov3D) {
    M3 S = make_m3(1.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f, 1.0f);
    S.c[0][0] = mod * scale.x; S.c[1][1] = mod * scale.y; S.c[2][2] = mod * scale.z;
    float r = q.x, x = q.y, y = q.z, z = q.w;
    // The nine entries share products, and which product of "a*b +- c*d" ends up inside the FMA decides the
    // last bit. Pinned with explicit intrinsics to the choice ptxas makes for the reference (its SASS).
    const float xz = __fmul_rn(x, z), rx = __fmul_rn(r, x), rz = __fmul_rn(r, z);
    const float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
    const float xz_p_ry = __fmaf_rn(r, y, xz), xz_m_ry = __fmaf_rn(-r, y, xz);
    const float yz_m_rx = __fmaf_rn(y, z, -rx), yz_p_rx = __fmaf_rn(y, z, rx);
    const float xy_m_rz = __fmaf_rn(x, y, -rz), xy_p_rz = __fmaf_rn(x, y, rz);
    const float xx_p_yy = __fmaf_rn(x, x, yy), yy_p_zz = __fadd_rn(yy, zz), xx_p_zz = __fmaf_rn(x, x, zz);
    M3 R = make_m3(__fsub_rn(1.f, __fadd_rn(yy_p_zz, yy_p_zz)), __fadd_rn(xy_m_rz, xy_m_rz), __fadd_rn(xz_p_ry, xz_p_ry),
                   __fadd_rn(xy_p_rz, xy_p_rz), __fsub_rn(1.f, __fadd_rn(xx_p_zz, xx_p_zz)), __fadd_rn(yz_m_rx, yz_m_rx),
                   __fadd_rn(xz_m_ry, xz_m_ry), __fadd_rn(yz_p_rx, yz_p_rx), __fsub_rn(1.f, __fadd_rn(xx_p_yy, xx_p_yy)));
    M3 Mm = mul(S, R);
    M3 Sigma = mul(transpose(Mm), Mm);
    cov3D[0] = Sigma.c[0][0]; cov3D[1] = Sigma.c[0][1]; cov3D[2] = Sigma.c[0][2];
    cov3D[3] = Sigma.c[1][1]; cov3D[4] = Sigma.c[1][2]; cov3D[5] = Sigma.c[2][2];
}

// EWA projection (forward.cu:74-113), WITHOUT the +0.3 low-pass (added by the caller).
__device__ __forceinline__ float3 cov2d_of(const float3& mean, float focal_x, float focal_y, float tan_fovx,
                                           float tan_fovy, const float* cov3D, const float* viewmatrix) {
    float3 t = xform43(mean, viewmatrix);
    const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
    const float txtz = t.x / t.z, tytz = t.y / t.z;
    t.x = min(limx, max(-limx, txtz)) * t.z;
    t.y = min(limy, max(-limy, tytz)) * t.z;
    M3 J = make_m3(focal_x / t.z, 0.0f, -(focal_x * t.x) / (t.z * t.z), 0.0f, focal_y / t.z,
                   -(focal_y * t.y) / (t.z * t.z), 0, 0, 0);
    M3 Wm = make_m3(viewmatrix[0], viewmatrix[4], viewmatrix[8], viewmatrix[1], viewmatrix[5], viewmatrix[9],
                    viewmatrix[2], viewmatrix[6], viewmatrix[10]);
    M3 T = mul(Wm, J);
    M3 Vrk = make_m3(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2], cov3D[4], cov3D[5]);
    M3 cov = mul(mul(transpose(T), transpose(Vrk)), T);
    return {cov.c[0][0], cov.c[0][1], cov.c[1][1]};
}

// Coefficient k of Gaussian idx lives at sh_ptr(...)[0..2]: in the fused layout all M coefficients are contiguous; with
// a separate degree-0 tensor (gsr_view::sh_dc) coefficient 0 comes from there and shs holds the other M-1.
__device__ __forceinline__ const float* sh_ptr(const float* shs, const float* dc, int idx, int M, int k) {
    if (dc == nullptr) return shs + ((size_t)idx * M + k) * 3;
    return k == 0 ? dc + (size_t)idx * 3 : shs + ((size_t)idx * (M - 1) + (k - 1)) * 3;
}
__device__ __forceinline__ float* sh_ptr(float* shs, float* dc, int idx, int M, int k) {
    if (dc == nullptr) return shs + ((size_t)idx * M + k) * 3;
    return k == 0 ? dc + (size_t)idx * 3 : shs + ((size_t)idx * (M - 1) + (k - 1)) * 3;
}

template <bool SH16>
__device__ __forceinline__ V3 sh_color(int idx, int deg, int max_coeffs, V3 pos, V3 campos, const float* shs,
                                       const float* dc, uint8_t& clamp_bits) {
    V3 dir = pos - campos;
    dir = dir / sqrtf(dot3(dir, dir));
    V3 sh[16];
    const int nb = (deg + 1) * (deg + 1);
    if (SH16) {
        const float4* p4 = reinterpret_cast<const float4*>(shs + (size_t)idx * 48);
        float f[48];
#pragma unroll
        for (int k = 0; k < 12; k++) {
            if (k * 4 < nb * 3) {
                float4 v = __ldg(p4 + k);
                f[4 * k] = v.x; f[4 * k + 1] = v.y; f[4 * k + 2] = v.z; f[4 * k + 3] = v.w;
            } else { f[4 * k] = f[4 * k + 1] = f[4 * k + 2] = f[4 * k + 3] = 0.f; }
        }
#pragma unroll
        for (int k = 0; k < 16; k++) sh[k] = {f[3 * k], f[3 * k + 1], f[3 * k + 2]};
    } else {
#pragma unroll
        for (int k = 0; k < 16; k++) {
            if (k < nb) {
                const float* p = sh_ptr(shs, dc, idx, max_coeffs, k);
                sh[k] = V3{__ldg(p), __ldg(p + 1), __ldg(p + 2)};
            } else sh[k] = V3{0.f, 0.f, 0.f};
        }
    }
    V3 result = kC0 * sh[0];
    if (deg > 0) {
        float x = dir.x, y = dir.y, z = dir.z;
        result = result - kC1 * y * sh[1] + kC1 * z * sh[2] - kC1 * x * sh[3];
        if (deg > 1) {
            float xx = x * x, yy = y * y, zz = z * z;
            float xy = x * y, yz = y * z, xz = x * z;
            result = result + kC2[0] * xy * sh[4] + kC2[1] * yz * sh[5] + kC2[2] * (2.0f * zz - xx - yy) * sh[6] +
                     kC2[3] * xz * sh[7] + kC2[4] * (xx - yy) * sh[8];
            if (deg > 2) {
                result = result + kC3[0] * y * (3.0f * xx - yy) * sh[9] + kC3[1] * xy * z * sh[10] +
                         kC3[2] * y * (4.0f * zz - xx - yy) * sh[11] +
                         kC3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[12] +
                         kC3[4] * x * (4.0f * zz - xx - yy) * sh[13] + kC3[5] * z * (xx - yy) * sh[14] +
                         kC3[6] * x * (xx - 3.0f * yy) * sh[15];
            }
        }
    }
    result = result + 0.5f;
    clamp_bits = (result.x < 0 ? 1 : 0) | (result.y < 0 ? 2 : 0) | (result.z < 0 ? 4 : 0);
    return {fmaxf(result.x, 0.0f), fmaxf(result.y, 0.0f), fmaxf(result.z, 0.0f)};
}

__device__ __forceinline__ void pre_prefetch_l2(const void* gmem) { asm volatile("prefetch.global.L2 [%0];" ::"l"(gmem)); }

struct PreArgs {
    int P, D, M, W, H;
    float tan_fovx, tan_fovy, focal_x, focal_y, scale_modifier;
    int prefiltered, antialiasing;
    const float *means3D, *shs, *sh_dc, *colors_precomp, *opacities, *scales, *rotations, *cov3D_precomp;
    const float *view, *proj, *campos;   // device pointers (16, 16, 3 floats)
    // Scratch tables the later stages need zero-filled (depth sort: histograms, tickets, look-back status; binning: the
    // same for the instance sort + scan words + difference array). The forward preprocess clears them on the side --
    // a fraction of a word per thread -- instead of two memset launches in front of latency-bound kernels.
    uint32_t* zero_a; long long zero_a_words;
    uint32_t* zero_b; long long zero_b_words;
};

#ifndef GSR_PFWD_CTAS
#define GSR_PFWD_CTAS 3           // (x 256 threads; CTAs are 128 threads) 85 registers: 0.130 ms against 0.131 at 4, 0.172 at 5
#endif
// GEOM_ONLY: the colour (SH evaluation, the 192-byte row per visible Gaussian) and the clamp bits are left to
// preprocess_color_kernel, which gsr_forward runs on a side stream UNDER the latency-bound depth sort / scan / instance
// sort: nothing before the blend needs the colours, and they are 60 % of this kernel's bytes.
template <bool SH16, bool GEOM_ONLY>
__global__ void __launch_bounds__(256, GSR_PFWD_CTAS) preprocess_fwd_kernel(const PreArgs a, int* __restrict__ radii,
                                                              float* __restrict__ depths, float4* __restrict__ rec,
                                                              float4* __restrict__ colr,
                                                              uint8_t* __restrict__ clamped,
                                                              uint32_t* __restrict__ tiles_touched,
                                                              uint32_t* __restrict__ dkey,
                                                              uint32_t* __restrict__ order,
                                                              uint32_t* __restrict__ rect_packed) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    {
        const long long nthreads = (long long)gridDim.x * blockDim.x;
        for (long long w = idx; w < a.zero_a_words; w += nthreads) a.zero_a[w] = 0u;
        for (long long w = idx; w < a.zero_b_words; w += nthreads) a.zero_b[w] = 0u;
    }
    if (idx >= a.P) return;
    int my_radii = 0;
    uint32_t my_tiles = 0, my_key = 0xFFFFFFFFu;   // culled Gaussians sort to the end and emit nothing
    uint32_t my_rect = 0;
    float my_depth = 0.0f;
    uint8_t my_clamp = 0;
    const dim3 grid((a.W + GSR_TILE - 1) / GSR_TILE, (a.H + GSR_TILE - 1) / GSR_TILE, 1);
    do {
        const float3 p_orig = {__ldg(a.means3D + 3 * idx), __ldg(a.means3D + 3 * idx + 1), __ldg(a.means3D + 3 * idx + 2)};
        // The covariance inputs are requested together with the position, before the cull is known: the arrays are dense,
        // so with half the Gaussians visible nearly every sector is fetched anyway, and the cull no longer sits between
        // two dependent memory round trips.
        V3 sc = {0.f, 0.f, 0.f};
        float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
        if (a.cov3D_precomp == nullptr) {
            sc = V3{__ldg(a.scales + 3 * idx), __ldg(a.scales + 3 * idx + 1), __ldg(a.scales + 3 * idx + 2)};
            q = __ldg(reinterpret_cast<const float4*>(a.rotations) + idx);
        }
        const float opacity_in = __ldg(a.opacities + idx);       // used last, requested first
        const float3 p_view = xform43(p_orig, a.view);
        if (p_view.z <= 0.2f) {                                  // auxiliary.h:154
            if (a.prefiltered) {
                printf("Point is filtered although prefiltered is set. This shouldn't happen!");
                __trap();
            }
            break;
        }
        const float4 p_hom = xform44(p_orig, a.proj);
        const float p_w = 1.0f / (p_hom.w + 0.0000001f);
        const float3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};
        // A centre that projects near the screen almost always survives the rectangle test below: pull its SH row
        // (192 bytes, two lines) into L2 now, while the covariance is computed. A hint only -- a large splat centred
        // far outside just reads its row from DRAM as before.
        // (margin 1.05: at C3 1.441M centres project onto the screen and 1.481M Gaussians are visible; a margin of 1.25
        // prefetched 1.712M rows -- 44 MB per view for Gaussians that are then culled -- 1.05 prefetches 1.503M)
        if (!GEOM_ONLY && SH16 && fabsf(p_proj.x) < 1.05f && fabsf(p_proj.y) < 1.05f) {
            pre_prefetch_l2(a.shs + (size_t)idx * 48);
            pre_prefetch_l2(a.shs + (size_t)idx * 48 + 32);
        }
        float c3[6];
        if (a.cov3D_precomp != nullptr) {
#pragma unroll
            for (int k = 0; k < 6; k++) c3[k] = __ldg(a.cov3D_precomp + 6 * (size_t)idx + k);
        } else {
            cov3d_of(sc, a.scale_modifier, q, c3);
        }
        float3 cov = cov2d_of(p_orig, a.focal_x, a.focal_y, a.tan_fovx, a.tan_fovy, c3, a.view);
        const float det_cov = cov.x * cov.z - cov.y * cov.y;       // only used by antialiasing
        cov.x += 0.3f;                                             // forward.cu:110-111
        cov.z += 0.3f;
        float det = (cov.x * cov.z - cov.y * cov.y);
        if (det == 0.0f) break;
        float det_inv = 1.f / det;
        float3 conic = {cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv};
        float mid = 0.5f * (cov.x + cov.z);
        float lambda1 = mid + sqrt(max(0.1f, mid * mid - det));
        float lambda2 = mid - sqrt(max(0.1f, mid * mid - det));
        float my_radius = ceil(3.f * sqrt(max(lambda1, lambda2)));
        float2 point_image = {ndc_to_pix(p_proj.x, a.W), ndc_to_pix(p_proj.y, a.H)};
        uint2 rect_min, rect_max;
        tile_rect(point_image, my_radius, rect_min, rect_max, grid);
        if ((rect_max.x - rect_min.x) * (rect_max.y - rect_min.y) == 0) break;

        V3 color = {0.f, 0.f, 0.f};
        uint8_t cb = 0;
        if (!GEOM_ONLY) {
            if (a.colors_precomp == nullptr) {      // SH input (fused, or dc + rest)
                color = sh_color<SH16>(idx, a.D, a.M, V3{p_orig.x, p_orig.y, p_orig.z},
                                       V3{__ldg(a.campos), __ldg(a.campos + 1), __ldg(a.campos + 2)}, a.shs, a.sh_dc, cb);
            } else {
                color = {__ldg(a.colors_precomp + 3 * idx), __ldg(a.colors_precomp + 3 * idx + 1),
                         __ldg(a.colors_precomp + 3 * idx + 2)};
            }
        }
        float opacity = opacity_in;
        if (a.antialiasing) opacity = opacity * sqrtf(fmaxf(0.000025f, det_cov / det));   // dr_aa (F1)

        // Conservative screen-space extent of the region where alpha >= 1/255 can hold, used by
        // the blend kernels to skip (sub-tile, Gaussian) pairs: |d| <= sqrt(2*tau*Sigma_dd) with
        // tau = ln(255*opacity) (+2% and absolute slack for fp32 rounding of `power`).
        float ex, ey;
        if (!(opacity >= 1.0f / 255.0f)) { ex = -1.0f; ey = -1.0f; }       // can never reach 1/255
        else if (det > 0.0f && cov.x > 0.0f && cov.z > 0.0f) {
            float tau = logf(opacity * 255.0f) * 1.02f + 0.02f;
            ex = sqrtf(2.0f * tau * cov.x) * 1.001f + 0.01f;
            ey = sqrtf(2.0f * tau * cov.z) * 1.001f + 0.01f;
            if (!(ex == ex) || !(ey == ey)) { ex = 1e30f; ey = 1e30f; }
        } else { ex = 1e30f; ey = 1e30f; }

        my_depth = p_view.z;
        my_radii = my_radius;
        rec[2 * (size_t)idx + 0] = make_float4(point_image.x, point_image.y, ex, ey);
        rec[2 * (size_t)idx + 1] = make_float4(conic.x, conic.y, conic.z, opacity);
        if (!GEOM_ONLY) colr[idx] = make_float4(color.x, color.y, color.z, 1.0f / p_view.z);
        my_clamp = cb;
        my_tiles = (rect_max.y - rect_min.y) * (rect_max.x - rect_min.x);
        // the rectangle itself, for the scan and the instance emission (binning): they then need neither the record nor
        // the radius of the Gaussian again. Only meaningful for grids up to 255 x 255 tiles (launch_binning checks).
        my_rect = rect_min.x | (rect_min.y << 8) | ((rect_max.x - rect_min.x) << 16) | ((rect_max.y - rect_min.y) << 24);
        my_key = __float_as_uint(p_view.z);                        // rasterizer_impl.cu:102-104 depth bits
    } while (0);
    // the small per-Gaussian arrays are written for EVERY Gaussian (zeros for culled ones): a warp then fills whole
    // sectors, where visible-only stores leave partially written sectors that cost a DRAM read-modify-write each
    radii[idx] = my_radii;
    if (tiles_touched != nullptr) tiles_touched[idx] = my_tiles;   // only kept when the rectangle cannot be packed
    dkey[idx] = my_key;
    order[idx] = (uint32_t)idx;
    depths[idx] = my_depth;
    if (!GEOM_ONLY) clamped[idx] = my_clamp;
    rect_packed[idx] = my_rect;
}

// The colour half of the forward preprocess (see GEOM_ONLY above): colour from SH (or the precomputed one) + inverse depth
// for every visible Gaussian, clamp bits for all. Visibility and depth come from the geometry pass (radii, depths), so the
// inverse depth has the same bits as in the one-kernel path.
template <bool SH16>
__global__ void __launch_bounds__(256, 4) preprocess_color_kernel(const PreArgs a, const int* __restrict__ radii,
                                                                  const float* __restrict__ depths, float4* __restrict__ colr,
                                                                  uint8_t* __restrict__ clamped) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= a.P) return;
    uint8_t cb = 0;
    if (__ldg(radii + idx) > 0) {
        const float depth = __ldg(depths + idx);
        V3 color;
        if (a.colors_precomp == nullptr) {
            const V3 p_orig = {__ldg(a.means3D + 3 * idx), __ldg(a.means3D + 3 * idx + 1), __ldg(a.means3D + 3 * idx + 2)};
            color = sh_color<SH16>(idx, a.D, a.M, p_orig, V3{__ldg(a.campos), __ldg(a.campos + 1), __ldg(a.campos + 2)}, a.shs,
                                   a.sh_dc, cb);
        } else {
            color = {__ldg(a.colors_precomp + 3 * idx), __ldg(a.colors_precomp + 3 * idx + 1),
                     __ldg(a.colors_precomp + 3 * idx + 2)};
        }
        colr[idx] = make_float4(color.x, color.y, color.z, 1.0f / depth);
    }
    clamped[idx] = cb;
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means, const float* __restrict__ view,
                                    uint8_t* __restrict__ present) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float3 p = {means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]};
    present[idx] = xform43(p, view).z > 0.2f;   // auxiliary.h:139-164, rasterizer_impl.cu:54-66
}

static void fill_pre_args(const gsr_view& v, PreArgs& a) {
    a.P = v.P; a.D = v.D; a.M = v.M; a.W = v.W; a.H = v.H;
    a.tan_fovx = v.tan_fovx; a.tan_fovy = v.tan_fovy;
    a.focal_y = v.H / (2.0f * v.tan_fovy);             // rasterizer_impl.cu:221-222
    a.focal_x = v.W / (2.0f * v.tan_fovx);
    a.scale_modifier = v.scale_modifier; a.prefiltered = v.prefiltered; a.antialiasing = v.antialiasing;
    a.means3D = v.means3D; a.shs = v.shs; a.sh_dc = v.sh_dc; a.colors_precomp = v.colors_precomp; a.opacities = v.opacities;
    a.scales = v.scales; a.rotations = v.rotations; a.cov3D_precomp = v.cov3D_precomp;
    a.view = v.viewmatrix; a.proj = v.projmatrix; a.campos = v.cam_pos;
    a.zero_a = nullptr; a.zero_a_words = 0; a.zero_b = nullptr; a.zero_b_words = 0;
}


// ---------------------------------------------------------------------------------------------
// Backward preprocess: backward.cu:144-274 (conic -> cov2D -> cov3D / mean), :346-396 (projection
// of dL/dmean2D, SH backward :20-139, cov3D -> scale / quaternion :278-341), fused into ONE kernel
// (the reference splits them only for code length, backward.cu:583-586). Reads the packed
// accumulator written by the backward blend; writes every gradient output in full (zeros for
// Gaussians with radii == 0) so no zero-fill pass is needed.
constexpr int kShRow = 52;     // floats per staged SH row: 16-byte reads of 8 consecutive rows fall into distinct banks
__device__ __forceinline__ void pre_cp_async16(void* smem, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void pre_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void pre_cp_async_wait() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

struct PreBwdArgs {
    PreArgs f;
    const int* radii;
    const uint8_t* clamped;
    float* acc;             // [P][GSR_ACC_FLOATS]
    int have_invdepth;
    int clear_acc;          // zero every accumulator row after reading it (gsr_grads::accum_clean)
    gsr_grads out;
};

#ifndef GSR_PBWD_CTAS
#define GSR_PBWD_CTAS 3           // 3 x 256 threads at 85 registers (no spills): 0.298 ms against 0.304 at 4 x 64 registers
#endif
#ifndef GSR_PBWD_KU
#define GSR_PBWD_KU 4
#endif
template <bool SH16>
__global__ void __launch_bounds__(256, GSR_PBWD_CTAS) preprocess_bwd_kernel(const PreBwdArgs b) {
    const PreArgs& a = b.f;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const bool inside = idx < a.P;
    if (!SH16 && !inside) return;                     // the SH16 path has CTA-wide barriers: everybody stays
    const gsr_grads& o = b.out;
    // position, scale and rotation are requested together with the radius that decides visibility (dense arrays: with
    // half the Gaussians visible nearly every sector is needed anyway), so they do not wait for that decision
    float3 mean = {0.f, 0.f, 0.f};
    float sc[3] = {0.f, 0.f, 0.f};
    float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
    int my_radius = 0;
    uint8_t cbits_in = 0;
    if (inside) {
        my_radius = b.radii[idx];
        cbits_in = b.clamped[idx];
        mean = make_float3(__ldg(a.means3D + 3 * idx), __ldg(a.means3D + 3 * idx + 1), __ldg(a.means3D + 3 * idx + 2));
        if (a.cov3D_precomp == nullptr) {
            sc[0] = __ldg(a.scales + 3 * idx); sc[1] = __ldg(a.scales + 3 * idx + 1); sc[2] = __ldg(a.scales + 3 * idx + 2);
            q = __ldg(reinterpret_cast<const float4*>(a.rotations) + idx);
        }
    }
    const bool vis = my_radius > 0;
    const int M = a.M;
    // SH16 path: the coefficient rows of the CTA's visible Gaussians are fetched COOPERATIVELY -- consecutive threads take
    // consecutive 16-byte pieces, so one warp-wide request covers 4 lines instead of the 32 a row-per-thread access
    // touches (the L1 tag stage, one line per cycle, was this kernel's busiest unit) -- with cp.async into shared
    // memory, where they wait while the covariance backward runs. The dL/dsh rows leave the same way.
    extern __shared__ __align__(16) float s_rows[];   // [256][kShRow]
    __shared__ uint8_t s_owner[256];                  // slot -> thread of the visible Gaussian
    __shared__ int s_warp[8];
    const int n_quads = ((a.D + 1) * (a.D + 1) * 3 + 3) / 4;      // 16-byte pieces of a row that hold active coefficients
    int slot = 0, n_vis = 0;
    if (SH16) {
        const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
        const unsigned m = __ballot_sync(0xffffffffu, vis);
        if (lane == 0) s_warp[w] = __popc(m);
        __syncthreads();
#pragma unroll
        for (int q = 0; q < 8; q++) {
            const int c = s_warp[q];
            slot += q < w ? c : 0;
            n_vis += c;
        }
        slot += __popc(m & ((1u << lane) - 1u));
        if (vis) s_owner[slot] = (uint8_t)threadIdx.x;
        __syncthreads();
        const float* src0 = a.shs + (size_t)blockIdx.x * 256 * 48;
        const bool rmw = o.accumulate != 0;
        const float* grad0 = o.dL_dsh + (size_t)blockIdx.x * 256 * 48;
        for (int f = threadIdx.x; f < n_vis * n_quads; f += 256) {
            const int sl = f / n_quads, k = f - sl * n_quads;
            const size_t at = (size_t)s_owner[sl] * 48 + 4 * k;
            pre_cp_async16(s_rows + sl * kShRow + 4 * k, src0 + at);
            // accumulate mode reads the gradient rows back at the very end: pull them into L2 now (one request per
            // 128 bytes of the row), so that those reads do not pay a DRAM round trip each
            if (rmw && (k & 7) == 0) pre_prefetch_l2(grad0 + at);
        }
        pre_cp_async_commit();
    }
    if (vis && o.accumulate != 0) {                    // same for the small per-Gaussian gradients
        pre_prefetch_l2(o.dL_dopacity + idx);
        pre_prefetch_l2(o.dL_dmeans3D + 3 * (size_t)idx);
        if (o.dL_dscales) pre_prefetch_l2(o.dL_dscales + 3 * (size_t)idx);
        if (o.dL_drotations) pre_prefetch_l2(o.dL_drotations + 4 * (size_t)idx);
    }
    float4* const my_row = reinterpret_cast<float4*>(s_rows + slot * kShRow);

    float dmean[3] = {0.f, 0.f, 0.f}, dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float dmean2d[2] = {0.f, 0.f}, dopac = 0.f, dcolor[3] = {0.f, 0.f, 0.f}, dconic[3] = {0.f, 0.f, 0.f};
    float dscale[3] = {0.f, 0.f, 0.f}, drot[4] = {0.f, 0.f, 0.f, 0.f};
    float sh_coef[16];   // basis value per coefficient (dL/dsh[k] = basis_k * dL/dRGB)
    float dRGB[3] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int k = 0; k < 16; k++) sh_coef[k] = 0.f;

    if (vis) {
        const float4* accp = reinterpret_cast<const float4*>(b.acc + (size_t)idx * GSR_ACC_FLOATS);
        const float4 a0 = accp[0], a1 = accp[1], a2 = accp[2];
        dmean2d[0] = a0.x; dmean2d[1] = a0.y;
        dconic[0] = a0.z; dconic[1] = a0.w; dconic[2] = a1.x;
        dopac = a1.y;
        dcolor[0] = a1.z; dcolor[1] = a1.w; dcolor[2] = a2.x;
        const float dinvd = b.have_invdepth ? a2.y : 0.f;

        const float* view = a.view;
        const float* proj = a.proj;
        float c3[6];
        if (a.cov3D_precomp != nullptr) {
#pragma unroll
            for (int k = 0; k < 6; k++) c3[k] = __ldg(a.cov3D_precomp + 6 * (size_t)idx + k);
        } else {
            cov3d_of(V3{sc[0], sc[1], sc[2]}, a.scale_modifier, q, c3);
        }
        // ---- recompute the projection (backward.cu:158-200)
        float3 t = xform43(mean, view);
        const float limx = 1.3f * a.tan_fovx, limy = 1.3f * a.tan_fovy;
        const float txtz = t.x / t.z, tytz = t.y / t.z;
        t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
        t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
        const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
        const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
        const float hx = a.focal_x, hy = a.focal_y;
        const float J00 = hx / t.z, J02 = -(hx * t.x) / (t.z * t.z);
        const float J11 = hy / t.z, J12 = -(hy * t.y) / (t.z * t.z);
        float T0[3], T1[3];   // rows of J * Rv ; Rv[i][j] = view[4*j+i]
#pragma unroll
        for (int j = 0; j < 3; j++) {
            T0[j] = J00 * view[4 * j + 0] + J02 * view[4 * j + 2];
            T1[j] = J11 * view[4 * j + 1] + J12 * view[4 * j + 2];
        }
        const float V[3][3] = {{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}};
        float TV0[3], TV1[3];
#pragma unroll
        for (int j = 0; j < 3; j++) {
            TV0[j] = T0[0] * V[j][0] + T0[1] * V[j][1] + T0[2] * V[j][2];
            TV1[j] = T1[0] * V[j][0] + T1[1] * V[j][1] + T1[2] * V[j][2];
        }
        const float cxx = TV0[0] * T0[0] + TV0[1] * T0[1] + TV0[2] * T0[2];
        const float cxy = TV0[0] * T1[0] + TV0[1] * T1[1] + TV0[2] * T1[2];
        const float cyy = TV1[0] * T1[0] + TV1[1] * T1[1] + TV1[2] * T1[2];
        // ---- antialiasing term (upstream dr_aa; F1)
        float aa_dxx = 0.f, aa_dxy = 0.f, aa_dyy = 0.f;
        if (a.antialiasing) {
            const float w = 0.3f;
            const float N = cxx * cyy - cxy * cxy, Dn = (cxx + w) * (cyy + w) - cxy * cxy;
            const float ratio = N / Dn;
            const float h = sqrtf(fmaxf(0.000025f, ratio));
            const float g_stored = dopac;
            dopac = g_stored * h;
            const float dL_dh = g_stored * __ldg(a.opacities + idx);
            const float dL_dratio = ratio <= 0.000025f ? 0.f : dL_dh / (2.f * h);
            const float f = dL_dratio / (Dn * Dn);
            aa_dxx = w * (w * cyy + cyy * cyy + cxy * cxy) * f;
            aa_dyy = w * (w * cxx + cxx * cxx + cxy * cxy) * f;
            aa_dxy = -2.f * w * cxy * (w + cxx + cyy) * f;
        }
        const float ca = cxx + 0.3f, cb = cxy, cc = cyy + 0.3f;
        const float denom = ca * cc - cb * cb;
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        if (denom2inv != 0) {
            dL_da = denom2inv * (-cc * cc * dconic[0] + 2 * cb * cc * dconic[1] + (denom - ca * cc) * dconic[2]);
            dL_dc = denom2inv * (-ca * ca * dconic[2] + 2 * ca * cb * dconic[1] + (denom - ca * cc) * dconic[0]);
            dL_db = denom2inv * 2 * (cb * cc * dconic[0] - (denom + 2 * cb * cb) * dconic[1] + ca * cb * dconic[2]);
        }
        dL_da += aa_dxx; dL_dc += aa_dyy; dL_db += aa_dxy;
        dcov[0] = (T0[0] * T0[0] * dL_da + T0[0] * T1[0] * dL_db + T1[0] * T1[0] * dL_dc);
        dcov[3] = (T0[1] * T0[1] * dL_da + T0[1] * T1[1] * dL_db + T1[1] * T1[1] * dL_dc);
        dcov[5] = (T0[2] * T0[2] * dL_da + T0[2] * T1[2] * dL_db + T1[2] * T1[2] * dL_dc);
        dcov[1] = 2 * T0[0] * T0[1] * dL_da + (T0[0] * T1[1] + T0[1] * T1[0]) * dL_db + 2 * T1[0] * T1[1] * dL_dc;
        dcov[2] = 2 * T0[0] * T0[2] * dL_da + (T0[0] * T1[2] + T0[2] * T1[0]) * dL_db + 2 * T1[0] * T1[2] * dL_dc;
        dcov[4] = 2 * T0[2] * T0[1] * dL_da + (T0[1] * T1[2] + T0[2] * T1[1]) * dL_db + 2 * T1[1] * T1[2] * dL_dc;
        float dT0[3], dT1[3];
#pragma unroll
        for (int j = 0; j < 3; j++) {
            dT0[j] = 2 * TV0[j] * dL_da + TV1[j] * dL_db;
            dT1[j] = 2 * TV1[j] * dL_dc + TV0[j] * dL_db;
        }
        const float dJ00 = view[0] * dT0[0] + view[4] * dT0[1] + view[8] * dT0[2];
        const float dJ02 = view[2] * dT0[0] + view[6] * dT0[1] + view[10] * dT0[2];
        const float dJ11 = view[1] * dT1[0] + view[5] * dT1[1] + view[9] * dT1[2];
        const float dJ12 = view[2] * dT1[0] + view[6] * dT1[1] + view[10] * dT1[2];
        const float tz = 1.f / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
        const float dtx = x_grad_mul * -hx * tz2 * dJ02;
        const float dty = y_grad_mul * -hy * tz2 * dJ12;
        float dtz = -hx * tz2 * dJ00 - hy * tz2 * dJ11 + (2 * hx * t.x) * tz3 * dJ02 + (2 * hy * t.y) * tz3 * dJ12;
        dtz += -dinvd * tz2;                                   // A-DEPTH: d(1/t.z)/dt.z
        dmean[0] = view[0] * dtx + view[1] * dty + view[2] * dtz;
        dmean[1] = view[4] * dtx + view[5] * dty + view[6] * dtz;
        dmean[2] = view[8] * dtx + view[9] * dty + view[10] * dtz;
        // ---- projection of the 2D mean gradient (backward.cu:366-382)
        const float4 mh = xform44(mean, proj);
        const float m_w = 1.0f / (mh.w + 0.0000001f);
        const float mul1 = mh.x * m_w * m_w, mul2 = mh.y * m_w * m_w;
        dmean[0] += (proj[0] * m_w - proj[3] * mul1) * dmean2d[0] + (proj[1] * m_w - proj[3] * mul2) * dmean2d[1];
        dmean[1] += (proj[4] * m_w - proj[7] * mul1) * dmean2d[0] + (proj[5] * m_w - proj[7] * mul2) * dmean2d[1];
        dmean[2] += (proj[8] * m_w - proj[11] * mul1) * dmean2d[0] + (proj[9] * m_w - proj[11] * mul2) * dmean2d[1];
        // ---- cov3D -> scale / rotation (backward.cu:278-341)
        if (a.scales != nullptr) {
            const float r = q.x, x = q.y, y = q.z, z = q.w;
            const float Rm[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                                    {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                                    {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
            const float s[3] = {a.scale_modifier * sc[0], a.scale_modifier * sc[1], a.scale_modifier * sc[2]};
            const float dS[3][3] = {{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]},
                                    {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                                    {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}};
            float Mt[3][3];   // reference's dL_dMt[a][b] (glm indexing) after scaling by s_a
#pragma unroll
            for (int i = 0; i < 3; i++) {
                float dMi[3];
#pragma unroll
                for (int j = 0; j < 3; j++)
                    dMi[j] = 2.0f * ((s[i] * Rm[0][i]) * dS[0][j] + (s[i] * Rm[1][i]) * dS[1][j] + (s[i] * Rm[2][i]) * dS[2][j]);
                dscale[i] = Rm[0][i] * dMi[0] + Rm[1][i] * dMi[1] + Rm[2][i] * dMi[2];
#pragma unroll
                for (int j = 0; j < 3; j++) Mt[i][j] = s[i] * dMi[j];
            }
            drot[0] = 2 * z * (Mt[0][1] - Mt[1][0]) + 2 * y * (Mt[2][0] - Mt[0][2]) + 2 * x * (Mt[1][2] - Mt[2][1]);
            drot[1] = 2 * y * (Mt[1][0] + Mt[0][1]) + 2 * z * (Mt[2][0] + Mt[0][2]) + 2 * r * (Mt[1][2] - Mt[2][1]) - 4 * x * (Mt[2][2] + Mt[1][1]);
            drot[2] = 2 * x * (Mt[1][0] + Mt[0][1]) + 2 * r * (Mt[2][0] - Mt[0][2]) + 2 * z * (Mt[1][2] + Mt[2][1]) - 4 * y * (Mt[2][2] + Mt[0][0]);
            drot[3] = 2 * r * (Mt[0][1] - Mt[1][0]) + 2 * x * (Mt[2][0] + Mt[0][2]) + 2 * y * (Mt[1][2] + Mt[2][1]) - 4 * z * (Mt[1][1] + Mt[0][0]);
        }
    }

    // ---- SH backward (backward.cu:20-139)
    if (SH16) {
        pre_cp_async_wait();
        __syncthreads();
    }
    if (vis) {
        if (a.shs != nullptr || a.sh_dc != nullptr) {
            const uint8_t cbits = cbits_in;
            dRGB[0] = (cbits & 1) ? 0.f : dcolor[0];
            dRGB[1] = (cbits & 2) ? 0.f : dcolor[1];
            dRGB[2] = (cbits & 4) ? 0.f : dcolor[2];
            const float d0x = mean.x - __ldg(a.campos), d0y = mean.y - __ldg(a.campos + 1), d0z = mean.z - __ldg(a.campos + 2);
            const float s2 = d0x * d0x + d0y * d0y + d0z * d0z;
            const float len = sqrtf(s2);
            const float x = d0x / len, y = d0y / len, z = d0z / len;
            const int deg = a.D;
            // dot(sh[k], dRGB) for the direction derivative
            float sd[16];
            const int nb = (deg + 1) * (deg + 1);
            if (SH16) {
                const float4* p4 = my_row;
                float f[48];
#pragma unroll
                for (int k = 0; k < 12; k++) {
                    if (k * 4 < nb * 3) {
                        float4 v = p4[k];
                        f[4 * k] = v.x; f[4 * k + 1] = v.y; f[4 * k + 2] = v.z; f[4 * k + 3] = v.w;
                    } else { f[4 * k] = f[4 * k + 1] = f[4 * k + 2] = f[4 * k + 3] = 0.f; }
                }
#pragma unroll
                for (int k = 0; k < 16; k++) sd[k] = f[3 * k] * dRGB[0] + f[3 * k + 1] * dRGB[1] + f[3 * k + 2] * dRGB[2];
            } else {
#pragma unroll
                for (int k = 0; k < 16; k++) {
                    if (k < nb) {
                        const float* p = sh_ptr(a.shs, a.sh_dc, idx, M, k);
                        sd[k] = __ldg(p) * dRGB[0] + __ldg(p + 1) * dRGB[1] + __ldg(p + 2) * dRGB[2];
                    } else sd[k] = 0.f;
                }
            }
            float ddx = 0.f, ddy = 0.f, ddz = 0.f;   // dL/ddir
            sh_coef[0] = kC0;
            if (deg > 0) {
                sh_coef[1] = -kC1 * y; sh_coef[2] = kC1 * z; sh_coef[3] = -kC1 * x;
                ddx = -kC1 * sd[3]; ddy = -kC1 * sd[1]; ddz = kC1 * sd[2];
                if (deg > 1) {
                    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                    sh_coef[4] = kC2[0] * xy; sh_coef[5] = kC2[1] * yz; sh_coef[6] = kC2[2] * (2.f * zz - xx - yy);
                    sh_coef[7] = kC2[3] * xz; sh_coef[8] = kC2[4] * (xx - yy);
                    ddx += kC2[0] * y * sd[4] + kC2[2] * 2.f * -x * sd[6] + kC2[3] * z * sd[7] + kC2[4] * 2.f * x * sd[8];
                    ddy += kC2[0] * x * sd[4] + kC2[1] * z * sd[5] + kC2[2] * 2.f * -y * sd[6] + kC2[4] * 2.f * -y * sd[8];
                    ddz += kC2[1] * y * sd[5] + kC2[2] * 2.f * 2.f * z * sd[6] + kC2[3] * x * sd[7];
                    if (deg > 2) {
                        sh_coef[9] = kC3[0] * y * (3.f * xx - yy);
                        sh_coef[10] = kC3[1] * xy * z;
                        sh_coef[11] = kC3[2] * y * (4.f * zz - xx - yy);
                        sh_coef[12] = kC3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy);
                        sh_coef[13] = kC3[4] * x * (4.f * zz - xx - yy);
                        sh_coef[14] = kC3[5] * z * (xx - yy);
                        sh_coef[15] = kC3[6] * x * (xx - 3.f * yy);
                        ddx += kC3[0] * sd[9] * 3.f * 2.f * xy + kC3[1] * sd[10] * yz + kC3[2] * sd[11] * -2.f * xy +
                               kC3[3] * sd[12] * -3.f * 2.f * xz + kC3[4] * sd[13] * (-3.f * xx + 4.f * zz - yy) +
                               kC3[5] * sd[14] * 2.f * xz + kC3[6] * sd[15] * 3.f * (xx - yy);
                        ddy += kC3[0] * sd[9] * 3.f * (xx - yy) + kC3[1] * sd[10] * xz +
                               kC3[2] * sd[11] * (-3.f * yy + 4.f * zz - xx) + kC3[3] * sd[12] * -3.f * 2.f * yz +
                               kC3[4] * sd[13] * -2.f * xy + kC3[5] * sd[14] * -2.f * yz + kC3[6] * sd[15] * -3.f * 2.f * xy;
                        ddz += kC3[1] * sd[10] * xy + kC3[2] * sd[11] * 4.f * 2.f * yz +
                               kC3[3] * sd[12] * 3.f * (2.f * zz - xx - yy) + kC3[4] * sd[13] * 4.f * 2.f * xz +
                               kC3[5] * sd[14] * (xx - yy);
                    }
                }
            }
            // normalisation backward (auxiliary.h:107-117)
            const float inv = 1.0f / sqrtf(s2 * s2 * s2);
            dmean[0] += ((+s2 - d0x * d0x) * ddx - d0y * d0x * ddy - d0z * d0x * ddz) * inv;
            dmean[1] += (-d0x * d0y * ddx + (s2 - d0y * d0y) * ddy - d0z * d0y * ddz) * inv;
            dmean[2] += (-d0x * d0z * ddx - d0y * d0z * ddy + (s2 - d0z * d0z) * ddz) * inv;
        }
    }

    // ---- write everything (zeros for invisible Gaussians); in accumulate mode add to what is there and leave the
    // rows of invisible Gaussians untouched (multi-view batches: no zero fill, no separate accumulation pass)
    const bool accum = o.accumulate != 0;
    if (inside) {
        o.dL_dmeans2D[3 * (size_t)idx + 0] = dmean2d[0]; // per-view sink of the caller (train.py:159): never accumulated
        o.dL_dmeans2D[3 * (size_t)idx + 1] = dmean2d[1];
        o.dL_dmeans2D[3 * (size_t)idx + 2] = 0.f;
        if (o.dL_dconic)
            reinterpret_cast<float4*>(o.dL_dconic)[idx] = make_float4(dconic[0], dconic[1], 0.f, dconic[2]);
    }
    if (inside && !(accum && !vis)) {
        // accumulate mode: every old value is loaded BEFORE the first store. Written one after the other (load, add,
        // store, next) the stores keep the later loads behind them -- the pointers may alias as far as the compiler
        // knows -- and a thread pays one memory round trip per output array.
        float old_op = 0.f, old_m[3] = {0.f, 0.f, 0.f}, old_c[3] = {0.f, 0.f, 0.f}, old_v[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        float old_s[3] = {0.f, 0.f, 0.f};
        float4 old_r = make_float4(0.f, 0.f, 0.f, 0.f);
        if (accum) {
            old_op = o.dL_dopacity[idx];
#pragma unroll
            for (int k = 0; k < 3; k++) old_m[k] = o.dL_dmeans3D[3 * (size_t)idx + k];
            if (o.dL_dcolors) {
#pragma unroll
                for (int k = 0; k < 3; k++) old_c[k] = o.dL_dcolors[3 * (size_t)idx + k];
            }
            if (o.dL_dcov3D) {
#pragma unroll
                for (int k = 0; k < 6; k++) old_v[k] = o.dL_dcov3D[6 * (size_t)idx + k];
            }
            if (o.dL_dscales) {
#pragma unroll
                for (int k = 0; k < 3; k++) old_s[k] = o.dL_dscales[3 * (size_t)idx + k];
            }
            if (o.dL_drotations) old_r = reinterpret_cast<const float4*>(o.dL_drotations)[idx];
        }
        auto put = [accum](float* p, float v) { *p = accum ? *p + v : v; };      // (SH fallback path below)
        o.dL_dopacity[idx] = old_op + dopac;
#pragma unroll
        for (int k = 0; k < 3; k++) o.dL_dmeans3D[3 * (size_t)idx + k] = old_m[k] + dmean[k];
        if (o.dL_dcolors) {
#pragma unroll
            for (int k = 0; k < 3; k++) o.dL_dcolors[3 * (size_t)idx + k] = old_c[k] + dcolor[k];
        }
        if (o.dL_dcov3D) {
#pragma unroll
            for (int k = 0; k < 6; k++) o.dL_dcov3D[6 * (size_t)idx + k] = old_v[k] + dcov[k];
        }
        if (o.dL_dscales) {
#pragma unroll
            for (int k = 0; k < 3; k++) o.dL_dscales[3 * (size_t)idx + k] = old_s[k] + dscale[k];
        }
        if (o.dL_drotations)
            reinterpret_cast<float4*>(o.dL_drotations)[idx] =
                make_float4(old_r.x + drot[0], old_r.y + drot[1], old_r.z + drot[2], old_r.w + drot[3]);
        if (o.dL_dsh || o.dL_ddc) {
            if (SH16) {
                if (vis) {
                    // the row's gradient replaces its coefficients in shared memory; the CTA writes all rows out together
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const float c0 = sh_coef[4 * k], c1 = sh_coef[4 * k + 1], c2 = sh_coef[4 * k + 2], c3 = sh_coef[4 * k + 3];
                        my_row[3 * k] = make_float4(c0 * dRGB[0], c0 * dRGB[1], c0 * dRGB[2], c1 * dRGB[0]);
                        my_row[3 * k + 1] = make_float4(c1 * dRGB[1], c1 * dRGB[2], c2 * dRGB[0], c2 * dRGB[1]);
                        my_row[3 * k + 2] = make_float4(c2 * dRGB[2], c3 * dRGB[0], c3 * dRGB[1], c3 * dRGB[2]);
                    }
                } else {                                   // culled, not accumulating: a zero row
                    float4* dst = reinterpret_cast<float4*>(o.dL_dsh + (size_t)idx * 48);
#pragma unroll
                    for (int k = 0; k < 12; k++) dst[k] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
            } else {
                // accumulate mode: coefficients above the active degree have zero gradient and need no traffic
                const int kmax = accum ? min(M, (a.D + 1) * (a.D + 1)) : M;
                for (int k = 0; k < kmax; k++) {
                    const float c = k < 16 ? sh_coef[k < 16 ? k : 0] : 0.f;
                    float* dst = sh_ptr(o.dL_dsh, o.dL_ddc, idx, M, k);
                    put(dst, c * dRGB[0]); put(dst + 1, c * dRGB[1]); put(dst + 2, c * dRGB[2]);
                }
            }
        }
    }
    if (vis && b.clear_acc) {
        // hand the accumulator back all zero (gsr_grads::accum_clean): no 48 B / Gaussian clear per view. Down here, behind
        // every load of the thread: a store up where the row is read keeps all later loads behind it (the pointers may
        // alias as far as the compiler knows) and cost 0.47 ms per view when it was tried there.
        float4* accw = reinterpret_cast<float4*>(b.acc + (size_t)idx * GSR_ACC_FLOATS);
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        accw[0] = z; accw[1] = z; accw[2] = z;
    }
    if (SH16) {
        __syncthreads();
        // coefficients above the active degree have zero gradient: in accumulate mode they need no traffic at all
        const int nq = accum ? n_quads : 12;
        float* dst0 = o.dL_dsh + (size_t)blockIdx.x * 256 * 48;
        const int total = n_vis * nq;
        constexpr int kU = GSR_PBWD_KU;                // old rows: kU 16-byte loads in flight per thread before any store
        for (int f0 = threadIdx.x; f0 < total; f0 += 256 * kU) {
            float4* d[kU];
            float4 c[kU], v[kU];
#pragma unroll
            for (int u = 0; u < kU; u++) {
                const int f = f0 + 256 * u;
                c[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (f < total) {
                    const int sl = f / nq, k = f - sl * nq;
                    d[u] = reinterpret_cast<float4*>(dst0 + (size_t)s_owner[sl] * 48 + 4 * k);
                    v[u] = *reinterpret_cast<const float4*>(s_rows + sl * kShRow + 4 * k);
                    if (accum) c[u] = *d[u];
                }
            }
#pragma unroll
            for (int u = 0; u < kU; u++) {
                if (f0 + 256 * u < total)
                    *d[u] = make_float4(v[u].x + c[u].x, v[u].y + c[u].y, v[u].z + c[u].z, v[u].w + c[u].w);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
static bool sh16_fast(const gsr_view& v) {
    return v.sh_dc == nullptr && v.shs != nullptr && v.M == 16 && (reinterpret_cast<uintptr_t>(v.shs) & 15) == 0;
}

int launch_preprocess(const gsr_view& v, GeomState& g, int* radii, cudaStream_t s, void* zero_a, size_t zero_a_bytes,
                      void* zero_b, size_t zero_b_bytes, bool geometry_only) {
    PreArgs a;
    fill_pre_args(v, a);
    a.zero_a = reinterpret_cast<uint32_t*>(zero_a); a.zero_a_words = (long long)(zero_a_bytes / 4);
    a.zero_b = reinterpret_cast<uint32_t*>(zero_b); a.zero_b_words = (long long)(zero_b_bytes / 4);
    constexpr int kThreads = 128;
    const int blocks = (v.P + kThreads - 1) / kThreads;
    uint32_t* tiles_touched = rect_packing_ok(v.W, v.H) ? nullptr : g.tiles_touched;
    if (geometry_only)
        GSR_LAUNCH("preprocess_geom", v.debug, s,
                   (preprocess_fwd_kernel<false, true><<<blocks, kThreads, 0, s>>>(a, radii, g.depths, g.rec, g.colr, g.clamped, tiles_touched, g.dkey, g.order, g.rect_packed)));
    else if (sh16_fast(v))
        GSR_LAUNCH("preprocess_fwd", v.debug, s,
                   (preprocess_fwd_kernel<true, false><<<blocks, kThreads, 0, s>>>(a, radii, g.depths, g.rec, g.colr, g.clamped, tiles_touched, g.dkey, g.order, g.rect_packed)));
    else
        GSR_LAUNCH("preprocess_fwd", v.debug, s,
                   (preprocess_fwd_kernel<false, false><<<blocks, kThreads, 0, s>>>(a, radii, g.depths, g.rec, g.colr, g.clamped, tiles_touched, g.dkey, g.order, g.rect_packed)));
    return 0;
}

// after launch_preprocess(..., geometry_only = true) on the same data: colours, inverse depths, clamp bits
int launch_preprocess_color(const gsr_view& v, GeomState& g, const int* radii, cudaStream_t s) {
    PreArgs a;
    fill_pre_args(v, a);
    a.zero_a = nullptr; a.zero_a_words = 0; a.zero_b = nullptr; a.zero_b_words = 0;
    constexpr int kThreads = 128;
    const int blocks = (v.P + kThreads - 1) / kThreads;
    if (sh16_fast(v))
        GSR_LAUNCH("preprocess_color", v.debug, s,
                   (preprocess_color_kernel<true><<<blocks, kThreads, 0, s>>>(a, radii, g.depths, g.colr, g.clamped)));
    else
        GSR_LAUNCH("preprocess_color", v.debug, s,
                   (preprocess_color_kernel<false><<<blocks, kThreads, 0, s>>>(a, radii, g.depths, g.colr, g.clamped)));
    return 0;
}

int launch_preprocess_backward(const gsr_view& v, const GeomState& g, const int* radii, float* acc,
                               bool have_invdepth, const gsr_grads& out, cudaStream_t s) {
    PreBwdArgs b;
    fill_pre_args(v, b.f);
    b.radii = radii; b.clamped = g.clamped; b.acc = acc; b.have_invdepth = have_invdepth ? 1 : 0; b.out = out;
    b.clear_acc = out.accum_clean ? 1 : 0;
    const int blocks = (v.P + 255) / 256;
    const bool fast = sh16_fast(v) && out.dL_dsh != nullptr && (reinterpret_cast<uintptr_t>(out.dL_dsh) & 15) == 0;
    if (fast) {
        const int smem = 256 * kShRow * (int)sizeof(float);
        static DeviceOnce once;                       // the opt-in is per device
        { const int e = ensure_dynamic_smem(preprocess_bwd_kernel<true>, smem, once); if (e) return e; }
        GSR_LAUNCH("preprocess_bwd", v.debug, s, (preprocess_bwd_kernel<true><<<blocks, 256, smem, s>>>(b)));
    }
    else
        GSR_LAUNCH("preprocess_bwd", v.debug, s, (preprocess_bwd_kernel<false><<<blocks, 256, 0, s>>>(b)));
    return 0;
}

int launch_mark_visible(int P, const float* means, const float* view, uint8_t* present, cudaStream_t s) {
    GSR_LAUNCH("mark_visible", false, s, (mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means, view, present)));
    return 0;
}

}  // namespace gsr
