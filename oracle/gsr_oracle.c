/*
 * gsr_oracle.c -- CPU restatement of the reference rasterizer hot path.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing in the product path (gaussian-splatting-slam_b200/) may
 * import, link or execute this file. It is used by tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py, as the checker.
 *
 * Pinning: the reference ships no golden vectors for the rasterizer.
 * This restatement is pinned (a) on the GPU box against the UNMODIFIED reference CUDA build
 * oracle/_ref/ref_dgr_C.so on identical inputs (tests/test_gpu_parity.py), and (b) here on the
 * CPU against an independent torch-autograd derivation (oracle/torch_oracle.py) and against the
 * reference's own Python sub-step restatements (utils/sh_utils.py, utils/general_utils.py), whose
 * outputs are committed under tests/golden/. Antialiasing and inverse depth are not in the
 * vendored reference: for those two switches parity is UNPINNED by reference
 * code and rests on the autograd oracle + finite differences only.
 *
 * Each function cites the reference file:line it follows (paths relative to
 * /root/reference/submodules/diff-gaussian-rasterization/cuda_rasterizer/ unless stated).
 * Arithmetic is plain fp32, one rounding per operation (-ffp-contract=off); the CUDA reference
 * contracts some a*b+c into FMAs, so floats agree to a few ulp and integer results (radii, tile
 * counts) can differ for the rare Gaussian that sits exactly on a ceil()/tile boundary.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define TILE 16 /* config.h:16-17 BLOCK_X, BLOCK_Y */

/* auxiliary.h:22-39 */
static const float C0 = 0.28209479177387814f;
static const float C1 = 0.4886025119029199f;
static const float C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                            -1.0925484305920792f, 0.5462742152960396f};
static const float C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                            0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                            -0.5900435899266435f};

typedef struct {
    int P, D, M, W, H;
    float tan_fovx, tan_fovy, scale_modifier;
    int antialiasing;
    const float *bg, *means3D, *shs, *colors_precomp, *opacities, *scales, *rotations,
        *cov3D_precomp, *view, *proj, *campos;
} gso_scene;

int gso_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void gso_set_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* auxiliary.h:58-77 : matrices are read column-major (= transposed torch matrices). */
/* The reference is built with nvcc's default -fmad=true; for these four-term sums its SASS is
   mul(second term), fma(first), fma(third), add(fourth). Written out with fmaf() so that view-space
   depth (the low half of the sort keys) and the projected means carry the reference's exact bits. */
static inline float xf_row(float a, float b, float c, float d, const float *p) {
    return fmaf(c, p[2], fmaf(a, p[0], b * p[1])) + d;
}
static inline void xf43(const float *p, const float *m, float *o) {
    o[0] = xf_row(m[0], m[4], m[8], m[12], p);
    o[1] = xf_row(m[1], m[5], m[9], m[13], p);
    o[2] = xf_row(m[2], m[6], m[10], m[14], p);
}
static inline void xf44(const float *p, const float *m, float *o) {
    xf43(p, m, o);
    o[3] = xf_row(m[3], m[7], m[11], m[15], p);
}

/* auxiliary.h:41-44 : evaluated in double because of the 1.0 / 0.5 literals. */
static inline float ndc2pix(float v, int S) { return (float)((((double)v + 1.0) * S - 1.0) * 0.5); }

/* auxiliary.h:46-56 */
static inline void get_rect(float px, float py, int r, int gx, int gy, int *mn, int *mx) {
    int a;
    a = (int)((px - r) / TILE); a = a < 0 ? 0 : a; mn[0] = a > gx ? gx : a;
    a = (int)((py - r) / TILE); a = a < 0 ? 0 : a; mn[1] = a > gy ? gy : a;
    a = (int)((px + r + TILE - 1) / TILE); a = a < 0 ? 0 : a; mx[0] = a > gx ? gx : a;
    a = (int)((py + r + TILE - 1) / TILE); a = a < 0 ? 0 : a; mx[1] = a > gy ? gy : a;
}

/* forward.cu:118-152. M = S*R in glm terms; Sigma = M^T M. Stored upper triangle. */
static void cov3d_from_scale_rot(const float *s, float mod, const float *q, float *cov) {
    float r = q[0], x = q[1], y = q[2], z = q[3];
    /* Rm[i][j]: row i, column j of the usual rotation matrix */
    float Rm[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                      {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                      {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
    float sc[3] = {mod * s[0], mod * s[1], mod * s[2]};
    /* Sigma = R diag(sc^2) R^T, accumulated in the order glm's transpose(M)*M produces:
       Sigma_ij = sum_k (Rm[i][k]*sc[k]) * (Rm[j][k]*sc[k]) */
    float A[3][3];
    for (int i = 0; i < 3; i++)
        for (int k = 0; k < 3; k++) A[i][k] = sc[k] * Rm[i][k];
    float S[3][3];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) S[i][j] = A[i][0] * A[j][0] + A[i][1] * A[j][1] + A[i][2] * A[j][2];
    cov[0] = S[0][0]; cov[1] = S[0][1]; cov[2] = S[0][2];
    cov[3] = S[1][1]; cov[4] = S[1][2]; cov[5] = S[2][2];
}

/* forward.cu:74-113 without the +0.3 (added by the caller so the AA ratio can be formed, as
   upstream dr_aa does). Returns cov2D (xx, xy, yy); optionally the 2x3 matrix T (rows) and t. */
static void cov2d_project(const float *mean, float fx, float fy, float tfx, float tfy, const float *c3,
                          const float *view, float *cov, float *Tout, float *tout, float *txtz_out,
                          float *tytz_out) {
    float t[3];
    xf43(mean, view, t);
    const float limx = 1.3f * tfx, limy = 1.3f * tfy;
    const float txtz = t[0] / t[2], tytz = t[1] / t[2];
    t[0] = fminf(limx, fmaxf(-limx, txtz)) * t[2];
    t[1] = fminf(limy, fmaxf(-limy, tytz)) * t[2];
    /* J (2x3, row form): [fx/tz, 0, -fx*tx/tz^2 ; 0, fy/tz, -fy*ty/tz^2] */
    float J00 = fx / t[2], J02 = -(fx * t[0]) / (t[2] * t[2]);
    float J11 = fy / t[2], J12 = -(fy * t[1]) / (t[2] * t[2]);
    /* Wr = rotation part of the view transform, Wr[i][j] = view[4*j+i] (camera row i, world col j) */
    /* T2 (2x3) = J * Wr */
    float T2[2][3];
    for (int j = 0; j < 3; j++) {
        T2[0][j] = J00 * view[4 * j + 0] + J02 * view[4 * j + 2];
        T2[1][j] = J11 * view[4 * j + 1] + J12 * view[4 * j + 2];
    }
    float V[3][3] = {{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}};
    float TV[2][3];
    for (int i = 0; i < 2; i++)
        for (int j = 0; j < 3; j++) TV[i][j] = T2[i][0] * V[0][j] + T2[i][1] * V[1][j] + T2[i][2] * V[2][j];
    cov[0] = TV[0][0] * T2[0][0] + TV[0][1] * T2[0][1] + TV[0][2] * T2[0][2];
    cov[1] = TV[0][0] * T2[1][0] + TV[0][1] * T2[1][1] + TV[0][2] * T2[1][2];
    cov[2] = TV[1][0] * T2[1][0] + TV[1][1] * T2[1][1] + TV[1][2] * T2[1][2];
    if (Tout) memcpy(Tout, T2, sizeof(T2));
    if (tout) { tout[0] = t[0]; tout[1] = t[1]; tout[2] = t[2]; }
    if (txtz_out) *txtz_out = txtz;
    if (tytz_out) *tytz_out = tytz;
}

/* SH basis for a unit direction, forward.cu:31-60 (same polynomial forms). */
static void sh_basis(int deg, float x, float y, float z, float *b) {
    b[0] = C0;
    if (deg > 0) {
        b[1] = -C1 * y; b[2] = C1 * z; b[3] = -C1 * x;
        if (deg > 1) {
            float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            b[4] = C2[0] * xy; b[5] = C2[1] * yz; b[6] = C2[2] * (2.0f * zz - xx - yy);
            b[7] = C2[3] * xz; b[8] = C2[4] * (xx - yy);
            if (deg > 2) {
                b[9] = C3[0] * y * (3.0f * xx - yy);
                b[10] = C3[1] * xy * z;
                b[11] = C3[2] * y * (4.0f * zz - xx - yy);
                b[12] = C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
                b[13] = C3[4] * x * (4.0f * zz - xx - yy);
                b[14] = C3[5] * z * (xx - yy);
                b[15] = C3[6] * x * (xx - 3.0f * yy);
            }
        }
    }
}

/*
 * Per-Gaussian forward preprocess. forward.cu:155-256 (+ upstream dr_aa for antialiasing, F1).
 * Outputs are the reference's GeometryState arrays (rasterizer_impl.cu:155-170).
 */
void gso_preprocess(const gso_scene *s, int *radii, float *depths, float *means2D, float *cov3D,
                    float *conic_opacity, float *rgb, uint8_t *clamped, uint32_t *tiles_touched) {
    const int P = s->P, W = s->W, H = s->H;
    const float fy = H / (2.0f * s->tan_fovy), fx = W / (2.0f * s->tan_fovx); /* rasterizer_impl.cu:221-222 */
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        radii[i] = 0;
        tiles_touched[i] = 0;
        const float *p = s->means3D + 3 * i;
        float pv[3], ph[4];
        xf43(p, s->view, pv);
        if (pv[2] <= 0.2f) continue; /* auxiliary.h:154 */
        xf44(p, s->proj, ph);
        float pw = 1.0f / (ph[3] + 0.0000001f);
        float pp[2] = {ph[0] * pw, ph[1] * pw};
        const float *c3;
        if (s->cov3D_precomp) c3 = s->cov3D_precomp + 6 * i;
        else {
            cov3d_from_scale_rot(s->scales + 3 * i, s->scale_modifier, s->rotations + 4 * i, cov3D + 6 * i);
            c3 = cov3D + 6 * i;
        }
        float cov[3];
        cov2d_project(p, fx, fy, s->tan_fovx, s->tan_fovy, c3, s->view, cov, 0, 0, 0, 0);
        const float det_cov = cov[0] * cov[2] - cov[1] * cov[1];
        cov[0] += 0.3f; cov[2] += 0.3f; /* forward.cu:110-111 */
        const float det = cov[0] * cov[2] - cov[1] * cov[1];
        float hscale = 1.0f;
        if (s->antialiasing) hscale = sqrtf(fmaxf(0.000025f, det_cov / det));
        if (det == 0.0f) continue;
        float det_inv = 1.f / det;
        float conic[3] = {cov[2] * det_inv, -cov[1] * det_inv, cov[0] * det_inv};
        float mid = 0.5f * (cov[0] + cov[2]);
        float l1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
        float l2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
        float my_radius = ceilf(3.f * sqrtf(fmaxf(l1, l2)));
        float pix[2] = {ndc2pix(pp[0], W), ndc2pix(pp[1], H)};
        int mn[2], mx[2];
        get_rect(pix[0], pix[1], (int)my_radius, gx, gy, mn, mx);
        if ((mx[0] - mn[0]) * (mx[1] - mn[1]) == 0) continue;
        if (!s->colors_precomp) { /* forward.cu:20-71 */
            float dir[3] = {p[0] - s->campos[0], p[1] - s->campos[1], p[2] - s->campos[2]};
            float len = sqrtf(dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2]);
            dir[0] /= len; dir[1] /= len; dir[2] /= len;
            float b[16];
            sh_basis(s->D, dir[0], dir[1], dir[2], b);
            int nb = (s->D + 1) * (s->D + 1);
            const float *sh = s->shs + (size_t)i * s->M * 3;
            for (int ch = 0; ch < 3; ch++) {
                float r = 0.f;
                for (int k = 0; k < nb; k++) r += b[k] * sh[3 * k + ch];
                r += 0.5f;
                clamped[3 * i + ch] = r < 0;
                rgb[3 * i + ch] = fmaxf(r, 0.0f);
            }
        }
        depths[i] = pv[2];
        radii[i] = (int)my_radius;
        means2D[2 * i] = pix[0]; means2D[2 * i + 1] = pix[1];
        conic_opacity[4 * i + 0] = conic[0]; conic_opacity[4 * i + 1] = conic[1];
        conic_opacity[4 * i + 2] = conic[2]; conic_opacity[4 * i + 3] = s->opacities[i] * hscale;
        tiles_touched[i] = (uint32_t)((mx[1] - mn[1]) * (mx[0] - mn[0]));
    }
}

/* rasterizer_impl.cu:35-50 */
uint32_t gso_higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

/* rasterizer_impl.cu:277 : inclusive prefix sum; returns the total (num_rendered, :281). */
uint32_t gso_scan(int P, const uint32_t *tiles_touched, uint32_t *offsets) {
    uint32_t acc = 0;
    for (int i = 0; i < P; i++) { acc += tiles_touched[i]; offsets[i] = acc; }
    return P ? offsets[P - 1] : 0;
}

/* rasterizer_impl.cu:70-111 */
void gso_duplicate_with_keys(int P, int W, int H, const float *means2D, const float *depths,
                             const uint32_t *offsets, const int *radii, uint64_t *keys, uint32_t *vals) {
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(dynamic, 4096)
    for (int i = 0; i < P; i++) {
        if (radii[i] <= 0) continue;
        uint32_t off = i == 0 ? 0 : offsets[i - 1];
        int mn[2], mx[2];
        get_rect(means2D[2 * i], means2D[2 * i + 1], radii[i], gx, gy, mn, mx);
        uint32_t dbits;
        memcpy(&dbits, depths + i, 4);
        for (int y = mn[1]; y < mx[1]; y++)
            for (int x = mn[0]; x < mx[0]; x++) {
                uint64_t key = (uint64_t)(y * gx + x);
                key <<= 32;
                key |= dbits;
                keys[off] = key;
                vals[off] = (uint32_t)i;
                off++;
            }
    }
}

/* rasterizer_impl.cu:303-308 : stable LSD radix sort of (u64 key, u32 value) on bits [0,end_bit). */
void gso_sort_pairs(size_t R, const uint64_t *keys_in, const uint32_t *vals_in, uint64_t *keys_out,
                    uint32_t *vals_out, int end_bit) {
    uint64_t *ka = (uint64_t *)malloc(R * 8 + 8), *kb = (uint64_t *)malloc(R * 8 + 8);
    uint32_t *va = (uint32_t *)malloc(R * 4 + 4), *vb = (uint32_t *)malloc(R * 4 + 4);
    memcpy(ka, keys_in, R * 8);
    memcpy(va, vals_in, R * 4);
    for (int shift = 0; shift < end_bit; shift += 8) {
        int bits = end_bit - shift < 8 ? end_bit - shift : 8;
        uint32_t mask = (1u << bits) - 1;
        size_t cnt[257];
        memset(cnt, 0, sizeof(cnt));
        for (size_t i = 0; i < R; i++) cnt[((ka[i] >> shift) & mask) + 1]++;
        for (int d = 0; d < 256; d++) cnt[d + 1] += cnt[d];
        for (size_t i = 0; i < R; i++) {
            size_t d = cnt[(ka[i] >> shift) & mask]++;
            kb[d] = ka[i];
            vb[d] = va[i];
        }
        uint64_t *tk = ka; ka = kb; kb = tk;
        uint32_t *tv = va; va = vb; vb = tv;
    }
    memcpy(keys_out, ka, R * 8);
    memcpy(vals_out, va, R * 4);
    free(ka); free(kb); free(va); free(vb);
}

/* rasterizer_impl.cu:116-138 after the memset at :310 ; ranges is uint2[tiles]. */
void gso_tile_ranges(size_t R, const uint64_t *keys, int tiles, uint32_t *ranges) {
    memset(ranges, 0, (size_t)tiles * 8);
    for (size_t i = 0; i < R; i++) {
        uint32_t cur = (uint32_t)(keys[i] >> 32);
        if (i == 0) ranges[2 * cur] = 0;
        else {
            uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
            if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)i; ranges[2 * cur] = (uint32_t)i; }
        }
        if (i == R - 1) ranges[2 * cur + 1] = (uint32_t)R;
    }
}

/*
 * Forward blend, forward.cu:261-374 (+ upstream dr_aa inverse depth, F1). One pixel at a time;
 * the per-pixel semantics do not depend on the reference's 256-wide batching.
 *   features: rgb[P,3] (either colors_precomp or the SH result).
 *   invdepth may be NULL.
 */
void gso_render(int W, int H, const uint32_t *ranges, const uint32_t *point_list, const float *means2D,
                const float *features, const float *conic_opacity, const float *depths, const float *bg,
                float *out_color, float *final_T, uint32_t *n_contrib, float *invdepth) {
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(dynamic, 1) collapse(2)
    for (int ty = 0; ty < gy; ty++)
        for (int tx = 0; tx < gx; tx++) {
            const uint32_t r0 = ranges[2 * (ty * gx + tx)], r1 = ranges[2 * (ty * gx + tx) + 1];
            for (int py = ty * TILE; py < (ty + 1) * TILE && py < H; py++)
                for (int px = tx * TILE; px < (tx + 1) * TILE && px < W; px++) {
                    float T = 1.0f, C[3] = {0, 0, 0}, D = 0.f;
                    uint32_t contributor = 0, last = 0;
                    const float fx = (float)px, fy = (float)py;
                    for (uint32_t k = r0; k < r1; k++) {
                        contributor++;
                        const uint32_t g = point_list[k];
                        float dx = means2D[2 * g] - fx, dy = means2D[2 * g + 1] - fy;
                        const float *co = conic_opacity + 4 * g;
                        float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
                        if (power > 0.0f) continue;
                        float alpha = fminf(0.99f, co[3] * expf(power));
                        if (alpha < 1.0f / 255.0f) continue;
                        float test_T = T * (1 - alpha);
                        if (test_T < 0.0001f) break; /* done = true; this Gaussian is not applied */
                        for (int ch = 0; ch < 3; ch++) C[ch] += features[3 * g + ch] * alpha * T;
                        if (invdepth) D += (1.0f / depths[g]) * alpha * T;
                        T = test_T;
                        last = contributor;
                    }
                    const size_t pid = (size_t)py * W + px;
                    final_T[pid] = T;
                    n_contrib[pid] = last;
                    for (int ch = 0; ch < 3; ch++) out_color[(size_t)ch * H * W + pid] = C[ch] + T * bg[ch];
                    if (invdepth) invdepth[pid] = D;
                }
        }
}

/*
 * Backward blend, backward.cu:399-557 (+ upstream dr_aa inverse-depth terms, F1).
 * Per pixel, back to front from the last contributor. Sums over pixels are taken in double
 * (the reference uses float atomics in arbitrary order, so the exact float sum is not defined).
 * Outputs (zero-initialised here): dL_dmean2D[P,3] (NDC-scaled, .z = 0), dL_dconic[P,4]
 * (.x, .y, .w used; .z stays 0), dL_dopacity[P], dL_dcolor[P,3], dL_dinvd[P] (may be NULL).
 */
void gso_render_backward(int P, int W, int H, const uint32_t *ranges, const uint32_t *point_list,
                         const float *bg, const float *means2D, const float *conic_opacity,
                         const float *colors, const float *depths, const float *final_T,
                         const uint32_t *n_contrib, const float *dL_dpix, const float *dL_dinvdepth_pix,
                         float *dL_dmean2D, float *dL_dconic, float *dL_dopacity, float *dL_dcolor,
                         float *dL_dinvd) {
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    double *acc = (double *)calloc((size_t)P * 10, sizeof(double));
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
#pragma omp parallel for schedule(dynamic, 1) collapse(2)
    for (int ty = 0; ty < gy; ty++)
        for (int tx = 0; tx < gx; tx++) {
            const uint32_t r0 = ranges[2 * (ty * gx + tx)];
            for (int py = ty * TILE; py < (ty + 1) * TILE && py < H; py++)
                for (int px = tx * TILE; px < (tx + 1) * TILE && px < W; px++) {
                    const size_t pid = (size_t)py * W + px;
                    const float T_final = final_T[pid];
                    float T = T_final;
                    const uint32_t last = n_contrib[pid];
                    float accum_rec[3] = {0, 0, 0}, last_color[3] = {0, 0, 0}, last_alpha = 0.f;
                    float accum_d = 0.f, last_d = 0.f;
                    float dLp[3];
                    for (int ch = 0; ch < 3; ch++) dLp[ch] = dL_dpix[(size_t)ch * H * W + pid];
                    const float dLd = dL_dinvdepth_pix ? dL_dinvdepth_pix[pid] : 0.f;
                    float bg_dot = 0.f;
                    for (int ch = 0; ch < 3; ch++) bg_dot += bg[ch] * dLp[ch];
                    const float fx = (float)px, fy = (float)py;
                    for (uint32_t c = last; c-- > 0;) {
                        const uint32_t g = point_list[r0 + c];
                        const float dx = means2D[2 * g] - fx, dy = means2D[2 * g + 1] - fy;
                        const float *co = conic_opacity + 4 * g;
                        const float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
                        if (power > 0.0f) continue;
                        const float G = expf(power);
                        const float alpha = fminf(0.99f, co[3] * G);
                        if (alpha < 1.0f / 255.0f) continue;
                        T = T / (1.f - alpha);
                        const float w = alpha * T;
                        float dL_dalpha = 0.f;
                        double *a = acc + (size_t)g * 10;
                        for (int ch = 0; ch < 3; ch++) {
                            const float cc = colors[3 * g + ch];
                            accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                            last_color[ch] = cc;
                            dL_dalpha += (cc - accum_rec[ch]) * dLp[ch];
                            const double v = (double)(w * dLp[ch]);
#pragma omp atomic
                            a[6 + ch] += v;
                        }
                        if (dL_dinvdepth_pix) {
                            const float invd = 1.f / depths[g];
                            accum_d = last_alpha * last_d + (1.f - last_alpha) * accum_d;
                            last_d = invd;
                            dL_dalpha += (invd - accum_d) * dLd;
                            const double v = (double)(w * dLd);
#pragma omp atomic
                            a[9] += v;
                        }
                        dL_dalpha *= T;
                        last_alpha = alpha;
                        dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
                        const float dL_dG = co[3] * dL_dalpha;
                        const float gdx = G * dx, gdy = G * dy;
                        const float dG_ddelx = -gdx * co[0] - gdy * co[1];
                        const float dG_ddely = -gdy * co[2] - gdx * co[1];
                        const double v0 = (double)(dL_dG * dG_ddelx * ddelx_dx), v1 = (double)(dL_dG * dG_ddely * ddely_dy);
                        const double v2 = (double)(-0.5f * gdx * dx * dL_dG), v3 = (double)(-0.5f * gdx * dy * dL_dG);
                        const double v4 = (double)(-0.5f * gdy * dy * dL_dG), v5 = (double)(G * dL_dalpha);
#pragma omp atomic
                        a[0] += v0;
#pragma omp atomic
                        a[1] += v1;
#pragma omp atomic
                        a[2] += v2;
#pragma omp atomic
                        a[3] += v3;
#pragma omp atomic
                        a[4] += v4;
#pragma omp atomic
                        a[5] += v5;
                    }
                }
        }
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        const double *a = acc + (size_t)i * 10;
        dL_dmean2D[3 * i] = (float)a[0]; dL_dmean2D[3 * i + 1] = (float)a[1]; dL_dmean2D[3 * i + 2] = 0.f;
        dL_dconic[4 * i] = (float)a[2]; dL_dconic[4 * i + 1] = (float)a[3]; dL_dconic[4 * i + 2] = 0.f;
        dL_dconic[4 * i + 3] = (float)a[4];
        dL_dopacity[i] = (float)a[5];
        for (int ch = 0; ch < 3; ch++) dL_dcolor[3 * i + ch] = (float)a[6 + ch];
        if (dL_dinvd) dL_dinvd[i] = (float)a[9];
    }
    free(acc);
}

/* backward.cu:20-139 : SH backward for one Gaussian. dmean gets += d(color)/d(mean) term. */
static void sh_backward(int deg, int M, const float *mean, const float *campos, const float *sh,
                        const uint8_t *clamped, const float *dL_dcolor, float *dL_dmean, float *dL_dsh) {
    float d0[3] = {mean[0] - campos[0], mean[1] - campos[1], mean[2] - campos[2]};
    float len = sqrtf(d0[0] * d0[0] + d0[1] * d0[1] + d0[2] * d0[2]);
    float x = d0[0] / len, y = d0[1] / len, z = d0[2] / len;
    float dRGB[3];
    for (int c = 0; c < 3; c++) dRGB[c] = dL_dcolor[c] * (clamped[c] ? 0.f : 1.f);
    float b[16];
    sh_basis(deg, x, y, z, b);
    int nb = (deg + 1) * (deg + 1);
    for (int k = 0; k < nb; k++)
        for (int c = 0; c < 3; c++) dL_dsh[3 * k + c] = b[k] * dRGB[c];
    (void)M;
    float dx[3] = {0, 0, 0}, dy[3] = {0, 0, 0}, dz[3] = {0, 0, 0};
#define SH(k) (sh[3 * (k) + c])
    for (int c = 0; c < 3; c++) {
        if (deg > 0) {
            dx[c] = -C1 * SH(3); dy[c] = -C1 * SH(1); dz[c] = C1 * SH(2);
            if (deg > 1) {
                float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                dx[c] += C2[0] * y * SH(4) + C2[2] * 2.f * -x * SH(6) + C2[3] * z * SH(7) + C2[4] * 2.f * x * SH(8);
                dy[c] += C2[0] * x * SH(4) + C2[1] * z * SH(5) + C2[2] * 2.f * -y * SH(6) + C2[4] * 2.f * -y * SH(8);
                dz[c] += C2[1] * y * SH(5) + C2[2] * 2.f * 2.f * z * SH(6) + C2[3] * x * SH(7);
                if (deg > 2) {
                    dx[c] += (C3[0] * SH(9) * 3.f * 2.f * xy + C3[1] * SH(10) * yz + C3[2] * SH(11) * -2.f * xy +
                              C3[3] * SH(12) * -3.f * 2.f * xz + C3[4] * SH(13) * (-3.f * xx + 4.f * zz - yy) +
                              C3[5] * SH(14) * 2.f * xz + C3[6] * SH(15) * 3.f * (xx - yy));
                    dy[c] += (C3[0] * SH(9) * 3.f * (xx - yy) + C3[1] * SH(10) * xz +
                              C3[2] * SH(11) * (-3.f * yy + 4.f * zz - xx) + C3[3] * SH(12) * -3.f * 2.f * yz +
                              C3[4] * SH(13) * -2.f * xy + C3[5] * SH(14) * -2.f * yz + C3[6] * SH(15) * -3.f * 2.f * xy);
                    dz[c] += (C3[1] * SH(10) * xy + C3[2] * SH(11) * 4.f * 2.f * yz +
                              C3[3] * SH(12) * 3.f * (2.f * zz - xx - yy) + C3[4] * SH(13) * 4.f * 2.f * xz +
                              C3[5] * SH(14) * (xx - yy));
                }
            }
        }
    }
#undef SH
    float ddir[3] = {dx[0] * dRGB[0] + dx[1] * dRGB[1] + dx[2] * dRGB[2],
                     dy[0] * dRGB[0] + dy[1] * dRGB[1] + dy[2] * dRGB[2],
                     dz[0] * dRGB[0] + dz[1] * dRGB[1] + dz[2] * dRGB[2]};
    /* auxiliary.h:107-117 dnormvdv */
    float s2 = d0[0] * d0[0] + d0[1] * d0[1] + d0[2] * d0[2];
    float inv = 1.0f / sqrtf(s2 * s2 * s2);
    dL_dmean[0] += ((+s2 - d0[0] * d0[0]) * ddir[0] - d0[1] * d0[0] * ddir[1] - d0[2] * d0[0] * ddir[2]) * inv;
    dL_dmean[1] += (-d0[0] * d0[1] * ddir[0] + (s2 - d0[1] * d0[1]) * ddir[1] - d0[2] * d0[1] * ddir[2]) * inv;
    dL_dmean[2] += (-d0[0] * d0[2] * ddir[0] - d0[1] * d0[2] * ddir[1] + (s2 - d0[2] * d0[2]) * ddir[2]) * inv;
}

/* backward.cu:278-341 : dL/dSigma -> dL/dscale, dL/dquaternion (no normalisation backward, :340;
   the scale gradient is w.r.t. mod*scale exactly as the reference leaves it). */
static void cov3d_backward(const float *scale, float mod, const float *q, const float *dcov, float *dscale,
                           float *drot) {
    float r = q[0], x = q[1], y = q[2], z = q[3];
    float Rm[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                      {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                      {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
    float s[3] = {mod * scale[0], mod * scale[1], mod * scale[2]};
    float dS[3][3] = {{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]},
                      {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                      {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}};
    /* Mm[i][k] = s_i * Rm[k][i];  dM = 2 * Mm * dS */
    float dM[3][3];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            float a = 0.f;
            for (int k = 0; k < 3; k++) a += (s[i] * Rm[k][i]) * dS[k][j];
            dM[i][j] = 2.0f * a;
        }
    for (int i = 0; i < 3; i++) dscale[i] = Rm[0][i] * dM[i][0] + Rm[1][i] * dM[i][1] + Rm[2][i] * dM[i][2];
    /* Mt[a][b] in the reference's glm indexing = s_a * dM[a][b] */
    float Mt[3][3];
    for (int a = 0; a < 3; a++)
        for (int b2 = 0; b2 < 3; b2++) Mt[a][b2] = s[a] * dM[a][b2];
    drot[0] = 2 * z * (Mt[0][1] - Mt[1][0]) + 2 * y * (Mt[2][0] - Mt[0][2]) + 2 * x * (Mt[1][2] - Mt[2][1]);
    drot[1] = 2 * y * (Mt[1][0] + Mt[0][1]) + 2 * z * (Mt[2][0] + Mt[0][2]) + 2 * r * (Mt[1][2] - Mt[2][1]) - 4 * x * (Mt[2][2] + Mt[1][1]);
    drot[2] = 2 * x * (Mt[1][0] + Mt[0][1]) + 2 * r * (Mt[2][0] - Mt[0][2]) + 2 * z * (Mt[1][2] + Mt[2][1]) - 4 * y * (Mt[2][2] + Mt[0][0]);
    drot[3] = 2 * r * (Mt[0][1] - Mt[1][0]) + 2 * x * (Mt[2][0] + Mt[0][2]) + 2 * y * (Mt[1][2] + Mt[2][1]) - 4 * z * (Mt[1][1] + Mt[0][0]);
}

/*
 * Backward preprocess: backward.cu:144-274 (cov2D/conic), :346-396 (projection, SH, cov3D),
 * + upstream dr_aa antialiasing / inverse-depth terms (F1).
 * In: dL_dmean2D[P,3], dL_dconic[P,4], dL_dopacity[P] (IN/OUT: scaled by the AA factor),
 *     dL_dcolor[P,3], dL_dinvd[P] or NULL. cov3D = precomputed or the forward's cov3D[P,6].
 * Out (all fully written, zeros for radii<=0): dL_dmean3D[P,3], dL_dcov3D[P,6], dL_dsh[P,M,3],
 *     dL_dscale[P,3], dL_drot[P,4].
 */
void gso_preprocess_backward(const gso_scene *s, const int *radii, const float *cov3D, const uint8_t *clamped,
                             const float *dL_dmean2D, const float *dL_dconic, float *dL_dopacity,
                             const float *dL_dcolor, const float *dL_dinvd, float *dL_dmean3D,
                             float *dL_dcov3D, float *dL_dsh, float *dL_dscale, float *dL_drot) {
    const int P = s->P, M = s->M;
    const float fy = s->H / (2.0f * s->tan_fovy), fx = s->W / (2.0f * s->tan_fovx);
    const float *view = s->view, *proj = s->proj;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        float *dm = dL_dmean3D + 3 * i, *dc = dL_dcov3D + 6 * i;
        dm[0] = dm[1] = dm[2] = 0.f;
        for (int k = 0; k < 6; k++) dc[k] = 0.f;
        if (s->shs) memset(dL_dsh + (size_t)i * M * 3, 0, sizeof(float) * M * 3);
        if (s->scales) {
            dL_dscale[3 * i] = dL_dscale[3 * i + 1] = dL_dscale[3 * i + 2] = 0.f;
            dL_drot[4 * i] = dL_drot[4 * i + 1] = dL_drot[4 * i + 2] = dL_drot[4 * i + 3] = 0.f;
        }
        if (!(radii[i] > 0)) continue;
        const float *mean = s->means3D + 3 * i;
        const float *c3 = cov3D + 6 * i;
        /* ---- backward.cu:144-274 ---- */
        float cov[3], T2[2][3], t[3], txtz, tytz;
        cov2d_project(mean, fx, fy, s->tan_fovx, s->tan_fovy, c3, view, cov, &T2[0][0], t, &txtz, &tytz);
        const float limx = 1.3f * s->tan_fovx, limy = 1.3f * s->tan_fovy;
        const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
        const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
        const float dcon[3] = {dL_dconic[4 * i], dL_dconic[4 * i + 1], dL_dconic[4 * i + 3]};
        float aa_dxx = 0.f, aa_dxy = 0.f, aa_dyy = 0.f;
        if (s->antialiasing) {
            /* h = sqrt(max(2.5e-5, N/D)), N = xy - z^2, D = (x+w)(y+w) - z^2, stored opacity = o*h */
            const float w = 0.3f, X = cov[0], Y = cov[2], Z = cov[1];
            const float N = X * Y - Z * Z, D = (X + w) * (Y + w) - Z * Z;
            const float ratio = N / D;
            const float h = sqrtf(fmaxf(0.000025f, ratio));
            const float g_stored = dL_dopacity[i];
            dL_dopacity[i] = g_stored * h;
            const float dL_dh = g_stored * s->opacities[i];
            const float dL_dratio = ratio <= 0.000025f ? 0.f : dL_dh / (2.f * h);
            const float f = dL_dratio / (D * D);
            aa_dxx = w * (w * Y + Y * Y + Z * Z) * f;
            aa_dyy = w * (w * X + X * X + Z * Z) * f;
            aa_dxy = -2.f * w * Z * (w + X + Y) * f;
        }
        const float a = cov[0] + 0.3f, b = cov[1], c = cov[2] + 0.3f;
        const float denom = a * c - b * b;
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        if (denom2inv != 0) {
            dL_da = denom2inv * (-c * c * dcon[0] + 2 * b * c * dcon[1] + (denom - a * c) * dcon[2]);
            dL_dc = denom2inv * (-a * a * dcon[2] + 2 * a * b * dcon[1] + (denom - a * c) * dcon[0]);
            dL_db = denom2inv * 2 * (b * c * dcon[0] - (denom + 2 * b * b) * dcon[1] + a * b * dcon[2]);
        }
        dL_da += aa_dxx; dL_dc += aa_dyy; dL_db += aa_dxy;
        if (denom2inv != 0 || s->antialiasing) {
            dc[0] = (T2[0][0] * T2[0][0] * dL_da + T2[0][0] * T2[1][0] * dL_db + T2[1][0] * T2[1][0] * dL_dc);
            dc[3] = (T2[0][1] * T2[0][1] * dL_da + T2[0][1] * T2[1][1] * dL_db + T2[1][1] * T2[1][1] * dL_dc);
            dc[5] = (T2[0][2] * T2[0][2] * dL_da + T2[0][2] * T2[1][2] * dL_db + T2[1][2] * T2[1][2] * dL_dc);
            dc[1] = 2 * T2[0][0] * T2[0][1] * dL_da + (T2[0][0] * T2[1][1] + T2[0][1] * T2[1][0]) * dL_db + 2 * T2[1][0] * T2[1][1] * dL_dc;
            dc[2] = 2 * T2[0][0] * T2[0][2] * dL_da + (T2[0][0] * T2[1][2] + T2[0][2] * T2[1][0]) * dL_db + 2 * T2[1][0] * T2[1][2] * dL_dc;
            dc[4] = 2 * T2[0][2] * T2[0][1] * dL_da + (T2[0][1] * T2[1][2] + T2[0][2] * T2[1][1]) * dL_db + 2 * T2[1][1] * T2[1][2] * dL_dc;
        }
        const float V[3][3] = {{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}};
        float dT[2][3];
        for (int j = 0; j < 3; j++) {
            const float tv0 = T2[0][0] * V[j][0] + T2[0][1] * V[j][1] + T2[0][2] * V[j][2];
            const float tv1 = T2[1][0] * V[j][0] + T2[1][1] * V[j][1] + T2[1][2] * V[j][2];
            dT[0][j] = 2 * tv0 * dL_da + tv1 * dL_db;
            dT[1][j] = 2 * tv1 * dL_dc + tv0 * dL_db;
        }
        /* Rv[i][j] = view[4*j+i] */
        const float dJ00 = view[0] * dT[0][0] + view[4] * dT[0][1] + view[8] * dT[0][2];
        const float dJ02 = view[2] * dT[0][0] + view[6] * dT[0][1] + view[10] * dT[0][2];
        const float dJ11 = view[1] * dT[1][0] + view[5] * dT[1][1] + view[9] * dT[1][2];
        const float dJ12 = view[2] * dT[1][0] + view[6] * dT[1][1] + view[10] * dT[1][2];
        const float tz = 1.f / t[2], tz2 = tz * tz, tz3 = tz2 * tz;
        float dtx = x_grad_mul * -fx * tz2 * dJ02;
        float dty = y_grad_mul * -fy * tz2 * dJ12;
        float dtz = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (2 * fx * t[0]) * tz3 * dJ02 + (2 * fy * t[1]) * tz3 * dJ12;
        if (dL_dinvd) dtz += -dL_dinvd[i] * tz2; /* d(1/t.z)/dt.z ; t.z is never clamped */
        dm[0] = view[0] * dtx + view[1] * dty + view[2] * dtz;
        dm[1] = view[4] * dtx + view[5] * dty + view[6] * dtz;
        dm[2] = view[8] * dtx + view[9] * dty + view[10] * dtz;
        /* ---- backward.cu:346-396 ---- */
        float mh[4];
        xf44(mean, proj, mh);
        const float m_w = 1.0f / (mh[3] + 0.0000001f);
        const float mul1 = mh[0] * m_w * m_w, mul2 = mh[1] * m_w * m_w;
        const float g0 = dL_dmean2D[3 * i], g1 = dL_dmean2D[3 * i + 1];
        dm[0] += (proj[0] * m_w - proj[3] * mul1) * g0 + (proj[1] * m_w - proj[3] * mul2) * g1;
        dm[1] += (proj[4] * m_w - proj[7] * mul1) * g0 + (proj[5] * m_w - proj[7] * mul2) * g1;
        dm[2] += (proj[8] * m_w - proj[11] * mul1) * g0 + (proj[9] * m_w - proj[11] * mul2) * g1;
        if (s->shs)
            sh_backward(s->D, M, mean, s->campos, s->shs + (size_t)i * M * 3, clamped + 3 * i, dL_dcolor + 3 * i, dm,
                        dL_dsh + (size_t)i * M * 3);
        if (s->scales)
            cov3d_backward(s->scales + 3 * i, s->scale_modifier, s->rotations + 4 * i, dc, dL_dscale + 3 * i,
                           dL_drot + 4 * i);
    }
}

/* rasterizer_impl.cu:54-66 */
void gso_mark_visible(int P, const float *means3D, const float *view, uint8_t *present) {
    for (int i = 0; i < P; i++) {
        float pv[3];
        xf43(means3D + 3 * i, view, pv);
        present[i] = pv[2] > 0.2f;
    }
}

/*
 * simple-knn: distCUDA2 result semantics (simple_knn.cu:132-184): mean of the three smallest
 * squared distances to OTHER points (by index; duplicates count). Brute force, O(P^2).
 */
void gso_dist2(int P, const float *pts, float *out) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
        const float rx = pts[3 * i], ry = pts[3 * i + 1], rz = pts[3 * i + 2];
        for (int j = 0; j < P; j++) {
            if (j == i) continue;
            float dx = pts[3 * j] - rx, dy = pts[3 * j + 1] - ry, dz = pts[3 * j + 2] - rz;
            float d = dx * dx + dy * dy + dz * dz;
            for (int k = 0; k < 3; k++)
                if (best[k] > d) { float t = best[k]; best[k] = d; d = t; }
        }
        out[i] = (best[0] + best[1] + best[2]) / 3.0f;
    }
}
