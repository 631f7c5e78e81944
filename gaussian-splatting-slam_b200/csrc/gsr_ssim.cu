// gsr_ssim.cu -- SSIM map forward / backward, optionally fused with the L1 term of the
// training loss. Replaces fusedssimCUDA / fusedssim_backwardCUDA (submodules/fused-ssim/ssim.cu:187-366).
//
// Same arithmetic as the reference: 11-tap Gaussian (sigma 1.5), separable, zero padding, horizontal pass first with
// the taps accumulated in ascending order, then the vertical pass; the map and its three partial derivatives
// (d/dmu1, d/dsigma1^2, d/dsigma12) are written in the forward and convolved again in the backward.
// What differs is the schedule. The reference runs the five convolutions of a channel one after another through one
// scratch plane with ~25 block-wide barriers per channel and loops over the channels inside the block. Here a CTA
// owns a 32x32 tile of ONE plane (grid.z = batch * channels), stages both images once, runs the five horizontal
// convolutions together out of registers (sliding window, 4 outputs per thread), and the five vertical ones after a
// single barrier: 2 barriers per tile. HBM-bound: 8 B/pixel read + 16 B/pixel written (forward, training), 20-24 B
// read + 4 B written (backward).
//
// Fused loss mode (the caller's loss = (1-l)*mean|x-y| + l*(1 - mean(ssim)), train.py:114-121): the forward also
// accumulates sum(ssim) and sum|x-y| (one double atomic per CTA), the backward takes dL/dmap as a scalar and adds
// the L1 term's sign(x-y), so neither the map nor dL/dmap ever travel through HBM.
#include "gsr_common.cuh"

namespace gsr {

constexpr int kST = 32;                 // tile edge (outputs)
constexpr int kSH = 5;                  // halo
constexpr int kSP = kST + 2 * kSH;      // staged edge: 42
constexpr int kSPad = 45;               // row stride of the staged tiles (odd: lanes walking down a column hit 32 banks)
constexpr int kHPad = kST + 1;          // row stride of the horizontally filtered planes (same reason)
constexpr int kSsimThreads = 256;

__constant__ float kG[11] = {0.001028380123898387f, 0.0075987582094967365f, 0.036000773310661316f, 0.10936068743467331f,
                             0.21300552785396576f,  0.26601171493530273f,   0.21300552785396576f,  0.10936068743467331f,
                             0.036000773310661316f, 0.0075987582094967365f, 0.001028380123898387f};

__device__ __forceinline__ float block_sum_256(float v, float* red) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = 0.f;
    if (threadIdx.x < 8) t = red[threadIdx.x];
    for (int o = 4; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    __syncthreads();
    return t;                            // valid in thread 0
}

template <bool TRAIN>
__global__ void __launch_bounds__(kSsimThreads) ssim_fwd_kernel(int H, int W, float C1, float C2,
                                                                 const float* __restrict__ img1,
                                                                 const float* __restrict__ img2,
                                                                 float* __restrict__ ssim_map, float* __restrict__ dm_dmu1,
                                                                 float* __restrict__ dm_dsigma1_sq,
                                                                 float* __restrict__ dm_dsigma12, double* __restrict__ sums) {
    __shared__ float a[kSP][kSPad], b[kSP][kSPad];
    __shared__ float h[5][kSP][kHPad];
    __shared__ float red[8];
    const int tid = threadIdx.x;
    const size_t plane = (size_t)blockIdx.z * H * W;
    const int x0 = blockIdx.x * kST, y0 = blockIdx.y * kST;
    for (int i = tid; i < kSP * kSP; i += kSsimThreads) {
        const int ly = i / kSP, lx = i - ly * kSP;
        const int gy = y0 + ly - kSH, gx = x0 + lx - kSH;
        const bool in = gx >= 0 && gx < W && gy >= 0 && gy < H;
        a[ly][lx] = in ? __ldg(img1 + plane + (size_t)gy * W + gx) : 0.0f;
        b[ly][lx] = in ? __ldg(img2 + plane + (size_t)gy * W + gx) : 0.0f;
    }
    __syncthreads();
    // horizontal: 8 segments of 4 outputs x 42 rows (consecutive lanes = consecutive rows: conflict-free with the odd
    // strides); 14-wide register window per image
    for (int item = tid; item < kSP * (kST / 4); item += kSsimThreads) {
        const int seg = item / kSP, row = item - seg * kSP, c0 = seg * 4;
        float va[14], vb[14];
#pragma unroll
        for (int k = 0; k < 14; k++) { va[k] = a[row][c0 + k]; vb[k] = b[row][c0 + k]; }
#pragma unroll
        for (int j = 0; j < 4; j++) {
            float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f, s4 = 0.f;
#pragma unroll
            for (int k = 0; k < 11; k++) {
                const float x = va[j + k], y = vb[j + k], g = kG[k];
                s0 += g * x; s1 += g * (x * x); s2 += g * y; s3 += g * (y * y); s4 += g * (x * y);
            }
            h[0][row][c0 + j] = s0; h[1][row][c0 + j] = s1; h[2][row][c0 + j] = s2; h[3][row][c0 + j] = s3;
            h[4][row][c0 + j] = s4;
        }
    }
    __syncthreads();
    // vertical: thread = (column, group of 4 rows)
    const int col = tid & 31, r0 = (tid >> 5) * 4;
    float out[5][4];
#pragma unroll
    for (int s = 0; s < 5; s++) {
        float v[14];
#pragma unroll
        for (int k = 0; k < 14; k++) v[k] = h[s][r0 + k][col];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            float acc = 0.f;
#pragma unroll
            for (int k = 0; k < 11; k++) acc += kG[k] * v[j + k];
            out[s][j] = acc;
        }
    }
    float part_ssim = 0.f, part_l1 = 0.f;
    const int gx = x0 + col;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int gy = y0 + r0 + j;
        if (gx < W && gy < H) {
            const float mu1 = out[0][j], mu2 = out[2][j];
            const float sigma1_sq = out[1][j] - mu1 * mu1;
            const float sigma2_sq = out[3][j] - mu2 * mu2;
            const float sigma12 = out[4][j] - mu1 * mu2;
            const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu1_mu2 = mu1 * mu2;
            const float C = (2.0f * mu1_mu2 + C1);
            const float D = (2.0f * sigma12 + C2);
            const float A = (mu1_sq + mu2_sq + C1);
            const float B = (sigma1_sq + sigma2_sq + C2);
            const float m = (C * D) / (A * B);
            const size_t gi = plane + (size_t)gy * W + gx;
            if (ssim_map) ssim_map[gi] = m;
            if (TRAIN) {
                dm_dmu1[gi] = ((mu2 * 2.0f * D) / (A * B) - (mu2 * 2.0f * C) / (A * B) - (mu1 * 2.0f * C * D) / (A * A * B) +
                               (mu1 * 2.0f * C * D) / (A * B * B));
                dm_dsigma1_sq[gi] = ((-C * D) / (A * B * B));
                dm_dsigma12[gi] = ((2 * C) / (A * B));
            }
            part_ssim += m;
            part_l1 += fabsf(a[r0 + j + kSH][col + kSH] - b[r0 + j + kSH][col + kSH]);
        }
    }
    if (sums) {
        const float ts = block_sum_256(part_ssim, red);
        const float tl = block_sum_256(part_l1, red);
        if (tid == 0) { atomicAdd(sums, (double)ts); atomicAdd(sums + 1, (double)tl); }
    }
}

// dL/dimg1 = conv(dL/dmap * dm/dmu1) + 2*img1*conv(dL/dmap * dm/dsigma1^2) + img2*conv(dL/dmap * dm/dsigma12)
// UNIFORM: dL/dmap == map_scale everywhere (a mean), and l1_scale * sign(img1 - img2) is added (fused L1 term).
template <bool UNIFORM>
__global__ void __launch_bounds__(kSsimThreads) ssim_bwd_kernel(int H, int W, const float* __restrict__ img1,
                                                                 const float* __restrict__ img2,
                                                                 const float* __restrict__ dL_dmap, float map_scale,
                                                                 float l1_scale, const float* __restrict__ upstream,
                                                                 const float* __restrict__ dm_dmu1,
                                                                 const float* __restrict__ dm_dsigma1_sq,
                                                                 const float* __restrict__ dm_dsigma12,
                                                                 float* __restrict__ dL_dimg1) {
    __shared__ float p[3][kSP][kSPad];
    __shared__ float h[3][kSP][kHPad];
    const int tid = threadIdx.x;
    const size_t plane = (size_t)blockIdx.z * H * W;
    const int x0 = blockIdx.x * kST, y0 = blockIdx.y * kST;
    if (UNIFORM && upstream) { const float u = __ldg(upstream); map_scale *= u; l1_scale *= u; }
    for (int i = tid; i < kSP * kSP; i += kSsimThreads) {
        const int ly = i / kSP, lx = i - ly * kSP;
        const int gy = y0 + ly - kSH, gx = x0 + lx - kSH;
        float v0 = 0.f, v1 = 0.f, v2 = 0.f;
        if (gx >= 0 && gx < W && gy >= 0 && gy < H) {
            const size_t gi = plane + (size_t)gy * W + gx;
            const float d = UNIFORM ? map_scale : __ldg(dL_dmap + gi);
            v0 = __ldg(dm_dmu1 + gi) * d; v1 = __ldg(dm_dsigma1_sq + gi) * d; v2 = __ldg(dm_dsigma12 + gi) * d;
        }
        p[0][ly][lx] = v0; p[1][ly][lx] = v1; p[2][ly][lx] = v2;
    }
    __syncthreads();
    for (int item = tid; item < kSP * (kST / 4); item += kSsimThreads) {
        const int seg = item / kSP, row = item - seg * kSP, c0 = seg * 4;
#pragma unroll
        for (int s = 0; s < 3; s++) {
            float v[14];
#pragma unroll
            for (int k = 0; k < 14; k++) v[k] = p[s][row][c0 + k];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                float acc = 0.f;
#pragma unroll
                for (int k = 0; k < 11; k++) acc += kG[k] * v[j + k];
                h[s][row][c0 + j] = acc;
            }
        }
    }
    __syncthreads();
    const int col = tid & 31, r0 = (tid >> 5) * 4;
    float out[3][4];
#pragma unroll
    for (int s = 0; s < 3; s++) {
        float v[14];
#pragma unroll
        for (int k = 0; k < 14; k++) v[k] = h[s][r0 + k][col];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            float acc = 0.f;
#pragma unroll
            for (int k = 0; k < 11; k++) acc += kG[k] * v[j + k];
            out[s][j] = acc;
        }
    }
    const int gx = x0 + col;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int gy = y0 + r0 + j;
        if (gx < W && gy < H) {
            const size_t gi = plane + (size_t)gy * W + gx;
            const float pix1 = __ldg(img1 + gi), pix2 = __ldg(img2 + gi);
            float g = 0.0f;
            g += out[0][j];
            g += pix1 * 2.0f * out[1][j];
            g += pix2 * out[2][j];
            if (UNIFORM) {
                const float d = pix1 - pix2;
                g += l1_scale * (d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f));
            }
            dL_dimg1[gi] = g;
        }
    }
}

int launch_ssim_forward(int planes, int H, int W, float C1, float C2, const float* img1, const float* img2, float* ssim_map,
                        float* dm_dmu1, float* dm_dsigma1_sq, float* dm_dsigma12, double* sums, cudaStream_t s) {
    const dim3 grid((W + kST - 1) / kST, (H + kST - 1) / kST, planes);
    if (dm_dmu1)
        GSR_LAUNCH("ssim_forward", false, s,
                   (ssim_fwd_kernel<true><<<grid, kSsimThreads, 0, s>>>(H, W, C1, C2, img1, img2, ssim_map, dm_dmu1,
                                                                        dm_dsigma1_sq, dm_dsigma12, sums)));
    else
        GSR_LAUNCH("ssim_forward", false, s,
                   (ssim_fwd_kernel<false><<<grid, kSsimThreads, 0, s>>>(H, W, C1, C2, img1, img2, ssim_map, nullptr, nullptr,
                                                                         nullptr, sums)));
    return 0;
}

int launch_ssim_backward(int planes, int H, int W, const float* img1, const float* img2, const float* dL_dmap,
                         float map_scale, float l1_scale, const float* upstream, const float* dm_dmu1,
                         const float* dm_dsigma1_sq, const float* dm_dsigma12, float* dL_dimg1, cudaStream_t s) {
    const dim3 grid((W + kST - 1) / kST, (H + kST - 1) / kST, planes);
    if (dL_dmap)
        GSR_LAUNCH("ssim_backward", false, s,
                   (ssim_bwd_kernel<false><<<grid, kSsimThreads, 0, s>>>(H, W, img1, img2, dL_dmap, 0.f, 0.f, nullptr, dm_dmu1,
                                                                         dm_dsigma1_sq, dm_dsigma12, dL_dimg1)));
    else
        GSR_LAUNCH("ssim_backward", false, s,
                   (ssim_bwd_kernel<true><<<grid, kSsimThreads, 0, s>>>(H, W, img1, img2, nullptr, map_scale, l1_scale, upstream, dm_dmu1,
                                                                        dm_dsigma1_sq, dm_dsigma12, dL_dimg1)));
    return 0;
}

}  // namespace gsr
