// gsr_adam.cu -- one-launch Adam step over all parameter groups of the Gaussian model (SURVEY.md 8f rank 2 and the
// optimiser half of rank 3). Replaces the per-group torch.optim.Adam step the reference runs after the rasterizer
// backward (scene/gaussian_model.py:155-176: six groups, eps 1e-15, per-group learning rates; train.py:171-176) --
// about 6 x 10 element-wise launches per step there. Pure HBM streaming: read p, g, m, v and write p, m, v =
// 28 B per element, 1652 B per Gaussian at SH degree 3.
//
// Dense mode follows torch.optim.Adam (betas, eps, bias correction by the per-group step count, no weight decay,
// no amsgrad). With a visibility mask (one byte per Gaussian, the radii > 0 of the view) rows of invisible Gaussians
// are skipped entirely -- parameters AND moments -- which is the SparseGaussianAdam contract the reference's caller
// switches to when the rasterizer exports it (train.py:37-41,173-176; gaussian_model.py:168-173).
#include "gsr_common.cuh"

namespace gsr {

struct AdamTable {
    gsr_adam_group g[GSR_ADAM_MAX_GROUPS];
    unsigned block_begin[GSR_ADAM_MAX_GROUPS + 1];   // first CTA of each group
    int n;
};

constexpr int kAdamThreads = 256;
constexpr int kAdamVec = 4;
constexpr int kAdamPerCta = kAdamThreads * kAdamVec * 4;     // 4096 elements per CTA

__global__ void __launch_bounds__(kAdamThreads) adam_kernel(const __grid_constant__ AdamTable t, const AdamHyper h,
                                                             const uint8_t* __restrict__ visible) {
    int gi = 0;
#pragma unroll
    for (int k = 1; k < GSR_ADAM_MAX_GROUPS; k++)
        if (k < t.n && blockIdx.x >= t.block_begin[k]) gi = k;
    const gsr_adam_group& G = t.g[gi];
    const long long base = (long long)(blockIdx.x - t.block_begin[gi]) * kAdamPerCta;
    const bool vec_ok = ((reinterpret_cast<uintptr_t>(G.param) | reinterpret_cast<uintptr_t>(G.grad) |
                          reinterpret_cast<uintptr_t>(G.exp_avg) | reinterpret_cast<uintptr_t>(G.exp_avg_sq)) & 15) == 0;
#pragma unroll
    for (int it = 0; it < 4; it++) {
        const long long i = base + ((long long)it * kAdamThreads + threadIdx.x) * kAdamVec;
        if (i >= G.numel) return;
        if (vec_ok && i + kAdamVec <= G.numel) {
            bool any = true;
            bool vis[4] = {true, true, true, true};
            if (visible) {
#pragma unroll
                for (int k = 0; k < 4; k++) vis[k] = visible[(i + k) / G.row_len] != 0;
                any = vis[0] || vis[1] || vis[2] || vis[3];
            }
            if (!any) continue;
            float4 p = *reinterpret_cast<float4*>(G.param + i);
            const float4 g = __ldg(reinterpret_cast<const float4*>(G.grad + i));
            float4 m = *reinterpret_cast<float4*>(G.exp_avg + i);
            float4 v = *reinterpret_cast<float4*>(G.exp_avg_sq + i);
            if (vis[0]) adam_one(p.x, g.x, m.x, v.x, h, G.step_size, G.inv_sqrt_bc2);
            if (vis[1]) adam_one(p.y, g.y, m.y, v.y, h, G.step_size, G.inv_sqrt_bc2);
            if (vis[2]) adam_one(p.z, g.z, m.z, v.z, h, G.step_size, G.inv_sqrt_bc2);
            if (vis[3]) adam_one(p.w, g.w, m.w, v.w, h, G.step_size, G.inv_sqrt_bc2);
            *reinterpret_cast<float4*>(G.param + i) = p;
            *reinterpret_cast<float4*>(G.exp_avg + i) = m;
            *reinterpret_cast<float4*>(G.exp_avg_sq + i) = v;
        } else {
            for (int k = 0; k < kAdamVec && i + k < G.numel; k++) {
                if (visible && visible[(i + k) / G.row_len] == 0) continue;
                float p = G.param[i + k], m = G.exp_avg[i + k], v = G.exp_avg_sq[i + k];
                adam_one(p, G.grad[i + k], m, v, h, G.step_size, G.inv_sqrt_bc2);
                G.param[i + k] = p; G.exp_avg[i + k] = m; G.exp_avg_sq[i + k] = v;
            }
        }
    }
}

int launch_adam(const gsr_adam_group* groups, int n, double b1, double b2, double eps, const uint8_t* visible, cudaStream_t s) {
    AdamTable t;
    t.n = n;
    unsigned blocks = 0;
    for (int k = 0; k < n; k++) {
        t.g[k] = groups[k];
        t.block_begin[k] = blocks;
        blocks += (unsigned)((groups[k].numel + kAdamPerCta - 1) / kAdamPerCta);
    }
    t.block_begin[n] = blocks;
    if (blocks == 0) return 0;
    GSR_LAUNCH("adam_step", false, s, (adam_kernel<<<blocks, kAdamThreads, 0, s>>>(t, AdamHyper{(float)(1.0 - b1), (float)b2, (float)(1.0 - b2), (float)eps}, visible)));
    return 0;
}

}  // namespace gsr
