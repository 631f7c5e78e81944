// gsr_densify.cu -- densification / pruning of the Gaussian map on the device.
//
// The reference grows and thins the map with ~30 torch ops per call (scene/gaussian_model.py:274-433): boolean-mask
// gathers and torch.cat over each of the 6 parameter tensors and their 12 Adam moments, three times over (clone,
// split, prune), with host synchronisations in between. Here the whole of densify_and_prune (:412-429) is
//   plan   classify every Gaussian once (clone / split / prune decisions of :393-425 folded together), count,
//          scan, and write each row's destination indices;
//   apply  ONE streaming pass that moves every row of every tensor (parameters and moments) to its final place,
//          computing the split children's positions and scales on the way.
// The result has the reference's exact row order: surviving originals, then surviving clones, then the split
// children, repetition-major (the .repeat(N, 1) of :380-387). prune_points (:307-322) is the same machinery with a
// caller-supplied mask. add_densification_stats (:431-433) and the caller's max_radii2D update (train.py:159) are
// one small kernel.
#include "gsr_common.cuh"

namespace gsr {

constexpr int kDenThreads = 256;

enum : uint8_t { kKeepOrig = 1, kCloneKept = 2, kSplitSel = 4, kSplitKept = 8 };

struct DensifyPlan {               // carved from the caller's plan blob
    uint32_t* counts;              // [8]: kept originals, kept clones, split-selected parents, kept split parents
    int4* dest;                    // [P] {row of the original, row of its clone, row of its first split child, noise row} or -1
    uint8_t* flags;                // [P]
    uint4* block_sums;             // [blocks + 1]
    static DensifyPlan from(void* blob, size_t P) {
        char* p = static_cast<char*>(blob);
        DensifyPlan d;
        d.counts = carve<uint32_t>(p, 8);
        d.dest = carve<int4>(p, P);
        d.flags = carve<uint8_t>(p, P);
        d.block_sums = carve<uint4>(p, (P + kDenThreads - 1) / kDenThreads + 1);
        return d;
    }
};

size_t densify_plan_bytes(size_t P) {          // layout measured from address 0, plus slack for the blob's own alignment
    DensifyPlan d = DensifyPlan::from(nullptr, P);
    return reinterpret_cast<size_t>(d.block_sums + (P + kDenThreads - 1) / kDenThreads + 1) + GSR_ALIGN;
}

// torch.sigmoid on a float tensor: 1 / (1 + exp(-x)) in fp32
__device__ __forceinline__ float sigmoid_ref(float x) { return __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-x))); }

__device__ __forceinline__ void block_totals(uint8_t f, uint4* out) {
    const int n0 = __syncthreads_count(f & kKeepOrig);
    const int n1 = __syncthreads_count(f & kCloneKept);
    const int n2 = __syncthreads_count(f & kSplitSel);
    const int n3 = __syncthreads_count(f & kSplitKept);
    if (threadIdx.x == 0) out[blockIdx.x] = make_uint4(n0, n1, n2, n3);
}

// Decisions of densify_and_clone (:393-398), densify_and_split (:368-378, 390-391) and the final prune (:419-425),
// evaluated per original Gaussian. A clone carries its parent's parameters, so the final prune treats it like the
// parent; the N children of a split share scale and opacity, so they survive or fall together. max_radii2D has been
// reset to zeros by densification_postfix (:365) before the reference reads it at :422, so the screen-size test
// compares 0 with max_screen_size.
__global__ void __launch_bounds__(kDenThreads) densify_classify_kernel(int P, const float* __restrict__ opacity,
                                                                       const float* __restrict__ scaling,
                                                                       const float* __restrict__ grad_accum,
                                                                       const float* __restrict__ denom,
                                                                       const gsr_densify_rule r, DensifyPlan plan) {
    const int i = blockIdx.x * kDenThreads + threadIdx.x;
    uint8_t f = 0;
    if (i < P) {
        float g = __fdiv_rn(grad_accum[i], denom[i]);
        if (isnan(g)) g = 0.0f;
        const float e0 = expf(scaling[3 * i]), e1 = expf(scaling[3 * i + 1]), e2 = expf(scaling[3 * i + 2]);
        const float mx = fmaxf(fmaxf(e0, e1), e2);
        const bool nan_scale = isnan(e0) || isnan(e1) || isnan(e2);         // torch.max propagates NaN: every comparison false
        const bool clone_sel = !nan_scale && sqrtf(__fmul_rn(g, g)) >= r.grad_threshold && mx <= r.dense_extent;
        const bool split_sel = !nan_scale && g >= r.grad_threshold && mx > r.dense_extent;
        const bool faint = sigmoid_ref(opacity[i]) < r.min_opacity;
        const bool screen = r.prune_big && 0.0f > r.max_screen_size;
        const bool pruned = faint || (r.prune_big && (screen || (!nan_scale && mx > r.big_extent)));
        // children: scaling_inverse_activation(get_scaling / (0.8 N)) read back through get_scaling (:384, :423)
        const float c0 = expf(logf(__fmul_rn(e0, r.split_shrink_inv))), c1 = expf(logf(__fmul_rn(e1, r.split_shrink_inv))),
                    c2 = expf(logf(__fmul_rn(e2, r.split_shrink_inv)));
        const float cmx = fmaxf(fmaxf(c0, c1), c2);
        const bool child_pruned = faint || (r.prune_big && (screen || (!nan_scale && cmx > r.big_extent)));
        if (!split_sel && !pruned) f |= kKeepOrig;
        if (clone_sel && !pruned) f |= kCloneKept;
        if (split_sel) f |= kSplitSel;
        if (split_sel && !child_pruned) f |= kSplitKept;
        plan.flags[i] = f;
    }
    block_totals(f, plan.block_sums);
}

__global__ void __launch_bounds__(kDenThreads) prune_classify_kernel(int P, const uint8_t* __restrict__ prune_mask,
                                                                     DensifyPlan plan) {
    const int i = blockIdx.x * kDenThreads + threadIdx.x;
    uint8_t f = 0;
    if (i < P) {
        f = prune_mask[i] ? 0 : kKeepOrig;
        plan.flags[i] = f;
    }
    block_totals(f, plan.block_sums);
}

__device__ __forceinline__ uint4 add4(uint4 a, uint4 b) { return make_uint4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }

// exclusive scan of the per-block totals, in place; totals to counts[0..3]. One CTA.
__global__ void __launch_bounds__(1024) densify_spine_kernel(int m, DensifyPlan plan) {
    __shared__ uint4 warp_tot[32];
    const int per = (m + 1023) / 1024;
    const int lo = min(m, (int)threadIdx.x * per), hi = min(m, lo + per);
    uint4 mine = make_uint4(0, 0, 0, 0);
    for (int k = lo; k < hi; k++) mine = add4(mine, plan.block_sums[k]);
    uint4 inc = mine;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint4 o;
        o.x = __shfl_up_sync(~0u, inc.x, d); o.y = __shfl_up_sync(~0u, inc.y, d);
        o.z = __shfl_up_sync(~0u, inc.z, d); o.w = __shfl_up_sync(~0u, inc.w, d);
        if (lane >= d) inc = add4(inc, o);
    }
    if (lane == 31) warp_tot[w] = inc;
    __syncthreads();
    if (w == 0) {
        uint4 t = warp_tot[lane], ti = t;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint4 o;
            o.x = __shfl_up_sync(~0u, ti.x, d); o.y = __shfl_up_sync(~0u, ti.y, d);
            o.z = __shfl_up_sync(~0u, ti.z, d); o.w = __shfl_up_sync(~0u, ti.w, d);
            if (lane >= d) ti = add4(ti, o);
        }
        warp_tot[lane] = make_uint4(ti.x - t.x, ti.y - t.y, ti.z - t.z, ti.w - t.w);
        if (lane == 31) {
            plan.counts[0] = ti.x; plan.counts[1] = ti.y; plan.counts[2] = ti.z; plan.counts[3] = ti.w;
        }
    }
    __syncthreads();
    const uint4 wb = warp_tot[w];
    uint4 run = make_uint4(wb.x + inc.x - mine.x, wb.y + inc.y - mine.y, wb.z + inc.z - mine.z, wb.w + inc.w - mine.w);
    for (int k = lo; k < hi; k++) {
        const uint4 v = plan.block_sums[k];
        plan.block_sums[k] = run;
        run = add4(run, v);
    }
}

// rank of every flagged row inside its class -> destination rows
__global__ void __launch_bounds__(kDenThreads) densify_rank_kernel(int P, DensifyPlan plan) {
    __shared__ uint4 warp_cnt[kDenThreads / 32];
    const int i = blockIdx.x * kDenThreads + threadIdx.x;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint8_t f = i < P ? plan.flags[i] : 0;
    const uint32_t lt = (1u << lane) - 1;
    const uint32_t b0 = __ballot_sync(~0u, f & kKeepOrig), b1 = __ballot_sync(~0u, f & kCloneKept),
                   b2 = __ballot_sync(~0u, f & kSplitSel), b3 = __ballot_sync(~0u, f & kSplitKept);
    if (lane == 0) warp_cnt[w] = make_uint4(__popc(b0), __popc(b1), __popc(b2), __popc(b3));
    __syncthreads();
    uint4 base = plan.block_sums[blockIdx.x];
    for (int k = 0; k < w; k++) base = add4(base, warp_cnt[k]);
    if (i >= P) return;
    const uint32_t n_orig = plan.counts[0], n_clone = plan.counts[1];
    int4 d;
    d.x = (f & kKeepOrig) ? (int)(base.x + __popc(b0 & lt)) : -1;
    d.y = (f & kCloneKept) ? (int)(n_orig + base.y + __popc(b1 & lt)) : -1;
    d.z = (f & kSplitKept) ? (int)(n_orig + n_clone + base.w + __popc(b3 & lt)) : -1;
    d.w = (f & kSplitSel) ? (int)(base.z + __popc(b2 & lt)) : -1;
    plan.dest[i] = d;
}

struct DensifyTable {
    gsr_densify_tensor t[GSR_DENSIFY_MAX_TENSORS];
    unsigned block_begin[GSR_DENSIFY_MAX_TENSORS + 1];
    int n;
};

constexpr int kDenPerCta = kDenThreads * 8;    // source elements per CTA of the apply kernel

// One component of build_rotation(q) @ sample + xyz (gaussian_model.py:379-383, general_utils.py:78-99). Every
// product and sum of the reference is its own torch kernel, hence individually rounded: no contraction here either.
__device__ __forceinline__ float split_position(int c, const float* __restrict__ rot, const float* __restrict__ scaling,
                                                const float* __restrict__ z, float center) {
    const float q0 = rot[0], q1 = rot[1], q2 = rot[2], q3 = rot[3];
    const float n = sqrtf(__fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(q0, q0), __fmul_rn(q1, q1)), __fmul_rn(q2, q2)), __fmul_rn(q3, q3)));
    const float r = __fdiv_rn(q0, n), x = __fdiv_rn(q1, n), y = __fdiv_rn(q2, n), zz = __fdiv_rn(q3, n);
    float R0, R1, R2;
    auto diag = [](float a, float b) { return __fsub_rn(1.0f, __fmul_rn(2.0f, __fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b)))); };
    auto offp = [](float a, float b, float c2, float d) { return __fmul_rn(2.0f, __fadd_rn(__fmul_rn(a, b), __fmul_rn(c2, d))); };
    auto offm = [](float a, float b, float c2, float d) { return __fmul_rn(2.0f, __fsub_rn(__fmul_rn(a, b), __fmul_rn(c2, d))); };
    if (c == 0)      { R0 = diag(y, zz);        R1 = offm(x, y, r, zz);  R2 = offp(x, zz, r, y); }
    else if (c == 1) { R0 = offp(x, y, r, zz);  R1 = diag(x, zz);        R2 = offm(y, zz, r, x); }
    else             { R0 = offm(x, zz, r, y);  R1 = offp(y, zz, r, x);  R2 = diag(x, y); }
    // samples = normal(mean 0, std get_scaling) = z * std + 0
    const float s0 = __fadd_rn(__fmul_rn(z[0], expf(scaling[0])), 0.0f), s1 = __fadd_rn(__fmul_rn(z[1], expf(scaling[1])), 0.0f),
                s2 = __fadd_rn(__fmul_rn(z[2], expf(scaling[2])), 0.0f);
    const float dot = fmaf(R2, s2, fmaf(R1, s1, __fmul_rn(R0, s0)));      // torch.bmm: accumulation order is the library's
    return __fadd_rn(dot, center);
}

// exp_avg / exp_avg_sq of every appended row are zero (cat_tensors_to_optimizer, :331-332): the appended rows are the
// contiguous tail of the output. Persistent grid, float4 stores once aligned.
__global__ void __launch_bounds__(kDenThreads) densify_zero_kernel(const __grid_constant__ DensifyTable tab, DensifyPlan plan,
                                                                   int n_split) {
    const long long first = plan.counts[0];
    const long long rows = (long long)plan.counts[1] + (long long)n_split * plan.counts[3];
    for (int k = 0; k < tab.n; k++) {
        const gsr_densify_tensor& T = tab.t[k];
        if (!T.dst_m) continue;
        const long long n = rows * T.row_len;
        for (int which = 0; which < 2; which++) {
            float* p = (which ? T.dst_v : T.dst_m) + first * T.row_len;
            const long long head = min(n, (long long)((16 - (reinterpret_cast<uintptr_t>(p) & 15)) & 15) / 4);
            const long long vec = (n - head) / 4;
            float4* p4 = reinterpret_cast<float4*>(p + head);
            for (long long j = (long long)blockIdx.x * kDenThreads + threadIdx.x; j < vec; j += (long long)gridDim.x * kDenThreads)
                p4[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (blockIdx.x == 0) {
                for (long long j = threadIdx.x; j < head; j += kDenThreads) p[j] = 0.0f;
                for (long long j = head + vec * 4 + threadIdx.x; j < n; j += kDenThreads) p[j] = 0.0f;
            }
        }
    }
}

// One CTA moves kDenPerCta consecutive source elements of one tensor (and of its two moment tensors):
//   A. every thread fetches two quads of consecutive elements of src / exp_avg / exp_avg_sq with 16-byte loads -- 96 bytes
//      in flight per thread, unconditionally, so that nothing waits on the plan -- and parks them in shared memory; the
//      destination records of the rows the chunk touches are fetched alongside (coalesced 16-byte loads);
//   B. element j of the chunk is then handled by thread j mod 256, so that the stores of a warp are 32 consecutive
//      words of the destination row(s).
// Rows that go nowhere (pruned, or split parents) are read for nothing: a few per cent of the map in a training run.
// Moments of NEW rows are not written here: they are the contiguous tail [kept originals, P_new) of dst_m / dst_v,
// cleared by densify_zero_kernel.
constexpr int kDenRows = kDenPerCta / 3 + 3;        // rows a chunk can touch when row_len >= 3; narrower tensors read the plan directly

__global__ void __launch_bounds__(kDenThreads, 5) densify_apply_kernel(const __grid_constant__ DensifyTable tab, int P,
                                                                       DensifyPlan plan, int n_split, float shrink_inv,
                                                                       const float* __restrict__ scaling,
                                                                       const float* __restrict__ rotation,
                                                                       const float* __restrict__ noise) {
    __shared__ int4 s_dest[kDenRows];
    __shared__ __align__(16) float s_val[3][kDenPerCta];
    int ti = 0;
#pragma unroll
    for (int k = 1; k < GSR_DENSIFY_MAX_TENSORS; k++)
        if (k < tab.n && blockIdx.x >= tab.block_begin[k]) ti = k;
    const gsr_densify_tensor& T = tab.t[ti];
    const uint32_t w = (uint32_t)T.row_len;
    const long long total = (long long)P * w;
    const long long e0 = (long long)(blockIdx.x - tab.block_begin[ti]) * kDenPerCta;
    const int count = (int)min((long long)kDenPerCta, total - e0);
    const int row0 = (int)(e0 / w);
    const uint32_t col0 = (uint32_t)(e0 - (long long)row0 * w);
    const uint32_t magic = w > 1 ? (uint32_t)((1ull << 32) / w + 1) : 0;      // x / w = umulhi(x, magic) while x * w < 2^32
    auto row_of = [&](uint32_t x) { return w > 1 ? __umulhi(x, magic) : x; };
    const bool staged = w >= 3;
    const bool moments = T.src_m != nullptr;
    const bool vec_ok = ((reinterpret_cast<uintptr_t>(T.src) | reinterpret_cast<uintptr_t>(T.src_m) |
                          reinterpret_cast<uintptr_t>(T.src_v)) & 15) == 0;
#pragma unroll
    for (int half = 0; half < kDenPerCta / (4 * kDenThreads); half++) {
        const int j = 4 * (half * kDenThreads + threadIdx.x);
        if (j >= count) break;
        if (vec_ok && j + 3 < count) {
            *reinterpret_cast<float4*>(&s_val[0][j]) = __ldg(reinterpret_cast<const float4*>(T.src + e0 + j));
            if (moments) {
                *reinterpret_cast<float4*>(&s_val[1][j]) = __ldg(reinterpret_cast<const float4*>(T.src_m + e0 + j));
                *reinterpret_cast<float4*>(&s_val[2][j]) = __ldg(reinterpret_cast<const float4*>(T.src_v + e0 + j));
            }
        } else {
            for (int k = j; k < min(j + 4, count); k++) {
                s_val[0][k] = __ldg(T.src + e0 + k);
                if (moments) { s_val[1][k] = __ldg(T.src_m + e0 + k); s_val[2][k] = __ldg(T.src_v + e0 + k); }
            }
        }
    }
    if (staged) {
        const int n_rows = (int)row_of(col0 + (uint32_t)count - 1) + 1;
        for (int r = threadIdx.x; r < n_rows; r += kDenThreads) s_dest[r] = __ldg(&plan.dest[row0 + r]);
    }
    const uint32_t n_sel = plan.counts[2], n_kept = plan.counts[3];
    __syncthreads();

#pragma unroll 2
    for (int it = 0; it < kDenPerCta / kDenThreads; it++) {
        const int j = it * kDenThreads + threadIdx.x;
        if (j >= count) break;
        const uint32_t r = row_of(col0 + j), c = col0 + j - r * w;
        const int4 d = staged ? s_dest[r] : __ldg(&plan.dest[row0 + r]);
        if ((d.x & d.y & d.z) < 0) continue;                              // all three negative: the row goes nowhere
        const float val = s_val[0][j];
        if (d.x >= 0) {
            const long long o = (long long)d.x * w + c;
            T.dst[o] = val;
            if (moments) { T.dst_m[o] = s_val[1][j]; T.dst_v[o] = s_val[2][j]; }
        }
        if (d.y >= 0) T.dst[(long long)d.y * w + c] = val;
        if (d.z >= 0) {
            const int i = row0 + (int)r;
            float child = val;
            if (T.role == GSR_DENSIFY_SCALING) child = logf(__fmul_rn(expf(val), shrink_inv));
            for (int rep = 0; rep < n_split; rep++) {
                if (T.role == GSR_DENSIFY_XYZ)
                    child = split_position((int)c, rotation + 4ll * i, scaling + 3ll * i,
                                           noise + 3ll * ((long long)rep * n_sel + d.w), val);
                T.dst[((long long)d.z + (long long)rep * n_kept) * w + c] = child;
            }
        }
    }
}

__global__ void __launch_bounds__(kDenThreads) densify_stats_kernel(int P, const float* __restrict__ dL_dmean2D,
                                                                    const uint8_t* __restrict__ update_filter,
                                                                    const int* __restrict__ radii,
                                                                    float* __restrict__ grad_accum, float* __restrict__ denom,
                                                                    float* __restrict__ max_radii2D) {
    const int i = blockIdx.x * kDenThreads + threadIdx.x;
    if (i >= P) return;
    const bool on = update_filter ? update_filter[i] != 0 : radii[i] > 0;
    if (!on) return;
    const float gx = dL_dmean2D[3 * i], gy = dL_dmean2D[3 * i + 1];
    // torch.norm over two elements: acc = acc + v * v from zero, which the reduction kernel contracts to an fma
    const float nrm = sqrtf(fmaf(gy, gy, __fmul_rn(gx, gx)));
    grad_accum[i] = __fadd_rn(grad_accum[i], nrm);
    denom[i] = __fadd_rn(denom[i], 1.0f);
    if (max_radii2D && radii) max_radii2D[i] = fmaxf(max_radii2D[i], (float)radii[i]);
}

static int run_plan_tail(int P, DensifyPlan& plan, cudaStream_t s) {
    const int blocks = (P + kDenThreads - 1) / kDenThreads;
    GSR_LAUNCH("densify_spine", false, s, (densify_spine_kernel<<<1, 1024, 0, s>>>(blocks, plan)));
    if (blocks) GSR_LAUNCH("densify_rank", false, s, (densify_rank_kernel<<<blocks, kDenThreads, 0, s>>>(P, plan)));
    return 0;
}

int launch_densify_plan(int P, const float* opacity, const float* scaling, const float* grad_accum, const float* denom,
                        const gsr_densify_rule& rule, void* blob, uint32_t* counts_out, cudaStream_t s) {
    DensifyPlan plan = DensifyPlan::from(blob, (size_t)P);
    const int blocks = (P + kDenThreads - 1) / kDenThreads;
    if (blocks)
        GSR_LAUNCH("densify_classify", false, s, (densify_classify_kernel<<<blocks, kDenThreads, 0, s>>>(
                       P, opacity, scaling, grad_accum, denom, rule, plan)));
    int e = run_plan_tail(P, plan, s);
    if (e) return e;
    if (counts_out) {
        cudaError_t ce = cudaMemcpyAsync(counts_out, plan.counts, 4 * sizeof(uint32_t), cudaMemcpyDefault, s);
        if (ce != cudaSuccess) return (int)ce;
    }
    return 0;
}

int launch_prune_plan(int P, const uint8_t* prune_mask, void* blob, uint32_t* counts_out, cudaStream_t s) {
    DensifyPlan plan = DensifyPlan::from(blob, (size_t)P);
    const int blocks = (P + kDenThreads - 1) / kDenThreads;
    if (blocks)
        GSR_LAUNCH("prune_classify", false, s, (prune_classify_kernel<<<blocks, kDenThreads, 0, s>>>(P, prune_mask, plan)));
    int e = run_plan_tail(P, plan, s);
    if (e) return e;
    if (counts_out) {
        cudaError_t ce = cudaMemcpyAsync(counts_out, plan.counts, 4 * sizeof(uint32_t), cudaMemcpyDefault, s);
        if (ce != cudaSuccess) return (int)ce;
    }
    return 0;
}

int launch_densify_apply(int P, const void* blob, const gsr_densify_rule* rule, const gsr_densify_tensor* tensors, int n,
                         const float* scaling, const float* rotation, const float* noise, cudaStream_t s) {
    DensifyPlan plan = DensifyPlan::from(const_cast<void*>(blob), (size_t)P);
    DensifyTable tab;
    tab.n = n;
    unsigned blocks = 0;
    for (int k = 0; k < n; k++) {
        tab.t[k] = tensors[k];
        tab.block_begin[k] = blocks;
        blocks += (unsigned)(((long long)P * tensors[k].row_len + kDenPerCta - 1) / kDenPerCta);
    }
    tab.block_begin[n] = blocks;
    if (blocks == 0) return 0;
    // zero moments for every new row: rows [counts[0], P_new) of dst_m / dst_v. The row counts live on the device, so
    // the clear is a kernel over the worst case rather than a sized memset.
    GSR_LAUNCH("densify_zero_moments", false, s, (densify_zero_kernel<<<148 * 8, kDenThreads, 0, s>>>(tab, plan, rule ? rule->n_split : 0)));
    GSR_LAUNCH("densify_apply", false, s, (densify_apply_kernel<<<blocks, kDenThreads, 0, s>>>(
                   tab, P, plan, rule ? rule->n_split : 0, rule ? rule->split_shrink_inv : 1.0f, scaling, rotation, noise)));
    return 0;
}

int launch_densify_stats(int P, const float* dL_dmean2D, const uint8_t* update_filter, const int* radii, float* grad_accum,
                         float* denom, float* max_radii2D, cudaStream_t s) {
    if (P == 0) return 0;
    GSR_LAUNCH("densify_stats", false, s, (densify_stats_kernel<<<(P + kDenThreads - 1) / kDenThreads, kDenThreads, 0, s>>>(
                   P, dL_dmean2D, update_filter, radii, grad_accum, denom, max_radii2D)));
    return 0;
}

}  // namespace gsr
