// gsr_comm.cu -- the exchange step of view-parallel training: the sum all-reduce of the per-Gaussian
// gradient bucket (59 floats = 236 B per Gaussian at SH degree 3; 708 MB at 3M), written directly against NVLink peer
// memory instead of a library collective. The reference is single-GPU (train.py never leaves cuda:0), so there is no
// reference function to replace; the contract is sum over ranks of the flat fp32 bucket, identical bits on all ranks.
//
// ONE launch per rank does the whole exchange:
//   barrier   every CTA meets its twin (same blockIdx) on every peer through a signal pad in peer memory, so
//             the peers' buckets are complete (their backward kernels precede this launch in stream order);
//   reduce    rank r owns the r-th 1/world slice of the bucket. With an NVSwitch multicast mapping one
//             multimem.ld_reduce pulls the slice with the addition done inside the switch and one multimem.st
//             pushes the sum into every replica (each byte crosses each link once per direction). Without
//             multicast, the slice is read from every peer's replica with 16-byte system-scope loads, added in
//             rank order (so every rank holds the same bits) and written to every replica with peer stores;
//   barrier   the same twins meet again once their stores are released, so when the kernel ends on a rank, all
//             of its bucket has arrived.
// No staging buffer, no copy kernels: the autograd .grad tensors ARE the peer-mapped bucket (view_parallel.py).
// Bound: NVLink -- per rank (world-1)/world of the bucket leaves and enters through the switch.
#include "gsr_common.cuh"
#include <cuda/atomic>
#include <algorithm>

namespace gsr {

constexpr int kCommThreads = 512;
constexpr int kCommUnroll = 4;          // 16-byte requests in flight per thread and peer (peer variant)
constexpr int kCommUnrollMc = 4;        // multicast variant. Measured at 8 ranks (tools/comm_sweep.py, profiles/r02_comm_sweep_n8.json):
                                        // 4 requests in flight on sms/2 CTAs 1.63 ms, 8 on 2*sms CTAs 1.70 ms, 16: 1.71-1.81 ms.
                                        // One stream alone (ld_reduce only / multimem.st only) moves the 708 MB in 1.06 ms
                                        // (665 GB/s per direction); together they reach 490 GB/s whether they alternate in
                                        // a thread or run in separate producer / consumer CTAs (variant 3): the fabric, not
                                        // the SMs' issue order, is what they share.

struct CommArgs {
    float4* buf[GSR_COMM_MAX_PEERS];
    uint32_t* sig[GSR_COMM_MAX_PEERS];
    float4* mc;
    int rank, world;
    long long lo, hi;                   // this rank's slice in 16-byte units
};

// signal pad protocol: slot (cta, source rank) on the destination's pad goes 0 -> 1 by the source, 1 -> 0 by the owner.
// The spins are BOUNDED: a peer that never arrives (crashed process, a rank whose grid is not fully resident) would
// otherwise hang every rank for ever. After ~2^26 polls with back-off (tens of seconds) the CTA sets the error word of
// its own pad (slot GSR_COMM_MAX_CTAS * GSR_COMM_MAX_PEERS) and traps; the host sees the launch failure on its next
// synchronisation instead of a silent hang.
constexpr uint32_t kSpinLimit = 1u << 26;
__device__ __forceinline__ void sig_fail(uint32_t* pad) {
    atomicExch(pad + (size_t)GSR_COMM_MAX_CTAS * GSR_COMM_MAX_PEERS, 1u);
    __threadfence_system();
    __trap();
}
__device__ __forceinline__ void sig_put(uint32_t* slot, uint32_t* my_pad) {
    cuda::atomic_ref<uint32_t, cuda::thread_scope_system> ref(*slot);
    uint32_t expect = 0, spins = 0;
    while (!ref.compare_exchange_strong(expect, 1u, cuda::std::memory_order_release, cuda::std::memory_order_relaxed)) {
        expect = 0;
        if (++spins > 1024u) __nanosleep(256);
        if (spins > kSpinLimit) sig_fail(my_pad);
    }
}
__device__ __forceinline__ void sig_wait(uint32_t* slot, uint32_t* my_pad) {
    cuda::atomic_ref<uint32_t, cuda::thread_scope_system> ref(*slot);
    uint32_t expect = 1, spins = 0;
    while (!ref.compare_exchange_strong(expect, 0u, cuda::std::memory_order_acquire, cuda::std::memory_order_relaxed)) {
        expect = 1;
        if (++spins > 1024u) __nanosleep(256);
        if (spins > kSpinLimit) sig_fail(my_pad);
    }
}
// cta = index of this CTA among the rank's CTAs (blockIdx.x of a real launch)
__device__ __forceinline__ void twin_barrier(const CommArgs& a, int cta) {
    __threadfence_system();
    __syncthreads();
    if ((int)threadIdx.x < a.world) {
        const int peer = threadIdx.x;
        sig_put(a.sig[peer] + (size_t)cta * GSR_COMM_MAX_PEERS + a.rank, a.sig[a.rank]);
        sig_wait(a.sig[a.rank] + (size_t)cta * GSR_COMM_MAX_PEERS + peer, a.sig[a.rank]);
    }
    __syncthreads();
}

__device__ __forceinline__ float4 mc_ld_reduce(const float4* p) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void mc_st(float4* p, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 peer_ld(const float4* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void peer_st(float4* p, float4 v) {
    asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

template <bool MULTICAST, int WORLD>
__device__ __forceinline__ void allreduce_sum_body(const CommArgs& a, int cta, int n_ctas) {
    twin_barrier(a, cta);
    const long long stride = (long long)n_ctas * kCommThreads;
    const int world = WORLD > 0 ? WORLD : a.world;
    constexpr int U = MULTICAST ? kCommUnrollMc : kCommUnroll;
    for (long long i0 = a.lo + (long long)cta * kCommThreads + threadIdx.x; i0 < a.hi; i0 += stride * U) {
        float4 v[U];
        if (MULTICAST) {
#pragma unroll
            for (int u = 0; u < U; u++)
                if (i0 + u * stride < a.hi) v[u] = mc_ld_reduce(a.mc + i0 + u * stride);
#pragma unroll
            for (int u = 0; u < U; u++)
                if (i0 + u * stride < a.hi) mc_st(a.mc + i0 + u * stride, v[u]);
        } else {
            float4 w[kCommUnroll];
#pragma unroll
            for (int u = 0; u < kCommUnroll; u++)
                if (i0 + u * stride < a.hi) v[u] = peer_ld(a.buf[0] + i0 + u * stride);
#pragma unroll
            for (int p = 1; p < (WORLD > 0 ? WORLD : GSR_COMM_MAX_PEERS); p++) {
                if (p < world) {
#pragma unroll
                    for (int u = 0; u < kCommUnroll; u++)
                        if (i0 + u * stride < a.hi) w[u] = peer_ld(a.buf[p] + i0 + u * stride);
#pragma unroll
                    for (int u = 0; u < kCommUnroll; u++) {      // rank order: the same bits on every rank
                        v[u].x = __fadd_rn(v[u].x, w[u].x); v[u].y = __fadd_rn(v[u].y, w[u].y);
                        v[u].z = __fadd_rn(v[u].z, w[u].z); v[u].w = __fadd_rn(v[u].w, w[u].w);
                    }
                }
            }
#pragma unroll
            for (int p = 0; p < (WORLD > 0 ? WORLD : GSR_COMM_MAX_PEERS); p++) {
                if (p < world) {
#pragma unroll
                    for (int u = 0; u < kCommUnroll; u++)
                        if (i0 + u * stride < a.hi) peer_st(a.buf[p] + i0 + u * stride, v[u]);
                }
            }
        }
    }
    twin_barrier(a, cta);
}

template <bool MULTICAST, int WORLD>
__global__ void __launch_bounds__(kCommThreads, 2) allreduce_sum_kernel(const __grid_constant__ CommArgs a) {
    allreduce_sum_body<MULTICAST, WORLD>(a, (int)blockIdx.x, (int)gridDim.x);
}
// ---- multicast variants (selected with GSR_COMM_VARIANT, measured with tools/comm_sweep.py) ---------------------------
// SPLIT: the rank's CTAs are pairs (producer c, consumer c + n/2) that share the c-th share of the slice. The producer
// streams multimem.ld_reduce and parks the sums in the rank's OWN replica (plain stores: they stay in L2); the consumer
// follows chunk by chunk behind a counter and streams multimem.st. Each link direction then carries one uninterrupted
// request stream (reduce: data leaves every GPU, broadcast: data enters every GPU) instead of alternating bursts.
constexpr int kSplitUnroll = 8;                                   // 512 threads x 8 x 16 B = 64 KB per chunk
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ uint32_t g_split_ready[GSR_COMM_MAX_CTAS];           // chunks parked per pair; zero between launches
__global__ void __launch_bounds__(kCommThreads, 2) allreduce_sum_split_kernel(const __grid_constant__ CommArgs a) {
    uint32_t* ready = g_split_ready;
    const int cta = (int)blockIdx.x, half = (int)gridDim.x / 2;
    twin_barrier(a, cta);
    const bool producer = cta < half;
    const int pair = producer ? cta : cta - half;
    float4* own = a.buf[a.rank];
    const long long chunk = (long long)kCommThreads * kSplitUnroll;
    uint32_t k = 0;
    for (long long base = a.lo + (long long)pair * chunk; base < a.hi; base += (long long)half * chunk, k++) {
        const long long i0 = base + threadIdx.x;
        if (producer) {
            float4 v[kSplitUnroll];
#pragma unroll
            for (int u = 0; u < kSplitUnroll; u++)
                if (i0 + u * kCommThreads < a.hi) v[u] = mc_ld_reduce(a.mc + i0 + u * kCommThreads);
#pragma unroll
            for (int u = 0; u < kSplitUnroll; u++)
                if (i0 + u * kCommThreads < a.hi) own[i0 + u * kCommThreads] = v[u];
            __threadfence();
            __syncthreads();
            if (threadIdx.x == 0) atomicAdd(ready + pair, 1u);
        } else {
            if (threadIdx.x == 0) {
                uint32_t spins = 0;
                while (ld_acquire_gpu(ready + pair) <= k) {
                    if (++spins > 64u) __nanosleep(64);
                    if (spins > kSpinLimit) sig_fail(a.sig[a.rank]);
                }
            }
            __syncthreads();
            float4 v[kSplitUnroll];
#pragma unroll
            for (int u = 0; u < kSplitUnroll; u++)
                if (i0 + u * kCommThreads < a.hi) v[u] = __ldcg(own + i0 + u * kCommThreads);
#pragma unroll
            for (int u = 0; u < kSplitUnroll; u++)
                if (i0 + u * kCommThreads < a.hi) mc_st(a.mc + i0 + u * kCommThreads, v[u]);
        }
    }
    if (!producer) {                                               // the producer wrote its last chunk: rearm the counter
        __syncthreads();
        if (threadIdx.x == 0) ready[pair] = 0u;
    }
    twin_barrier(a, cta);
}
// PROBES (not an all-reduce: they measure what one request stream alone reaches; the bucket's contents are garbage
// afterwards). mode 1: multimem.ld_reduce only, mode 2: multimem.st only.
template <int MODE, int U>
__global__ void __launch_bounds__(kCommThreads, 2) comm_probe_kernel(const __grid_constant__ CommArgs a) {
    const int cta = (int)blockIdx.x;
    twin_barrier(a, cta);
    const long long stride = (long long)gridDim.x * kCommThreads;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (long long i0 = a.lo + (long long)cta * kCommThreads + threadIdx.x; i0 < a.hi; i0 += stride * U) {
        if (MODE == 1) {
            float4 v[U];
#pragma unroll
            for (int u = 0; u < U; u++)
                if (i0 + u * stride < a.hi) v[u] = mc_ld_reduce(a.mc + i0 + u * stride);
#pragma unroll
            for (int u = 0; u < U; u++)
                if (i0 + u * stride < a.hi) { acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w; }
        } else {
#pragma unroll
            for (int u = 0; u < U; u++)
                if (i0 + u * stride < a.hi) mc_st(a.mc + i0 + u * stride, make_float4((float)u, 1.f, 2.f, 3.f));
        }
    }
    if (MODE == 1 && acc.x + acc.y + acc.z + acc.w == 123.456f) a.buf[a.rank][a.lo] = acc;      // keeps the loads alive
    twin_barrier(a, cta);
}
template <int U>
__global__ void __launch_bounds__(kCommThreads, 2) allreduce_sum_mc_unroll_kernel(const __grid_constant__ CommArgs a) {
    const int cta = (int)blockIdx.x;
    twin_barrier(a, cta);
    const long long stride = (long long)gridDim.x * kCommThreads;
    for (long long i0 = a.lo + (long long)cta * kCommThreads + threadIdx.x; i0 < a.hi; i0 += stride * U) {
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; u++)
            if (i0 + u * stride < a.hi) v[u] = mc_ld_reduce(a.mc + i0 + u * stride);
#pragma unroll
        for (int u = 0; u < U; u++)
            if (i0 + u * stride < a.hi) mc_st(a.mc + i0 + u * stride, v[u]);
    }
    twin_barrier(a, cta);
}

// Emulated ranks on ONE device (tests without a second GPU): a single cooperative launch whose blockIdx.y is the rank.
// Separate launches that wait for one another are not guaranteed to run at the same time; one cooperative grid is.
__global__ void __launch_bounds__(kCommThreads, 2) allreduce_sum_emulated_kernel(const CommArgs* __restrict__ all) {
    allreduce_sum_body<false, 0>(all[blockIdx.y], (int)blockIdx.x, (int)gridDim.x);
}

// ---------------------------------------------------------------------------------------------
// Exchange + Adam in one launch (gsr_exchange_adam_step): the reduce-scatter half of the exchange feeds the optimiser
// directly and the all-gather half carries PARAMETERS instead of gradients. Rank r is the only one that touches slice r
// of the moments, so they are stored (and streamed) once per box instead of once per rank.
struct ExchangeAdamArgs {
    CommArgs g;                                  // gradient buckets + signal pads + this rank's slice
    float4* pbuf[GSR_COMM_MAX_PEERS];            // parameter buckets, same layout
    float4* pmc;
    float4* m;                                   // this rank's slice of exp_avg / exp_avg_sq: element i at [i - lo]
    float4* v;
    long long sb[GSR_ADAM_MAX_SPANS], se[GSR_ADAM_MAX_SPANS];      // parameter spans in 16-byte units
    float step_size[GSR_ADAM_MAX_SPANS], inv_sqrt_bc2[GSR_ADAM_MAX_SPANS];
    int n_spans;
    AdamHyper h;
};

template <bool MULTICAST, int WORLD>
__device__ __forceinline__ void exchange_adam_body(const ExchangeAdamArgs& a, int cta, int n_ctas) {
    twin_barrier(a.g, cta);
    const long long stride = (long long)n_ctas * kCommThreads;
    const int world = WORLD > 0 ? WORLD : a.g.world;
    constexpr int U = 2;
    for (long long i0 = a.g.lo + (long long)cta * kCommThreads + threadIdx.x; i0 < a.g.hi; i0 += stride * U) {
        float4 g[U], p[U], m[U], v[U];
        int span[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const long long i = i0 + u * stride;
            span[u] = -1;
            if (i < a.g.hi) {
#pragma unroll
                for (int k = 0; k < GSR_ADAM_MAX_SPANS; k++)
                    if (k < a.n_spans && i >= a.sb[k] && i < a.se[k]) span[u] = k;
            }
            if (span[u] >= 0) {
                if (MULTICAST) g[u] = mc_ld_reduce(a.g.mc + i);
                else g[u] = peer_ld(a.g.buf[0] + i);
                p[u] = a.pbuf[a.g.rank][i];
                m[u] = a.m[i - a.g.lo];
                v[u] = a.v[i - a.g.lo];
            }
        }
        if (!MULTICAST) {
#pragma unroll
            for (int q = 1; q < (WORLD > 0 ? WORLD : GSR_COMM_MAX_PEERS); q++) {
                if (q < world) {
                    float4 w[U];
#pragma unroll
                    for (int u = 0; u < U; u++)
                        if (span[u] >= 0) w[u] = peer_ld(a.g.buf[q] + i0 + u * stride);
#pragma unroll
                    for (int u = 0; u < U; u++) {            // rank order, as gsr_allreduce_sum
                        if (span[u] >= 0) {
                            g[u].x = __fadd_rn(g[u].x, w[u].x); g[u].y = __fadd_rn(g[u].y, w[u].y);
                            g[u].z = __fadd_rn(g[u].z, w[u].z); g[u].w = __fadd_rn(g[u].w, w[u].w);
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (span[u] < 0) continue;
            const long long i = i0 + u * stride;
            const float ss = a.step_size[span[u]], ib = a.inv_sqrt_bc2[span[u]];
            adam_one(p[u].x, g[u].x, m[u].x, v[u].x, a.h, ss, ib);
            adam_one(p[u].y, g[u].y, m[u].y, v[u].y, a.h, ss, ib);
            adam_one(p[u].z, g[u].z, m[u].z, v[u].z, a.h, ss, ib);
            adam_one(p[u].w, g[u].w, m[u].w, v[u].w, a.h, ss, ib);
            a.m[i - a.g.lo] = m[u];
            a.v[i - a.g.lo] = v[u];
            if (MULTICAST) mc_st(a.pmc + i, p[u]);
            else {
#pragma unroll
                for (int q = 0; q < (WORLD > 0 ? WORLD : GSR_COMM_MAX_PEERS); q++)
                    if (q < world) peer_st(a.pbuf[q] + i, p[u]);
            }
        }
    }
    twin_barrier(a.g, cta);
}

template <bool MULTICAST, int WORLD>
__global__ void __launch_bounds__(kCommThreads, 2) exchange_adam_kernel(const __grid_constant__ ExchangeAdamArgs a) {
    exchange_adam_body<MULTICAST, WORLD>(a, (int)blockIdx.x, (int)gridDim.x);
}
__global__ void __launch_bounds__(kCommThreads, 2) exchange_adam_emulated_kernel(const ExchangeAdamArgs* __restrict__ all) {
    exchange_adam_body<false, 0>(all[blockIdx.y], (int)blockIdx.x, (int)gridDim.x);
}

static int g_comm_variant = 0;           // gsr_debug_set_comm_variant: experiments of the multicast path, 0 = product kernel
int set_comm_variant(int v) { g_comm_variant = v; return 0; }

void comm_slice(long long n16, int rank, int world, long long* lo, long long* hi) {
    const long long per = (n16 + world - 1) / world;
    const long long l = per * rank < n16 ? per * rank : n16;
    const long long h = l + per < n16 ? l + per : n16;
    *lo = l;
    *hi = h;
}

static void fill_comm_args(const gsr_peer_group& g, long long numel, CommArgs& a) {
    for (int p = 0; p < GSR_COMM_MAX_PEERS; p++) {
        a.buf[p] = p < g.world ? reinterpret_cast<float4*>(g.buf[p]) : nullptr;
        a.sig[p] = p < g.world ? g.signal[p] : nullptr;
    }
    a.mc = reinterpret_cast<float4*>(g.multicast);
    a.rank = g.rank;
    a.world = g.world;
    comm_slice(numel / 4, g.rank, g.world, &a.lo, &a.hi);
}
// the grid must be the same on every rank (CTAs meet their twins): derive it from the largest slice only
static int comm_grid(long long numel, int world, int max_ctas, int per_cta_units, unsigned* grid, bool multicast = false) {
    int sms = 0;
    { const int e = sm_count(&sms); if (e) return e; }
    const long long n16 = numel / 4;
    const long long per = (n16 + world - 1) / world;
    long long ctas = (per + (long long)kCommThreads * per_cta_units - 1) / ((long long)kCommThreads * per_cta_units);
    // multicast requests are served by the switch: half an SM count of CTAs already saturates the links (and leaves the
    // other SMs to whatever else the stream's neighbours run); the peer path needs every SM's load/store slots
    long long cap = max_ctas > 0 ? max_ctas : (multicast ? std::max(1, sms / 2) : 2 * sms);
    if (cap > GSR_COMM_MAX_CTAS) cap = GSR_COMM_MAX_CTAS;
    if (ctas > cap) ctas = cap;
    if (ctas < 1) ctas = 1;
    *grid = (unsigned)ctas;
    return 0;
}

int launch_allreduce_sum(const gsr_peer_group& g, long long numel, int max_ctas, cudaStream_t s) {
    CommArgs a;
    fill_comm_args(g, numel, a);
    unsigned grid = 1;
    { const int e = comm_grid(numel, g.world, max_ctas, kCommUnroll, &grid, a.mc != nullptr); if (e) return e; }
    const int variant = g_comm_variant;
    if (a.mc != nullptr && variant != 0) {
        // experiments (tools/comm_sweep.py): 1 / 2 = probes (NOT an all-reduce), 3 = producer/consumer split,
        // 4 / 16 = unroll of the fused loop
        switch (variant) {
            case 1: GSR_LAUNCH("comm_probe_ld", false, s, (comm_probe_kernel<1, 8><<<grid, kCommThreads, 0, s>>>(a))); return 0;
            case 2: GSR_LAUNCH("comm_probe_st", false, s, (comm_probe_kernel<2, 8><<<grid, kCommThreads, 0, s>>>(a))); return 0;
            case 3:
                if (grid < 2) break;
                GSR_LAUNCH("allreduce_sum_split", false, s, (allreduce_sum_split_kernel<<<grid & ~1u, kCommThreads, 0, s>>>(a)));
                return 0;
            case 4: GSR_LAUNCH("allreduce_sum_mc_u4", false, s, (allreduce_sum_mc_unroll_kernel<4><<<grid, kCommThreads, 0, s>>>(a))); return 0;
            case 16: GSR_LAUNCH("allreduce_sum_mc_u16", false, s, (allreduce_sum_mc_unroll_kernel<16><<<grid, kCommThreads, 0, s>>>(a))); return 0;
            default: break;
        }
    }
    if (a.mc != nullptr)
        GSR_LAUNCH("allreduce_sum_multimem", false, s, (allreduce_sum_kernel<true, 0><<<grid, kCommThreads, 0, s>>>(a)));
    else if (g.world == 2)
        GSR_LAUNCH("allreduce_sum_peer", false, s, (allreduce_sum_kernel<false, 2><<<grid, kCommThreads, 0, s>>>(a)));
    else if (g.world == 4)
        GSR_LAUNCH("allreduce_sum_peer", false, s, (allreduce_sum_kernel<false, 4><<<grid, kCommThreads, 0, s>>>(a)));
    else if (g.world == 8)
        GSR_LAUNCH("allreduce_sum_peer", false, s, (allreduce_sum_kernel<false, 8><<<grid, kCommThreads, 0, s>>>(a)));
    else
        GSR_LAUNCH("allreduce_sum_peer", false, s, (allreduce_sum_kernel<false, 0><<<grid, kCommThreads, 0, s>>>(a)));
    return 0;
}

// All ranks of an emulated group in ONE cooperative launch on the current device (blockIdx.y = rank).
size_t comm_emulation_scratch_bytes() { return GSR_COMM_MAX_PEERS * (sizeof(ExchangeAdamArgs) > sizeof(CommArgs) ? sizeof(ExchangeAdamArgs) : sizeof(CommArgs)); }
static int coop_launch(const void* kernel, unsigned grid_x, int world, void* args_dev, cudaStream_t s) {
    int dev = 0, coop = 0, sms = 0;
    { const int e = current_device(&dev); if (e) return e; }
    cudaError_t ce = cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    if (ce != cudaSuccess) return (int)ce;
    { const int e = sm_count(&sms); if (e) return e; }
    if (!coop || (long long)grid_x * world > 2ll * sms) return GSR_E_BADARG;     // every CTA of every rank must be resident
    void* params[] = {&args_dev};
    ce = cudaLaunchCooperativeKernel(kernel, dim3(grid_x, (unsigned)world, 1), dim3(kCommThreads, 1, 1), params, 0, s);
    return (int)ce;
}
int launch_allreduce_sum_emulated(const gsr_peer_group* groups, int world, long long numel, int max_ctas, void* scratch_dev,
                                  cudaStream_t s) {
    CommArgs host[GSR_COMM_MAX_PEERS];
    for (int r = 0; r < world; r++) {
        fill_comm_args(groups[r], numel, host[r]);
        host[r].mc = nullptr;                                       // one device has no multicast group
    }
    unsigned grid = 1;
    { const int e = comm_grid(numel, world, max_ctas, kCommUnroll, &grid); if (e) return e; }
    cudaError_t ce = cudaMemcpyAsync(scratch_dev, host, sizeof(CommArgs) * world, cudaMemcpyHostToDevice, s);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);           // `host` lives on this stack frame
    if (ce != cudaSuccess) return (int)ce;
    int le = 0;
    GSR_LAUNCH("allreduce_sum_emulated", false, s, (le = coop_launch((const void*)allreduce_sum_emulated_kernel, grid, world, scratch_dev, s)));
    return le;
}

static void fill_exchange_args(const gsr_peer_group& grads, const gsr_peer_group& params, float* exp_avg, float* exp_avg_sq,
                               const gsr_adam_span* spans, int n_spans, double b1, double b2, double eps, long long numel,
                               ExchangeAdamArgs& a) {
    fill_comm_args(grads, numel, a.g);
    for (int p = 0; p < GSR_COMM_MAX_PEERS; p++) a.pbuf[p] = p < params.world ? reinterpret_cast<float4*>(params.buf[p]) : nullptr;
    a.pmc = reinterpret_cast<float4*>(params.multicast);
    a.m = reinterpret_cast<float4*>(exp_avg);
    a.v = reinterpret_cast<float4*>(exp_avg_sq);
    a.n_spans = n_spans;
    for (int k = 0; k < GSR_ADAM_MAX_SPANS; k++) {
        a.sb[k] = k < n_spans ? spans[k].begin / 4 : 0;
        a.se[k] = k < n_spans ? (spans[k].end + 3) / 4 : 0;
        a.step_size[k] = k < n_spans ? spans[k].step_size : 0.f;
        a.inv_sqrt_bc2[k] = k < n_spans ? spans[k].inv_sqrt_bc2 : 0.f;
    }
    a.h = AdamHyper{(float)(1.0 - b1), (float)b2, (float)(1.0 - b2), (float)eps};
}

int launch_exchange_adam(const gsr_peer_group& grads, const gsr_peer_group& params, float* exp_avg, float* exp_avg_sq,
                         const gsr_adam_span* spans, int n_spans, double b1, double b2, double eps, long long numel, int max_ctas,
                         cudaStream_t s) {
    ExchangeAdamArgs a;
    fill_exchange_args(grads, params, exp_avg, exp_avg_sq, spans, n_spans, b1, b2, eps, numel, a);
    unsigned grid = 1;
    { const int e = comm_grid(numel, grads.world, max_ctas, 2, &grid); if (e) return e; }
    if (a.g.mc != nullptr)
        GSR_LAUNCH("exchange_adam_multimem", false, s, (exchange_adam_kernel<true, 0><<<grid, kCommThreads, 0, s>>>(a)));
    else if (grads.world == 2)
        GSR_LAUNCH("exchange_adam_peer", false, s, (exchange_adam_kernel<false, 2><<<grid, kCommThreads, 0, s>>>(a)));
    else if (grads.world == 4)
        GSR_LAUNCH("exchange_adam_peer", false, s, (exchange_adam_kernel<false, 4><<<grid, kCommThreads, 0, s>>>(a)));
    else if (grads.world == 8)
        GSR_LAUNCH("exchange_adam_peer", false, s, (exchange_adam_kernel<false, 8><<<grid, kCommThreads, 0, s>>>(a)));
    else
        GSR_LAUNCH("exchange_adam_peer", false, s, (exchange_adam_kernel<false, 0><<<grid, kCommThreads, 0, s>>>(a)));
    return 0;
}

int launch_exchange_adam_emulated(const gsr_peer_group* grads, const gsr_peer_group* params, float* const* exp_avg,
                                  float* const* exp_avg_sq, int world, const gsr_adam_span* spans, int n_spans, double b1, double b2,
                                  double eps, long long numel, int max_ctas, void* scratch_dev, cudaStream_t s) {
    static ExchangeAdamArgs host[GSR_COMM_MAX_PEERS];               // too large for the stack of a small thread
    for (int r = 0; r < world; r++) {
        fill_exchange_args(grads[r], params[r], exp_avg[r], exp_avg_sq[r], spans, n_spans, b1, b2, eps, numel, host[r]);
        host[r].g.mc = nullptr;
        host[r].pmc = nullptr;
    }
    unsigned grid = 1;
    { const int e = comm_grid(numel, world, max_ctas, 2, &grid); if (e) return e; }
    cudaError_t ce = cudaMemcpyAsync(scratch_dev, host, sizeof(ExchangeAdamArgs) * world, cudaMemcpyHostToDevice, s);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(s);
    if (ce != cudaSuccess) return (int)ce;
    int le = 0;
    GSR_LAUNCH("exchange_adam_emulated", false, s, (le = coop_launch((const void*)exchange_adam_emulated_kernel, grid, world, scratch_dev, s)));
    return le;
}

}  // namespace gsr
