"""GPU tests (-m gpu) of the peer-memory gradient exchange (csrc/gsr_comm.cu) on ONE device: the "ranks"
are buffers and signal pads on the same GPU, and ALL of them run in ONE cooperative launch (blockIdx.y = rank,
gsr_allreduce_sum_emulated / gsr_exchange_adam_step_emulated: the same device code as the per-rank kernels). Separate
launches that wait for one another are not guaranteed to be resident together, so they are never used. The twin
barriers, the slice ownership and the rank-ordered summation are exercised exactly as across NVLink (the multicast
variant needs a real NVSwitch group: tests/test_gpu_multirank.py and tools/comm_check.py under torchrun cover it).
Oracle: the sum in rank order, computed by torch on the same device, bit for bit."""
import ctypes as C
import math

import pytest
import torch

from view_parallel import _PeerGroup, comm_slice, emulated_all_reduce, emulated_exchange_adam, peer_all_reduce, _gsr

pytestmark = pytest.mark.gpu


def _group(bufs, pads, rank):
    g = _PeerGroup()
    g.rank, g.world = rank, len(bufs)
    for p, (b, s) in enumerate(zip(bufs, pads)):
        g.buf[p], g.signal[p] = b.data_ptr(), s.data_ptr()
    g.multicast = None
    return g


@pytest.mark.parametrize("world,numel,max_ctas", [(1, 4096, 0), (2, 4, 8), (2, 1_000_000, 32), (3, 65_540, 16), (4, 3_000_004, 24),
                                                  (8, 262_144, 12)])
@pytest.mark.timeout(120)
def test_emulated_ranks_on_one_device_sum_in_rank_order(world, numel, max_ctas):
    L, _ = _gsr()
    dev = torch.device("cuda:0")
    gen = torch.Generator(device=dev).manual_seed(world * 1000 + numel % 997)
    bufs = [torch.randn(numel, device=dev, generator=gen) * (10.0 ** (r % 3)) for r in range(world)]
    pads = [torch.zeros(L.gsr_comm_signal_bytes() // 4, dtype=torch.int32, device=dev) for _ in range(world)]
    torch.cuda.synchronize()
    for rep in range(3):                                        # the pads reset themselves: the launch is repeatable
        want = bufs[0].clone()
        for r in range(1, world):
            want = want + bufs[r]                               # rank order, fp32, round to nearest
        if world == 1:
            peer_all_reduce(_group(bufs, pads, 0), numel, max_ctas)       # a single rank waits for nobody
        else:
            # one cooperative launch for all ranks: max_ctas x world stays far below 2 x 148 resident CTAs
            emulated_all_reduce([_group(bufs, pads, r) for r in range(world)], numel, max_ctas)
        torch.cuda.synchronize()
        for r in range(world):
            assert torch.equal(bufs[r], want), f"rank {r} rep {rep}: max diff {(bufs[r] - want).abs().max().item()}"
        assert all(int(p.abs().sum()) == 0 for p in pads)       # every signal consumed, no error word set


def test_slices_match_the_kernel_ownership():
    assert comm_slice(36, 0, 2) == (0, 20) and comm_slice(36, 1, 2) == (20, 36)
    assert comm_slice(4, 1, 2) == (4, 4)                        # fewer 16-byte units than ranks: empty slices are fine


@pytest.mark.parametrize("world,sizes,max_ctas", [(2, (9003, 3001, 144048, 12004), 16), (3, (300, 100, 4800, 400), 4),
                                                  (4, (30000, 10000, 480000, 40000), 24)])
@pytest.mark.timeout(120)
def test_exchange_adam_on_emulated_ranks_equals_allreduce_then_adam(world, sizes, max_ctas):
    """gsr_exchange_adam_step (slice-wise gradient sum -> Adam on the slice -> parameters written into every replica)
    against the two-step path it replaces: rank-ordered sum of the gradient buckets, then gsr_adam_step over the whole
    bucket. Parameters on every rank and the moment slices must match bit for bit, over two consecutive steps; the
    gradient buckets stay untouched; alignment padding between the parameter spans is left alone."""
    import gsr_optim
    L, _ = _gsr()
    dev = torch.device("cuda:0")
    gen = torch.Generator(device=dev).manual_seed(world)
    offs, off = [], 0
    for n in sizes:
        offs.append(off)
        off += (n + 31) // 32 * 32
    numel = off
    lrs = [1.6e-4, 5e-2, 2.5e-3, 1e-3]
    p0 = torch.randn(numel, device=dev, generator=gen)
    pbufs = [p0.clone() for _ in range(world)]
    pads = [torch.zeros(L.gsr_comm_signal_bytes() // 4, dtype=torch.int32, device=dev) for _ in range(world)]
    slices = [comm_slice(numel, r, world) for r in range(world)]
    ms = [torch.zeros(max(hi - lo, 4), device=dev) for lo, hi in slices]
    vs = [torch.zeros(max(hi - lo, 4), device=dev) for lo, hi in slices]
    ref_p, ref_m, ref_v = p0.clone(), torch.zeros(numel, device=dev), torch.zeros(numel, device=dev)
    b1, b2, eps = 0.9, 0.999, 1e-15
    for t in (1, 2):
        gbufs = [torch.randn(numel, device=dev, generator=gen) * (0.1 + r) for r in range(world)]
        for o, n in zip(offs, sizes):                       # padding carries no gradient
            for g in gbufs:
                g[o + n:(o + n + 31) // 32 * 32] = 0
        keep = [g.clone() for g in gbufs]
        gsum = gbufs[0].clone()
        for r in range(1, world):
            gsum = gsum + gbufs[r]
        spans = [(o, o + n, lr / (1 - b1 ** t), 1 / math.sqrt(1 - b2 ** t)) for o, n, lr in zip(offs, sizes, lrs)]
        gsr_optim.adam_step_raw([(ref_p[b:e], gsum[b:e], ref_m[b:e], ref_v[b:e], 1, ss, ib) for b, e, ss, ib in spans], b1, b2, eps)
        torch.cuda.synchronize()
        emulated_exchange_adam([_group(gbufs, pads, r) for r in range(world)], [_group(pbufs, pads, r) for r in range(world)],
                               ms, vs, spans, b1, b2, eps, numel, max_ctas)
        torch.cuda.synchronize()
        for r in range(world):
            assert torch.equal(pbufs[r], ref_p), f"step {t} rank {r}: max diff {(pbufs[r] - ref_p).abs().max().item()}"
            assert torch.equal(gbufs[r], keep[r]), "gradient bucket was modified"
            lo, hi = slices[r]
            assert torch.equal(ms[r][:hi - lo], ref_m[lo:hi]) and torch.equal(vs[r][:hi - lo], ref_v[lo:hi])
        assert all(int(p.abs().sum()) == 0 for p in pads)
    for o, n in zip(offs, sizes):                           # padding untouched
        assert torch.equal(pbufs[0][o + n:(o + n + 31) // 32 * 32], p0[o + n:(o + n + 31) // 32 * 32])
