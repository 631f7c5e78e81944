"""GPU tests (-m gpu) of the peer-memory gradient exchange (csrc/gsr_comm.cu) on ONE device: the "ranks"
are buffers and signal pads on the same GPU and their kernels run concurrently on separate streams, so the twin
barriers, the slice ownership and the rank-ordered summation are exercised exactly as across NVLink (the multicast
variant needs a real NVSwitch group: tools/comm_check.py under torchrun covers it). Oracle: the sum in rank order,
computed by torch on the same device, bit for bit."""
import ctypes as C

import pytest
import torch

from view_parallel import _PeerGroup, comm_slice, peer_all_reduce, _gsr

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
    streams = [torch.cuda.Stream(dev) for _ in range(world)]
    torch.cuda.synchronize()
    for rep in range(3):                                        # the pads reset themselves: the launch is repeatable
        want = bufs[0].clone()
        for r in range(1, world):
            want = want + bufs[r]                               # rank order, fp32, round to nearest
        # all ranks' CTAs must be co-resident (they wait for their twins): max_ctas x world stays far below 2 x 148
        for r in range(world):
            peer_all_reduce(_group(bufs, pads, r), numel, max_ctas, stream=streams[r].cuda_stream)
        torch.cuda.synchronize()
        for r in range(world):
            assert torch.equal(bufs[r], want), f"rank {r} rep {rep}: max diff {(bufs[r] - want).abs().max().item()}"
        assert all(int(p.abs().sum()) == 0 for p in pads)       # every signal consumed


def test_slices_match_the_kernel_ownership():
    assert comm_slice(36, 0, 2) == (0, 20) and comm_slice(36, 1, 2) == (20, 36)
    assert comm_slice(4, 1, 2) == (4, 4)                        # fewer 16-byte units than ranks: empty slices are fine
