"""CPU tests (-m "not gpu") of the multi-GPU host logic with the gloo backend, world_size 2:
view sharding, the flat gradient bucket, the single exchange step, and 1-vs-2-rank gradient equality."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from view_parallel import GradBucket, reduce_densification_stats, shard_views


def test_shard_views_partitions_every_batch():
    for n in (0, 1, 7, 8, 64):
        for world in (1, 2, 3, 8):
            parts = [shard_views(n, r, world) for r in range(world)]
            assert sorted(sum(parts, [])) == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
    assert shard_views(8, 1, 4) == [1, 5]
    with pytest.raises(ValueError):
        shard_views(8, 4, 4)


def test_grad_bucket_views_alias_flat_storage_and_autograd_accumulates_in_place():
    p = {"a": torch.randn(5, 3, requires_grad=True), "b": torch.randn(7, requires_grad=True)}
    bk = GradBucket(p)
    assert bk.flat.numel() % 32 == 0 and all(o % 32 == 0 for o in bk.offsets.values())
    bk.attach()
    for _ in range(3):
        ((p["a"] * 2).sum() + (p["b"] * 3).sum()).backward()
    assert p["a"].grad.data_ptr() == bk.views["a"].data_ptr()
    assert torch.allclose(bk.views["a"], torch.full((5, 3), 6.0)) and torch.allclose(bk.views["b"], torch.full((7,), 9.0))
    bk.zero()
    assert float(bk.flat.abs().sum()) == 0.0 and float(p["a"].grad.abs().sum()) == 0.0
    assert bk.all_reduce() is None          # no process group: a no-op
    with pytest.raises(ValueError):
        GradBucket({})
    with pytest.raises(ValueError):
        GradBucket({"h": torch.zeros(3, dtype=torch.float16)})


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _toy_view_loss(params, view):
    # a differentiable stand-in for "render view v": each view weights the parameters differently
    w = torch.linspace(0.5, 1.5, params["x"].numel()).view_as(params["x"]) * (view + 1)
    return (params["x"] * w).sin().sum() + (params["y"] ** 2).sum() * (view + 1)


def _worker(rank, world, port, n_views, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(0)                       # replicated parameters
        params = {"x": torch.randn(6, 3, generator=g).requires_grad_(True),
                  "y": torch.randn(11, generator=g).requires_grad_(True)}
        bk = GradBucket(params)
        bk.zero(); bk.attach()
        for v in shard_views(n_views, rank, world):
            _toy_view_loss(params, v).backward()
        bk.all_reduce()
        acc, den, mx = torch.full((4,), float(rank + 1)), torch.ones(4), torch.tensor([1.0, 5.0, 2.0, 0.0]) * (rank + 1)
        reduce_densification_stats(acc, den, mx)
        if rank == 0:
            torch.save({"flat": bk.flat.clone(), "acc": acc, "den": den, "mx": mx}, out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_views", [8, 5])
def test_two_ranks_equal_one_rank(tmp_path, n_views):
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), n_views, out), nprocs=2, join=True)
    got = torch.load(out)
    g = torch.Generator().manual_seed(0)
    params = {"x": torch.randn(6, 3, generator=g).requires_grad_(True), "y": torch.randn(11, generator=g).requires_grad_(True)}
    bk = GradBucket(params)
    bk.attach()
    for v in range(n_views):
        _toy_view_loss(params, v).backward()
    assert torch.allclose(got["flat"], bk.flat, rtol=1e-5, atol=1e-6)
    assert torch.equal(got["acc"], torch.full((4,), 3.0)) and torch.equal(got["den"], torch.full((4,), 2.0))
    assert torch.equal(got["mx"], torch.tensor([2.0, 10.0, 4.0, 0.0]))


def test_comm_slices_partition_the_bucket_and_bad_arguments_are_refused(gsr_lib):
    """Host side of the peer-memory exchange (include/gsr.h: gsr_comm_slice, gsr_allreduce_sum): every rank's slice is
    16-byte granular, the slices are disjoint and cover the bucket; malformed groups never reach a launch."""
    import ctypes as C
    from view_parallel import _PeerGroup, comm_slice, _gsr
    L, _ = _gsr()
    for numel in (0, 4, 36, 4096, 3_000_000 * 59 // 4 * 4, 177_000_032):
        for world in (1, 2, 3, 4, 8, 16):
            parts = [comm_slice(numel, r, world) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == numel
            assert all(lo % 4 == 0 and hi % 4 == 0 and lo <= hi for lo, hi in parts)
            assert all(parts[r][1] == parts[r + 1][0] for r in range(world - 1))
            assert max(hi - lo for lo, hi in parts) <= (numel // 4 + world - 1) // world * 4
    lo, hi = C.c_longlong(), C.c_longlong()
    assert L.gsr_comm_slice(6, 0, 2, C.byref(lo), C.byref(hi)) == -1          # not a multiple of 4
    assert L.gsr_comm_slice(8, 2, 2, C.byref(lo), C.byref(hi)) == -1          # rank outside the group
    assert L.gsr_comm_slice(8, 0, 17, C.byref(lo), C.byref(hi)) == -1         # more peers than GSR_COMM_MAX_PEERS
    assert L.gsr_comm_signal_bytes() == 4 * (512 * 16 + 16)               # slots + the error area
    g = _PeerGroup()
    g.rank, g.world = 0, 2
    assert L.gsr_allreduce_sum(C.byref(g), 1024, 0, None) == -1               # NULL peer mappings
    assert L.gsr_allreduce_sum(None, 1024, 0, None) == -1
    g.buf[0], g.buf[1], g.signal[0], g.signal[1] = 4096, 8200, 4096, 4096
    assert L.gsr_allreduce_sum(C.byref(g), 1024, 0, None) == -1               # buf[1] not 16-byte aligned
    g.buf[1] = 8192
    assert L.gsr_allreduce_sum(C.byref(g), 1022, 0, None) == -1               # numel not a multiple of 4
    assert L.gsr_allreduce_sum(C.byref(g), 0, 0, None) == 0                   # empty bucket: nothing to launch


def test_grad_bucket_exchange_mode_is_validated():
    p = {"a": torch.randn(4, requires_grad=True)}
    assert GradBucket(p).exchange == "collective"
    with pytest.raises(ValueError):
        GradBucket(p, exchange="carrier-pigeon")


def test_exchange_adam_arguments_are_validated(gsr_lib):
    """gsr_exchange_adam_step never launches on malformed input (include/gsr.h): mismatched groups, overlapping or
    unordered spans, misaligned moments."""
    import ctypes as C
    from view_parallel import _AdamSpan, _PeerGroup, _gsr
    L, _ = _gsr()

    def grp(rank=0, world=2, mc=None):
        g = _PeerGroup()
        g.rank, g.world = rank, world
        for p in range(world):
            g.buf[p], g.signal[p] = 4096 * (p + 1), 65536 * (p + 1)
        g.multicast = mc
        return g
    spans = (_AdamSpan * 2)(_AdamSpan(0, 100, 1e-3, 1.0), _AdamSpan(128, 256, 1e-3, 1.0))
    call = lambda g, p, m, v, sp, n, numel: L.gsr_exchange_adam_step(C.byref(g), C.byref(p), C.c_void_p(m), C.c_void_p(v), sp, n,
                                                                     0.9, 0.999, 1e-15, numel, 0, None)
    assert call(grp(), grp(), 4096, 8192, spans, 2, 0) == 0                        # empty bucket: nothing to do
    assert call(grp(), grp(rank=1), 4096, 8192, spans, 2, 256) == -1               # the two buckets disagree on the rank
    assert call(grp(), grp(mc=1 << 20), 4096, 8192, spans, 2, 256) == -1           # one multicast, one not
    assert call(grp(), grp(), 4100, 8192, spans, 2, 256) == -1                     # moments not 16-byte aligned
    assert call(grp(), grp(), 4096, 8192, spans, 2, 200) == -1                     # span beyond the bucket
    bad = (_AdamSpan * 2)(_AdamSpan(0, 130, 1e-3, 1.0), _AdamSpan(128, 256, 1e-3, 1.0))
    assert call(grp(), grp(), 4096, 8192, bad, 2, 256) == -1                       # overlapping spans
    assert call(grp(), grp(), 4096, 8192, spans, 9, 256) == -1                     # more spans than GSR_ADAM_MAX_SPANS


def test_param_bucket_aliases_flat_storage_with_the_gradient_layout():
    """Host logic of the exchange+Adam path without peer memory: the parameter bucket copies the values, shares the
    gradient bucket's offsets, and hands out LEAF tensors that alias its flat storage (what the rasterizer must be fed)."""
    from view_parallel import ExchangeAdam, ParamBucket, bind_rank_to_gpu_numa
    p = {"a": torch.randn(5, 3, requires_grad=True), "b": torch.randn(7, requires_grad=True)}
    gb = GradBucket(p)
    pb = ParamBucket(p, gb)
    assert pb.peer is None and pb.offsets == gb.offsets and pb.flat.numel() == gb.flat.numel()
    for k in p:
        t = pb.tensors[k]
        assert t.is_leaf and t.requires_grad and torch.equal(t.detach(), p[k].detach())
        o = pb.offsets[k]
        assert t.data_ptr() == pb.flat[o:].data_ptr()
    with torch.no_grad():
        pb.flat[pb.offsets["b"]] = 42.0
    assert float(pb.tensors["b"].detach()[0]) == 42.0
    gb.params = pb.tensors                      # gradients now attach to the aliased leaves
    gb.attach()
    (pb.tensors["a"].sum() * 2).backward()
    assert torch.equal(gb.views["a"], torch.full((5, 3), 2.0))
    with pytest.raises(ValueError):
        ParamBucket({"b": p["b"], "a": p["a"]}, gb)          # order differs from the gradient bucket
    opt = ExchangeAdam(gb, pb, {"a": 1e-3, "b": 1e-2})
    assert not opt.fused and (opt.lo, opt.hi) == (0, gb.flat.numel())
    opt.t = 3
    spans = opt._spans()
    assert [(b, e) for b, e, _, _ in spans] == [(gb.offsets["a"], gb.offsets["a"] + 15), (gb.offsets["b"], gb.offsets["b"] + 7)]
    assert abs(spans[0][2] - 1e-3 / (1 - 0.9 ** 3)) < 1e-12 and abs(spans[1][3] - (1 - 0.999 ** 3) ** -0.5) < 1e-9
    assert bind_rank_to_gpu_numa(0) is None     # no NVML device here: the affinity is left alone
