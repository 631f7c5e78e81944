"""View-parallel data parallelism for the rasterizer (the reference is single-GPU, so this axis is new).

One process per GPU. Every rank holds a full replica of the Gaussian parameters, renders a disjoint subset of the
batch's cameras, accumulates dL/dtheta over its views locally, and then takes part in ONE exchange step: a sum
all-reduce of the per-Gaussian parameter gradients (59 floats = 236 B per Gaussian at SH degree 3; 708 MB at 3M)
before the (identical) optimiser step. There is no data-path collective inside the rasterizer itself.

The gradients live in one flat fp32 bucket (`GradBucket`): the parameters' .grad tensors are views into it, so the
exchange is a single step over one contiguous buffer with no packing copies, and zeroing is one memset.

The exchange itself (`GradBucket.all_reduce`):
  * on CUDA ranks of one NVLink box the bucket is allocated as symmetric memory (every rank maps every other rank's
    bucket, plus the NVSwitch multicast mapping when the fabric offers one) and the step is ONE launch of libgsr's own
    kernel per rank (`gsr_allreduce_sum`, csrc/gsr_comm.cu): in-switch reduction with multimem.ld_reduce /
    multimem.st, or peer loads + peer stores without multicast. The autograd .grad tensors ARE that peer-mapped memory;
  * otherwise (gloo in the CPU tests, `GSR_EXCHANGE=nccl`, or a box where peer mapping cannot be set up) one
    torch.distributed all-reduce over the same flat buffer.
`GSR_EXCHANGE` = auto (default) | multimem | peer | nccl selects the path; anything but auto fails loudly when the
requested path is unavailable.
"""
import ctypes as C
import os
import sys
from typing import Dict, List, Sequence

import torch
import torch.distributed as dist

GSR_COMM_MAX_PEERS = 16


class _PeerGroup(C.Structure):      # gsr_peer_group (include/gsr.h)
    _fields_ = [("rank", C.c_int), ("world", C.c_int), ("buf", C.c_void_p * GSR_COMM_MAX_PEERS),
                ("signal", C.c_void_p * GSR_COMM_MAX_PEERS), ("multicast", C.c_void_p)]


class _AdamSpan(C.Structure):       # gsr_adam_span (include/gsr.h)
    _fields_ = [("begin", C.c_longlong), ("end", C.c_longlong), ("step_size", C.c_float), ("inv_sqrt_bc2", C.c_float)]


def _gsr():
    from diff_gaussian_rasterization import _C
    L = _C.lib()
    if not getattr(L, "_comm_bound", False):
        L.gsr_exchange_adam_step.restype = C.c_int
        L.gsr_exchange_adam_step.argtypes = [C.POINTER(_PeerGroup), C.POINTER(_PeerGroup), C.c_void_p, C.c_void_p,
                                             C.POINTER(_AdamSpan), C.c_int, C.c_double, C.c_double, C.c_double,
                                             C.c_longlong, C.c_int, C.c_void_p]
        L.gsr_comm_signal_bytes.restype = C.c_size_t
        L.gsr_comm_slice.restype = C.c_int
        L.gsr_comm_slice.argtypes = [C.c_longlong, C.c_int, C.c_int, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]
        L.gsr_allreduce_sum.restype = C.c_int
        L.gsr_allreduce_sum.argtypes = [C.POINTER(_PeerGroup), C.c_longlong, C.c_int, C.c_void_p]
        L.gsr_comm_emulation_scratch_bytes.restype = C.c_size_t
        L.gsr_allreduce_sum_emulated.restype = C.c_int
        L.gsr_allreduce_sum_emulated.argtypes = [C.POINTER(_PeerGroup), C.c_int, C.c_longlong, C.c_int, C.c_void_p, C.c_void_p]
        L.gsr_exchange_adam_step_emulated.restype = C.c_int
        L.gsr_exchange_adam_step_emulated.argtypes = [C.POINTER(_PeerGroup), C.POINTER(_PeerGroup), C.POINTER(C.c_void_p),
                                                      C.POINTER(C.c_void_p), C.c_int, C.POINTER(_AdamSpan), C.c_int, C.c_double,
                                                      C.c_double, C.c_double, C.c_longlong, C.c_int, C.c_void_p, C.c_void_p]
        L._comm_bound = True
    return L, _C


def comm_slice(numel: int, rank: int, world: int):
    """[lo, hi) of the bucket that `rank` reduces in gsr_allreduce_sum (host helper of the C ABI)."""
    L, _C = _gsr()
    lo, hi = C.c_longlong(0), C.c_longlong(0)
    _C._check(L.gsr_comm_slice(numel, rank, world, C.byref(lo), C.byref(hi)), "gsr_comm_slice")
    return lo.value, hi.value


def peer_all_reduce(group_desc: "_PeerGroup", numel: int, max_ctas: int = 0, stream=None):
    """One launch of the exchange kernel on `stream` (default: torch's current stream)."""
    L, _C = _gsr()
    st = torch.cuda.current_stream().cuda_stream if stream is None else stream
    _C._check(L.gsr_allreduce_sum(C.byref(group_desc), numel, max_ctas, C.c_void_p(st)), "gsr_allreduce_sum")


def _emulation_scratch(device):
    L, _ = _gsr()
    return torch.empty(int(L.gsr_comm_emulation_scratch_bytes()), dtype=torch.uint8, device=device)


def emulated_all_reduce(groups, numel: int, max_ctas: int = 0, stream=None):
    """Tests without a second GPU: every rank of an EMULATED group (all buffers and pads on the current device) in ONE
    cooperative launch -- kernels that wait for one another must never be separate launches on one GPU."""
    L, _C = _gsr()
    world = len(groups)
    arr = (_PeerGroup * world)(*groups)
    dev = torch.device("cuda", torch.cuda.current_device())
    scratch = _emulation_scratch(dev)
    st = torch.cuda.current_stream().cuda_stream if stream is None else stream
    _C._check(L.gsr_allreduce_sum_emulated(arr, world, numel, max_ctas, C.c_void_p(scratch.data_ptr()), C.c_void_p(st)),
              "gsr_allreduce_sum_emulated")
    torch.cuda.current_stream().synchronize()       # scratch must outlive the launch


def emulated_exchange_adam(grads, params, exp_avg, exp_avg_sq, spans, beta1, beta2, eps, numel, max_ctas: int = 0):
    """gsr_exchange_adam_step for all ranks of an emulated group in one cooperative launch (see emulated_all_reduce)."""
    L, _C = _gsr()
    world = len(grads)
    ga, pa = (_PeerGroup * world)(*grads), (_PeerGroup * world)(*params)
    ma = (C.c_void_p * world)(*[t.data_ptr() for t in exp_avg])
    va = (C.c_void_p * world)(*[t.data_ptr() for t in exp_avg_sq])
    arr = (_AdamSpan * max(1, len(spans)))()
    for k, (b, e, ss, ib) in enumerate(spans):
        arr[k] = _AdamSpan(int(b), int(e), float(ss), float(ib))
    dev = torch.device("cuda", torch.cuda.current_device())
    scratch = _emulation_scratch(dev)
    st = torch.cuda.current_stream().cuda_stream
    _C._check(L.gsr_exchange_adam_step_emulated(ga, pa, ma, va, world, arr, len(spans), beta1, beta2, eps, numel, max_ctas,
                                                C.c_void_p(scratch.data_ptr()), C.c_void_p(st)),
              "gsr_exchange_adam_step_emulated")
    torch.cuda.current_stream().synchronize()


class PeerExchange:
    """Flat fp32 buffer in symmetric memory + the descriptor gsr_allreduce_sum needs. Collective constructor."""

    def __init__(self, numel: int, device, group=None, multicast=True, agree=None):
        """agree(ok: bool) -> bool, optional: a collective AND over the ranks, called between the local and the collective
        stages of the set-up. With it, a rank whose local allocation failed and the healthy ranks all learn so BEFORE
        anybody enters the rendezvous, and every rank raises together instead of running mismatched collectives
        (ADVICE r1). Without it a local failure simply raises on that rank."""
        import torch.distributed._symmetric_memory as symm_mem
        L, _ = _gsr()
        group = group if group is not None else dist.group.WORLD
        if numel % 4:
            raise ValueError("peer exchange needs a multiple of 4 elements")
        # ---- stage 1: local (no collective): the two symmetric allocations
        err = None
        try:
            self.flat = symm_mem.empty(numel, dtype=torch.float32, device=device)
            self.signal = symm_mem.empty(L.gsr_comm_signal_bytes() // 4, dtype=torch.int32, device=device)
        except Exception as e:          # noqa: BLE001
            err = e
        if agree is not None:
            if not agree(err is None):
                raise RuntimeError(f"symmetric allocation failed on some rank ({err})")
        elif err is not None:
            raise err
        # ---- stage 2: collective on every rank (rendezvous x 2, barrier)
        self.hdl = symm_mem.rendezvous(self.flat, group)
        self.signal.zero_()
        self.sig_hdl = symm_mem.rendezvous(self.signal, group)
        self.flat.zero_()
        torch.cuda.synchronize(device)
        dist.barrier(group)                        # every pad is zero before anybody signals
        h, sh = self.hdl, self.sig_hdl
        if h.world_size > GSR_COMM_MAX_PEERS:
            raise RuntimeError(f"peer exchange supports up to {GSR_COMM_MAX_PEERS} ranks")
        g = _PeerGroup()
        g.rank, g.world = h.rank, h.world_size
        off = self.flat.data_ptr() - int(h.buffer_ptrs[h.rank])         # the tensor may sit inside a pooled block
        soff = self.signal.data_ptr() - int(sh.buffer_ptrs[sh.rank])
        for p in range(h.world_size):
            g.buf[p] = int(h.buffer_ptrs[p]) + off
            g.signal[p] = int(sh.buffer_ptrs[p]) + soff
        # link bytes per rank and direction for a bucket of S bytes: multicast (world + 1) / world x S (the switch also
        # pulls the local contribution), peer loads + stores 2 (world - 1) / world x S -- multicast wins from 4 ranks on
        # (measured at 2 ranks, 708 MB: peer 1.06 ms, multicast 1.83 ms, NCCL 1.31 ms; profiles/r01e_comm_check_n2.json)
        if multicast == "auto":
            multicast = h.world_size >= 4
        mc = int(h.multicast_ptr) if multicast else 0
        g.multicast = (mc + off) if mc else None
        self.desc, self.numel, self.has_multicast = g, numel, bool(mc)
        self.kind = "multimem" if mc else "peer"

    def all_reduce(self, max_ctas: int = 0):
        peer_all_reduce(self.desc, self.numel, max_ctas)


def exchange_adam_raw(grads: "_PeerGroup", params: "_PeerGroup", exp_avg, exp_avg_sq, spans, beta1, beta2, eps, numel,
                      max_ctas: int = 0, stream=None):
    """One launch of gsr_exchange_adam_step. spans: list of (begin, end, step_size, inv_sqrt_bc2)."""
    L, _C = _gsr()
    arr = (_AdamSpan * max(1, len(spans)))()
    for k, (b, e, ss, ib) in enumerate(spans):
        arr[k] = _AdamSpan(int(b), int(e), float(ss), float(ib))
    st = torch.cuda.current_stream().cuda_stream if stream is None else stream
    _C._check(L.gsr_exchange_adam_step(C.byref(grads), C.byref(params), C.c_void_p(exp_avg.data_ptr()),
                                       C.c_void_p(exp_avg_sq.data_ptr()), arr, len(spans), beta1, beta2, eps, numel,
                                       max_ctas, C.c_void_p(st)), "gsr_exchange_adam_step")


class ParamBucket:
    """The Gaussian parameters as views of ONE flat buffer with the GradBucket's layout. When the gradient bucket lives
    in peer memory so does this one, which is what lets `ExchangeAdam` write a rank's updated slice into every replica.
    `tensors[name]` are leaf tensors (requires_grad) aliasing the flat storage; use them instead of the originals."""

    def __init__(self, params: Dict[str, torch.Tensor], grads: "GradBucket", group=None):
        if list(params.keys()) != list(grads.params.keys()):
            raise ValueError("ParamBucket: parameter names / order differ from the gradient bucket")
        self.peer = None
        numel = grads.flat.numel()
        if grads.peer is not None:
            self.peer = PeerExchange(numel, grads.device, group, multicast=grads.peer.has_multicast)
            if self.peer.has_multicast != grads.peer.has_multicast:
                raise RuntimeError("ParamBucket: parameter and gradient buckets must use the same data path")
            self.flat = self.peer.flat
        else:
            self.flat = torch.zeros(numel, dtype=torch.float32, device=grads.device)
        self.offsets = dict(grads.offsets)
        self.tensors = {}
        with torch.no_grad():
            for name, p in params.items():
                o = self.offsets[name]
                view = self.flat[o:o + p.numel()].view_as(p)
                view.copy_(p)
                self.tensors[name] = view.requires_grad_(True)
        if self.peer is not None:
            torch.cuda.synchronize(grads.device)
            dist.barrier(group)                    # every replica initialised before anybody's kernel writes into it


class ExchangeAdam:
    """The exchange step and the optimiser step of view-parallel training as ONE launch per rank
    (gsr_exchange_adam_step): rank r sums slice r of the gradient buckets, runs Adam on slice r of the parameters with
    its own slice of the moments and writes the result into every replica. Same update as
    `GradBucket.all_reduce()` followed by `gsr_optim.FusedAdam.step()` on every rank (torch.optim.Adam semantics:
    bias correction by the step count, eps added after the square root; scene/gaussian_model.py:168 uses eps 1e-15),
    bit for bit on the peer path. Moments are stored for the own slice only (1 / world of FusedAdam's state).
    Without peer memory (one rank, gloo, GSR_EXCHANGE=nccl) it does exactly those two steps with full-size moments."""

    def __init__(self, grads: "GradBucket", params: ParamBucket, lrs: Dict[str, float], betas=(0.9, 0.999), eps=1e-15,
                 group=None):
        self.grads, self.params, self.lrs = grads, params, dict(lrs)
        self.betas, self.eps, self.group, self.t = betas, eps, group, 0
        numel = grads.flat.numel()
        self.fused = grads.peer is not None and params.peer is not None
        if self.fused:
            d = grads.peer.desc
            lo, hi = comm_slice(numel, d.rank, d.world)
        else:
            lo, hi = 0, numel
        self.lo, self.hi = lo, hi
        n = max(hi - lo, 4)
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=grads.device)
        self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=grads.device)

    def _spans(self):
        import math
        b1, b2 = self.betas
        out = []
        for name, p in self.grads.params.items():
            o = self.grads.offsets[name]
            lr = self.lrs[name]
            out.append((o, o + p.numel(), lr / (1.0 - b1 ** self.t), 1.0 / math.sqrt(1.0 - b2 ** self.t)))
        return out

    @torch.no_grad()
    def step(self):
        self.t += 1
        spans = self._spans()
        b1, b2 = self.betas
        if self.fused:
            exchange_adam_raw(self.grads.peer.desc, self.params.peer.desc, self.exp_avg, self.exp_avg_sq, spans, b1, b2,
                              self.eps, self.grads.flat.numel())
            return
        import gsr_optim
        self.grads.all_reduce(self.group)
        entries = []
        for (name, p), (b, e, ss, ib) in zip(self.grads.params.items(), spans):
            row_len = p.numel() // p.shape[0] if p.dim() > 0 and p.shape[0] > 0 else 1
            entries.append((self.params.flat[b:e], self.grads.flat[b:e], self.exp_avg[b:e], self.exp_avg_sq[b:e], row_len, ss, ib))
        gsr_optim.adam_step_raw(entries, b1, b2, self.eps)


def bind_rank_to_gpu_numa(device_index: int):
    """Pins the calling process to the CPUs NVML reports as local to GPU `device_index` (its NUMA node), so that the
    rank's pinned host buffers are first-touched next to its GPU and its H2D / D2H copies do not cross the socket
    interconnect. With 8 ranks sharing one host this is what keeps the end-to-end (host-buffer) path scaling.
    Returns the CPU list it bound to, or None when NVML has no answer (the process then stays as it was)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return sorted(cpus)
    except Exception:       # noqa: BLE001 - no NVML / no permission: leave the affinity alone
        return None


def shard_views(num_views: int, rank: int, world_size: int) -> List[int]:
    """Views r, r+n, r+2n, ... of a batch go to rank r. Ragged batches are allowed."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError(f"bad rank/world_size {rank}/{world_size}")
    return list(range(rank, num_views, world_size))


class GradBucket:
    """Flat gradient storage for a dict of leaf parameters."""

    def __init__(self, params: Dict[str, torch.Tensor], exchange: str = None, group=None):
        if not params:
            raise ValueError("GradBucket needs at least one parameter")
        dev = {p.device for p in params.values()}
        if len(dev) != 1:
            raise ValueError("all parameters must live on one device")
        self.params = dict(params)
        self.device = dev.pop()
        sizes = [p.numel() for p in self.params.values()]
        # 128-byte aligned slices so vectorised kernels may write them
        self.offsets, off = {}, 0
        for (name, p), n in zip(self.params.items(), sizes):
            if p.dtype != torch.float32:
                raise ValueError(f"{name}: gradients are exchanged in fp32")
            self.offsets[name] = off
            off += (n + 31) // 32 * 32
        self.peer, self.exchange = None, "collective"
        mode = (exchange or os.environ.get("GSR_EXCHANGE", "auto")).lower()
        if mode not in ("auto", "multimem", "peer", "nccl"):
            raise ValueError(f"GSR_EXCHANGE={mode}: expected auto, multimem, peer or nccl")
        multi = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
        if multi and self.device.type == "cuda" and mode != "nccl":
            self.peer = self._try_peer(off, mode, group)
        if self.peer is not None:
            self.flat, self.exchange = self.peer.flat, self.peer.kind
        else:
            self.flat = torch.zeros(off, dtype=torch.float32, device=self.device)
        self.views = {name: self.flat[o:o + p.numel()].view_as(p) for (name, p), o in
                      zip(self.params.items(), self.offsets.values())}

    def _try_peer(self, numel, mode, group):
        """Sets the peer-memory exchange up on all ranks or on none. Every rank goes through the same collectives in the
        same order whatever fails where: agreement (a MIN all-reduce) after the local allocations, the rendezvous only if
        all of them succeeded, agreement again after it."""
        def agree(ok: bool) -> bool:
            t = torch.tensor([1 if ok else 0], device=self.device, dtype=torch.int32)
            dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
            return int(t.item()) == 1

        px, err = None, None
        try:
            px = PeerExchange(numel, self.device, group, multicast={"peer": False, "multimem": True}.get(mode, "auto"),
                              agree=agree)
            if mode == "multimem" and not px.has_multicast:
                raise RuntimeError("no NVSwitch multicast mapping for the bucket")
        except Exception as e:          # noqa: BLE001 - any set-up failure means "not available here"
            px, err = None, e
        # a failure in stage 1 was agreed on inside the constructor and raised on every rank alike, so all ranks arrive
        # here having run the same collectives; this second agreement covers asymmetric failures after the rendezvous
        if agree(px is not None):
            return px
        if mode != "auto":
            raise RuntimeError(f"GSR_EXCHANGE={mode} requested but peer memory could not be set up: {err}")
        if dist.get_rank(group) == 0:
            print(f"[view_parallel] peer-memory exchange unavailable ({err}); using the backend's all-reduce", file=sys.stderr)
        return None

    def attach(self):
        """Points every parameter's .grad at its slice of the bucket (autograd then accumulates in place)."""
        for name, p in self.params.items():
            p.grad = self.views[name]

    def zero(self):
        self.flat.zero_()

    @property
    def views_first(self):
        """grad_sinks for the FIRST view of a step: the backward writes the bucket in full (zeros for culled Gaussians)
        instead of adding to it, so the step needs no zero() -- 708 MB of stores at 3M Gaussians."""
        d = dict(self.views)
        d["overwrite"] = True
        return d

    def nbytes(self) -> int:
        return self.flat.numel() * 4

    def all_reduce(self, group=None, async_op=False, average=False):
        """The exchange step: sum (optionally mean) of the bucket over all ranks."""
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return None
        if self.peer is not None:
            self.peer.all_reduce()                  # one launch on the current stream; stream order is the only sync
            if average:
                self.flat.div_(dist.get_world_size(group))
            return None
        work = dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
        if average and not async_op:
            self.flat.div_(dist.get_world_size(group))
        return work


def reduce_densification_stats(grad_accum: torch.Tensor, denom: torch.Tensor, max_radii2D: torch.Tensor, group=None):
    """Every densification interval: xyz_gradient_accum / denom are summed, max_radii2D is max-reduced
    (scene/gaussian_model.py:431-433, train.py:159 on one GPU)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    dist.all_reduce(grad_accum, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(denom, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(max_radii2D, op=dist.ReduceOp.MAX, group=group)


def render_views_accumulate(rasterize_one, views: Sequence, bucket: GradBucket):
    """Renders `views` one after the other with `rasterize_one(view) -> (outputs, grad_outputs)` and backpropagates
    each into the bucket. Returns the number of views rendered."""
    n = 0
    for v in views:
        outs, gouts = rasterize_one(v)
        torch.autograd.backward(list(outs), list(gouts))
        n += 1
    return n
