"""Multi-rank check + timing of the peer-memory gradient exchange (csrc/gsr_comm.cu) against NCCL on a real NVLink box.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
      tools/comm_check.py [--gaussians 3000000] [--json gpurun_out/comm_check.json]

For each available path (multimem, peer, nccl): correctness on random per-rank buckets (peer: bit-exact against the
rank-ordered sum; multimem / nccl: within fp32 summation-order tolerance), then the time of ONE exchange of the C3 bucket
(59 floats per Gaussian) measured with CUDA events on the launching stream, max over ranks, median of `--iters`.
Rank 0 prints one JSON object. Not part of the product path."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-slam_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gaussians", type=int, default=3_000_000)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--json", default="")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from view_parallel import PeerExchange
    numel = (args.gaussians * 59 + 31) // 32 * 32
    out = {"world": world, "numel": numel, "bytes": numel * 4, "paths": {}}

    def contrib(r):
        g = torch.Generator(device=dev).manual_seed(1234 + r)
        return torch.randn(numel, device=dev, generator=g) * (1.0 + r)

    def time_it(fn):
        ms = []
        for _ in range(3):
            fn()
        for _ in range(args.iters):
            torch.cuda.synchronize()
            dist.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            t = torch.tensor([a.elapsed_time(b)], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms.append(float(t.item()))
        ms.sort()
        return ms[len(ms) // 2], ms[0]

    # the rank-ordered fp32 sum (the peer path's contract), built locally from the seeds
    want = contrib(0)
    for r in range(1, world):
        want += contrib(r)
    mine = contrib(rank)

    for kind in ("multimem", "peer"):
        try:
            px = PeerExchange(numel, dev, multicast=(kind == "multimem"))
            if kind == "multimem" and not px.has_multicast:
                raise RuntimeError("handle has no multicast mapping")
        except Exception as e:      # noqa: BLE001
            out["paths"][kind] = {"available": False, "why": repr(e)[:300]}
            continue
        px.flat.copy_(mine)
        torch.cuda.synchronize(); dist.barrier()
        px.all_reduce()
        torch.cuda.synchronize(); dist.barrier()
        diff = (px.flat - want).abs().max()
        bit_exact = bool(torch.equal(px.flat, want))
        rel = float((px.flat - want).double().norm() / want.double().norm())
        same = px.flat.clone()
        dist.broadcast(same, 0)
        identical_across_ranks = torch.tensor([int(torch.equal(same, px.flat))], device=dev)
        dist.all_reduce(identical_across_ranks, op=dist.ReduceOp.MIN)
        med, best = time_it(px.all_reduce)
        # bytes through one rank's links, per direction (see view_parallel.PeerExchange)
        per_rank_link_bytes = numel * 4 * ((world + 1) / world if kind == "multimem" else 2 * (world - 1) / world)
        out["paths"][kind] = {"available": True, "max_abs_diff": float(diff), "rel_l2": rel, "bit_exact_vs_rank_order": bit_exact,
                              "identical_on_all_ranks": bool(identical_across_ranks.item()), "ms_median": med, "ms_best": best,
                              "link_GBps_per_rank_per_dir": per_rank_link_bytes / med / 1e6}
        del px
    flat = mine.clone()
    dist.all_reduce(flat)
    torch.cuda.synchronize()
    rel = float((flat - want).double().norm() / want.double().norm())
    med, best = time_it(lambda: dist.all_reduce(flat))
    out["paths"]["nccl"] = {"available": True, "rel_l2": rel, "ms_median": med, "ms_best": best,
                            "link_GBps_per_rank_per_dir": numel * 4 * 2 * (world - 1) / world / med / 1e6}
    if rank == 0:
        line = json.dumps(out)
        print(line)
        if args.json:
            os.makedirs(os.path.dirname(args.json) or ".", exist_ok=True)
            open(args.json, "w").write(line + "\n")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
