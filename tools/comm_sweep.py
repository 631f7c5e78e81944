"""GPU tool (torchrun, one rank per GPU): sweep of the multicast exchange's experimental variants
(gsr_debug_set_comm_variant, csrc/gsr_comm.cu) and CTA counts on the 708 MB gradient bucket. Prints one JSON object on
rank 0. Variants 1 / 2 are probes of one request stream (ld_reduce only / multimem.st only), not all-reduces.
usage: python -m torch.distributed.run --nproc-per-node N tools/comm_sweep.py [--json out.json]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "gaussian-splatting-slam_b200")]
import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gaussians", type=int, default=3_000_000)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--json", default="")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from diff_gaussian_rasterization import _C
    from view_parallel import PeerExchange
    L = _C.lib()
    numel = (args.gaussians * 59 + 31) // 32 * 32

    def contrib(r):
        g = torch.Generator(device=dev).manual_seed(1234 + r)
        return torch.randn(numel, device=dev, generator=g) * (1.0 + r)

    want = contrib(0)
    for r in range(1, world):
        want += contrib(r)
    mine = contrib(rank)
    px = PeerExchange(numel, dev, multicast=True)
    out = {"world": world, "bytes": numel * 4, "has_multicast": bool(px.has_multicast), "runs": []}

    def time_it(fn):
        ms = []
        for _ in range(2):
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
        return ms[len(ms) // 2]

    if px.has_multicast:
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        for variant in (0, 4, 16, 3, 1, 2):
            for ctas in (sms // 2, sms, 2 * sms):
                L.gsr_debug_set_comm_variant(variant)
                row = {"variant": variant, "max_ctas": ctas}
                if variant in (0, 3, 4, 16):
                    px.flat.copy_(mine)
                    torch.cuda.synchronize(); dist.barrier()
                    px.all_reduce(ctas)
                    torch.cuda.synchronize(); dist.barrier()
                    row["rel_l2"] = float((px.flat - want).double().norm() / want.double().norm())
                row["ms_median"] = time_it(lambda: px.all_reduce(ctas))
                # bytes through one rank's links per direction: all-reduce (world+1)/world of the bucket, a probe 1x
                per_dir = numel * 4 * ((world + 1) / world if variant in (0, 3, 4, 16) else 1.0)
                row["link_GBps_per_rank_per_dir"] = per_dir / row["ms_median"] / 1e6
                out["runs"].append(row)
                if rank == 0:
                    print(json.dumps(row), flush=True)
        L.gsr_debug_set_comm_variant(0)
    if rank == 0:
        if args.json:
            with open(args.json, "w") as f:
                json.dump(out, f, indent=1)
        print(json.dumps(out))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
