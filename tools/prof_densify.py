"""Times one densify_and_prune at the bench size (3M Gaussians, SH degree 3, Adam moments present) on one GPU:
  ours   gsr_densify.densify_raw (plan + 16-byte read-back + one apply launch), per-kernel times from gsr_profile_*;
  torch  the reference's op sequence (scene/gaussian_model.py:287-429: mask gathers, torch.cat per tensor and moment,
         zero statistics, two prune passes) restated with the same torch ops on the same device -- the reference itself
         cannot run here (its module does not parse, make_golden_densify.py) and /root/reference is not on the GPU box.
Prints one JSON line. usage: python tools/prof_densify.py [--points 3000000] [--reps 5]"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gaussian-splatting-slam_b200"))
import gsr_densify as gd  # noqa: E402
from diff_gaussian_rasterization import _C  # noqa: E402

NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


def make(P, dev):
    g = torch.Generator(device=dev).manual_seed(0)
    p = {"xyz": torch.rand(P, 3, generator=g, device=dev) * 10 - 5, "f_dc": torch.randn(P, 1, 3, generator=g, device=dev),
         "f_rest": torch.randn(P, 15, 3, generator=g, device=dev), "opacity": torch.randn(P, 1, generator=g, device=dev) * 3 - 1,
         "scaling": torch.log(torch.rand(P, 3, generator=g, device=dev) * 0.065 + 1e-3),
         "rotation": torch.randn(P, 4, generator=g, device=dev)}
    mom = {n: (torch.randn(t.shape, generator=g, device=dev), torch.rand(t.shape, generator=g, device=dev)) for n, t in p.items()}
    denom = torch.randint(0, 4, (P, 1), generator=g, device=dev).float()
    accum = torch.rand(P, 1, generator=g, device=dev) * 0.0006 * denom
    return p, mom, accum, denom


def build_rotation(r):
    q = r / torch.sqrt((r * r).sum(dim=1))[:, None]
    R = torch.zeros((q.size(0), 3, 3), device=r.device)
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R[:, 0, 0] = 1 - 2 * (y * y + z * z); R[:, 0, 1] = 2 * (x * y - w * z); R[:, 0, 2] = 2 * (x * z + w * y)
    R[:, 1, 0] = 2 * (x * y + w * z); R[:, 1, 1] = 1 - 2 * (x * x + z * z); R[:, 1, 2] = 2 * (y * z - w * x)
    R[:, 2, 0] = 2 * (x * z - w * y); R[:, 2, 1] = 2 * (y * z + w * x); R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def torch_sequence(p, mom, accum, denom, max_grad, min_opacity, extent, mss, pd, noise, N=2):
    def cat(p, mom, new):
        p2 = {n: torch.cat((p[n], new[n]), dim=0) for n in NAMES}
        m2 = {n: (torch.cat((mom[n][0], torch.zeros_like(new[n])), dim=0), torch.cat((mom[n][1], torch.zeros_like(new[n])), dim=0))
              for n in NAMES}
        P = p2["xyz"].shape[0]
        return p2, m2, [torch.zeros((P, 1), device=dev), torch.zeros((P, 1), device=dev), torch.zeros(P, device=dev)]

    def prune(p, mom, stats, mask):
        keep = ~mask
        return ({n: p[n][keep] for n in NAMES}, {n: (mom[n][0][keep], mom[n][1][keep]) for n in NAMES}, [s[keep] for s in stats])

    dev = accum.device
    grads = accum / denom
    grads[grads.isnan()] = 0.0
    sel = torch.where(torch.norm(grads, dim=-1) >= max_grad, True, False)
    sel = torch.logical_and(sel, torch.max(torch.exp(p["scaling"]), dim=1).values <= pd * extent)
    p, mom, stats = cat(p, mom, {n: p[n][sel] for n in NAMES})
    n_init = p["xyz"].shape[0]
    padded = torch.zeros(n_init, device=dev)
    padded[:grads.shape[0]] = grads.squeeze()
    sel = torch.where(padded >= max_grad, True, False)
    sel = torch.logical_and(sel, torch.max(torch.exp(p["scaling"]), dim=1).values > pd * extent)
    stds = torch.exp(p["scaling"])[sel].repeat(N, 1)
    samples = noise[:stds.shape[0]] * stds
    rots = build_rotation(p["rotation"][sel]).repeat(N, 1, 1)
    new = {"xyz": torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + p["xyz"][sel].repeat(N, 1),
           "scaling": torch.log(torch.exp(p["scaling"])[sel].repeat(N, 1) / (0.8 * N)), "rotation": p["rotation"][sel].repeat(N, 1),
           "f_dc": p["f_dc"][sel].repeat(N, 1, 1), "f_rest": p["f_rest"][sel].repeat(N, 1, 1), "opacity": p["opacity"][sel].repeat(N, 1)}
    p, mom, stats = cat(p, mom, new)
    p, mom, stats = prune(p, mom, stats, torch.cat((sel, torch.zeros(N * int(sel.sum()), device=dev, dtype=bool))))
    mask = (torch.sigmoid(p["opacity"]) < min_opacity).squeeze()
    if mss:
        mask = torch.logical_or(torch.logical_or(mask, stats[2] > mss), torch.exp(p["scaling"]).max(dim=1).values > 0.1 * extent)
    return prune(p, mom, stats, mask)


def timed(fn, reps):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        out = fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
        del out
    return sorted(ts)[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=3_000_000)
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    p, mom, accum, denom = make(a.points, dev)
    rule = gd.make_rule(0.0002, 0.005, 5.0, 20, 0.01)
    _, _, counts = gd.densify_raw(p, mom, accum, denom, rule)
    noise = torch.randn(2 * counts[2], 3, device=dev)
    ours = timed(lambda: gd.densify_raw(p, mom, accum, denom, rule, noise), a.reps)
    ref = timed(lambda: torch_sequence(p, mom, accum, denom, 0.0002, 0.005, 5.0, 20, 0.01, noise), a.reps)
    _C.profile_enable(True)
    gd.densify_raw(p, mom, accum, denom, rule, noise)
    kern = {}
    for ln in _C.profile_report().splitlines():
        name, n, ms = ln.split()
        kern[name] = {"launches": int(n), "total_ms": float(ms)}
    _C.profile_enable(False)
    P_new = counts[0] + counts[1] + 2 * counts[3]
    floats = 59
    moved = 4 * floats * 3 * (counts[0] + P_new)            # kept rows read + every output row written, p / m / v
    line = {"points": a.points, "points_after": P_new, "counts": counts, "ours_ms": round(ours, 3), "torch_sequence_ms": round(ref, 3),
            "speedup": round(ref / ours, 2), "kernels": kern, "algorithmic_bytes": moved}
    if kern and "densify_apply" in kern:
        ms = kern["densify_apply"]["total_ms"] + kern.get("densify_zero_moments", {"total_ms": 0.0})["total_ms"]
        line["apply_plus_zero_GBps"] = round(moved / ms / 1e6, 1)          # algorithmic bytes over the two data-moving kernels
    print(json.dumps(line))


if __name__ == "__main__":
    main()
