"""GPU probe: where does the run-to-run (float-atomic) noise of the accumulated gradients sit? Two identical steps of the
C3 bench on one GPU; per tensor: rel-L2 of the difference, and the same with the k worst rows left out."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "gaussian-splatting-slam_b200")]
import torch, bench
sys.argv = [sys.argv[0]]
args = bench.parse_args()
dev = torch.device("cuda", 0)
sc, cams, dpix, _all = bench.build_workload(args, 0, 1, dev)
arm = bench.OursArm(args, sc, cams, dpix, dev)
runs = []
for _ in range(4):
    arm.step(); torch.cuda.synchronize()
    runs.append({k: v.detach().clone() for k, v in arm.bucket.views.items()})
out = {}
for k in runs[0]:
    a = runs[0][k].double().reshape(runs[0][k].shape[0], -1)
    res = []
    for r in runs[1:]:
        b = r[k].double().reshape(a.shape)
        e = ((a - b) ** 2).sum(1)
        tot = float(e.sum().sqrt() / a.norm())
        srt, _ = torch.sort(e, descending=True)
        row = {"rel": float("%.3e" % tot)}
        for kk in (8, 32, 128, 1024):
            row["drop%d" % kk] = float("%.3e" % float((srt[kk:].sum()).sqrt() / a.norm()))
        res.append(row)
    out[k] = res
print(json.dumps(out, indent=1))
