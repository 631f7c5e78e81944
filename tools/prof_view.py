"""Profiling driver (GPU box): a few forward(+backward) views of the C3 bench workload, nothing else."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "gaussian-splatting-slam_b200")]
import torch, bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
sys.argv = [sys.argv[0], "--views-per-step", "1"]
args = bench.parse_args()
dev = torch.device("cuda", 0)
sc, cams, dpix, _all = bench.build_workload(args, 0, 1, dev)
arm = bench.OursArm(args, sc, cams, dpix, dev)
for _ in range(n):
    arm.step()
torch.cuda.synchronize()
print("ok")
