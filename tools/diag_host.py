"""Diagnostic (GPU box): pure host overhead of one forward+backward through the public API (tiny scene, GPU idle)."""
import sys, os, time, cProfile, pstats
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "gaussian-splatting-slam_b200")]
import torch, bench
sys.argv = [sys.argv[0], "--gaussians", "2000", "--width", "64", "--height", "64"]
args = bench.parse_args()
dev = torch.device("cuda", 0)
sc, cams, dpix, _all = bench.build_workload(args, 0, 1, dev)
arm = bench.OursArm(args, sc, cams, dpix, dev)
for _ in range(5):
    arm.step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20):
    arm.step()
torch.cuda.synchronize()
print("host ms per view:", (time.perf_counter() - t0) * 1e3 / (20 * 8))
pr = cProfile.Profile(); pr.enable()
for _ in range(10):
    arm.step()
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(22)
