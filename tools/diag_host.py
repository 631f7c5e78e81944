"""Diagnostic (GPU box): host profile of the small workloads (bench.py --workload c1 | slam), where the step is bound by
Python / launch overhead rather than by the kernels. usage: python tools/diag_host.py [c1|slam] [ours|reference]"""
import sys, os, time, cProfile, pstats
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "gaussian-splatting-slam_b200")]
wl = sys.argv[1] if len(sys.argv) > 1 else "c1"
impl = sys.argv[2] if len(sys.argv) > 2 else "ours"
import torch, bench, bench_extra
sys.argv = [sys.argv[0], "--workload", wl, "--impl", impl]
args = bench.parse_args()
dev = torch.device("cuda", 0)
arm = bench_extra.SmallArm(args, impl, dev)
for _ in range(5):
    arm.step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20):
    arm.step()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
n = 20 * args.views_per_step
print(f"{wl} {impl}: host enqueue {1e3 * (t1 - t0) / n:.4f} ms/view, with final sync {1e3 * (t2 - t0) / n:.4f} ms/view")
pr = cProfile.Profile(); pr.enable()
for _ in range(10):
    arm.step()
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(25)
