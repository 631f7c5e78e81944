"""Diagnostic (GPU box): run-to-run stability of the bench step (per-step device + wall time over many steps)."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "gaussian-splatting-slam_b200")]
import torch, bench
sys.argv = [sys.argv[0]] + sys.argv[1:]
args = bench.parse_args()
dev = torch.device("cuda", 0)
sc, cams, dpix = bench.build_workload(args, 0, 1, dev)
arm = bench.OursArm(args, sc, cams, dpix, dev)
for it in range(14):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    arm.step()
    t_enq = (time.perf_counter() - t0) * 1e3
    e1.record(); torch.cuda.synchronize()
    print(f"step {it}: device {e0.elapsed_time(e1)/args.views_per_step:.3f} ms/view, wall {(time.perf_counter()-t0)*1e3/args.views_per_step:.3f}, host enqueue {t_enq/args.views_per_step:.3f}", flush=True)
