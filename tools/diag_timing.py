"""Diagnostic (GPU box): per-step device time + allocator activity over many steps of the bench workload."""
import sys, os, time, gc
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "gaussian-splatting-slam_b200")]
import torch, bench
args = bench.parse_args()
dev = torch.device("cuda", 0)
sc, cams, dpix, _all = bench.build_workload(args, 0, 1, dev)
arm = bench.OursArm(args, sc, cams, dpix, dev)
def stat(k): return torch.cuda.memory_stats().get(k, 0)
for it in range(40):
    torch.cuda.synchronize()
    a0, s0 = stat("num_device_alloc"), stat("num_device_free")
    g0 = gc.get_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    arm.step()
    e1.record(); torch.cuda.synchronize()
    print(f"step {it}: device {e0.elapsed_time(e1):.2f} ms wall {(time.perf_counter()-t0)*1e3:.2f} cudaMalloc +{stat('num_device_alloc')-a0} cudaFree +{stat('num_device_free')-s0} reserved {torch.cuda.memory_reserved()>>20} MiB gc {g0}", flush=True)
