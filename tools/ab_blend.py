"""A/B of the blend kernels' list walk on the C3 bench view set (GPU box): per-kernel device times (CUDA events inside
libgsr) for every gsr_debug_set_blend_variant setting, plus a bit-identity check of the rendered images between them.
usage: python tools/ab_blend.py [variants ...]   (default: 0 1)"""
import ctypes as C
import json
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "gaussian-splatting-slam_b200")]
import torch, bench
variants = [int(a) for a in sys.argv[1:]] or [0, 1]
sys.argv = [sys.argv[0]]
args = bench.parse_args()
dev = torch.device("cuda", 0)
sc, cams, dpix, _all = bench.build_workload(args, 0, 1, dev)
arm = bench.OursArm(args, sc, cams, dpix, dev)
L = arm._C.lib()
L.gsr_debug_set_blend_variant.argtypes = [C.c_int]
out, imgs = {}, {}
for v in variants:
    L.gsr_debug_set_blend_variant(v)
    for _ in range(3):
        arm.step()
    torch.cuda.synchronize()
    arm._C.profile_enable(True)
    for _ in range(5):
        arm.step()
    torch.cuda.synchronize()
    arm._C.profile_enable(False)
    tab = bench.kernel_table(arm._C.profile_report(), 5 * args.views_per_step)
    out[v] = {k: round(r["ms_per_view"], 4) for k, r in tab.items() if k.startswith("blend")}
    imgs[v] = arm.last[0].detach().clone()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        arm.step()
    e1.record()
    torch.cuda.synchronize()
    out[v]["ms_per_view_total"] = round(e0.elapsed_time(e1) / (5 * args.views_per_step), 4)
base = imgs[variants[0]]
out["images_bit_identical"] = all(bool(torch.equal(base, imgs[v])) for v in variants)
print(json.dumps(out))
