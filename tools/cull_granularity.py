"""Design study (GPU box): how many (region, splat) evaluations the blend kernels would do at finer culling granularity.
Prints the work counters of one C3 view: sub-tile (8x8, what the kernels cull at today) and half sub-tile (8x4)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "gaussian-splatting-slam_b200")]
import torch, bench
sys.argv = [sys.argv[0], "--views-per-step", "1"]
args = bench.parse_args()
dev = torch.device("cuda", 0)
sc, cams, dpix, _all = bench.build_workload(args, 0, 1, dev)
arm = bench.OursArm(args, sc, cams, dpix, dev)
R, radii, st = arm.state_of_view(0)
trav, contrib = st["pair_counts"]
box, ell, anyc, missed, half = arm._C.blend_stats.last_cull[:5]
out = {"num_rendered": R, "traversed_pixel_pairs": trav, "contributing_pixel_pairs": contrib,
       "subtile_box_pass": box, "subtile_ellipse_pass": ell, "subtile_with_contribution": anyc, "ellipse_false_drops": missed,
       "half_subtile_ellipse_pass": half,
       "pixel_pairs_evaluated_8x8": 64 * ell, "pixel_pairs_evaluated_8x4": 32 * half,
       "lane_efficiency_8x8": contrib / (64 * ell) if ell else None, "lane_efficiency_8x4": contrib / (32 * half) if half else None}
print(json.dumps(out))
