"""GPU probe: distance of our needle-scene gradients from the float64 derivation, run to run, per blend variant and binding
(test_gpu_parity.py::test_needles_and_faint_splats_exercise_the_ellipse_cull is the gate; this prints the spread)."""
import math
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "gaussian-splatting-slam_b200"), os.path.join(ROOT, "tests")]
import numpy as np
import torch
import scenes
import test_gpu_parity as T
from diff_gaussian_rasterization import _C
from oracle import ref_cuda
from util import rel_l2

g = torch.Generator().manual_seed(21)
P = 4000
sc = scenes.small_scene(P=P, deg=2, spread=1.4, seed=21)
long_axis = torch.randint(0, 3, (P,), generator=g)
scales = torch.exp(math.log(0.002) + torch.rand(P, 3, generator=g) * (math.log(0.008) - math.log(0.002)))
scales[torch.arange(P), long_axis] = 0.15 + 0.5 * torch.rand(P, generator=g)
opac = torch.sigmoid(torch.randn(P, 1, generator=g) * 3.0 - 3.0)
opac[:200] = (1.0 / 255.0) * (0.9 + 0.3 * torch.rand(200, 1, generator=g))
sc = scenes.Scene(sc.means3D, scales, sc.rotations, opac, sc.shs, sc.sh_degree)
cam = scenes.look_at_camera(320, 200, math.radians(65), (0.2, 0.1, -3.2))
names = (("means3D", "dL_dmeans3D", "dL_dmean3D"), ("scales", "dL_dscales", "dL_dscale"), ("rotations", "dL_drotations", "dL_drot"))
print("binding", _C.binding_name())
for bg, seed in (((0.0, 0.0, 0.0), 5), ((0.3, 0.6, 0.1), 6)):
    dpix = np.random.default_rng(seed).standard_normal((3, cam.H, cam.W)).astype(np.float32)
    g64 = T._f64_autograd_grads(sc, cam, bg, False, dpix, None)
    fw = T._ref_forward(sc, cam, bg)
    for rep in range(2):
        gref = {k: v.cpu().numpy() for k, v in ref_cuda.backward(fw, torch.as_tensor(dpix, device=T.DEV)).items()}
        print("bg", bg, "reference", {o: float("%.3e" % rel_l2(gref[r].reshape(np.asarray(g64[n]).shape), np.asarray(g64[n]))) for o, r, n in names})
    for variant in (2, 0):
        _C.lib().gsr_debug_set_blend_variant(variant)
        for rep in range(3):
            ours = T._run_ours(sc, cam, bg, dpix=dpix)
            print("bg", bg, "variant", variant, {o: float("%.3e" % rel_l2(ours["grads"][o].reshape(np.asarray(g64[n]).shape), np.asarray(g64[n]))) for o, r, n in names})
_C.lib().gsr_debug_set_blend_variant(2)
