import sys, os, math
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "gaussian-splatting-slam_b200"), os.path.join(ROOT, "tests")]
import numpy as np, torch
import test_gpu_parity as T
import scenes
from diff_gaussian_rasterization import _C
from util import rel_l2
g = torch.Generator().manual_seed(21)
P = 4000
sc = scenes.small_scene(P=P, deg=2, spread=1.4, seed=21)
long_axis = torch.randint(0, 3, (P,), generator=g)
scales = torch.exp(math.log(0.002) + torch.rand(P, 3, generator=g) * (math.log(0.008) - math.log(0.002)))
scales[torch.arange(P), long_axis] = 0.15 + 0.5 * torch.rand(P, generator=g)
opac = torch.sigmoid(torch.randn(P, 1, generator=g) * 3.0 - 3.0)
sc = scenes.Scene(sc.means3D, scales, sc.rotations, opac, sc.shs, sc.sh_degree)
cam = scenes.look_at_camera(320, 200, math.radians(65), (0.2, 0.1, -3.2))
dpix = np.random.default_rng(5).standard_normal((3, cam.H, cam.W)).astype(np.float32)
runs = {}
for tag, on in (("on1", True), ("on2", True), ("off1", False), ("off2", False)):
    _C.set_exact_cull(on)
    runs[tag] = T._run_ours(sc, cam, (0, 0, 0), dpix=dpix)["grads"]
_C.set_exact_cull(True)
for a, b in (("on1", "on2"), ("off1", "off2"), ("on1", "off1")):
    print(a, b, {k: float("%.2e" % rel_l2(runs[a][k], runs[b][k])) for k in runs[a] if runs[a][k] is not None and np.linalg.norm(runs[b][k]) > 0})
