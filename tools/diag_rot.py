"""Diagnostic (GPU box): where does the rotation-gradient error of the AA / inverse-depth case come from?"""
import sys, os, math
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "gaussian-splatting-slam_b200"), os.path.join(ROOT, "tests")]
import numpy as np, torch, scenes
from oracle import c_oracle
from util import oracle_scene, rel_l2
from test_gpu_parity import _run_ours, _f64_autograd_grads
sc = scenes.small_scene(P=2500, deg=3, spread=1.2)
cam = scenes.c5_camera(3)
sc = sc._replace(means3D=sc.means3D + torch.tensor([0.0, 0.0, 3.0]))
for aa, bg, seed in ((True, (0., 0., 0.), 0), (False, (0.2, 0.3, 0.4), 5)):
    rng = np.random.default_rng(seed)
    dpix = rng.standard_normal((3, cam.H, cam.W)).astype(np.float32)
    dinv = rng.standard_normal((1, cam.H, cam.W)).astype(np.float32)
    ours = _run_ours(sc, cam, bg, antialiasing=aa, dpix=dpix, dinv=dinv)
    S = oracle_scene(sc, cam, bg, antialiasing=aa)
    fw = c_oracle.forward(S); g = c_oracle.backward(S, fw, dpix, dinv)
    g64 = _f64_autograd_grads(sc, cam, bg, aa, dpix, dinv)
    for k, ok in (("dL_drot", "rotations"), ("dL_dscale", "scales"), ("dL_dmean3D", "means3D"), ("dL_dopacity", "opacities")):
        t = g64[k]; a = ours["grads"][ok].reshape(t.shape); c = g[k].reshape(t.shape)
        print(f"aa={aa} {k}: ours-vs-f64 {rel_l2(a, t):.3e}  C-vs-f64 {rel_l2(c, t):.3e}  ours-vs-C {rel_l2(a, c):.3e}")
    t = g64["dL_drot"]; a = ours["grads"]["rotations"]; c = g["dL_drot"]
    ea = np.linalg.norm(a - t, axis=1); ec = np.linalg.norm(c - t, axis=1)
    top = np.argsort(-ea)[:6]
    print(" norm truth", np.linalg.norm(t), "top ours errs", ea[top], "C errs same idx", ec[top], "idx", top)
    print(" radii", ours["radii"][top], "scales", sc.scales.numpy()[top].max(1), "depth", fw["depths"][top])
    print(" err energy in top6:", (ea[top] ** 2).sum() / (ea ** 2).sum())
