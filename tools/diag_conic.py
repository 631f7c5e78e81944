"""Diagnostic (GPU box): where do our preprocess outputs differ from the reference CUDA build?"""
import sys, os, math
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "gaussian-splatting-slam_b200"), os.path.join(ROOT, "tests")]
import numpy as np, torch, scenes
from oracle import ref_cuda
from test_gpu_parity import _run_ours, _ref_forward
for name, sc, cam in (("c1", scenes.make_c1(), scenes.c1_camera()),
                      ("sh3", scenes.small_scene(P=3000, deg=3, spread=1.5), scenes.look_at_camera(200, 123, math.radians(70), (0.3, -0.2, -3.5)))):
    ours = _run_ours(sc, cam, (0, 0, 0))
    fw = _ref_forward(sc, cam, (0, 0, 0))
    rd = {k: v.cpu().numpy() for k, v in ref_cuda.decode(fw, sc.means3D.shape[0], cam.W, cam.H).items()}
    st = ours["state"]
    vis = ours["radii"] > 0
    for key in ("conic_opacity", "rgb", "means2D", "depths"):
        a, b = st[key][vis], rd[key][vis]
        a = a.reshape(len(a), -1); b = b.reshape(len(b), -1)
        ai, bi = a.view(np.int32).astype(np.int64), b.view(np.int32).astype(np.int64)
        d = np.abs(ai - bi)
        print(name, key, "cols mism frac:", (d > 0).mean(0), "max ulp:", d.max(0))
    d = np.abs(ours["color"] - fw["color"].cpu().numpy())
    print(name, "colour max-abs", d.max(), "n_contrib equal", np.array_equal(st["n_contrib"], rd["n_contrib"]))
    bad = np.flatnonzero((st["conic_opacity"][vis] != rd["conic_opacity"][vis]).any(1))[:5]
    for i in bad:
        print(" ours", st["conic_opacity"][vis][i].view(np.uint32), "ref", rd["conic_opacity"][vis][i].view(np.uint32))
