"""Generates tests/golden/ref_python_substeps.npz by RUNNING the reference's own Python restatements of
rasterizer sub-steps (they exist in-tree because pipe.convert_SHs_python / compute_cov3D_python can
replace the CUDA versions, gaussian_renderer/__init__.py:63-79):
    utils/sh_utils.py:57-112        eval_sh            -> pins forward.cu:20-71
    utils/general_utils.py:78-110   build_rotation / build_scaling_rotation / strip_symmetric
    scene/gaussian_model.py:32-36   build_covariance_from_scaling_rotation -> pins forward.cu:118-152
    utils/graphics_utils.py:38-71   getWorld2View2 / getProjectionMatrix   -> pins scenes.py cameras
Run here (needs /root/reference, CPU only):  python tests/golden/make_golden.py
The reference hard-codes device="cuda" in general_utils; torch.zeros is wrapped to drop that keyword so
the unmodified functions run on the CPU. The vectors are committed; tests never read /root/reference.
"""
import importlib.util
import math
import os
import sys

import numpy as np
import torch

REF = os.environ.get("GSR_REFERENCE", "/root/reference")


def load(name, rel):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def main():
    sh_utils = load("ref_sh_utils", "utils/sh_utils.py")
    gen = load("ref_general_utils", "utils/general_utils.py")
    gfx = load("ref_graphics_utils", "utils/graphics_utils.py")
    _zeros = torch.zeros

    def zeros_cpu(*a, **k):
        k.pop("device", None)
        return _zeros(*a, **k)
    gen.torch.zeros = zeros_cpu      # general_utils.py:61,79,105 pass device="cuda"

    g = torch.Generator().manual_seed(1234)
    n = 64
    dirs = torch.randn(n, 3, generator=g)
    dirs = dirs / dirs.norm(dim=1, keepdim=True)
    shs = torch.randn(n, 16, 3, generator=g)                       # rasterizer layout [P,M,3]
    out = {"dirs": dirs.numpy(), "shs": shs.numpy()}
    for deg in range(4):
        # caller layout: shs_view = features.transpose(1,2) -> [P,3,M]  (gaussian_renderer/__init__.py:75)
        out[f"sh2rgb_deg{deg}"] = sh_utils.eval_sh(deg, shs.transpose(1, 2), dirs).numpy()
    scales = torch.exp(torch.randn(n, 3, generator=g) * 0.7 - 2.0)
    rots = torch.randn(n, 4, generator=g)
    rots = rots / rots.norm(dim=1, keepdim=True)
    for mod in (1.0, 0.7):
        L = gen.build_scaling_rotation(mod * scales, rots)        # gaussian_model.py:33
        out[f"cov3D_mod{mod}"] = gen.strip_symmetric(L @ L.transpose(1, 2)).numpy()
    out["scales"], out["rots"] = scales.numpy(), rots.numpy()
    gen.torch.zeros = _zeros
    fovx, fovy = math.radians(60.0), math.radians(35.0)
    out["proj"] = gfx.getProjectionMatrix(znear=0.01, zfar=100.0, fovX=fovx, fovY=fovy).numpy()
    out["fov"] = np.array([fovx, fovy])
    R = np.array([[0.36, 0.48, -0.8], [-0.8, 0.6, 0.0], [0.48, 0.64, 0.6]])
    T = np.array([0.3, -0.2, 4.0])
    out["w2v_R"], out["w2v_T"] = R, T
    out["w2v"] = gfx.getWorld2View2(R, T)
    np.savez(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_python_substeps.npz"), **out)
    print("wrote", len(out), "arrays")


if __name__ == "__main__":
    main()
