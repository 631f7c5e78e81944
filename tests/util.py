"""Shared helpers for the tests: scene -> oracle / CUDA argument plumbing and error metrics."""
import numpy as np
import torch

import scenes
from oracle import c_oracle


def rel_l2(a, b):
    a = np.asarray(a, dtype=np.float64).reshape(-1)
    b = np.asarray(b, dtype=np.float64).reshape(-1)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def oracle_scene(sc: scenes.Scene, cam: scenes.Camera, bg, antialiasing=False, colors_precomp=None,
                 cov3D_precomp=None, scale_modifier=1.0):
    kw = dict(W=cam.W, H=cam.H, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=np.asarray(bg, np.float32),
              means3D=sc.means3D.numpy(), opacities=sc.opacities.numpy(), viewmatrix=cam.viewmatrix.numpy(),
              projmatrix=cam.projmatrix.numpy(), campos=cam.campos.numpy(), sh_degree=sc.sh_degree,
              antialiasing=antialiasing, scale_modifier=scale_modifier)
    if colors_precomp is None:
        kw["shs"] = sc.shs.numpy()
    else:
        kw["colors_precomp"] = colors_precomp.numpy()
    if cov3D_precomp is None:
        kw.update(scales=sc.scales.numpy(), rotations=sc.rotations.numpy())
    else:
        kw["cov3D_precomp"] = cov3D_precomp.numpy()
    return c_oracle.Scene(**kw)


def cuda_settings(cam: scenes.Camera, bg, sh_degree, dev, antialiasing=False, debug=False, scale_modifier=1.0):
    from diff_gaussian_rasterization import GaussianRasterizationSettings
    return GaussianRasterizationSettings(
        image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
        bg=torch.as_tensor(bg, dtype=torch.float32, device=dev), scale_modifier=scale_modifier,
        viewmatrix=cam.viewmatrix.to(dev), projmatrix=cam.projmatrix.to(dev), sh_degree=sh_degree,
        campos=cam.campos.to(dev), prefiltered=False, debug=debug, antialiasing=antialiasing)


def summarize_mismatch(name, a, b):
    a = np.asarray(a); b = np.asarray(b)
    bad = np.flatnonzero(a.reshape(-1) != b.reshape(-1))
    return f"{name}: {bad.size}/{a.size} differ; first {bad[:8].tolist()} ours={a.reshape(-1)[bad[:8]].tolist()} ref={b.reshape(-1)[bad[:8]].tolist()}"
