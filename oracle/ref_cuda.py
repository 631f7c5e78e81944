"""Harness around the UNMODIFIED reference CUDA build (oracle/_ref/ref_dgr_C.so, ref_knn_C.so).

TEST INFRASTRUCTURE ONLY (checker + the `--impl reference` arm of bench.py). The .so files are
produced by `make -C oracle ref` from the sources where they lie under /root/reference (never copied
into this repo); oracle/_ref/ is git-ignored but travels to the GPU box. This module calls the
reference's pybind entry points (rasterize_points.cu:35-217, spatial.cu:15-26) directly and decodes
the returned scratch blobs with the layout of rasterizer_impl.cu:155-194 -- no reference source needed.
"""
import importlib.util
import os
import sys

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF = os.path.join(_HERE, "_ref")
_mods = {}


def available():
    return os.path.exists(os.path.join(_REF, "ref_dgr_C.so"))


def _load(name):
    if name not in _mods:
        path = os.path.join(_REF, name + ".so")
        if not os.path.exists(path):
            raise ImportError(f"{path} missing: run `make -C oracle ref` where /root/reference exists")
        spec = importlib.util.spec_from_file_location(name, path)
        m = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(m)
        _mods[name] = m
    return _mods[name]


def dgr():
    return _load("ref_dgr_C")


def knn():
    return _load("ref_knn_C")


def _e(dev):
    return torch.empty(0, dtype=torch.float32, device=dev)


def forward(*, bg, means3D, opacities, viewmatrix, projmatrix, campos, tanfovx, tanfovy, H, W, sh_degree=0, shs=None,
            colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None, scale_modifier=1.0,
            prefiltered=False, debug=False):
    """Calls the reference's rasterize_gaussians; returns a dict incl. the three scratch blobs."""
    dev = means3D.device
    args = (bg, means3D, colors_precomp if colors_precomp is not None else _e(dev), opacities,
            scales if scales is not None else _e(dev), rotations if rotations is not None else _e(dev),
            float(scale_modifier), cov3D_precomp if cov3D_precomp is not None else _e(dev), viewmatrix, projmatrix,
            float(tanfovx), float(tanfovy), int(H), int(W), shs if shs is not None else _e(dev), int(sh_degree), campos,
            bool(prefiltered), bool(debug))
    n, color, radii, geom, binning, img = dgr().rasterize_gaussians(*args)
    return dict(num_rendered=n, color=color, radii=radii, geom=geom, binning=binning, img=img, args=args)


def backward(fw, dL_dpix):
    """Calls the reference's rasterize_gaussians_backward for a forward() result."""
    a = fw["args"]
    (bg, means3D, colors, opac, scales, rots, smod, cov3D, view, proj, tfx, tfy, H, W, sh, deg, campos, _pf, debug) = a
    out = dgr().rasterize_gaussians_backward(bg, means3D, fw["radii"], colors, scales, rots, smod, cov3D, view, proj,
                                             tfx, tfy, dL_dpix, sh, deg, campos, fw["geom"], fw["num_rendered"],
                                             fw["binning"], fw["img"], debug)
    names = ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
             "dL_drotations")
    return dict(zip(names, out))


def _carve(buf, pos, nbytes, dtype, shape):
    base = buf.data_ptr()
    start = ((base + pos + 127) & ~127) - base
    t = buf[start:start + nbytes].view(dtype).view(*shape)
    return t, start + nbytes


def decode(fw, P, W, H):
    """Slices the reference's blobs (rasterizer_impl.cu:155-194) into named tensors."""
    out = {}
    g, pos = fw["geom"], 0
    out["depths"], pos = _carve(g, pos, 4 * P, torch.float32, (P,))
    out["clamped"], pos = _carve(g, pos, 3 * P, torch.uint8, (P, 3))
    out["internal_radii"], pos = _carve(g, pos, 4 * P, torch.int32, (P,))
    out["means2D"], pos = _carve(g, pos, 8 * P, torch.float32, (P, 2))
    out["cov3D"], pos = _carve(g, pos, 24 * P, torch.float32, (P, 6))
    out["conic_opacity"], pos = _carve(g, pos, 16 * P, torch.float32, (P, 4))
    out["rgb"], pos = _carve(g, pos, 12 * P, torch.float32, (P, 3))
    out["tiles_touched"], pos = _carve(g, pos, 4 * P, torch.int32, (P,))
    N = W * H
    im, pos = fw["img"], 0
    out["final_T"], pos = _carve(im, pos, 4 * N, torch.float32, (N,))
    out["n_contrib"], pos = _carve(im, pos, 4 * N, torch.int32, (N,))
    out["ranges"], pos = _carve(im, pos, 8 * N, torch.int32, (N, 2))
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    out["ranges"] = out["ranges"][:tiles]
    R = fw["num_rendered"]
    if R > 0:
        b, pos = fw["binning"], 0
        out["point_list"], pos = _carve(b, pos, 4 * R, torch.int32, (R,))
        out["point_list_unsorted"], pos = _carve(b, pos, 4 * R, torch.int32, (R,))
        out["keys"], pos = _carve(b, pos, 8 * R, torch.int64, (R,))
        out["keys_unsorted"], pos = _carve(b, pos, 8 * R, torch.int64, (R,))
    return out


def mark_visible(means3D, viewmatrix, projmatrix):
    return dgr().mark_visible(means3D, viewmatrix, projmatrix)


def dist2(points):
    return knn().distCUDA2(points)


def autograd_function():
    """torch.autograd.Function around the reference's two pybind entry points, for bench.py's `--impl reference`
    arm (harness code: same call order as the reference's own Python wrapper, written for keyword settings)."""
    class _RefRasterize(torch.autograd.Function):
        @staticmethod
        def forward(ctx, means3D, means2D, shs, opacities, scales, rotations, st):
            fw = forward(means3D=means3D, shs=shs, opacities=opacities, scales=scales, rotations=rotations, **st)
            ctx.fw = fw
            ctx.mark_non_differentiable(fw["radii"])
            return fw["color"], fw["radii"]

        @staticmethod
        def backward(ctx, dL_dcolor, _):
            g = backward(ctx.fw, dL_dcolor.contiguous())
            ctx.fw = None
            return (g["dL_dmeans3D"], g["dL_dmeans2D"], g["dL_dsh"], g["dL_dopacity"], g["dL_dscales"],
                    g["dL_drotations"], None)
    return _RefRasterize


def ssim_available():
    return os.path.exists(os.path.join(_REF, "ref_fused_ssim_cuda.so"))


def fused_ssim_cuda():
    """The UNMODIFIED reference fused-ssim extension (ssim.cu + ext.cpp): fusedssim / fusedssim_backward."""
    return _load("ref_fused_ssim_cuda")


def fused_ssim_function():
    """mean-SSIM callable over the reference extension, shaped like the reference's fused_ssim/__init__.py:8-41 (harness)."""
    ext = fused_ssim_cuda()

    class _Map(torch.autograd.Function):
        @staticmethod
        def forward(ctx, C1, C2, img1, img2):
            m, d1, d2, d3 = ext.fusedssim(C1, C2, img1, img2, True)
            ctx.save_for_backward(img1.detach(), img2, d1, d2, d3)
            ctx.C1, ctx.C2 = C1, C2
            return m

        @staticmethod
        def backward(ctx, g):
            img1, img2, d1, d2, d3 = ctx.saved_tensors
            return None, None, ext.fusedssim_backward(ctx.C1, ctx.C2, img1, img2, g.contiguous(), d1, d2, d3), None

    return lambda a, b: _Map.apply(0.01 ** 2, 0.03 ** 2, a, b).mean()
