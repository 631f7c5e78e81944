"""Drop-in replacement for the `diff_gaussian_rasterization` package used by
gaussian_renderer/__init__.py:14,36-109 of Runnlion/gaussian-splatting-slam, backed by hand-written
sm_100a kernels (libgsr.so, include/gsr.h) instead of the vendored CUDA rasterizer.

API parity (reference: submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py):
  * GaussianRasterizationSettings  (:157-169)  same 12 fields in the same order, plus a 13th
    `antialiasing: bool = False`, which the reference's caller passes (gaussian_renderer/__init__.py:49)
    although the vendored package lacks it.
  * GaussianRasterizer             (:171-221)  same constructor / markVisible / forward keywords and the
    same two validation errors.
  * _RasterizeGaussians            (:44-155)   same positional forward signature; backward returns the
    same 9-tuple in input order.
  * forward returns (color[3,H,W], radii[P] int32, invdepth[1,H,W]) -- the 3-tuple the caller unpacks
    (gaussian_renderer/__init__.py:101). Set RETURN_INVDEPTH = False (or env GSR_RETURN_INVDEPTH=0)
    to get the vendored 2-tuple (color, radii).
  * SparseGaussianAdam             exported (lazily, from gsr_optim): the caller probes for it (train.py:37-41,
    gaussian_model.py:24-27) and, when present, renders with the separate_sh call shape
    (gaussian_renderer/__init__.py:81-99: dc= and shs= as two tensors), which forward() takes natively.
    GSR_EXPORT_SPARSE_ADAM=0 hides it again (callers then pass the concatenated features).
"""
import os
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _C

RETURN_INVDEPTH = os.environ.get("GSR_RETURN_INVDEPTH", "1") != "0"
EXPORT_SPARSE_ADAM = os.environ.get("GSR_EXPORT_SPARSE_ADAM", "1") != "0"


def __getattr__(name):          # lazy: gsr_optim imports this package for the library handle
    if name == "SparseGaussianAdam" and EXPORT_SPARSE_ADAM:
        from gsr_optim import SparseGaussianAdam
        return SparseGaussianAdam
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")


def cpu_deep_copy_tuple(input_tuple):
    return tuple(item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings, grad_sinks=None, dc=None):
    out = _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                    cov3Ds_precomp, raster_settings, grad_sinks, dc)
    return out if RETURN_INVDEPTH else out[:2]


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings, grad_sinks=None, dc=None):
        rs = raster_settings
        if dc is None:
            dc = torch.Tensor([])
        antialiasing = bool(getattr(rs, "antialiasing", False))
        lease = _C.ScratchLease()                    # scratch blobs go back to the pool when this context dies
        args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh,
                rs.sh_degree, rs.campos, rs.prefiltered, rs.debug, antialiasing, True, lease, dc)
        if rs.debug:
            cpu_args = cpu_deep_copy_tuple(args[:-3] + (dc,))    # snapshot before anything can be corrupted
            try:
                res = _C.rasterize_gaussians(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise ex
        else:
            res = _C.rasterize_gaussians(*args)
        num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, invdepth = res

        ctx.raster_settings = rs
        ctx.antialiasing = antialiasing
        ctx.num_rendered = num_rendered
        ctx.grad_sinks = grad_sinks
        ctx.scratch_lease = lease
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, opacities, dc,
                              geomBuffer, binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii)
        ctx.set_materialize_grads(False)
        return color, radii, invdepth

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii=None, grad_out_invdepth=None):
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, opacities, dc, geomBuffer, binningBuffer,
         imgBuffer) = ctx.saved_tensors
        if grad_out_color is None:
            grad_out_color = torch.zeros((3, rs.image_height, rs.image_width), dtype=torch.float32,
                                         device=means3D.device)
        args = (rs.bg, means3D, radii, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color, grad_out_invdepth, sh,
                rs.sh_degree, rs.campos, geomBuffer, ctx.num_rendered, binningBuffer, imgBuffer, ctx.antialiasing,
                rs.debug, False, ctx.grad_sinks, ctx.scratch_lease, dc)
        if rs.debug:
            cpu_args = cpu_deep_copy_tuple(args[:-4] + (dc,))
            try:
                res = _C.rasterize_gaussians_backward(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise ex
        else:
            res = _C.rasterize_gaussians_backward(*args)
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh, grad_scales,
         grad_rotations) = res[:8]
        grad_dc = res[8] if dc.numel() != 0 else None
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities, grad_scales, grad_rotations,
                grad_cov3Ds_precomp, None, None, grad_dc)


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    antialiasing: bool = False


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """Boolean mask of the points that pass the near-plane test of this camera."""
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, grad_sinks=None, dc=None):
        """Reference keywords (diff_gaussian_rasterization/__init__.py:187). Supersets:
        `dc` -- the caller's separate_sh call shape (gaussian_renderer/__init__.py:81-99): the degree-0 SH coefficient as
        its own [P,1,3] tensor and `shs` = the higher coefficients [P,M-1,3], read in place by the kernels (no
        torch.cat of the features, gaussian_model.py:113-117, and its backward split);
        `grad_sinks` -- a dict of fp32 tensors keyed by input name; the backward then ADDS this view's gradients into
        them inside the kernel and autograd receives None for those inputs (multi-view batches); with the extra entry
        `"overwrite": True` the sinks are written in full instead (first view of a batch: no cleared sinks needed)."""
        rs = self.raster_settings
        if ((shs is None and dc is None) and colors_precomp is None) or \
                ((shs is not None or dc is not None) and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        empty = torch.Tensor([])
        shs = empty if shs is None else shs
        colors_precomp = empty if colors_precomp is None else colors_precomp
        scales = empty if scales is None else scales
        rotations = empty if rotations is None else rotations
        cov3D_precomp = empty if cov3D_precomp is None else cov3D_precomp
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                                   rs, grad_sinks, dc)
