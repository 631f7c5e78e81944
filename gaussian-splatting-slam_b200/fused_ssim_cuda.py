"""`fused_ssim_cuda` with the reference extension's two entry points (submodules/fused-ssim/ext.cpp:5-6,
ssim.cu:368-444): fusedssim(C1, C2, img1, img2, train) -> (ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12) and
fusedssim_backward(C1, C2, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12) -> dL_dimg1.
The reference's own fused_ssim/__init__.py imports exactly these, so it runs unchanged on top of this module.
Everything runs in libgsr.so (gsr_ssim_forward / gsr_ssim_backward); CUDA tensors only, no fallback."""
import ctypes as C

import torch
from diff_gaussian_rasterization import _C as _gsr

_ready = False


def _lib():
    global _ready
    L = _gsr.lib()
    if not _ready:
        L.gsr_ssim_forward.restype = C.c_int
        L.gsr_ssim_forward.argtypes = [C.c_int] * 4 + [C.c_float] * 2 + [C.c_void_p] * 8
        L.gsr_ssim_backward.restype = C.c_int
        L.gsr_ssim_backward.argtypes = [C.c_int] * 4 + [C.c_float] * 2 + [C.c_void_p] * 3 + [C.c_float] * 2 + [C.c_void_p] * 6
        _ready = True
    return L


def _check(rc, what):
    if rc != 0:
        raise RuntimeError(f"{what}: {_gsr.lib().gsr_error_string(rc).decode()} (code {rc})")


def _prep(t, name):
    if not t.is_cuda:
        raise RuntimeError(f"fused_ssim: {name} must be a CUDA tensor (there is no CPU path)")
    if t.dim() != 4:
        raise RuntimeError(f"fused_ssim: {name} must be [B, CH, H, W]")
    return t.float().contiguous()


def _ptr(t):
    return t.data_ptr() if t is not None and t.numel() > 0 else None


def ssim_forward(C1, C2, img1, img2, train=True, want_map=True, sums=None):
    """gsr_ssim_forward; `sums` (cuda float64[2], caller-zeroed) additionally receives sum(ssim) and sum|img1-img2|."""
    img1, img2 = _prep(img1, "img1"), _prep(img2, "img2")
    if img1.shape != img2.shape:
        raise RuntimeError("fused_ssim: img1 and img2 must have the same shape")
    B, CH, H, W = img1.shape
    dev = img1.device
    ssim_map = torch.empty_like(img1) if want_map else None
    d1 = torch.empty_like(img1) if train else torch.empty(0)
    d2 = torch.empty_like(img1) if train else torch.empty(0)
    d3 = torch.empty_like(img1) if train else torch.empty(0)
    if img1.numel() > 0:
        with torch.cuda.device(dev):
            _check(_lib().gsr_ssim_forward(B, CH, H, W, C1, C2, img1.data_ptr(), img2.data_ptr(), _ptr(ssim_map), _ptr(d1),
                                           _ptr(d2), _ptr(d3), _ptr(sums), torch.cuda.current_stream(dev).cuda_stream),
                   "gsr_ssim_forward")
    return ssim_map, d1, d2, d3


def ssim_backward(C1, C2, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12, map_scale=0.0, l1_scale=0.0,
                  upstream=None):
    """gsr_ssim_backward; dL_dmap=None selects the uniform (fused-loss) mode with map_scale / l1_scale, both multiplied
    on the device by `upstream` (a one-element cuda float tensor) when given."""
    img1, img2 = _prep(img1, "img1"), _prep(img2, "img2")
    B, CH, H, W = img1.shape
    dev = img1.device
    if dL_dmap is not None:
        dL_dmap = _prep(dL_dmap, "dL_dmap")
        if dL_dmap.shape != img1.shape:
            raise RuntimeError("fused_ssim: dL_dmap must have the shape of img1")
    out = torch.empty_like(img1)
    if img1.numel() > 0:
        with torch.cuda.device(dev):
            _check(_lib().gsr_ssim_backward(B, CH, H, W, C1, C2, img1.data_ptr(), img2.data_ptr(), _ptr(dL_dmap), map_scale,
                                            l1_scale, _ptr(upstream), dm_dmu1.contiguous().data_ptr(), dm_dsigma1_sq.contiguous().data_ptr(),
                                            dm_dsigma12.contiguous().data_ptr(), out.data_ptr(),
                                            torch.cuda.current_stream(dev).cuda_stream), "gsr_ssim_backward")
    return out


# ---- the reference extension's surface
def fusedssim(C1, C2, img1, img2, train):
    return ssim_forward(C1, C2, img1, img2, train=bool(train), want_map=True)


def fusedssim_backward(C1, C2, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12):
    return ssim_backward(C1, C2, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12)
