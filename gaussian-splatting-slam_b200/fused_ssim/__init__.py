"""`fused_ssim` with the reference package's surface (submodules/fused-ssim/fused_ssim/__init__.py:1-41):
allowed_padding, FusedSSIMMap (autograd over the SSIM map, "same" / "valid" padding) and
fused_ssim(img1, img2, padding="same", train=True) -> mean SSIM.

Added on top: fused_l1_ssim_loss(image, gt, lambda_dssim) -- the training loss of
train.py:114-121, (1-l)*mean|image-gt| + l*(1-mean(ssim)), in one forward and one backward kernel: the SSIM map and
dL/dmap never touch HBM and the L1 term rides along.
"""
import torch

from fused_ssim_cuda import fusedssim, fusedssim_backward, ssim_backward, ssim_forward

allowed_padding = ["same", "valid"]


class FusedSSIMMap(torch.autograd.Function):
    @staticmethod
    def forward(ctx, C1, C2, img1, img2, padding="same", train=True):
        ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12 = fusedssim(C1, C2, img1, img2, train)
        if padding == "valid":
            ssim_map = ssim_map[:, :, 5:-5, 5:-5]
        ctx.save_for_backward(img1.detach(), img2, dm_dmu1, dm_dsigma1_sq, dm_dsigma12)
        ctx.C1, ctx.C2, ctx.padding = C1, C2, padding
        return ssim_map

    @staticmethod
    def backward(ctx, opt_grad):
        img1, img2, dm_dmu1, dm_dsigma1_sq, dm_dsigma12 = ctx.saved_tensors
        dL_dmap = opt_grad
        if ctx.padding == "valid":
            dL_dmap = torch.zeros_like(img1)
            dL_dmap[:, :, 5:-5, 5:-5] = opt_grad
        grad = fusedssim_backward(ctx.C1, ctx.C2, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12)
        return None, None, grad, None, None, None


def fused_ssim(img1, img2, padding="same", train=True):
    C1 = 0.01 ** 2
    C2 = 0.03 ** 2
    assert padding in allowed_padding
    map = FusedSSIMMap.apply(C1, C2, img1, img2, padding, train)
    return map.mean()


class _FusedL1SSIMLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, gt, lambda_dssim):
        img1 = image if image.dim() == 4 else image.unsqueeze(0)
        img2 = gt if gt.dim() == 4 else gt.unsqueeze(0)
        sums = torch.zeros(2, dtype=torch.float64, device=img1.device)
        _, d1, d2, d3 = ssim_forward(0.01 ** 2, 0.03 ** 2, img1, img2, train=True, want_map=False, sums=sums)
        n = float(img1.numel())
        ctx.save_for_backward(img1.detach(), img2, d1, d2, d3)
        ctx.lam, ctx.n, ctx.squeeze = float(lambda_dssim), n, image.dim() != 4
        mean = sums / n                                   # [mean ssim, mean |image - gt|]
        loss = (1.0 - ctx.lam) * mean[1] + ctx.lam * (1.0 - mean[0])
        return loss.to(torch.float32)

    @staticmethod
    def backward(ctx, g):
        img1, img2, d1, d2, d3 = ctx.saved_tensors
        up = g.detach().to(torch.float32).reshape(1).contiguous()     # stays on the device: no host sync
        grad = ssim_backward(0.01 ** 2, 0.03 ** 2, img1, img2, None, d1, d2, d3, map_scale=-ctx.lam / ctx.n,
                             l1_scale=(1.0 - ctx.lam) / ctx.n, upstream=up)
        return (grad[0] if ctx.squeeze else grad), None, None


def fused_l1_ssim_loss(image, gt, lambda_dssim=0.2):
    """(1 - lambda) * l1_loss(image, gt) + lambda * (1 - fused_ssim(image, gt)) with gradient w.r.t. `image` only."""
    return _FusedL1SSIMLoss.apply(image, gt, lambda_dssim)
