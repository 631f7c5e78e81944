"""CPU tests for the SSIM row: the numpy restatement oracle/ssim_oracle.py is pinned against golden
vectors produced by the reference's own Python SSIM + autograd (tests/golden/make_golden_ssim.py) and against torch
autograd of the conv2d formulation; the Python surface mirrors the reference package."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import ssim_oracle

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_ssim.npz"))


def test_window_matches_reference_constants():
    # the reference's Python window (loss_utils.py:66-80, fp32) and the constants of ssim.cu:9-19 are the same bits;
    # an exact exp/normalise in float64 rounds two taps one ulp differently
    assert np.array_equal(ssim_oracle.gauss_window(np.float32), GOLD["window"])
    x = np.arange(11, dtype=np.float64)
    g = np.exp(-((x - 5.0) ** 2) / (2.0 * 1.5 ** 2))
    assert np.allclose(g / g.sum(), GOLD["window"].astype(np.float64), rtol=0, atol=3e-8)


@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_oracle_matches_reference_python_ssim_value_and_gradient(tag):
    img1, img2 = GOLD[f"{tag}_img1"], GOLD[f"{tag}_img2"]
    val64, grad64 = ssim_oracle.fused_ssim(img1, img2, dtype=np.float64)
    # float64 on both sides; what is left is the reference building its 2-D window as an fp32 outer product
    # (loss_utils.py:95-96) where the oracle (like ssim.cu) applies the fp32 taps separably
    assert abs(val64 - float(GOLD[f"{tag}_ssim"])) < 5e-8
    assert np.allclose(grad64, GOLD[f"{tag}_grad"], rtol=1e-5, atol=1e-9)          # golden stored as float32
    val32, grad32 = ssim_oracle.fused_ssim(img1, img2, dtype=np.float32)
    # the bar of the reference's own test (torch.isclose defaults, fused-ssim/tests/test.py:82,90)
    assert np.isclose(val32, float(GOLD[f"{tag}_ssim"]), rtol=1e-5, atol=1e-8)
    assert np.allclose(grad32, GOLD[f"{tag}_grad"], rtol=1e-3, atol=2e-8)


@pytest.mark.parametrize("tag", ["a", "c"])
def test_oracle_training_loss_matches_reference_formula(tag):
    img1, img2 = GOLD[f"{tag}_img1"][0], GOLD[f"{tag}_img2"][0]
    loss, grad = ssim_oracle.l1_ssim_loss(img1, img2, 0.2, dtype=np.float64)
    assert abs(loss - float(GOLD[f"{tag}_loss"])) < 5e-8
    assert np.allclose(grad, GOLD[f"{tag}_loss_grad"], rtol=1e-5, atol=1e-9)


def test_oracle_valid_padding_and_arbitrary_upstream_vs_autograd_float64():
    g = torch.Generator().manual_seed(5)
    B, CH, H, W = 1, 2, 29, 31
    x = torch.rand(B, CH, H, W, generator=g, dtype=torch.float64, requires_grad=True)
    y = torch.rand(B, CH, H, W, generator=g, dtype=torch.float64)
    up = torch.randn(B, CH, H, W, generator=g, dtype=torch.float64)
    w1 = torch.tensor(ssim_oracle.gauss_window(np.float64))
    win = (w1[:, None] * w1[None, :]).expand(CH, 1, 11, 11).contiguous()

    def conv(t):
        return F.conv2d(t, win, padding=5, groups=CH)
    mu1, mu2 = conv(x), conv(y)
    s1, s2, s12 = conv(x * x) - mu1 * mu1, conv(y * y) - mu2 * mu2, conv(x * y) - mu1 * mu2
    m = ((2 * mu1 * mu2 + 1e-4) * (2 * s12 + 9e-4)) / ((mu1 * mu1 + mu2 * mu2 + 1e-4) * (s1 + s2 + 9e-4))
    (m * up).sum().backward()
    mo, d1, d2, d3 = ssim_oracle.ssim_forward(x.detach().numpy(), y.numpy(), dtype=np.float64)
    assert np.allclose(mo, m.detach().numpy(), rtol=1e-10, atol=1e-12)
    go = ssim_oracle.ssim_backward(x.detach().numpy(), y.numpy(), up.numpy(), d1, d2, d3, dtype=np.float64)
    assert np.allclose(go, x.grad.numpy(), rtol=1e-8, atol=1e-11)
    # "valid" = crop 5 pixels per side (fused_ssim/__init__.py:13-14,27-29)
    val, _ = ssim_oracle.fused_ssim(x.detach().numpy(), y.numpy(), "valid", dtype=np.float64)
    assert abs(val - float(m.detach()[:, :, 5:-5, 5:-5].mean())) < 1e-12


def test_python_surface_matches_reference_package():
    import fused_ssim
    import fused_ssim_cuda
    assert fused_ssim.allowed_padding == ["same", "valid"]
    assert callable(fused_ssim.fused_ssim) and issubclass(fused_ssim.FusedSSIMMap, torch.autograd.Function)
    assert callable(fused_ssim_cuda.fusedssim) and callable(fused_ssim_cuda.fusedssim_backward)
    with pytest.raises(RuntimeError, match="CUDA tensor"):       # no CPU fallback
        fused_ssim.fused_ssim(torch.rand(1, 3, 16, 16), torch.rand(1, 3, 16, 16))
