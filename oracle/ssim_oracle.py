"""CPU restatement of the reference's fused SSIM (submodules/fused-ssim/ssim.cu) -- TEST INFRASTRUCTURE ONLY.

numpy, float32 by default (float64 on request), written from the algorithm: 11-tap Gaussian window of sigma 1.5
(ssim.cu:9-19 holds the same numbers as constants; utils/loss_utils.py:66-80 `gaussian` builds them), separable
convolution with zero padding (ssim.cu:36-42 `get_pix_value`, :105-185), the per-pixel statistics and the SSIM map
(ssim.cu:225-258), the three partial-derivative maps written for training (ssim.cu:262-271) and the backward
(ssim.cu:272-366). Pinned by tests/test_cpu_ssim.py against (1) golden vectors produced by the reference's own Python
`ssim` (utils/loss_utils.py:99-124, the oracle of the reference's test, fused-ssim/tests/test.py:57-91) and its autograd
gradient, and (2) torch autograd of the conv2d formulation in float64.
"""
import numpy as np


# The 11 taps as both reference implementations hold them: the constants of ssim.cu:9-19 are bit-for-bit the fp32 window
# that loss_utils.py:66-80 builds (exp / sum evaluated in fp32); tests/test_cpu_ssim.py checks both statements.
_TAPS = (0.001028380123898387, 0.0075987582094967365, 0.036000773310661316, 0.10936068743467331, 0.21300552785396576,
         0.26601171493530273, 0.21300552785396576, 0.10936068743467331, 0.036000773310661316, 0.0075987582094967365,
         0.001028380123898387)


def gauss_window(dtype=np.float32):
    return np.array(_TAPS, dtype=np.float32).astype(dtype)


def _conv_sep(x, g):
    """'same' separable correlation with zero padding over the last two axes: horizontal first, then vertical,
    taps accumulated in ascending order (ssim.cu:105-185)."""
    H, W = x.shape[-2:]
    xp = np.zeros(x.shape[:-1] + (W + 10,), dtype=x.dtype)
    xp[..., 5:5 + W] = x
    h = np.zeros_like(x)
    for k in range(11):
        h = h + g[k] * xp[..., k:k + W]
    hp = np.zeros(x.shape[:-2] + (H + 10, W), dtype=x.dtype)
    hp[..., 5:5 + H, :] = h
    v = np.zeros_like(x)
    for k in range(11):
        v = v + g[k] * hp[..., k:k + H, :]
    return v


def ssim_forward(img1, img2, C1=0.01 ** 2, C2=0.03 ** 2, dtype=np.float32):
    """-> (ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12), each [B, CH, H, W]."""
    a, b = np.asarray(img1, dtype), np.asarray(img2, dtype)
    g = gauss_window(dtype)
    C1, C2 = dtype(C1), dtype(C2)
    mu1, mu2 = _conv_sep(a, g), _conv_sep(b, g)
    sigma1_sq = _conv_sep(a * a, g) - mu1 * mu1
    sigma2_sq = _conv_sep(b * b, g) - mu2 * mu2
    sigma12 = _conv_sep(a * b, g) - mu1 * mu2
    two = dtype(2.0)
    C = two * mu1 * mu2 + C1
    D = two * sigma12 + C2
    A = mu1 * mu1 + mu2 * mu2 + C1
    B = sigma1_sq + sigma2_sq + C2
    m = (C * D) / (A * B)
    dm_dmu1 = (mu2 * two * D) / (A * B) - (mu2 * two * C) / (A * B) - (mu1 * two * C * D) / (A * A * B) + \
              (mu1 * two * C * D) / (A * B * B)
    dm_dsigma1_sq = (-C * D) / (A * B * B)
    dm_dsigma12 = (two * C) / (A * B)
    return m, dm_dmu1, dm_dsigma1_sq, dm_dsigma12


def ssim_backward(img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12, dtype=np.float32):
    """dL/dimg1 (ssim.cu:301-365)."""
    a, b, d = np.asarray(img1, dtype), np.asarray(img2, dtype), np.asarray(dL_dmap, dtype)
    g = gauss_window(dtype)
    t1 = _conv_sep(np.asarray(dm_dmu1, dtype) * d, g)
    t2 = _conv_sep(np.asarray(dm_dsigma1_sq, dtype) * d, g)
    t3 = _conv_sep(np.asarray(dm_dsigma12, dtype) * d, g)
    return t1 + a * dtype(2.0) * t2 + b * t3


def fused_ssim(img1, img2, padding="same", dtype=np.float32):
    """mean SSIM and its gradient w.r.t. img1 (fused_ssim/__init__.py:8-41)."""
    m, d1, d2, d3 = ssim_forward(img1, img2, dtype=dtype)
    if padding == "valid":
        mv = m[:, :, 5:-5, 5:-5]
        dL = np.zeros_like(m)
        dL[:, :, 5:-5, 5:-5] = dtype(1.0 / mv.size)
        val = mv.astype(np.float64).mean()
    else:
        dL = np.full_like(m, dtype(1.0 / m.size))
        val = m.astype(np.float64).mean()
    return val, ssim_backward(img1, img2, dL, d1, d2, d3, dtype=dtype)


def l1_ssim_loss(image, gt, lambda_dssim=0.2, dtype=np.float32):
    """train.py:114-121: (1-l)*mean|image-gt| + l*(1-mean ssim); -> (loss, dL/dimage)."""
    a, b = np.asarray(image, dtype), np.asarray(gt, dtype)
    if a.ndim == 3:
        a, b = a[None], b[None]
    val, gs = fused_ssim(a, b, dtype=dtype)
    l1 = np.abs(a.astype(np.float64) - b.astype(np.float64)).mean()
    loss = (1.0 - lambda_dssim) * l1 + lambda_dssim * (1.0 - val)
    grad = dtype((1.0 - lambda_dssim) / a.size) * np.sign(a - b) - dtype(lambda_dssim) * gs
    return loss, grad.reshape(np.asarray(image).shape)
