"""Generates tests/golden/ref_ssim.npz by RUNNING the reference's own Python SSIM (utils/loss_utils.py:99-124 `ssim` /
`_ssim`, the oracle of the reference's fused-ssim test, submodules/fused-ssim/tests/test.py:24-54,77) and l1_loss
(utils/loss_utils.py:41-53) on seeded inputs, with torch autograd for the gradients. Needs /root/reference (this
container only); the vectors it writes are what travels.   usage: python tests/golden/make_golden_ssim.py"""
import importlib.util
import os

import numpy as np
import torch

REF = "/root/reference/utils/loss_utils.py"
spec = importlib.util.spec_from_file_location("ref_loss_utils", REF)
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

out = {}
for tag, (B, CH, H, W), seed in (("a", (2, 3, 37, 53), 0), ("b", (1, 1, 11, 64), 1), ("c", (1, 3, 70, 33), 2)):
    g = torch.Generator().manual_seed(seed)
    # float32-representable inputs, evaluated by the reference code in float64 (stored as float32 to keep the file small)
    img1 = torch.rand(B, CH, H, W, generator=g, dtype=torch.float32).double().requires_grad_(True)
    img2 = torch.rand(B, CH, H, W, generator=g, dtype=torch.float32).double()
    val = ref.ssim(img1, img2)
    val.backward()
    out[f"{tag}_img1"], out[f"{tag}_img2"] = img1.detach().numpy().astype(np.float32), img2.numpy().astype(np.float32)
    out[f"{tag}_ssim"], out[f"{tag}_grad"] = val.detach().numpy(), img1.grad.numpy().astype(np.float32)
    # the training loss of train.py:114-121 with lambda_dssim = 0.2 (arguments/__init__.py), per image (CH, H, W)
    x = img1.detach()[0].clone().requires_grad_(True)
    loss = (1.0 - 0.2) * ref.l1_loss(x, img2[0]) + 0.2 * (1.0 - ref.ssim(x, img2[0]))
    loss.backward()
    out[f"{tag}_loss"], out[f"{tag}_loss_grad"] = loss.detach().numpy(), x.grad.numpy().astype(np.float32)
# the window itself
out["window"] = ref.gaussian(11, 1.5).numpy()
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_ssim.npz"), **out)
print({k: v.shape for k, v in out.items()})
