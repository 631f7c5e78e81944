"""Profiling driver (GPU box): a few fused L1/SSIM loss forward+backward passes at 1080p, nothing else."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "gaussian-splatting-slam_b200")]
import torch
from fused_ssim import fused_l1_ssim_loss
x = torch.rand(3, 1080, 1920, device="cuda", requires_grad=True)
y = torch.rand(3, 1080, 1920, device="cuda")
for _ in range(3):
    fused_l1_ssim_loss(x, y, 0.2).backward()
torch.cuda.synchronize()
print("ok")
