"""Generates tests/golden/ref_densify.npz by RUNNING the reference's own GaussianModel (scene/gaussian_model.py) on the
CPU: training_setup (:155-183), a few torch.optim.Adam steps to fill the moments, add_densification_stats (:431-433)
with the caller's max_radii2D update (train.py:159), densify_and_prune (:412-429), prune_points (:307-322) and
reset_opacity (:226-229). The reference source is imported unmodified from /root/reference (this container only; the
vectors are what travels). Two shims make it importable and CPU-runnable, neither touches its arithmetic:
  * stub modules for plyfile / simple_knn._C (imported at module top, not used by these methods);
  * torch.zeros / torch.ones / torch.empty ignore device="cuda" (the model hard-codes it, :357-365, general_utils.py:83);
  * torch.normal(mean=, std=) is replaced by randn * std + mean with the randn recorded, after asserting that this is
    bit-identical to the real torch.normal under the same seed -- the recorded noise is what the device path is fed.
usage: python tests/golden/make_golden_densify.py"""
import os
import sys
import types
from argparse import Namespace

import numpy as np
import torch
from torch import nn

REF = "/root/reference"
sys.path.insert(0, REF)
for name in ("plyfile", "simple_knn", "simple_knn._C"):
    sys.modules[name] = types.ModuleType(name)
sys.modules["plyfile"].PlyData = sys.modules["plyfile"].PlyElement = object
sys.modules["simple_knn._C"].distCUDA2 = None


def _cpu(fn):
    def wrapped(*a, **k):
        if str(k.get("device", "")).startswith("cuda"):
            k["device"] = "cpu"
        return fn(*a, **k)
    return wrapped


torch.zeros, torch.ones, torch.empty = _cpu(torch.zeros), _cpu(torch.ones), _cpu(torch.empty)

_real_normal = torch.normal
_noise_log = []


def _normal(mean, std):
    state = torch.get_rng_state()
    real = _real_normal(mean=mean, std=std)
    torch.set_rng_state(state)
    z = torch.empty_like(std).normal_()
    assert torch.equal(real, z * std + mean), "torch.normal(mean, std) is no longer randn * std + mean"
    _noise_log.append(z.clone())
    return real


torch.normal = _normal

# scene/gaussian_model.py does not parse as shipped: :410-411 is a `def densify_and_prune(...)` header whose body is only
# a comment, directly followed by the real definition at :412. The generator drops those two lines IN MEMORY (checked
# verbatim first) and executes the rest of the file untouched; nothing of the reference is written anywhere.
_src = open(os.path.join(REF, "scene", "gaussian_model.py")).read().split("\n")
assert _src[409].strip() == "def densify_and_prune(self, max_grad, min_opacity, extent, max_screen_size):", _src[409]
assert _src[410].strip().startswith("#"), _src[410]
assert _src[411].strip().startswith("def densify_and_prune(self, max_grad, min_opacity, extent, max_screen_size, radii):")
_src[409] = _src[410] = ""
_mod = types.ModuleType("ref_gaussian_model")
exec(compile("\n".join(_src), os.path.join(REF, "scene", "gaussian_model.py"), "exec"), _mod.__dict__)
GaussianModel = _mod.GaussianModel

OPT = Namespace(percent_dense=0.01, position_lr_init=0.00016, position_lr_final=0.0000016, position_lr_delay_mult=0.01,
                position_lr_max_steps=30000, feature_lr=0.0025, opacity_lr=0.025, scaling_lr=0.005, rotation_lr=0.001)
NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


def make_model(P, seed, sh_degree, extent):
    g = torch.Generator().manual_seed(seed)
    gm = GaussianModel(sh_degree)
    M = (sh_degree + 1) ** 2
    gm._xyz = nn.Parameter((torch.rand(P, 3, generator=g) * 2 - 1) * extent)
    gm._features_dc = nn.Parameter(torch.randn(P, 1, 3, generator=g))
    gm._features_rest = nn.Parameter(torch.randn(P, M - 1, 3, generator=g) * 0.1)
    gm._opacity = nn.Parameter(torch.randn(P, 1, generator=g) * 3.0 - 1.0)             # sigmoid: many below 0.005
    big = (torch.rand(P, 1, generator=g) < 0.06).float() * 9.0 + 1.0                   # a few beyond 0.1 * extent
    gm._scaling = nn.Parameter(torch.log((torch.rand(P, 3, generator=g) * 0.013 * extent + 1e-3) * big))
    gm._rotation = nn.Parameter(torch.randn(P, 4, generator=g))
    gm._exposure = nn.Parameter(torch.eye(3, 4)[None])
    gm.spatial_lr_scale = extent
    gm.max_radii2D = torch.zeros(P)
    gm.training_setup(OPT)
    for _ in range(3):                                                                    # fill exp_avg / exp_avg_sq
        for grp in gm.optimizer.param_groups:
            p = grp["params"][0]
            p.grad = torch.randn(p.shape, generator=g) * 1e-3
        gm.optimizer.step()
    return gm, g


def snapshot(gm, tag, out):
    for grp in gm.optimizer.param_groups:
        p = grp["params"][0]
        st = gm.optimizer.state[p]
        out[f"{tag}_{grp['name']}"] = p.detach().numpy().copy()
        out[f"{tag}_{grp['name']}_m"] = st["exp_avg"].numpy().copy()
        out[f"{tag}_{grp['name']}_v"] = st["exp_avg_sq"].numpy().copy()
    out[f"{tag}_accum"], out[f"{tag}_denom"] = gm.xyz_gradient_accum.numpy().copy(), gm.denom.numpy().copy()
    out[f"{tag}_max_radii2D"] = gm.max_radii2D.numpy().copy()


out = {}
# cases: (tag, P, seed, sh_degree, max_grad, min_opacity, extent, max_screen_size)
CASES = (("a", 700, 0, 3, 0.0002, 0.005, 5.0, 20),      # everything happens: clone, split, opacity and world-size pruning
         ("b", 600, 1, 1, 0.0002, 0.005, 3.0, None),    # before opacity_reset_interval: size_threshold is None (train.py:163)
         ("c", 300, 2, 0, 1e9, 0.0, 5.0, None),         # nothing selected, nothing pruned; f_rest is [P, 0, 3]
         ("d", 257, 3, 1, 0.0001, 0.9, 4.0, 20))        # nearly everything pruned
for tag, P, seed, deg, max_grad, min_opacity, extent, mss in CASES:
    gm, g = make_model(P, seed, deg, extent)
    # three views' worth of add_densification_stats + max_radii2D (train.py:157-160)
    for v in range(3):
        vsp = nn.Parameter(torch.zeros(P, 3))
        vsp.grad = torch.randn(P, 3, generator=g) * 0.00018
        radii = (torch.rand(P, generator=g) * 60 - 20).clamp_min(0).to(torch.int32)
        vis = radii > 0
        out[f"{tag}_stats{v}_grad"], out[f"{tag}_stats{v}_radii"] = vsp.grad.numpy().copy(), radii.numpy().copy()
        gm.max_radii2D[vis] = torch.max(gm.max_radii2D[vis], radii[vis])
        gm.add_densification_stats(vsp, vis)
    snapshot(gm, f"{tag}_in", out)
    _noise_log.clear()
    torch.manual_seed(100 + seed)
    gm.densify_and_prune(max_grad, min_opacity, extent, mss, radii)
    assert len(_noise_log) == 1
    out[f"{tag}_noise"] = _noise_log[0].numpy().copy()
    snapshot(gm, f"{tag}_out", out)
    out[f"{tag}_args"] = np.array([max_grad, min_opacity, extent, -1.0 if mss is None else mss, OPT.percent_dense], dtype=np.float64)
    # prune_points with an arbitrary mask, then reset_opacity, on the densified model
    P2 = gm.get_xyz.shape[0]
    mask = torch.rand(P2, generator=g) < 0.3
    out[f"{tag}_prune_mask"] = mask.numpy().copy()
    gm.tmp_radii = torch.zeros(P2)
    gm.prune_points(mask)
    snapshot(gm, f"{tag}_pruned", out)
    gm.reset_opacity()
    snapshot(gm, f"{tag}_reset", out)
    print(tag, "P", P, "->", P2, "->", gm.get_xyz.shape[0], "noise rows", out[f"{tag}_noise"].shape[0])

np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_densify.npz"), **out)
print(sum(v.nbytes for v in out.values()) / 1e6, "MB raw")
