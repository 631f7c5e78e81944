"""CPU tests (-m "not gpu"): pin the checker.

  1. the C restatement (oracle/gsr_oracle.c) against the golden vectors produced by RUNNING the reference's
     own Python restatements of rasterizer sub-steps (tests/golden/make_golden.py: utils/sh_utils.py:57-112,
     utils/general_utils.py:78-110, utils/graphics_utils.py:38-71);
  2. the C restatement against the independent torch-autograd derivation (oracle/torch_oracle.py): every
     forward intermediate and all 8 gradient tensors, with and without antialiasing / inverse depth;
  3. the autograd derivation against central finite differences in float64 (the only pin for the two
     switches the vendored reference lacks);
  4. integer stages (scan, keys, stable sort, ranges) against numpy on ragged / empty inputs;
  5. distCUDA2 semantics against a brute-force 3-NN.
Nothing here needs a GPU or /root/reference.
"""
import math
import os

import numpy as np
import pytest
import torch

import scenes
from oracle import c_oracle, torch_oracle
from util import oracle_scene, rel_l2

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_python_substeps.npz")


def _torch_run(sc, cam, bg, dtype=torch.float32, aa=False, dpix=None, dinv=None, colors=None, cov=None):
    cast = lambda t: t.to(dtype).clone().requires_grad_(True)
    leaves = dict(means3D=cast(sc.means3D), means2D=torch.zeros(sc.means3D.shape, dtype=dtype, requires_grad=True),
                  opacities=cast(sc.opacities))
    kw = {}
    if colors is None:
        leaves["shs"] = cast(sc.shs); kw["shs"] = leaves["shs"]
    else:
        leaves["colors_precomp"] = cast(colors); kw["colors_precomp"] = leaves["colors_precomp"]
    if cov is None:
        leaves["scales"], leaves["rotations"] = cast(sc.scales), cast(sc.rotations)
        kw.update(scales=leaves["scales"], rotations=leaves["rotations"])
    else:
        leaves["cov3D_precomp"] = cast(cov); kw["cov3D_precomp"] = leaves["cov3D_precomp"]
    color, radii, invd, aux = torch_oracle.rasterize(
        leaves["means3D"], leaves["means2D"], leaves["opacities"], W=cam.W, H=cam.H, tanfovx=cam.tanfovx,
        tanfovy=cam.tanfovy, bg=torch.as_tensor(bg, dtype=dtype), viewmatrix=cam.viewmatrix, projmatrix=cam.projmatrix,
        campos=cam.campos, sh_degree=sc.sh_degree, antialiasing=aa, return_aux=True, **kw)
    grads = None
    if dpix is not None:
        loss = (color * torch.as_tensor(dpix, dtype=dtype)).sum()
        if dinv is not None:
            loss = loss + (invd * torch.as_tensor(dinv, dtype=dtype)).sum()
        loss.backward()
        grads = {k: v.grad.numpy() for k, v in leaves.items() if v.grad is not None}
    return color.detach().numpy(), radii.numpy(), invd.detach().numpy(), aux, grads


# ------------------------------------------------------------------------------------ 1. golden vectors
def test_golden_sh_to_rgb_all_degrees():
    g = np.load(GOLD)
    dirs, shs = g["dirs"], g["shs"]
    n = dirs.shape[0]
    cam = scenes.look_at_camera(128, 128, math.radians(60), (0.0, 0.0, -4.0))
    for deg in range(4):
        S = c_oracle.Scene(W=cam.W, H=cam.H, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=np.zeros(3, np.float32),
                           means3D=dirs * 0.5, opacities=np.full(n, 0.5, np.float32), viewmatrix=cam.viewmatrix.numpy(),
                           projmatrix=cam.projmatrix.numpy(), campos=np.zeros(3, np.float32), sh_degree=deg, shs=shs,
                           scales=np.full((n, 3), 0.05, np.float32),
                           rotations=np.tile(np.array([1, 0, 0, 0], np.float32), (n, 1)))
        fw = c_oracle.forward(S)
        assert (fw["radii"] > 0).all()
        want = g[f"sh2rgb_deg{deg}"] + 0.5                      # forward.cu:62-70 adds 0.5 and clamps
        assert np.allclose(fw["rgb"], np.maximum(want, 0.0), rtol=1e-5, atol=2e-6), f"deg {deg}"
        assert np.array_equal(fw["clamped"].astype(bool), want < 0) or \
            np.abs(want[fw["clamped"].astype(bool) != (want < 0)]).max() < 1e-6


def test_golden_cov3d_from_scale_rotation():
    g = np.load(GOLD)
    n = g["scales"].shape[0]
    cam = scenes.look_at_camera(128, 128, math.radians(60), (0.0, 0.0, -4.0))
    gen = torch.Generator().manual_seed(5)
    means = ((torch.rand(n, 3, generator=gen) * 2 - 1) * 0.3).numpy()
    for mod in (1.0, 0.7):
        S = c_oracle.Scene(W=cam.W, H=cam.H, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=np.zeros(3, np.float32),
                           means3D=means, opacities=np.full(n, 0.5, np.float32), viewmatrix=cam.viewmatrix.numpy(),
                           projmatrix=cam.projmatrix.numpy(), campos=cam.campos.numpy(), sh_degree=0,
                           shs=np.zeros((n, 1, 3), np.float32), scales=g["scales"], rotations=g["rots"],
                           scale_modifier=mod)
        fw = c_oracle.forward(S)
        vis = fw["radii"] > 0
        assert vis.sum() > n // 2
        want = g[f"cov3D_mod{mod}"]
        assert np.allclose(fw["cov3D"][vis], want[vis], rtol=2e-5, atol=1e-7)
        # and the torch derivation
        tc = torch_oracle.cov3d(torch.from_numpy(g["scales"]), mod, torch.from_numpy(g["rots"])).numpy()
        assert np.allclose(tc, want, rtol=2e-5, atol=1e-7)


def test_golden_camera_matrices():
    g = np.load(GOLD)
    fovx, fovy = g["fov"]
    assert np.allclose(scenes.projection_matrix(0.01, 100.0, fovx, fovy).numpy(), g["proj"], rtol=1e-6, atol=1e-7)
    # getWorld2View2(R, t): Rt[:3,:3] = R^T, Rt[:3,3] = t  (graphics_utils.py:38-49)
    Rt = np.eye(4, dtype=np.float32)
    Rt[:3, :3] = g["w2v_R"].T
    Rt[:3, 3] = g["w2v_T"]
    assert np.allclose(Rt, g["w2v"], atol=1e-6)
    # our look_at cameras obey the same convention: view^T maps the eye to the origin, looks down +z
    cam = scenes.look_at_camera(64, 48, math.radians(50), (0.5, -0.3, -3.0))
    eye_h = np.array([0.5, -0.3, -3.0, 1.0], np.float32)
    assert np.allclose(eye_h @ cam.viewmatrix.numpy(), [0, 0, 0, 1], atol=1e-5)
    tgt = np.array([0, 0, 0, 1.0], np.float32) @ cam.viewmatrix.numpy()
    assert abs(tgt[0]) < 1e-5 and abs(tgt[1]) < 1e-5 and tgt[2] > 0
    assert np.allclose(cam.campos.numpy(), eye_h[:3], atol=1e-5)


# ------------------------------------------------------------------- 2. C restatement vs autograd derivation
CASES = [
    dict(P=250, deg=3, W=64, H=48, aa=False, depth=False, bg=(0.0, 0.0, 0.0)),
    dict(P=250, deg=3, W=72, H=40, aa=True, depth=True, bg=(0.3, 0.6, 0.1)),
    dict(P=200, deg=1, W=50, H=35, aa=False, depth=True, bg=(1.0, 1.0, 1.0)),
    dict(P=150, deg=0, W=33, H=33, aa=True, depth=False, bg=(0.5, 0.5, 0.5)),
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: f"P{c['P']}d{c['deg']}_{c['W']}x{c['H']}_aa{int(c['aa'])}dep{int(c['depth'])}")
def test_c_oracle_matches_autograd_oracle(case):
    sc = scenes.small_scene(P=case["P"], deg=case["deg"], seed=case["P"])
    cam = scenes.look_at_camera(case["W"], case["H"], math.radians(60), (0.1, -0.05, -3.0))
    rng = np.random.default_rng(1)
    dpix = rng.standard_normal((3, cam.H, cam.W)).astype(np.float32)
    dinv = rng.standard_normal((1, cam.H, cam.W)).astype(np.float32) if case["depth"] else None
    S = oracle_scene(sc, cam, case["bg"], antialiasing=case["aa"])
    fw = c_oracle.forward(S)
    g = c_oracle.backward(S, fw, dpix, dinv)
    color, radii, invd, aux, tg = _torch_run(sc, cam, case["bg"], aa=case["aa"], dpix=dpix, dinv=dinv)
    assert np.array_equal(fw["radii"], radii)
    assert np.array_equal(fw["tiles_touched"], aux["pre"]["tiles_touched"].numpy().astype(np.uint32))
    assert fw["num_rendered"] == aux["bins"]["num_rendered"]
    # view-space depth differs by an ulp between torch's matmul and the C expression order, so the depth half of
    # the key is compared to a few ulp; tile halves, ranges and the per-tile membership are exact
    kc, kt = fw["keys"], aux["bins"]["keys"]
    assert np.array_equal(kc >> np.uint64(32), kt >> np.uint64(32))
    dc, dtb = (kc & np.uint64(0xFFFFFFFF)).astype(np.int64), (kt & np.uint64(0xFFFFFFFF)).astype(np.int64)
    assert np.abs(dc - dtb).max() <= 8
    assert np.array_equal(fw["ranges"], aux["bins"]["ranges"])
    assert float((fw["point_list"] != aux["bins"]["point_list"]).mean()) <= 0.01
    for a, b in fw["ranges"]:
        assert np.array_equal(np.sort(fw["point_list"][a:b]), np.sort(aux["bins"]["point_list"][a:b]))
    assert np.abs(fw["color"] - color).max() <= 1e-5
    assert np.abs(fw["invdepth"] - invd).max() <= 1e-5
    assert np.abs(fw["final_T"].reshape(cam.H, cam.W) - aux["final_T"].numpy()).max() <= 1e-6
    assert np.array_equal(fw["n_contrib"].reshape(cam.H, cam.W), aux["n_contrib"].numpy().astype(np.uint32))
    for ours, theirs in (("dL_dmean3D", "means3D"), ("dL_dmean2D", "means2D"), ("dL_dopacity", "opacities"),
                         ("dL_dsh", "shs"), ("dL_dscale", "scales"), ("dL_drot", "rotations")):
        a, b = g[ours], tg[theirs]
        if ours == "dL_dmean2D":
            a, b = a[:, :2], b[:, :2]
        assert rel_l2(a.reshape(b.shape), b) <= 1e-4, ours


def test_c_oracle_precomputed_inputs_match_autograd():
    sc = scenes.small_scene(P=180, deg=0, seed=3)
    cam = scenes.look_at_camera(48, 48, math.radians(55), (0.0, 0.0, -3.2))
    gen = torch.Generator().manual_seed(3)
    colors = torch.rand(180, 3, generator=gen)
    cov = torch_oracle.cov3d(sc.scales, 1.0, sc.rotations)
    rng = np.random.default_rng(2)
    dpix = rng.standard_normal((3, cam.H, cam.W)).astype(np.float32)
    S = oracle_scene(sc, cam, (0.2, 0.2, 0.2), colors_precomp=colors, cov3D_precomp=cov)
    fw = c_oracle.forward(S)
    g = c_oracle.backward(S, fw, dpix, None)
    color, radii, _, _, tg = _torch_run(sc, cam, (0.2, 0.2, 0.2), dpix=dpix, colors=colors, cov=cov)
    assert np.array_equal(fw["radii"], radii)
    assert np.abs(fw["color"] - color).max() <= 1e-5
    assert rel_l2(g["dL_dcolor"], tg["colors_precomp"]) <= 1e-4
    assert rel_l2(g["dL_dcov3D"], tg["cov3D_precomp"]) <= 1e-4
    assert rel_l2(g["dL_dmean3D"], tg["means3D"]) <= 1e-4


# ------------------------------------------------------------------------- 3. autograd vs finite differences
def test_autograd_oracle_vs_finite_differences_float64():
    """float64 central differences through the complete pipeline, antialiasing + inverse depth on (the switches
    with no reference CUDA code). The blend's skip/stop decisions are piecewise constant, so a small step on a
    smooth scene measures the same branch."""
    sc = scenes.small_scene(P=24, deg=2, seed=11, scale_lo=0.15, scale_hi=0.4)
    cam = scenes.look_at_camera(24, 16, math.radians(60), (0.0, 0.0, -3.0))
    rng = np.random.default_rng(3)
    dpix = rng.standard_normal((3, cam.H, cam.W))
    dinv = rng.standard_normal((1, cam.H, cam.W))
    bg = (0.1, 0.2, 0.3)
    dt = torch.float64
    _, _, _, _, grads = _torch_run(sc, cam, bg, dtype=dt, aa=True, dpix=dpix, dinv=dinv)

    def loss_of(**over):
        vals = dict(means3D=sc.means3D.to(dt), opacities=sc.opacities.to(dt), shs=sc.shs.to(dt),
                    scales=sc.scales.to(dt), rotations=sc.rotations.to(dt))
        vals.update(over)
        with torch.no_grad():
            c, _, d = torch_oracle.rasterize(vals["means3D"], None, vals["opacities"], shs=vals["shs"],
                                             scales=vals["scales"], rotations=vals["rotations"], W=cam.W, H=cam.H,
                                             tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=torch.tensor(bg, dtype=dt),
                                             viewmatrix=cam.viewmatrix, projmatrix=cam.projmatrix, campos=cam.campos,
                                             sh_degree=sc.sh_degree, antialiasing=True)
        return float((c * torch.as_tensor(dpix)).sum() + (d * torch.as_tensor(dinv)).sum())

    eps = 1e-6
    probe = np.random.default_rng(4)
    for name in ("means3D", "opacities", "shs", "scales", "rotations"):
        base = getattr(sc, name).to(dt)
        flat_n = base.numel()
        worst = 0.0
        for j in probe.choice(flat_n, size=min(12, flat_n), replace=False):
            e = torch.zeros(flat_n, dtype=dt); e[j] = eps
            e = e.view_as(base)
            fd = (loss_of(**{name: base + e}) - loss_of(**{name: base - e})) / (2 * eps)
            an = float(grads[name].reshape(-1)[j])
            worst = max(worst, abs(fd - an) / max(1e-6, abs(fd), abs(an)))
        assert worst < 2e-4, f"{name}: finite-difference mismatch {worst:.2e}"


# ---------------------------------------------------------------------------------------- 4. integer stages
def test_integer_stages_scan_sort_ranges_ragged_and_empty():
    L = c_oracle.lib()
    import ctypes as C
    rng = np.random.default_rng(0)
    for P in (0, 1, 7, 1000):
        tiles = rng.integers(0, 9, size=P).astype(np.uint32)
        if P > 3:
            tiles[:3] = 0
        off = np.zeros(P, np.uint32)
        R = int(L.gso_scan(P, c_oracle._p(tiles), c_oracle._p(off)))
        assert R == int(tiles.sum()) and np.array_equal(off, np.cumsum(tiles, dtype=np.uint64).astype(np.uint32))
    assert [int(L.gso_higher_msb(C.c_uint32(t))) for t in (256, 1200, 8160, 1, 2)] == [9, 11, 13, 1, 2]
    for R, ntiles in ((0, 16), (1, 16), (5000, 300)):
        bit = int(L.gso_higher_msb(C.c_uint32(ntiles)))
        tile = rng.integers(0, ntiles, size=R).astype(np.uint64)
        depth = rng.random(R).astype(np.float32).view(np.uint32).astype(np.uint64)
        if R > 10:
            depth[:R // 2] = depth[0]                        # many ties: stability decides the order
        keys = (tile << np.uint64(32)) | depth
        vals = np.arange(R, dtype=np.uint32)
        ko, vo = np.zeros(R, np.uint64), np.zeros(R, np.uint32)
        L.gso_sort_pairs(C.c_size_t(R), c_oracle._p(keys), c_oracle._p(vals), c_oracle._p(ko), c_oracle._p(vo), 32 + bit)
        order = np.argsort(keys, kind="stable")
        assert np.array_equal(ko, keys[order]) and np.array_equal(vo, vals[order])
        ranges = np.zeros((ntiles, 2), np.uint32)
        L.gso_tile_ranges(C.c_size_t(R), c_oracle._p(ko), ntiles, c_oracle._p(ranges))
        t_sorted = (ko >> np.uint64(32)).astype(np.int64)
        for t in range(ntiles):
            idx = np.flatnonzero(t_sorted == t)
            want = (idx[0], idx[-1] + 1) if idx.size else (0, 0)
            assert tuple(ranges[t]) == want


def test_all_culled_and_empty_scene_render_background():
    cam = scenes.c1_camera(40, 24)
    sc = scenes.make_c1(P=50)
    sc = sc._replace(means3D=sc.means3D - torch.tensor([0.0, 0.0, 20.0]))
    fw = c_oracle.forward(oracle_scene(sc, cam, (0.25, 0.5, 0.75)))
    assert fw["num_rendered"] == 0 and (fw["radii"] == 0).all() and (fw["ranges"] == 0).all()
    assert np.allclose(fw["color"], np.array([0.25, 0.5, 0.75], np.float32).reshape(3, 1, 1))
    assert (fw["final_T"] == 1).all() and (fw["n_contrib"] == 0).all()
    vis = c_oracle.mark_visible(sc.means3D.numpy(), cam.viewmatrix.numpy())
    assert not vis.any()


# ---------------------------------------------------------------------------------------------- 5. distCUDA2
def test_dist2_oracle_is_exact_3nn_mean():
    gen = torch.Generator().manual_seed(9)
    for P in (4, 50, 3000):
        pts = torch.randn(P, 3, generator=gen) * 2 + torch.tensor([1.0, 2.0, -3.0])
        if P >= 50:
            pts[7] = pts[3]                                  # duplicates count with distance 0
        want = torch_oracle.dist2_bruteforce(pts).numpy()
        got = c_oracle.dist2(pts.numpy())
        assert np.allclose(got, want, rtol=1e-5, atol=1e-12)
