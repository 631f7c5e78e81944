"""GPU parity tests (-m gpu): the sm_100a kernels, called through the Python operator API -> ctypes ->
C ABI (include/gsr.h), against (1) the C restatement oracle/gsr_oracle.c and (2) the UNMODIFIED
reference CUDA build oracle/_ref/ref_dgr_C.so on the same seeded inputs.

Tolerances (BASELINE.json north_star / SURVEY.md 8c):
  bit-exact : radii, tiles_touched, sorted (tile|depth) keys, point_list, ranges, n_contrib
  max-abs   : colour, inverse depth, final_T  <= 1e-5
  rel-L2    : every gradient tensor           <= 1e-4
Against the reference CUDA build the forward is expected to be bit-identical; against the CPU oracle the
integer outputs may differ for the odd Gaussian sitting on a ceil()/tile boundary (the oracle does not
contract FMAs), so a tiny mismatch fraction is tolerated THERE and reported.
"""
import math

import numpy as np
import pytest
import torch

import scenes
from oracle import c_oracle, ref_cuda
from util import cuda_settings, oracle_scene, rel_l2, summarize_mismatch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
IMG_TOL = 1e-5
GRAD_TOL = 1e-4


def _run_ours(sc, cam, bg, antialiasing=False, colors_precomp=None, cov3D_precomp=None, dpix=None, dinv=None,
              want_state=True):
    from diff_gaussian_rasterization import GaussianRasterizer, _C
    dev = DEV
    st = cuda_settings(cam, bg, sc.sh_degree, dev, antialiasing=antialiasing)
    leaves = dict(means3D=sc.means3D.to(dev).requires_grad_(True),
                  means2D=torch.zeros_like(sc.means3D, device=dev).requires_grad_(True),
                  opacities=sc.opacities.to(dev).requires_grad_(True))
    kw = {}
    if colors_precomp is None:
        leaves["shs"] = sc.shs.to(dev).requires_grad_(True); kw["shs"] = leaves["shs"]
    else:
        leaves["colors_precomp"] = colors_precomp.to(dev).requires_grad_(True); kw["colors_precomp"] = leaves["colors_precomp"]
    if cov3D_precomp is None:
        leaves["scales"] = sc.scales.to(dev).requires_grad_(True)
        leaves["rotations"] = sc.rotations.to(dev).requires_grad_(True)
        kw.update(scales=leaves["scales"], rotations=leaves["rotations"])
    else:
        leaves["cov3D_precomp"] = cov3D_precomp.to(dev).requires_grad_(True); kw["cov3D_precomp"] = leaves["cov3D_precomp"]
    color, radii, invd = GaussianRasterizer(st)(means3D=leaves["means3D"], means2D=leaves["means2D"],
                                                opacities=leaves["opacities"], **kw)
    out = dict(color=color.detach().cpu().numpy(), radii=radii.cpu().numpy(), invdepth=invd.detach().cpu().numpy())
    if want_state:
        fn = color.grad_fn
        geom, binning, img = fn.saved_tensors[-3:]
        R = fn.num_rendered
        P = sc.means3D.shape[0]
        s = _C.decode_state(P, R, cam.W, cam.H, geom, binning, img)
        out["state"] = {k: v.cpu().numpy() for k, v in s.items()}
        out["keys"] = _C.sorted_keys(P, R, cam.W, cam.H, geom, binning, img).cpu().numpy().view(np.uint64)
        out["num_rendered"] = R
    if dpix is not None:
        loss = (color * torch.as_tensor(dpix, device=dev)).sum()
        if dinv is not None:
            loss = loss + (invd * torch.as_tensor(dinv, device=dev)).sum()
        loss.backward()
        out["grads"] = {k: v.grad.detach().cpu().numpy() for k, v in leaves.items() if v.grad is not None}
    torch.cuda.synchronize()
    return out


def _check_forward_vs_oracle(ours, fw, tag, int_slack=2e-4):
    st = ours["state"]
    P = len(fw["radii"])
    mism = float((ours["radii"] != fw["radii"]).mean())
    assert mism <= int_slack, summarize_mismatch(f"{tag} radii", ours["radii"], fw["radii"])
    exact = mism == 0.0 and np.array_equal(st["tiles_touched"].astype(np.uint32), fw["tiles_touched"])
    if exact:   # then the binning must be bit-identical too
        assert ours["num_rendered"] == fw["num_rendered"]
        ko, kf = ours["keys"], fw["keys"]
        assert np.array_equal(ko >> np.uint64(32), kf >> np.uint64(32)), f"{tag}: tile halves of the keys differ"
        if not np.array_equal(ko, kf):      # CPU depth can be an ulp off the CUDA value; order then may swap ties
            lo = np.uint64(0xFFFFFFFF)
            assert np.abs((ko & lo).astype(np.int64) - (kf & lo).astype(np.int64)).max() <= 4, \
                summarize_mismatch(f"{tag} keys", ko, kf)
            assert float((st["point_list"].astype(np.uint32) != fw["point_list"]).mean()) <= 0.01
        else:
            assert np.array_equal(st["point_list"].astype(np.uint32), fw["point_list"]), \
                summarize_mismatch(f"{tag} point_list", st["point_list"], fw["point_list"])
        assert np.array_equal(st["ranges"].astype(np.uint32), fw["ranges"]), \
            summarize_mismatch(f"{tag} ranges", st["ranges"], fw["ranges"])
        nc = float((st["n_contrib"].astype(np.uint32) != fw["n_contrib"]).mean())
        assert nc <= 1e-4, summarize_mismatch(f"{tag} n_contrib", st["n_contrib"], fw["n_contrib"])
    vis = fw["radii"] > 0
    assert np.allclose(st["means2D"][vis], fw["means2D"][vis], rtol=1e-5, atol=1e-3)
    assert np.allclose(st["conic_opacity"][vis], fw["conic_opacity"][vis], rtol=2e-4, atol=1e-6)
    d = np.abs(ours["color"] - fw["color"])
    frac_bad = float((d > IMG_TOL).mean())
    # a flipped alpha-threshold decision (CPU libm expf vs CUDA expf) moves a pixel by up to ~4e-3; the
    # reference-CUDA comparison below is the strict one
    assert frac_bad <= 2e-4 and d.max() < 2e-2, f"{tag} colour: max {d.max():.3e}, frac>{IMG_TOL} = {frac_bad:.2e}"
    return exact


def _check_grads(g_ours, g_ref, tag, names, flip_rows=0.0):
    """rel-L2 <= 1e-4 per gradient tensor. flip_rows > 0 (CPU-oracle comparisons only): the CPU's libm expf and CUDA's
    expf differ in the last bit, so an isolated (pixel, Gaussian) pair can fall on the other side of the alpha >= 1/255
    or T < 1e-4 test; one such flip moves ONE Gaussian's gradient by far more than 1e-4 of the tensor norm. Up to that
    fraction of Gaussians (rows) with the largest error is therefore set aside, like the pixel slack of the colour
    check. Against the reference CUDA build the forward is bit-identical, no rows are set aside."""
    for ours_name, ref_name in names:
        if ref_name not in g_ref or g_ref[ref_name] is None or ours_name not in g_ours:
            continue
        a, b = g_ours[ours_name], np.asarray(g_ref[ref_name])
        if np.linalg.norm(b) == 0:
            assert np.linalg.norm(a) == 0, f"{tag} {ours_name}: expected zero gradient"
            continue
        a = a.reshape(b.shape)
        if flip_rows > 0:
            P = b.shape[0]
            err = np.linalg.norm((a.astype(np.float64) - b).reshape(P, -1), axis=1)
            drop = max(1, int(flip_rows * P))
            keep = np.argsort(err)[:P - drop]
            a, b = a[keep], b[keep]
        e = rel_l2(a, b)
        assert e <= GRAD_TOL, f"{tag} grad {ours_name}: rel-L2 {e:.3e}"


ORACLE_GRADS = [("means3D", "dL_dmean3D"), ("means2D", "dL_dmean2D"), ("opacities", "dL_dopacity"), ("shs", "dL_dsh"),
                ("scales", "dL_dscale"), ("rotations", "dL_drot"), ("colors_precomp", "dL_dcolor"),
                ("cov3D_precomp", "dL_dcov3D")]
REF_GRADS = [("means3D", "dL_dmeans3D"), ("means2D", "dL_dmeans2D"), ("opacities", "dL_dopacity"), ("shs", "dL_dsh"),
             ("scales", "dL_dscales"), ("rotations", "dL_drotations"), ("colors_precomp", "dL_dcolors"),
             ("cov3D_precomp", "dL_dcov3D")]


def _ref_forward(sc, cam, bg, colors_precomp=None, cov3D_precomp=None):
    dev = DEV
    kw = dict(bg=torch.as_tensor(bg, dtype=torch.float32, device=dev), means3D=sc.means3D.to(dev),
              opacities=sc.opacities.to(dev), viewmatrix=cam.viewmatrix.to(dev), projmatrix=cam.projmatrix.to(dev),
              campos=cam.campos.to(dev), tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, H=cam.H, W=cam.W,
              sh_degree=sc.sh_degree)
    if colors_precomp is None:
        kw["shs"] = sc.shs.to(dev)
    else:
        kw["colors_precomp"] = colors_precomp.to(dev)
    if cov3D_precomp is None:
        kw.update(scales=sc.scales.to(dev), rotations=sc.rotations.to(dev))
    else:
        kw["cov3D_precomp"] = cov3D_precomp.to(dev)
    return ref_cuda.forward(**kw)


def _check_vs_reference_cuda(ours, sc, cam, bg, dpix, tag, **kw):
    """Strict comparison against the unmodified reference CUDA rasterizer (AA off, depth ignored)."""
    if not ref_cuda.available():
        pytest.skip("oracle/_ref/ref_dgr_C.so not built")
    P = sc.means3D.shape[0]
    fw = _ref_forward(sc, cam, bg, **kw)
    rd = {k: v.cpu().numpy() for k, v in ref_cuda.decode(fw, P, cam.W, cam.H).items()}
    st = ours["state"]
    assert np.array_equal(ours["radii"], fw["radii"].cpu().numpy()), \
        summarize_mismatch(f"{tag} radii vs ref", ours["radii"], fw["radii"].cpu().numpy())
    assert np.array_equal(st["tiles_touched"], rd["tiles_touched"]), \
        summarize_mismatch(f"{tag} tiles", st["tiles_touched"], rd["tiles_touched"])
    assert ours["num_rendered"] == fw["num_rendered"]
    if fw["num_rendered"] > 0:
        assert np.array_equal(ours["keys"].view(np.int64), rd["keys"]), \
            summarize_mismatch(f"{tag} keys vs ref", ours["keys"].view(np.int64), rd["keys"])
        assert np.array_equal(st["point_list"], rd["point_list"]), \
            summarize_mismatch(f"{tag} point_list vs ref", st["point_list"], rd["point_list"])
    assert np.array_equal(st["ranges"], rd["ranges"]), summarize_mismatch(f"{tag} ranges", st["ranges"], rd["ranges"])
    vis = ours["radii"] > 0
    assert np.array_equal(st["means2D"][vis], rd["means2D"][vis]), f"{tag}: means2D not bit-identical"
    assert np.array_equal(st["depths"][vis], rd["depths"][vis]), f"{tag}: depths not bit-identical"
    assert np.array_equal(st["conic_opacity"][vis], rd["conic_opacity"][vis]), f"{tag}: conic/opacity not bit-identical"
    assert np.array_equal(st["n_contrib"], rd["n_contrib"]), \
        summarize_mismatch(f"{tag} n_contrib vs ref", st["n_contrib"], rd["n_contrib"])
    assert np.abs(st["final_T"] - rd["final_T"]).max() <= 1e-6
    d = np.abs(ours["color"] - fw["color"].cpu().numpy())
    assert d.max() <= IMG_TOL, f"{tag} colour vs reference CUDA: max-abs {d.max():.3e}"
    if dpix is not None:
        g = ref_cuda.backward(fw, torch.as_tensor(dpix, device=DEV))
        g = {k: v.cpu().numpy() for k, v in g.items()}
        _check_grads(ours["grads"], g, tag + " vs ref", REF_GRADS)


def _f64_autograd_grads(sc, cam, bg, aa, dpix, dinv):
    """Gradients of the float64 torch-autograd derivation: the arbiter when two fp32 implementations disagree at
    the 1e-4 level (rotation / scale gradients are sums with heavy cancellation)."""
    from test_cpu_oracle import _torch_run
    *_, tg = _torch_run(sc, cam, bg, dtype=torch.float64, aa=aa, dpix=dpix, dinv=dinv)
    return {"dL_dmean3D": tg["means3D"], "dL_dopacity": tg["opacities"], "dL_dsh": tg["shs"], "dL_dscale": tg["scales"],
            "dL_drot": tg["rotations"]}


def _case(sc, cam, bg, seed=0, aa=False, depth_grad=False, vs_ref=True, f64_arbiter=False, **kw):
    rng = np.random.default_rng(seed)
    dpix = rng.standard_normal((3, cam.H, cam.W)).astype(np.float32)
    dinv = rng.standard_normal((1, cam.H, cam.W)).astype(np.float32) if depth_grad else None
    ours = _run_ours(sc, cam, bg, antialiasing=aa, dpix=dpix, dinv=dinv, **kw)
    S = oracle_scene(sc, cam, bg, antialiasing=aa, **kw)
    fw = c_oracle.forward(S)
    _check_forward_vs_oracle(ours, fw, "oracle")
    d = np.abs(ours["invdepth"] - fw["invdepth"])
    assert float((d > IMG_TOL).mean()) <= 2e-4, f"invdepth max {d.max():.3e}"
    g = c_oracle.backward(S, fw, dpix, dinv)
    if f64_arbiter:
        # truth for the switches without reference CUDA code = float64 autograd
        g64 = _f64_autograd_grads(sc, cam, bg, aa, dpix, dinv)
        _check_grads({k: g[k] for k in g64}, g64, "C oracle vs float64 autograd", [(k, k) for k in g64], flip_rows=0.002)
        _check_grads(ours["grads"], g64, "float64 autograd", [(o, r) for o, r in ORACLE_GRADS if r in g64], flip_rows=0.002)
    _check_grads(ours["grads"], g, "oracle", ORACLE_GRADS, flip_rows=0.002)
    if vs_ref and not aa and not depth_grad:
        _check_vs_reference_cuda(ours, sc, cam, bg, dpix, "C", **kw)
    return ours, fw


def test_c1_sh0_black_and_white_bg():
    sc, cam = scenes.make_c1(), scenes.c1_camera()
    _case(sc, cam, (0.0, 0.0, 0.0))
    _case(sc, cam, (1.0, 1.0, 1.0), seed=1)


def test_sh3_wide_image_ragged_edges():
    # 200x123 is not a multiple of 16 in either direction: partial tiles on both borders
    sc = scenes.small_scene(P=3000, deg=3, spread=1.5)
    cam = scenes.look_at_camera(200, 123, math.radians(70), (0.3, -0.2, -3.5))
    _case(sc, cam, (0.1, 0.4, 0.9))


def test_strip_wider_than_255_tiles_takes_the_unpacked_rectangle_path():
    """Binning packs a Gaussian's tile rectangle into 4 x 8 bits when the grid has at most 255 x 255 tiles; a 4128-pixel
    wide strip (258 tiles) must take the other path (rectangle recomputed from record + radius at emission) and give
    the same bit-identical binning against the reference CUDA build."""
    sc = scenes.small_scene(P=4000, deg=1, spread=2.5, scale_lo=0.02, scale_hi=0.4)
    cam = scenes.look_at_camera(4128, 48, math.radians(120), (0.1, 0.05, -2.0))
    rng = np.random.default_rng(11)
    dpix = rng.standard_normal((3, cam.H, cam.W)).astype(np.float32)
    ours = _run_ours(sc, cam, (0.2, 0.1, 0.0), dpix=dpix)
    assert ours["num_rendered"] > 1000
    assert int(ours["state"]["tiles_touched"].astype(np.int64).sum()) == ours["num_rendered"]
    _check_vs_reference_cuda(ours, sc, cam, (0.2, 0.1, 0.0), dpix, "strip")


@pytest.mark.parametrize("deg", [1, 2])
def test_partial_sh_degree(deg):
    sc = scenes.small_scene(P=1500, deg=3)
    sc = sc._replace(sh_degree=deg)           # M = 16 allocated, active degree lower (train.py:92-93 ramps it)
    cam = scenes.look_at_camera(128, 96, math.radians(60), (0.0, 0.0, -3.0))
    _case(sc, cam, (0.5, 0.5, 0.5))


def _aa_depth_case(sc, cam, bg, aa, seed):
    """Antialiasing / inverse depth have no reference CUDA code ('parity unpinned'). The pins are the C
    restatement, the float64 autograd derivation and finite differences (next test). NO Gaussian is set aside here:
    instead of dropping the rows with the largest error, the comparison is restricted to the pixels where the CUDA
    kernels and the CPU oracle took IDENTICAL decisions (same last contributor, colour / transmittance / inverse depth
    equal to 1e-5 -- CPU libm expf and CUDA expf differ in the last bit, so an isolated pair can fall on the other side
    of the alpha >= 1/255 or T < 1e-4 test). Those pixels must be >= 99.98 % of the image; dL/dpixel is zeroed on the
    others for BOTH sides and then every row of every gradient tensor is compared."""
    rng = np.random.default_rng(seed)
    dpix = rng.standard_normal((3, cam.H, cam.W)).astype(np.float32)
    dinv = rng.standard_normal((1, cam.H, cam.W)).astype(np.float32)
    ours = _run_ours(sc, cam, bg, antialiasing=aa)
    S = oracle_scene(sc, cam, bg, antialiasing=aa)
    fw = c_oracle.forward(S)
    st = ours["state"]
    same_radii = float((ours["radii"] == fw["radii"]).mean())
    assert same_radii >= 0.9998, f"radii equal on {same_radii:.6f} of the Gaussians"
    ok = (st["n_contrib"].astype(np.uint32) == fw["n_contrib"].reshape(-1)).reshape(cam.H, cam.W)
    ok &= np.abs(ours["color"] - fw["color"]).max(axis=0) <= IMG_TOL
    ok &= np.abs(ours["invdepth"][0] - fw["invdepth"].reshape(cam.H, cam.W)) <= IMG_TOL
    ok &= np.abs(st["final_T"].reshape(cam.H, cam.W) - fw["final_T"].reshape(cam.H, cam.W)) <= 1e-6
    assert float(ok.mean()) >= 0.9998, f"identical decisions on {float(ok.mean()):.6f} of the pixels"
    dpix = dpix * ok[None].astype(np.float32)
    dinv = dinv * ok[None].astype(np.float32)
    ours = _run_ours(sc, cam, bg, antialiasing=aa, dpix=dpix, dinv=dinv, want_state=False)
    g = c_oracle.backward(S, fw, dpix, dinv)
    g64 = _f64_autograd_grads(sc, cam, bg, aa, dpix, dinv)
    _check_grads({k: g[k] for k in g64}, g64, "C oracle vs float64 autograd", [(k, k) for k in g64])
    _check_grads(ours["grads"], g64, "float64 autograd", [(o, r) for o, r in ORACLE_GRADS if r in g64])
    _check_grads(ours["grads"], g, "oracle", ORACLE_GRADS)


def test_antialiasing_and_inverse_depth_vs_oracle():
    sc = scenes.small_scene(P=2500, deg=3, spread=1.2)
    cam = scenes.c5_camera(3)
    sc = sc._replace(means3D=sc.means3D + torch.tensor([0.0, 0.0, 3.0]))
    _aa_depth_case(sc, cam, (0.0, 0.0, 0.0), True, 0)
    _aa_depth_case(sc, cam, (0.2, 0.3, 0.4), False, 5)


def test_antialiasing_and_inverse_depth_finite_differences_on_the_gpu():
    """A pin that is not another implementation of the same formulas: fp32 central differences THROUGH THE CUDA FORWARD
    (antialiasing on, loss over colour and inverse depth, accumulated in float64) on 64 random parameters of a
    500-Gaussian scene against the CUDA backward. The loss is piecewise smooth (alpha >= 1/255, T < 1e-4, radius ceil),
    so a difference quotient that straddles a jump is off by far more than rounding: at least 56 of the 64 must agree to
    3 % (+ 2e-3 of the tensor's largest gradient), and the median relative error must be below 1 %."""
    from diff_gaussian_rasterization import GaussianRasterizer
    sc = scenes.small_scene(P=500, deg=2, spread=1.0, scale_lo=0.05, scale_hi=0.3)
    cam = scenes.look_at_camera(160, 112, math.radians(60), (0.1, -0.1, -3.0))
    st = cuda_settings(cam, (0.1, 0.2, 0.3), sc.sh_degree, DEV, antialiasing=True)
    rng = np.random.default_rng(12)
    # smooth, positive-ish weights keep the loss well conditioned
    yy, xx = np.meshgrid(np.linspace(0, 1, cam.H), np.linspace(0, 1, cam.W), indexing="ij")
    wc = torch.as_tensor(np.stack([1 + 0.5 * np.sin(3 * xx + k) * np.cos(2 * yy) for k in range(3)]).astype(np.float32), device=DEV)
    wd = torch.as_tensor((0.5 + 0.3 * np.cos(4 * xx * yy))[None].astype(np.float32), device=DEV)
    names = ("means3D", "opacities", "shs", "scales", "rotations")
    base = {n: getattr(sc, n).to(DEV).contiguous() for n in names}

    def loss_of(p, need_grad=False):
        q = {n: t.clone().requires_grad_(need_grad) for n, t in p.items()}
        color, radii, invd = GaussianRasterizer(st)(means3D=q["means3D"], means2D=torch.zeros_like(q["means3D"]),
                                                    opacities=q["opacities"], shs=q["shs"], scales=q["scales"],
                                                    rotations=q["rotations"])
        loss = (color.double() * wc.double()).sum() + (invd.double() * wd.double()).sum()
        if need_grad:
            loss.backward()
            return float(loss), {n: t.grad.detach().cpu().numpy() for n, t in q.items()}
        return float(loss)

    _, grads = loss_of(base, need_grad=True)
    radii = GaussianRasterizer(st)(means3D=base["means3D"], means2D=torch.zeros_like(base["means3D"]), opacities=base["opacities"],
                                   shs=base["shs"], scales=base["scales"], rotations=base["rotations"])[1].cpu().numpy()
    vis = np.flatnonzero(radii > 0)
    assert vis.size > 200
    rel_err, good, total = [], 0, 0
    steps = {"means3D": 2e-3, "opacities": 2e-3, "shs": 5e-3, "scales": 1e-3, "rotations": 2e-3}
    for n in names:
        flat_n = base[n].reshape(base[n].shape[0], -1).shape[1]
        gmax = float(np.abs(grads[n]).max())
        for _ in range(13 if n != "opacities" else 12):
            i, j = int(rng.choice(vis)), int(rng.integers(flat_n))
            h = steps[n]
            lp, lm = {}, {}
            for sgn, store in ((+1, lp), (-1, lm)):
                p = {k: v for k, v in base.items()}
                t = base[n].clone()
                t.view(t.shape[0], -1)[i, j] += sgn * h
                p[n] = t
                store["v"] = loss_of(p)
            fd = (lp["v"] - lm["v"]) / (2 * h)
            an = float(grads[n].reshape(grads[n].shape[0], -1)[i, j])
            err = abs(fd - an)
            total += 1
            if err <= 3e-2 * abs(an) + 2e-3 * gmax:
                good += 1
            rel_err.append(err / max(abs(an), 2e-3 * gmax))
    assert total == 64
    assert good >= 56, f"finite differences agree for {good} of {total} parameters; relative errors {sorted(rel_err)[-10:]}"
    assert float(np.median(rel_err)) < 1e-2, f"median relative error {np.median(rel_err):.3e}"


def test_precomputed_colors_and_cov3d():
    sc = scenes.small_scene(P=2000, deg=0)
    cam = scenes.look_at_camera(160, 160, math.radians(55), (0.0, 0.0, -3.2))
    g = torch.Generator().manual_seed(3)
    colors = torch.rand(2000, 3, generator=g)
    S = oracle_scene(sc, cam, (0, 0, 0))
    cov = torch.from_numpy(c_oracle.forward(S)["cov3D"].copy())
    vis_fix = cov.abs().sum(1) == 0           # culled Gaussians never got a cov3D: give them any valid one
    cov[vis_fix] = torch.tensor([1e-4, 0, 0, 1e-4, 0, 1e-4])
    _case(sc, cam, (0.3, 0.3, 0.3), colors_precomp=colors, cov3D_precomp=cov)


def test_huge_gaussians_cover_every_tile_and_behind_camera():
    sc = scenes.small_scene(P=600, deg=1, spread=2.5, scale_lo=0.3, scale_hi=3.0)
    cam = scenes.look_at_camera(256, 144, math.radians(80), (0.0, 0.0, -1.0))   # many points behind / near plane
    ours, fw = _case(sc, cam, (0.0, 0.0, 0.0))
    assert (fw["radii"] == 0).any() and fw["tiles_touched"].max() >= 16 * 9 // 2


def test_needles_and_faint_splats_exercise_the_ellipse_cull():
    # The blend kernels drop a splat for a whole 8x8 sub-tile when the alpha >= 1/255 ellipse misses it (exact test after
    # the bounding box). Needle-shaped, randomly rotated splats make box and ellipse differ the most; opacities around
    # 1/255 put the ellipse's size at its limit. Everything must still match the reference CUDA build and the oracle.
    g = torch.Generator().manual_seed(21)
    P = 4000
    sc = scenes.small_scene(P=P, deg=2, spread=1.4, seed=21)
    long_axis = torch.randint(0, 3, (P,), generator=g)
    scales = torch.exp(math.log(0.002) + torch.rand(P, 3, generator=g) * (math.log(0.008) - math.log(0.002)))
    scales[torch.arange(P), long_axis] = 0.15 + 0.5 * torch.rand(P, generator=g)
    opac = torch.sigmoid(torch.randn(P, 1, generator=g) * 3.0 - 3.0)
    opac[:200] = (1.0 / 255.0) * (0.9 + 0.3 * torch.rand(200, 1, generator=g))          # straddling the alpha threshold
    sc = scenes.Scene(sc.means3D, scales, sc.rotations, opac, sc.shs, sc.sh_degree)
    cam = scenes.look_at_camera(320, 200, math.radians(65), (0.2, 0.1, -3.2))
    # Checker 1: the unmodified reference CUDA build -- state bit-identical (n_contrib included, i.e. no contributing
    # splat was dropped anywhere), colour <= 1e-5.
    # Checker 2 (gradients): needles make the gradient sums ill-conditioned -- two fp32 implementations that add the same
    # terms in a different order differ by 1e-3 .. 1e-2 on means3D / scales / rotations here, and so do two RUNS of either
    # one (float atomics; tools/needle_probe.py prints the spread: the reference CUDA build lands between 0.9e-3 and
    # 2.9e-3 from the float64 means3D gradient from run to run, ours between 0.9e-3 and 2.8e-3). Comparing two noisy fp32
    # results with each other says little, so the arbiter is the float64 autograd derivation, and both sides are SAMPLED:
    # the median of our distances over 5 runs must not exceed 1.5 x the reference CUDA build's largest over 5 runs, or
    # 3 x its median (floor 1e-4 = the north_star bar). The second term is there because the noise is bimodal and
    # sticky within a process: profiles/r02_needle_probe_ctypes.txt shows a process in which the reference drew 1.1e-3
    # and 1.3e-3 on means3D while ours drew 2.4e-3 .. 2.8e-3 three times in a row (and 1.0e-3 .. 1.7e-3 with the other
    # backward variant), profiles/r02_needle_probe_torch_extension.txt one in which the reference drew 2.9e-3 and ours
    # 0.9e-3. A wrong term shows up at 1e-2 .. 1e-1 on this scene; the well-conditioned scenes hold 1e-4 strictly.
    # Checker 3: the same kernels with the cull switched off (every splat whose box touches the sub-tile is blended,
    # alpha = 0 where it does not contribute) -- forward bit-identical, directly accumulated gradients <= 1e-6.
    from diff_gaussian_rasterization import _C
    for bg, seed in (((0.0, 0.0, 0.0), 5), ((0.3, 0.6, 0.1), 6)):
        dpix = np.random.default_rng(seed).standard_normal((3, cam.H, cam.W)).astype(np.float32)
        ours = _run_ours(sc, cam, bg, dpix=dpix)
        assert ours["state"]["n_contrib"].max() > 40 and (ours["radii"] > 0).sum() > 3000
        _check_vs_reference_cuda(ours, sc, cam, bg, None, "needles")          # forward state + colour, strict
        if ref_cuda.available():
            fw = _ref_forward(sc, cam, bg)
            g64 = _f64_autograd_grads(sc, cam, bg, False, dpix, None)
            runs_ref = [{k: v.cpu().numpy() for k, v in ref_cuda.backward(fw, torch.as_tensor(dpix, device=DEV)).items()}
                        for _ in range(5)]
            runs_ours = [ours["grads"]] + [_run_ours(sc, cam, bg, dpix=dpix)["grads"] for _ in range(4)]
            pairs = (("means3D", "dL_dmeans3D", "dL_dmean3D"), ("opacities", "dL_dopacity", "dL_dopacity"),
                     ("shs", "dL_dsh", "dL_dsh"), ("scales", "dL_dscales", "dL_dscale"), ("rotations", "dL_drotations", "dL_drot"))
            for ours_name, ref_name, name64 in pairs:
                want = np.asarray(g64[name64])
                e_ours = float(np.median([rel_l2(g[ours_name].reshape(want.shape), want) for g in runs_ours]))
                d_ref = [rel_l2(g[ref_name].reshape(want.shape), want) for g in runs_ref]
                e_ref = max(d_ref)
                assert e_ours <= max(1.5 * e_ref, 3.0 * float(np.median(d_ref)), GRAD_TOL), \
                    f"needles {ours_name}: ours {e_ours:.3e} from the float64 gradients, reference CUDA build {e_ref:.3e}"
        _C.set_exact_cull(False)
        try:
            plain = _run_ours(sc, cam, bg, dpix=dpix)
        finally:
            _C.set_exact_cull(True)
        assert np.array_equal(ours["color"], plain["color"]) and np.array_equal(ours["invdepth"], plain["invdepth"])
        assert np.array_equal(ours["state"]["n_contrib"], plain["state"]["n_contrib"])
        assert np.array_equal(ours["state"]["final_T"], plain["state"]["final_T"])
        # the gradients the blend accumulates directly agree to the noise of its float atomics; those derived from
        # dL/dconic through the needle covariances are covered by checker 2
        for k, g in ours["grads"].items():
            if g is not None and np.linalg.norm(plain["grads"][k]) > 0 and k not in ("means3D", "scales", "rotations", "cov3D_precomp"):
                assert rel_l2(g, plain["grads"][k]) <= 1e-6, f"cull on/off: grad {k} rel-L2 {rel_l2(g, plain['grads'][k]):.3e}"


def test_huge_screen_space_needles_keep_every_contributing_pixel():
    """the exact sub-tile cull evaluates the ellipse in fp32; for splats hundreds to thousands of pixels long,
    A x^2 reaches 1e5 .. 1e6 and the rounding of the two candidate values exceeds a fixed margin. The margin now scales
    with the magnitude of the terms. 1080p, screen-space sigmas up to ~2500 px, rotated needles: n_contrib and colour
    must stay bit-identical to the reference CUDA build (a dropped fringe pixel shows in n_contrib / colour)."""
    g = torch.Generator().manual_seed(33)
    P = 1500
    sc = scenes.small_scene(P=P, deg=1, spread=1.2, seed=33)
    long_axis = torch.randint(0, 3, (P,), generator=g)
    scales = torch.exp(math.log(0.001) + torch.rand(P, 3, generator=g) * (math.log(0.004) - math.log(0.001)))
    scales[torch.arange(P), long_axis] = 0.3 + 3.0 * torch.rand(P, generator=g)            # ~150 .. 2500 px at this camera
    opac = torch.sigmoid(torch.randn(P, 1, generator=g) * 2.0 - 1.0)
    sc = scenes.Scene(sc.means3D, scales, sc.rotations, opac, sc.shs, sc.sh_degree)
    cam = scenes.look_at_camera(1920, 1080, math.radians(60), (0.1, 0.05, -2.2))
    from diff_gaussian_rasterization import _C
    for bg in ((0.0, 0.0, 0.0), (0.5, 0.2, 0.7)):
        ours = _run_ours(sc, cam, bg)
        assert ours["radii"].max() > 1500, "scene does not reach the large-splat regime"
        _check_vs_reference_cuda(ours, sc, cam, bg, None, "huge needles")
        _C.set_exact_cull(False)
        try:
            plain = _run_ours(sc, cam, bg)
        finally:
            _C.set_exact_cull(True)
        assert np.array_equal(ours["state"]["n_contrib"], plain["state"]["n_contrib"])
        assert np.array_equal(ours["color"], plain["color"])


def test_empty_and_all_culled():
    from diff_gaussian_rasterization import GaussianRasterizer
    cam = scenes.c1_camera(64, 48)
    st = cuda_settings(cam, (0.25, 0.5, 0.75), 0, DEV)
    z = torch.zeros(0, 3, device=DEV)
    color, radii, invd = GaussianRasterizer(st)(means3D=z, means2D=z, opacities=torch.zeros(0, 1, device=DEV),
                                                shs=torch.zeros(0, 1, 3, device=DEV), scales=z,
                                                rotations=torch.zeros(0, 4, device=DEV))
    assert color.shape == (3, 48, 64) and float(color.abs().max()) == 0.0 and radii.numel() == 0   # rasterize_points.cu:81
    sc = scenes.make_c1(P=500)
    sc = sc._replace(means3D=sc.means3D - torch.tensor([0.0, 0.0, 20.0]))      # everything behind the camera
    ours = _run_ours(sc, cam, (0.25, 0.5, 0.75), dpix=np.ones((3, 48, 64), np.float32))
    assert ours["num_rendered"] == 0 and (ours["radii"] == 0).all()
    assert np.allclose(ours["color"], np.array([0.25, 0.5, 0.75], np.float32).reshape(3, 1, 1))
    assert all(np.abs(v).max() == 0 for v in ours["grads"].values())


def test_packed_exp_is_bit_identical_to_expf_for_every_float():
    """The blend kernels' two-value exp (FFMA2 restatement of CUDA expf) vs expf over all 2^32 bit patterns."""
    from diff_gaussian_rasterization import _C
    assert _C.exp_mismatches(DEV) == 0


def test_mark_visible_and_validation_errors():
    from diff_gaussian_rasterization import GaussianRasterizer
    sc, cam = scenes.make_c1(P=4000), scenes.c1_camera()
    sc = sc._replace(means3D=sc.means3D * 4.0)
    r = GaussianRasterizer(cuda_settings(cam, (0, 0, 0), 0, DEV))
    vis = r.markVisible(sc.means3D.to(DEV)).cpu().numpy()
    assert np.array_equal(vis, c_oracle.mark_visible(sc.means3D.numpy(), cam.viewmatrix.numpy()))
    if ref_cuda.available():
        rv = ref_cuda.mark_visible(sc.means3D.to(DEV), cam.viewmatrix.to(DEV), cam.projmatrix.to(DEV))
        assert np.array_equal(vis, rv.cpu().numpy())
    m = sc.means3D.to(DEV)
    with pytest.raises(Exception, match="excatly one of either SHs"):
        r(means3D=m, means2D=m, opacities=sc.opacities.to(DEV), scales=sc.scales.to(DEV), rotations=sc.rotations.to(DEV))
    with pytest.raises(Exception, match="scale/rotation pair"):
        r(means3D=m, means2D=m, opacities=sc.opacities.to(DEV), shs=sc.shs.to(DEV), scales=sc.scales.to(DEV))


def test_dist2_vs_oracle_and_reference():
    from simple_knn._C import distCUDA2
    g = torch.Generator().manual_seed(11)
    for P, spread in ((1, 1.0), (3, 1.0), (5, 1.0), (1000, 1.0), (20000, 3.0)):
        pts = torch.randn(P, 3, generator=g) * spread + torch.tensor([2.0, -1.0, 0.5])
        if P >= 1000:
            pts[17] = pts[5]                                   # duplicate point: distance 0 counts
        ours = distCUDA2(pts.to(DEV)).cpu().numpy()
        if P >= 4:
            want = c_oracle.dist2(pts.numpy())
            assert np.allclose(ours, want, rtol=2e-6, atol=1e-12), f"P={P}: max rel {np.abs(ours / want - 1).max()}"
        if ref_cuda.available():
            ref = ref_cuda.dist2(pts.to(DEV)).cpu().numpy()
            assert np.allclose(ours, ref, rtol=2e-6, atol=1e-12, equal_nan=True), f"P={P} vs reference"


@pytest.mark.parametrize("P", [1_000_000, 3_000_000])
def test_dist2_at_full_size_vs_reference_cuda(P):
    """create_from_pcd runs distCUDA2 over the whole cloud (scene/gaussian_model.py:140): parity at 1M and 3M points
    against the unmodified simple-knn build (exact 3-NN on both sides => only the rounding of the mean differs)."""
    if not ref_cuda.available():
        pytest.skip("oracle/_ref/ref_knn_C.so not built")
    from simple_knn._C import distCUDA2
    pts = scenes.make_c3(P=P).means3D.to(DEV)
    ours = distCUDA2(pts).cpu().numpy()
    ref = ref_cuda.dist2(pts).cpu().numpy()
    assert ours.shape == ref.shape and np.isfinite(ours).all()
    assert np.allclose(ours, ref, rtol=2e-6, atol=1e-12), f"P={P}: max rel {np.abs(ours / np.maximum(ref, 1e-30) - 1).max():.3e}"


def test_dist2_duplicate_heavy_and_collinear_clouds():
    """Edge cases of the box-pruned search (simple_knn.cu:132-184): every point 4 times (all three neighbours at distance
    0), a cloud on one line (Morton codes vary along one curve only, boxes degenerate), a planar cloud, and a cloud with
    one far outlier (the bounding box that normalises the Morton codes is dominated by it)."""
    from simple_knn._C import distCUDA2
    g = torch.Generator().manual_seed(5)
    base = torch.randn(5000, 3, generator=g)
    t = torch.rand(20000, 1, generator=g)
    clouds = {
        "duplicates x4": base.repeat(4, 1)[torch.randperm(20000, generator=g)],
        "collinear": torch.tensor([[1.0, -2.0, 0.5]]) + t * torch.tensor([[3.0, 1.0, -2.0]]),
        "planar": torch.cat([torch.randn(20000, 2, generator=g), torch.full((20000, 1), 0.25)], dim=1),
        "outlier": torch.cat([torch.randn(19999, 3, generator=g) * 0.01, torch.tensor([[1000.0, -500.0, 250.0]])]),
    }
    for name, pts in clouds.items():
        ours = distCUDA2(pts.to(DEV)).cpu().numpy()
        want = c_oracle.dist2(pts.numpy())
        assert np.allclose(ours, want, rtol=2e-6, atol=1e-12), f"{name}: max abs {np.abs(ours - want).max():.3e}"
        if name == "duplicates x4":
            assert float(np.abs(ours).max()) == 0.0
        if ref_cuda.available():
            ref = ref_cuda.dist2(pts.to(DEV)).cpu().numpy()
            assert np.allclose(ours, ref, rtol=2e-6, atol=1e-12, equal_nan=True), f"{name} vs reference"


def test_c2_scale_vs_reference_cuda_and_properties():
    """500k Gaussians, SH3, 1080p (BASELINE config 2): strict parity with the reference CUDA build plus the
    size-independent invariants (sorted keys, ranges partition the list, sum(tiles) == R, determinism)."""
    sc = scenes.make_c2()
    cam = scenes.orbit_camera(1920, 1080, 70.0, 4.0, 35.0)
    rng = np.random.default_rng(7)
    dpix = rng.standard_normal((3, cam.H, cam.W)).astype(np.float32)
    ours = _run_ours(sc, cam, (0.0, 0.0, 0.0), dpix=dpix)
    st, keys, R = ours["state"], ours["keys"], ours["num_rendered"]
    assert int(st["tiles_touched"].astype(np.int64).sum()) == R
    assert np.all(keys[1:] >= keys[:-1]), "keys not sorted"
    rg = st["ranges"].astype(np.int64)
    nz = rg[:, 1] > rg[:, 0]
    assert int((rg[nz, 1] - rg[nz, 0]).sum()) == R
    tiles_of_keys = (keys >> np.uint64(32)).astype(np.int64)
    first = rg[nz, 0]
    assert np.array_equal(tiles_of_keys[first], np.flatnonzero(nz))
    again = _run_ours(sc, cam, (0.0, 0.0, 0.0), want_state=False)
    assert np.array_equal(again["color"], ours["color"]), "forward is not deterministic"
    _check_vs_reference_cuda(ours, sc, cam, (0.0, 0.0, 0.0), dpix, "C2")


def test_grad_sinks_accumulate_two_views_equals_sum_of_separate_backwards():
    """Superset API (multi-view batches): in-kernel accumulation into caller-owned sinks == autograd's own sum."""
    from diff_gaussian_rasterization import GaussianRasterizer
    sc = scenes.small_scene(P=3000, deg=3, spread=1.6)
    cams = [scenes.look_at_camera(160, 96, math.radians(65), e) for e in ((0.3, -0.2, -3.5), (-2.5, 0.1, -2.0))]
    rng = np.random.default_rng(3)
    dpix = torch.as_tensor(rng.standard_normal((3, 96, 160)).astype(np.float32), device=DEV)
    names = ("means3D", "opacities", "shs", "scales", "rotations")

    def leaves():
        return {n: getattr(sc, n).to(DEV).clone().requires_grad_(True) for n in names}

    def run(p, sinks):
        m2 = []
        for cam in cams:
            st = cuda_settings(cam, (0.1, 0.2, 0.3), 3, DEV)
            s2 = torch.zeros_like(p["means3D"], requires_grad=True)
            color, radii, _ = GaussianRasterizer(st)(means3D=p["means3D"], means2D=s2, opacities=p["opacities"],
                                                     shs=p["shs"], scales=p["scales"], rotations=p["rotations"],
                                                     grad_sinks=sinks)
            torch.autograd.backward([color], [dpix])
            m2.append(s2.grad.clone())
        return m2
    pa = leaves()
    m2a = run(pa, None)
    pb = leaves()
    sinks = {n: torch.zeros_like(t) for n, t in pb.items()}
    m2b = run(pb, sinks)
    assert all(t.grad is None for t in pb.values())
    for n in names:
        assert rel_l2(sinks[n].cpu().numpy(), pa[n].grad.cpu().numpy()) <= 1e-5, n
    for a, b in zip(m2a, m2b):
        assert rel_l2(b.cpu().numpy(), a.cpu().numpy()) <= 1e-5          # means2D stays per view
    with pytest.raises(RuntimeError, match="all-or-nothing"):
        run(leaves(), {"means3D": sinks["means3D"]})
