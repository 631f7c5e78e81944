"""Pure-torch, autograd-differentiable restatement of the rasterizer (CPU-capable, any float dtype).

TEST INFRASTRUCTURE ONLY (see gsr_oracle.c header). This is the INDEPENDENT derivation: the
forward follows the reference stage by stage, the backward comes from torch.autograd, not from a
transcription of backward.cu. It pins the hand-derived C restatement (gsr_oracle.c) and, for the
two switches the vendored reference lacks (antialiasing, inverse depth: SURVEY.md F1), it is the
only pin ("parity unpinned" by reference code).

Reference behaviours reproduced on purpose so that autograd equals backward.cu:
  * alpha = min(0.99, o*G) is differentiated as if unclamped (backward.cu:497-530 never masks
    the clamp)  -> straight-through estimator below.
  * grad of means2D is d/d(NDC-scaled pixel) (backward.cu:460-461,545-546) -> the dummy
    `means2D` input is added in NDC space before ndc2Pix.
  * the frustum clamp of t.x/t.z, t.y/t.z (forward.cu:84-88) is backpropagated as the reference does: a clamped
    coordinate is a constant (backward.cu:174-175,258-260 drop the d(lim*t.z)/dt.z term).
  * dL/dscale is taken w.r.t. (scale_modifier*scale) (backward.cu:278-341); compare gradients
    with scale_modifier == 1.
Citations are relative to submodules/diff-gaussian-rasterization/cuda_rasterizer/.
"""
import math
import numpy as np
import torch

C0 = 0.28209479177387814
C1 = 0.4886025119029199
C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
      -0.4570457994644658, 1.445305721320277, -0.5900435899266435]
TILE = 16


def sh_basis(deg, d):
    """forward.cu:31-60 ; d [P,3] unit directions -> [P,(deg+1)^2]."""
    x, y, z = d[:, 0], d[:, 1], d[:, 2]
    b = [torch.full_like(x, C0)]
    if deg > 0:
        b += [-C1 * y, C1 * z, -C1 * x]
    if deg > 1:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        b += [C2[0] * xy, C2[1] * yz, C2[2] * (2.0 * zz - xx - yy), C2[3] * xz, C2[4] * (xx - yy)]
    if deg > 2:
        b += [C3[0] * y * (3.0 * xx - yy), C3[1] * xy * z, C3[2] * y * (4.0 * zz - xx - yy),
              C3[3] * z * (2.0 * zz - 3.0 * xx - 3.0 * yy), C3[4] * x * (4.0 * zz - xx - yy),
              C3[5] * z * (xx - yy), C3[6] * x * (xx - 3.0 * yy)]
    return torch.stack(b, dim=1)


def sh_to_rgb(deg, shs, means, campos):
    """forward.cu:20-71 ; returns rgb [P,3] and the clamped mask."""
    d = means - campos
    d = d / d.norm(dim=1, keepdim=True)
    nb = (deg + 1) ** 2
    res = (sh_basis(deg, d).unsqueeze(2) * shs[:, :nb]).sum(1) + 0.5
    return torch.clamp_min(res, 0.0), res < 0


def quat_to_rot(q):
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    return torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=1).view(-1, 3, 3)


def cov3d(scales, mod, rots):
    """forward.cu:118-152 ; upper triangle [P,6]. Quaternion is NOT normalised (caller does)."""
    R = quat_to_rot(rots)
    A = R * (mod * scales).unsqueeze(1)
    S = A @ A.transpose(1, 2)
    return torch.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], dim=1)


def preprocess(means3D, means2D, opacities, shs, colors_precomp, scales, rotations, cov3D_precomp, *, W, H,
               tanfovx, tanfovy, viewmatrix, projmatrix, campos, sh_degree, scale_modifier=1.0,
               antialiasing=False):
    """forward.cu:155-256 (+ dr_aa antialiasing). Returns a dict of per-Gaussian quantities."""
    dt = means3D.dtype
    P = means3D.shape[0]
    view, proj = viewmatrix.to(dt), projmatrix.to(dt)
    fx, fy = W / (2.0 * tanfovx), H / (2.0 * tanfovy)
    ones = torch.ones(P, 1, dtype=dt)
    ph = torch.cat([means3D, ones], 1)
    pv = ph @ view[:, :3]
    phom = ph @ proj
    pw = 1.0 / (phom[:, 3] + 0.0000001)
    ndc = phom[:, :2] * pw.unsqueeze(1)
    if means2D is not None:
        ndc = ndc + means2D[:, :2]
    c3 = cov3D_precomp if cov3D_precomp is not None else cov3d(scales, scale_modifier, rotations)
    # forward.cu:74-113
    tz = pv[:, 2]
    tz_safe = torch.where(tz.abs() > 1e-12, tz, torch.full_like(tz, 1e-12))
    limx, limy = 1.3 * tanfovx, 1.3 * tanfovy
    # backward.cu:168-177,258-260: a clamped t.x / t.y is a CONSTANT for the reference's backward (x_grad_mul = 0
    # and no d(lim*t.z)/dt.z term); an unclamped one has gradient 1 w.r.t. view-space x / y only.
    rx, ry = pv[:, 0] / tz_safe, pv[:, 1] / tz_safe
    in_x = ((rx >= -limx) & (rx <= limx)).to(dt).detach()
    in_y = ((ry >= -limy) & (ry <= limy)).to(dt).detach()
    tx = (torch.clamp(rx, -limx, limx) * tz).detach() + (pv[:, 0] - pv[:, 0].detach()) * in_x
    ty = (torch.clamp(ry, -limy, limy) * tz).detach() + (pv[:, 1] - pv[:, 1].detach()) * in_y
    z0 = torch.zeros_like(tz)
    J = torch.stack([fx / tz_safe, z0, -(fx * tx) / (tz_safe * tz_safe),
                     z0, fy / tz_safe, -(fy * ty) / (tz_safe * tz_safe)], dim=1).view(P, 2, 3)
    Rv = view[:3, :3].t()
    T2 = J @ Rv
    V = torch.stack([c3[:, 0], c3[:, 1], c3[:, 2], c3[:, 1], c3[:, 3], c3[:, 4], c3[:, 2], c3[:, 4], c3[:, 5]],
                    dim=1).view(P, 3, 3)
    cov = T2 @ V @ T2.transpose(1, 2)
    a0, b0, c0 = cov[:, 0, 0], cov[:, 0, 1], cov[:, 1, 1]
    det_cov = a0 * c0 - b0 * b0
    a, c = a0 + 0.3, c0 + 0.3
    det = a * c - b0 * b0
    det_safe = torch.where(det != 0, det, torch.ones_like(det))
    hscale = torch.ones_like(det)
    if antialiasing:
        hscale = torch.sqrt(torch.clamp_min(det_cov / det_safe, 0.000025))
    conic = torch.stack([c / det_safe, -b0 / det_safe, a / det_safe], dim=1)
    with torch.no_grad():
        mid = 0.5 * (a + c)
        l1 = mid + torch.sqrt(torch.clamp_min(mid * mid - det, 0.1))
        l2 = mid - torch.sqrt(torch.clamp_min(mid * mid - det, 0.1))
        radius = torch.ceil(3.0 * torch.sqrt(torch.maximum(l1, l2)))
    # auxiliary.h:41-44 ndc2Pix in double
    pix = torch.stack([(((ndc[:, 0].double() + 1.0) * W - 1.0) * 0.5).to(dt),
                       (((ndc[:, 1].double() + 1.0) * H - 1.0) * 0.5).to(dt)], dim=1)
    gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    with torch.no_grad():
        def tdiv(v):  # C-style float -> int truncation of v / 16
            return torch.trunc(v / TILE).to(torch.int64)
        r = radius
        pxd, pyd = pix[:, 0].detach(), pix[:, 1].detach()
        mnx = tdiv(pxd - r).clamp(0, gx); mny = tdiv(pyd - r).clamp(0, gy)
        mxx = tdiv(pxd + r + (TILE - 1)).clamp(0, gx); mxy = tdiv(pyd + r + (TILE - 1)).clamp(0, gy)
        tiles = (mxy - mny) * (mxx - mnx)
        visible = (tz > 0.2) & (det != 0) & (tiles > 0) & torch.isfinite(radius)
        radii = torch.where(visible, radius, torch.zeros_like(radius)).to(torch.int32)
        tiles = torch.where(visible, tiles, torch.zeros_like(tiles))
    if colors_precomp is not None:
        rgb, clamped = colors_precomp, None
    else:
        rgb, clamped = sh_to_rgb(sh_degree, shs, means3D, campos.to(dt))
    return dict(depths=tz, means2D=pix, cov3D=c3, conic=conic, opacity=opacities.reshape(-1) * hscale,
                rgb=rgb, clamped=clamped, radii=radii, tiles_touched=tiles, visible=visible,
                rect=(mnx, mny, mxx, mxy), hscale=hscale)


def binning(pre, W, H):
    """rasterizer_impl.cu:70-138,277-317 : offsets, keys, stable sort, ranges (numpy ints)."""
    gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    tiles = pre["tiles_touched"].numpy().astype(np.int64)
    offsets = np.cumsum(tiles)
    R = int(offsets[-1]) if len(offsets) else 0
    mnx, mny, mxx, mxy = [t.numpy() for t in pre["rect"]]
    idx = np.repeat(np.arange(len(tiles)), tiles)
    start = np.repeat(offsets - tiles, tiles)
    local = np.arange(R) - start
    wdt = (mxx - mnx)[idx]
    ty = mny[idx] + local // np.maximum(wdt, 1)
    tx = mnx[idx] + local % np.maximum(wdt, 1)
    dbits = pre["depths"].detach().to(torch.float32).numpy().view(np.uint32).astype(np.uint64)
    keys = ((ty * gx + tx).astype(np.uint64) << np.uint64(32)) | dbits[idx]
    order = np.argsort(keys, kind="stable")
    skeys, plist = keys[order], idx[order].astype(np.uint32)
    ranges = np.zeros((gx * gy, 2), np.uint32)
    if R:
        t = (skeys >> np.uint64(32)).astype(np.int64)
        b = np.flatnonzero(np.diff(t)) + 1
        starts = np.concatenate([[0], b]); ends = np.concatenate([b, [R]])
        ranges[t[starts], 0] = starts; ranges[t[starts], 1] = ends
    return dict(point_offsets=offsets.astype(np.uint32), num_rendered=R, keys_unsorted=keys,
                vals_unsorted=idx.astype(np.uint32), keys=skeys, point_list=plist, ranges=ranges)


def blend(pre, bins, bg, W, H, want_invdepth=True):
    """forward.cu:261-374 (+ dr_aa inverse depth), vectorised per tile, autograd-differentiable."""
    dt = pre["means2D"].dtype
    gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    color = torch.zeros(3, H, W, dtype=dt)
    invd = torch.zeros(1, H, W, dtype=dt)
    final_T = torch.ones(H, W, dtype=dt)
    n_contrib = torch.zeros(H, W, dtype=torch.int64)
    xy, conic, opac, rgb, depth = pre["means2D"], pre["conic"], pre["opacity"], pre["rgb"], pre["depths"]
    plist = torch.from_numpy(bins["point_list"].astype(np.int64))
    bg = bg.to(dt)
    col_parts, inv_parts = [], []
    for t in range(gx * gy):
        r0, r1 = int(bins["ranges"][t, 0]), int(bins["ranges"][t, 1])
        tx, ty = t % gx, t // gx
        x0, y0, x1, y1 = tx * TILE, ty * TILE, min(tx * TILE + TILE, W), min(ty * TILE + TILE, H)
        ys, xs = torch.meshgrid(torch.arange(y0, y1), torch.arange(x0, x1), indexing="ij")
        npx = ys.numel()
        if r1 <= r0:
            col_parts.append((y0, y1, x0, x1, bg.view(3, 1, 1).expand(3, y1 - y0, x1 - x0)))
            continue
        g = plist[r0:r1]
        pxf, pyf = xs.reshape(-1, 1).to(dt), ys.reshape(-1, 1).to(dt)
        dx = xy[g, 0].unsqueeze(0) - pxf
        dy = xy[g, 1].unsqueeze(0) - pyf
        A, B, Cc = conic[g, 0], conic[g, 1], conic[g, 2]
        power = -0.5 * (A * dx * dx + Cc * dy * dy) - B * dx * dy
        raw = opac[g] * torch.exp(torch.clamp_max(power, 0.0))
        alpha = raw + (torch.clamp_max(raw, 0.99) - raw).detach()       # STE, see module docstring
        with torch.no_grad():
            valid = (power <= 0.0) & (alpha >= 1.0 / 255.0)
        a_eff = torch.where(valid, alpha, torch.zeros_like(alpha))
        one_m = 1.0 - a_eff
        T_incl = torch.cumprod(one_m, dim=1)
        T_excl = torch.cat([torch.ones(npx, 1, dtype=dt), T_incl[:, :-1]], dim=1)
        with torch.no_grad():
            stop = valid & (T_incl < 0.0001)
            stopped = torch.cumsum(stop.to(torch.int64), dim=1) > 0      # this and all later entries
            keep = valid & ~stopped
        w = torch.where(keep, a_eff * T_excl, torch.zeros_like(a_eff))
        # final T = product over kept entries only
        Tf = torch.where(keep, one_m, torch.ones_like(one_m)).prod(dim=1)
        c = w @ rgb[g] + Tf.unsqueeze(1) * bg.unsqueeze(0)
        col_parts.append((y0, y1, x0, x1, c.t().reshape(3, y1 - y0, x1 - x0)))
        if want_invdepth:
            inv_parts.append((y0, y1, x0, x1, (w @ (1.0 / depth[g])).reshape(1, y1 - y0, x1 - x0)))
        with torch.no_grad():
            L = keep.shape[1]
            pos = torch.arange(1, L + 1).unsqueeze(0) * keep.to(torch.int64)
            n_contrib[y0:y1, x0:x1] = pos.max(dim=1).values.reshape(y1 - y0, x1 - x0)
            final_T[y0:y1, x0:x1] = Tf.detach().reshape(y1 - y0, x1 - x0)
    # assemble without in-place writes on a leaf so autograd stays simple
    rows = []
    for tyi in range(gy):
        rows.append(torch.cat([p[4] for p in col_parts[tyi * gx:(tyi + 1) * gx]], dim=2))
    color = torch.cat(rows, dim=1)
    if want_invdepth:
        full = {(p[0], p[2]): p[4] for p in inv_parts}
        rows = []
        for tyi in range(gy):
            row = []
            for txi in range(gx):
                y0, x0 = tyi * TILE, txi * TILE
                y1, x1 = min(y0 + TILE, H), min(x0 + TILE, W)
                row.append(full.get((y0, x0), torch.zeros(1, y1 - y0, x1 - x0, dtype=dt)))
            rows.append(torch.cat(row, dim=2))
        invd = torch.cat(rows, dim=1)
    return color, invd, final_T, n_contrib


def rasterize(means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
              cov3D_precomp=None, *, W, H, tanfovx, tanfovy, bg, viewmatrix, projmatrix, campos, sh_degree=0,
              scale_modifier=1.0, antialiasing=False, return_aux=False):
    """Same call shape as GaussianRasterizer.forward (diff_gaussian_rasterization/__init__.py:187);
    returns (color[3,H,W], radii[P], invdepth[1,H,W]) and optionally every intermediate."""
    pre = preprocess(means3D, means2D, opacities, shs, colors_precomp, scales, rotations, cov3D_precomp, W=W, H=H,
                     tanfovx=tanfovx, tanfovy=tanfovy, viewmatrix=viewmatrix, projmatrix=projmatrix,
                     campos=campos, sh_degree=sh_degree, scale_modifier=scale_modifier,
                     antialiasing=antialiasing)
    bins = binning(pre, W, H)
    color, invd, final_T, n_contrib = blend(pre, bins, bg, W, H)
    if return_aux:
        return color, pre["radii"], invd, dict(pre=pre, bins=bins, final_T=final_T, n_contrib=n_contrib)
    return color, pre["radii"], invd


def dist2_bruteforce(points):
    """simple_knn.cu:132-184 result semantics: mean of 3 smallest squared distances to other points."""
    d = torch.cdist(points.double(), points.double()) ** 2
    d.fill_diagonal_(float("inf"))
    return d.topk(3, dim=1, largest=False).values.mean(dim=1).to(points.dtype)
