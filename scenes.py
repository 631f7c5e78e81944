"""Seeded synthetic scenes and cameras for tests and bench.py.

Pure CPU torch; nothing here touches the oracle or the CUDA library. Camera matrices follow the
reference's conventions (scene/cameras.py:69-72, utils/graphics_utils.py:38-71): viewmatrix and
projmatrix are the TRANSPOSED (row-vector) 4x4 matrices, projmatrix = view^T-form @ proj^T-form.
"""
import math
from typing import NamedTuple, Optional
import torch


class Camera(NamedTuple):
    W: int
    H: int
    tanfovx: float
    tanfovy: float
    viewmatrix: torch.Tensor   # [4,4] world_view_transform (transposed W2C)
    projmatrix: torch.Tensor   # [4,4] full_proj_transform
    campos: torch.Tensor       # [3]


class Scene(NamedTuple):
    means3D: torch.Tensor      # [P,3]
    scales: torch.Tensor       # [P,3]  (already activated: exp)
    rotations: torch.Tensor    # [P,4]  (already normalised; r,x,y,z)
    opacities: torch.Tensor    # [P,1]  (already sigmoid)
    shs: torch.Tensor          # [P,M,3]
    sh_degree: int


def projection_matrix(znear, zfar, fovx, fovy):
    """utils/graphics_utils.py:51-71 (getProjectionMatrix), untransposed."""
    ty, tx = math.tan(fovy / 2), math.tan(fovx / 2)
    top, right = ty * znear, tx * znear
    Pm = torch.zeros(4, 4)
    Pm[0, 0] = 2.0 * znear / (2 * right)
    Pm[1, 1] = 2.0 * znear / (2 * top)
    Pm[3, 2] = 1.0
    Pm[2, 2] = zfar / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    return Pm


def look_at_camera(W, H, fovx, eye, target=(0.0, 0.0, 0.0), up=(0.0, -1.0, 0.0), fovy=None,
                   znear=0.01, zfar=100.0) -> Camera:
    """Camera at `eye` looking at `target`; view space is x right, y down, z forward (COLMAP)."""
    eye = torch.tensor(eye, dtype=torch.float64)
    f = torch.tensor(target, dtype=torch.float64) - eye
    f = f / f.norm()
    r = torch.linalg.cross(f, torch.tensor(up, dtype=torch.float64))
    if r.norm() < 1e-6:
        r = torch.linalg.cross(f, torch.tensor((1.0, 0.0, 0.0), dtype=torch.float64))
    r = r / r.norm()
    d = torch.linalg.cross(f, r)
    Rw2c = torch.stack([r, d, f])                      # rows
    w2c = torch.eye(4, dtype=torch.float64)
    w2c[:3, :3] = Rw2c
    w2c[:3, 3] = -Rw2c @ eye
    if fovy is None:
        fovy = 2 * math.atan(math.tan(fovx / 2) * H / W)
    view = w2c.float().t().contiguous()
    proj = projection_matrix(znear, zfar, fovx, fovy).t()
    full = (view.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0).contiguous()
    campos = view.inverse()[3, :3].contiguous()
    return Camera(W, H, math.tan(fovx / 2), math.tan(fovy / 2), view, full, campos)


def pinhole_camera(W, H, fx, fy, eye, yaw=0.0, znear=0.01, zfar=100.0) -> Camera:
    """C5: 640x480 PINHOLE intrinsics of scene/dataset_readers.py:322-324; forward = +z rotated by yaw."""
    fovx, fovy = 2 * math.atan(W / (2 * fx)), 2 * math.atan(H / (2 * fy))
    tgt = (eye[0] + math.sin(yaw), eye[1], eye[2] + math.cos(yaw))
    return look_at_camera(W, H, fovx, eye, tgt, fovy=fovy, znear=znear, zfar=zfar)


def _quat(g, P):
    q = torch.randn(P, 4, generator=g)
    return q / q.norm(dim=1, keepdim=True)


def _shs(g, P, deg, dc_range=1.5, rest_std=0.1):
    M = (deg + 1) ** 2
    shs = torch.empty(P, M, 3)
    shs[:, 0] = (torch.rand(P, 3, generator=g) * 2 - 1) * dc_range
    if M > 1:
        shs[:, 1:] = torch.randn(P, M - 1, 3, generator=g) * rest_std
    return shs


def make_c1(P=10_000, seed=0) -> Scene:
    g = torch.Generator().manual_seed(seed)
    means = torch.rand(P, 3, generator=g) * 2 - 1
    scales = torch.exp(math.log(0.01) + torch.rand(P, 3, generator=g) * (math.log(0.08) - math.log(0.01)))
    opac = torch.sigmoid(torch.randn(P, 1, generator=g) * 1.5)
    return Scene(means, scales, _quat(g, P), opac, _shs(g, P, 0), 0)


def c1_camera(W=256, H=256) -> Camera:
    return look_at_camera(W, H, math.radians(60), (0.0, 0.0, -4.0), fovy=math.radians(60))


def make_c2(P=500_000, seed=1, deg=3) -> Scene:
    g = torch.Generator().manual_seed(seed)
    means = torch.randn(P, 3, generator=g) * 1.5
    scales = torch.exp(math.log(0.01) + torch.randn(P, 3, generator=g) * 0.6)
    opac = torch.sigmoid(torch.randn(P, 1, generator=g) * 1.5)
    return Scene(means, scales, _quat(g, P), opac, _shs(g, P, deg), deg)


def orbit_camera(W, H, fovx_deg, radius, azimuth_deg, height=0.5) -> Camera:
    a = math.radians(azimuth_deg)
    return look_at_camera(W, H, math.radians(fovx_deg), (radius * math.sin(a), -height, -radius * math.cos(a)))


def make_c3(P=3_000_000, seed=2, deg=3) -> Scene:
    """Mip-NeRF360-scale mixture: 70% foreground blob, 30% background shell."""
    g = torch.Generator().manual_seed(seed)
    nf = int(P * 0.7)
    nb = P - nf
    fg = torch.randn(nf, 3, generator=g) * 1.2
    fg_s = torch.exp(math.log(0.006) + torch.randn(nf, 3, generator=g) * 0.7)
    r = 5 + torch.rand(nb, 1, generator=g) * 25
    dirv = torch.randn(nb, 3, generator=g)
    dirv = dirv / dirv.norm(dim=1, keepdim=True)
    bgp = dirv * r
    bg_s = r * torch.exp(math.log(0.004) + torch.randn(nb, 3, generator=g) * 0.5)
    means = torch.cat([fg, bgp])
    scales = torch.cat([fg_s, bg_s])
    perm = torch.randperm(P, generator=g)
    means, scales = means[perm].contiguous(), scales[perm].contiguous()
    opac = torch.sigmoid(torch.randn(P, 1, generator=g) * 2.0)
    return Scene(means, scales, _quat(g, P), opac, _shs(g, P, deg), deg)


def c3_cameras(n=8, W=1920, H=1080, seed=3, jitter=0.0):
    g = torch.Generator().manual_seed(seed)
    cams = []
    for i in range(n):
        az = 360.0 * i / max(n, 1)
        h = 0.5
        if jitter > 0:
            az += float(torch.randn(1, generator=g)) * jitter * 10
            h += float(torch.randn(1, generator=g)) * jitter
        cams.append(orbit_camera(W, H, 60.0, 4.0, az, h))
    return cams


def make_c5(P=200_000, seed=4, deg=3, z0=0.0) -> Scene:
    """SLAM-style corridor slab ahead of the camera, depth U[0.5,10] m."""
    g = torch.Generator().manual_seed(seed)
    z = z0 + 0.5 + torch.rand(P, 1, generator=g) * 9.5
    xy = (torch.rand(P, 2, generator=g) * 2 - 1) * torch.tensor([3.0, 1.5])
    means = torch.cat([xy, z], dim=1)
    scales = torch.exp(math.log(0.01) + torch.randn(P, 3, generator=g) * 0.5)
    opac = torch.sigmoid(torch.randn(P, 1, generator=g) * 1.5)
    return Scene(means, scales, _quat(g, P), opac, _shs(g, P, deg), deg)


def c5_camera(k=0) -> Camera:
    return pinhole_camera(640, 480, 615.6607, 615.7676, (0.02 * math.sin(0.3 * k), 0.0, 0.05 * k),
                          yaw=0.05 * math.sin(0.7 * k))


def small_scene(P=300, seed=7, deg=3, spread=1.0, scale_lo=0.02, scale_hi=0.25) -> Scene:
    """Small, fat-Gaussian scene for exhaustive CPU tests (autograd / finite differences)."""
    g = torch.Generator().manual_seed(seed)
    means = (torch.rand(P, 3, generator=g) * 2 - 1) * spread
    scales = torch.exp(math.log(scale_lo) + torch.rand(P, 3, generator=g) * (math.log(scale_hi) - math.log(scale_lo)))
    opac = torch.sigmoid(torch.randn(P, 1, generator=g) * 1.5)
    return Scene(means, scales, _quat(g, P), opac, _shs(g, P, deg, rest_std=0.3), deg)
