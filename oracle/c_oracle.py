"""ctypes front-end for oracle/libgsr_oracle.so (the C restatement in gsr_oracle.c).

TEST INFRASTRUCTURE ONLY -- see the header of gsr_oracle.c. Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs may import this module.

All arrays are numpy, fp32 / int32 / uint32 / uint64, C-contiguous. The function names follow the
reference stages: preprocess (A4/A5), scan (A6), duplicate_with_keys (A7),
sort (A8), tile_ranges (A9), render (A10), render_backward (A11), preprocess_backward (A12/A13).
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libgsr_oracle.so")


def build(force=False):
    src = os.path.join(_HERE, "gsr_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(
            ["gcc", "-O2", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off", "-o", _SO, src, "-lm"])
    return _SO


class _Scene(C.Structure):
    _fields_ = [("P", C.c_int), ("D", C.c_int), ("M", C.c_int), ("W", C.c_int), ("H", C.c_int),
                ("tan_fovx", C.c_float), ("tan_fovy", C.c_float), ("scale_modifier", C.c_float),
                ("antialiasing", C.c_int)] + [(n, C.c_void_p) for n in (
                    "bg", "means3D", "shs", "colors_precomp", "opacities", "scales", "rotations",
                    "cov3D_precomp", "view", "proj", "campos")]


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.gso_scan.restype = C.c_uint32
        _lib.gso_higher_msb.restype = C.c_uint32
        _lib.gso_num_threads.restype = C.c_int
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(np.asarray(a, dtype=np.float32))


class Scene:
    """Holds the inputs of one rasterizer call (same fields as rasterizer.h:35-58)."""

    def __init__(self, *, W, H, tanfovx, tanfovy, bg, means3D, opacities, viewmatrix, projmatrix, campos,
                 sh_degree=0, shs=None, colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None,
                 scale_modifier=1.0, antialiasing=False):
        self.W, self.H = int(W), int(H)
        self.means3D = _f32(means3D).reshape(-1, 3)
        self.P = self.means3D.shape[0]
        self.shs = _f32(shs)
        self.M = 0 if shs is None else self.shs.shape[1]
        self.D = int(sh_degree)
        self.colors_precomp = _f32(colors_precomp)
        self.opacities = _f32(opacities).reshape(-1)
        self.scales, self.rotations, self.cov3D_precomp = _f32(scales), _f32(rotations), _f32(cov3D_precomp)
        self.bg, self.view, self.proj, self.campos = _f32(bg), _f32(viewmatrix).reshape(16), \
            _f32(projmatrix).reshape(16), _f32(campos)
        self.antialiasing = bool(antialiasing)
        self.c = _Scene(self.P, self.D, self.M, self.W, self.H, float(tanfovx), float(tanfovy),
                        float(scale_modifier), int(self.antialiasing), _p(self.bg), _p(self.means3D),
                        _p(self.shs), _p(self.colors_precomp), _p(self.opacities), _p(self.scales),
                        _p(self.rotations), _p(self.cov3D_precomp), _p(self.view), _p(self.proj),
                        _p(self.campos))
        self.tiles_x, self.tiles_y = (self.W + 15) // 16, (self.H + 15) // 16


def forward(sc: Scene, want_invdepth=True):
    """Full forward; returns a dict with every intermediate buffer of SURVEY.md section 8a/A0."""
    L, P = lib(), sc.P
    o = dict(radii=np.zeros(P, np.int32), depths=np.zeros(P, np.float32), means2D=np.zeros((P, 2), np.float32),
             cov3D=np.zeros((P, 6), np.float32), conic_opacity=np.zeros((P, 4), np.float32),
             rgb=np.zeros((P, 3), np.float32), clamped=np.zeros((P, 3), np.uint8),
             tiles_touched=np.zeros(P, np.uint32), point_offsets=np.zeros(P, np.uint32))
    L.gso_preprocess(C.byref(sc.c), _p(o["radii"]), _p(o["depths"]), _p(o["means2D"]), _p(o["cov3D"]),
                     _p(o["conic_opacity"]), _p(o["rgb"]), _p(o["clamped"]), _p(o["tiles_touched"]))
    R = int(L.gso_scan(P, _p(o["tiles_touched"]), _p(o["point_offsets"])))
    o["num_rendered"] = R
    o["keys_unsorted"], o["vals_unsorted"] = np.zeros(R, np.uint64), np.zeros(R, np.uint32)
    o["keys"], o["point_list"] = np.zeros(R, np.uint64), np.zeros(R, np.uint32)
    L.gso_duplicate_with_keys(P, sc.W, sc.H, _p(o["means2D"]), _p(o["depths"]), _p(o["point_offsets"]),
                              _p(o["radii"]), _p(o["keys_unsorted"]), _p(o["vals_unsorted"]))
    tiles = sc.tiles_x * sc.tiles_y
    bit = int(L.gso_higher_msb(C.c_uint32(tiles)))
    o["key_bits"] = 32 + bit
    L.gso_sort_pairs(C.c_size_t(R), _p(o["keys_unsorted"]), _p(o["vals_unsorted"]), _p(o["keys"]),
                     _p(o["point_list"]), 32 + bit)
    o["ranges"] = np.zeros((tiles, 2), np.uint32)
    L.gso_tile_ranges(C.c_size_t(R), _p(o["keys"]), tiles, _p(o["ranges"]))
    feats = sc.colors_precomp if sc.colors_precomp is not None else o["rgb"]
    o["color"] = np.zeros((3, sc.H, sc.W), np.float32)
    o["final_T"] = np.zeros(sc.H * sc.W, np.float32)
    o["n_contrib"] = np.zeros(sc.H * sc.W, np.uint32)
    o["invdepth"] = np.zeros((1, sc.H, sc.W), np.float32) if want_invdepth else None
    L.gso_render(sc.W, sc.H, _p(o["ranges"]), _p(o["point_list"]), _p(o["means2D"]), _p(feats),
                 _p(o["conic_opacity"]), _p(o["depths"]), _p(sc.bg), _p(o["color"]), _p(o["final_T"]),
                 _p(o["n_contrib"]), _p(o["invdepth"]))
    return o


def backward(sc: Scene, fw: dict, dL_dpix, dL_dinvdepth=None):
    """Backward for the forward state `fw`; returns the 8 reference gradients (+ internals)."""
    L, P = lib(), sc.P
    dL_dpix = _f32(dL_dpix)
    dL_dinvdepth = _f32(dL_dinvdepth)
    g = dict(dL_dmean2D=np.zeros((P, 3), np.float32), dL_dconic=np.zeros((P, 4), np.float32),
             dL_dopacity=np.zeros((P, 1), np.float32), dL_dcolor=np.zeros((P, 3), np.float32),
             dL_dinvd=np.zeros(P, np.float32) if dL_dinvdepth is not None else None,
             dL_dmean3D=np.zeros((P, 3), np.float32), dL_dcov3D=np.zeros((P, 6), np.float32),
             dL_dsh=np.zeros((P, sc.M, 3), np.float32), dL_dscale=np.zeros((P, 3), np.float32),
             dL_drot=np.zeros((P, 4), np.float32))
    feats = sc.colors_precomp if sc.colors_precomp is not None else fw["rgb"]
    L.gso_render_backward(P, sc.W, sc.H, _p(fw["ranges"]), _p(fw["point_list"]), _p(sc.bg), _p(fw["means2D"]),
                          _p(fw["conic_opacity"]), _p(feats), _p(fw["depths"]), _p(fw["final_T"]),
                          _p(fw["n_contrib"]), _p(dL_dpix), _p(dL_dinvdepth), _p(g["dL_dmean2D"]),
                          _p(g["dL_dconic"]), _p(g["dL_dopacity"]), _p(g["dL_dcolor"]), _p(g["dL_dinvd"]))
    cov3D = sc.cov3D_precomp if sc.cov3D_precomp is not None else fw["cov3D"]
    L.gso_preprocess_backward(C.byref(sc.c), _p(fw["radii"]), _p(cov3D), _p(fw["clamped"]), _p(g["dL_dmean2D"]),
                              _p(g["dL_dconic"]), _p(g["dL_dopacity"]), _p(g["dL_dcolor"]), _p(g["dL_dinvd"]),
                              _p(g["dL_dmean3D"]), _p(g["dL_dcov3D"]), _p(g["dL_dsh"]), _p(g["dL_dscale"]),
                              _p(g["dL_drot"]))
    return g


def mark_visible(means3D, viewmatrix):
    m = _f32(means3D).reshape(-1, 3)
    v = _f32(viewmatrix).reshape(16)
    out = np.zeros(m.shape[0], np.uint8)
    lib().gso_mark_visible(m.shape[0], _p(m), _p(v), _p(out))
    return out.astype(bool)


def dist2(points):
    pts = _f32(points).reshape(-1, 3)
    out = np.zeros(pts.shape[0], np.float32)
    lib().gso_dist2(pts.shape[0], _p(pts), _p(out))
    return out


def num_threads():
    return int(lib().gso_num_threads())


def set_threads(n):
    lib().gso_set_threads(int(n))
