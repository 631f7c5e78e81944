"""ctypes binding of libgsr.so with the call shapes of the reference's pybind module `_C`
(submodules/diff-gaussian-rasterization/rasterize_points.cu:35-217, ext.cpp:15-19).

torch is used for device memory and the current stream only; every computation happens in the
hand-written sm_100a kernels behind include/gsr.h. There is NO fallback: if libgsr.so is missing or a
tensor is not on a CUDA device, this raises.
"""
import ctypes as C
import os
import torch

_PKG_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB_PATH = os.environ.get("GSR_LIB") or os.path.join(_PKG_ROOT, "libgsr.so")      # GSR_LIB: an experimental build (build.py --variant)


class _View(C.Structure):
    _fields_ = [("P", C.c_int), ("D", C.c_int), ("M", C.c_int), ("W", C.c_int), ("H", C.c_int),
                ("tan_fovx", C.c_float), ("tan_fovy", C.c_float), ("scale_modifier", C.c_float),
                ("prefiltered", C.c_int), ("antialiasing", C.c_int), ("debug", C.c_int)] + \
               [(n, C.c_void_p) for n in ("background", "means3D", "shs", "colors_precomp", "opacities", "scales",
                                          "rotations", "cov3D_precomp", "viewmatrix", "projmatrix", "cam_pos", "sh_dc")]


class _Grads(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("dL_dmeans2D", "dL_dopacity", "dL_dmeans3D", "dL_dcolors", "dL_dcov3D",
                                          "dL_dsh", "dL_dscales", "dL_drotations", "dL_dconic", "dL_ddc")] + \
               [("accumulate", C.c_int), ("accum_clean", C.c_int)]


_ALLOC_FN = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_size_t)
_lib = None
_ext = None         # the compiled torch binding (csrc/torch_binding.cpp -> _gsr_torch.so), False when absent / disabled


def torch_ext():
    """The thin PyTorch C++ extension over the same C ABI, preferred for the three hot entry points when it has been
    built (build.py builds it next to libgsr.so): one pybind transition per call instead of ~60 Python-level operations.
    GSR_BINDING=ctypes forces the ctypes binding below; both call the same gsr_forward / gsr_backward."""
    global _ext
    if _ext is None:
        _ext = False
        if os.environ.get("GSR_BINDING", "").lower() != "ctypes" and not os.environ.get("GSR_LIB"):
            path = os.path.join(_PKG_ROOT, "_gsr_torch.so")
            if os.path.exists(path):
                import importlib.util
                lib()                                   # libgsr.so first: the extension links it
                spec = importlib.util.spec_from_file_location("_gsr_torch", path)
                m = importlib.util.module_from_spec(spec)
                spec.loader.exec_module(m)
                _ext = m
    return _ext


def binding_name():
    return "torch-extension" if torch_ext() else "ctypes"


def lib():
    """Loads libgsr.so once. Fails loudly when the CUDA library has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise ImportError(f"{_LIB_PATH} not found: build it with `python gaussian-splatting-slam_b200/build.py` "
                              "(there is no CPU or PyTorch fallback for the rasterizer)")
        L = C.CDLL(_LIB_PATH)
        L.gsr_forward.restype = C.c_int
        L.gsr_forward.argtypes = [C.POINTER(_View), _ALLOC_FN, C.c_void_p, _ALLOC_FN, C.c_void_p, _ALLOC_FN, C.c_void_p,
                                  C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int), C.c_void_p]
        L.gsr_backward.restype = C.c_int
        L.gsr_backward.argtypes = [C.POINTER(_View), C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_void_p, C.POINTER(_Grads), C.c_void_p]
        L.gsr_mark_visible.restype = C.c_int
        L.gsr_mark_visible.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.gsr_dist2.restype = C.c_int
        L.gsr_dist2.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.gsr_dist2_noalloc.restype = C.c_int
        L.gsr_dist2_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.gsr_debug_blend_stats.restype = C.c_int
        L.gsr_debug_blend_stats.argtypes = [C.POINTER(_View), C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.gsr_debug_exp_mismatches.restype = C.c_int
        L.gsr_debug_exp_mismatches.argtypes = [C.c_void_p, C.c_void_p]
        L.gsr_debug_sorted_keys.restype = C.c_int
        L.gsr_debug_sorted_keys.argtypes = [C.POINTER(_View), C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.gsr_set_speculation.restype = C.c_int
        L.gsr_set_speculation.argtypes = [C.c_int]
        L.gsr_reset_capacity_hints.restype = C.c_int
        L.gsr_speculation_stats.restype = C.c_int
        L.gsr_speculation_stats.argtypes = [C.POINTER(C.c_ulonglong)]
        L.gsr_debug_fp32_peak.restype = C.c_int
        L.gsr_debug_fp32_peak.argtypes = [C.c_int, C.c_void_p, C.POINTER(C.c_int), C.c_void_p]
        for n in ("gsr_required_geom", "gsr_required_image", "gsr_backward_scratch_bytes", "gsr_dist2_temp_bytes"):
            getattr(L, n).restype = C.c_size_t
        L.gsr_required_binning.restype = C.c_size_t
        L.gsr_required_binning.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int]
        L.gsr_launch_count.restype = C.c_uint64
        L.gsr_profile_enable.restype = C.c_int
        L.gsr_profile_enable.argtypes = [C.c_int]
        L.gsr_profile_report.restype = C.c_long
        L.gsr_profile_report.argtypes = [C.c_char_p, C.c_size_t]
        L.gsr_error_string.restype = C.c_char_p
        L.gsr_version.restype = C.c_char_p
        for n in ("gsr_geom_layout", "gsr_image_layout"):
            getattr(L, n).restype = C.c_int
        L.gsr_binning_layout.restype = C.c_int
        _lib = L
    return _lib


_EMPTY = torch.empty(0)


def _check(code, what):
    if code != 0:
        raise RuntimeError(f"{what}: {lib().gsr_error_string(code).decode()} (code {code})")


def _dev(t, name):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (the rasterizer has no CPU path)")
    return t


def _f32(t):
    """contiguous fp32 view of t (rasterize_points.cu:94-111 calls .contiguous() on every input)."""
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _ptr(t):
    return None if t is None or t.numel() == 0 else t.data_ptr()


def _round_capacity(nbytes):
    """Scratch sizes follow num_rendered, which changes from view to view; rounding large requests up to
    {1, 1.25, 1.5, 1.75} x 2^k keeps them within a handful of size classes, so torch's caching allocator serves them
    from its pool instead of calling cudaMalloc (a device-wide stall) for every new camera."""
    if nbytes <= (1 << 20):
        return nbytes
    k = nbytes.bit_length() - 1
    for m in (4, 5, 6, 7, 8):
        c = (m << k) >> 2
        if c >= nbytes:
            return c
    return nbytes


_acc_clean = {}     # (device index, stream id, P) -> zero-filled accumulator of the backward (kept all zero by the library)
_pool = {}          # (device index, stream id) -> free scratch tensors
_DEBUG_POOL = os.environ.get("GSR_DEBUG_POOL", "0") == "1"
_POOL_MAX = 12


class ScratchLease:
    """Scratch blobs of one forward, checked out of a per-(device, stream) pool and handed back when the autograd
    context that owns the lease dies (after its backward, or when the graph is dropped). The blob sizes follow
    num_rendered, i.e. change with every camera; serving them from torch's caching allocator directly makes it
    split and re-malloc for dozens of iterations (13 GB reserved, cudaMalloc stalls). Reuse is stream-ordered: a
    buffer only ever goes back to the pool of the stream its forward ran on."""

    def __init__(self):
        self.key, self.bufs = None, []

    def take(self, device, nbytes):
        if self.key is None:
            self.key = (device.index, torch.cuda.current_stream(device).cuda_stream)
        free = _pool.setdefault(self.key, [])
        best = -1
        for i, t in enumerate(free):
            if t.numel() >= nbytes and (best < 0 or t.numel() < free[best].numel()):
                best = i
        if best >= 0 and free[best].numel() <= 2 * nbytes + (8 << 20):
            t = free.pop(best)
        else:
            t = torch.empty(_round_capacity(nbytes), dtype=torch.uint8, device=device)
            if _DEBUG_POOL:
                print(f"[gsr pool] miss: want {nbytes >> 20} MiB, free {[x.numel() >> 20 for x in free]} -> new {t.numel() >> 20} MiB")
        self.bufs.append(t)
        return t

    def release(self):
        if self.key is not None and _pool is not None:
            free = _pool.setdefault(self.key, [])
            for t in self.bufs:
                if len(free) < _POOL_MAX:
                    free.append(t)
        self.bufs, self.key = [], None

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass


def empty_scratch_pool():
    """Drops every pooled scratch buffer (the torch caching allocator gets the memory back)."""
    _pool.clear()
    _acc_clean.clear()
    if torch_ext():
        torch_ext().drop_cached_scratch()


class _Scratch:
    """The three resizable byte buffers of rasterize_points.cu:68-78, handed out through callbacks."""

    def __init__(self, device, lease=None):
        self.device = device
        self.bufs = {}
        self.err = None
        self.lease = lease

        def make(key):
            def cb(_ctx, nbytes):
                try:
                    if self.lease is not None:
                        t = self.lease.take(self.device, int(nbytes))
                    else:
                        t = torch.empty(_round_capacity(int(nbytes)), dtype=torch.uint8, device=self.device)
                    self.bufs[key] = t
                    return t.data_ptr()
                except Exception as ex:   # surfaces as GSR_E_ALLOC, re-raised by the caller
                    self.err = ex
                    return 0
            return _ALLOC_FN(cb)
        self.cbs = [make(k) for k in ("geom", "binning", "image")]

    def get(self, key):
        t = self.bufs.get(key)
        return t if t is not None else torch.empty(0, dtype=torch.uint8, device=self.device)

    def close(self):
        """The ctypes callbacks close over self: break that cycle, or the blobs (hundreds of MB) and the lease stay
        alive until Python's cyclic collector happens to run."""
        out = self.get("geom"), self.get("binning"), self.get("image")
        self.cbs, self.bufs, self.lease = None, {}, None
        return out


def _make_view(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
               projmatrix, tan_fovx, tan_fovy, H, W, sh, degree, campos, prefiltered, debug, antialiasing, keep, dc=None):
    M = sh.size(1) if sh.numel() != 0 else 0          # rasterize_points.cu:84-88
    if dc is not None and dc.numel() != 0:            # separate_sh: dc [P,1,3] + the M-1 higher coefficients in `sh`
        M += 1
    ts = dict(background=bg, means3D=means3D, shs=sh, colors_precomp=colors, opacities=opacity, scales=scales,
              rotations=rotations, cov3D_precomp=cov3D_precomp, viewmatrix=viewmatrix, projmatrix=projmatrix,
              cam_pos=campos, sh_dc=dc)
    ptrs = {}
    for k, t in ts.items():
        if t is None or t.numel() == 0:
            ptrs[k] = None
        else:
            t = _f32(_dev(t, k))
            keep.append(t)
            ptrs[k] = t.data_ptr()
    return _View(int(means3D.size(0)), int(degree), int(M), int(W), int(H), float(tan_fovx), float(tan_fovy),
                 float(scale_modifier), int(bool(prefiltered)), int(bool(antialiasing)), int(bool(debug)),
                 ptrs["background"], ptrs["means3D"], ptrs["shs"], ptrs["colors_precomp"], ptrs["opacities"],
                 ptrs["scales"], ptrs["rotations"], ptrs["cov3D_precomp"], ptrs["viewmatrix"], ptrs["projmatrix"],
                 ptrs["cam_pos"], ptrs["sh_dc"])


def rasterize_gaussians(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                        projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos, prefiltered,
                        debug, antialiasing=False, want_invdepth=True, lease=None, dc=None):
    """rasterize_points.cu:35-115 + superset outputs (`dc`: the degree-0 SH coefficient as its own [P,1,3] tensor,
    `sh` then holds the higher ones -- the caller's separate_sh call shape). Returns
    (num_rendered, color[3,H,W], radii[P] int32, geomBuffer, binningBuffer, imgBuffer, invdepth[1,H,W])."""
    if means3D.ndim != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")   # rasterize_points.cu:57-59
    _dev(means3D, "means3D")
    ext = torch_ext()
    if ext:
        n, color, radii, geom, binning, image, invdepth = ext.rasterize_gaussians(
            bg, means3D, colors, opacity, scales, rotations, float(scale_modifier), cov3D_precomp, viewmatrix, projmatrix,
            float(tan_fovx), float(tan_fovy), int(image_height), int(image_width), sh, int(degree), campos, bool(prefiltered),
            bool(debug), bool(antialiasing), bool(want_invdepth), dc if dc is not None else _EMPTY)
        return n, color, radii, geom, binning, image, (invdepth if want_invdepth else None)
    H, W, P = int(image_height), int(image_width), int(means3D.size(0))
    dev = means3D.device
    with torch.cuda.device(dev):
        keep = []
        view = _make_view(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                          projmatrix, tan_fovx, tan_fovy, H, W, sh, degree, campos, prefiltered, debug, antialiasing, keep,
                          dc=dc)
        color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
        radii = torch.empty((P,), dtype=torch.int32, device=dev)
        invdepth = torch.empty((1, H, W), dtype=torch.float32, device=dev) if want_invdepth else None
        scratch = _Scratch(dev, lease)
        n = C.c_int(0)
        rc = lib().gsr_forward(C.byref(view), scratch.cbs[0], None, scratch.cbs[1], None, scratch.cbs[2], None,
                               color.data_ptr(), _ptr(radii), _ptr(invdepth), C.byref(n),
                               torch.cuda.current_stream(dev).cuda_stream)
        err = scratch.err
        geom, binning, image = scratch.close()
        if err is not None:
            raise err
        _check(rc, "gsr_forward")
    return n.value, color, radii, geom, binning, image, invdepth


def rasterize_gaussians_backward(bg, means3D, radii, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                                 viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, dL_dout_invdepth, sh,
                                 degree, campos, geomBuffer, R, binningBuffer, imageBuffer, antialiasing, debug,
                                 want_conic=False, grad_sinks=None, lease=None, dc=None):
    """rasterize_points.cu:117-196. Returns (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh,
    dL_dscales, dL_drotations); gradients of inputs that were not supplied are None. With want_conic=True a 9th
    tensor dL_dconic[P,2,2] is appended (parity tests). With `dc` (separate_sh) one more tensor, dL_ddc[P,1,3], comes
    last and dL_dsh covers the higher coefficients only (grad_sinks key: "dc").
    grad_sinks (superset, multi-view batches): dict of fp32 contiguous CUDA tensors keyed by input name ("means3D",
    "opacities", "shs" | "colors_precomp", "scales" + "rotations" | "cov3D_precomp"), one for EVERY supplied input. The
    gradients of this view are then ADDED into them inside the kernel (gsr_grads.accumulate) and the corresponding
    returned entries are None; dL_dmeans2D is still returned per view. With the extra entry `"overwrite": True` the sinks
    are WRITTEN in full instead (zeros for culled Gaussians) -- the first view of a batch then needs no cleared sinks."""
    overwrite = bool(grad_sinks.get("overwrite", False)) if grad_sinks is not None else False
    ext = torch_ext()
    if ext:
        has_dc = dc is not None and dc.numel() != 0
        sinks = []
        if grad_sinks is not None:
            sinks = [grad_sinks.get(k) for k in ("means3D", "opacities", "colors_precomp", "cov3D_precomp", "shs", "scales",
                                                 "rotations", "dc")]
        res = ext.rasterize_gaussians_backward(
            bg, means3D, radii, colors, opacity, scales, rotations, float(scale_modifier), cov3D_precomp, viewmatrix,
            projmatrix, float(tan_fovx), float(tan_fovy), dL_dout_color, dL_dout_invdepth, sh, int(degree), campos, geomBuffer,
            int(R), binningBuffer, imageBuffer, bool(antialiasing), bool(debug), bool(want_conic), grad_sinks is not None, sinks,
            dc if dc is not None else _EMPTY, overwrite)
        out = tuple(res[:8])
        if want_conic:
            out = out + (res[8],)
        return out + (res[9],) if has_dc else out
    P = int(means3D.size(0))
    H, W = int(dL_dout_color.size(1)), int(dL_dout_color.size(2))
    dev = means3D.device
    with torch.cuda.device(dev):
        keep = []
        view = _make_view(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                          projmatrix, tan_fovx, tan_fovy, H, W, sh, degree, campos, False, debug, antialiasing, keep,
                          dc=dc)
        has_dc = dc is not None and dc.numel() != 0
        M = view.M - (1 if has_dc else 0)             # rows of `sh` / dL_dsh
        f = dict(dtype=torch.float32, device=dev)
        dL_dmeans2D = torch.empty((P, 3), **f)
        dL_dconic = torch.empty((P, 2, 2), **f) if want_conic else None
        shapes = dict(means3D=(P, 3), opacities=(P, 1), colors_precomp=(P, 3) if colors.numel() != 0 else None,
                      cov3D_precomp=(P, 6) if cov3D_precomp.numel() != 0 else None,
                      shs=(P, M, 3) if sh.numel() != 0 else None, dc=(P, 1, 3) if has_dc else None,
                      scales=(P, 3) if scales.numel() != 0 else None,
                      rotations=(P, 4) if rotations.numel() != 0 else None)
        if grad_sinks is None:
            o = {k: (torch.empty(shp, **f) if shp is not None else None) for k, shp in shapes.items()}
        else:
            o = {}
            for k, shp in shapes.items():
                if shp is None:
                    o[k] = None
                    continue
                t = grad_sinks.get(k)
                if t is None:
                    raise RuntimeError(f"grad_sinks: no sink for supplied input '{k}' (accumulation is all-or-nothing)")
                if (not t.is_cuda or t.device != dev or t.dtype != torch.float32 or not t.is_contiguous()
                        or t.numel() != torch.Size(shp).numel()):
                    raise RuntimeError(f"grad_sinks['{k}'] must be a contiguous fp32 tensor of {shp} on {dev}")
                o[k] = t
        dL_dmeans3D, dL_dopacity, dL_dcolors, dL_dcov3D = o["means3D"], o["opacities"], o["colors_precomp"], o["cov3D_precomp"]
        dL_dsh, dL_dscales, dL_drot, dL_ddc = o["shs"], o["scales"], o["rotations"], o["dc"]
        if P != 0:
            dpix = _f32(dL_dout_color)
            dinv = _f32(dL_dout_invdepth) if dL_dout_invdepth is not None else None
            stream = torch.cuda.current_stream(dev).cuda_stream
            # The packed gradient accumulator (48 B per Gaussian) is kept per (device, stream, P) and stays ALL ZERO between
            # calls: the backward zeroes the rows it read (gsr_grads.accum_clean), so no clear per view.
            key = (dev.index, stream, P)
            acc = _acc_clean.get(key)
            if acc is None:
                if len(_acc_clean) >= 8:
                    _acc_clean.clear()
                acc = torch.zeros(int(lib().gsr_backward_scratch_bytes(P)), dtype=torch.uint8, device=dev)
                _acc_clean[key] = acc
            g = _Grads(_ptr(dL_dmeans2D), _ptr(dL_dopacity), _ptr(dL_dmeans3D), _ptr(dL_dcolors), _ptr(dL_dcov3D),
                       _ptr(dL_dsh), _ptr(dL_dscales), _ptr(dL_drot), _ptr(dL_dconic), _ptr(dL_ddc),
                       0 if (grad_sinks is None or overwrite) else 1, 1)
            rc = lib().gsr_backward(C.byref(view), int(R), _ptr(radii), _ptr(geomBuffer), _ptr(binningBuffer),
                                    _ptr(imageBuffer), dpix.data_ptr(), _ptr(dinv), acc.data_ptr(), C.byref(g), stream)
            if rc != 0:
                _acc_clean.pop(key, None)               # possibly dirty now
            _check(rc, "gsr_backward")
    if grad_sinks is not None:
        dL_dcolors = dL_dopacity = dL_dmeans3D = dL_dcov3D = dL_dsh = dL_dscales = dL_drot = dL_ddc = None
    out = (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drot)
    if want_conic:
        out = out + (dL_dconic,)
    return out + (dL_ddc,) if has_dc else out


def mark_visible(means3D, viewmatrix, projmatrix):
    """rasterize_points.cu:198-217"""
    _dev(means3D, "means3D")
    ext = torch_ext()
    if ext:
        return ext.mark_visible(means3D, viewmatrix, projmatrix)
    P = int(means3D.size(0))
    dev = means3D.device
    present = torch.zeros((P,), dtype=torch.bool, device=dev)
    if P != 0:
        with torch.cuda.device(dev):
            m, v, p = _f32(means3D), _f32(_dev(viewmatrix, "viewmatrix")), _f32(_dev(projmatrix, "projmatrix"))
            _check(lib().gsr_mark_visible(P, m.data_ptr(), v.data_ptr(), p.data_ptr(), present.data_ptr(),
                                          torch.cuda.current_stream(dev).cuda_stream), "gsr_mark_visible")
    return present


def launch_count():
    return int(lib().gsr_launch_count())


def profile_enable(on):
    """Per-launch CUDA-event timing inside libgsr.so (bench.py's roofline leg); enabling clears the log."""
    _check(lib().gsr_profile_enable(1 if on else 0), "gsr_profile_enable")


def profile_report():
    """'<kernel> <launches> <total_ms>' lines for everything launched since profile_enable(True)."""
    buf = C.create_string_buffer(1 << 16)
    n = lib().gsr_profile_report(buf, len(buf))
    if n < 0:
        _check(int(-n), "gsr_profile_report")
    return buf.value.decode()


def decode_state(P, R, W, H, geomBuffer, binningBuffer, imgBuffer, view_args=None):
    """Parity helper: slices the private scratch blobs into named tensors (layouts from gsr_*_layout)."""
    L = lib()
    off = (C.c_size_t * 8)()

    def base(t):
        return (-t.data_ptr()) % 128

    def take(buf, o, nbytes, dtype, shape):
        b0 = base(buf)
        return buf[b0 + o: b0 + o + nbytes].view(dtype).view(*shape)
    out = {}
    n = L.gsr_geom_layout(C.c_int(P), off, 8)
    go = [off[i] for i in range(n)]
    out["depths"] = take(geomBuffer, go[0], 4 * P, torch.float32, (P,))
    # splat records: cull part {x,y,ex,ey}{a,b,c,opacity} as [P,8], then the colour part {r,g,b,1/depth} as [P,4]
    cull = take(geomBuffer, go[1], 32 * P, torch.float32, (P, 8))
    colr = take(geomBuffer, go[1] + 32 * P, 16 * P, torch.float32, (P, 4))
    out["means2D"], out["extent"] = cull[:, 0:2], cull[:, 2:4]
    out["conic_opacity"], out["rgb"], out["invdepth_g"] = cull[:, 4:8], colr[:, 0:3], colr[:, 3]
    out["clamped"] = take(geomBuffer, go[2], P, torch.uint8, (P,))
    if (W + 15) // 16 <= 255 and (H + 15) // 16 <= 255 and len(go) > 7:
        # grids up to 255 x 255 tiles: the library keeps the packed tile rectangle (x0 | y0 << 8 | w << 16 | h << 24)
        # instead of the count; the reference's tiles_touched is its width x height
        rect = take(geomBuffer, go[7], 4 * P, torch.int32, (P,))
        out["tile_rect"] = rect
        out["tiles_touched"] = ((rect >> 16) & 255) * ((rect >> 24) & 255)
    else:
        out["tiles_touched"] = take(geomBuffer, go[3], 4 * P, torch.int32, (P,))
    out["point_offsets"] = take(geomBuffer, go[4], 4 * P, torch.int32, (P,))
    out["order"] = take(geomBuffer, go[5], 4 * P, torch.int32, (P,))
    n = L.gsr_image_layout(C.c_int(W), C.c_int(H), off, 8)
    io = [off[i] for i in range(n)]
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    out["final_T"] = take(imgBuffer, io[0], 4 * W * H, torch.float32, (H * W,))
    out["n_contrib"] = take(imgBuffer, io[1], 4 * W * H, torch.int32, (H * W,))
    out["ranges"] = take(imgBuffer, io[2], 8 * tiles, torch.int32, (tiles, 2))
    if R > 0:
        n = L.gsr_binning_layout(C.c_int(R), C.c_int(W), C.c_int(H), off, 8)
        bo = [off[i] for i in range(n)]
        out["point_list"] = take(binningBuffer, bo[0], 4 * R, torch.int32, (R,))
    return out


def sorted_keys(P, R, W, H, geomBuffer, binningBuffer, imgBuffer):
    """Parity helper: the reference's sorted u64 keys (as int64) for the last forward (tile id from the ranges the
    instance falls into, depth bits of its Gaussian)."""
    dev = geomBuffer.device
    keys = torch.empty((R,), dtype=torch.int64, device=dev)
    if R > 0:
        view = _View()
        view.P, view.W, view.H = P, W, H
        with torch.cuda.device(dev):
            _check(lib().gsr_debug_sorted_keys(C.byref(view), R, geomBuffer.data_ptr(), binningBuffer.data_ptr(),
                                               imgBuffer.data_ptr(), keys.data_ptr(),
                                               torch.cuda.current_stream(dev).cuda_stream), "gsr_debug_sorted_keys")
    return keys


def set_speculation(on):
    """gsr_set_speculation: False = size the binning blob from the counted num_rendered (one mid-forward sync, the
    reference's route); True (default) = size it from the largest count seen for this image size."""
    _check(lib().gsr_set_speculation(1 if on else 0), "gsr_set_speculation")


def reset_capacity_hints():
    _check(lib().gsr_reset_capacity_hints(), "gsr_reset_capacity_hints")


def speculation_stats():
    """(forwards whose capacity estimate held, forwards that repeated binning + blend, forwards that counted first)"""
    out = (C.c_ulonglong * 3)()
    _check(lib().gsr_speculation_stats(out), "gsr_speculation_stats")
    return int(out[0]), int(out[1]), int(out[2])


def fp32_peak_tflops(device="cuda", iters=4000, reps=5):
    """Measured FP32 SIMT ceiling of the device (gsr_debug_fp32_peak, timed with CUDA events; best of reps)."""
    dev = torch.device(device)
    with torch.cuda.device(dev):
        sms = C.c_int(0)
        out = torch.empty(256 * 8 * 256, dtype=torch.float32, device=dev)       # up to 256 SMs
        st = torch.cuda.current_stream(dev).cuda_stream
        best = None
        for _ in range(reps + 1):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            _check(lib().gsr_debug_fp32_peak(iters, out.data_ptr(), C.byref(sms), st), "gsr_debug_fp32_peak")
            e1.record()
            e1.synchronize()
            ms = e0.elapsed_time(e1)
            best = ms if best is None else min(best, ms)
        flop = 2.0 * 16 * iters * 256 * 8 * sms.value
    return flop / (best * 1e-3) / 1e12


def exp_mismatches(device="cuda"):
    """Parity helper: how many of the 2^32 float bit patterns the blend kernels' packed exp gets differently from expf."""
    dev = torch.device(device)
    n = torch.zeros(1, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        _check(lib().gsr_debug_exp_mismatches(n.data_ptr(), torch.cuda.current_stream(dev).cuda_stream),
               "gsr_debug_exp_mismatches")
    return int(n.item())


def set_exact_cull(on):
    """Parity helper (gsr_debug_set_exact_cull): switch the blend kernels' exact sub-tile cull off / on."""
    _check(lib().gsr_debug_set_exact_cull(1 if on else 0), "gsr_debug_set_exact_cull")


def blend_stats(P, R, W, H, geomBuffer, binningBuffer, imgBuffer):
    """Work counters of a forward: (pairs up to each pixel's last contributor, contributing pairs)."""
    dev = geomBuffer.device
    n = torch.zeros(8, dtype=torch.int64, device=dev)
    if R > 0:
        view = _View()
        view.P, view.W, view.H = P, W, H
        with torch.cuda.device(dev):
            _check(lib().gsr_debug_blend_stats(C.byref(view), R, geomBuffer.data_ptr(), binningBuffer.data_ptr(),
                                               imgBuffer.data_ptr(), n.data_ptr(),
                                               torch.cuda.current_stream(dev).cuda_stream), "gsr_debug_blend_stats")
    t = n.tolist()
    blend_stats.last_cull = [int(x) for x in t[2:]]
    return int(t[0]), int(t[1])
