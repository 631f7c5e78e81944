"""Map densification / pruning on the device, on top of libgsr.so's gsr_densify_* entry points.

Host-side mirror of the GaussianModel methods the reference's training loop calls (scene/gaussian_model.py):
  add_densification_stats(model, viewspace_point_tensor, update_filter, radii=None)   :431-433 (+ train.py:159)
  densify_and_prune(model, max_grad, min_opacity, extent, max_screen_size, radii)     :412-429 (clone :393-408,
                                                                                       split :368-391, prune :419-425)
  prune_points(model, mask)                                                            :307-322
  reset_opacity(model)                                                                 :226-229
`model` is the reference's GaussianModel or anything with the same attributes (_xyz, _features_dc, _features_rest,
_opacity, _scaling, _rotation, optimizer with named param_groups, xyz_gradient_accum, denom, max_radii2D,
percent_dense). DeviceDensifyMixin carries the same four methods under the reference's names and signatures, to be
mixed into GaussianModel.

Where the reference runs ~30 mask-gather / cat ops over 18 tensors with a host sync at each, this is: one plan (3 small
kernels), ONE 16-byte read-back of the new row count, one apply launch that moves parameters and Adam moments together.
Row order, the optimiser's param_groups / state keys and the reset of the statistics are the reference's.
CUDA tensors only; there is no CPU path.
"""
import ctypes as C

import torch
from torch import nn

from diff_gaussian_rasterization import _C as _gsr

ROLE_COPY, ROLE_XYZ, ROLE_SCALING = 0, 1, 2
MAX_TENSORS = 8
GROUPS = (("xyz", "_xyz", ROLE_XYZ), ("f_dc", "_features_dc", ROLE_COPY), ("f_rest", "_features_rest", ROLE_COPY),
          ("opacity", "_opacity", ROLE_COPY), ("scaling", "_scaling", ROLE_SCALING), ("rotation", "_rotation", ROLE_COPY))


class _Rule(C.Structure):
    _fields_ = [("grad_threshold", C.c_float), ("min_opacity", C.c_float), ("dense_extent", C.c_float),
                ("big_extent", C.c_float), ("max_screen_size", C.c_float), ("prune_big", C.c_int), ("n_split", C.c_int),
                ("split_shrink_inv", C.c_float)]


class _Tensor(C.Structure):
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("src_m", C.c_void_p), ("dst_m", C.c_void_p),
                ("src_v", C.c_void_p), ("dst_v", C.c_void_p), ("row_len", C.c_int), ("role", C.c_int)]


_ready = False


def _lib():
    global _ready
    L = _gsr.lib()
    if not _ready:
        L.gsr_densify_plan_bytes.restype = C.c_size_t
        L.gsr_densify_plan_bytes.argtypes = [C.c_int]
        L.gsr_densify_plan.restype = C.c_int
        L.gsr_densify_plan.argtypes = [C.c_int] + [C.c_void_p] * 4 + [C.POINTER(_Rule), C.c_void_p, C.c_void_p, C.c_void_p]
        L.gsr_prune_plan.restype = C.c_int
        L.gsr_prune_plan.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.gsr_densify_apply.restype = C.c_int
        L.gsr_densify_apply.argtypes = [C.c_int, C.c_void_p, C.POINTER(_Rule), C.POINTER(_Tensor), C.c_int, C.c_void_p,
                                        C.c_void_p, C.c_void_p, C.c_void_p]
        L.gsr_densify_stats.restype = C.c_int
        L.gsr_densify_stats.argtypes = [C.c_int] + [C.c_void_p] * 7
        _ready = True
    return L


def _check(rc, what):
    if rc != 0:
        raise RuntimeError(f"{what}: {_gsr.lib().gsr_error_string(rc).decode()} (code {rc})")


def _need_cuda(t, what, dtype=torch.float32):
    if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == dtype and t.is_contiguous()):
        raise RuntimeError(f"gsr densify: {what} must be a contiguous {dtype} CUDA tensor (there is no CPU path)")
    return t


def _np32(x):
    """Python double -> the float32 a torch comparison against a float tensor uses."""
    return float(torch.tensor(x, dtype=torch.float64).to(torch.float32))


def make_rule(max_grad, min_opacity, extent, max_screen_size, percent_dense, N=2):
    if not max_grad > 0:
        raise ValueError("gsr densify: max_grad must be positive (clones, whose gradient is zero, must not be split)")
    r = _Rule()
    r.grad_threshold, r.min_opacity = float(max_grad), float(min_opacity)
    r.dense_extent, r.big_extent = float(percent_dense * extent), float(0.1 * extent)
    r.prune_big = 1 if max_screen_size else 0
    r.max_screen_size = float(max_screen_size) if max_screen_size else 0.0
    r.n_split = int(N)
    r.split_shrink_inv = 1.0 / _np32(0.8 * N)          # c_float rounds the quotient to fp32, as the CUDA division kernel does
    return r


def stats_raw(dL_dmean2D, update_filter, radii, grad_accum, denom, max_radii2D=None):
    """gsr_densify_stats on raw tensors. update_filter: bool/uint8 [P] or None (then radii > 0)."""
    P = grad_accum.shape[0]
    _need_cuda(dL_dmean2D, "viewspace gradient"), _need_cuda(grad_accum, "xyz_gradient_accum"), _need_cuda(denom, "denom")
    if dL_dmean2D.shape != (P, 3):
        raise RuntimeError("gsr densify: viewspace gradient must be [P, 3]")
    if update_filter is not None:
        if update_filter.dtype == torch.bool:
            update_filter = update_filter.view(torch.uint8)
        _need_cuda(update_filter, "update_filter", torch.uint8)
    if radii is not None:
        _need_cuda(radii, "radii", torch.int32)
    if max_radii2D is not None:
        _need_cuda(max_radii2D, "max_radii2D")
    dev = grad_accum.device
    with torch.cuda.device(dev):
        rc = _lib().gsr_densify_stats(P, dL_dmean2D.data_ptr(), update_filter.data_ptr() if update_filter is not None else None,
                                      radii.data_ptr() if radii is not None else None, grad_accum.data_ptr(), denom.data_ptr(),
                                      max_radii2D.data_ptr() if max_radii2D is not None else None,
                                      torch.cuda.current_stream(dev).cuda_stream)
    _check(rc, "gsr_densify_stats")


def _apply(P, plan, rule, counts, tensors, scaling, rotation, noise, dev):
    """tensors: list of (src, m, v, role); returns list of (dst, dst_m, dst_v)."""
    c0, c1, c2, c3 = counts
    N = rule.n_split if rule is not None else 0
    P_new = c0 + c1 + N * c3
    out, table = [], []
    for src, m, v, role in tensors:
        shape = (P_new,) + tuple(src.shape[1:])
        dst = torch.empty(shape, dtype=torch.float32, device=dev)
        dm = torch.empty(shape, dtype=torch.float32, device=dev) if m is not None else None
        dv = torch.empty(shape, dtype=torch.float32, device=dev) if v is not None else None
        out.append((dst, dm, dv))
        row_len = 1
        for d in src.shape[1:]:
            row_len *= int(d)
        if row_len == 0 or P == 0 or P_new == 0:
            continue                                       # e.g. f_rest at SH degree 0: [P, 0, 3]
        table.append(_Tensor(src.data_ptr(), dst.data_ptr(), m.data_ptr() if m is not None else None,
                             dm.data_ptr() if dm is not None else None, v.data_ptr() if v is not None else None,
                             dv.data_ptr() if dv is not None else None, row_len, role))
    L = _lib()
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        for i in range(0, len(table), MAX_TENSORS):
            chunk = table[i:i + MAX_TENSORS]
            arr = (_Tensor * len(chunk))(*chunk)
            rc = L.gsr_densify_apply(P, plan.data_ptr(), C.byref(rule) if rule is not None else None, arr, len(chunk),
                                     scaling.data_ptr() if scaling is not None else None,
                                     rotation.data_ptr() if rotation is not None else None,
                                     noise.data_ptr() if noise is not None and noise.numel() else None, stream)
            _check(rc, "gsr_densify_apply")
    return out, P_new


def _plan_blob(P, dev):
    return torch.empty(int(_lib().gsr_densify_plan_bytes(P)), dtype=torch.uint8, device=dev)


def densify_raw(params, moments, grad_accum, denom, rule, noise=None):
    """params: dict name -> tensor for the six groups of GROUPS; moments: dict name -> (exp_avg, exp_avg_sq) or None.
    Returns (new params, new moments, counts). noise: [N * counts[2], 3] standard normal, drawn here when None."""
    xyz = params["xyz"]
    dev, P = xyz.device, xyz.shape[0]
    for n, _, _ in GROUPS:
        _need_cuda(params[n], n)
        mv = moments.get(n) if moments else None
        if mv is not None:
            _need_cuda(mv[0], n + " exp_avg"), _need_cuda(mv[1], n + " exp_avg_sq")
    _need_cuda(grad_accum, "xyz_gradient_accum"), _need_cuda(denom, "denom")
    plan = _plan_blob(P, dev)
    counts_dev = torch.zeros(4, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = _lib().gsr_densify_plan(P, params["opacity"].data_ptr(), params["scaling"].data_ptr(), grad_accum.data_ptr(),
                                     denom.data_ptr(), C.byref(rule), plan.data_ptr(), counts_dev.data_ptr(),
                                     torch.cuda.current_stream(dev).cuda_stream)
    _check(rc, "gsr_densify_plan")
    counts = [int(c) for c in counts_dev.tolist()]           # the one host synchronisation
    rows = rule.n_split * counts[2]
    if noise is None:
        noise = torch.empty((rows, 3), dtype=torch.float32, device=dev).normal_()     # = torch.normal(mean 0, std 1)
    _need_cuda(noise, "noise")
    if noise.numel() != rows * 3:
        raise RuntimeError(f"gsr densify: noise must hold {rows} x 3 values (n_split x split-selected parents), got {tuple(noise.shape)}")
    tensors = []
    for n, _, role in GROUPS:
        mv = moments.get(n) if moments else None
        tensors.append((params[n], mv[0] if mv else None, mv[1] if mv else None, role))
    out, P_new = _apply(P, plan, rule, counts, tensors, params["scaling"], params["rotation"], noise, dev)
    new_p = {n: out[k][0] for k, (n, _, _) in enumerate(GROUPS)}
    new_m = {n: (out[k][1], out[k][2]) for k, (n, _, _) in enumerate(GROUPS) if out[k][1] is not None}
    return new_p, new_m, counts


def prune_raw(tensors, mask):
    """tensors: list of (tensor, exp_avg or None, exp_avg_sq or None), all with P rows; mask: bool [P], True = remove.
    Returns (list of (tensor, exp_avg, exp_avg_sq) with the kept rows, number kept)."""
    dev, P = mask.device, mask.shape[0]
    if mask.dtype == torch.bool:
        mask = mask.view(torch.uint8)
    _need_cuda(mask, "mask", torch.uint8)
    plan = _plan_blob(P, dev)
    counts_dev = torch.zeros(4, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = _lib().gsr_prune_plan(P, mask.data_ptr(), plan.data_ptr(), counts_dev.data_ptr(),
                                   torch.cuda.current_stream(dev).cuda_stream)
    _check(rc, "gsr_prune_plan")
    counts = [int(c) for c in counts_dev.tolist()]
    for t, m, v in tensors:
        _need_cuda(t, "tensor")
        if t.shape[0] != P:
            raise RuntimeError("gsr densify: every tensor needs one row per mask entry")
    out, P_new = _apply(P, plan, None, counts, [(t, m, v, ROLE_COPY) for t, m, v in tensors], None, None, None, dev)
    return out, P_new


# ------------------------------------------------------------------------------------------------ model level
def _groups_of(model):
    by_name = {g["name"]: g for g in model.optimizer.param_groups}
    return [(n, attr, role, by_name[n]) for n, attr, role in GROUPS]


def _install(model, name, attr, group, tensor, m, v):
    """What _prune_optimizer / cat_tensors_to_optimizer do per group (:287-344): new Parameter, state carried over."""
    old = group["params"][0]
    state = model.optimizer.state.get(old, None)
    new = nn.Parameter(tensor.requires_grad_(True))
    if state is not None:
        state["exp_avg"], state["exp_avg_sq"] = m, v
        del model.optimizer.state[old]
        model.optimizer.state[new] = state
    group["params"][0] = new
    setattr(model, attr, new)


def _moments(model, group):
    st = model.optimizer.state.get(group["params"][0], None)
    if st is None or "exp_avg" not in st:
        return None
    return st["exp_avg"], st["exp_avg_sq"]


def add_densification_stats(model, viewspace_point_tensor, update_filter, radii=None):
    """gaussian_model.py:431-433; pass `radii` to fold in train.py:159's max_radii2D update."""
    stats_raw(viewspace_point_tensor.grad, update_filter, radii, model.xyz_gradient_accum, model.denom,
              model.max_radii2D if radii is not None else None)


def densify_and_prune(model, max_grad, min_opacity, extent, max_screen_size, radii=None, noise=None, N=2):
    """gaussian_model.py:412-429. `radii` is accepted for signature parity (the reference only parks it in tmp_radii).
    Returns the four plan counters."""
    rule = make_rule(max_grad, min_opacity, extent, max_screen_size, model.percent_dense, N)
    groups = _groups_of(model)
    params = {n: g["params"][0].detach() for n, _, _, g in groups}
    moments = {n: _moments(model, g) for n, _, _, g in groups}
    new_p, new_m, counts = densify_raw(params, moments, model.xyz_gradient_accum, model.denom, rule, noise)
    for n, attr, _, g in groups:
        m, v = new_m.get(n, (None, None))
        _install(model, n, attr, g, new_p[n], m, v)
    P_new, dev = new_p["xyz"].shape[0], new_p["xyz"].device
    model.xyz_gradient_accum = torch.zeros((P_new, 1), device=dev)
    model.denom = torch.zeros((P_new, 1), device=dev)
    model.max_radii2D = torch.zeros((P_new,), device=dev)
    model.tmp_radii = None
    return counts


def prune_points(model, mask):
    """gaussian_model.py:307-322 (mask: True = remove)."""
    groups = _groups_of(model)
    tensors = []
    for n, _, _, g in groups:
        mv = _moments(model, g)
        tensors.append((g["params"][0].detach(), mv[0] if mv else None, mv[1] if mv else None))
    extra = [model.xyz_gradient_accum, model.denom, model.max_radii2D]
    tmp = getattr(model, "tmp_radii", None)
    if isinstance(tmp, torch.Tensor) and tmp.element_size() == 4 and tmp.is_cuda:
        extra.append(tmp.contiguous().view(torch.float32))   # rows are moved bit for bit: int32 radii ride along as 4-byte words
    out, _ = prune_raw(tensors + [(t, None, None) for t in extra], mask)
    for k, (n, attr, _, g) in enumerate(groups):
        _install(model, n, attr, g, out[k][0], out[k][1], out[k][2])
    model.xyz_gradient_accum, model.denom, model.max_radii2D = (out[len(groups) + k][0] for k in range(3))
    if len(extra) == 4:
        model.tmp_radii = out[len(groups) + 3][0].view(tmp.dtype)


def reset_opacity(model):
    """gaussian_model.py:226-229 with replace_tensor_to_optimizer (:274-285). Three element-wise torch ops on [P, 1]
    once every opacity_reset_interval iterations: left to torch."""
    x = torch.min(torch.sigmoid(model._opacity.detach()), torch.ones_like(model._opacity) * 0.01)
    new = torch.log(x / (1 - x))
    for g in model.optimizer.param_groups:
        if g["name"] == "opacity":
            _install(model, "opacity", "_opacity", g, new, torch.zeros_like(new), torch.zeros_like(new))


class DeviceDensifyMixin:
    """Mix into the reference's GaussianModel (listed first) to route its densification through libgsr.so."""

    def add_densification_stats(self, viewspace_point_tensor, update_filter, radii=None):
        add_densification_stats(self, viewspace_point_tensor, update_filter, radii)

    def densify_and_prune(self, max_grad, min_opacity, extent, max_screen_size, radii=None):
        densify_and_prune(self, max_grad, min_opacity, extent, max_screen_size, radii)
        torch.cuda.empty_cache()                      # the reference does (:429)

    def prune_points(self, mask):
        prune_points(self, mask)

    def reset_opacity(self):
        reset_opacity(self)
