"""Optimisers for the Gaussian model on top of libgsr.so's one-launch Adam (gsr_adam_step).

FusedAdam            drop-in for the reference's `torch.optim.Adam(l, lr=0.0, eps=1e-15)` (scene/gaussian_model.py:168):
                     same constructor, same param_groups (per-group 'lr' and 'name', which update_learning_rate and the
                     densification helpers index, gaussian_model.py:184-191,300-372), same per-parameter state keys
                     ('step', 'exp_avg', 'exp_avg_sq') so cat_tensors_to_optimizer / _prune_optimizer /
                     replace_tensor_to_optimizer keep working, and state_dict()s interchange with torch.optim.Adam.
SparseGaussianAdam   the class the reference's caller looks for in the rasterizer package (train.py:37-41,
                     gaussian_model.py:24-27,170): step(visibility, N) updates only the Gaussians the view saw and, like
                     the upstream implementation it names, applies no bias correction. Not vendored in the reference tree
                     (SURVEY F1), so its parity is pinned by the oracle only.
All groups of one step go to the device in ONE kernel launch; CUDA tensors only, no fallback.
"""
import ctypes as C
import math

import torch
from diff_gaussian_rasterization import _C as _gsr

MAX_GROUPS = 8


class _Group(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("numel", C.c_longlong), ("row_len", C.c_int), ("step_size", C.c_float), ("inv_sqrt_bc2", C.c_float)]


_ready = False


def _lib():
    global _ready
    L = _gsr.lib()
    if not _ready:
        L.gsr_adam_step.restype = C.c_int
        L.gsr_adam_step.argtypes = [C.POINTER(_Group), C.c_int, C.c_double, C.c_double, C.c_double, C.c_void_p, C.c_void_p]
        _ready = True
    return L


def adam_step_raw(entries, beta1, beta2, eps, visible=None):
    """entries: list of (param, grad, exp_avg, exp_avg_sq, row_len, step_size, inv_sqrt_bc2), contiguous fp32 CUDA tensors
    on one device. Launches gsr_adam_step once per MAX_GROUPS entries."""
    if not entries:
        return
    dev = entries[0][0].device
    if dev.type != "cuda":
        raise RuntimeError("gsr Adam: parameters must be CUDA tensors (there is no CPU path)")
    L = _lib()
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        for i in range(0, len(entries), MAX_GROUPS):
            chunk = entries[i:i + MAX_GROUPS]
            arr = (_Group * len(chunk))()
            for k, (p, g, m, v, row_len, step_size, inv_bc2) in enumerate(chunk):
                for t, nm in ((p, "param"), (g, "grad"), (m, "exp_avg"), (v, "exp_avg_sq")):
                    if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.device == dev):
                        raise RuntimeError(f"gsr Adam: {nm} must be a contiguous fp32 CUDA tensor on {dev} (there is no CPU path)")
                if not (p.numel() == g.numel() == m.numel() == v.numel()):
                    raise RuntimeError("gsr Adam: param / grad / moment sizes differ")
                arr[k] = _Group(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(), int(row_len),
                                float(step_size), float(inv_bc2))
            vis = visible.data_ptr() if visible is not None else None
            rc = L.gsr_adam_step(arr, len(chunk), beta1, beta2, eps, vis, stream)
            if rc != 0:
                raise RuntimeError(f"gsr_adam_step: {L.gsr_error_string(rc).decode()} (code {rc})")


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        if lr < 0.0 or eps < 0.0 or not (0.0 <= betas[0] < 1.0) or not (0.0 <= betas[1] < 1.0):
            raise ValueError("FusedAdam: invalid hyper-parameter")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))

    bias_correction = True

    def _state_of(self, p):
        st = self.state[p]
        if len(st) == 0:
            st["step"] = torch.tensor(0.0)
            st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return st

    def _entries(self, group):
        out = []
        b1, b2 = group["betas"]
        for p in group["params"]:
            if p.grad is None:
                continue
            if p.grad.is_sparse:
                raise RuntimeError("FusedAdam does not support sparse gradients")
            st = self._state_of(p)
            st["step"] += 1
            t = float(st["step"])
            if self.bias_correction:
                step_size, inv_bc2 = group["lr"] / (1.0 - b1 ** t), 1.0 / math.sqrt(1.0 - b2 ** t)
            else:
                step_size, inv_bc2 = group["lr"], 1.0
            row_len = p.numel() // p.shape[0] if p.dim() > 0 and p.shape[0] > 0 else 1
            g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
            out.append((p.data, g, st["exp_avg"], st["exp_avg_sq"], row_len, step_size, inv_bc2))
        return out

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        by_hyper = {}
        for group in self.param_groups:
            key = (group["betas"][0], group["betas"][1], group["eps"])
            by_hyper.setdefault(key, []).extend(self._entries(group))
        for (b1, b2, eps), entries in by_hyper.items():
            adam_step_raw(entries, b1, b2, eps)
        return loss


class SparseGaussianAdam(FusedAdam):
    """step(visibility, N): `visibility` is the view's radii > 0 mask over the N Gaussians (train.py:173-176)."""
    bias_correction = False

    def __init__(self, params, lr, eps):
        super().__init__(params, lr=lr, eps=eps)

    @torch.no_grad()
    def step(self, visibility, N):
        vis = visibility.to(torch.uint8).contiguous()
        if vis.numel() != N:
            raise RuntimeError("SparseGaussianAdam: visibility must have one entry per Gaussian")
        by_hyper = {}
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is not None and p.shape[0] != N:
                    raise RuntimeError("SparseGaussianAdam: every parameter must have N rows")
            key = (group["betas"][0], group["betas"][1], group["eps"])
            by_hyper.setdefault(key, []).extend(self._entries(group))
        for (b1, b2, eps), entries in by_hyper.items():
            adam_step_raw(entries, b1, b2, eps, visible=vis)
