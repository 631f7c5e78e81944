"""CPU restatement of the optimiser step the reference runs after the rasterizer backward -- TEST INFRASTRUCTURE ONLY.

The reference uses `torch.optim.Adam(l, lr=0.0, eps=1e-15)` (scene/gaussian_model.py:168): a third-party dependency
(PyTorch; 2.11.0 in this image), whose published algorithm (Kingma & Ba 2015, Algorithm 1, as implemented by
torch/optim/adam.py `_single_tensor_adam`: lerp for the first moment, mul/addcmul for the second, bias correction folded
into step_size and the denominator, no weight decay, no amsgrad) is restated here in numpy.
tests/test_cpu_adam.py pins it against torch.optim.Adam itself running on the CPU in this container.

`sparse_adam_step` restates SparseGaussianAdam, the visibility-masked variant the caller switches to when the rasterizer
package exports it (train.py:37-41,173-176). It is NOT in the reference tree: semantics from the upstream
implementation the reference names (no bias correction; rows of invisible Gaussians untouched). Parity unpinned.
"""
import math

import numpy as np


def adam_step(p, g, m, v, step, lr, beta1=0.9, beta2=0.999, eps=1e-8, dtype=np.float32):
    """One Adam step; `step` is the 1-based step count. Returns new (p, m, v)."""
    p, g, m, v = (np.asarray(a, dtype) for a in (p, g, m, v))
    m = m + (g - m) * dtype(1.0 - beta1)
    v = v * dtype(beta2) + dtype(1.0 - beta2) * g * g
    step_size = dtype(lr / (1.0 - beta1 ** step))
    inv_sqrt_bc2 = dtype(1.0 / math.sqrt(1.0 - beta2 ** step))
    denom = np.sqrt(v) * inv_sqrt_bc2 + dtype(eps)
    return p - step_size * (m / denom), m, v


def sparse_adam_step(p, g, m, v, visible, lr, beta1=0.9, beta2=0.999, eps=1e-15, dtype=np.float32):
    """Rows (first axis) with visible == 0 are untouched; no bias correction."""
    p, g, m, v = (np.asarray(a, dtype).copy() for a in (p, g, m, v))
    vis = np.asarray(visible).astype(bool)
    m2 = m[vis] + (g[vis] - m[vis]) * dtype(1.0 - beta1)
    v2 = v[vis] * dtype(beta2) + dtype(1.0 - beta2) * g[vis] * g[vis]
    p[vis] = p[vis] - dtype(lr) * (m2 / (np.sqrt(v2) + dtype(eps)))
    m[vis], v[vis] = m2, v2
    return p, m, v
