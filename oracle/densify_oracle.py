"""CPU restatement of the reference's map densification / pruning -- TEST INFRASTRUCTURE ONLY (imported by tests/ and
bench tooling as the checker; the product path, gsr_densify.py -> libgsr.so, never touches it).

Follows scene/gaussian_model.py of Runnlion/gaussian-splatting-slam step by step, in numpy float32, with the model held
as a dict of arrays {"xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation"} plus the Adam moments under
"<name>_m" / "<name>_v" and the statistics "accum", "denom", "max_radii2D":
  add_densification_stats   :431-433 (+ the caller's max_radii2D update, train.py:159)
  prune_points              :307-322 with _prune_optimizer :287-305
  densification_postfix     :346-365 with cat_tensors_to_optimizer :324-344
  densify_and_clone         :393-408
  densify_and_split         :368-391 (build_rotation: utils/general_utils.py:78-99)
  densify_and_prune         :412-429
  reset_opacity             :226-229 with replace_tensor_to_optimizer :274-285
It deliberately keeps the reference's three-stage structure (clone, then split, then prune, each a mask-gather and a
concatenation) so that it checks the device path's single folded pass independently.
Pinned by tests/test_cpu_densify.py against tests/golden/ref_densify.npz, which was produced by running the reference's
own GaussianModel methods (tests/golden/make_golden_densify.py).

Two places follow the reference's CUDA habitat rather than the CPU run that produced the golden file (differences are
1 ulp, inside the stated tolerances): a CUDA tensor divided by a Python scalar is multiplied by the float reciprocal
(:384), and torch.norm's reduction fuses v*v into the running sum.
"""
import numpy as np

NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")
f32 = np.float32


def _sigmoid(x):
    return (f32(1) / (f32(1) + np.exp(-x.astype(f32)))).astype(f32)


def add_densification_stats(model, viewspace_grad, update_filter, radii=None):
    """:431-433; with `radii`, also max_radii2D[filter] = max(max_radii2D[filter], radii[filter]) (train.py:159)."""
    m = dict(model)
    f = np.asarray(update_filter, bool)
    g = np.asarray(viewspace_grad, f32)
    nrm = np.sqrt(g[:, 0] * g[:, 0] + g[:, 1] * g[:, 1]).astype(f32)
    accum, denom = m["accum"].copy(), m["denom"].copy()
    accum[f, 0] += nrm[f]
    denom[f, 0] += f32(1)
    m["accum"], m["denom"] = accum, denom
    if radii is not None:
        mr = m["max_radii2D"].copy()
        mr[f] = np.maximum(mr[f], np.asarray(radii)[f].astype(f32))
        m["max_radii2D"] = mr
    return m


def prune_points(model, mask):
    """:307-322. mask: True = remove."""
    keep = ~np.asarray(mask, bool)
    m = {}
    for n in NAMES:
        for suffix in ("", "_m", "_v"):
            if n + suffix in model:
                m[n + suffix] = model[n + suffix][keep]
    for k in ("accum", "denom", "max_radii2D"):
        m[k] = model[k][keep]
    return m


def _postfix(model, new):
    """densification_postfix :346-365: append rows, zero moments for them, reset ALL statistics to zeros."""
    m = {}
    for n in NAMES:
        m[n] = np.concatenate([model[n], new[n]], axis=0)
        for suffix in ("_m", "_v"):
            if n + suffix in model:
                m[n + suffix] = np.concatenate([model[n + suffix], np.zeros_like(new[n])], axis=0)
    P = m["xyz"].shape[0]
    m["accum"], m["denom"], m["max_radii2D"] = np.zeros((P, 1), f32), np.zeros((P, 1), f32), np.zeros(P, f32)
    return m


def build_rotation(r):
    """utils/general_utils.py:78-99, each product / sum rounded on its own as separate torch ops do."""
    r = np.asarray(r, f32)
    norm = np.sqrt(r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1] + r[:, 2] * r[:, 2] + r[:, 3] * r[:, 3])
    q = r / norm[:, None]
    R = np.zeros((q.shape[0], 3, 3), f32)
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    two, one = f32(2), f32(1)
    R[:, 0, 0] = one - two * (y * y + z * z)
    R[:, 0, 1] = two * (x * y - w * z)
    R[:, 0, 2] = two * (x * z + w * y)
    R[:, 1, 0] = two * (x * y + w * z)
    R[:, 1, 1] = one - two * (x * x + z * z)
    R[:, 1, 2] = two * (y * z - w * x)
    R[:, 2, 0] = two * (x * z - w * y)
    R[:, 2, 1] = two * (y * z + w * x)
    R[:, 2, 2] = one - two * (x * x + y * y)
    return R


def densify_and_clone(model, grads, grad_threshold, scene_extent, percent_dense):
    """:393-408."""
    sel = np.sqrt(grads[:, 0] * grads[:, 0]) >= f32(grad_threshold)
    sel &= np.exp(model["scaling"]).max(axis=1) <= f32(percent_dense * scene_extent)
    return _postfix(model, {n: model[n][sel] for n in NAMES}), int(sel.sum())


def densify_and_split(model, grads, grad_threshold, scene_extent, percent_dense, noise, N=2):
    """:368-391. noise: [N * n_selected, 3] standard normal (torch.normal(mean 0, std) = noise * std + 0)."""
    n_init = model["xyz"].shape[0]
    padded = np.zeros(n_init, f32)
    padded[:grads.shape[0]] = grads[:, 0]
    sel = padded >= f32(grad_threshold)
    scale = np.exp(model["scaling"]).astype(f32)
    sel &= scale.max(axis=1) > f32(percent_dense * scene_extent)
    n_sel = int(sel.sum())
    noise = np.asarray(noise, f32).reshape(-1, 3)
    assert noise.shape[0] == N * n_sel, (noise.shape, N, n_sel)
    stds = np.tile(scale[sel], (N, 1))
    samples = (noise * stds + f32(0)).astype(f32)
    rots = np.tile(build_rotation(model["rotation"][sel]), (N, 1, 1))
    new = {"xyz": (np.einsum("nij,nj->ni", rots.astype(np.float64), samples.astype(np.float64)).astype(f32)
                   + np.tile(model["xyz"][sel], (N, 1))).astype(f32),
           "scaling": np.log((stds * (f32(1) / f32(0.8 * N))).astype(f32)).astype(f32),
           "rotation": np.tile(model["rotation"][sel], (N, 1)),
           "f_dc": np.tile(model["f_dc"][sel], (N, 1, 1)),
           "f_rest": np.tile(model["f_rest"][sel], (N, 1, 1)),
           "opacity": np.tile(model["opacity"][sel], (N, 1))}
    m = _postfix(model, new)
    prune = np.concatenate([sel, np.zeros(N * n_sel, bool)])
    return prune_points(m, prune), n_sel


def densify_and_prune(model, max_grad, min_opacity, extent, max_screen_size, percent_dense, noise, N=2):
    """:412-429. Returns (new model, number of split-selected parents)."""
    with np.errstate(divide="ignore", invalid="ignore"):
        grads = (model["accum"] / model["denom"]).astype(f32)
    grads[np.isnan(grads)] = f32(0)
    m, _ = densify_and_clone(model, grads, max_grad, extent, percent_dense)
    m, n_sel = densify_and_split(m, grads, max_grad, extent, percent_dense, noise, N)
    prune = _sigmoid(m["opacity"])[:, 0] < f32(min_opacity)
    if max_screen_size:
        big_vs = m["max_radii2D"] > f32(max_screen_size)           # zeros since _postfix: never true for a positive cut
        big_ws = np.exp(m["scaling"]).max(axis=1) > f32(0.1 * extent)
        prune = prune | big_vs | big_ws
    return prune_points(m, prune), n_sel


def count_split_selected(model, max_grad, extent, percent_dense):
    """How many noise rows / N densify_and_split will ask for (the host needs it before drawing the noise)."""
    with np.errstate(divide="ignore", invalid="ignore"):
        g = (model["accum"] / model["denom"]).astype(f32)[:, 0]
    g[np.isnan(g)] = f32(0)
    return int(((g >= f32(max_grad)) & (np.exp(model["scaling"]).max(axis=1) > f32(percent_dense * extent))).sum())


def reset_opacity(model):
    """:226-229: opacity = inverse_sigmoid(min(sigmoid(opacity), 0.01)); its moments zeroed (:274-285)."""
    m = dict(model)
    x = np.minimum(_sigmoid(model["opacity"]), f32(1) * f32(0.01)).astype(f32)
    m["opacity"] = np.log(x / (f32(1) - x)).astype(f32)
    m["opacity_m"], m["opacity_v"] = np.zeros_like(m["opacity"]), np.zeros_like(m["opacity"])
    return m
