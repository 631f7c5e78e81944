"""CPU tests for the Adam row: the numpy restatement is pinned against torch.optim.Adam itself (the
reference's optimiser, scene/gaussian_model.py:168) running on the CPU; the Python surface mirrors what the reference's
model code touches."""
import numpy as np
import pytest
import torch

from oracle import adam_oracle


def test_oracle_matches_torch_adam_over_several_steps_with_reference_hyperparameters():
    g = torch.Generator().manual_seed(0)
    shapes = {"xyz": (50, 3), "f_dc": (50, 1, 3), "f_rest": (50, 15, 3), "opacity": (50, 1), "scaling": (50, 3), "rotation": (50, 4)}
    lrs = {"xyz": 0.00016, "f_dc": 0.0025, "f_rest": 0.0025 / 20.0, "opacity": 0.025, "scaling": 0.005, "rotation": 0.001}
    params = {k: torch.nn.Parameter(torch.randn(s, generator=g)) for k, s in shapes.items()}
    opt = torch.optim.Adam([{"params": [params[k]], "lr": lrs[k], "name": k} for k in shapes], lr=0.0, eps=1e-15)
    mine = {k: (params[k].detach().numpy().copy(), np.zeros(s, np.float32), np.zeros(s, np.float32)) for k, s in shapes.items()}
    for step in range(1, 6):
        for k in shapes:
            params[k].grad = torch.randn(shapes[k], generator=g) * (10.0 ** (-step))     # wide dynamic range incl. tiny grads
        opt.step()
        for k in shapes:
            p, m, v = mine[k]
            mine[k] = adam_oracle.adam_step(p, params[k].grad.numpy(), m, v, step, lrs[k], eps=1e-15)
            st = opt.state[params[k]]
            # fp32 on both sides, one rounding apart per operation (torch fuses lerp / addcmul differently)
            for ours, ref in ((mine[k][1], st["exp_avg"].numpy()), (mine[k][2], st["exp_avg_sq"].numpy()),
                              (mine[k][0], params[k].detach().numpy())):
                assert np.allclose(ours, ref, rtol=1e-5, atol=1e-6 * np.abs(ref).max()), (k, step)


def test_sparse_oracle_leaves_invisible_rows_untouched():
    rng = np.random.default_rng(1)
    p, g = rng.standard_normal((20, 4)).astype(np.float32), rng.standard_normal((20, 4)).astype(np.float32)
    m, v = rng.standard_normal((20, 4)).astype(np.float32) * 0.1, rng.random((20, 4)).astype(np.float32) * 0.01
    vis = rng.random(20) < 0.5
    p2, m2, v2 = adam_oracle.sparse_adam_step(p, g, m, v, vis, 0.01)
    assert np.array_equal(p2[~vis], p[~vis]) and np.array_equal(m2[~vis], m[~vis]) and np.array_equal(v2[~vis], v[~vis])
    assert not np.array_equal(p2[vis], p[vis])
    # no bias correction: equals the dense formula with step_size = lr and denominator sqrt(v) + eps
    mm = m[vis] + (g[vis] - m[vis]) * np.float32(0.1)
    vv = v[vis] * np.float32(0.999) + np.float32(0.001) * g[vis] * g[vis]
    assert np.allclose(p2[vis], p[vis] - np.float32(0.01) * mm / (np.sqrt(vv) + np.float32(1e-15)), rtol=1e-6)


def test_optimizer_surface_matches_what_the_reference_model_code_touches():
    import gsr_optim
    p = torch.nn.Parameter(torch.zeros(4, 3))
    opt = gsr_optim.FusedAdam([{"params": [p], "lr": 0.1, "name": "xyz"}], lr=0.0, eps=1e-15)
    assert opt.param_groups[0]["name"] == "xyz" and opt.param_groups[0]["lr"] == 0.1 and opt.param_groups[0]["eps"] == 1e-15
    opt.param_groups[0]["lr"] = 0.05                                    # update_learning_rate (gaussian_model.py:184-191)
    assert opt.state.get(p, None) is None                               # lazily created, like torch.optim.Adam
    p.grad = torch.ones_like(p)
    with pytest.raises(RuntimeError, match="CUDA"):                      # no CPU fallback
        opt.step()
    sp = gsr_optim.SparseGaussianAdam([{"params": [p], "lr": 0.1, "name": "xyz"}], lr=0.0, eps=1e-15)
    assert sp.bias_correction is False
    # state_dict interchange with torch.optim.Adam (checkpoint capture/restore, gaussian_model.py:67-99)
    ref = torch.optim.Adam([{"params": [p], "lr": 0.1, "name": "xyz"}], lr=0.0, eps=1e-15)
    p.grad = torch.ones_like(p)
    ref.step()
    opt.load_state_dict(ref.state_dict())
    assert set(opt.state[p].keys()) >= {"step", "exp_avg", "exp_avg_sq"}
