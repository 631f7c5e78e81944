"""GPU parity tests (-m gpu) for the Adam row: gsr_adam_step through gsr_optim against
  (1) the reference's optimiser itself, torch.optim.Adam(l, lr=0.0, eps=1e-15) (scene/gaussian_model.py:168), on CUDA,
  (2) the numpy restatement oracle/adam_oracle.py,
within fp32 rounding (rtol 1e-5, atol 1e-6 of the largest entry); the sparse variant against its oracle."""
import numpy as np
import pytest
import torch

from oracle import adam_oracle

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
SHAPES = {"xyz": (3, 0.00016), "f_dc": (3, 0.0025), "f_rest": (45, 0.0025 / 20.0), "opacity": (1, 0.025), "scaling": (3, 0.005),
          "rotation": (4, 0.001)}


def _model(P, seed):
    g = torch.Generator().manual_seed(seed)
    shape = {"xyz": (P, 3), "f_dc": (P, 1, 3), "f_rest": (P, 15, 3), "opacity": (P, 1), "scaling": (P, 3), "rotation": (P, 4)}
    return {k: torch.randn(shape[k], generator=g) for k in SHAPES}


def _close(a, b, what):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    assert np.allclose(a, b, rtol=1e-5, atol=1e-6 * max(np.abs(b).max(), 1e-30)), f"{what}: max-abs {np.abs(a - b).max()}"


@pytest.mark.parametrize("P", [1000, 4097])
def test_fused_adam_matches_torch_adam_and_oracle(P):
    import gsr_optim
    init = _model(P, 0)
    ours = {k: torch.nn.Parameter(v.clone().to(DEV)) for k, v in init.items()}
    ref = {k: torch.nn.Parameter(v.clone().to(DEV)) for k, v in init.items()}
    groups = lambda d: [{"params": [d[k]], "lr": SHAPES[k][1], "name": k} for k in SHAPES]
    o1 = gsr_optim.FusedAdam(groups(ours), lr=0.0, eps=1e-15)
    o2 = torch.optim.Adam(groups(ref), lr=0.0, eps=1e-15)
    orc = {k: (v.numpy().copy(), np.zeros(v.shape, np.float32), np.zeros(v.shape, np.float32)) for k, v in init.items()}
    g = torch.Generator().manual_seed(1)
    for step in range(1, 5):
        for k in SHAPES:
            gr = torch.randn(init[k].shape, generator=g) * (10.0 ** (1 - 2 * step))      # 1e-1 .. 1e-7
            if k == "xyz" and step == 2:
                gr[::3] = 0.0                                                             # exact zeros: m decays, v decays
            ours[k].grad, ref[k].grad = gr.to(DEV), gr.to(DEV)
            p, m, v = orc[k]
            lr = 0.0001 if (k == "xyz" and step >= 3) else SHAPES[k][1]                   # update_learning_rate at step 3
            orc[k] = adam_oracle.adam_step(p, gr.numpy(), m, v, step, lr, eps=1e-15)
        if step == 3:
            o1.param_groups[0]["lr"] = o2.param_groups[0]["lr"] = 0.0001
        n0 = gsr_optim._gsr.launch_count()
        o1.step()
        assert gsr_optim._gsr.launch_count() - n0 == 1, "all six groups must go in ONE launch"
        o2.step()
        for k in SHAPES:
            _close(ours[k].detach().cpu().numpy(), ref[k].detach().cpu().numpy(), f"{k} param step {step} vs torch")
            _close(o1.state[ours[k]]["exp_avg"].cpu().numpy(), o2.state[ref[k]]["exp_avg"].cpu().numpy(), f"{k} m vs torch")
            _close(o1.state[ours[k]]["exp_avg_sq"].cpu().numpy(), o2.state[ref[k]]["exp_avg_sq"].cpu().numpy(), f"{k} v vs torch")
            _close(ours[k].detach().cpu().numpy(), orc[k][0], f"{k} param step {step} vs oracle")


def test_sparse_gaussian_adam_updates_only_visible_rows():
    import gsr_optim
    P = 3001
    init = _model(P, 3)
    ours = {k: torch.nn.Parameter(v.clone().to(DEV)) for k, v in init.items()}
    opt = gsr_optim.SparseGaussianAdam([{"params": [ours[k]], "lr": SHAPES[k][1], "name": k} for k in SHAPES], lr=0.0, eps=1e-15)
    orc = {k: (v.numpy().copy(), np.zeros(v.shape, np.float32), np.zeros(v.shape, np.float32)) for k, v in init.items()}
    g = torch.Generator().manual_seed(4)
    for step in range(3):
        vis = torch.rand(P, generator=g) < 0.4
        for k in SHAPES:
            gr = torch.randn(init[k].shape, generator=g) * 0.01
            ours[k].grad = gr.to(DEV)
            p, m, v = orc[k]
            orc[k] = adam_oracle.sparse_adam_step(p, gr.numpy(), m, v, vis.numpy(), SHAPES[k][1])
        before = {k: ours[k].detach().clone() for k in SHAPES}
        opt.step(vis.to(DEV), P)
        for k in SHAPES:
            assert torch.equal(ours[k].detach()[~vis.to(DEV)], before[k][~vis.to(DEV)]), f"{k}: invisible rows moved"
            _close(ours[k].detach().cpu().numpy(), orc[k][0], f"{k} param")
            _close(opt.state[ours[k]]["exp_avg"].cpu().numpy(), orc[k][1], f"{k} m")
            _close(opt.state[ours[k]]["exp_avg_sq"].cpu().numpy(), orc[k][2], f"{k} v")


def test_full_size_properties_3m_gaussians():
    """BASELINE size (3M Gaussians, 59 floats each): zero gradients leave fresh parameters unchanged (m = v = 0 ->
    0 / (0 + eps)); a constant gradient moves every element by exactly -lr * sign(g) on step 1 (bias-corrected Adam)."""
    import gsr_optim
    P = 3_000_000
    p = torch.nn.Parameter(torch.zeros(P, 59, device=DEV))
    opt = gsr_optim.FusedAdam([{"params": [p], "lr": 0.01, "name": "all"}], lr=0.0, eps=1e-15)
    p.grad = torch.zeros_like(p)
    opt.step()
    assert p.detach().abs().max().item() == 0.0
    opt2 = gsr_optim.FusedAdam([{"params": [p], "lr": 0.01, "name": "all"}], lr=0.0, eps=1e-15)
    p.grad = torch.full_like(p, -3.0)
    p.grad[::2] = 2.0
    opt2.step()
    d = p.detach()
    assert torch.allclose(d[::2], torch.full_like(d[::2], -0.01), rtol=1e-5, atol=0)
    assert torch.allclose(d[1::2], torch.full_like(d[1::2], 0.01), rtol=1e-5, atol=0)
