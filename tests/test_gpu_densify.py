"""GPU parity of the device-side densification (gsr_densify.py -> libgsr.so gsr_densify_*).

 * against tests/golden/ref_densify.npz, vectors from the reference's own GaussianModel methods
   (scene/gaussian_model.py:226-229, 307-322, 412-433);
 * against oracle/densify_oracle.py on larger seeded models;
 * at the bench size (3M Gaussians) through size-independent properties: count conservation, kept originals =
   torch's boolean gather of the same mask bit for bit, prune with an empty mask is the identity, moments of new rows zero.
Bar: row counts, row order, every copied value and the Adam moments bit-exact; computed values (children's xyz and
scaling, statistics, reset opacity) within 2e-6 of the tensor's scale.
"""
import os
import sys

import numpy as np
import pytest
import torch
from torch import nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gaussian-splatting-slam_b200"))

pytestmark = pytest.mark.gpu
TOL = 2e-6
NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")
ATTR = dict(xyz="_xyz", f_dc="_features_dc", f_rest="_features_rest", opacity="_opacity", scaling="_scaling", rotation="_rotation")


def _mods():
    import gsr_densify
    from oracle import densify_oracle
    return gsr_densify, densify_oracle


class Model:
    """The attributes of the reference's GaussianModel that densification touches (training_setup, :155-170)."""

    def __init__(self, arrays, dev="cuda", percent_dense=0.01, with_state=True):
        self.percent_dense = percent_dense
        groups = []
        for n in NAMES:
            p = nn.Parameter(torch.from_numpy(np.ascontiguousarray(arrays[n])).to(dev))
            setattr(self, ATTR[n], p)
            groups.append({"params": [p], "lr": 0.0, "name": n})
        self.optimizer = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
        if with_state:
            for g in self.optimizer.param_groups:
                n = g["name"]
                self.optimizer.state[g["params"][0]] = {
                    "step": torch.tensor(3.0),
                    "exp_avg": torch.from_numpy(np.ascontiguousarray(arrays[n + "_m"])).to(dev),
                    "exp_avg_sq": torch.from_numpy(np.ascontiguousarray(arrays[n + "_v"])).to(dev)}
        self.xyz_gradient_accum = torch.from_numpy(arrays["accum"].copy()).to(dev)
        self.denom = torch.from_numpy(arrays["denom"].copy()).to(dev)
        self.max_radii2D = torch.from_numpy(arrays["max_radii2D"].copy()).to(dev)

    def arrays(self):
        out = {}
        for g in self.optimizer.param_groups:
            n, p = g["name"], g["params"][0]
            assert getattr(self, ATTR[n]) is p and p.requires_grad
            out[n] = p.detach().cpu().numpy()
            st = self.optimizer.state.get(p)
            if st is not None:
                out[n + "_m"], out[n + "_v"] = st["exp_avg"].cpu().numpy(), st["exp_avg_sq"].cpu().numpy()
        out["accum"], out["denom"] = self.xyz_gradient_accum.cpu().numpy(), self.denom.cpu().numpy()
        out["max_radii2D"] = self.max_radii2D.cpu().numpy()
        return out


def close(a, b, what):
    assert a.shape == b.shape, (what, a.shape, b.shape)
    if a.size:
        scale = max(1.0, float(np.abs(b).max()))
        err = float(np.abs(a.astype(np.float64) - b).max())
        assert err <= TOL * scale, (what, err)


def check(got, want, what, computed=()):
    for k in want:
        if k not in got:
            continue
        if k in computed or k in ("accum",):
            close(got[k], want[k], f"{what}:{k}")
        else:
            assert got[k].shape == want[k].shape, (what, k, got[k].shape, want[k].shape)
            assert np.array_equal(got[k], want[k]), f"{what}:{k} not bit-exact"


def gold(tag):
    G = np.load(os.path.join(ROOT, "tests", "golden", "ref_densify.npz"))
    return G, lambda t: {k[len(t) + 1:]: G[k] for k in G.files if k.startswith(t + "_")}


@pytest.mark.parametrize("tag", ["a", "b", "c", "d"])
def test_golden_reference_vectors(tag):
    gd, _ = _mods()
    G, load = gold(tag)
    m_in = load(f"{tag}_in")
    P = m_in["xyz"].shape[0]
    zero = dict(m_in, accum=np.zeros((P, 1), np.float32), denom=np.zeros((P, 1), np.float32), max_radii2D=np.zeros(P, np.float32))
    model = Model(zero)
    for v in range(3):
        vsp = nn.Parameter(torch.zeros(P, 3, device="cuda"))
        vsp.grad = torch.from_numpy(G[f"{tag}_stats{v}_grad"]).cuda()
        radii = torch.from_numpy(G[f"{tag}_stats{v}_radii"]).cuda()
        gd.add_densification_stats(model, vsp, radii > 0, radii)
    check(model.arrays(), m_in, tag + ":stats")
    max_grad, min_opacity, extent, mss, pd = G[f"{tag}_args"]
    model.percent_dense = float(pd)
    noise = torch.from_numpy(G[f"{tag}_noise"]).cuda()
    counts = gd.densify_and_prune(model, float(max_grad), float(min_opacity), float(extent), None if mss < 0 else float(mss),
                                  None, noise=noise)
    assert counts[2] * 2 == noise.shape[0]
    want = load(f"{tag}_out")
    got = model.arrays()
    check(got, want, tag + ":densify", computed=("xyz", "scaling"))
    assert np.array_equal(got["xyz"][:counts[0] + counts[1]], want["xyz"][:counts[0] + counts[1]])       # copies: exact
    assert np.array_equal(got["scaling"][:counts[0] + counts[1]], want["scaling"][:counts[0] + counts[1]])
    assert model.optimizer.state[model._xyz]["step"].item() == 3.0
    model.tmp_radii = torch.arange(got["xyz"].shape[0], dtype=torch.int32, device="cuda")
    mask = torch.from_numpy(G[f"{tag}_prune_mask"]).cuda()
    gd.prune_points(model, mask)
    # the children's xyz / scaling were computed on the device a step earlier: they carry that step's tolerance along
    check(model.arrays(), load(f"{tag}_pruned"), tag + ":prune", computed=("xyz", "scaling"))
    assert torch.equal(model.tmp_radii, torch.arange(mask.shape[0], dtype=torch.int32, device="cuda")[~mask])
    gd.reset_opacity(model)
    check(model.arrays(), load(f"{tag}_reset"), tag + ":reset", computed=("xyz", "scaling", "opacity"))


def random_model(P, seed, deg=3, extent=5.0):
    r = np.random.default_rng(seed)
    M = (deg + 1) ** 2
    f = np.float32
    big = (r.random((P, 1)) < 0.05) * 9.0 + 1.0
    a = {"xyz": ((r.random((P, 3)) * 2 - 1) * extent).astype(f), "f_dc": r.standard_normal((P, 1, 3)).astype(f),
         "f_rest": (r.standard_normal((P, M - 1, 3)) * 0.1).astype(f), "opacity": (r.standard_normal((P, 1)) * 3 - 1).astype(f),
         "scaling": np.log((r.random((P, 3)) * 0.013 * extent + 1e-3) * big).astype(f),
         "rotation": r.standard_normal((P, 4)).astype(f)}
    for n in NAMES:
        a[n + "_m"] = (r.standard_normal(a[n].shape) * 1e-3).astype(f)
        a[n + "_v"] = (r.random(a[n].shape) * 1e-6).astype(f)
    denom = r.integers(0, 4, (P, 1)).astype(f)
    a["accum"] = (r.random((P, 1)) * 0.0006 * denom).astype(f)          # denom 0 -> accum 0 -> 0/0 = NaN -> 0
    a["denom"] = denom
    a["max_radii2D"] = (r.random(P) * 40).astype(f)
    return a


@pytest.mark.parametrize("P,deg,mss,seed", [(100_003, 3, 20, 0), (40_000, 1, None, 1), (5_000, 0, 20, 2), (1, 3, 20, 3), (255, 2, None, 4),
                                            (257, 2, 20, 5)])
def test_against_oracle(P, deg, mss, seed):
    gd, D = _mods()
    a = random_model(P, seed, deg)
    n_sel = D.count_split_selected(a, 0.0002, 5.0, 0.01)
    noise = np.random.default_rng(seed + 50).standard_normal((2 * n_sel, 3)).astype(np.float32)
    want, _ = D.densify_and_prune(a, 0.0002, 0.005, 5.0, mss, 0.01, noise)
    model = Model(a)
    counts = gd.densify_and_prune(model, 0.0002, 0.005, 5.0, mss, None, noise=torch.from_numpy(noise).cuda())
    assert counts[2] == n_sel
    got = model.arrays()
    check(got, want, f"P={P}", computed=("xyz", "scaling"))
    assert got["xyz"].shape[0] == counts[0] + counts[1] + 2 * counts[3]
    # prune_points against the oracle on the result
    mask = np.random.default_rng(seed + 60).random(got["xyz"].shape[0]) < 0.4
    gd.prune_points(model, torch.from_numpy(mask).cuda())
    check(model.arrays(), D.prune_points(got, mask), f"P={P}:prune")


def test_edges():
    gd, D = _mods()
    # empty model
    a = random_model(0, 0)
    model = Model(a)
    assert gd.densify_and_prune(model, 0.0002, 0.005, 5.0, 20) == [0, 0, 0, 0]
    assert model._xyz.shape == (0, 3) and model._features_rest.shape == (0, 15, 3) and model.denom.shape == (0, 1)
    gd.prune_points(model, torch.zeros(0, dtype=torch.bool, device="cuda"))
    # no optimiser state yet (stored_state is None, :297-303 / :337-339)
    a = random_model(3000, 7)
    model = Model(a, with_state=False)
    n_sel = D.count_split_selected(a, 0.0002, 5.0, 0.01)
    noise = np.random.default_rng(1).standard_normal((2 * n_sel, 3)).astype(np.float32)
    gd.densify_and_prune(model, 0.0002, 0.005, 5.0, 20, noise=torch.from_numpy(noise).cuda())
    want, _ = D.densify_and_prune({k: v for k, v in a.items() if not k.endswith(("_m", "_v"))}, 0.0002, 0.005, 5.0, 20, 0.01, noise)
    check(model.arrays(), want, "nostate", computed=("xyz", "scaling"))
    assert len(model.optimizer.state) == 0
    # everything pruned
    model = Model(a)
    gd.prune_points(model, torch.ones(3000, dtype=torch.bool, device="cuda"))
    assert model._xyz.shape == (0, 3) and model.optimizer.state[model._rotation]["exp_avg"].shape == (0, 4)
    # noise drawn here: same generator consumption as torch.normal(mean 0, std) of the reference (randn of that shape)
    model = Model(a)
    torch.manual_seed(11)
    counts = gd.densify_and_prune(model, 0.0002, 0.005, 5.0, 20)
    torch.manual_seed(11)
    z = torch.empty((2 * counts[2], 3), device="cuda").normal_()
    model2 = Model(a)
    gd.densify_and_prune(model2, 0.0002, 0.005, 5.0, 20, noise=z)
    assert torch.equal(model._xyz, model2._xyz)
    # loud failures
    with pytest.raises(ValueError):
        gd.densify_and_prune(Model(a), 0.0, 0.005, 5.0, 20)
    with pytest.raises(RuntimeError):
        gd.densify_and_prune(Model(a, dev="cpu"), 0.0002, 0.005, 5.0, 20)
    with pytest.raises(RuntimeError):
        gd.densify_and_prune(Model(a), 0.0002, 0.005, 5.0, 20, noise=torch.zeros(5, 3, device="cuda"))


def test_bench_size_properties():
    gd, _ = _mods()
    P, dev = 3_000_000, "cuda"
    g = torch.Generator(device=dev).manual_seed(0)
    arrays = {"xyz": torch.rand(P, 3, generator=g, device=dev) * 10 - 5, "f_dc": torch.randn(P, 1, 3, generator=g, device=dev),
              "f_rest": torch.randn(P, 15, 3, generator=g, device=dev), "opacity": torch.randn(P, 1, generator=g, device=dev) * 3 - 1,
              "scaling": torch.log(torch.rand(P, 3, generator=g, device=dev) * 0.065 + 1e-3),
              "rotation": torch.randn(P, 4, generator=g, device=dev)}
    moments = {n: (torch.randn(t.shape, generator=g, device=dev), torch.rand(t.shape, generator=g, device=dev)) for n, t in arrays.items()}
    denom = torch.randint(0, 4, (P, 1), generator=g, device=dev).float()
    accum = torch.rand(P, 1, generator=g, device=dev) * 0.0006 * denom
    rule = gd.make_rule(0.0002, 0.005, 5.0, 20, 0.01)
    new_p, new_m, c = gd.densify_raw(arrays, moments, accum, denom, rule)
    P_new = c[0] + c[1] + 2 * c[3]
    assert all(t.shape[0] == P_new for t in new_p.values()) and c[3] <= c[2] and c[0] > 0 and c[1] > 0 and c[3] > 0
    # the same decisions restated with torch ops on the device (checker only)
    grads = accum / denom
    grads[grads.isnan()] = 0.0
    mx = torch.exp(arrays["scaling"]).max(dim=1).values
    split = (grads[:, 0] >= 0.0002) & (mx > 0.05)
    clone = (grads[:, 0] >= 0.0002) & (mx <= 0.05)
    pruned = (torch.sigmoid(arrays["opacity"])[:, 0] < 0.005) | (mx > 0.5)
    keep, kept_clone = ~split & ~pruned, clone & ~pruned
    assert int(keep.sum()) == c[0] and int(kept_clone.sum()) == c[1] and int(split.sum()) == c[2]
    for n in NAMES:
        assert torch.equal(new_p[n][:c[0]], arrays[n][keep]), n
        assert torch.equal(new_m[n][0][:c[0]], moments[n][0][keep]) and torch.equal(new_m[n][1][:c[0]], moments[n][1][keep]), n
        assert torch.equal(new_p[n][c[0]:c[0] + c[1]], arrays[n][kept_clone]), n
        assert not new_m[n][0][c[0]:].any() and not new_m[n][1][c[0]:].any(), n
        if n not in ("xyz", "scaling"):       # both repetitions of the children carry the parent's row
            kids = new_p[n][c[0] + c[1]:]
            assert torch.equal(kids[:c[3]], kids[c[3]:]), n
    # prune with an empty mask is the identity; with a mask it is torch's boolean gather
    out, n = gd.prune_raw([(arrays["f_rest"], *moments["f_rest"])], torch.zeros(P, dtype=torch.bool, device=dev))
    assert n == P and torch.equal(out[0][0], arrays["f_rest"]) and torch.equal(out[0][1], moments["f_rest"][0])
    assert torch.equal(out[0][2], moments["f_rest"][1])
    with pytest.raises(RuntimeError):           # exp_avg without exp_avg_sq: rejected by the C ABI
        gd.prune_raw([(arrays["f_rest"], moments["f_rest"][0], None)], torch.zeros(P, dtype=torch.bool, device=dev))
    mask = torch.rand(P, generator=g, device=dev) < 0.37
    out, n = gd.prune_raw([(arrays["rotation"], None, None), (arrays["f_rest"], None, None)], mask)
    assert n == int((~mask).sum()) and torch.equal(out[0][0], arrays["rotation"][~mask]) and torch.equal(out[1][0], arrays["f_rest"][~mask])
