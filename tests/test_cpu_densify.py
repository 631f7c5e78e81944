"""Pins oracle/densify_oracle.py against tests/golden/ref_densify.npz -- vectors produced by the reference's own
GaussianModel methods (scene/gaussian_model.py:226-229, 307-322, 412-433; generator: tests/golden/make_golden_densify.py).

Bar: row counts, row order and every copied value bit-exact; the values the reference COMPUTES (split children's xyz
and scaling, the statistics, reset opacities) within 2e-6 relative to the tensor's scale (fp32 libm / summation order).
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import densify_oracle as D  # noqa: E402

GOLD = np.load(os.path.join(ROOT, "tests", "golden", "ref_densify.npz"))
CASES = ("a", "b", "c", "d")
TOL = 2e-6


def load(tag):
    m = {}
    for n in D.NAMES:
        for s in ("", "_m", "_v"):
            m[n + s] = GOLD[f"{tag}_{n}{s}"]
    for k in ("accum", "denom", "max_radii2D"):
        m[k] = GOLD[f"{tag}_{k}"]
    return m


def close(a, b, what):
    assert a.shape == b.shape, (what, a.shape, b.shape)
    if a.size:
        scale = max(1.0, float(np.abs(b).max()))
        assert float(np.abs(a.astype(np.float64) - b).max()) <= TOL * scale, what


def check_model(got, want, what, computed=()):
    for n in D.NAMES:
        for s in ("", "_m", "_v"):
            k = n + s
            if k in computed:
                close(got[k], want[k], f"{what}:{k}")
            else:
                assert got[k].shape == want[k].shape and np.array_equal(got[k], want[k]), f"{what}:{k}"
    for k in ("accum", "denom", "max_radii2D"):
        close(got[k], want[k], f"{what}:{k}")


@pytest.mark.parametrize("tag", CASES)
def test_stats(tag):
    want = load(f"{tag}_in")
    P = want["xyz"].shape[0]
    m = dict(want, accum=np.zeros((P, 1), np.float32), denom=np.zeros((P, 1), np.float32), max_radii2D=np.zeros(P, np.float32))
    for v in range(3):
        radii = GOLD[f"{tag}_stats{v}_radii"]
        m = D.add_densification_stats(m, GOLD[f"{tag}_stats{v}_grad"], radii > 0, radii)
    close(m["accum"], want["accum"], "accum")
    assert np.array_equal(m["denom"], want["denom"]) and np.array_equal(m["max_radii2D"], want["max_radii2D"])


@pytest.mark.parametrize("tag", CASES)
def test_densify_and_prune(tag):
    max_grad, min_opacity, extent, mss, pd = GOLD[f"{tag}_args"]
    mss = None if mss < 0 else mss
    m_in = load(f"{tag}_in")
    noise = GOLD[f"{tag}_noise"]
    assert D.count_split_selected(m_in, max_grad, extent, pd) * 2 == noise.shape[0]
    got, n_sel = D.densify_and_prune(m_in, max_grad, min_opacity, extent, mss, pd, noise)
    assert n_sel * 2 == noise.shape[0]
    want = load(f"{tag}_out")
    # split children are the only computed rows; everything before them must match to the bit
    check_model(got, want, tag, computed=("xyz", "scaling"))
    n_old = int((np.abs(want["xyz_m"]).sum(axis=1) + np.abs(want["xyz_v"]).sum(axis=1) > 0).sum())
    assert np.array_equal(got["xyz"][:n_old], want["xyz"][:n_old]) and np.array_equal(got["scaling"][:n_old], want["scaling"][:n_old])
    assert not want["accum"].any() and not want["denom"].any() and not want["max_radii2D"].any()


@pytest.mark.parametrize("tag", CASES)
def test_prune_points_and_reset(tag):
    m = load(f"{tag}_out")
    got = D.prune_points(m, GOLD[f"{tag}_prune_mask"])
    check_model(got, load(f"{tag}_pruned"), tag + ":pruned")
    got = D.reset_opacity(got)
    check_model(got, load(f"{tag}_reset"), tag + ":reset", computed=("opacity",))
    assert not got["opacity_m"].any() and not got["opacity_v"].any()
