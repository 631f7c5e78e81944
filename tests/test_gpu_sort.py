"""GPU tests (-m gpu) of the library's radix sort (csrc/gsr_sort.cu) through its test hook gsr_debug_sort_pairs: what
replaces cub::DeviceRadixSort::SortPairs (rasterizer_impl.cu:303, simple_knn.cu:211). The checker is numpy's STABLE sort:
the binning's final order (tile, depth, Gaussian id) relies on stability. Sizes straddle the one-CTA path (<= 12288 pairs)
and the onesweep passes; keys with few distinct values exercise the ranking of equal digits; drop mode is the depth sort's
(culled Gaussians carry the key 0xFFFFFFFF)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _lib():
    from diff_gaussian_rasterization import _C
    L = _C.lib()
    L.gsr_debug_sort_scratch_bytes.restype = C.c_size_t
    L.gsr_debug_sort_scratch_bytes.argtypes = [C.c_int]
    L.gsr_debug_sort_pairs.restype = C.c_int
    L.gsr_debug_sort_pairs.argtypes = [C.c_void_p] * 4 + [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    return L


def _sort(L, keys_np, bits, drop):
    n = len(keys_np)
    k = torch.from_numpy(keys_np.view(np.int32).copy()).to(DEV)
    v = torch.arange(n, dtype=torch.int32, device=DEV)
    kt, vt = torch.empty_like(k), torch.empty_like(v)
    scratch = torch.zeros(int(L.gsr_debug_sort_scratch_bytes(n)) // 4 + 16, dtype=torch.int32, device=DEV)
    nv = torch.full((1,), -1, dtype=torch.int32, device=DEV)
    rc = L.gsr_debug_sort_pairs(k.data_ptr(), v.data_ptr(), kt.data_ptr(), vt.data_ptr(), n, bits, int(drop), scratch.data_ptr(),
                                nv.data_ptr() if drop else None, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L.gsr_error_string(rc).decode()
    torch.cuda.synchronize()
    return k.cpu().numpy().view(np.uint32), v.cpu().numpy(), int(nv.item())


@pytest.mark.parametrize("n", [1, 31, 33, 1000, 4097, 12287, 12288, 12289, 50_000, 300_000])
@pytest.mark.parametrize("kind", ["random32", "few_values", "sorted_desc"])
def test_sort_pairs_is_numpys_stable_sort(n, kind):
    L = _lib()
    rng = np.random.default_rng(n * 7 + len(kind))
    if kind == "random32":
        keys, bits = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32), 32
    elif kind == "few_values":                          # 13-bit tile ids, long runs of equal keys: stability matters
        keys, bits = rng.integers(0, 37, n, dtype=np.uint64).astype(np.uint32) * 211 % 8160, 13
    else:
        keys, bits = np.arange(n, 0, -1, dtype=np.uint32) * 3 % (1 << 30), 30
    ks, vs, _ = _sort(L, keys, bits, drop=False)
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(vs, order.astype(np.int32)), f"n={n} {kind}: permutation differs from the stable sort"
    assert np.array_equal(ks, keys[order])


@pytest.mark.parametrize("n", [5, 4096, 10_000, 12288, 12289, 200_000])
def test_drop_mode_compacts_the_culled_keys_away(n):
    L = _lib()
    rng = np.random.default_rng(n)
    depth = (0.2 + rng.random(n) * 50.0).astype(np.float32)
    keys = depth.view(np.uint32).copy()
    culled = rng.random(n) < 0.45
    keys[culled] = 0xFFFFFFFF
    ks, vs, n_valid = _sort(L, keys, 32, drop=True)
    kept = np.flatnonzero(~culled)
    assert n_valid == len(kept)
    order = kept[np.argsort(keys[kept], kind="stable")]
    assert np.array_equal(vs[:n_valid], order.astype(np.int32))
    assert np.array_equal(ks[:n_valid], keys[order])
