"""GPU test (-m gpu) of gsr_dist2_noalloc (include/gsr.h): the distCUDA2 chain with its scratch in a fixed device array
inside the library, for callers that cannot allocate device memory at the call. It has no counterpart in the reference;
the checker is the existing entry point gsr_dist2 (itself pinned to the
reference CUDA build and the C oracle in test_gpu_parity.py), called through the same C ABI.

The file name sorts last on purpose (a failure here hides nothing)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
MAX_POINTS = 131072                      # GSR_DIST2_NOALLOC_MAX_POINTS


def _dist2_both(L, pts):
    P = int(pts.size(0))
    stream = torch.cuda.current_stream(pts.device).cuda_stream
    want = torch.full((P,), -1.0, dtype=torch.float32, device=pts.device)
    temp = torch.empty(int(L.gsr_dist2_temp_bytes(P)), dtype=torch.uint8, device=pts.device)
    assert L.gsr_dist2(P, pts.data_ptr(), want.data_ptr(), temp.data_ptr(), stream) == 0
    got = torch.full((P,), -2.0, dtype=torch.float32, device=pts.device)       # allocated BEFORE the call under test
    torch.cuda.synchronize()
    allocs = torch.cuda.memory_stats(pts.device).get("num_device_alloc", 0)
    rc = L.gsr_dist2_noalloc(P, pts.data_ptr(), got.data_ptr(), stream)
    assert rc == 0, L.gsr_error_string(rc).decode()
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats(pts.device).get("num_device_alloc", 0) == allocs
    return got.cpu().numpy(), want.cpu().numpy()


def test_dist2_noalloc_matches_gsr_dist2_on_two_calls_in_a_row():
    from diff_gaussian_rasterization import _C
    L = _C.lib()
    g = torch.Generator().manual_seed(23)
    # two calls in a row on the shared scratch, the second SMALLER than the first (stale contents behind it) and then
    # the largest size the entry point takes
    for P, spread in ((50_000, 2.0), (1_234, 0.5), (MAX_POINTS, 3.0), (5, 1.0)):
        pts = (torch.randn(P, 3, generator=g) * spread + torch.tensor([1.0, -2.0, 0.5])).to(DEV).contiguous()
        if P > 100:
            pts[17] = pts[5]                                       # a duplicate: distance 0 counts
        for call in range(2):
            got, want = _dist2_both(L, pts)
            # same kernels on the same input: the tolerance is test_dist2_vs_oracle_and_reference's, the result is in
            # fact identical
            assert np.allclose(got, want, rtol=2e-6, atol=1e-12), f"P={P} call {call}: max rel {np.abs(got / want - 1).max()}"
            assert np.array_equal(got, want), f"P={P} call {call}"


def test_dist2_noalloc_refuses_sizes_beyond_its_scratch_and_bad_arguments():
    from diff_gaussian_rasterization import _C
    L = _C.lib()
    P = MAX_POINTS + 1
    pts = torch.randn(P, 3, device=DEV)
    out = torch.full((P,), 7.0, device=DEV)
    rc = L.gsr_dist2_noalloc(P, pts.data_ptr(), out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == -4                                                # GSR_E_STATIC_LIMIT
    msg = L.gsr_error_string(rc).decode()
    assert "131072" in msg and "gsr_dist2_noalloc" in msg
    torch.cuda.synchronize()
    assert bool((out == 7.0).all()), "a refused call must not launch anything"
    assert L.gsr_dist2_noalloc(10, None, out.data_ptr(), None) == -1          # GSR_E_BADARG
    assert L.gsr_dist2_noalloc(0, None, None, None) == 0                      # empty cloud: nothing to do
    # and the entry point still works afterwards
    small = pts[:4000].contiguous()
    got, want = _dist2_both(L, small)
    assert np.array_equal(got, want)
