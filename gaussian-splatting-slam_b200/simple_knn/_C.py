"""distCUDA2 with the reference's import path and signature (submodules/simple-knn/spatial.cu:15-26,
ext.cpp:15-17): distCUDA2(points: f32[P,3] cuda) -> f32[P], the mean of the squared distances to the three
nearest other points. Runs entirely in libgsr.so (gsr_dist2); no host synchronisation, no fallback."""
import torch
from diff_gaussian_rasterization import _C as _gsr


def distCUDA2(points):
    if not points.is_cuda:
        raise RuntimeError("distCUDA2: points must be a CUDA tensor (there is no CPU path)")
    P = int(points.size(0))
    dev = points.device
    means = torch.zeros((P,), dtype=torch.float32, device=dev)
    if P == 0:
        return means
    pts = points.float().contiguous()
    L = _gsr.lib()
    with torch.cuda.device(dev):
        temp = torch.empty(int(L.gsr_dist2_temp_bytes(P)), dtype=torch.uint8, device=dev)
        rc = L.gsr_dist2(P, pts.data_ptr(), means.data_ptr(), temp.data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
        if rc != 0:
            raise RuntimeError(f"gsr_dist2: {L.gsr_error_string(rc).decode()} (code {rc})")
    return means
