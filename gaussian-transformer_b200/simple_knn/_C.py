"""`simple_knn._C`: the one function the reference's pybind module exports (submodules/simple-knn/ext.cpp:15-17)."""
import torch

from diff_gaussian_rasterization import _lib

import ctypes


def distCUDA2(points):
    """(P,) float32: mean of the squared distances to the 3 nearest neighbours of every point
    (submodules/simple-knn/spatial.cu:15-26 -> SimpleKNN::knn).  `points`: (P,3) float32 CUDA tensor."""
    if not points.is_cuda:
        raise RuntimeError("distCUDA2: points must be a CUDA tensor (there is no CPU path)")
    if points.dim() != 2 or points.size(1) != 3:
        raise RuntimeError("distCUDA2: points must have dimensions (num_points, 3)")
    dev = points.device
    P = points.size(0)
    pts = points.detach().to(torch.float32).contiguous()
    out = torch.zeros(P, dtype=torch.float32, device=dev)  # the reference returns torch::full({P}, 0.0)
    if P == 0:
        return out
    with torch.cuda.device(dev):
        nbytes = ctypes.c_size_t()
        _lib.check(_lib.lib.gsr_knn_scratch_bytes(P, ctypes.byref(nbytes)))
        scratch = torch.empty(nbytes.value, dtype=torch.uint8, device=dev)
        _lib.check(_lib.lib.gsr_knn_dist2(P, pts.data_ptr(), out.data_ptr(), scratch.data_ptr(),
                                          torch.cuda.current_stream(dev).cuda_stream))
    return out
