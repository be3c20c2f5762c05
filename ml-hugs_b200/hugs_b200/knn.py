"""distCUDA2: mean squared distance of every point to its 3 nearest neighbours.

Same name, argument and result as ``simple_knn._C.distCUDA2`` (submodules/simple-knn/ext.cpp,
spatial.cu), which hugs/models/scene.py:181 calls to initialise the scene Gaussians' scales:
``dist2 = torch.clamp_min(distCUDA2(points.float().cuda()), 1e-7)``.
"""
import ctypes as C

import torch

from . import _native as N


def distCUDA2(points):
    if not points.is_cuda:
        raise RuntimeError("hugs_b200 has no CPU path: distCUDA2 needs a CUDA tensor")
    if points.dim() != 2 or points.shape[1] != 3:
        raise RuntimeError("distCUDA2: points must be [P,3]")
    device = points.device
    pts = points.detach().float().contiguous()
    P = int(pts.shape[0])
    means = torch.zeros(P, dtype=torch.float32, device=device)   # spatial.cu: torch::full({P}, 0.0)
    if P == 0:
        return means
    nbytes = int(N.knn_temp_bytes(P))
    temp = torch.empty(nbytes, dtype=torch.uint8, device=device)
    with torch.cuda.device(device):
        rc = N.knn_mean_dist2(P, pts.data_ptr(), means.data_ptr(), temp.data_ptr(), nbytes,
                              C.c_void_p(torch.cuda.current_stream(device).cuda_stream))
    N.check(rc, "distCUDA2")
    return means
