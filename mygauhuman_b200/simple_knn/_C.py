"""Drop-in for ``simple_knn._C`` (/root/reference/submodules/simple-knn/ext.cpp:16,
spatial.cu:15-26): ``distCUDA2(points[P,3]) -> float32[P]``, the mean of the
three smallest squared distances from each point to the other points, on the
C ABI of libghr.so (no per-call cudaMalloc, no host synchronisation)."""
import torch

from .._lib import check, lib


def distCUDA2(points):
    if not points.is_cuda:
        raise RuntimeError("distCUDA2: points must be a CUDA tensor")
    P = points.size(0)
    out = torch.zeros(P, dtype=torch.float32, device=points.device)
    if P == 0:
        return out
    pts = points.float().contiguous()
    with torch.cuda.device(points.device):
        l = lib()
        ws = torch.empty(l.ghr_knn_workspace_bytes(P), dtype=torch.uint8, device=points.device)
        check(l.ghr_dist2_knn3(P, pts.data_ptr(), out.data_ptr(), ws.data_ptr(), ws.numel(),
                               torch.cuda.current_stream().cuda_stream), "ghr_dist2_knn3")
    return out
