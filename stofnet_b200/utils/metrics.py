"""ToA metrics with the reference's surface (/root/reference/utils/metrics.py), computed by
libstofnet_b200.so (stof_toa_rmse: one warp per row instead of the reference's Python row loop, :16-34)."""
import torch

from .._lib import lib, check
from ..engine import require_cuda, stream_ptr


def toa_rmse(gt_samples, es_samples, tol=1):
    """(N, 7) = rmse, precision, recall, jaccard [%], tp, fp, fn per row (reference :9-41).
    `tol` (config `etol`) compares with the SQUARED distance (:27-33)."""
    require_cuda(gt_samples, "gt_samples")
    require_cuda(es_samples, "es_samples")
    n = gt_samples.shape[0]
    gt = gt_samples.detach().to(torch.float32).reshape(n, -1).contiguous()
    es = es_samples.detach().to(torch.float32).reshape(es_samples.shape[0], -1).contiguous()
    if es.shape[0] != n:
        raise RuntimeError("stofnet_b200: toa_rmse got %d ground-truth rows and %d estimate rows" % (n, es.shape[0]))
    out = torch.empty((n, 7), dtype=torch.float32, device=gt.device)
    with torch.cuda.device(gt.device):
        check(lib.stof_toa_rmse(gt.data_ptr(), gt.shape[1], es.data_ptr(), es.shape[1], n, float(tol), out.data_ptr(),
                                stream_ptr(gt.device)))
    return out
