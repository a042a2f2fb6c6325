"""Ideal-threshold estimate with the reference's surface (/root/reference/utils/threshold.py:6-28).

The reference copies the whole mask to the host and calls sklearn.metrics.roc_curve on batch element 0
(main.py:330); here the ROC (descending sort, cumulative tp / fp per distinct score, drop_intermediate,
(0,0) point at +inf) and the arg-max of the G-mean run on the device (stof_find_threshold); only the
8-byte result crosses the bus."""
import torch

from .._lib import lib, check
from ..engine import require_cuda, stream_ptr


def find_threshold(masks_pred, masks_true, window_size=None, norm_opt=False):
    require_cuda(masks_pred, "masks_pred")
    require_cuda(masks_true, "masks_true")
    n = masks_pred.shape[0]
    pred = masks_pred.detach().to(torch.float32).reshape(n, -1).contiguous()
    true = masks_true.detach().to(torch.float32).reshape(n, -1).contiguous()
    if pred.shape != true.shape:
        raise RuntimeError("stofnet_b200: find_threshold shapes differ: %s vs %s" % (tuple(pred.shape), tuple(true.shape)))
    dev = pred.device
    m = pred.shape[1]
    th = torch.empty((1,), dtype=torch.float64, device=dev)
    scratch = torch.empty((lib.stof_threshold_scratch_bytes(n, m),), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib.stof_find_threshold(pred.data_ptr(), true.data_ptr(), n, m, int(bool(norm_opt)), th.data_ptr(),
                                      scratch.data_ptr(), stream_ptr(dev)))
    return float(th.item())
