"""Training / validation loss of the reference's step (/root/reference/main.py:183-188,228-232) as one fused
device pass (stof_mask_loss):

    masks_true = coords2mask(gt_true, masks_pred)                       utils/mask2samples.py:139-148
    blur       = F.conv1d(masks_true, gauss(kernel_size, sigma), padding=kernel_size // 2)
    blur      /= blur.max();  blur *= mask_amplitude                    (max over the WHOLE batch, :230)
    loss       = MSE(pred, blur) + lambda_value * L1(pred, 0)

The one-hot mask is never materialised: the blurred target is evaluated from each row's ground-truth list.
`MaskLoss.apply` is differentiable w.r.t. masks_pred (the kernel writes d loss / d pred in the same pass), so
`loss.backward()` in a main.py-style loop reaches StofNet's backward kernels.
"""
import numpy as np
import torch

from ._lib import lib, check
from .engine import require_cuda, stream_ptr
from .utils.gaussian import gaussian_kernel


_GAUSS_CACHE = {}


def _gauss_on(gauss, dev):
    """The (tiny) blur kernel as a device tensor, uploaded once per (values, device) - not once per step."""
    a = np.asarray(gauss, dtype=np.float32).reshape(-1)
    key = (a.tobytes(), dev.type, dev.index)
    g = _GAUSS_CACHE.get(key)
    if g is None:
        if len(_GAUSS_CACHE) > 64:
            _GAUSS_CACHE.clear()
        g = _GAUSS_CACHE[key] = torch.as_tensor(a, device=dev).contiguous()
    return g


def _prep(masks_pred, gt_true):
    require_cuda(masks_pred, "masks_pred")
    require_cuda(gt_true, "gt_true")
    n, m = masks_pred.shape[0], masks_pred.shape[-1]
    pred = masks_pred.detach().to(torch.float32).reshape(n, m).contiguous()
    smp = gt_true.detach().to(torch.int64).reshape(n, -1).contiguous()
    return pred, smp, n, m


def blur_max(gt_true, m, gauss, device=None):
    """Max of the un-normalised blurred target over these rows (phase 1 of the sharded loss: all-reduce it with MAX)."""
    require_cuda(gt_true, "gt_true")
    n = gt_true.shape[0]
    smp = gt_true.detach().to(torch.int64).reshape(n, -1).contiguous()
    dev = smp.device
    g = _gauss_on(gauss, dev)
    mx = torch.zeros((1,), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(lib.stof_mask_loss(None, smp.data_ptr(), n, smp.shape[1], int(m), g.data_ptr(), g.numel(), 0.0, 0.0, 1,
                                 mx.data_ptr(), None, None, None, stream_ptr(dev)))
    return mx


def mask_loss_raw(masks_pred, gt_true, gauss, mask_amplitude=20.0, lambda_value=1e-2, want_grad=False, blur_max_value=None):
    """Returns (loss fp64 tensor (1,), grad (N, M) fp32 or None).  blur_max_value: (1,) fp32 tensor holding a global
    (all-reduced) blur max, or None to compute it over these rows like the single-GPU reference."""
    pred, smp, n, m = _prep(masks_pred, gt_true)
    dev = pred.device
    g = _gauss_on(gauss, dev)
    mx = blur_max_value if blur_max_value is not None else torch.zeros((1,), dtype=torch.float32, device=dev)
    loss = torch.empty((1,), dtype=torch.float64, device=dev)
    grad = torch.empty((n, m), dtype=torch.float32, device=dev) if want_grad else None
    scratch = torch.empty((lib.stof_loss_scratch_bytes(n, m),), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib.stof_mask_loss(pred.data_ptr(), smp.data_ptr(), n, smp.shape[1], m, g.data_ptr(), g.numel(),
                                 float(mask_amplitude), float(lambda_value), 2 if blur_max_value is not None else 0,
                                 mx.data_ptr(), loss.data_ptr(), grad.data_ptr() if want_grad else None,
                                 scratch.data_ptr(), stream_ptr(dev)))
    return loss, grad


class MaskLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, masks_pred, gt_true, gauss, mask_amplitude, lambda_value, blur_max_value):
        loss, grad = mask_loss_raw(masks_pred, gt_true, gauss, mask_amplitude, lambda_value, want_grad=True,
                                   blur_max_value=blur_max_value)
        ctx.save_for_backward(grad)
        ctx.shape = masks_pred.shape
        return loss.to(torch.float32).reshape(())

    @staticmethod
    def backward(ctx, g_out):
        (grad,) = ctx.saved_tensors
        return (grad * g_out).reshape(ctx.shape), None, None, None, None, None


def mask_loss(masks_pred, gt_true, kernel_size=7, sigma=1.0, mask_amplitude=20.0, lambda_value=1e-2, blur_max_value=None):
    """Scalar loss of main.py:232 (differentiable w.r.t. masks_pred)."""
    return MaskLoss.apply(masks_pred, gt_true, gaussian_kernel(kernel_size, sigma), mask_amplitude, lambda_value,
                          blur_max_value)
