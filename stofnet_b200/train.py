"""Training step of the reference's loop (/root/reference/main.py:199-248,179-180) on the B200 kernels.

    masks_pred = model(frame)                      stof_forward_train   (activations kept)
    loss       = MSE(pred, blur) + lambda L1(pred) stof_mask_loss       (loss + d loss / d pred in one pass)
    loss.backward()                                stof_backward        (all parameter gradients -> ONE flat buffer)
    [all_reduce of the flat gradient bucket]       torch.distributed    (NCCL over NVLink; the only collective)
    optimizer.step()  (AdamW)                      stof_adamw_step      (fused, on the flat parameter buffer)

Parameters stay ordinary nn.Parameters (state_dict / checkpoints unchanged) but are re-pointed into one flat fp32
buffer so that the optimizer and the gradient all-reduce are single launches.  Data parallelism: every
rank holds the same weights and its own rows; the loss is the mean over the GLOBAL batch, so per-rank gradients
(means over equal local batches) are averaged, and the target normalisation `blur.max()` (main.py:230, a
batch-wide maximum) is MAX-all-reduced first so that every rank sees the single-GPU target.
"""
import math

import torch
import torch.distributed as dist

from ._lib import lib, check
from .engine import stream_ptr
from .loss import mask_loss_raw, blur_max
from .utils.gaussian import gaussian_kernel


def cosine_lr(base_lr, epoch, t_max, eta_min=0.0):
    """optim.lr_scheduler.CosineAnnealingLR closed form (main.py:180)."""
    return eta_min + (base_lr - eta_min) * (1.0 + math.cos(math.pi * epoch / t_max)) / 2.0


def _world(group):
    return dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1


def average_gradients(flat, group=None):
    """Data-parallel gradient exchange: ONE all-reduce of the flat bucket, then the mean."""
    w = _world(group)
    if w > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        flat.div_(w)
    return flat


def global_blur_max(local_max, group=None):
    if _world(group) > 1:
        dist.all_reduce(local_max, op=dist.ReduceOp.MAX, group=group)
    return local_max


class Trainer:
    """AdamW training of a stofnet_b200 StofNet (lr / weight_decay defaults: config.yaml:14,16)."""

    def __init__(self, model, lr=5e-4, weight_decay=1e-8, betas=(0.9, 0.999), eps=1e-8, kernel_size=7, sigma=1.0,
                 mask_amplitude=20.0, lambda_value=1e-2, group=None):
        self.model = model
        self.group = group
        self.lr, self.weight_decay, self.betas, self.eps = lr, weight_decay, betas, eps
        self.gauss = gaussian_kernel(kernel_size, sigma)
        self.mask_amplitude, self.lambda_value = mask_amplitude, lambda_value
        self.names = [k for k, _ in model.named_parameters()]
        params = [p for _, p in model.named_parameters()]
        dev = params[0].device
        if dev.type != "cuda":
            raise RuntimeError("stofnet_b200: Trainer needs the model on a CUDA device (no CPU path)")
        self.device = dev
        self.sizes = [p.numel() for p in params]
        self.pad = [(-s) % 4 for s in self.sizes]
        n = sum(self.sizes) + sum(self.pad)
        self.flat = torch.zeros((n,), dtype=torch.float32, device=dev)
        o = 0
        with torch.no_grad():
            for p, s, pd in zip(params, self.sizes, self.pad):
                self.flat[o:o + s].copy_(p.detach().reshape(-1))
                p.data = self.flat[o:o + s].view(p.shape)        # same layout as Engine.backward's gradient bucket
                o += s + pd
        self.exp_avg = torch.zeros_like(self.flat)
        self.exp_avg_sq = torch.zeros_like(self.flat)
        self.step_count = 0

    def step(self, frame, gt_true, lr=None):
        """One optimisation step on this rank's rows.  frame (N,1,L) fp32 CUDA, gt_true (N,1,G) int64 mask indices
        (main.py:218).  Returns the (global-mean) loss as a 0-dim fp64 CUDA tensor."""
        model, eng = self.model, self.model._engine
        mask, xc, ws = eng.forward_train(model, frame)
        mx = global_blur_max(blur_max(gt_true, mask.shape[-1], self.gauss), self.group)
        loss, g_mask = mask_loss_raw(mask, gt_true, self.gauss, self.mask_amplitude, self.lambda_value, want_grad=True,
                                     blur_max_value=mx)
        grads, _ = eng.backward(model, xc, g_mask, ws, self.names)
        average_gradients(grads, self.group)
        w = _world(self.group)
        if w > 1:
            dist.all_reduce(loss, op=dist.ReduceOp.SUM, group=self.group)
            loss = loss / w
        self.step_count += 1
        with torch.cuda.device(self.device):
            check(lib.stof_adamw_step(self.flat.data_ptr(), grads.data_ptr(), self.exp_avg.data_ptr(),
                                      self.exp_avg_sq.data_ptr(), self.flat.numel(), self.step_count,
                                      float(self.lr if lr is None else lr), float(self.betas[0]), float(self.betas[1]),
                                      float(self.eps), float(self.weight_decay), stream_ptr(self.device)))
        # parameters changed in place through the flat buffer: the kernels' layouts follow on the device, same stream
        eng.sync_device(model, self.device)
        self.last_grads = grads
        return loss.reshape(())
