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

import numpy as np
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
        self._params = params
        self.exp_avg = torch.zeros_like(self.flat)
        self.exp_avg_sq = torch.zeros_like(self.flat)
        self.step_count = 0

    def step(self, frame, gt_true, lr=None):
        """One optimisation step on this rank's rows.  frame (N,1,L) fp32 CUDA, gt_true (N,1,G) int64 mask indices
        (main.py:218).  Returns the (global-mean) loss as a 0-dim fp64 CUDA tensor."""
        self.step_count += 1
        return self._step_impl(frame, gt_true, float(self.lr if lr is None else lr), None)

    def _step_impl(self, frame, gt_true, lr, hyper_dev):
        model, eng = self.model, self.model._engine
        # the handle follows the parameters through sync_device below; a full sync is needed only when somebody else wrote them
        # (load_state_dict, a manual edit: torch bumps the tensors' version counters)
        ver = tuple(p._version for p in self._params)
        synced = (ver == getattr(self, "_seen_versions", None))
        mask, xc, ws = eng.forward_train(model, frame, synced=synced)
        self._seen_versions = ver
        mx = global_blur_max(blur_max(gt_true, mask.shape[-1], self.gauss), self.group)
        loss, g_mask = mask_loss_raw(mask, gt_true, self.gauss, self.mask_amplitude, self.lambda_value, want_grad=True,
                                     blur_max_value=mx)
        grads, _ = eng.backward(model, xc, g_mask, ws, self.names, reuse=True)
        average_gradients(grads, self.group)
        w = _world(self.group)
        if w > 1:
            dist.all_reduce(loss, op=dist.ReduceOp.SUM, group=self.group)
            loss = loss / w
        with torch.cuda.device(self.device):
            if hyper_dev is None:
                check(lib.stof_adamw_step(self.flat.data_ptr(), grads.data_ptr(), self.exp_avg.data_ptr(),
                                          self.exp_avg_sq.data_ptr(), self.flat.numel(), self.step_count, lr,
                                          float(self.betas[0]), float(self.betas[1]), float(self.eps), float(self.weight_decay),
                                          stream_ptr(self.device)))
            else:
                check(lib.stof_adamw_step_dev(self.flat.data_ptr(), grads.data_ptr(), self.exp_avg.data_ptr(),
                                              self.exp_avg_sq.data_ptr(), self.flat.numel(), hyper_dev.data_ptr(),
                                              stream_ptr(self.device)))
        # parameters changed in place through the flat buffer: the kernels' layouts follow on the device, same stream
        eng.sync_device(model, self.device, cache_key=self.flat.data_ptr())
        self.last_grads = grads                      # the Trainer's gradient bucket: overwritten by the next step
        return loss.reshape(())

    # -- the whole step as ONE CUDA graph --------------------------------------------------------------------------------
    def capture(self, frame, gt_true):
        """Capture forward + loss + backward + AdamW + weight re-layout for batches of this shape in a CUDA graph (single
        process).  At the reference's batch size of 4 (config.yaml:11) the step is ~140 small launches and launch-bound;
        replaying the graph removes the per-launch host cost.  Nothing step-dependent is baked in: the inputs are copied
        into static buffers and the AdamW scalars (lr, bias corrections) into a 7-float device buffer before each replay."""
        if _world(self.group) > 1:
            raise RuntimeError("stofnet_b200: Trainer.capture supports single-process training (no collective in the graph)")
        dev = self.device
        self._g_frame, self._g_gt = frame.detach().clone(), gt_true.detach().clone()
        self._g_hyper = torch.zeros((8,), dtype=torch.float32, device=dev)
        # the warm-up steps are real optimizer steps: snapshot the training state and put it back afterwards
        snap = (self.flat.clone(), self.exp_avg.clone(), self.exp_avg_sq.clone(), self.step_count)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):                                    # warm-up: every lazy allocation happens here
            for _ in range(2):
                self._fill_hyper(self.lr)
                self._step_impl(self._g_frame, self._g_gt, self.lr, self._g_hyper)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self._graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._graph):
            self._g_loss = self._step_impl(self._g_frame, self._g_gt, self.lr, self._g_hyper)
        self.flat.copy_(snap[0]); self.exp_avg.copy_(snap[1]); self.exp_avg_sq.copy_(snap[2]); self.step_count = snap[3]
        self.model._engine.sync_device(self.model, dev)
        return self

    def _fill_hyper(self, lr):
        self.step_count += 1
        # the bias corrections exactly as stof_adamw_step forms them: pow() in double of the FLOAT betas
        b1, b2 = float(np.float32(self.betas[0])), float(np.float32(self.betas[1]))
        # a fresh pinned block per step: the caching host allocator keeps it alive until the asynchronous copy has run
        h = torch.tensor([float(lr), b1, b2, float(self.eps), float(self.weight_decay), 1.0 - b1 ** self.step_count,
                          math.sqrt(1.0 - b2 ** self.step_count), 0.0], dtype=torch.float32).pin_memory()
        self._g_hyper.copy_(h, non_blocking=True)

    def step_graph(self, frame, gt_true, lr=None):
        """Replay the captured step on a new batch of the captured shape; returns the loss (a static 0-dim CUDA tensor that
        the next replay overwrites)."""
        self._fill_hyper(self.lr if lr is None else lr)
        self._g_frame.copy_(frame, non_blocking=True)
        self._g_gt.copy_(gt_true, non_blocking=True)
        self._graph.replay()
        check(lib.stof_model_mark_weights_changed(self.model._engine.handle))
        return self._g_loss
