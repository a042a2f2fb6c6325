"""StofNet with the reference's module surface, computed by libstofnet_b200.so.

Mirrors /root/reference/models/stofnet.py: same constructor signature and attribute names
(:11-40), same state_dict keys (conv1, conv2..conv{num_blocks-1}, conv_last,
semi_global_block.{contract_conv,expand_conv}), so the shipped .pth checkpoints load with
``load_state_dict(strict=True)`` (main.py:173-177).  ``forward`` (:42-67) is one C-ABI call
(stof_forward): conv stack, semi-global block with the max-pool fused into the contract conv,
and the SampleShuffle1D fused into conv_last's store.  Parameters are ordinary fp32
``nn.Parameter``s in PyTorch's (Cout, Cin, K) layout; they are repacked and uploaded to the
kernels' layout whenever they change.  There is no CPU path: CPU inputs raise.
"""
import torch
import torch.nn as nn

from .. import _lib
from ..engine import Engine
from ..utils.sample_shuffle import SampleShuffle1D


class SemiGlobalBlock(nn.Module):
    """Semi-global block with the reference's surface (:80-117).  Inside StofNet.forward its arithmetic (contract conv +
    LeakyReLU + MaxPool(80) -> expand conv + LeakyReLU -> nearest x80 -> centred pad -> skip add) runs in the fused
    tensor-core kernels; called on its own, ``forward`` runs the same stages as fp32 FFMA kernels on the caller's
    (N, 64, L) tensor (stof_semi_global_block).  Inference only (no autograd through the standalone call)."""

    def __init__(self, in_channels, out_channels, sample_scale=2, kernel_size=5):
        super().__init__()
        self.sample_scale = sample_scale
        self.feat_scale = max(1, sample_scale // 10)
        self.contract_conv = nn.Conv1d(in_channels, self.feat_scale * out_channels, kernel_size=kernel_size,
                                       stride=1, padding=kernel_size // 2)
        self.contract_relu = nn.LeakyReLU()
        self.contract_pool = nn.MaxPool1d(kernel_size=sample_scale, stride=sample_scale)
        self.expand_conv = nn.Conv1d(self.feat_scale * out_channels, out_channels, kernel_size=kernel_size,
                                     stride=1, padding=kernel_size // 2)
        self.expand_relu = nn.LeakyReLU()
        self.expand_upsample = nn.Upsample(scale_factor=sample_scale, mode="nearest")
        self._engine = None

    def forward(self, x):
        cw = self.contract_conv.weight
        if tuple(cw.shape) != (512, 64, 5) or tuple(self.expand_conv.weight.shape) != (64, 512, 5) or self.sample_scale != 80:
            raise ValueError("stofnet_b200: the semi-global block kernels are built for 64 channels, sample_scale=80, "
                             "kernel_size=5 (got contract_conv %s, sample_scale %d)" % (tuple(cw.shape), self.sample_scale))
        if self._engine is None:
            self._engine = Engine(_lib.KIND_STOFNET)
        return self._engine.semi_global_block(self, x)


class _StofNetFunction(torch.autograd.Function):
    """masks_pred = model(frame) with a backward: stof_forward_train keeps the activations, stof_backward turns
    d loss / d masks_pred into every parameter's gradient (main.py:221,247)."""

    @staticmethod
    def forward(ctx, module, names, x, *params):
        mask, xc, ws = module._engine.forward_train(module, x)
        ctx.module, ctx.names, ctx.xc, ctx.ws = module, names, xc, ws
        return mask

    @staticmethod
    def backward(ctx, g_mask):
        _, views = ctx.module._engine.backward(ctx.module, ctx.xc, g_mask, ctx.ws, ctx.names)
        ctx.ws = None
        return (None, None, None) + tuple(views)


class StofNet(nn.Module):

    def __init__(self, upsample_factor=4, num_features=64, num_blocks=13, kernel_sizes=[9, 7, 3], in_channels=1,
                 semi_global_scale=80, weights_init=False):
        super().__init__()
        self.num_blocks = num_blocks
        self.in_channels = in_channels
        self.num_features = num_features
        self.kernel_sizes = kernel_sizes
        self.upsample_factor = upsample_factor
        self.semi_global_scale = semi_global_scale

        self.conv1 = nn.Conv1d(in_channels, num_features, kernel_sizes[0], 1, 4)
        self.conv_last = nn.Conv1d(num_features, upsample_factor, kernel_sizes[-1], 1, 1)
        self.semi_global_block = (SemiGlobalBlock(num_features, num_features, semi_global_scale)
                                  if semi_global_scale != 1 else None)
        for i in range(2, num_blocks):
            setattr(self, "conv%d" % i, nn.Conv1d(num_features, num_features, kernel_sizes[1], 1, padding="same"))
        self.sample_shuffle = SampleShuffle1D(upsample_factor)
        self.residual_layers = list(range(3, num_blocks - 1, 2)) + [num_blocks - 1, num_blocks]
        if weights_init:
            self._initialize_weights()
        self._engine = Engine(_lib.KIND_STOFNET)

    def forward(self, x):
        if torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()):
            # training step (main.py:221,247): fp32 forward that keeps the activations + hand-written backward
            named = [(k, p) for k, p in self.named_parameters()]
            if not all(p.requires_grad for _, p in named):
                raise NotImplementedError("stofnet_b200: the backward kernels compute every parameter's gradient; "
                                          "freezing a subset is not supported")
            return _StofNetFunction.apply(self, tuple(k for k, _ in named), x, *[p for _, p in named])
        return self._engine.forward(self, x)

    def set_precision(self, precision="fp32"):
        """Arithmetic of the conv stack (inference): "fp32" (default) = fp32-faithful hi/lo split on the tensor cores,
        index-exact against the reference; "fp16" = ONE tensor-core pass on fp16-rounded operands with fp32 accumulation
        (about a third of the tensor work; detections agree with the exact mode only to the tolerance stated in
        DESIGN.md); "simt" = plain fp32 FFMA kernels (cross-check path)."""
        path = {"fp32": _lib.PATH_TC, "fp16": _lib.PATH_TC_FP16, "simt": _lib.PATH_SIMT}.get(precision)
        if path is None:
            raise ValueError("stofnet_b200: precision must be 'fp32', 'fp16' or 'simt', got %r" % (precision,))
        self._engine.set_path(path)
        return self

    def infer_host(self, x_host, window_size=20, threshold=None, cap=64, want_mask=False):
        """Eval step of main.py:307-320 on host tensors in one C-ABI call (stof_infer_host)."""
        return self._engine.infer_host(self, x_host, window_size, threshold, cap, want_mask)

    def _initialize_weights(self):
        # same scheme as the reference (:69-77); never enabled there (weights_init=False)
        for i in range(1, self.num_blocks):
            conv = getattr(self, "conv%d" % i)
            if i not in self.residual_layers:
                nn.init.orthogonal_(conv.weight, nn.init.calculate_gain("relu"))
            else:
                nn.init.orthogonal_(conv.weight)
        nn.init.orthogonal_(self.conv_last.weight)
