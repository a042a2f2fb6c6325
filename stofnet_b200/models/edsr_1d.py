"""EDSR_1D with the reference's module surface (/root/reference/models/edsr_1d.py:7-45), computed by
libstofnet_b200.so (comparison model with shipped checkpoints proud-cherry / snowy-dragon).

The whole stack runs on the tensor-core chain kernel: relu(conv_input) is the x0 kernel with K = 3, the 2*num_blocks + 1
k3 convs are chain layers (ReLU / short-skip / long-skip epilogues), and SampleShuffle1D + conv_output is folded into an
equivalent (4, 64, 3) conv + shuffle store (stof_api.cu: edsr_tail_weights)."""
import torch.nn as nn

from .. import _lib
from ..engine import Engine
from ..utils.sample_shuffle import SampleShuffle1D


class ResidualBlock(nn.Module):
    """Parameter container with the reference's names (models/edsr_1d.py:7-19); the arithmetic is in the chain kernel."""

    def __init__(self, channels):
        super().__init__()
        self.conv1 = nn.Conv1d(channels, channels, kernel_size=3, stride=1, padding=1)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = nn.Conv1d(channels, channels, kernel_size=3, stride=1, padding=1)


class EDSR_1D(nn.Module):
    def __init__(self, num_channels=1, num_features=64, num_blocks=8, upscale_factor=4):
        super().__init__()
        self.upsample_factor = upscale_factor
        self.conv_input = nn.Conv1d(num_channels, num_features, kernel_size=3, stride=1, padding=1)
        self.relu = nn.ReLU(inplace=True)
        self.residual_blocks = nn.ModuleList([ResidualBlock(num_features) for _ in range(num_blocks)])
        self.conv_mid = nn.Conv1d(num_features, num_features, kernel_size=3, stride=1, padding=1)
        self.upscale = SampleShuffle1D(upscale_factor)
        self.conv_output = nn.Conv1d(num_features // upscale_factor, num_channels, kernel_size=3, stride=1, padding=1)
        self._engine = Engine(_lib.KIND_EDSR)

    def forward(self, x):
        return self._engine.forward(self, x)

    def set_precision(self, mode="fp32"):
        """'fp32' (default): fp16 hi/lo split operands, fp32-faithful.  'fp16': one tensor-core pass (stated tolerance)."""
        if mode not in ("fp32", "fp16"):
            raise ValueError("stofnet_b200: precision %r (expected 'fp32' or 'fp16')" % (mode,))
        self._engine.set_path(_lib.PATH_TC_FP16 if mode == "fp16" else _lib.PATH_TC)

    def infer_host(self, x_host, window_size=20, threshold=None, cap=64, want_mask=False):
        return self._engine.infer_host(self, x_host, window_size, threshold, cap, want_mask)
