"""ESPCN_1D with the reference's module surface (/root/reference/models/espcn_1d.py:8-36),
computed by libstofnet_b200.so (BASELINE config 2 names its checkpoint)."""
import torch.nn as nn

from .. import _lib
from ..engine import Engine
from ..utils.sample_shuffle import SampleShuffle1D


class ESPCN_1D(nn.Module):
    def __init__(self, upscale_factor):
        super().__init__()
        self.upsample_factor = upscale_factor
        self.conv1 = nn.Conv1d(1, 64, 5, 1, 2)
        self.conv2 = nn.Conv1d(64, 32, 3, 1, 1)
        self.conv3 = nn.Conv1d(32, upscale_factor, 3, 1, 1)
        self.sample_shuffle = SampleShuffle1D(upscale_factor)
        for module in self.modules():
            if isinstance(module, nn.Conv1d):
                if module.in_channels == 32:
                    nn.init.normal_(module.weight.data, 0.0, 0.001)
                else:
                    nn.init.normal_(module.weight.data, 0.0,
                                    (2 / (module.out_channels * module.weight.data[0][0].numel())) ** .5)
                nn.init.zeros_(module.bias.data)
        self._engine = Engine(_lib.KIND_ESPCN)

    def forward(self, x):
        return self._engine.forward(self, x)

    def infer_host(self, x_host, window_size=20, threshold=None, cap=64, want_mask=False):
        return self._engine.infer_host(self, x_host, window_size, threshold, cap, want_mask)
