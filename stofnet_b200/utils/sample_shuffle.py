"""SampleShuffle1D with the reference's surface (/root/reference/utils/sample_shuffle.py:5-28).

Inside StofNet / ESPCN_1D the shuffle is fused into the last conv's store; this standalone module
calls the same library (stof_sample_shuffle) for callers that use the operator on its own."""
import torch
import torch.nn as nn

from .._lib import lib, check
from ..engine import require_cuda, stream_ptr


class SampleShuffle1D(nn.Module):
    def __init__(self, upsample_factor):
        super().__init__()
        self.upsample_factor = upsample_factor

    def forward(self, x):
        require_cuda(x, "input")
        b, cs, w = x.shape
        r = self.upsample_factor
        x = x.detach().to(torch.float32).contiguous()
        out = torch.empty((b, cs // r, r * w), dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            check(lib.stof_sample_shuffle(x.data_ptr(), b, cs, w, r, out.data_ptr(), stream_ptr(x.device)))
        return out
