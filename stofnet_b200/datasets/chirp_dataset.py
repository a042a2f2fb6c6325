"""rf_scale_factor pre-interpolation with the reference's function names
(/root/reference/datasets/chirp_dataset.py:10-17 upscale_1d, :79-91 ChirpDataset.iq2rf), computed on the
device by stof_rf_upscale: linear interpolation on np.linspace(endpoint=True) grids in fp64 (scipy interp1d
semantics), optional IQ -> RF carrier modulation, fp32 output = the (N, L) waveforms StofNet.forward takes
after main.py:213's `.float()`.  The dataset loader itself (txt files, natsort, omegaconf) is out of scope."""
import torch

from .._lib import lib, check
from ..engine import require_cuda, stream_ptr


def _rows(t, what):
    require_cuda(t, what)
    t = t.detach()
    return (t if t.dim() == 2 else t.reshape(1, -1)), t.dim() == 1


def upscale_1d(data, rescale_factor, fs=1):
    """data (n,) or (N, n) -> (int(n * rescale_factor),) or (N, .) fp32."""
    rows, squeeze = _rows(data, "data")
    re = rows.to(torch.float32).contiguous()
    n_rows, n = re.shape
    out = torch.empty((n_rows, int(n * rescale_factor)), dtype=torch.float32, device=re.device)
    with torch.cuda.device(re.device):
        check(lib.stof_rf_upscale(re.data_ptr(), None, n_rows, n, float(rescale_factor), float(fs), 0.0, out.data_ptr(),
                                  stream_ptr(re.device)))
    return out[0] if squeeze else out


def iq2rf(iq_data, fc, fs, rescale_factor=1):
    """Complex IQ rows -> real RF rows: Re(interp(iq)(t) * exp(2j pi fc t))."""
    rows, squeeze = _rows(iq_data, "iq_data")
    if not rows.is_complex():
        raise RuntimeError("stofnet_b200: iq2rf expects a complex tensor")
    re = rows.real.to(torch.float32).contiguous()
    im = rows.imag.to(torch.float32).contiguous()
    n_rows, n = re.shape
    out = torch.empty((n_rows, int(n * rescale_factor)), dtype=torch.float32, device=re.device)
    with torch.cuda.device(re.device):
        check(lib.stof_rf_upscale(re.data_ptr(), im.data_ptr(), n_rows, n, float(rescale_factor), float(fs), float(fc),
                                  out.data_ptr(), stream_ptr(re.device)))
    return out[0] if squeeze else out
