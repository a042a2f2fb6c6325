"""Torch-CPU restatement of the reference path - the timed CPU baseline ("port").

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  The reference is Python that calls
PyTorch; /root/reference does not exist on the GPU box, so this module re-issues the same
ATen calls in the same order from the state_dict alone (functional form, no nn.Module), and
therefore runs the very CPU kernels (oneDNN convolution, ATen max_pool1d / nonzero /
bincount) the reference would run there.  Checked against tests/golden/*.npz in
tests/test_oracle.py (bit-identical to the reference's fp32 output in the build container).

Citations are relative to /root/reference.
"""
import numpy as np
import torch
import torch.nn.functional as F


def to_torch_sd(sd):
    return {k: (v if isinstance(v, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(v)))
            for k, v in sd.items()}


def sample_shuffle(x, r):
    """utils/sample_shuffle.py:10-28."""
    b, cs, w = x.shape
    cl = cs // r
    return x.contiguous().view(b, r, cl, w).permute(0, 2, 3, 1).contiguous().view(b, cl, r * w)


def semi_global_block(sd, x, scale=80):
    """models/stofnet.py:98-117."""
    p = "semi_global_block."
    s = F.conv1d(x, sd[p + "contract_conv.weight"], sd[p + "contract_conv.bias"], padding=2)
    s = F.leaky_relu(s)
    s = F.max_pool1d(s, scale, scale)
    s = F.conv1d(s, sd[p + "expand_conv.weight"], sd[p + "expand_conv.bias"], padding=2)
    s = F.leaky_relu(s)
    s = F.interpolate(s, scale_factor=scale, mode="nearest")
    pad = max(0, x.size(-1) - s.size(-1))
    s = F.pad(s, (pad // 2, pad // 2))
    return torch.add(x, s)


def stofnet_forward(sd, x, num_blocks=13):
    """models/stofnet.py:42-67.  x (N,1,L) -> (N,1,r*L)."""
    r = sd["conv_last.weight"].shape[0]
    x = F.relu(F.conv1d(x, sd["conv1.weight"], sd["conv1.bias"], padding=4))
    if "semi_global_block.contract_conv.weight" in sd:
        x = semi_global_block(sd, x)
    res, res1 = x, x
    for i in range(2, num_blocks - 1):
        y = F.conv1d(x, sd["conv%d.weight" % i], sd["conv%d.bias" % i], padding="same")
        if i >= 3 and i % 2 == 1:
            x = torch.add(res, y)
            res = x
        else:
            x = F.leaky_relu(y)
    i = num_blocks - 1
    x = torch.add(res1, F.conv1d(x, sd["conv%d.weight" % i], sd["conv%d.bias" % i], padding="same"))
    x = F.conv1d(x, sd["conv_last.weight"], sd["conv_last.bias"], padding=1)
    return sample_shuffle(x, r)


def espcn_forward(sd, x):
    """models/espcn_1d.py:31-36."""
    r = sd["conv3.weight"].shape[0]
    x = torch.tanh(F.conv1d(x, sd["conv1.weight"], sd["conv1.bias"], padding=2))
    x = torch.tanh(F.conv1d(x, sd["conv2.weight"], sd["conv2.bias"], padding=1))
    return torch.sigmoid(sample_shuffle(F.conv1d(x, sd["conv3.weight"], sd["conv3.bias"], padding=1), r))


def edsr_forward(sd, x):
    """models/edsr_1d.py:14-19,32-45."""
    out = F.relu(F.conv1d(x, sd["conv_input.weight"], sd["conv_input.bias"], padding=1))
    residual = out
    nb = 1 + max(int(k.split(".")[1]) for k in sd if k.startswith("residual_blocks."))
    for i in range(nb):
        p = "residual_blocks.%d." % i
        t = F.relu(F.conv1d(out, sd[p + "conv1.weight"], sd[p + "conv1.bias"], padding=1))
        out = F.conv1d(t, sd[p + "conv2.weight"], sd[p + "conv2.bias"], padding=1) + out
    out = F.conv1d(out, sd["conv_mid.weight"], sd["conv_mid.bias"], padding=1) + residual
    r = sd["conv_input.weight"].shape[0] // sd["conv_output.weight"].shape[1]
    out = sample_shuffle(out, r)
    return F.conv1d(out, sd["conv_output.weight"], sd["conv_output.bias"], padding=1)


def get_maxima_positions(scores, window_size, threshold=None):
    """utils/mask2samples.py:5-34."""
    s = scores.clone().detach()
    w = window_size // 2 * 2 + 1
    m = F.max_pool1d(s, w, stride=1, padding=(w - 1) // 2).squeeze(0)
    s = (s == m).float() * s
    if threshold:
        s[s < threshold] = 0
    else:
        s = s.masked_fill(s < torch.max(s, dim=-1, keepdim=True)[0], 0)
    return torch.nonzero(s.squeeze(1), as_tuple=False).long()


def thresholding(scores, threshold=None):
    """utils/mask2samples.py:14-23 (float threshold: in place)."""
    if threshold:
        scores[scores < threshold] = 0
    else:
        scores = scores.masked_fill(scores < torch.max(scores, dim=-1, keepdim=True)[0], 0)
    return scores


def get_amplitudes(frames, samples):
    """utils/mask2samples.py:135-136."""
    return torch.gather(frames.squeeze(), -1, torch.round(samples).long())


def reduce_echoes(pairs, echo_max):
    """utils/mask2samples.py:117-132: (N, K, 2) [position, amplitude] -> the echo_max largest amplitudes, by position."""
    if pairs.shape[1] > echo_max:
        c = pairs.shape[-1]
        order = torch.argsort(pairs[..., 1], descending=True, dim=1, stable=True)
        pairs = torch.gather(pairs, 1, order[..., None].expand(-1, -1, c))[:, :echo_max]
        order = torch.argsort(pairs[..., 0], descending=False, dim=1, stable=True)
        pairs = torch.gather(pairs, 1, order[..., None].expand(-1, -1, c))
    return pairs


def mask2coords(scores, window_size, threshold=None, upsample_factor=1, echo_max=None):
    """utils/mask2samples.py:81-114, including its per-row Python pack loop and the echo_max branch (:103-110)."""
    idx = get_maxima_positions(scores, window_size, threshold)
    if idx.numel() == 0:
        return torch.zeros((scores.shape[0], scores.shape[1], 1))
    n = scores.shape[0]
    counts = torch.bincount(idx[:, 0], minlength=n)
    kmax = int(max(counts))
    rows = []
    for i, c in enumerate(counts):
        rows.append(torch.cat((torch.arange(0, c) + i * kmax, -1 * torch.ones(kmax - c)), dim=-1))
    flat = torch.vstack(rows).long().flatten()
    flat = flat[flat >= 0]
    coords = torch.zeros((n, kmax))
    coords.view(-1)[flat] = idx[:, 1].float()
    if echo_max and echo_max < coords.shape[-1]:
        amps = torch.gather(scores.squeeze(), -1, torch.round(coords).long())          # get_amplitudes :135-136
        coords = reduce_echoes(torch.dstack([coords, amps]), echo_max)[..., 0]
    elif echo_max and echo_max > coords.shape[-1]:
        coords = torch.cat([coords, torch.zeros(coords.shape[0], echo_max - coords.shape[-1])], dim=-1)
    coords /= upsample_factor
    return coords


def infer(sd, x, window_size=20, threshold=None, espcn=False):
    """The reference's eval step (main.py:313-320): forward then mask2coords."""
    with torch.no_grad():
        m = espcn_forward(sd, x) if espcn else stofnet_forward(sd, x)
        r = m.shape[-1] // x.shape[-1]
        return m, mask2coords(m, window_size, threshold, upsample_factor=r)
