"""Time-of-flight extraction with the reference's operator surface
(/root/reference/utils/mask2samples.py), computed by libstofnet_b200.so.

``get_maxima_positions`` / ``mask2coords`` run ONE fused kernel (stof_peaks: sliding-window NMS,
threshold / row-max selection, ordered per-row compaction, division by ``upsample_factor``)
instead of the reference's max_pool1d + elementwise passes + torch.nonzero + bincount + per-row
Python pack loop (:92-102).  Return shapes follow the reference exactly, including the
``zeros((N, C, 1))`` result when nothing is detected anywhere (:87-88) and right zero padding.
"""
import ctypes as C

import torch

from .._lib import lib, check
from ..engine import require_cuda, stream_ptr

_DEFAULT_CAP = 64


def _as_rows(scores):
    require_cuda(scores, "scores")
    s = scores.detach().to(torch.float32)
    if s.dim() == 3:
        if s.shape[1] != 1:
            raise ValueError("stofnet_b200: scores must have a single channel, got shape %s" % (tuple(s.shape),))
        rows = s[:, 0, :]
    elif s.dim() == 2:
        rows = s
    else:
        raise ValueError("stofnet_b200: scores must be (N, 1, M), got shape %s" % (tuple(s.shape),))
    return rows.contiguous()


def _th(threshold):
    none = (threshold is None) or (not threshold)      # `if threshold:` (reference :16)
    return int(none), 0.0 if none else float(threshold)


def _run_peaks(rows, window_size, threshold, upsample_factor, want_pos, cap):
    N, M = rows.shape
    dev = rows.device
    none, th = _th(threshold)
    while True:
        coords = torch.empty((N, cap), dtype=torch.float32, device=dev)
        pos = torch.empty((N, cap), dtype=torch.int32, device=dev) if want_pos else None
        counts = torch.empty((N,), dtype=torch.int32, device=dev)
        stats = torch.zeros((2,), dtype=torch.int64, device=dev)
        with torch.cuda.device(dev):
            check(lib.stof_peaks(rows.data_ptr(), N, M, int(window_size), none, th, int(upsample_factor),
                                 coords.data_ptr(), pos.data_ptr() if want_pos else None, cap,
                                 counts.data_ptr(), stats.data_ptr(), stream_ptr(dev)))
        kmax, total = (int(v) for v in stats.tolist())        # the one host sync (reference :93)
        if kmax <= cap:
            return coords, pos, counts, kmax, total
        cap = kmax


def nms_1d(scores, window_size):
    """(scores == max_pool1d(scores, odd(window), 1, pad)).float() * scores   (reference :5-11)."""
    rows = _as_rows(scores)
    out = torch.empty_like(rows)
    with torch.cuda.device(rows.device):
        check(lib.stof_nms_1d(rows.data_ptr(), rows.shape[0], rows.shape[1], int(window_size), out.data_ptr(),
                              stream_ptr(rows.device)))
    return out.view(scores.shape)


def thresholding(scores, threshold=None):
    """Reference :14-23 as one kernel (stof_thresholding).  A float threshold zeroes the entries below it IN PLACE and returns
    the caller's tensor, like `scores[scores < threshold] = 0`; None / 0.0 returns a new tensor in which every entry below
    its row maximum is zero (masked_fill)."""
    require_cuda(scores, "scores")
    if scores.dtype != torch.float32:
        raise RuntimeError("stofnet_b200: thresholding expects float32 scores, got %s" % scores.dtype)
    none, th = _th(threshold)
    m = scores.shape[-1]
    src = scores if scores.is_contiguous() else scores.contiguous()
    out = torch.empty_like(src) if none else src
    if src.numel():
        with torch.cuda.device(src.device):
            check(lib.stof_thresholding(src.data_ptr(), src.numel() // m, m, none, th, out.data_ptr(), stream_ptr(src.device)))
    if none:
        return out
    if src is not scores:
        scores.copy_(src)
    return scores


def get_maxima_positions(scores, window_size, threshold=None):
    """(K, 2) int64 [row, position], row-major ascending (reference :26-34)."""
    rows = _as_rows(scores)
    N = rows.shape[0]
    _, pos, counts, kmax, total = _run_peaks(rows, window_size, threshold, 1, True, _DEFAULT_CAP)
    idx = torch.empty((total, 2), dtype=torch.int64, device=rows.device)
    if total:
        offsets = torch.empty((N,), dtype=torch.int64, device=rows.device)
        with torch.cuda.device(rows.device):
            check(lib.stof_peaks_indices(pos.data_ptr(), counts.data_ptr(), N, pos.shape[1], offsets.data_ptr(),
                                         idx.data_ptr(), stream_ptr(rows.device)))
    return idx


def mask2coords(scores, window_size, threshold=None, upsample_factor=1, echo_max=None):
    """(N, K_max) fp32 time-of-flight table: ascending positions / upsample_factor, zero padded
    (reference :81-114)."""
    rows = _as_rows(scores)
    N, M = rows.shape
    int_r = isinstance(upsample_factor, int) or float(upsample_factor).is_integer()
    if echo_max:
        # reference :103-111: the amplitude selection works on the undivided positions, the division comes last.
        # One kernel (stof_reduce_echoes): amplitudes, keep the echo_max largest, re-order by position, zero pad, divide.
        pos, _, _, kmax, _ = _run_peaks(rows, window_size, threshold, 1, False, _DEFAULT_CAP)
        if kmax == 0:
            return torch.zeros((scores.shape[0], scores.shape[1] if scores.dim() == 3 else 1, 1), device=scores.device)
        echo_max = int(echo_max)
        out = torch.empty((N, echo_max), dtype=torch.float32, device=rows.device)
        with torch.cuda.device(rows.device):
            check(lib.stof_reduce_echoes(rows.data_ptr(), N, M, pos.data_ptr(), pos.shape[1], kmax, echo_max,
                                         float(upsample_factor), out.data_ptr(), stream_ptr(rows.device)))
        return out
    coords, _, counts, kmax, _ = _run_peaks(rows, window_size, threshold, int(upsample_factor) if int_r else 1,
                                            False, _DEFAULT_CAP)
    if kmax == 0:
        return torch.zeros((scores.shape[0], scores.shape[1] if scores.dim() == 3 else 1, 1), device=scores.device)
    coords = coords[:, :kmax].contiguous()
    if not int_r:
        coords /= upsample_factor
    return coords


def reduce_echoes(samples_and_amps, echo_max=100):
    """(N, K, C) entries [time sample, amplitude, ...] -> the echo_max entries with the largest amplitude, ordered by time
    sample (reference :117-132).  K <= echo_max returns the input unchanged (the reference's function raises UnboundLocalError
    there - `echoes` is only bound inside its `if` - and mask2coords never calls it that way).  One kernel
    (stof_reduce_echoes_pairs)."""
    require_cuda(samples_and_amps, "samples_and_amps")
    n, k, c = samples_and_amps.shape
    if k <= echo_max:
        return samples_and_amps
    src = samples_and_amps.detach().to(torch.float32).contiguous()
    out = torch.empty((n, int(echo_max), c), dtype=torch.float32, device=src.device)
    with torch.cuda.device(src.device):
        check(lib.stof_reduce_echoes_pairs(src.data_ptr(), n, k, c, int(echo_max), out.data_ptr(), stream_ptr(src.device)))
    return out


def get_amplitudes(frames, samples):
    """frames (N, 1, M) or (N, M), samples (N, K) float positions -> frames[n, round(samples[n, k])] (reference :135-136)."""
    rows = _as_rows(frames)
    require_cuda(samples, "samples")
    smp = samples.detach().to(torch.float32).contiguous()
    n, k = smp.shape
    out = torch.empty((n, k), dtype=torch.float32, device=rows.device)
    oob = torch.zeros((1,), dtype=torch.int32, device=rows.device)
    with torch.cuda.device(rows.device):
        check(lib.stof_get_amplitudes(rows.data_ptr(), smp.data_ptr(), n, rows.shape[1], k, out.data_ptr(), oob.data_ptr(),
                                      stream_ptr(rows.device)))
    if int(oob.item()):
        raise RuntimeError("stofnet_b200: get_amplitudes: index out of range (torch.gather raises here as well)")
    return out


def coords2mask(samples, ref):
    """One-hot ground-truth mask for the loss (reference :139-148): ones at the (clamped, in place like the
    reference) sample indices, column 0 cleared (placeholder for absent echoes).  One memset + one scatter
    kernel (stof_coords2mask).  samples (N, 1, G) int64, ref (N, 1, M)."""
    require_cuda(ref, "ref")
    require_cuda(samples, "samples")
    if samples.dtype != torch.int64:
        raise RuntimeError("stofnet_b200: coords2mask expects int64 sample indices (scatter_ index), got %s" % samples.dtype)
    n, m = ref.shape[0], ref.shape[-1]
    work = samples if samples.is_contiguous() else samples.contiguous()
    g = work.numel() // n
    mask = torch.empty(ref.shape, dtype=torch.float32, device=ref.device)
    with torch.cuda.device(ref.device):
        check(lib.stof_coords2mask(work.data_ptr(), n, g, m, mask.data_ptr(), stream_ptr(ref.device)))
    if work is not samples:
        samples.copy_(work)
    return mask if ref.dtype == torch.float32 else mask.to(ref.dtype)
