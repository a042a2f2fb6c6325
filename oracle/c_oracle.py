"""ctypes binding of oracle/stof_oracle.c (test infrastructure only, see oracle/__init__.py)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libstof_oracle.so")
_lib = None

MAX_CONVS = 16


class Weights(C.Structure):
    _fields_ = [
        ("conv1_w", C.c_void_p), ("conv1_b", C.c_void_p),
        ("convs_w", C.c_void_p * MAX_CONVS), ("convs_b", C.c_void_p * MAX_CONVS),
        ("last_w", C.c_void_p), ("last_b", C.c_void_p),
        ("contract_w", C.c_void_p), ("contract_b", C.c_void_p),
        ("expand_w", C.c_void_p), ("expand_b", C.c_void_p),
        ("num_features", C.c_int), ("num_blocks", C.c_int), ("upsample_factor", C.c_int),
        ("k_first", C.c_int), ("k_mid", C.c_int), ("k_last", C.c_int),
        ("sgb_scale", C.c_int), ("sgb_feat", C.c_int), ("sgb_k", C.c_int),
    ]


def build(force=False):
    src = os.path.join(_HERE, "stof_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.stof_oracle_maxima.restype = C.c_int64
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def pack_weights(sd):
    """sd: dict name -> np.ndarray with the reference state_dict key names."""
    keep = {k: _f32(v) for k, v in sd.items()}
    W = Weights()
    W.conv1_w, W.conv1_b = _p(keep["conv1.weight"]), _p(keep["conv1.bias"])
    F = keep["conv1.weight"].shape[0]
    nb = 2 + sum(1 for k in keep if k.startswith("conv") and k.endswith(".weight")
                 and k not in ("conv1.weight", "conv_last.weight"))
    for i in range(2, nb):
        W.convs_w[i - 2] = _p(keep["conv%d.weight" % i])
        W.convs_b[i - 2] = _p(keep["conv%d.bias" % i])
    W.last_w, W.last_b = _p(keep["conv_last.weight"]), _p(keep["conv_last.bias"])
    W.num_features, W.num_blocks = F, nb
    W.upsample_factor = keep["conv_last.weight"].shape[0]
    W.k_first = keep["conv1.weight"].shape[2]
    W.k_mid = keep["conv2.weight"].shape[2]
    W.k_last = keep["conv_last.weight"].shape[2]
    if "semi_global_block.contract_conv.weight" in keep:
        cw = keep["semi_global_block.contract_conv.weight"]
        W.contract_w, W.contract_b = _p(cw), _p(keep["semi_global_block.contract_conv.bias"])
        W.expand_w = _p(keep["semi_global_block.expand_conv.weight"])
        W.expand_b = _p(keep["semi_global_block.expand_conv.bias"])
        W.sgb_feat, W.sgb_k = cw.shape[0], cw.shape[2]
        W.sgb_scale = 10 * (cw.shape[0] // F)   # feat_scale = scale // 10 (stofnet.py:85)
    else:
        W.contract_w = W.contract_b = W.expand_w = W.expand_b = None
        W.sgb_scale, W.sgb_feat, W.sgb_k = 1, 0, 0
    W._keep = keep
    return W


def forward(sd, x, acc64=True):
    """x (N, L) fp32 -> mask (N, r*L) fp32; raises like the reference on a bad L."""
    x = _f32(x)
    N, L = x.shape
    W = sd if isinstance(sd, Weights) else pack_weights(sd)
    out = np.empty((N, L * W.upsample_factor), dtype=np.float32)
    rc = lib().stof_oracle_forward(C.byref(W), _p(x), N, L, _p(out), int(acc64))
    if rc:
        raise RuntimeError("stof_oracle_forward: bad input length L=%d (rc=%d)" % (L, rc))
    return out


def espcn_forward(sd, x, acc64=True):
    x = _f32(x)
    N, L = x.shape
    k = {n: _f32(v) for n, v in sd.items()}
    r = k["conv3.weight"].shape[0]
    out = np.empty((N, L * r), dtype=np.float32)
    lib().stof_oracle_espcn_forward(_p(k["conv1.weight"]), _p(k["conv1.bias"]), _p(k["conv2.weight"]),
                                    _p(k["conv2.bias"]), _p(k["conv3.weight"]), _p(k["conv3.bias"]),
                                    r, _p(x), N, L, _p(out), int(acc64))
    return out


def conv1d(x, w, b, pad, acc64=True):
    x, w, b = _f32(x), _f32(w), _f32(b)
    Cin, L = x.shape
    Cout, _, K = w.shape
    y = np.empty((Cout, L), dtype=np.float32)
    lib().stof_oracle_conv1d(_p(x), Cin, L, _p(w), _p(b), Cout, K, pad, _p(y), int(acc64))
    return y


def edsr_forward(sd, x, acc64=True):
    """EDSR_1D.forward (models/edsr_1d.py:14-19,32-45) composed from the C conv1d / shuffle primitives, row by row."""
    x = _f32(x)
    k = {n: _f32(v) for n, v in sd.items()}
    nb = 1 + max(int(n.split(".")[1]) for n in k if n.startswith("residual_blocks."))
    r = k["conv_input.weight"].shape[0] // k["conv_output.weight"].shape[1]
    rows = []
    for row in x:
        out = np.maximum(conv1d(row[None, :], k["conv_input.weight"], k["conv_input.bias"], 1, acc64), 0)
        residual = out
        for i in range(nb):
            p = "residual_blocks.%d." % i
            t = np.maximum(conv1d(out, k[p + "conv1.weight"], k[p + "conv1.bias"], 1, acc64), 0)
            out = conv1d(t, k[p + "conv2.weight"], k[p + "conv2.bias"], 1, acc64) + out
        out = conv1d(out, k["conv_mid.weight"], k["conv_mid.bias"], 1, acc64) + residual
        s = sample_shuffle(out[None], r)[0]
        rows.append(conv1d(s, k["conv_output.weight"], k["conv_output.bias"], 1, acc64)[0])
    return np.stack(rows)


def _th(threshold):
    none = (threshold is None) or (not threshold)     # "if threshold:" (mask2samples.py:16)
    return int(none), C.c_float(0.0 if none else float(threshold))


def nms_1d(scores, window):
    s = _f32(scores)
    flat = s.reshape(-1, s.shape[-1])
    out = np.empty_like(flat)
    for i in range(flat.shape[0]):
        lib().stof_oracle_nms_1d(_p(flat[i]), flat.shape[1], int(window), _p(out[i]))
    return out.reshape(s.shape)


def get_maxima_positions(scores, window, threshold=None):
    """scores (N, M) or (N,1,M) -> (K,2) int64 [row, position]."""
    s = _f32(scores)
    s = s.reshape(s.shape[0], s.shape[-1])
    N, M = s.shape
    none, th = _th(threshold)
    counts = np.zeros(N, dtype=np.int32)
    K = lib().stof_oracle_maxima(_p(s), N, M, int(window), none, th, None, C.c_int64(0), _p(counts))
    idx = np.zeros((K, 2), dtype=np.int64)
    lib().stof_oracle_maxima(_p(s), N, M, int(window), none, th, _p(idx), C.c_int64(K), _p(counts))
    return idx, counts


def mask2coords(scores, window, threshold=None, upsample_factor=1):
    """Reference return shapes: (N, kmax) fp32, or zeros((N,1,1)) when nothing is found."""
    s = _f32(scores)
    C1 = s.shape[1] if s.ndim == 3 else 1
    s = s.reshape(s.shape[0], s.shape[-1])
    N, M = s.shape
    none, th = _th(threshold)
    counts = np.zeros(N, dtype=np.int32)
    kmax = lib().stof_oracle_mask2coords(_p(s), N, M, int(window), none, th, int(upsample_factor),
                                         None, 0, _p(counts))
    if kmax == 0:
        return np.zeros((N, C1, 1), dtype=np.float32)
    coords = np.zeros((N, kmax), dtype=np.float32)
    lib().stof_oracle_mask2coords(_p(s), N, M, int(window), none, th, int(upsample_factor),
                                  _p(coords), kmax, _p(counts))
    return coords


def mask2coords_echo_max(scores, window, threshold, upsample_factor, echo_max):
    """mask2coords with echo_max (utils/mask2samples.py:103-111, reduce_echoes :117-132, get_amplitudes :135-136), numpy
    on top of the C stage-4 oracle: amplitudes at the rounded (undivided) positions, stable descending sort, first echo_max,
    stable ascending sort by position, zero padding when echo_max > K_max, division last."""
    s = _f32(scores)
    coords = mask2coords(s, window, threshold, 1)
    if coords.ndim == 3:
        return coords
    rows = s.reshape(s.shape[0], s.shape[-1])
    K = coords.shape[1]
    if echo_max < K:
        amps = np.take_along_axis(rows, np.rint(coords).astype(np.int64), axis=1)
        key = np.where(np.isnan(amps), np.float32(np.inf), amps)                    # torch.sort: NaN is the largest value
        nanfirst = np.isnan(amps)
        order = np.lexsort((np.arange(K)[None, :].repeat(len(rows), 0), -key.astype(np.float64), ~nanfirst), axis=1)[:, :echo_max]
        kept = np.take_along_axis(coords, order, axis=1)
        coords = np.take_along_axis(kept, np.argsort(kept, axis=1, kind="stable"), axis=1)
    elif echo_max > K:
        coords = np.concatenate([coords, np.zeros((coords.shape[0], echo_max - K), np.float32)], axis=1)
    return (coords / np.float32(upsample_factor)).astype(np.float32)


def coords2mask(samples, M):
    s = np.ascontiguousarray(samples, dtype=np.int64)
    s2 = s.reshape(s.shape[0], s.shape[-1])
    out = np.empty((s2.shape[0], M), dtype=np.float32)
    lib().stof_oracle_coords2mask(_p(s2), s2.shape[0], s2.shape[1], int(M), _p(out))
    return out.reshape(s.shape[:-1] + (M,))


def toa_rmse(gt, es, tol=1.0):
    gt, es = _f32(gt), _f32(es)
    out = np.empty((gt.shape[0], 7), dtype=np.float32)
    lib().stof_oracle_toa_rmse(_p(gt), gt.shape[1], _p(es), es.shape[1], gt.shape[0],
                               C.c_float(tol), _p(out))
    return out


def sample_shuffle(x, r):
    x = _f32(x)
    B, Cs, W = x.shape
    out = np.empty((B, Cs // r, W * r), dtype=np.float32)
    lib().stof_oracle_sample_shuffle(_p(x), B, Cs, W, int(r), _p(out))
    return out


def mask_loss(pred, samples, gauss, amp=20.0, lam=1e-2, want_grad=True):
    """main.py:228-232 -> (loss, grad (N,M) or None, blur_max)."""
    pred = _f32(pred)
    n, m = pred.shape[0], pred.shape[-1]
    pred2 = pred.reshape(n, m)
    s2 = np.ascontiguousarray(samples, dtype=np.int64).reshape(n, -1)
    g = _f32(gauss).reshape(-1)
    grad = np.empty((n, m), dtype=np.float32) if want_grad else None
    mx = C.c_float(0)
    f = lib().stof_oracle_mask_loss
    f.restype = C.c_double
    loss = f(_p(pred2), _p(s2), n, s2.shape[1], m, _p(g), g.size, C.c_float(amp), C.c_float(lam),
             _p(grad) if want_grad else None, C.byref(mx))
    return float(loss), grad, float(mx.value)


def find_threshold(pred, truth, norm_opt=False):
    """utils/threshold.py:6-28 (sklearn roc_curve restated)."""
    pred = _f32(pred); truth = _f32(truth)
    n = pred.shape[0]
    p2 = pred.reshape(n, -1); t2 = truth.reshape(n, -1)
    f = lib().stof_oracle_find_threshold
    f.restype = C.c_double
    return float(f(_p(p2), _p(t2), n, p2.shape[1], int(bool(norm_opt))))


def rf_upscale(re, im, rf, fs=1.0, fc=0.0):
    """datasets/chirp_dataset.py:10-17,79-91."""
    re = _f32(re)
    re2 = re.reshape(-1, re.shape[-1])
    im2 = _f32(im).reshape(re2.shape) if im is not None else None
    n = re2.shape[1]
    out = np.empty((re2.shape[0], int(n * rf)), dtype=np.float32)
    lib().stof_oracle_rf_upscale(_p(re2), _p(im2) if im2 is not None else None, re2.shape[0], n, C.c_double(rf),
                                 C.c_double(fs), C.c_double(fc), _p(out))
    return out


def attention(x, wq, bq, wk, bk, wv, bv):
    """utils/attention_block.py:13-25, dropout disabled."""
    x = _f32(x)
    b, s, d = x.shape
    ws = [_f32(a) for a in (wq, bq, wk, bk, wv, bv)]
    h = ws[0].shape[0]
    out = np.empty((b, s, h), dtype=np.float32)
    lib().stof_oracle_attention(_p(x), b, s, d, h, *[_p(a) for a in ws], _p(out))
    return out
