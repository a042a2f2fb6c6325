#!/usr/bin/env python
"""Generate the golden fixtures in this directory FROM THE REFERENCE ITSELF.

Run in the build container only (needs /root/reference, read-only):

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

The reference ships no tests or golden vectors, so parity is pinned on
outputs of the unmodified reference code run here with this environment's torch CPU
kernels (torch 2.11.0, oneDNN, fp32; fp64 re-run stored beside it as the tie-breaker):

  weights_<ckpt>.npz   the shipped checkpoints' tensors, verbatim fp32, original key names
                       (/root/reference/ckpts/*.pth; data, not source - they must travel
                       to the GPU box where /root/reference does not exist)
  forward_<ckpt>.npz   seeded inputs -> StofNet.forward masks (fp32 + fp64) and
                       mask2coords / get_maxima_positions outputs for th in
                       {None, checkpoint th, 0.5}
  peaks_cases.npz      hand-made stage-4 edge cases -> reference indices / coords
  espcn_<ckpt>.npz     ESPCN_1D forward masks
  edsr_<ckpt>.npz      EDSR_1D forward masks, fp32 + fp64
  metrics_cases.npz    toa_rmse / coords2mask / find_threshold known answers
  echo_cases.npz       mask2coords(echo_max=...) : amplitude selection / zero padding (utils/mask2samples.py:103-132)
  train_<ckpt>.npz     three training steps (model -> loss -> autograd -> AdamW, main.py:217-248): gradients
                       of every parameter after the first backward, losses and parameters after each step
  scale_<ckpt>.npz     the SURVEY 8c golden set: N = 64 rows x seeds {3008, 0, 1} x L in {2560, 2578, 5120, 160} per StofNet
                       checkpoint; inputs come from the host-independent generator (synth_waveforms(q15=True)) and are NOT
                       stored - only their SHA-256, the SHA-256 of the fp32 mask, per-row mask maxima, and the index lists /
                       ToF tables for th in {None, checkpoint th, 0.5}
  surface_cases.npz    standalone operators: thresholding, reduce_echoes, get_amplitudes, SemiGlobalBlock.forward
  aux_cases.npz        loss + gradient (main.py:228-232 through the reference's ops and autograd),
                       find_threshold, upscale_1d / iq2rf, AttentionBlock (dropout off)

Nothing here is imported by the product; tests/ read the .npz files only.
"""
import os
import sys
import glob

sys.dont_write_bytecode = True
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"
sys.path.insert(0, REF)

import numpy as np
import torch

from stofnet_b200.synth import synth_waveforms  # our generator (inputs only)

from models.stofnet import StofNet  # noqa: E402  (reference)
from models.espcn_1d import ESPCN_1D  # noqa: E402
from models.edsr_1d import EDSR_1D  # noqa: E402
from utils.mask2samples import mask2coords, get_maxima_positions, coords2mask, nms_1d  # noqa: E402
from utils.metrics import toa_rmse  # noqa: E402
from utils.threshold import find_threshold  # noqa: E402
from utils.sample_shuffle import SampleShuffle1D  # noqa: E402

torch.manual_seed(0)
torch.set_num_threads(8)

# checkpoint -> (rf_scale_factor, benchmark threshold) from bash_scripts/array_*_params.txt
STOF = {
    "different-armadillo": (10, None),
    "clean-serenity": (10, None),
    "graceful-snow": (20, 0.015),
    "breezy-violet": (20, 2.3),
}
ESPCN = {"vital-puddle": (10, None), "wobbly-sponge": (20, 0.2)}
EDSR = {"proud-cherry": (10, None), "snowy-dragon": (20, 1.5)}


def ckpt_path(prefix):
    (p,) = glob.glob(os.path.join(REF, "ckpts", prefix + "*.pth"))
    return p


def load_sd(prefix):
    return torch.load(ckpt_path(prefix), weights_only=True, map_location="cpu")


def th_key(th):
    return "none" if th is None else ("%g" % th).replace(".", "p").replace("-", "m")


def peaks_dict(prefix, masks, th, out):
    """Store reference stage-4 outputs for one (masks, th)."""
    idx = get_maxima_positions(masks, 20, th)
    co = mask2coords(masks, 20, th, upsample_factor=4)
    out[prefix + "_idx_" + th_key(th)] = idx.numpy().astype(np.int64)
    out[prefix + "_coords_" + th_key(th)] = co.numpy().astype(np.float32)


def gen_forward():
    for name, (rf, th) in STOF.items():
        sd = load_sd(name)
        np.savez(os.path.join(HERE, "weights_%s.npz" % name),
                 **{k: v.numpy() for k, v in sd.items()})
        has_sgb = any(k.startswith("semi_global_block") for k in sd)
        model = StofNet(upsample_factor=4, semi_global_scale=80 if has_sgb else 1).eval()
        print(name, model.load_state_dict(sd))
        model64 = StofNet(upsample_factor=4, semi_global_scale=80 if has_sgb else 1).double().eval()
        model64.load_state_dict({k: v.double() for k, v in sd.items()})
        out = {"rf": np.int64(rf), "th": np.float64(np.nan if th is None else th)}
        cases = [(3, 160, 11), (2, 2578, 3008), (2, 256 * rf, 3008), (1, 320, 1)]
        out["cases"] = np.array(cases, dtype=np.int64)
        for ci, (N, L, seed) in enumerate(cases):
            x = torch.from_numpy(synth_waveforms(N, L, seed=seed, rf=rf)).unsqueeze(1)
            with torch.no_grad():
                m = model(x)
                m64 = model64(x.double())
            p = "c%d" % ci
            out[p + "_x"] = x.squeeze(1).numpy()
            out[p + "_mask"] = m.squeeze(1).numpy()
            out[p + "_mask64"] = m64.squeeze(1).numpy()
            for t in sorted({None, th, 0.5}, key=lambda v: (v is not None, v or 0)):
                peaks_dict(p, m, t, out)
            print("  case", ci, N, L, "mask range", float(m.min()), float(m.max()),
                  "max|fp32-fp64|", float((m.double() - m64).abs().max()))
        np.savez_compressed(os.path.join(HERE, "forward_%s.npz" % name), **out)

    for name, (rf, th) in ESPCN.items():
        sd = load_sd(name)
        np.savez(os.path.join(HERE, "weights_%s.npz" % name),
                 **{k: v.numpy() for k, v in sd.items()})
        model = ESPCN_1D(4).eval()
        print(name, model.load_state_dict(sd))
        out = {"rf": np.int64(rf), "th": np.float64(np.nan if th is None else th)}
        cases = [(3, 160, 11), (2, 256 * rf, 3008), (1, 37, 2)]
        out["cases"] = np.array(cases, dtype=np.int64)
        for ci, (N, L, seed) in enumerate(cases):
            x = torch.from_numpy(synth_waveforms(N, L, seed=seed, rf=rf)).unsqueeze(1)
            with torch.no_grad():
                m = model(x)
            p = "c%d" % ci
            out[p + "_x"] = x.squeeze(1).numpy()
            out[p + "_mask"] = m.squeeze(1).numpy()
            for t in sorted({None, th, 0.5}, key=lambda v: (v is not None, v or 0)):
                peaks_dict(p, m, t, out)
        np.savez_compressed(os.path.join(HERE, "espcn_%s.npz" % name), **out)


def gen_edsr():
    """edsr_<ckpt>.npz: EDSR_1D forward masks (fp32 + fp64) and stage-4 outputs."""
    for name, (rf, th) in EDSR.items():
        sd = load_sd(name)
        np.savez(os.path.join(HERE, "weights_%s.npz" % name), **{k: v.numpy() for k, v in sd.items()})
        model = EDSR_1D().eval()
        print(name, model.load_state_dict(sd))
        model64 = EDSR_1D().double().eval()
        model64.load_state_dict({k: v.double() for k, v in sd.items()})
        out = {"rf": np.int64(rf), "th": np.float64(np.nan if th is None else th)}
        cases = [(3, 160, 11), (2, 256 * rf, 3008), (1, 37, 2), (2, 601, 5)]
        out["cases"] = np.array(cases, dtype=np.int64)
        for ci, (N, L, seed) in enumerate(cases):
            x = torch.from_numpy(synth_waveforms(N, L, seed=seed, rf=rf)).unsqueeze(1)
            with torch.no_grad():
                m = model(x)
                m64 = model64(x.double())
            p = "c%d" % ci
            out[p + "_x"] = x.squeeze(1).numpy()
            out[p + "_mask"] = m.squeeze(1).numpy()
            out[p + "_mask64"] = m64.squeeze(1).numpy()
            for t in sorted({None, th, 0.5}, key=lambda v: (v is not None, v or 0)):
                peaks_dict(p, m, t, out)
            print("  case", ci, N, L, "mask range", float(m.min()), float(m.max()),
                  "max|fp32-fp64|", float((m.double() - m64).abs().max()))
        np.savez_compressed(os.path.join(HERE, "edsr_%s.npz" % name), **out)


def gen_peaks_cases():
    rng = np.random.default_rng(7)
    cases = {}

    def add(name, scores, th, win=20, r=4):
        s = torch.from_numpy(np.ascontiguousarray(scores, dtype=np.float32))
        idx = get_maxima_positions(s.clone(), win, th)
        co = mask2coords(s.clone(), win, th, upsample_factor=r)
        sup = nms_1d(s.clone(), win)
        cases[name + "_scores"] = s.numpy()
        cases[name + "_th"] = np.float64(np.nan if th is None else th)
        cases[name + "_win"] = np.int64(win)
        cases[name + "_r"] = np.int64(r)
        cases[name + "_idx"] = idx.numpy().astype(np.int64)
        cases[name + "_coords"] = co.numpy().astype(np.float32)
        cases[name + "_nms"] = sup.numpy().astype(np.float32)
        names.append(name)

    names = []
    add("allzero_none", np.zeros((3, 1, 64)), None)
    add("allzero_th", np.zeros((3, 1, 64)), 0.5)
    add("allneg_none", -2.5 * np.ones((2, 1, 48)), None)
    add("allneg_th", -2.5 * np.ones((2, 1, 48)), 0.5)
    add("allpos_const_th", 1.5 * np.ones((2, 1, 48)), 0.5)
    x = np.zeros((2, 1, 100)); x[0, 0, 30] = 2.0; x[0, 0, 36] = 2.0; x[1, 0, 50] = 1.0; x[1, 0, 70] = 1.0
    add("ties_in_window_th", x, 0.5)
    add("ties_in_window_none", x, None)
    x = np.zeros((2, 1, 100)); x[0, 0, 0] = 3.0; x[0, 0, 99] = 3.0; x[1, 0, 10] = 0.7; x[1, 0, 20] = 0.7; x[1, 0, 31] = 0.6
    add("edges_th", x, 0.5)
    add("edges_none", x, None)
    add("n1_th", rng.normal(size=(1, 1, 400)), 0.5)
    add("n1_none", rng.normal(size=(1, 1, 400)), None)
    add("th_zero_is_none", rng.normal(size=(3, 1, 257)), 0.0)
    add("neg_th", rng.normal(size=(3, 1, 257)), -0.25)
    x = rng.normal(size=(4, 1, 300)).astype(np.float32)
    x[0, 0, 17] = np.nan; x[1, 0, 40] = np.inf; x[2, 0, 5] = -np.inf; x[3, 0, 299] = np.nan; x[3, 0, 100] = np.inf
    add("nonfinite_th", x, 0.5)
    add("nonfinite_none", x, None)
    add("noise_th", rng.normal(size=(16, 1, 4096)), 1.0)
    add("noise_none", rng.normal(size=(16, 1, 4096)), None)
    add("noise_win4", rng.normal(size=(5, 1, 1000)), 0.3, win=4, r=1)
    add("noise_win1", rng.normal(size=(5, 1, 333)), 0.3, win=1, r=2)
    add("noise_win63", rng.normal(size=(5, 1, 777)), 0.1, win=63, r=10)
    add("short_row", rng.normal(size=(4, 1, 7)), 0.1)
    x = np.abs(rng.normal(size=(6, 1, 2048))); x[2] = 0; x[4] = -x[4]
    add("mixed_rows_th", x, 1.5)
    add("mixed_rows_none", x, None)
    # plateaus and quantised values: many exact ties
    add("quantised_th", np.round(rng.normal(size=(8, 1, 1500)) * 2) / 2, 0.5)
    add("quantised_none", np.round(rng.normal(size=(8, 1, 1500)) * 2) / 2, None)
    cases["names"] = np.array(names)
    np.savez_compressed(os.path.join(HERE, "peaks_cases.npz"), **cases)
    print("peaks cases:", len(names))


def gen_echo_cases():
    """mask2coords with echo_max (reference utils/mask2samples.py:103-111): fewer, equal and more echoes than K_max,
    rows of very different counts (zero pads compete with their amplitude scores[row][0]), r = 4 / 3 / 2.5."""
    rng = np.random.default_rng(33)
    cases, names = {}, []

    def add(name, scores, th, echo_max, win=20, r=4):
        s = torch.from_numpy(np.ascontiguousarray(scores, dtype=np.float32))
        co = mask2coords(s.clone(), win, th, upsample_factor=r, echo_max=echo_max)
        cases[name + "_scores"] = s.numpy()
        cases[name + "_th"] = np.float64(np.nan if th is None else th)
        cases[name + "_win"] = np.int64(win)
        cases[name + "_r"] = np.float64(r)
        cases[name + "_echo_max"] = np.int64(echo_max)
        cases[name + "_coords"] = co.numpy().astype(np.float32)
        names.append(name)

    x = rng.normal(size=(6, 1, 2000))
    kmax = mask2coords(torch.from_numpy(x.astype(np.float32)), 20, 1.0).shape[-1]
    for e in (1, 5, kmax - 1, kmax, kmax + 7):
        add("noise_e%d" % e, x, 1.0, e)
    add("noise_r3", x, 1.0, 9, r=3)
    add("noise_rfloat", x, 1.0, 9, r=2.5)
    add("none_pad", x, None, 3)
    y = np.abs(rng.normal(size=(5, 1, 1500)))
    y[1] *= 0.05                      # a row with almost no detections: its zero pads enter the selection ...
    y[1, 0, 0] = 9.0                  # ... and win it, because their amplitude is scores[row][0]
    y[3, 0, 40:1400] = 0.0
    add("ragged_e4", y, 1.2, 4)
    add("ragged_e12", y, 1.2, 12, win=8, r=1)
    # (equal amplitudes across the cut are left out: the reference's argsort is the unstable one, so which of two equal
    #  echoes survives is implementation-defined there; the oracle and the kernel keep table order)
    cases["names"] = np.array(names)
    np.savez_compressed(os.path.join(HERE, "echo_cases.npz"), **cases)
    print("echo cases:", len(names), "kmax", kmax)


def gen_metrics_cases():
    rng = np.random.default_rng(21)
    out = {}
    # toa_rmse (utils/metrics.py:9-41): etol compares with the SQUARED distance
    gt = rng.uniform(10, 500, size=(6, 4)).astype(np.float32)
    gt[1, 2:] = 0; gt[3, :] = 0; gt[4, 1] = np.nan
    es = gt + rng.normal(0, 0.8, size=gt.shape).astype(np.float32)
    es = np.concatenate([es, rng.uniform(10, 500, size=(6, 2)).astype(np.float32)], axis=1)
    es[2, :] = 0; es[5, 4] = np.inf
    for tol in (1.0, 0.25, 4.0):
        r = toa_rmse(torch.from_numpy(gt.copy()), torch.from_numpy(es.copy()), tol=tol)
        out["rmse_tol%s" % th_key(tol)] = r.numpy()
    out["rmse_gt"] = gt; out["rmse_es"] = es
    # coords2mask (utils/mask2samples.py:139-148)
    ref = torch.zeros((3, 1, 64))
    samples = torch.tensor([[[5, 9, 0]], [[-3, 63, 63]], [[0, 0, 0]]], dtype=torch.long)
    out["c2m_samples"] = samples.numpy().copy()
    out["c2m_mask"] = coords2mask(samples.clone(), ref).numpy()
    # find_threshold (utils/threshold.py:6-28): row 0 only
    pred = torch.from_numpy(rng.normal(size=(2, 1, 512)).astype(np.float32))
    true = torch.zeros((2, 1, 512)); true[0, 0, [40, 200, 333]] = 1; true[1, 0, [5]] = 1
    pred[0, 0, [40, 200, 333]] += 3
    out["ft_pred"] = pred.numpy(); out["ft_true"] = true.numpy()
    out["ft_th"] = np.float64(find_threshold(pred, true, 20))
    out["ft_th_norm"] = np.float64(find_threshold(pred, true, 20, norm_opt=True))
    # SampleShuffle1D (utils/sample_shuffle.py:10-28)
    x = torch.arange(2 * 8 * 5, dtype=torch.float32).reshape(2, 8, 5)
    out["shuffle_x"] = x.numpy()
    out["shuffle_r4"] = SampleShuffle1D(4)(x).numpy()
    out["shuffle_r2"] = SampleShuffle1D(2)(x).numpy()
    np.savez_compressed(os.path.join(HERE, "metrics_cases.npz"), **out)


def gen_aux_cases():
    """loss (main.py:228-232), find_threshold, rf pre-interpolation, AttentionBlock -> aux_cases.npz"""
    import types
    import torch.nn.functional as F
    from utils.gaussian import gaussian_kernel
    rng = np.random.default_rng(33)
    out = {}
    # ---- loss + d loss / d pred through the reference's own ops and autograd
    gk = torch.tensor(gaussian_kernel(size=7, sigma=1), dtype=torch.float32).unsqueeze(0).unsqueeze(0)
    out["gauss7"] = gk.numpy().reshape(-1)
    loss_cases = {
        "a": (rng.normal(size=(3, 1, 400)).astype(np.float32), [[37, 120, 0, 0], [5, 8, 399, 250], [-4, 0, 0, 0]], 1e-2),
        "b": (rng.normal(scale=3, size=(2, 1, 9000)).astype(np.float32), [[100, 100, 103], [8999, 1, 4500]], 1.0),
        "c": (np.zeros((2, 1, 64), dtype=np.float32), [[10], [20]], 1e-2),
    }
    for name, (pred, gt, lam) in loss_cases.items():
        mp = torch.from_numpy(pred.copy()).requires_grad_(True)
        gt_true = torch.tensor(gt, dtype=torch.long).unsqueeze(1)
        mt = coords2mask(gt_true.clone(), mp.detach())
        blur = F.conv1d(mt, gk, padding=3)
        blur /= blur.max()
        blur *= 20
        loss = torch.nn.MSELoss()(mp.squeeze(1), blur.squeeze(1).float()) + \
            torch.nn.L1Loss()(mp.squeeze(1), torch.zeros_like(mp.squeeze(1))) * lam
        loss.backward()
        out["loss_%s_pred" % name] = pred
        out["loss_%s_gt" % name] = gt_true.numpy()
        out["loss_%s_lambda" % name] = np.float64(lam)
        out["loss_%s_value" % name] = np.float64(loss.item())
        out["loss_%s_grad" % name] = mp.grad.numpy()
        out["loss_%s_blur" % name] = blur.numpy()
    out["loss_names"] = np.array(sorted(loss_cases))
    # ---- find_threshold (utils/threshold.py:6-28)
    ft = {}
    pred = rng.normal(size=(2, 1, 10240)).astype(np.float32)
    true = np.zeros((2, 1, 10240), dtype=np.float32); true[0, 0, [100, 5000, 9999]] = 1; pred[0, 0, [100, 5000]] += 4
    ft["long"] = (pred, true)
    pred = (np.round(rng.normal(size=(1, 1, 700)) * 2) / 2).astype(np.float32)
    true = np.zeros((1, 1, 700), dtype=np.float32); true[0, 0, ::50] = 1
    ft["ties"] = (pred, true)
    ft["nopos"] = (rng.normal(size=(2, 1, 300)).astype(np.float32), np.zeros((2, 1, 300), dtype=np.float32))
    pred = rng.normal(size=(1, 1, 5000)).astype(np.float32)
    true = np.zeros((1, 1, 5000), dtype=np.float32); true[0, 0, 77] = 20.0; pred[0, 0, 77] = 9
    ft["amp20"] = (pred, true)
    pred = np.abs(rng.normal(size=(1, 1, 4096))).astype(np.float32)
    true = (rng.uniform(size=(1, 1, 4096)) < 0.3).astype(np.float32); pred += true
    ft["dense"] = (pred, true)
    import warnings
    for name, (pred, true) in ft.items():
        out["ft_%s_pred" % name] = pred
        out["ft_%s_true" % name] = true
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            out["ft_%s_th" % name] = np.float64(find_threshold(torch.from_numpy(pred), torch.from_numpy(true), 20))
            out["ft_%s_th_norm" % name] = np.float64(find_threshold(torch.from_numpy(pred), torch.from_numpy(true), 20, norm_opt=True))
    out["ft_names"] = np.array(sorted(ft))
    # ---- rf_scale_factor pre-interpolation: import datasets/chirp_dataset.py unmodified; natsort / omegaconf are not
    # installed here and are only used by the file loader, so empty stand-in modules satisfy its import lines
    for modname, attr in (("natsort", "natsorted"), ("omegaconf", "OmegaConf")):
        if modname not in sys.modules:
            mod = types.ModuleType(modname)
            setattr(mod, attr, None)
            sys.modules[modname] = mod
    import importlib.util
    spec = importlib.util.spec_from_file_location("ref_chirp_dataset", os.path.join(REF, "datasets", "chirp_dataset.py"))
    cd = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(cd)                      # the reference file, loaded by path (no package __init__ there)
    upscale_1d, ChirpDataset = cd.upscale_1d, cd.ChirpDataset
    env = np.abs(rng.normal(size=(3, 257))).astype(np.float32)
    iq = (rng.normal(size=(3, 200)) + 1j * rng.normal(size=(3, 200))).astype(np.complex64)
    out["up_env"] = env
    out["up_iq_re"] = iq.real.copy(); out["up_iq_im"] = iq.imag.copy()
    for rf in (10, 20, 2.5):
        out["up_env_rf%s" % th_key(rf)] = np.stack([upscale_1d(e.astype(np.float64), rf) for e in env]).astype(np.float32)
    fs, fc = 20e6, 5e6
    out["up_fs"] = np.float64(fs); out["up_fc"] = np.float64(fc)
    for rf in (10, 20):
        out["up_iq_rf%s" % th_key(rf)] = np.stack([ChirpDataset.iq2rf(z.astype(np.complex128), fc=fc, fs=fs, rescale_factor=rf)
                                                   for z in iq]).astype(np.float32)
    # ---- AttentionBlock (utils/attention_block.py; dead code: F is never imported -> inject it, dropout forced to 0)
    import utils.attention_block as ab

    class _F:
        @staticmethod
        def scaled_dot_product_attention(q, k, v, dropout_p=0.0):
            return F.scaled_dot_product_attention(q, k, v, dropout_p=0.0)
    ab.F = _F
    torch.manual_seed(5)
    blk = ab.AttentionBlock(48, 40).eval()
    x = torch.randn(2, 70, 48)
    with torch.no_grad():
        y = blk(x)
    out["att_x"] = x.numpy(); out["att_y"] = y.numpy()
    for nm in ("query", "key", "value"):
        out["att_%s_w" % nm] = getattr(blk, nm).weight.detach().numpy()
        out["att_%s_b" % nm] = getattr(blk, nm).bias.detach().numpy()
    np.savez_compressed(os.path.join(HERE, "aux_cases.npz"), **out)
    print("aux cases done")


def gen_train_cases():
    """Training step of main.py:217-248 run with the reference model, torch autograd and torch.optim.AdamW:
    gradients of every parameter after the first backward, and losses / parameters over three steps."""
    import torch.nn.functional as F
    from utils.gaussian import gaussian_kernel
    gk = torch.tensor(gaussian_kernel(size=7, sigma=1), dtype=torch.float32).unsqueeze(0).unsqueeze(0)
    for name, L, lam in (("different-armadillo", 338, 1e-2), ("clean-serenity", 320, 1.0)):
        sd = load_sd(name)
        has_sgb = any(k.startswith("semi_global_block") for k in sd)
        torch.manual_seed(1)
        model = StofNet(upsample_factor=4, semi_global_scale=80 if has_sgb else 1)
        model.load_state_dict(sd)
        model.train()
        opt = torch.optim.AdamW(model.parameters(), lr=5e-4, weight_decay=1e-8)
        N = 3
        x = torch.from_numpy(synth_waveforms(N, L, seed=77, rf=10)).unsqueeze(1)
        gt = torch.tensor([[[4 * 100 + 1, 4 * 200, 0]], [[4 * 50, 0, 0]], [[4 * 300 + 2, 4 * 310, 4 * 20]]], dtype=torch.long)
        out = {"x": x.squeeze(1).numpy(), "gt": gt.numpy(), "lambda": np.float64(lam), "L": np.int64(L)}
        losses = []
        for step in range(3):
            masks_pred = model(x)
            mt = coords2mask(gt.clone(), masks_pred)
            blur = F.conv1d(mt, gk, padding=3)
            blur /= blur.max()
            blur *= 20
            loss = torch.nn.MSELoss()(masks_pred.squeeze(1), blur.squeeze(1).float()) + \
                torch.nn.L1Loss()(masks_pred.squeeze(1), torch.zeros_like(masks_pred.squeeze(1))) * lam
            opt.zero_grad()
            loss.backward()
            if step == 0:
                out["mask0"] = masks_pred.detach().squeeze(1).numpy()
                for k, p in model.named_parameters():
                    out["grad0/" + k] = p.grad.detach().numpy().copy()
            opt.step()
            losses.append(loss.item())
            if step == 2:
                for k, p in model.named_parameters():
                    out["param3/" + k] = p.detach().numpy().copy()
        out["losses"] = np.array(losses, dtype=np.float64)
        out["names"] = np.array([k for k, _ in model.named_parameters()])
        np.savez_compressed(os.path.join(HERE, "train_%s.npz" % name), **out)
        print("train case", name, losses)



def sha(a):
    import hashlib
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def gen_scale():
    """SURVEY 8c golden set at N = 64 x 3 seeds x 4 lengths per StofNet checkpoint (hashes + index lists, no tensors)."""
    for name, (rf, th) in STOF.items():
        sd = load_sd(name)
        has_sgb = any(k.startswith("semi_global_block") for k in sd)
        model = StofNet(upsample_factor=4, semi_global_scale=80 if has_sgb else 1).eval()
        model.load_state_dict(sd)
        out = {"rf": np.int64(rf), "th": np.float64(np.nan if th is None else th)}
        cases = [(64, L, seed) for seed in (3008, 0, 1) for L in (2560, 2578, 5120, 160)]
        out["cases"] = np.array(cases, dtype=np.int64)
        for ci, (N, L, seed) in enumerate(cases):
            xn = synth_waveforms(N, L, seed=seed, rf=rf, q15=True)
            x = torch.from_numpy(xn).unsqueeze(1)
            with torch.no_grad():
                m = model(x)
            p = "c%d" % ci
            out[p + "_x_sha"] = np.array(sha(xn))
            out[p + "_mask_sha"] = np.array(sha(m.numpy()))
            out[p + "_rowmax"] = m.squeeze(1).max(dim=1)[0].numpy()
            out[p + "_rowabsmax"] = m.squeeze(1).abs().max(dim=1)[0].numpy()
            for t in sorted({None, th, 0.5}, key=lambda v: (v is not None, v or 0)):
                peaks_dict(p, m, t, out)
            print("  scale", name, ci, N, L, seed, "K(th=0.5) =", out[p + "_idx_0p5"].shape[0])
        np.savez_compressed(os.path.join(HERE, "scale_%s.npz" % name), **out)


def gen_surface_cases():
    """Standalone operators of the reference's surface: thresholding (:14-23), reduce_echoes (:117-132), get_amplitudes
    (:135-136) of utils/mask2samples.py and SemiGlobalBlock.forward (models/stofnet.py:98-117)."""
    from utils.mask2samples import thresholding, reduce_echoes, get_amplitudes
    rng = np.random.default_rng(23)
    out = {}
    s = rng.normal(size=(5, 1, 333)).astype(np.float32)
    s[1, 0, 7] = np.nan
    s[3, 0, :] = -1.25
    for key, th in (("th_float", 0.4), ("th_none", None), ("th_zero", 0.0), ("th_neg", -0.5)):
        t = torch.from_numpy(s.copy())
        r = thresholding(t, th)
        out[key + "_in"] = s
        out[key + "_th"] = np.float64(np.nan if th is None else th)
        out[key + "_out"] = r.numpy()
        out[key + "_inplace"] = np.int64(r.data_ptr() == t.data_ptr())
    # reduce_echoes: distinct amplitudes (torch.argsort's tie order is unspecified), K > echo_max and K <= echo_max
    K = 37
    pos = np.sort(np.stack([rng.choice(4000, size=K, replace=False) for _ in range(6)]).astype(np.float32), axis=1)   # distinct: tie order is unspecified in the reference
    amp = rng.permutation(6 * K).reshape(6, K).astype(np.float32) / 7.0
    pairs = np.dstack([pos, amp]).astype(np.float32)
    out["re_in"] = pairs
    for e in (5, 36):       # the reference raises UnboundLocalError for K <= echo_max (`echoes` is only bound inside the if, :123-130)
        out["re_out_%d" % e] = reduce_echoes(torch.from_numpy(pairs.copy()), echo_max=e).numpy()
    triple = np.dstack([pos, amp, rng.normal(size=(6, K)).astype(np.float32)]).astype(np.float32)
    out["re3_in"] = triple
    out["re3_out_9"] = reduce_echoes(torch.from_numpy(triple.copy()), echo_max=9).numpy()
    # get_amplitudes: half-way positions round to even
    frames = rng.normal(size=(4, 1, 500)).astype(np.float32)
    smp = rng.uniform(0, 499, size=(4, 21)).astype(np.float32)
    smp[:, :4] = np.array([0.5, 1.5, 2.5, 498.5], np.float32)
    out["ga_frames"], out["ga_samples"] = frames, smp
    out["ga_out"] = get_amplitudes(torch.from_numpy(frames), torch.from_numpy(smp)).numpy()
    # SemiGlobalBlock.forward standalone, weights of different-armadillo, L with and without the padded branch
    sd = load_sd("different-armadillo")
    model = StofNet(upsample_factor=4, semi_global_scale=80).eval()
    model.load_state_dict(sd)
    for key, (N, L) in (("sgb_a", (2, 400)), ("sgb_b", (1, 2578)), ("sgb_c", (1, 80))):
        # inputs on a coarse grid that float16 holds exactly (half the fixture bytes)
        x = torch.from_numpy((rng.integers(-2048, 2049, size=(N, 64, L)) / 512.0).astype(np.float32))
        with torch.no_grad():
            y = model.semi_global_block(x)
        out[key + "_x"], out[key + "_y"] = x.numpy().astype(np.float16), y.numpy()
    np.savez_compressed(os.path.join(HERE, "surface_cases.npz"), **out)
    print("surface cases done")

def gen_train_b4():
    """The reference's own training configuration (config.yaml:11 batch_size 4, rf 10: L = 2560) for one step of
    main.py:217-248: loss, mask and every parameter's gradient from the reference model and torch autograd."""
    import torch.nn.functional as F
    from utils.gaussian import gaussian_kernel
    gk = torch.tensor(gaussian_kernel(size=7, sigma=1), dtype=torch.float32).unsqueeze(0).unsqueeze(0)
    name, N, L, lam = "different-armadillo", 4, 2560, 1e-2
    sd = load_sd(name)
    torch.manual_seed(1)
    model = StofNet(upsample_factor=4, semi_global_scale=80)
    model.load_state_dict(sd)
    model.train()
    x = torch.from_numpy(synth_waveforms(N, L, seed=91, rf=10)).unsqueeze(1)
    rng = np.random.default_rng(17)
    gt = torch.from_numpy(np.sort(rng.integers(40, 4 * L - 40, size=(N, 1, 3)), axis=-1))
    gt[1, 0, 2] = 0
    gt[3, 0, 1:] = 0                                              # absent echoes, as the batches pad them
    masks_pred = model(x)
    mt = coords2mask(gt.clone(), masks_pred)
    blur = F.conv1d(mt, gk, padding=3)
    blur /= blur.max()
    blur *= 20
    loss = torch.nn.MSELoss()(masks_pred.squeeze(1), blur.squeeze(1).float()) + \
        torch.nn.L1Loss()(masks_pred.squeeze(1), torch.zeros_like(masks_pred.squeeze(1))) * lam
    loss.backward()
    out = {"x": x.squeeze(1).numpy(), "gt": gt.numpy(), "lambda": np.float64(lam), "loss": np.float64(loss.item()),
           "mask0": masks_pred.detach().squeeze(1).numpy(), "names": np.array([k for k, _ in model.named_parameters()])}
    for k, p in model.named_parameters():
        out["grad0/" + k] = p.grad.detach().numpy().copy()
    np.savez_compressed(os.path.join(HERE, "train_b4_%s.npz" % name), **out)
    print("train b4 case", name, loss.item())


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "train_b4":
        gen_train_b4()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "scale":
        gen_scale()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "surface":
        gen_surface_cases()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "train":
        gen_train_cases()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "edsr":
        gen_edsr()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "echo":
        gen_echo_cases()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "aux":
        gen_aux_cases()
        sys.exit(0)
    gen_forward()
    gen_edsr()
    gen_peaks_cases()
    gen_echo_cases()
    gen_metrics_cases()
    gen_aux_cases()
    gen_train_cases()
    gen_train_b4()
    gen_scale()
    gen_surface_cases()
    print("done")
