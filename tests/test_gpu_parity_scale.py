"""Index-exactness at the sizes the bench runs (SURVEY.md 8c parity protocol; VERDICT round 1, task 1).

Part A - the committed N = 64 x 3 seeds x 4 lengths golden set (tests/golden/scale_<ckpt>.npz, made by the UNMODIFIED
reference in the build container): inputs are regenerated from the host-independent generator and checked by SHA-256,
index lists and ToF tables must equal the reference's.

Part B - every StofNet checkpoint at its BASELINE shape, 256 rows x seeds {3008, 0, 1}, through model(...) +
get_maxima_positions / mask2coords on the GPU and the same rows through oracle/torch_port (the reference's call sequence
on this box's CPU): C2 graceful-snow L = 5120 th = 0.015; C3 different-armadillo L = 2560 on PALA-shaped frames with
th = None and 0.5; the two checkpoints without semi-global block at their array_pala_params thresholds.

Bar: identical index lists.  A differing (row, position) is tolerated only when the fp64 forward shows the decision
(|score - th|, |score - window max|, |score - row max|) to lie within the reference's own fp32 rounding noise
(tests/parity_util.py); the count is printed.  Masks: max |err| <= 1e-4 of the mask range.
"""
import hashlib

import numpy as np
import pytest
import torch

from conftest import STOF_CKPTS, th_key, th_list
from parity_util import assert_index_parity, port_masks, to_sd

pytestmark = pytest.mark.gpu

FWD_TOL = 1e-4
SEEDS = (3008, 0, 1)
# checkpoint -> (rf, L, thresholds, input kind) at its BASELINE shape (bash_scripts/array_*_params.txt of the reference)
SHAPES = {
    "graceful-snow": (20, 5120, (0.015,), "chirp"),            # C2
    "different-armadillo": (10, 2560, (None, 0.5), "pala"),    # C3
    "clean-serenity": (10, 2560, (None,), "chirp"),
    "breezy-violet": (20, 5120, (2.3,), "chirp"),
}
ROWS = 256


@pytest.fixture(scope="module")
def sb():
    import stofnet_b200
    assert torch.cuda.is_available()
    return stofnet_b200


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize("name", STOF_CKPTS)
def test_scale_golden_set(sb, dev, golden, name):
    """Part A: 64 rows x 3 seeds x L in {2560, 2578, 5120, 160} against the reference's committed index lists."""
    from stofnet_b200.synth import synth_waveforms
    g = golden("scale_%s.npz" % name)
    wsd = golden("weights_%s.npz" % name)
    sd = to_sd(wsd)
    model = sb.build_model(sd, dev)
    rf = int(g["rf"])
    flips = 0
    for ci, (N, L, seed) in enumerate(g["cases"]):
        x = synth_waveforms(int(N), int(L), seed=int(seed), rf=rf, q15=True)
        assert _sha(x) == str(g["c%d_x_sha" % ci]), "input generator is not host independent (case %d)" % ci
        with torch.no_grad():
            m = model(torch.from_numpy(x).unsqueeze(1).to(dev))
        got = m.squeeze(1).cpu().numpy()
        scale = float(g["c%d_rowabsmax" % ci].max())
        assert np.abs(got.max(axis=1) - g["c%d_rowmax" % ci]).max() <= FWD_TOL * scale, (name, ci)
        for th in th_list(g):
            ref_idx = g["c%d_idx_%s" % (ci, th_key(th))]
            idx = sb.get_maxima_positions(m, 20, th).cpu().numpy()
            n = 0
            if not np.array_equal(idx, ref_idx):
                n = assert_index_parity(sd, x, idx, ref_idx, th, what="%s case %d th=%r" % (name, ci, th))
            flips += n
            if n == 0:
                co = sb.mask2coords(m, 20, th, upsample_factor=4).cpu().numpy()
                ref_co = g["c%d_coords_%s" % (ci, th_key(th))]
                assert co.shape == ref_co.shape and np.array_equal(co, ref_co), (name, ci, th)
    print("scale golden set %s: %d cases, %d boundary flips" % (name, len(g["cases"]), flips))


@pytest.mark.parametrize("name", STOF_CKPTS)
def test_baseline_shape_vs_port(sb, dev, golden, name):
    """Part B: 256 rows x 3 seeds at the checkpoint's BASELINE shape, GPU vs the reference port on this box's CPU."""
    from oracle import torch_port
    from stofnet_b200.synth import synth_waveforms, pala_shaped
    rf, L, ths, kind = SHAPES[name]
    sd = to_sd(golden("weights_%s.npz" % name))
    model = sb.build_model(sd, dev)
    total_k, flips, worst = 0, 0, 0.0
    for seed in SEEDS:
        if kind == "pala":
            x = pala_shaped(128, ROWS // 128, L, seed=seed, rf=rf, q15=True)     # 2 frames x 128 channels folded into rows
        else:
            x = synth_waveforms(ROWS, L, seed=seed, rf=rf, q15=True)
        xt = torch.from_numpy(x).unsqueeze(1)
        ref = port_masks(sd, x)
        with torch.no_grad():
            m = model(xt.to(dev))
        got = m.squeeze(1).cpu().numpy()
        err = float(np.abs(got - ref).max() / np.abs(ref).max())
        worst = max(worst, err)
        assert err <= FWD_TOL, (name, seed, err)
        ref_t = torch.from_numpy(ref).unsqueeze(1)
        for th in ths:
            ref_idx = torch_port.get_maxima_positions(ref_t, 20, th).numpy()
            idx = sb.get_maxima_positions(m, 20, th).cpu().numpy()
            total_k += ref_idx.shape[0]
            n = 0
            if not np.array_equal(idx, ref_idx):
                n = assert_index_parity(sd, x, idx, ref_idx, th, ref_masks32=ref,
                                        what="%s seed %d th=%r" % (name, seed, th))
            flips += n
            if n == 0:
                co = sb.mask2coords(m, 20, th, upsample_factor=4).cpu().numpy()
                ref_co = torch_port.mask2coords(ref_t, 20, th, upsample_factor=4).numpy()
                assert co.shape == ref_co.shape and np.array_equal(co, ref_co), (name, seed, th)
    print("baseline shape %s: %d rows x L=%d, %d detections compared, %d boundary flips, max mask err %.2e of range"
          % (name, ROWS * len(SEEDS), L, total_k, flips, worst))


def test_c2_bench_batch_slice(sb, dev, golden):
    """The bench's own batch construction (256 base rows tiled to 4096 with per-row roll + gain): a 192-row slice spread
    over the batch, GPU (processed inside the full 4096-row forward, i.e. in their real chunks) vs the port."""
    import importlib.util
    import os
    from oracle import torch_port
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("bench", os.path.join(root, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    sd = to_sd(golden("weights_graceful-snow.npz"))
    model = sb.build_model(sd, dev)
    base = bench.make_inputs(256, 3008)
    xb = bench.expand_batch(base, bench.ROWS, 7)
    with torch.no_grad():
        m = model(xb.unsqueeze(1).to(dev))
    pick = np.r_[0:64, 2000:2064, 4032:4096]
    x = xb[pick].numpy()
    ref = port_masks(sd, x)
    got = m.squeeze(1)[torch.from_numpy(pick).to(dev)].contiguous()
    assert float(np.abs(got.cpu().numpy() - ref).max() / np.abs(ref).max()) <= FWD_TOL
    ref_idx = torch_port.get_maxima_positions(torch.from_numpy(ref).unsqueeze(1), 20, bench.TH).numpy()
    idx = sb.get_maxima_positions(got.unsqueeze(1), 20, bench.TH).cpu().numpy()
    if not np.array_equal(idx, ref_idx):
        assert_index_parity(sd, x, idx, ref_idx, bench.TH, ref_masks32=ref, what="bench batch slice")
    co = sb.mask2coords(m, 20, bench.TH, upsample_factor=4)
    print("bench batch: K_max = %d over %d rows (bench.py widens its table from CAP=%d when needed)" % (co.shape[1], bench.ROWS, bench.CAP))


@pytest.mark.parametrize("name,th", [("wobbly-sponge", 0.2), ("vital-puddle", 0.2)])
def test_espcn_baseline_shape_vs_port(sb, dev, golden, name, th):
    """ESPCN_1D (the checkpoint BASELINE config 2 names is one) on the tensor-core kernel at the C2 shape, 256 rows x seeds
    {3008, 0, 1}, against the reference's call sequence on this box's CPU (oracle/torch_port.espcn_forward): masks within 2e-5
    (sigmoid outputs); index lists identical, or the differing sample within fp32 rounding of a decision boundary when judged by
    the fp64 re-run (tolerance 4 x the port's own fp32-vs-fp64 distance on that row, as tests/parity_util.py)."""
    from oracle import torch_port
    from stofnet_b200.synth import synth_waveforms
    sd = to_sd(golden("weights_%s.npz" % name))
    sd64 = {k: v.double() for k, v in sd.items()}
    model = sb.build_model(sd, dev)
    total = flips = 0
    for seed in SEEDS:
        x = synth_waveforms(ROWS, 5120, seed=seed, rf=20, q15=True)
        xt = torch.from_numpy(x).unsqueeze(1)
        with torch.no_grad():
            got = model(xt.to(dev))
            ref = torch_port.espcn_forward(sd, xt)
        assert float((got.cpu() - ref).abs().max()) <= 2e-5
        idx = sb.get_maxima_positions(got, 20, th).cpu().numpy()
        ref_idx = torch_port.get_maxima_positions(ref, 20, th).numpy()
        total += len(ref_idx)
        diff = sorted(set(map(tuple, idx.tolist())) ^ set(map(tuple, ref_idx.tolist())))
        if diff:
            rows = sorted({r for r, _ in diff})
            with torch.no_grad():
                m64 = torch_port.espcn_forward(sd64, xt[rows].double()).squeeze(1).numpy()
            m32 = ref.squeeze(1).numpy()[rows]
            for r, pos in diff:
                k = rows.index(r)
                tol = max(4.0 * float(np.abs(m32[k].astype(np.float64) - m64[k]).max()), 8 * np.finfo(np.float32).eps)
                lo, hi = max(0, pos - 10), min(m64.shape[1], pos + 11)
                others = np.delete(m64[k][lo:hi], pos - lo)
                margin = min(abs(m64[k][pos] - th), abs(m64[k][pos] - others.max()) if others.size else np.inf)
                assert margin <= tol, (name, seed, r, pos, float(margin), tol)
            flips += len(diff)
    print("espcn %s: %d detections over %d rows, %d boundary flips" % (name, total, ROWS * len(SEEDS), flips))
