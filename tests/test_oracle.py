"""Pin the CPU oracle (oracle/) against the reference's own outputs (tests/golden/*.npz)."""
import numpy as np
import pytest
import torch

from conftest import STOF_CKPTS, ESPCN_CKPTS, EDSR_CKPTS, th_key, th_list
from oracle import c_oracle, torch_port


@pytest.mark.parametrize("name", STOF_CKPTS)
def test_c_forward_matches_reference(golden, name):
    sd = golden("weights_%s.npz" % name)
    g = golden("forward_%s.npz" % name)
    W = c_oracle.pack_weights(sd)
    for ci, (N, L, seed) in enumerate(g["cases"]):
        if L > 3000 and ci != 2:
            continue
        x = g["c%d_x" % ci]
        m = c_oracle.forward(W, x, acc64=True)
        ref64 = g["c%d_mask64" % ci]
        ref32 = g["c%d_mask" % ci]
        scale = np.abs(ref64).max()
        # double-accumulating C restatement vs the reference run in fp64: fp32 storage noise only
        assert np.abs(m - ref64).max() <= 2e-5 * scale
        assert np.abs(m - ref32).max() <= 1e-5 * scale + np.abs(ref32 - ref64).max() * 2
        # the detections the reference makes on its own mask are reproduced on the oracle's mask
        for th in th_list(g):
            idx, _ = c_oracle.get_maxima_positions(m, 20, th)
            assert np.array_equal(idx, g["c%d_idx_%s" % (ci, th_key(th))]), (name, ci, th)


@pytest.mark.parametrize("name", STOF_CKPTS)
def test_c_forward_fp32_accumulate(golden, name):
    sd = golden("weights_%s.npz" % name)
    g = golden("forward_%s.npz" % name)
    x = g["c0_x"]
    m = c_oracle.forward(sd, x, acc64=False)
    ref = g["c0_mask"]
    assert np.abs(m - ref).max() <= 2e-5 * np.abs(ref).max()


@pytest.mark.parametrize("name", STOF_CKPTS)
def test_torch_port_is_bitwise_reference(golden, name):
    sd = torch_port.to_torch_sd(golden("weights_%s.npz" % name))
    g = golden("forward_%s.npz" % name)
    for ci, (N, L, seed) in enumerate(g["cases"]):
        x = torch.from_numpy(g["c%d_x" % ci]).unsqueeze(1)
        with torch.no_grad():
            m = torch_port.stofnet_forward(sd, x)
        ref = torch.from_numpy(g["c%d_mask" % ci])
        # same ATen kernels in the same order; thread-count dependent blocking may differ in the last bits
        assert (m.squeeze(1) - ref).abs().max() <= 2e-5 * ref.abs().max()
        for th in th_list(g):
            co = torch_port.mask2coords(ref.unsqueeze(1), 20, th, 4)
            assert np.array_equal(co.numpy(), g["c%d_coords_%s" % (ci, th_key(th))])
            idx = torch_port.get_maxima_positions(ref.unsqueeze(1), 20, th)
            assert np.array_equal(idx.numpy(), g["c%d_idx_%s" % (ci, th_key(th))])


@pytest.mark.parametrize("name", ESPCN_CKPTS)
def test_espcn_oracles(golden, name):
    sd = golden("weights_%s.npz" % name)
    g = golden("espcn_%s.npz" % name)
    tsd = torch_port.to_torch_sd(sd)
    for ci, (N, L, seed) in enumerate(g["cases"]):
        x = g["c%d_x" % ci]
        ref = g["c%d_mask" % ci]
        m = c_oracle.espcn_forward(sd, x)
        assert np.abs(m - ref).max() <= 5e-6      # sigmoid output in (0,1); tanhf/expf libm vs ATen vectorised
        with torch.no_grad():
            mt = torch_port.espcn_forward(tsd, torch.from_numpy(x).unsqueeze(1)).squeeze(1).numpy()
        assert np.abs(mt - ref).max() <= 1e-6


@pytest.mark.parametrize("name", EDSR_CKPTS)
def test_edsr_oracles(golden, name):
    """EDSR_1D restatements against the reference's own outputs (tests/golden/make_golden.py edsr)."""
    sd = golden("weights_%s.npz" % name)
    g = golden("edsr_%s.npz" % name)
    tsd = torch_port.to_torch_sd(sd)
    for ci, (N, L, seed) in enumerate(g["cases"]):
        x = g["c%d_x" % ci]
        ref, ref64 = g["c%d_mask" % ci], g["c%d_mask64" % ci]
        with torch.no_grad():
            mt = torch_port.edsr_forward(tsd, torch.from_numpy(x).unsqueeze(1)).squeeze(1).numpy()
        assert np.array_equal(mt, ref)            # same ATen kernels, same call sequence
        if L <= 640:                              # the C composition loops rows / channels in Python-free C, small cases only
            m = c_oracle.edsr_forward(sd, x, acc64=True)
            assert np.abs(m - ref64).max() <= 2e-6 * max(1.0, np.abs(ref64).max())


def test_stage4_bit_exact_on_reference_masks(golden):
    """Stage 4 involves no rounding: oracle indices/coords must equal the reference's exactly."""
    for name in STOF_CKPTS:
        g = golden("forward_%s.npz" % name)
        for ci in range(len(g["cases"])):
            ref = g["c%d_mask" % ci]
            for th in th_list(g):
                idx, counts = c_oracle.get_maxima_positions(ref, 20, th)
                assert np.array_equal(idx, g["c%d_idx_%s" % (ci, th_key(th))])
                co = c_oracle.mask2coords(ref[:, None, :], 20, th, 4)
                gc = g["c%d_coords_%s" % (ci, th_key(th))]
                assert co.shape == gc.shape and np.array_equal(co, gc)


def test_stage4_edge_cases(golden):
    g = golden("peaks_cases.npz")
    for name in g["names"]:
        s = g[name + "_scores"]
        th = float(g[name + "_th"])
        th = None if np.isnan(th) else th
        win, r = int(g[name + "_win"]), int(g[name + "_r"])
        nms = c_oracle.nms_1d(s, win)
        assert np.array_equal(nms, g[name + "_nms"], equal_nan=True), name
        idx, _ = c_oracle.get_maxima_positions(s, win, th)
        assert np.array_equal(idx, g[name + "_idx"]), name
        co = c_oracle.mask2coords(s, win, th, r)
        assert co.shape == g[name + "_coords"].shape, name
        assert np.array_equal(co, g[name + "_coords"]), name
        ts = torch.from_numpy(s)
        assert np.array_equal(torch_port.mask2coords(ts, win, th, r).numpy(), g[name + "_coords"]), name


def test_echo_max_cases(golden):
    """mask2coords(echo_max=...) of the reference (utils/mask2samples.py:103-132) against both oracle restatements."""
    g = golden("echo_cases.npz")
    for name in g["names"]:
        s = g[name + "_scores"]
        th = float(g[name + "_th"])
        th = None if np.isnan(th) else th
        win, r, e = int(g[name + "_win"]), float(g[name + "_r"]), int(g[name + "_echo_max"])
        r = int(r) if r.is_integer() else r
        want = g[name + "_coords"]
        got = torch_port.mask2coords(torch.from_numpy(s), win, th, r, echo_max=e).numpy()
        assert got.shape == want.shape and np.array_equal(got, want), name
        got = c_oracle.mask2coords_echo_max(s, win, th, r, e)
        assert got.shape == want.shape and np.array_equal(got, want), name


def test_bad_lengths_raise(golden):
    sd = golden("weights_different-armadillo.npz")
    for L in (79, 161, 1001):
        with pytest.raises(RuntimeError):
            c_oracle.forward(sd, np.zeros((1, L), np.float32))
    # without the semi-global block any length works (reference: no pooling constraint)
    sd2 = golden("weights_clean-serenity.npz")
    assert c_oracle.forward(sd2, np.zeros((1, 37), np.float32)).shape == (1, 148)


def test_aux_ops(golden):
    g = golden("metrics_cases.npz")
    assert np.array_equal(c_oracle.sample_shuffle(g["shuffle_x"], 4), g["shuffle_r4"])
    assert np.array_equal(c_oracle.sample_shuffle(g["shuffle_x"], 2), g["shuffle_r2"])
    assert np.array_equal(c_oracle.coords2mask(g["c2m_samples"], 64), g["c2m_mask"])
    for tol in (1.0, 0.25, 4.0):
        ref = g["rmse_tol%s" % th_key(tol)]
        out = c_oracle.toa_rmse(g["rmse_gt"], g["rmse_es"], tol)
        assert np.allclose(out, ref, rtol=1e-5, atol=1e-6, equal_nan=True), tol


@pytest.mark.parametrize("name", STOF_CKPTS)
def test_scale_golden_set_port(golden, name):
    """The N = 64 x 3 seeds golden set: the host-independent generator reproduces the inputs (SHA-256), and the
    port reproduces the reference's index lists on a subset small enough for the CPU suite (L = 160 and one L = 2560 case)."""
    import hashlib
    from stofnet_b200.synth import synth_waveforms
    g = golden("scale_%s.npz" % name)
    sd = torch_port.to_torch_sd(golden("weights_%s.npz" % name))
    rf = int(g["rf"])
    assert len(g["cases"]) == 12 and all(int(c[0]) == 64 for c in g["cases"])
    for ci, (N, L, seed) in enumerate(g["cases"]):
        if not (L == 160 or (L == 2560 and seed == 3008)):
            continue
        x = synth_waveforms(int(N), int(L), seed=int(seed), rf=rf, q15=True)
        assert hashlib.sha256(x.tobytes()).hexdigest() == str(g["c%d_x_sha" % ci])
        with torch.no_grad():
            m = torch_port.stofnet_forward(sd, torch.from_numpy(x).unsqueeze(1))
        assert np.abs(m.squeeze(1).max(dim=1)[0].numpy() - g["c%d_rowmax" % ci]).max() <= 2e-5 * g["c%d_rowabsmax" % ci].max()
        for th in th_list(g):
            idx = torch_port.get_maxima_positions(m, 20, th).numpy()
            ref = g["c%d_idx_%s" % (ci, th_key(th))]
            if not np.array_equal(idx, ref):        # thread-count dependent last bits: only boundary flips are tolerated
                from parity_util import assert_index_parity
                assert_index_parity(sd, x, idx, ref, th, what="%s case %d" % (name, ci))


def test_surface_ops_port(golden):
    """thresholding / reduce_echoes / get_amplitudes / SemiGlobalBlock restatements against the reference's outputs."""
    g = golden("surface_cases.npz")
    for key in ("th_float", "th_none", "th_zero", "th_neg"):
        th = float(g[key + "_th"]); th = None if np.isnan(th) else th
        t = torch.from_numpy(g[key + "_in"].copy())
        r = torch_port.thresholding(t, th)
        assert np.array_equal(r.numpy(), g[key + "_out"], equal_nan=True), key
        assert int(r.data_ptr() == t.data_ptr()) == int(g[key + "_inplace"]), key
    for e in (5, 36):
        r = torch_port.reduce_echoes(torch.from_numpy(g["re_in"].copy()), e)
        assert np.array_equal(r.numpy(), g["re_out_%d" % e])
    assert np.array_equal(torch_port.reduce_echoes(torch.from_numpy(g["re3_in"].copy()), 9).numpy(), g["re3_out_9"])
    a = torch_port.get_amplitudes(torch.from_numpy(g["ga_frames"]), torch.from_numpy(g["ga_samples"]))
    assert np.array_equal(a.numpy(), g["ga_out"])
    sd = torch_port.to_torch_sd(golden("weights_different-armadillo.npz"))
    for key in ("sgb_a", "sgb_b", "sgb_c"):
        x = torch.from_numpy(g[key + "_x"].astype(np.float32))
        with torch.no_grad():
            y = torch_port.semi_global_block(sd, x)
        assert np.abs(y.numpy() - g[key + "_y"]).max() <= 2e-5 * np.abs(g[key + "_y"]).max(), key


def test_port_training_step_matches_reference_autograd(golden):
    """The CPU baseline of the training step (bench.py cpu_train_rate: oracle/torch_port.stofnet_forward + torch autograd) against
    the reference model's own autograd at the reference's batch shape (train_b4 golden: 4 x 2560): same ATen kernels in the same
    order, so loss, mask and every gradient agree to fp32 noise."""
    import torch.nn.functional as F
    g = golden("train_b4_different-armadillo.npz")
    sd = golden("weights_different-armadillo.npz")
    torch.manual_seed(0)
    params = {k: torch.from_numpy(np.ascontiguousarray(v)).clone().requires_grad_(True) for k, v in sd.items()}
    x = torch.from_numpy(g["x"]).unsqueeze(1)
    gt = torch.from_numpy(g["gt"]).clone()
    pred = torch_port.stofnet_forward(params, x)
    assert np.abs(pred.detach().squeeze(1).numpy() - g["mask0"]).max() <= 1e-6 * np.abs(g["mask0"]).max()
    mask = torch.zeros_like(pred)
    gt[gt < 0] = 0
    mask.scatter_(2, gt, 1)
    mask[..., :1] = 0                                               # utils/mask2samples.py:139-148
    gk = torch.tensor([0.00443185, 0.05399097, 0.24197072, 0.39894228, 0.24197072, 0.05399097, 0.00443185]).view(1, 1, -1)
    blur = F.conv1d(mask, gk, padding=3)
    blur = blur / blur.max() * 20
    loss = F.mse_loss(pred.squeeze(1), blur.squeeze(1)) + float(g["lambda"]) * pred.abs().mean()
    assert abs(float(loss.detach()) - float(g["loss"])) <= 1e-5 * abs(float(g["loss"]))
    loss.backward()
    for k in g["names"]:
        k = str(k)
        ref = g["grad0/" + k]
        err = np.abs(params[k].grad.numpy() - ref).max()
        assert err <= 2e-5 * np.abs(ref).max() + 1e-9, (k, err)
