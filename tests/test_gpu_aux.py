"""GPU parity of the operators either side of the path (through the C ABI) against the golden vectors generated
from the reference and against the oracle on seeded inputs.

Bars: counts / indices / thresholds (integer and comparison work) bit-exact; fp32 sums within the tolerance stated
at each assert (summation order differs from torch's)."""
import numpy as np
import pytest
import torch

from oracle import c_oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sb():
    import stofnet_b200
    assert torch.cuda.is_available()
    return stofnet_b200


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


def _close_nan(a, b, tol):
    assert a.shape == b.shape
    assert np.array_equal(np.isnan(a), np.isnan(b))
    m = ~np.isnan(a)
    assert np.abs(a[m] - b[m]).max(initial=0.0) <= tol


def test_toa_rmse_golden(sb, dev, golden):
    g = golden("metrics_cases.npz")
    gt, es = torch.from_numpy(g["rmse_gt"]).to(dev), torch.from_numpy(g["rmse_es"]).to(dev)
    for key, tol in (("1", 1.0), ("0p25", 0.25), ("4", 4.0)):
        got = sb.toa_rmse(gt, es, tol=tol).cpu().numpy()
        ref = g["rmse_tol" + key]
        _close_nan(got[:, :4], ref[:, :4], 1e-4)          # rmse / percentages: fp32 sum order
        assert np.array_equal(got[:, 4:], ref[:, 4:])     # tp, fp, fn counts: exact


def test_toa_rmse_random_vs_oracle(sb, dev):
    rng = np.random.default_rng(4)
    gt = rng.uniform(1, 2000, size=(300, 70)).astype(np.float32)
    gt[rng.uniform(size=gt.shape) < 0.3] = 0
    es = (gt[:, :50] + rng.normal(0, 0.7, size=(300, 50))).astype(np.float32)
    es[rng.uniform(size=es.shape) < 0.2] = 0
    gt[5] = 0; es[9] = 0; es[11, 3] = np.nan; gt[12, 1] = np.inf
    got = sb.toa_rmse(torch.from_numpy(gt).to(dev), torch.from_numpy(es).to(dev), tol=1).cpu().numpy()
    ref = c_oracle.toa_rmse(gt, es, 1.0)
    _close_nan(got[:, :4], ref[:, :4], 1e-4)
    assert np.array_equal(got[:, 4:], ref[:, 4:])


def test_coords2mask_golden(sb, dev, golden):
    g = golden("metrics_cases.npz")
    smp = torch.from_numpy(g["c2m_samples"].copy()).to(dev)
    mask = sb.coords2mask(smp, torch.zeros((3, 1, 64), device=dev))
    assert np.array_equal(mask.cpu().numpy(), g["c2m_mask"])
    assert int(smp.min()) == 0                              # negatives clamped in place like the reference


def test_mask_loss_golden(sb, dev, golden):
    from stofnet_b200.loss import mask_loss_raw
    g = golden("aux_cases.npz")
    for name in g["loss_names"]:
        p = "loss_%s_" % name
        pred = torch.from_numpy(g[p + "pred"]).to(dev)
        gt = torch.from_numpy(g[p + "gt"]).to(dev)
        loss, grad = mask_loss_raw(pred, gt, g["gauss7"], 20.0, float(g[p + "lambda"]), want_grad=True)
        ref = float(g[p + "value"])
        assert abs(float(loss.item()) - ref) <= 2e-6 * abs(ref), name
        rg = g[p + "grad"].reshape(grad.shape)
        assert np.abs(grad.cpu().numpy() - rg).max() <= 1e-6 * np.abs(rg).max() + 1e-12, name
        # differentiable wrapper: autograd reaches masks_pred
        pr = pred.clone().requires_grad_(True)
        l2 = sb.mask_loss(pr, gt, 7, 1.0, 20.0, float(g[p + "lambda"]))
        l2.backward()
        assert torch.equal(pr.grad.reshape(grad.shape), grad)


def test_mask_loss_sharded_blur_max(sb, dev, golden):
    """two-phase form used under data parallelism: blur max computed separately (MAX all-reduce point) gives the
    same loss as the single call"""
    from stofnet_b200.loss import mask_loss_raw, blur_max
    g = golden("aux_cases.npz")
    pred = torch.from_numpy(g["loss_b_pred"]).to(dev)
    gt = torch.from_numpy(g["loss_b_gt"]).to(dev)
    mx = blur_max(gt, pred.shape[-1], g["gauss7"])
    l1, g1 = mask_loss_raw(pred, gt, g["gauss7"], 20.0, 1.0, True)
    l2, g2 = mask_loss_raw(pred, gt, g["gauss7"], 20.0, 1.0, True, blur_max_value=mx)
    assert float(l1) == float(l2) and torch.equal(g1, g2)


def test_find_threshold_golden(sb, dev, golden):
    g = golden("aux_cases.npz")
    for name in g["ft_names"]:
        p = "ft_%s_" % name
        pred, true = torch.from_numpy(g[p + "pred"]).to(dev), torch.from_numpy(g[p + "true"]).to(dev)
        for key, norm in (("th", False), ("th_norm", True)):
            got = sb.find_threshold(pred, true, 20, norm_opt=norm)
            ref = float(g[p + key])
            assert got == ref or (np.isinf(got) and np.isinf(ref)), (name, key, got, ref)
    m = golden("metrics_cases.npz")
    pred, true = torch.from_numpy(m["ft_pred"]).to(dev), torch.from_numpy(m["ft_true"]).to(dev)
    assert sb.find_threshold(pred, true, 20) == float(m["ft_th"])
    assert sb.find_threshold(pred, true, 20, norm_opt=True) == float(m["ft_th_norm"])


@pytest.mark.parametrize("M", [3, 4096, 4097, 20480, 100001])
def test_find_threshold_sizes_vs_oracle(sb, dev, M):
    rng = np.random.default_rng(M)
    pred = rng.normal(size=(2, 1, M)).astype(np.float32)
    true = (rng.uniform(size=(2, 1, M)) < 0.01).astype(np.float32)
    true[0, 0, 0] = 1; pred[0, 0, 0] += 2
    if M > 3:
        pred[0, 0, 1:M:7] = 0.25            # ties
    got = sb.find_threshold(torch.from_numpy(pred).to(dev), torch.from_numpy(true).to(dev), 20)
    assert got == c_oracle.find_threshold(pred, true)


def test_rf_upscale_golden(sb, dev, golden):
    g = golden("aux_cases.npz")
    env = torch.from_numpy(g["up_env"]).to(dev)
    for rf, key in ((10, "10"), (20, "20"), (2.5, "2p5")):
        got = sb.upscale_1d(env, rf).cpu().numpy()
        ref = g["up_env_rf" + key]
        assert got.shape == ref.shape
        assert np.abs(got - ref).max() <= 1e-6 * np.abs(ref).max()
    iq = torch.complex(torch.from_numpy(g["up_iq_re"]), torch.from_numpy(g["up_iq_im"])).to(dev)
    for rf in (10, 20):
        got = sb.iq2rf(iq, fc=float(g["up_fc"]), fs=float(g["up_fs"]), rescale_factor=rf).cpu().numpy()
        ref = g["up_iq_rf%d" % rf]
        assert got.shape == ref.shape
        assert np.abs(got - ref).max() <= 2e-6 * np.abs(ref).max()
    one = sb.upscale_1d(env[0], 10)
    assert one.dim() == 1 and torch.equal(one, sb.upscale_1d(env, 10)[0])


def test_attention_block(sb, dev, golden):
    g = golden("aux_cases.npz")
    blk = sb.AttentionBlock(48, 40, dropout_p=0.0)
    with torch.no_grad():
        for nm in ("query", "key", "value"):
            getattr(blk, nm).weight.copy_(torch.from_numpy(g["att_%s_w" % nm]))
            getattr(blk, nm).bias.copy_(torch.from_numpy(g["att_%s_b" % nm]))
    blk = blk.to(dev)
    x = torch.from_numpy(g["att_x"]).to(dev)
    y = blk(x).cpu().numpy()
    assert np.abs(y - g["att_y"]).max() <= 2e-5 * np.abs(g["att_y"]).max()     # fp32 vs torch's fp32 SDPA
    # the reference's hard-coded dropout_p = 0.5: statistical check (unbiased, about half the weights dropped)
    blk.dropout_p = 0.5
    ys = torch.stack([blk(x, seed=s) for s in range(64)]).mean(0).cpu().numpy()
    assert np.abs(ys - g["att_y"]).mean() <= 0.1 * np.abs(g["att_y"]).mean() + 0.05
    assert not torch.equal(blk(x, seed=1), blk(x, seed=2)) and torch.equal(blk(x, seed=3), blk(x, seed=3))
