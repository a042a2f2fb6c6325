"""GPU parity of the training step (forward with saved activations, backward, AdamW) against the reference's own
autograd / torch.optim.AdamW run (tests/golden/train_<ckpt>.npz, generated from /root/reference by make_golden.py).

Tolerances (fp32, different summation orders): forward mask 1e-4 of its range (as the inference tests); every
parameter gradient within 2e-4 of that tensor's max |grad| (+ a 1e-7 absolute floor); loss within 1e-4 relative;
parameters after three AdamW steps within 3e-5 absolute (lr = 5e-4: a sign flip of a ~zero gradient moves a weight
by up to lr per step, so elementwise agreement is checked at the 99.9th percentile and the max is bounded by 3 lr)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

CASES = ["different-armadillo", "clean-serenity"]


@pytest.fixture(scope="module")
def sb():
    import stofnet_b200
    assert torch.cuda.is_available()
    return stofnet_b200


def _model(sb, golden, name, dev):
    sd = {k: torch.from_numpy(v) for k, v in golden("weights_%s.npz" % name).items()}
    return sb.build_model(sd, dev)


@pytest.mark.parametrize("name", CASES)
def test_backward_matches_reference_autograd(sb, golden, name):
    dev = torch.device("cuda:0")
    g = golden("train_%s.npz" % name)
    model = _model(sb, golden, name, dev).train()
    x = torch.from_numpy(g["x"]).unsqueeze(1).to(dev)
    gt = torch.from_numpy(g["gt"]).to(dev)
    masks_pred = model(x)                                       # autograd path: stof_forward_train
    ref_mask = g["mask0"]
    assert np.abs(masks_pred.detach().squeeze(1).cpu().numpy() - ref_mask).max() <= 1e-4 * np.abs(ref_mask).max()
    loss = sb.mask_loss(masks_pred, gt, 7, 1.0, 20.0, float(g["lambda"]))
    assert abs(float(loss.detach()) - g["losses"][0]) <= 1e-4 * abs(g["losses"][0])
    loss.backward()                                             # stof_backward
    names = [str(k) for k in g["names"]]
    assert names == [k for k, _ in model.named_parameters()]
    for k, p in model.named_parameters():
        ref = g["grad0/" + k]
        got = p.grad.cpu().numpy()
        assert got.shape == ref.shape
        err = np.abs(got - ref).max()
        assert err <= 2e-4 * np.abs(ref).max() + 1e-7, (k, err, np.abs(ref).max())


@pytest.mark.parametrize("name", CASES)
def test_backward_is_deterministic(sb, golden, name):
    dev = torch.device("cuda:0")
    g = golden("train_%s.npz" % name)
    model = _model(sb, golden, name, dev).train()
    x = torch.from_numpy(g["x"]).unsqueeze(1).to(dev)
    gt = torch.from_numpy(g["gt"]).to(dev)
    grads = []
    for _ in range(2):
        model.zero_grad()
        sb.mask_loss(model(x), gt, 7, 1.0, 20.0, float(g["lambda"])).backward()
        grads.append(torch.cat([p.grad.reshape(-1) for p in model.parameters()]).clone())
    assert torch.equal(grads[0], grads[1])


@pytest.mark.parametrize("name", CASES)
def test_three_adamw_steps(sb, golden, name):
    dev = torch.device("cuda:0")
    g = golden("train_%s.npz" % name)
    model = _model(sb, golden, name, dev).train()
    tr = sb.Trainer(model, lr=5e-4, weight_decay=1e-8, lambda_value=float(g["lambda"]))
    x = torch.from_numpy(g["x"]).unsqueeze(1).to(dev)
    gt = torch.from_numpy(g["gt"]).to(dev)
    losses = [float(tr.step(x, gt)) for _ in range(3)]
    ref = g["losses"]
    assert abs(losses[0] - ref[0]) <= 1e-4 * abs(ref[0])
    assert abs(losses[1] - ref[1]) <= 5e-3 * abs(ref[1]) and abs(losses[2] - ref[2]) <= 5e-3 * abs(ref[2])
    for k, p in model.named_parameters():
        d = np.abs(p.detach().cpu().numpy() - g["param3/" + k])
        assert np.quantile(d, 0.999) <= 3e-5, (k, np.quantile(d, 0.999))
        assert d.max() <= 3 * 5e-4 + 1e-6, (k, d.max())
    # the reference-style loop (torch.optim.AdamW on our module + loss.backward()) gives the same trajectory
    model2 = _model(sb, golden, name, dev).train()
    opt = torch.optim.AdamW(model2.parameters(), lr=5e-4, weight_decay=1e-8)
    l2 = []
    for _ in range(3):
        loss = sb.mask_loss(model2(x), gt, 7, 1.0, 20.0, float(g["lambda"]))
        opt.zero_grad()
        loss.backward()
        opt.step()
        l2.append(float(loss))
    assert np.allclose(l2, losses, rtol=2e-3)


def test_inference_after_training_uses_new_weights(sb, golden):
    dev = torch.device("cuda:0")
    g = golden("train_clean-serenity.npz")
    model = _model(sb, golden, "clean-serenity", dev).train()
    tr = sb.Trainer(model, lambda_value=float(g["lambda"]))
    x = torch.from_numpy(g["x"]).unsqueeze(1).to(dev)
    gt = torch.from_numpy(g["gt"]).to(dev)
    with torch.no_grad():
        before = model(x).clone()
    tr.step(x, gt)
    with torch.no_grad():
        after = model(x)
    assert not torch.equal(before, after)
    sd = model.state_dict()
    assert all(torch.isfinite(v).all() for v in sd.values())
    # the optimizer step re-derives the kernels' weight layouts on the device (stof_model_set_weights_device) and the
    # tensor-core packs are rebuilt from a device-to-host copy: a fresh handle created from the state_dict through the
    # host path must give the same masks bit for bit - on the tensor-core path, the fp32 path and the training forward
    fresh = sb.build_model({k: v.detach().cpu().clone() for k, v in sd.items()}, dev)
    with torch.no_grad():
        assert torch.equal(after, fresh(x))
        model._engine.set_path(sb._lib.PATH_SIMT)
        fresh._engine.set_path(sb._lib.PATH_SIMT)
        assert torch.equal(model(x), fresh(x))
    model.train(); fresh.train()
    assert torch.equal(model(x).detach(), fresh(x).detach())
