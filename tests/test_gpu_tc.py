"""GPU tests of the tcgen05 chain kernel (conv_tc.cu) through the C ABI.

Layer-level checks feed split fp16 hi/lo planes into stof_chain_segment and compare with a float64 torch restatement of
the same layers (models/stofnet.py:51-65 of the reference); end-to-end checks live in test_gpu_parity.py, which runs the
tensor-core path by default."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from conftest import STOF_CKPTS

pytestmark = pytest.mark.gpu

REL_TOL = 2e-5      # of max |reference|: fp32 accumulation-order noise + 2^-22 operand split


@pytest.fixture(scope="module")
def sb():
    import stofnet_b200
    assert torch.cuda.is_available()
    return stofnet_b200


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


def to_planes(a):
    """(N, 64, L) fp32 -> (N, 2, 8, L, 8) fp16 hi/lo planes."""
    N, C, L = a.shape
    c = a.clamp(-65504.0, 65504.0)
    hi = c.half()
    lo = ((c - hi.float()) * 2048.0).half()

    def lay(h):
        return h.view(N, 8, 8, L).permute(0, 1, 3, 2)
    return torch.stack([lay(hi), lay(lo)], 1).contiguous()


def from_planes(p):
    N, _, _, L, _ = p.shape
    v = p[:, 0].float() + p[:, 1].float() / 2048.0
    return v.permute(0, 1, 3, 2).reshape(N, 64, L)


def ref_layers(model, x, x0, first, count):
    """float64 reference of layers [first, first+count) on CPU; x, x0: (N, 64, L)."""
    n_mid = model.num_blocks - 2
    x = x.double().cpu()
    x0 = x0.double().cpu() if x0 is not None else None
    res = x.clone()          # only used when the range starts on an even (LeakyReLU) layer: res = its input
    for j in range(first, first + count):
        if j == n_mid:
            w, b = model.conv_last.weight.double().cpu(), model.conv_last.bias.double().cpu()
            y = F.conv1d(x, w, b, padding=1)                     # (N, 4, L)
            return y.permute(0, 2, 1).reshape(y.shape[0], 1, -1)  # sample shuffle, C_long = 1
        conv = getattr(model, "conv%d" % (j + 2))
        y = F.conv1d(x, conv.weight.double().cpu(), conv.bias.double().cpu(), padding=3)
        if j == n_mid - 1:
            x = x0 + y
        elif j % 2 == 0:
            res = x
            x = F.leaky_relu(y, 0.01)
        else:
            x = res + y
    return x


def _model(sb, golden, name, dev):
    sd = {k: torch.from_numpy(v) for k, v in golden("weights_%s.npz" % name).items()}
    return sb.build_model(sd, dev)


def _check(got, ref, what):
    ref = ref.float()
    err = (got.cpu() - ref).abs().max().item() / ref.abs().max().item()
    assert err <= REL_TOL, (what, err)


@pytest.mark.parametrize("first,count,L", [(0, 1, 700), (0, 2, 700), (0, 4, 1300), (4, 4, 256), (2, 2, 250), (0, 6, 999)])
def test_chain_mid_layers(sb, dev, golden, first, count, L):
    model = _model(sb, golden, "different-armadillo", dev)
    g = torch.Generator().manual_seed(100 * first + count)
    x = torch.randn(3, 64, L, generator=g) * 2.0
    out = model._engine.chain_segment(model, first, count, to_planes(x.to(dev)))
    torch.cuda.synchronize()
    _check(from_planes(out), ref_layers(model, x, None, first, count), (first, count, L))


@pytest.mark.parametrize("name", ["different-armadillo", "breezy-violet"])
@pytest.mark.parametrize("first,count", [(10, 2), (8, 4), (10, 1), (11, 1)])
def test_chain_tail_with_long_skip_and_shuffle(sb, dev, golden, name, first, count):
    model = _model(sb, golden, name, dev)
    L = 900
    g = torch.Generator().manual_seed(7)
    x = torch.randn(2, 64, L, generator=g)
    x0 = torch.randn(2, 64, L, generator=g)
    out = model._engine.chain_segment(model, first, count, to_planes(x.to(dev)), to_planes(x0.to(dev)))
    torch.cuda.synchronize()
    ref = ref_layers(model, x, x0, first, count)
    if first + count == model.num_blocks - 1:
        assert out.shape == (2, 1, 4 * L)
        _check(out, ref, (name, first, count))
    else:
        _check(from_planes(out), ref, (name, first, count))


def test_chain_large_magnitude_and_edges(sb, dev, golden):
    """values beyond the fp16 normal range of the lo plane, exact zeros, and a row shorter than one tile"""
    model = _model(sb, golden, "graceful-snow", dev)
    for L in (80, 129, 239):
        g = torch.Generator().manual_seed(L)
        x = torch.randn(5, 64, L, generator=g) * torch.tensor([1e-4, 1.0, 50.0, 0.0, 3.0]).view(5, 1, 1)
        out = model._engine.chain_segment(model, 0, 4, to_planes(x.to(dev)))
        torch.cuda.synchronize()
        _check(from_planes(out), ref_layers(model, x, None, 0, 4), L)


@pytest.mark.parametrize("name", STOF_CKPTS)
@pytest.mark.parametrize("seg", [1, 2, 4, 6])
def test_forward_tc_matches_simt(sb, dev, golden, name, seg):
    from stofnet_b200 import _lib
    from stofnet_b200.synth import synth_waveforms
    model = _model(sb, golden, name, dev)
    x = torch.from_numpy(synth_waveforms(6, 1280, seed=3, rf=10)).unsqueeze(1).to(dev)
    with torch.no_grad():
        model._engine.set_path(_lib.PATH_SIMT)
        a = model(x)
        model._engine.set_path(_lib.PATH_TC, seg)
        b = model(x)
    torch.cuda.synchronize()
    err = (a - b).abs().max().item() / a.abs().max().item()
    # two fp32-faithful paths with different accumulation orders (fp32 FFMA vs fp16 hi/lo split on tcgen05, the semi-global
    # block included): each is within FWD_TOL of the fp64 oracle (test_gpu_parity.py), so they differ by at most twice that
    assert err <= 2 * REL_TOL, (name, seg, err)


@pytest.mark.parametrize("N,P", [(1, 1), (3, 2), (5, 31), (7, 64), (2, 65), (3, 700), (300, 64)])
def test_sgb_expand_tc(sb, dev, golden, N, P):
    """expand_conv (512 -> 64, k5, pad 2) + LeakyReLU on tcgen05 (expand_tc.cu) against float64 torch, and against the
    fp32 FFMA kernel; row ends are zero padded per waveform row (no leakage between rows of a tile)."""
    from stofnet_b200._lib import lib, check, PATH_TC, PATH_SIMT
    from stofnet_b200.engine import stream_ptr
    model = _model(sb, golden, "graceful-snow", dev)
    model._engine.sync(model, dev)
    g = torch.Generator().manual_seed(N * 1000 + P)
    pooled = (torch.randn(N, 512, P, generator=g) * torch.rand(N, 512, 1, generator=g) * 3).to(dev)
    w = model.semi_global_block.expand_conv.weight.detach().double()
    b = model.semi_global_block.expand_conv.bias.detach().double()
    ref = F.leaky_relu(F.conv1d(pooled.double(), w, b, padding=2))
    out = {}
    for path in (PATH_TC, PATH_SIMT):
        e = torch.full((N, 64, P), float("nan"), device=dev)
        check(lib.stof_sgb_expand(model._engine.handle, pooled.data_ptr(), N, P, e.data_ptr(), path, stream_ptr(dev)))
        torch.cuda.synchronize()
        out[path] = e
        err = (e.double() - ref).abs().max().item() / ref.abs().max().item()
        assert err <= 1e-5, (path, N, P, err)      # K = 2560 products, fp32 accumulation (either path measures 1e-6 .. 4e-6)


def test_fp16_one_pass_mode_tolerance(sb, dev, golden):
    """STOF_PATH_TC_FP16: one tensor-core pass on fp16-rounded operands.  Not index-exact by design; stated tolerance
    (DESIGN.md): mask error <= 2e-2 of the mask range (rms <= 5e-4), >= 97 % of the exact mode's detections reproduced
    at the identical mask sample, and the exact mode is restored by set_precision('fp32')."""
    from stofnet_b200.synth import synth_waveforms
    for name, rf, th in (("different-armadillo", 10, 0.5), ("graceful-snow", 20, 0.015)):
        sd = {k: torch.from_numpy(v) for k, v in golden("weights_%s.npz" % name).items()}
        model = sb.build_model(sd, dev)
        x = torch.from_numpy(synth_waveforms(96, 256 * rf, seed=5, rf=rf)).unsqueeze(1).to(dev)
        with torch.no_grad():
            ref = model(x)
            fast = model.set_precision("fp16")(x)
            back = model.set_precision("fp32")(x)
        assert torch.equal(ref, back)
        rng_ = float(ref.abs().max())
        assert float((fast - ref).abs().max()) <= 2e-2 * rng_
        assert float((fast - ref).pow(2).mean().sqrt()) <= 5e-4 * rng_
        a = set(map(tuple, sb.get_maxima_positions(ref, 20, th).cpu().numpy()))
        b = set(map(tuple, sb.get_maxima_positions(fast, 20, th).cpu().numpy()))
        assert len(a & b) >= 0.97 * len(a) and len(a & b) >= 0.97 * len(b), (name, len(a), len(b), len(a & b))


@pytest.mark.parametrize("name", ["graceful-snow", "clean-serenity", "proud-cherry"])
def test_pair_chain_kernel_bit_identical(sb, dev, golden, name):
    """conv_tc2.cu (2-CTA clusters, tcgen05 cta_group::2, each CTA holding half of every weight tap) against conv_tc.cu: same
    MMAs in the same order per accumulator, so the masks must be bit-identical - including odd tile counts (the last pair's
    second CTA runs a phantom tile), both precisions, and EDSR_1D's k3 layers."""
    from stofnet_b200._lib import lib, check
    from stofnet_b200.synth import synth_waveforms
    sd = {k: torch.from_numpy(v) for k, v in golden("weights_%s.npz" % name).items()}
    model = sb.build_model(sd, dev)
    for N, L in ((1, 160), (3, 640), (5, 2578), (41, 5120)):
        x = torch.from_numpy(synth_waveforms(N, L, seed=N, rf=10)).unsqueeze(1).to(dev)
        for prec in ("fp32", "fp16"):
            model.set_precision(prec)
            outs = []
            for pair in (0, 1):
                with torch.no_grad():
                    model(x[:1])                                            # creates / syncs the handle
                    check(lib.stof_model_set_chain_kernel(model._engine.handle, pair))
                    outs.append(model(x).clone())
            check(lib.stof_model_set_chain_kernel(model._engine.handle, 0))
            assert torch.equal(outs[0], outs[1]), (name, N, L, prec, float((outs[0] - outs[1]).abs().max()))
    model.set_precision("fp32")
