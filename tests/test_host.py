"""CPU-side tests: the C-ABI library loads and exports every symbol the header declares, host logic
(checkpoint / config loading, module surface) behaves like the reference, and compute fails loudly
without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT, STOF_CKPTS

import stofnet_b200 as sb
from stofnet_b200 import _lib


def header_functions():
    txt = open(os.path.join(ROOT, "include", "stofnet_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(stof_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    names = header_functions()
    assert len(names) >= 15
    lib = C.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), "libstofnet_b200.so does not export %s" % n
    # and the Python binding covers all of them
    assert set(names) == set(_lib.EXPORTS)
    assert b"sm_100a" in _lib.lib.stof_version()


def test_library_is_built_from_this_tree():
    """stof_version() embeds the SHA-256 of the sources the .so was compiled from (csrc/Makefile): a stale pre-built
    library (the .so is git-ignored and shipped to the GPU box as built) cannot pass the suite silently."""
    import hashlib
    import glob
    csrc = os.path.join(ROOT, "stofnet_b200", "csrc")
    files = sorted(glob.glob(os.path.join(csrc, "*.cu")) + glob.glob(os.path.join(csrc, "*.cuh")), key=os.path.basename)
    h = hashlib.sha256()
    for f in files + [os.path.join(ROOT, "include", "stofnet_b200.h")]:
        h.update(open(f, "rb").read())
    want = h.hexdigest()[:16]
    got = _lib.lib.stof_version().decode()
    assert got.endswith("src:" + want), "libstofnet_b200.so is stale: built from %s, tree is %s - run __graft_entry__.build()" % (got, want)


def test_library_is_sm100a_with_lineinfo():
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out


@pytest.mark.parametrize("name", STOF_CKPTS)
def test_checkpoints_load_strict_and_roundtrip(golden, name, tmp_path):
    sd = {k: torch.from_numpy(v) for k, v in golden("weights_%s.npz" % name).items()}
    model = sb.build_model(sd)
    assert isinstance(model, sb.StofNet)
    has_sgb = any(k.startswith("semi_global_block") for k in sd)
    assert (model.semi_global_block is not None) == has_sgb
    assert sum(p.numel() for p in model.parameters()) == (645764 if has_sgb else 317508)
    assert model.residual_layers == [3, 5, 7, 9, 11, 12, 13] and model.num_blocks == 13
    # .pth round trip exactly like main.py:427 / :176-177
    p = str(tmp_path / (name + "_rf-scale10_epoch_1.pth"))
    torch.save(model.state_dict(), p)
    sd2 = sb.load_checkpoint(p)
    assert set(sd2) == set(sd)
    for k in sd:
        assert torch.equal(sd2[k], sd[k])
    assert sb.checkpoint.find_checkpoint(str(tmp_path), name + "_rf-scale10") == p
    with pytest.raises(RuntimeError):          # strict load: wrong topology must fail like the reference
        sb.StofNet(semi_global_scale=1 if has_sgb else 80).load_state_dict(sd)


def test_espcn_checkpoint_detected(golden):
    sd = {k: torch.from_numpy(v) for k, v in golden("weights_wobbly-sponge.npz").items()}
    model = sb.build_model(sd)
    assert isinstance(model, sb.ESPCN_1D) and model.upsample_factor == 4
    with pytest.raises(ValueError):
        sb.build_model({"foo.weight": torch.zeros(1)})


def test_edsr_checkpoint_detected(golden):
    sd = {k: torch.from_numpy(v) for k, v in golden("weights_proud-cherry.npz").items()}
    model = sb.build_model(sd)
    assert isinstance(model, sb.EDSR_1D) and model.upsample_factor == 4 and len(model.residual_blocks) == 8
    assert set(model.state_dict()) == set(sd)
    with pytest.raises(RuntimeError, match="no CPU path"):
        model(torch.zeros(1, 1, 160))


def test_no_cpu_fallback(golden):
    sd = {k: torch.from_numpy(v) for k, v in golden("weights_clean-serenity.npz").items()}
    model = sb.build_model(sd)
    with pytest.raises(RuntimeError, match="no CPU path"):
        model(torch.zeros(1, 1, 160))
    with pytest.raises(RuntimeError, match="no CPU path"):
        sb.mask2coords(torch.zeros(1, 1, 64), 20, 0.5)
    if not torch.cuda.is_available():
        # the C ABI itself reports the missing device instead of computing anything on the host
        W, keep = model._engine._pack(dict(model.state_dict()))
        h = C.c_void_p()
        rc = _lib.lib.stof_model_create(C.byref(W), 0, C.byref(h))
        assert rc == _lib.E_CUDA and h.value is None
        assert "CUDA" in _lib.last_error()


def test_abi_argument_validation():
    rc = _lib.lib.stof_model_create(None, 0, None)
    assert rc == _lib.E_INVALID
    W = _lib.Weights()
    h = C.c_void_p()
    assert _lib.lib.stof_model_create(C.byref(W), 0, C.byref(h)) == _lib.E_INVALID
    sz = C.c_size_t()
    assert _lib.lib.stof_forward_workspace_bytes(None, 4, 160, 0, C.byref(sz)) == _lib.E_INVALID
    assert _lib.lib.stof_peaks(None, 1, 1, 20, 0, 0.5, 4, None, None, 0, None, None, None) == _lib.E_INVALID
    assert _lib.lib.stof_model_set_weights_device(None, C.byref(W), None) == _lib.E_INVALID
    assert "NULL" in _lib.last_error()
    with pytest.raises(ValueError):
        _lib.check(_lib.E_INVALID)
    with pytest.raises(RuntimeError):
        _lib.check(_lib.E_LENGTH_ODD)


def test_config_loader_keeps_reference_options(tmp_path):
    p = tmp_path / "config.yaml"
    p.write_text("seed: 3008\nlr: 5e-4\nweight_decay: 1e-8\nth: 0.5\nnms_win_size: 20\nupsample_factor: 4\n"
                 "rf_scale_factor: 10\netol: 1\nmodel_file: Null\ndata_path: PALA/\nubx_dir: /x/${data_path}\n")
    cfg = sb.load_config(str(p), ["th=Null", "rf_scale_factor=20", "model_file=graceful-snow", "etol=0.25"])
    assert cfg.lr == 5e-4 and cfg.weight_decay == 1e-8 and isinstance(cfg.lr, float)
    assert cfg.th is None and cfg.rf_scale_factor == 20 and cfg.etol == 0.25
    assert cfg.nms_win_size == 20 and cfg.upsample_factor == 4 and cfg.model_file == "graceful-snow"
    assert cfg.ubx_dir == "/x/PALA/"


def test_synth_is_deterministic_and_normalised():
    from stofnet_b200.synth import synth_waveforms, pala_shaped, valid_length
    a, b = synth_waveforms(5, 640, seed=3), synth_waveforms(5, 640, seed=3)
    assert np.array_equal(a, b) and a.dtype == np.float32
    assert np.allclose(np.abs(a).max(1), 1.0)
    p = pala_shaped(8, 2, 800, seed=1)
    assert p.shape == (16, 800) and np.allclose(np.abs(p).max(1), 1.0)
    assert valid_length(2560) and valid_length(2578) and not valid_length(161) and not valid_length(79)


def test_product_does_not_import_oracle():
    """The oracle is test infrastructure: nothing under stofnet_b200/ may reference it."""
    for dp, _, files in os.walk(os.path.join(ROOT, "stofnet_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert "oracle" not in txt.replace("oracle/", "").lower() or f == "__init__.py", (dp, f)
