"""Checkpoint and config loading with the reference's conventions.

* ``load_checkpoint``: the shipped ``.pth`` files are plain fp32 state_dicts (main.py:174-177,
  425-427); also accepts the ``.npz`` dumps of tests/golden (same key names).
* ``build_model``: detects the topology from the keys (with / without ``semi_global_block.*``,
  StofNet vs ESPCN_1D) like main.py:132-167 does from ``cfg.model``.
* ``load_config``: config.yaml + ``key=value`` overrides (OmegaConf in the reference, main.py:31-34);
  PyYAML parses ``5e-4`` / ``1e-8`` as strings and ``Null`` as None, so numeric strings are coerced.
"""
import glob
import os
import re

import numpy as np
import torch


def load_checkpoint(path):
    if path.endswith(".npz"):
        with np.load(path) as z:
            return {k: torch.from_numpy(np.ascontiguousarray(z[k])) for k in z.files}
    sd = torch.load(path, map_location="cpu", weights_only=True)
    if not isinstance(sd, dict):
        raise ValueError("%s does not hold a state_dict" % path)
    return sd


def find_checkpoint(ckpt_dir, model_file):
    """First file whose name starts with model_file.split('_')[0] (main.py:174)."""
    prefix = model_file.split("_")[0]
    for p in sorted(glob.glob(os.path.join(ckpt_dir, "*"))):
        if os.path.basename(p).startswith(prefix):
            return p
    raise FileNotFoundError("no checkpoint starting with %r in %s" % (prefix, ckpt_dir))


def build_model(sd, device=None):
    """state_dict -> (model in eval mode with the weights loaded strictly)."""
    from .models.stofnet import StofNet
    from .models.espcn_1d import ESPCN_1D
    from .models.edsr_1d import EDSR_1D
    keys = set(sd)
    if "conv_last.weight" in keys:
        nb = 2 + sum(1 for k in keys if re.fullmatch(r"conv\d+\.weight", k) and k != "conv1.weight")
        r = sd["conv_last.weight"].shape[0]
        has_sgb = any(k.startswith("semi_global_block.") for k in keys)
        scale = 1
        if has_sgb:
            cw = sd["semi_global_block.contract_conv.weight"]
            scale = 10 * (cw.shape[0] // cw.shape[1])      # feat_scale = scale // 10 (stofnet.py:85)
        model = StofNet(upsample_factor=r, num_features=sd["conv1.weight"].shape[0], num_blocks=nb,
                        semi_global_scale=scale)
    elif {"conv1.weight", "conv2.weight", "conv3.weight"} <= keys and len(keys) == 6:
        model = ESPCN_1D(sd["conv3.weight"].shape[0])
    elif "conv_input.weight" in keys and "conv_output.weight" in keys:
        nb = 1 + max(int(k.split(".")[1]) for k in keys if k.startswith("residual_blocks."))
        f = sd["conv_input.weight"].shape[0]
        model = EDSR_1D(num_channels=sd["conv_input.weight"].shape[1], num_features=f, num_blocks=nb,
                        upscale_factor=f // sd["conv_output.weight"].shape[1])
    else:
        raise ValueError("Model not recognized")          # main.py:167
    model.load_state_dict(sd, strict=True)
    model.eval()
    if device is not None:
        model = model.to(device)
    return model


_NUM = re.compile(r"^[+-]?(\d+\.?\d*|\.\d+)([eE][+-]?\d+)?$")


def _coerce(v):
    if isinstance(v, str):
        s = v.strip()
        if s.lower() in ("null", "none", "~"):
            return None
        if s.lower() in ("true", "false"):
            return s.lower() == "true"
        if _NUM.match(s):
            f = float(s)
            return int(f) if re.fullmatch(r"[+-]?\d+", s) else f
    return v


class Config(dict):
    __getattr__ = dict.get

    def __setattr__(self, k, v):
        self[k] = v


def load_config(path=None, overrides=()):
    """config.yaml options keep their names and meaning: rf_scale_factor, th, etol, upsample_factor,
    nms_win_size, model_file ... (config.yaml:1-45)."""
    import yaml
    cfg = Config()
    if path is not None:
        with open(path) as f:
            raw = yaml.safe_load(f) or {}
        for k, v in raw.items():
            cfg[k] = _coerce(v)
    for ov in overrides:
        k, _, v = ov.partition("=")
        cfg[k.strip()] = _coerce(yaml.safe_load(v) if v.strip() else None)
    # ${key} interpolation (OmegaConf resolver semantics for the simple case used in config.yaml:41-45)
    for k, v in list(cfg.items()):
        if isinstance(v, str):
            cfg[k] = re.sub(r"\$\{(\w+)\}", lambda m: str(cfg.get(m.group(1), m.group(0))), v)
    return cfg
