#!/usr/bin/env python
"""One-pass fp16 mode (STOF_PATH_TC_FP16) against the exact mode: mask error, detection agreement, ToF error, speed."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import stofnet_b200 as sb
from stofnet_b200.synth import synth_waveforms

dev = torch.device("cuda:0")
GOLD = os.path.join(ROOT, "tests", "golden")
CK = {"different-armadillo": (10, None), "clean-serenity": (10, None), "graceful-snow": (20, 0.015), "breezy-violet": (20, 2.3)}
for name, (rf, th0) in CK.items():
    model = sb.build_model(sb.load_checkpoint(os.path.join(GOLD, "weights_%s.npz" % name)), dev)
    L = 256 * rf
    x = torch.from_numpy(synth_waveforms(512, L, seed=3008, rf=rf)).unsqueeze(1).to(dev)
    with torch.no_grad():
        model.set_precision("fp32"); ref = model(x)
        model.set_precision("fp16"); fast = model(x)
    rng_ = float(ref.abs().max())
    err = float((fast - ref).abs().max())
    out = {"ckpt": name, "L": L, "rows": 512, "mask_max_abs_err_over_range": err / rng_, "mask_rms_err_over_range": float((fast - ref).pow(2).mean().sqrt()) / rng_}
    for th in sorted({th0, 0.5}, key=lambda v: (v is not None, v or 0)):
        a = sb.get_maxima_positions(ref, 20, th).cpu().numpy()
        b = sb.get_maxima_positions(fast, 20, th).cpu().numpy()
        sa = set(map(tuple, a)); sbb = set(map(tuple, b))
        tp = len(sa & sbb)
        # matched within +-1 mask sample (0.25 input samples)
        near = 0
        bd = {}
        for r, i in b: bd.setdefault(r, []).append(i)
        for r, i in a:
            if any(abs(i - j) <= 1 for j in bd.get(r, ())): near += 1
        out["th=%s" % th] = {"ref": len(sa), "fast": len(sbb), "identical": tp, "within_1_mask_sample": near,
                             "precision": tp / max(1, len(sbb)), "recall": tp / max(1, len(sa))}
    print(json.dumps(out), flush=True)

# speed on the bench workload
model = sb.build_model(sb.load_checkpoint(os.path.join(GOLD, "weights_graceful-snow.npz")), dev)
x = torch.from_numpy(synth_waveforms(256, 5120, seed=1, rf=20)).repeat(16, 1).unsqueeze(1).contiguous().to(dev)
for prec in ("fp32", "fp16"):
    model.set_precision(prec)
    with torch.no_grad():
        for _ in range(3): m = model(x)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3): m = model(x)
        e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print(json.dumps({"precision": prec, "rows": 4096, "L": 5120, "ms_forward": ms, "waveforms_per_s": 4096 / ms * 1e3}), flush=True)
