#!/usr/bin/env python
"""Per-tensor difference of the tensor-core and the all-fp32 training step: tools/train_paths_diff.py ckpt N L"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import stofnet_b200 as sb
from stofnet_b200.synth import synth_waveforms
name, N, L = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
dev = torch.device("cuda:0")
sd = sb.load_checkpoint(os.path.join(ROOT, "tests", "golden", "weights_%s.npz" % name))
x = torch.from_numpy(synth_waveforms(N, L, seed=17, rf=10)).unsqueeze(1).to(dev)
gt = torch.from_numpy(np.random.default_rng(3).integers(50, 4 * L - 50, size=(N, 1, 4))).to(dev)
out = {}
for path in ("tc", "simt", "tc2"):
    model = sb.build_model(sd, dev).train()
    model._engine.set_path(sb._lib.PATH_SIMT if path == "simt" else sb._lib.PATH_TC)
    mask = model(x)
    loss = sb.mask_loss(mask, gt, 7, 1.0, 20.0, 1e-2)
    loss.backward()
    out[path] = (mask.detach().clone(), {k: p.grad.clone() for k, p in model.named_parameters()})
print("mask rel diff", float((out["tc"][0] - out["simt"][0]).abs().max() / out["simt"][0].abs().max()), "tc deterministic",
      all(torch.equal(out["tc"][1][k], out["tc2"][1][k]) for k in out["tc"][1]))
for k in out["simt"][1]:
    a, b = out["tc"][1][k], out["simt"][1][k]
    d = (a - b).abs().flatten().float()
    mx = float(b.abs().max())
    print("%-40s max|g| %.3e  maxdiff/max %.2e  q99.9/max %.2e  median/max %.2e" % (k, mx, float(d.max()) / mx, float(torch.quantile(d, 0.999)) / mx if d.numel() > 1000 else float("nan"), float(d.median()) / mx))
