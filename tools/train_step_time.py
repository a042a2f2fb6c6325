#!/usr/bin/env python
"""Device time of the training step (CUDA events around 20 steps after 5 warm-up steps): tools/train_step_time.py [rows] [L]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import stofnet_b200 as sb
from stofnet_b200.synth import synth_waveforms
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 256
L = int(sys.argv[2]) if len(sys.argv) > 2 else 2560
dev = torch.device('cuda:0')
sd = sb.load_checkpoint(os.path.join(ROOT, 'tests', 'golden', 'weights_different-armadillo.npz'))
model = sb.build_model(sd, dev); model.train()
tr = sb.Trainer(model, lr=5e-4, weight_decay=1e-8, lambda_value=1e-2)
x = torch.from_numpy(synth_waveforms(rows, L, seed=11, rf=10)).unsqueeze(1).to(dev)
gt = torch.from_numpy(np.random.default_rng(5).integers(100, 4 * L - 100, size=(rows, 1, 3))).to(dev)
for _ in range(5):
    tr.step(x, gt)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    loss = tr.step(x, gt)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
print("rows %d L %d: %.3f ms per step = %.1f rows/s, loss %.6g" % (rows, L, ms, rows / ms * 1e3, float(loss)))
