#!/usr/bin/env python
"""One training step at a given batch (default: the reference's 4 rows x 2560) for a per-kernel launch list under ncu:
   ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv python tools/train_step_kernels.py [rows] [L]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import stofnet_b200 as sb
from stofnet_b200.synth import synth_waveforms
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 4
L = int(sys.argv[2]) if len(sys.argv) > 2 else 2560
dev = torch.device('cuda:0')
sd = sb.load_checkpoint(os.path.join(ROOT, 'tests', 'golden', 'weights_different-armadillo.npz'))
model = sb.build_model(sd, dev); model.train()
tr = sb.Trainer(model, lr=5e-4, weight_decay=1e-8, lambda_value=1e-2)
x = torch.from_numpy(synth_waveforms(rows, L, seed=11, rf=10)).unsqueeze(1).to(dev)
gt = torch.from_numpy(np.random.default_rng(5).integers(100, 4 * L - 100, size=(rows, 1, 3))).to(dev)
for _ in range(2):
    tr.step(x, gt)
torch.cuda.synchronize()
print("ok")
