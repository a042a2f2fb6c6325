#!/usr/bin/env python
"""Host-side profile of one training step at the reference's batch size (config.yaml:11 batch_size 4): step time, the
weight re-upload (engine.sync) alone, and a cProfile of ten steps."""
import sys, time, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import stofnet_b200 as sb
from stofnet_b200.synth import synth_waveforms
dev = torch.device('cuda:0')
sd = sb.load_checkpoint(os.path.join(ROOT, 'tests', 'golden', 'weights_different-armadillo.npz'))
model = sb.build_model(sd, dev); model.train()
tr = sb.Trainer(model, lr=5e-4, weight_decay=1e-8, lambda_value=1e-2)
rows, L = 4, 2560
x = torch.from_numpy(synth_waveforms(rows, L, seed=11, rf=10)).unsqueeze(1).to(dev)
gt = torch.from_numpy(np.random.default_rng(5).integers(100, 4*L-100, size=(rows,1,3))).to(dev)
for _ in range(3): tr.step(x, gt)
torch.cuda.synchronize()
eng = model._engine
t0=time.perf_counter()
for _ in range(10): tr.step(x, gt)
torch.cuda.synchronize(); t1=time.perf_counter()
print('step ms', (t1-t0)/10*1e3)
# sync alone
ts=[]
for _ in range(10):
    eng.key=None; torch.cuda.synchronize(); a=time.perf_counter(); eng.sync(model, dev); torch.cuda.synchronize(); ts.append(time.perf_counter()-a)
print('sync ms', np.mean(ts)*1e3)
import cProfile, pstats
pr=cProfile.Profile(); pr.enable()
for _ in range(10): tr.step(x, gt)
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats('cumulative').print_stats(18)
