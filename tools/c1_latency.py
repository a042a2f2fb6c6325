#!/usr/bin/env python
"""C1 (one waveform, L = 2560, host buffers in / ToF table out): latency of stof_infer_host and of the device-side pieces.
   python tools/c1_latency.py            (under ncu --metrics gpu__time_duration.sum for the per-kernel list)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import stofnet_b200 as sb
from stofnet_b200.synth import synth_waveforms
dev = torch.device("cuda:0")
sd = sb.load_checkpoint(os.path.join(ROOT, "tests", "golden", "weights_different-armadillo.npz"))
model = sb.build_model(sd, dev)
x = torch.from_numpy(synth_waveforms(1, 2560, seed=3, rf=10)).unsqueeze(1).pin_memory()
for _ in range(5):
    out = model._engine.infer_host(model, x, 20, None)
ts = []
for _ in range(50):
    t0 = time.perf_counter(); out = model._engine.infer_host(model, x, 20, None); ts.append(time.perf_counter() - t0)
print("infer_host N=1 L=2560: median %.1f us, min %.1f us" % (np.median(ts) * 1e6, np.min(ts) * 1e6))
xd = x.to(dev)
with torch.no_grad():
    for _ in range(3): m = model(xd)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): m = model(xd)
    e1.record(); torch.cuda.synchronize()
    print("device forward only: %.1f us" % (e0.elapsed_time(e1) / 20 * 1e3))
