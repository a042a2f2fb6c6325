#!/usr/bin/env python
"""ESPCN_1D at the C2 shape (BASELINE config 2 names the ESPCN checkpoint wobbly-sponge): waveforms/s of
forward + mask2coords on device-resident batches, CUDA-event timed.

    python tools/espcn_bench.py [rows] [L] [steps] [tc|ffma]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import stofnet_b200 as sb
from stofnet_b200.synth import synth_waveforms

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
L = int(sys.argv[2]) if len(sys.argv) > 2 else 5120
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
dev = torch.device("cuda:0")
sd = sb.load_checkpoint(os.path.join(ROOT, "tests", "golden", "weights_wobbly-sponge.npz"))
model = sb.build_model(sd, dev)
path = sys.argv[4] if len(sys.argv) > 4 else "tc"
model._engine.set_path(sb._lib.PATH_TC if path == "tc" else sb._lib.PATH_SIMT, 0 if path == "tc" else 4)
base = torch.from_numpy(synth_waveforms(256, L, seed=3008, rf=20))
xs = [torch.roll(base.repeat((rows + 255) // 256, 1)[:rows], s, 0).unsqueeze(1).contiguous().to(dev) for s in (0, 7, 13)]
with torch.no_grad():
    for i in range(3):
        mask = model(xs[i % 3])
        co = sb.mask2coords(mask, 20, 0.2, upsample_factor=4)
    torch.cuda.synchronize()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record()
    for i in range(steps):
        mask = model(xs[i % 3])
    e1.record()
    for i in range(steps):
        co = sb.mask2coords(mask, 20, 0.2, upsample_factor=4)
    e2.record()
    torch.cuda.synchronize()
fwd, pk = e0.elapsed_time(e1) / steps, e1.elapsed_time(e2) / steps
print(json.dumps({"model": "ESPCN_1D wobbly-sponge", "path": path, "rows": rows, "L": L, "forward_ms": round(fwd, 3), "mask2coords_ms": round(pk, 3),
                  "waveforms_per_s": round(rows / ((fwd + pk) * 1e-3)), "kmax": int(co.shape[-1])}))
