#!/usr/bin/env python
"""Small fixed case for ncu: a few forward + peaks passes of the bench workload's shape on fewer rows.

    python tools/prof_case.py [rows] [L] [passes]

Default 296 rows x L=5120 (two persistent waves of the chain kernel), graceful-snow, th=0.015."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import stofnet_b200 as sb
from stofnet_b200.synth import synth_waveforms

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 296
L = int(sys.argv[2]) if len(sys.argv) > 2 else 5120
passes = int(sys.argv[3]) if len(sys.argv) > 3 else 2
dev = torch.device("cuda:0")
sd = sb.load_checkpoint(os.path.join(ROOT, "tests", "golden", "weights_graceful-snow.npz"))
model = sb.build_model(sd, dev)
x = torch.from_numpy(synth_waveforms(rows, L, seed=3008, rf=20)).unsqueeze(1).to(dev)
with torch.no_grad():
    for _ in range(passes):
        mask = model(x)
        coords = sb.mask2coords(mask, 20, 0.015, upsample_factor=4)
torch.cuda.synchronize()
print("prof_case ok", tuple(mask.shape), tuple(coords.shape), "launches", sb._lib.lib.stof_launch_count())
