#!/usr/bin/env python
"""CTA-pair chain kernel (conv_tc2.cu) against the one-CTA kernel (conv_tc.cu): bit-identical masks, and kernel times.
    python tools/pair_check.py [rows] [L]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import stofnet_b200 as sb
from stofnet_b200._lib import lib, check
from stofnet_b200.synth import synth_waveforms

dev = torch.device("cuda:0")
gold = os.path.join(ROOT, "tests", "golden")
for name, rf in (("graceful-snow", 20), ("clean-serenity", 10), ("proud-cherry", 10)):
    model = sb.build_model(sb.load_checkpoint(os.path.join(gold, "weights_%s.npz" % name)), dev)
    for (N, L) in ((1, 160), (3, 640), (5, 2578), (int(sys.argv[1]) if len(sys.argv) > 1 else 37, int(sys.argv[2]) if len(sys.argv) > 2 else 5120)):
        x = torch.from_numpy(synth_waveforms(N, L, seed=N, rf=rf)).unsqueeze(1).to(dev)
        for prec in ("fp32", "fp16"):
            model.set_precision(prec)
            outs = []
            for pair in (0, 1):
                with torch.no_grad():
                    model(x[:1])                                  # creates / syncs the handle
                    check(lib.stof_model_set_chain_kernel(model._engine.handle, pair))
                    outs.append(model(x).clone())
                torch.cuda.synchronize()
            same = torch.equal(outs[0], outs[1])
            print(name, N, L, prec, "identical" if same else "DIFFERENT max|d|=%g" % float((outs[0] - outs[1]).abs().max()), flush=True)
            assert torch.isfinite(outs[1]).all()
print("pair_check done")
