#!/usr/bin/env python
"""Layer-granularity timeline of CTA 0 of one chain segment (few events: low perturbation).
usage: chain_timeline2.py first count rows precision"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import stofnet_b200 as sb
from stofnet_b200._lib import lib, check
first, count, N = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
prec = sys.argv[4] if len(sys.argv) > 4 else "fp32"
L = 5120
dev = torch.device("cuda:0")
gold = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
model = sb.build_model(sb.load_checkpoint(os.path.join(gold, "weights_graceful-snow.npz")), dev)
model.set_precision(prec)
model._engine.sync(model, dev); model._engine._apply_path()
x = (torch.randn(N, 2, 8, L, 8, device=dev)).half()
x0 = x.clone()
eng = model._engine
def run():
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); eng.chain_segment(model, first, count, x, x0); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)
for _ in range(2): run()
plain = run()
cap = 40000
tl = torch.zeros(4 + 2 * cap, dtype=torch.int64, device=dev)
check(lib.stof_debug_set_timeline(tl.data_ptr(), cap))
timed = run()
check(lib.stof_debug_set_timeline(None, 0))
t = tl.cpu().numpy(); capr = cap // 4
ev = []
waits = {}
for role in range(4):
    for i in range(min(int(t[role]), capr)):
        o = 4 + 2 * (role * capr + i)
        code, val = int(t[o]), int(t[o + 1])
        if (code & 255) in (40, 41):          # per-layer wait cycles of an MMA issuer (activations / weight stages)
            waits.setdefault((role, (code >> 16) & 255, code & 255), []).append(val)
        else:
            ev.append((val, code))
ev.sort()
print("precision", prec, "kernel ms plain", plain, "with timeline", timed, "events", len(ev))
for (role, k, what), v in sorted(waits.items()):
    v = v[len(v) // 4:]
    print("# issuer %d layer %d: mean cycles waiting for %s = %.0f (over %d tiles)" % (role - 1, k, "activations" if what == 40 else "weights", sum(v) / max(len(v), 1), len(v)))
t0 = ev[0][0]
names = {0: "PROD", 1: "MMA0", 2: "MMA1", 3: "EPI"}
# skip the first tile: print tiles 3-4 of CTA 0
lim = int(sys.argv[5]) if len(sys.argv) > 5 else 160
start = 0
cnt = 0
for i, (c, code) in enumerate(ev):
    if (code >> 24) == 1 and (code & 255) == 1:
        cnt += 1
        if cnt == 4: start = i; break
for c, code in ev[start:start + lim]:
    role, k, b, what = code >> 24, (code >> 16) & 255, (code >> 8) & 255, code & 255
    print("%9d %s k=%d b=%d what=%d" % (c - ev[start][0], names.get(role, "?"), k, b, what))
