#!/usr/bin/env python
"""Run one chain segment on random planes and print CTA 0's event timeline (cycles)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import stofnet_b200 as sb
from stofnet_b200._lib import lib, check
first, count = int(sys.argv[1]), int(sys.argv[2])
N, L = int(sys.argv[3]) if len(sys.argv) > 3 else 592, 5120
dev = torch.device("cuda:0")
gold = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
model = sb.build_model(sb.load_checkpoint(os.path.join(gold, "weights_graceful-snow.npz")), dev)
x = (torch.randn(N, 2, 8, L, 8, device=dev)).half()
x0 = x.clone()
eng = model._engine
for _ in range(2):
    eng.chain_segment(model, first, count, x, x0)
torch.cuda.synchronize()
cap = 40000
tl = torch.zeros(4 + 2 * cap, dtype=torch.int64, device=dev)
check(lib.stof_debug_set_timeline(tl.data_ptr(), cap))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); eng.chain_segment(model, first, count, x, x0); e1.record()
torch.cuda.synchronize()
check(lib.stof_debug_set_timeline(None, 0))
t = tl.cpu().numpy()
capr = cap // 4
ev = []
for role in range(4):
    for i in range(min(int(t[role]), capr)):
        o = 4 + 2 * (role * capr + i)
        ev.append((int(t[o + 1]), int(t[o])))
ev.sort()
n = len(ev)
print("kernel ms", e0.elapsed_time(e1), "events", n)
t0 = ev[0][0]
names = {0: "PROD", 1: "MMA0", 2: "MMA1", 3: "EPI"}
lim = int(sys.argv[4]) if len(sys.argv) > 4 else 400
for c, code in ev[:lim]:
    role, k, b, what = code >> 24, (code >> 16) & 255, (code >> 8) & 255, code & 255
    print("%9d %s k=%d b=%d what=%d" % (c - t0, names.get(role, "?"), k, b, what))
