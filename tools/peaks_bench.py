#!/usr/bin/env python
"""Stage-4 (nms + threshold + ToF compaction) alone at the bench workload's size: achieved HBM GB/s.

    python tools/peaks_bench.py [rows] [M] [iters]

Default 4096 rows x M = 20480 (C2: L = 5120, upsample_factor 4).  Three distinct masks are rotated (1 GB > 126 MB L2), so
every launch reads its rows from HBM.  Algorithmic bytes = 4 M per row (read once with a float threshold; the row-max mode
reads each row twice, the second time from L2 - reported against the same 4 M)."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import stofnet_b200 as sb
from stofnet_b200._lib import lib, check
from stofnet_b200.engine import stream_ptr

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
M = int(sys.argv[2]) if len(sys.argv) > 2 else 20480
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 30
dev = torch.device("cuda:0")
g = torch.Generator(device="cuda").manual_seed(1)
# mask-like scores: small noise with a few echoes per row
masks = []
for i in range(3):
    m = torch.randn(rows, M, device=dev, generator=g) * 0.004
    idx = torch.randint(0, M, (rows, 3), device=dev, generator=g)
    m.scatter_(1, idx, torch.rand(rows, 3, device=dev, generator=g) * 5 + 1)
    masks.append(m)
cap = 64
coords = torch.empty((rows, cap), dtype=torch.float32, device=dev)
counts = torch.empty((rows,), dtype=torch.int32, device=dev)
stats = torch.zeros((2,), dtype=torch.int64, device=dev)
peak = None
try:
    peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
for name, none, th in (("th=0.5", 0, 0.5), ("th=None (row max)", 1, 0.0)):
    def run(m):
        stats.zero_()
        check(lib.stof_peaks(m.data_ptr(), rows, M, 20, none, C.c_float(th), 4, coords.data_ptr(), None, cap,
                             counts.data_ptr(), stats.data_ptr(), stream_ptr(dev)))
    for i in range(3):
        run(masks[i % 3])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        run(masks[i % 3])
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    gbs = rows * M * 4 / (ms * 1e-3) / 1e9
    print(json.dumps({"kernel": "peaks_kernel", "mode": name, "rows": rows, "M": M, "ms": round(ms, 4),
                      "achieved_gbs": round(gbs, 1), "peak_gbs": peak, "frac": round(gbs / peak, 3) if peak else None,
                      "kmax": int(stats[0].item()), "rows_per_s": round(rows / (ms * 1e-3))}))
