#!/usr/bin/env python
"""Top SASS instructions of a kernel by warp-stall samples: tools/ncu_hot.py rep.ncu-rep kernel_regex [topN]"""
import csv, subprocess, sys
rep, k = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-id", "::regex:%s:1" % k],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hi + 1:] if len(r) == len(hdr) and r[0] != "Address"]
tot = sum(int(r[col["# Samples"]] or 0) for r in data)
print("total samples", tot, "instructions", len(data))
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
order = sorted(range(len(data)), key=lambda i: -int(data[i][col["# Samples"]] or 0))[:top]
for i in sorted(order):
    r = data[i]
    s = int(r[col["# Samples"]] or 0)
    why = sorted(((int(r[col[h]] or 0), h[6:]) for h in stalls), reverse=True)[:2]
    print("%5d %5.1f%% exec=%-9s %-70s %s" % (i, 100.0 * s / max(tot, 1), r[col["Instructions Executed"]], r[col["Source"]][:70],
                                          " ".join("%s:%d" % (n, c) for c, n in why if c)))
