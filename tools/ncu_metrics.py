#!/usr/bin/env python
"""Print selected metrics of an .ncu-rep (raw page) by regex:  tools/ncu_metrics.py rep.ncu-rep 'regex' """
import csv, re, subprocess, sys
rep, pat = sys.argv[1], re.compile(sys.argv[2])
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[0]
for i, h in enumerate(hdr):
    if pat.search(h):
        print("%-90s %s" % (h, " | ".join(r[i] for r in rows[1:])))
