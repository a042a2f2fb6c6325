#!/usr/bin/env python
"""Top stall lines of one kernel from `ncu -i rep --page source --csv`: tools/ncu_src_top.py file.csv kernel_substring [n]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2]; topn = int(sys.argv[3]) if len(sys.argv) > 3 else 40
starts = [i for i, r in enumerate(rows) if r and r[0] == 'Kernel Name']
for si, s0 in enumerate(starts):
    if want not in rows[s0][1]: continue
    end = starts[si + 1] if si + 1 < len(starts) else len(rows)
    h = rows[s0 + 1]
    src = h.index('Source'); samp = h.index('# Samples'); ie = h.index('Instructions Executed')
    keys = [k for k in h if k.startswith('stall_') and 'Not Issued' not in k]
    cols = {k: h.index(k) for k in keys}
    data = [r for r in rows[s0 + 2:end] if len(r) > max(cols.values())]
    tot = sum(int(r[samp] or 0) for r in data)
    print('kernel', rows[s0][1][:60], 'total samples', tot, 'sass lines', len(data), 'inst', sum(int(r[ie] or 0) for r in data))
    idx = sorted(range(len(data)), key=lambda i: -int(data[i][samp] or 0))[:topn]
    for i in sorted(idx):
        r = data[i]
        st = ' '.join('%s=%s' % (k.replace('stall_', ''), r[c]) for k, c in cols.items() if r[c] not in ('0', ''))
        print(i, r[samp], r[ie], r[src][:60], '|', st)
    break
