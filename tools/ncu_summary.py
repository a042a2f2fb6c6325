#!/usr/bin/env python
"""Table of the headline metrics of every kernel in an .ncu-rep (kernels as columns):  tools/ncu_summary.py rep.ncu-rep"""
import csv
import subprocess
import sys

METRICS = """gpu__time_duration.sum sm__cycles_elapsed.max launch__grid_size launch__block_size launch__registers_per_thread
launch__shared_mem_per_block_dynamic sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active
sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed sm__throughput.avg.pct_of_peak_sustained_elapsed
sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active smsp__issue_active.avg.pct_of_peak_sustained_active
sm__warps_active.avg.pct_of_peak_sustained_active smsp__inst_executed.sum dram__bytes_read.sum dram__bytes_write.sum
dram__bytes.sum.per_second gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed lts__t_sectors_srcunit_tex_op_read.sum
lts__t_sector_hit_rate.pct l1tex__data_pipe_lsu_wavefronts_mem_shared.sum l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum""".split()

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
kcol = hdr.index("Kernel Name")
names = [r[kcol].split("(")[0][:18] for r in data]
print("%-78s %-16s | %s" % ("metric", "unit", " | ".join("%-18s" % n for n in names)))
for m in METRICS:
    if m in hdr:
        i = hdr.index(m)
        print("%-78s %-16s | %s" % (m, units[i], " | ".join("%-18s" % r[i] for r in data)))
