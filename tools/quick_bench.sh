#!/bin/bash
# usage: tools/quick_bench.sh <tag>  -- GPU parity tests + exact / fp16 bench lines (no CPU baseline) into gpurun_out/<tag>_*
tag=${1:-q}
timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -12
for prec in fp32 fp16; do
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --precision $prec > gpurun_out/${tag}_bench_$prec.json 2>gpurun_out/${tag}_bench_$prec.err
python -c "
import json; d=json.load(open('gpurun_out/${tag}_bench_$prec.json')); print('$prec', round(d['value']), round(d['ms_per_step'],2), d['kernel_ms_per_step'], round(d['roofline']['issued_frac'],3), d['clocks']['sm_mhz'], d['clocks']['reasons'])"
done
