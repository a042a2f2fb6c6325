#!/bin/bash
# usage: tools/ab_env.sh <tag> VAR=a VAR=b ...  -- isolated chain-segment times and a short bench line per environment setting
tag=$1; shift
mkdir -p gpurun_out
for rep in 1 2; do
for kv in "$@"; do
  m=$(env $kv timeout 120 python tools/chain_timeline2.py 0 4 592 fp32 0 2>&1 | head -1)
  t=$(env $kv timeout 120 python tools/chain_timeline2.py 8 4 592 fp32 0 2>&1 | head -1)
  echo "$kv rep$rep mid: $m | tail: $t"
done
done 2>&1 | tee gpurun_out/${tag}_ab_segments.txt
for kv in "$@"; do
  name=$(echo $kv | tr '=' '_')
  env $kv timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-other-configs > gpurun_out/${tag}_bench_$name.json 2> gpurun_out/${tag}_bench_$name.err
  python -c "
import json; d=json.load(open('gpurun_out/${tag}_bench_$name.json')); print('$kv', round(d['value']), round(d['ms_per_step'],2), d['kernel_ms_per_step'], round(d['roofline']['issued_frac'],3), d['clocks']['sm_mhz'], d['parity_check']['mismatches'], d['parity_check']['boundary_flips'])"
done 2>&1 | tee gpurun_out/${tag}_ab_bench.txt
