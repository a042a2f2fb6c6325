#!/bin/bash
# usage: tools/ab_chain.sh <tag> <lib A> <lib B> ...   -- isolated chain-segment kernel times (mid segment, tail segment) and a short
# bench line per library build on ONE box (A/B of kernel variants; libraries are selected with STOFNET_B200_LIB)
tag=$1; shift
mkdir -p gpurun_out
for rep in 1 2; do
for lib in "$@"; do
  name=$(basename $lib .so)
  export STOFNET_B200_LIB=$PWD/$lib
  m=$(timeout 120 python tools/chain_timeline2.py 0 4 592 fp32 0 2>&1 | head -1)
  t=$(timeout 120 python tools/chain_timeline2.py 8 4 592 fp32 0 2>&1 | head -1)
  echo "$name rep$rep mid: $m | tail: $t"
done
done 2>&1 | tee gpurun_out/${tag}_ab_segments.txt
for lib in "$@"; do
  name=$(basename $lib .so)
  export STOFNET_B200_LIB=$PWD/$lib
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-other-configs > gpurun_out/${tag}_bench_$name.json 2> gpurun_out/${tag}_bench_$name.err
  python -c "
import json; d=json.load(open('gpurun_out/${tag}_bench_$name.json')); print('$name', round(d['value']), round(d['ms_per_step'],2), d['kernel_ms_per_step'], round(d['roofline']['issued_frac'],3), d['clocks']['sm_mhz'], d['parity_check']['mismatches'], d['parity_check']['boundary_flips'])"
done 2>&1 | tee gpurun_out/${tag}_ab_bench.txt
