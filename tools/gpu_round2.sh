#!/bin/bash
# usage: tools/gpu_round2.sh <tag> [bench steps] -- GPU tests + smoke + bench (exact, fp16) + ncu launch list + ncu full set of the hot kernels
tag=${1:-run}; steps=${2:-5}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${tag}_pytest_gpu.log
tail -3 gpurun_out/${tag}_pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py --steps $steps --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --steps $steps --warmup 3 --precision fp16 --no-cpu-baseline > gpurun_out/${tag}_bench_fp16.json 2> gpurun_out/${tag}_bench_fp16.err; echo "bench fp16 rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2> gpurun_out/${tag}_bench_reference.err; echo "bench ref rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/${tag}_ncu_launches.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'chain_tc|contract_tc|peaks|expand|x0_split' -c 10 -f -o gpurun_out/${tag}_full python tools/prof_case.py 296 5120 1 > gpurun_out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out
