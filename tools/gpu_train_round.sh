#!/bin/bash
# usage: tools/gpu_train_round.sh <tag> [ncu kernel regex] -- training tests, step time at 256 x 2560 and 4 x 2560, per-kernel launch
# list under ncu, and (optionally) one ncu --set full capture of the named kernels of the 256 x 2560 step
tag=${1:-run}; mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_train.py tests/test_sharded.py -x -q -m gpu 2>&1 | tail -4
timeout 200 python tools/train_step_time.py 256 2560
timeout 200 python tools/train_step_time.py 4 2560
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${tag}_train_launches_256x2560.csv python tools/train_step_kernels.py 256 2560 > gpurun_out/${tag}_ncu.log 2>&1; echo ncu rc=$?
if [ -n "$2" ]; then
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:"$2" -c 4 -f -o gpurun_out/${tag}_full python tools/train_step_kernels.py 256 2560 > gpurun_out/${tag}_ncu_full.log 2>&1; echo ncu full rc=$?
fi
