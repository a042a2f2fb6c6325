#!/bin/bash
# usage: tools/gpu_round.sh <tag> [bench steps]   -- GPU tests + smoke + bench + chain timeline into gpurun_out/<tag>_*
tag=${1:-run}; steps=${2:-3}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log
tail -3 gpurun_out/${tag}_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py --steps $steps --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
cat gpurun_out/${tag}_bench.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e']['value'], d['kernel_ms_per_step'], d['roofline']['issued_frac'])"
timeout 120 python tools/chain_timeline.py 0 4 592 700 > gpurun_out/${tag}_tl_0_4.txt 2>&1
timeout 120 python tools/chain_timeline.py 8 4 592 700 > gpurun_out/${tag}_tl_8_4.txt 2>&1
head -1 gpurun_out/${tag}_tl_0_4.txt gpurun_out/${tag}_tl_8_4.txt
