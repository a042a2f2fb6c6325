#!/bin/bash
# usage: tools/gpu_chain_ab.sh <tag>  -- chain-kernel correctness (tensor-path tests + goldens) and an A/B of the chain segments
# against the previous build (stofnet_b200/libstofnet_b200_base.so) on ONE box
tag=${1:-ab}
mkdir -p gpurun_out
timeout 120 python tools/chain_timeline2.py 0 4 592 fp32 120 > gpurun_out/${tag}_tl_mid.txt 2>&1; echo "timeline mid rc=$?"
head -1 gpurun_out/${tag}_tl_mid.txt
timeout 120 python tools/chain_timeline2.py 8 4 592 fp32 120 > gpurun_out/${tag}_tl_tail.txt 2>&1; echo "timeline tail rc=$?"
head -1 gpurun_out/${tag}_tl_tail.txt
timeout 600 python -m pytest tests/test_gpu_tc.py tests/test_gpu_parity.py -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/${tag}_pytest.log
tools/ab_chain.sh $tag stofnet_b200/libstofnet_b200_base.so stofnet_b200/libstofnet_b200.so
