#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py -x -q -m gpu -k "layernorm or block or model or fp32 or bf16" > gpurun_out/t47.log 2>&1
echo "rc=$?" >> gpurun_out/t47.log
B="timeout 300 python bench.py --no-cpu-baseline --no-ref-eager --no-graph --no-module-path --no-row0 --no-e2e"
VITB200_LNB_EXACT=0 $B > gpurun_out/bench47_a.json 2> gpurun_out/bench47_a.err
$B > gpurun_out/bench47_b.json 2> gpurun_out/bench47_b.err
VITB200_LNB_EXACT=0 $B > gpurun_out/bench47_a2.json 2>> gpurun_out/bench47_a.err
$B > gpurun_out/bench47_b2.json 2>> gpurun_out/bench47_b.err
