#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py tests/test_trainer_gpu.py -x -q -m gpu -k "layernorm or block or model or trainer or step or oracle" > gpurun_out/t49.log 2>&1
echo "rc=$?" >> gpurun_out/t49.log
B="timeout 120 python bench.py --steps 12 --warmup 4 --no-cpu-baseline --no-ref-eager --no-graph --no-module-path --no-row0 --no-e2e"
VITB200_LNB_HOT=0 $B > gpurun_out/bench49_a.json 2> gpurun_out/bench49_a.err
$B > gpurun_out/bench49_b.json 2> gpurun_out/bench49_b.err
