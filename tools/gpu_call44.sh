#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_trainer_gpu.py -x -q -m gpu > gpurun_out/t44.log 2>&1
echo "rc=$?" >> gpurun_out/t44.log
timeout 400 python bench.py > gpurun_out/bench44.json 2> gpurun_out/bench44.err
echo "bench rc=$?" >> gpurun_out/bench44.err
