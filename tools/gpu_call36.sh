#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_trainer_gpu.py -x -q -m gpu -k "fused_head_dot or row0 or gemm" > gpurun_out/t36.log 2>&1
echo "rc=$?" >> gpurun_out/t36.log
