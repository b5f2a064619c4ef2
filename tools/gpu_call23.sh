#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_reference_gpu.py -m gpu -q -k "trajectory" > gpurun_out/t23.log 2>&1
echo "rc=$?" >> gpurun_out/t23.log
