#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_trainer_gpu.py -x -q -m gpu -k "fused_head_dot or row0 or gemm" > gpurun_out/t37.log 2>&1
echo "rc=$?" >> gpurun_out/t37.log
NOREF=1 SITES="bwd dO" timeout 120 python tools/gemm_sites.py > gpurun_out/sites37.log 2>&1
NOREF=1 SITES="bwd dO" VITB200_GEMM_EPI_FLAGS=8 timeout 120 python tools/gemm_sites.py >> gpurun_out/sites37.log 2>&1
B=256 timeout 120 python tools/attn_probe.py > gpurun_out/attn37.log 2>&1
