#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu -k "attention or noalloc" > gpurun_out/t41.log 2>&1
echo "rc=$?" >> gpurun_out/t41.log
B=256 timeout 120 python tools/attn_probe.py > gpurun_out/attn41.log 2>&1
B=256 timeout 120 python tools/attn_probe.py >> gpurun_out/attn41.log 2>&1
