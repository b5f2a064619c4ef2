#!/bin/bash
# split-K rule A/B: GEMM unit tests, per-site probe with the round-1 rule and the new one, then the default bench
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_trainer_gpu.py -m gpu -q -x -k "gemm or trainer or step" > gpurun_out/t24.log 2>&1
echo "rc=$?" >> gpurun_out/t24.log
VITB200_GEMM_SPLITS_OLD=1 timeout 300 python tools/gemm_sites.py > gpurun_out/sites24_old.log 2>&1
timeout 300 python tools/gemm_sites.py > gpurun_out/sites24_new.log 2>&1
VITB200_GEMM_SPLITS_OLD=1 timeout 300 python bench.py --no-cpu-baseline --no-ref-eager --no-graph > gpurun_out/bench24_old.json 2> gpurun_out/bench24_old.err
timeout 300 python bench.py --no-cpu-baseline --no-ref-eager --no-graph > gpurun_out/bench24_new.json 2> gpurun_out/bench24_new.err
