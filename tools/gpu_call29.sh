#!/bin/bash
# LayerNorm forward ring kernel: unit tests, model parity, in-step A/B
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py tests/test_fullsize_gpu.py -x -q -m gpu > gpurun_out/t29.log 2>&1
echo "rc=$?" >> gpurun_out/t29.log
B="timeout 300 python bench.py --no-cpu-baseline --no-ref-eager --no-graph --no-module-path --no-row0 --no-e2e"
VITB200_LN_FWD_RING=0 $B > gpurun_out/bench29_reg.json 2> gpurun_out/bench29_reg.err
$B > gpurun_out/bench29_ring.json 2> gpurun_out/bench29_ring.err
VITB200_LN_FWD_RING=0 $B > gpurun_out/bench29_reg2.json 2>> gpurun_out/bench29_reg.err
$B > gpurun_out/bench29_ring2.json 2>> gpurun_out/bench29_ring.err
