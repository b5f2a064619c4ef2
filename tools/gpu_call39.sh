#!/bin/bash
# ncu --set full of the eight backward GEMM call sites of one block (final code), same command as the launch list
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-row0 --no-module-path --no-ref-eager --no-graph"
timeout 300 $CMD > gpurun_out/plain39.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 350 -c 8 -o gpurun_out/gemm_bwd_r02 -f $CMD > gpurun_out/ncu39.log 2>&1
echo "ncu rc=$?" >> gpurun_out/ncu39.log
